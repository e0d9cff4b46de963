"""Synthetic scenes for BASELINE.json configs[2] and configs[4], derived from the
arm_around_table fixture.  The reference ships no such fixtures; these are generated deterministically from a seed
and are run through the oracle and the CUDA path alike.

  cluttered_scene(seed)  C3: the PR2 right arm among 64 random convex hulls (<= 32 vertices each), T = 50
  dual_arm_scene(seed)   C5: two copies of the arm chain (the second mirrored in y and shifted by +0.6 m) riding on a
                         2-axis positioner (revolute z, then revolute y), n_dof = 16, every arm link split into two
                         convex sub-shapes (compound links: shape_id 0 / 1), the 64 hulls and the table as obstacles
"""
import numpy as np
from scipy.spatial import ConvexHull

from tests import helpers

capi = helpers.capi

IDENT = np.array([1.0, 0, 0, 0, 0, 1.0, 0, 0, 0, 0, 1.0, 0])


def _pose(R=None, t=(0, 0, 0)):
    T = np.zeros((3, 4))
    T[:, :3] = np.eye(3) if R is None else R
    T[:, 3] = t
    return T.reshape(12)


def random_hulls(rng, n_hulls, box_lo, box_hi, max_verts=32):
    """hull k = convex hull of 64 points N(0, (0.05 s)^2), s ~ U(0.5, 2), pruned to <= max_verts vertices,
    centre U(box)"""
    out = []
    for _ in range(n_hulls):
        s = rng.uniform(0.5, 2.0)
        pts = rng.normal(0.0, 0.05 * s, (64, 3))
        hv = pts[ConvexHull(pts).vertices]
        while len(hv) > max_verts:  # drop the vertex closest to the centroid and re-hull
            d = np.linalg.norm(hv - hv.mean(0), axis=1)
            hv = np.delete(hv, int(np.argmin(d)), 0)
            hv = hv[ConvexHull(hv).vertices]
        out.append((hv, rng.uniform(box_lo, box_hi)))
    return out


def _append_static_hulls(s, hulls):
    """adds one static link + one hull shape per entry; ACM: checked against every active link"""
    n_links0 = len(s["link_frame"])
    n_new = len(hulls)
    s["link_frame"] = np.concatenate([s["link_frame"], -np.ones(n_new, np.int32)])
    s["link_pose"] = np.concatenate([s["link_pose"], np.stack([_pose(t=c) for _, c in hulls])])
    s["link_names"] = np.concatenate([s["link_names"], np.array(["hull_%d" % i for i in range(n_new)])])
    voff = len(s["verts"])
    for i, (hv, _) in enumerate(hulls):
        s["shape_type"] = np.append(s["shape_type"], np.int32(1))
        s["shape_link"] = np.append(s["shape_link"], np.int32(n_links0 + i))
        s["shape_id"] = np.append(s["shape_id"], np.int32(0))
        s["shape_params"] = np.concatenate([s["shape_params"], np.zeros((1, 4))])
        s["shape_pose"] = np.concatenate([s["shape_pose"], IDENT[None]])
        s["shape_vert_offset"] = np.append(s["shape_vert_offset"], np.int32(voff))
        s["shape_vert_count"] = np.append(s["shape_vert_count"], np.int32(len(hv)))
        s["verts"] = np.concatenate([s["verts"], hv])
        voff += len(hv)
    n = n_links0 + n_new
    acm = np.ones((n, n), np.uint8)
    acm[:n_links0, :n_links0] = s["acm"]
    na = int(s["n_active_links"])
    acm[:na, n_links0:] = 0
    acm[n_links0:, :na] = 0
    s["acm"] = acm
    return s


def cluttered_scene(seed=3):
    rng = np.random.default_rng(seed)
    s = {k: np.array(v) for k, v in helpers.load_scene("arm_around_table").items()}
    hulls = random_hulls(rng, 64, np.array([0.2, -1.0, 0.3]), np.array([1.2, 0.2, 1.3]))
    s = _append_static_hulls(s, hulls)
    # T = 50: start / goal of the fixture path through 48 interior points = linear + N(0, 0.05^2) joint noise
    w = np.linspace(0, 1, 50)[:, None]
    traj = s["seed_traj"][0] * (1 - w) + s["seed_traj"][-1] * w
    traj[1:-1] += rng.normal(0, 0.05, (48, traj.shape[1]))
    lim = s["joint_limits"]
    s["traj"] = np.clip(traj, lim[:, 0], lim[:, 1])
    return s


def _split_hull(pts):
    """two convex sub-shapes of a vertex set: halves along the longest extent, sharing the middle slab"""
    ax = int(np.argmax(pts.max(0) - pts.min(0)))
    order = np.argsort(pts[:, ax], kind="stable")
    n = len(order)
    a, b = pts[order[: n // 2 + 2]], pts[order[n // 2 - 2:]]
    return a[ConvexHull(a).vertices], b[ConvexHull(b).vertices]


def dual_arm_scene(seed=5):
    rng = np.random.default_rng(seed)
    a = {k: np.array(v) for k, v in helpers.load_scene("arm_around_table").items()}
    nf, na, nl = len(a["frame_parent"]), int(a["n_active_links"]), len(a["link_frame"])
    nd = int(a["n_dof"])
    mirror = np.diag([1.0, -1.0, 1.0])

    def mirrored_pose(p):
        T = p.reshape(3, 4)
        return _pose(mirror @ T[:, :3] @ mirror, mirror @ T[:, 3])

    # frames: 0 = positioner z, 1 = positioner y, then arm A (offset 2), then arm B (offset 2 + nf)
    fp = [-1, 0] + [2 + p if p >= 0 else 1 for p in a["frame_parent"]] + [2 + nf + p if p >= 0 else 1 for p in a["frame_parent"]]
    ft = [1, 1] + list(a["frame_joint_type"]) * 2
    fd = [0, 1] + [2 + d if d >= 0 else -1 for d in a["frame_dof"]] + [2 + nd + d if d >= 0 else -1 for d in a["frame_dof"]]
    fax = np.concatenate([[[0, 0, 1.0], [0, 1.0, 0]], a["frame_axis"], a["frame_axis"] @ mirror])
    forig_a = a["frame_origin"].copy()
    forig_b = np.stack([mirrored_pose(p) for p in a["frame_origin"]])
    base = a["base_pose"].reshape(3, 4)
    # the arm roots hang off the positioner: A at the fixture's base offset, B mirrored and shifted by +0.6 in y
    root_a, root_b = forig_a[0].reshape(3, 4).copy(), forig_b[0].reshape(3, 4).copy()
    root_a[:, 3] = base[:, :3] @ root_a[:, 3] + base[:, 3] - np.array([0, 0, 0.4])
    root_b[:, 3] = mirror @ (root_a[:, 3]) + np.array([0, 0.6, 0])
    forig_a[0], forig_b[0] = root_a.reshape(12), root_b.reshape(12)
    forig = np.concatenate([[_pose(t=(0, 0, 0.2)), _pose(t=(0, 0, 0.2))], forig_a, forig_b])
    s = dict(n_dof=np.int32(2 + 2 * nd), frame_parent=np.array(fp, np.int32), frame_joint_type=np.array(ft, np.int32),
             frame_dof=np.array(fd, np.int32), frame_axis=fax, frame_origin=forig, base_pose=IDENT.copy())
    lim = np.concatenate([[[-0.5, 0.5], [-0.3, 0.3]], a["joint_limits"], a["joint_limits"]])
    s["joint_limits"] = np.clip(lim, -np.pi, np.pi)
    # links: active = positioner table (frame 1), arm A links, arm B links; static = table_link only + hulls later
    table = int(np.where(a["link_names"] == "table_link")[0][0])
    link_frame = [1] + [2 + f for f in a["link_frame"][:na]] + [2 + nf + f for f in a["link_frame"][:na]] + [-1]
    names = ["positioner"] + ["a_" + n for n in a["link_names"][:na]] + ["b_" + n for n in a["link_names"][:na]] + ["table_link"]
    link_pose = [IDENT] * (1 + 2 * na) + [a["link_pose"][table]]
    shape_type, shape_link, shape_id, shape_params, shape_pose, voff, vcnt, verts = [], [], [], [], [], [], [], []

    def add_hull(link, sid, pts, pose=IDENT):
        shape_type.append(1), shape_link.append(link), shape_id.append(sid), shape_params.append(np.zeros(4))
        shape_pose.append(pose), voff.append(sum(len(v) for v in verts)), vcnt.append(len(pts)), verts.append(pts)

    add_hull(0, 0, np.array([[x, y, z] for x in (-0.15, 0.15) for y in (-0.15, 0.15) for z in (-0.02, 0.02)]))
    for arm, flip in ((0, False), (1, True)):
        for l in range(na):
            sh = int(np.where(a["shape_link"] == l)[0][0])
            pts = a["verts"][a["shape_vert_offset"][sh]: a["shape_vert_offset"][sh] + a["shape_vert_count"][sh]]
            sp = a["shape_pose"][sh]
            if flip:
                pts, sp = pts @ mirror, mirrored_pose(sp)
            if a["shape_type"][sh] != 1:
                raise ValueError("expected hull shapes on the arm links")
            for sid, part in enumerate(_split_hull(pts)):
                add_hull(1 + arm * na + l, sid, part, sp)
    tsh = int(np.where(a["shape_link"] == table)[0][0])
    add_hull(1 + 2 * na, 0, a["verts"][a["shape_vert_offset"][tsh]: a["shape_vert_offset"][tsh] + a["shape_vert_count"][tsh]],
             a["shape_pose"][tsh])
    n_links = 2 + 2 * na
    acm = np.zeros((n_links, n_links), np.uint8)
    for arm in (0, 1):  # each arm keeps the fixture's self-collision exclusions
        o = 1 + arm * na
        acm[o:o + na, o:o + na] = a["acm"][:na, :na]
    acm[0, 1:1 + 2 * na] = acm[1:1 + 2 * na, 0] = 1  # the positioner plate carries the arms
    np.fill_diagonal(acm, 1)
    s.update(link_names=np.array(names), link_frame=np.array(link_frame, np.int32), link_pose=np.stack(link_pose),
             n_active_links=np.int32(1 + 2 * na), shape_type=np.array(shape_type, np.int32),
             shape_link=np.array(shape_link, np.int32), shape_id=np.array(shape_id, np.int32),
             shape_params=np.stack(shape_params), shape_pose=np.stack(shape_pose),
             shape_vert_offset=np.array(voff, np.int32), shape_vert_count=np.array(vcnt, np.int32),
             verts=np.concatenate(verts), acm=acm)
    s = _append_static_hulls(s, random_hulls(rng, 64, np.array([0.1, -0.9, 0.4]), np.array([1.1, 1.5, 1.6])))
    # T = 100 smooth path: cubic through 5 knots drawn inside the (clipped) limits
    knots = rng.uniform(0.4 * s["joint_limits"][:, 0], 0.4 * s["joint_limits"][:, 1], (5, 2 + 2 * nd))
    tk, tt = np.linspace(0, 1, 5), np.linspace(0, 1, 100)
    coef = [np.polyfit(tk, knots[:, j], 3) for j in range(knots.shape[1])]
    s["traj"] = np.stack([np.polyval(c, tt) for c in coef], axis=1)
    return s


def primitives_scene():
    """spherebot's kinematics (two prismatic joints) carrying a capsule and a small box hull; obstacles: cylinder, cone,
    sphere, capsule and a box hull, each with a tilted pose — every shape type of include/trajopt_cuda.h meets every
    other through the out-of-line primitive support path"""
    s = {k: np.array(v) for k, v in helpers.load_scene("spherebot").items()}
    rng = np.random.default_rng(21)

    def tilt(t):
        a = rng.normal(size=3)
        a /= np.linalg.norm(a)
        th = rng.uniform(0.2, 1.2)
        K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
        return _pose(np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K, t)

    box = np.array([[x, y, z] for x in (-0.2, 0.2) for y in (-0.3, 0.3) for z in (-0.25, 0.25)], dtype=np.float64)
    n_links = 6  # 0 = robot link, 1..5 obstacles
    s["link_names"] = np.array(["bot", "cyl", "cone", "ball", "cap", "crate"])
    s["link_frame"] = np.array([2, -1, -1, -1, -1, -1], np.int32)
    s["link_pose"] = np.stack([IDENT, _pose(t=(0.0, 0.0, 0.0)), _pose(t=(-1.2, 0.3, 0.1)), _pose(t=(0.4, 1.3, -0.1)),
                               _pose(t=(1.4, -0.6, 0.2)), _pose(t=(-0.5, -1.4, 0.0))]).reshape(n_links, 12)
    #            type                 link  id  params (r, half length)
    shapes = [(capi.SHAPE_CAPSULE, 0, 0, (0.25, 0.4)), (capi.SHAPE_HULL, 0, 1, (0, 0)),
              (capi.SHAPE_CYLINDER, 1, 0, (0.35, 0.5)), (capi.SHAPE_CONE, 2, 0, (0.4, 0.45)),
              (capi.SHAPE_SPHERE, 3, 0, (0.45, 0)), (capi.SHAPE_CAPSULE, 4, 0, (0.2, 0.6)), (capi.SHAPE_HULL, 5, 0, (0, 0))]
    s["shape_type"] = np.array([t for t, *_ in shapes], np.int32)
    s["shape_link"] = np.array([l for _, l, *_ in shapes], np.int32)
    s["shape_id"] = np.array([i for _, _, i, _ in shapes], np.int32)
    s["shape_params"] = np.array([[p[0], p[1], 0, 0] for *_, p in shapes], dtype=np.float64)
    s["shape_pose"] = np.stack([tilt((0, 0, 0)), tilt((0.3, 0.1, 0.0))] + [tilt((0, 0, 0)) for _ in range(5)]).reshape(-1, 12)
    s["shape_vert_offset"] = np.array([0, 0, 0, 0, 0, 0, 8], np.int32)
    s["shape_vert_count"] = np.array([0, 8, 0, 0, 0, 0, 8], np.int32)
    s["verts"] = np.concatenate([box, 1.5 * box])
    s["acm"] = np.zeros((n_links, n_links), np.uint8)
    return s


def arm_with_curved_obstacles_scene(seed=13):
    """the PR2 arm of the fixture (REVOLUTE joints, hull links) among static cylinders and cones placed along its seed
    path: curved cores met by shapes whose poses come from a sin/cos FK chain (primitives_scene has prismatic joints)"""
    rng = np.random.default_rng(seed)
    s = {k: np.array(v) for k, v in helpers.load_scene("arm_around_table").items()}
    n_links0, n_new = len(s["link_frame"]), 6
    centres = rng.uniform(np.array([0.35, -0.7, 0.55]), np.array([0.95, 0.1, 1.15]), (n_new, 3))
    s["link_frame"] = np.concatenate([s["link_frame"], -np.ones(n_new, np.int32)])
    s["link_names"] = np.concatenate([s["link_names"], np.array(["curved_%d" % i for i in range(n_new)])])
    poses = []
    for i in range(n_new):
        a = rng.normal(size=3)
        a /= np.linalg.norm(a)
        th = rng.uniform(0.0, np.pi)
        K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
        poses.append(_pose(np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K, centres[i]))
        s["shape_type"] = np.append(s["shape_type"], np.int32(capi.SHAPE_CYLINDER if i % 2 == 0 else capi.SHAPE_CONE))
        s["shape_link"] = np.append(s["shape_link"], np.int32(n_links0 + i))
        s["shape_id"] = np.append(s["shape_id"], np.int32(0))
        s["shape_params"] = np.concatenate([s["shape_params"], [[rng.uniform(0.04, 0.09), rng.uniform(0.06, 0.15), 0, 0]]])
        s["shape_pose"] = np.concatenate([s["shape_pose"], IDENT[None]])
        s["shape_vert_offset"] = np.append(s["shape_vert_offset"], np.int32(0))
        s["shape_vert_count"] = np.append(s["shape_vert_count"], np.int32(0))
    s["link_pose"] = np.concatenate([s["link_pose"], np.stack(poses)])
    n = n_links0 + n_new
    acm = np.ones((n, n), np.uint8)
    acm[:n_links0, :n_links0] = s["acm"]
    na = int(s["n_active_links"])
    acm[:na, n_links0:] = 0
    acm[n_links0:, :na] = 0
    s["acm"] = acm
    return s


def dynamic_env_scene(seed=17):
    """arm_around_table plus two links that move WITHOUT belonging to the planning group (a cart and a tool on it;
    trajopt/src/collision_terms.cpp:399-411 dynamic_environment): they directly follow the group's links
    (n_env_active_links = 2), are checked against the arm, against the static scene and against EACH OTHER."""
    s = {k: np.array(v) for k, v in helpers.load_scene("arm_around_table").items()}
    na, nl0 = int(s["n_active_links"]), len(s["link_frame"])
    ins = [("cart_link", np.array([[x, y, z] for x in (-0.15, 0.15) for y in (-0.2, 0.2) for z in (-0.1, 0.1)], dtype=np.float64),
            _pose(t=(0.55, -0.45, 0.95))),
           ("tool_link", np.array([[x, y, z] for x in (-0.05, 0.05) for y in (-0.05, 0.05) for z in (-0.12, 0.12)], dtype=np.float64),
            _pose(t=(0.55, -0.45, 1.16)))]
    k = len(ins)
    remap = np.concatenate([np.arange(na), np.arange(na, nl0) + k])  # old link index -> new
    s["link_names"] = np.concatenate([s["link_names"][:na], np.array([n for n, _, _ in ins]), s["link_names"][na:]])
    s["link_frame"] = np.concatenate([s["link_frame"][:na], -np.ones(k, np.int32), s["link_frame"][na:]]).astype(np.int32)
    lp = np.asarray(s["link_pose"], dtype=np.float64).reshape(nl0, 12)
    s["link_pose"] = np.concatenate([lp[:na], np.stack([p for _, _, p in ins]), lp[na:]])
    s["shape_link"] = remap[s["shape_link"]].astype(np.int32)
    voff = len(s["verts"])
    for i, (_, hv, _) in enumerate(ins):
        s["shape_type"] = np.append(s["shape_type"], np.int32(1))
        s["shape_link"] = np.append(s["shape_link"], np.int32(na + i))
        s["shape_id"] = np.append(s["shape_id"], np.int32(0))
        s["shape_params"] = np.concatenate([s["shape_params"], np.zeros((1, 4))])
        s["shape_pose"] = np.concatenate([s["shape_pose"], IDENT[None]])
        s["shape_vert_offset"] = np.append(s["shape_vert_offset"], np.int32(voff))
        s["shape_vert_count"] = np.append(s["shape_vert_count"], np.int32(len(hv)))
        s["verts"] = np.concatenate([s["verts"], hv])
        voff += len(hv)
    nl = nl0 + k
    acm = np.zeros((nl, nl), np.uint8)
    old = np.asarray(s["acm"])
    acm[np.ix_(remap, remap)] = old
    np.fill_diagonal(acm, 1)
    s["acm"] = acm
    s["n_env_active_links"] = np.int32(k)
    return s
