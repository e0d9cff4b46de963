#!/usr/bin/env python3
"""Generate scene fixtures under trajopt-1_b200/data/scenes/ from the reference's data files.

Run in the build container only (needs /root/reference).  The GPU box never sees
/root/reference, so everything tests/bench need from it is frozen here as small .npz
files: flattened kinematic tree of the planning group, collision shapes (meshes replaced
by their convex hulls because the fixtures carry tesseract:make_convex="true"), static
link poses at the test's environment state, the allowed-collision matrix, and the seed
trajectories / term parameters from the JSON problem files.

Sources (reference file -> what we take):
  trajopt_common/data/boxbot.urdf,.srdf                 -> boxbot scene
  trajopt_common/data/spherebot.urdf,.srdf              -> spherebot scene
  trajopt_common/data/arm_around_table.urdf, pr2.srdf   -> PR2 arm_around_table scene
  trajopt_common/data/config/*.json                     -> seed trajectories, margins, coeffs
  trajopt_common/data/meshes/Table.stl, pr2/meshes/**.stl -> hull vertices

Scene layout written (all float64 / int32, see include/trajopt_cuda.h for meaning):
  n_dof, frame_parent[F], frame_joint_type[F], frame_dof[F], frame_axis[F,3],
  frame_origin[F,12], base_pose[12], joint_limits[n_dof,2],
  link_names[L], link_frame[L] (-1 static), link_pose[L,12], n_active_links,
  shape_type[S], shape_link[S], shape_id[S], shape_params[S,4], shape_pose[S,12],
  shape_vert_offset[S], shape_vert_count[S], verts[V,3],
  acm[L,L] uint8 (1 = collision allowed/disabled for that pair)
"""
import json
import os
import struct
import sys
import xml.etree.ElementTree as ET

import numpy as np
from scipy.spatial import ConvexHull

REF = "/root/reference/trajopt_common/data"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "trajopt-1_b200", "data", "scenes")

JOINT_FIXED, JOINT_REVOLUTE, JOINT_PRISMATIC = 0, 1, 2
SHAPE_SPHERE, SHAPE_HULL, SHAPE_CAPSULE, SHAPE_CYLINDER, SHAPE_CONE = 0, 1, 2, 3, 4


def rpy_to_R(r, p, y):
    cr, sr, cp, sp, cy, sy = np.cos(r), np.sin(r), np.cos(p), np.sin(p), np.cos(y), np.sin(y)
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def parse_origin(el):
    T = np.eye(4)
    if el is None:
        return T
    xyz = [float(v) for v in el.get("xyz", "0 0 0").split()]
    rpy = [float(v) for v in el.get("rpy", "0 0 0").split()]
    T[:3, :3] = rpy_to_R(*rpy)
    T[:3, 3] = xyz
    return T


def axis_angle_T(axis, q):
    a = np.asarray(axis, float)
    a = a / np.linalg.norm(a)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    R = np.eye(3) + np.sin(q) * K + (1 - np.cos(q)) * (K @ K)
    T = np.eye(4)
    T[:3, :3] = R
    return T


def read_stl(path):
    with open(path, "rb") as f:
        data = f.read()
    if data[:5] == b"solid" and b"facet" in data[:1000]:
        pts = []
        for line in data.decode("ascii", "ignore").splitlines():
            line = line.strip()
            if line.startswith("vertex"):
                pts.append([float(v) for v in line.split()[1:4]])
        return np.array(pts, float)
    (ntri,) = struct.unpack_from("<I", data, 80)
    rec = np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("attr", "<u2")])
    tris = np.frombuffer(data, dtype=rec, count=ntri, offset=84)
    return tris["v"].reshape(-1, 3).astype(float)


def hull_vertices(pts):
    pts = np.unique(np.round(pts, 12), axis=0)
    h = ConvexHull(pts)
    return pts[np.sort(h.vertices)]


def resolve(fn):
    assert fn.startswith("package://trajopt_common/data/"), fn
    return os.path.join(REF, fn[len("package://trajopt_common/data/"):])


def T12(T):
    return np.asarray(T, float)[:3, :4].reshape(12).copy()


class Urdf:
    def __init__(self, path):
        text = open(path).read().replace("tesseract:", "tesseract_")  # attribute prefix has no xmlns binding
        root = ET.fromstring(text)
        self.links = {}
        self.link_order = []
        for l in root.findall("link"):
            cols = []
            for c in l.findall("collision"):
                T = parse_origin(c.find("origin"))
                g = c.find("geometry")
                child = list(g)[0]
                cols.append((T, child))
            self.links[l.get("name")] = cols
            self.link_order.append(l.get("name"))
        self.joints = {}
        self.child_joint = {}
        self.children = {}
        for j in root.findall("joint"):
            if j.find("parent") is None:
                continue  # transmission-style <joint name=...> tags
            name = j.get("name")
            jt = j.get("type")
            ax = j.find("axis")
            axis = [float(v) for v in ax.get("xyz").split()] if ax is not None else [1.0, 0, 0]
            lim = j.find("limit")
            lo, hi = -1e9, 1e9
            if lim is not None and jt in ("revolute", "prismatic"):
                lo, hi = float(lim.get("lower", -1e9)), float(lim.get("upper", 1e9))
            rec = dict(name=name, type=jt, parent=j.find("parent").get("link"), child=j.find("child").get("link"),
                       origin=parse_origin(j.find("origin")), axis=np.array(axis, float), limits=(lo, hi))
            self.joints[name] = rec
            self.child_joint[rec["child"]] = rec
            self.children.setdefault(rec["parent"], []).append(rec)
        roots = [l for l in self.link_order if l not in self.child_joint]
        assert len(roots) == 1, roots
        self.root = roots[0]

    def chain_joints(self, base, tip):
        out = []
        l = tip
        while l != base:
            j = self.child_joint[l]
            out.append(j)
            l = j["parent"]
        out.reverse()
        return [j for j in out if j["type"] != "fixed"]

    def fk_all(self, qmap):
        """World pose of every link with joint values from qmap (default 0)."""
        poses = {self.root: np.eye(4)}
        stack = [self.root]
        while stack:
            l = stack.pop()
            for j in self.children.get(l, []):
                T = poses[l] @ j["origin"]
                q = qmap.get(j["name"], 0.0)
                if j["type"] in ("revolute", "continuous"):
                    T = T @ axis_angle_T(j["axis"], q)
                elif j["type"] == "prismatic":
                    Tq = np.eye(4)
                    Tq[:3, 3] = j["axis"] / np.linalg.norm(j["axis"]) * q
                    T = T @ Tq
                poses[j["child"]] = T
                stack.append(j["child"])
        return poses


def parse_srdf(path):
    root = ET.parse(path).getroot()
    groups = {}
    for g in root.findall("group"):
        ch = g.find("chain")
        if ch is not None:
            groups[g.get("name")] = (ch.get("base_link"), ch.get("tip_link"))
    acm = set()
    for d in root.findall("disable_collisions"):
        acm.add((d.get("link1"), d.get("link2")))
    return groups, acm


def build_scene(urdf_path, srdf_path, group, env_state=None):
    env_state = env_state or {}
    u = Urdf(urdf_path)
    groups, acm_pairs = parse_srdf(srdf_path)
    base, tip = groups[group]
    dof_joints = u.chain_joints(base, tip)
    dof_of = {j["name"]: i for i, j in enumerate(dof_joints)}
    n_dof = len(dof_joints)
    world = u.fk_all(env_state)

    # Frames: every link strictly below the first dof joint's parent, in DFS order.
    # parent = -1 refers to the chain base link whose world pose is base_pose.
    frame_of = {}
    f_parent, f_type, f_dof, f_axis, f_origin = [], [], [], [], []

    def add_subtree(link, parent_frame):
        for j in u.children.get(link, []):
            jt = j["type"]
            origin = j["origin"].copy()
            if j["name"] in dof_of:
                t = JOINT_REVOLUTE if jt in ("revolute", "continuous") else JOINT_PRISMATIC
                d = dof_of[j["name"]]
            else:
                # not a group joint: frozen at the environment state
                q = env_state.get(j["name"], 0.0)
                if jt in ("revolute", "continuous"):
                    origin = origin @ axis_angle_T(j["axis"], q)
                elif jt == "prismatic":
                    Tq = np.eye(4)
                    Tq[:3, 3] = j["axis"] / np.linalg.norm(j["axis"]) * q
                    origin = origin @ Tq
                t, d = JOINT_FIXED, -1
            idx = len(f_parent)
            frame_of[j["child"]] = idx
            f_parent.append(parent_frame)
            f_type.append(t)
            f_dof.append(d)
            f_axis.append(j["axis"] / np.linalg.norm(j["axis"]))
            f_origin.append(T12(origin))
            add_subtree(j["child"], idx)

    # Only descend from links on the chain so siblings of the base (left arm etc) stay static.
    first = dof_joints[0]
    # walk from base down to the parent link of the first dof joint through fixed joints
    assert first["parent"] == base or True
    # generic: build frames for the whole subtree of `base`, then mark which move
    add_subtree(base, -1)
    F = len(f_parent)
    moves = [False] * F
    for f in range(F):
        p = f_parent[f]
        moves[f] = (f_dof[f] >= 0) or (p >= 0 and moves[p])

    # prune frames that never move and are not ancestors of moving frames
    needed = [False] * F
    for f in range(F):
        if moves[f]:
            g = f
            while g >= 0 and not needed[g]:
                needed[g] = True
                g = f_parent[g]
    remap = {}
    for f in range(F):
        if needed[f]:
            remap[f] = len(remap)
    frame_of = {l: remap[f] for l, f in frame_of.items() if needed[f] and moves[f]}
    keep = [f for f in range(F) if needed[f]]
    f_parent = [remap[f_parent[f]] if f_parent[f] >= 0 else -1 for f in keep]
    f_type = [f_type[f] for f in keep]
    f_dof = [f_dof[f] for f in keep]
    f_axis = [f_axis[f] for f in keep]
    f_origin = [f_origin[f] for f in keep]

    # collision links: active first (in frame order) then static (urdf order)
    col_links = [l for l in u.link_order if u.links[l]]
    active = sorted([l for l in col_links if l in frame_of], key=lambda l: frame_of[l])
    static = [l for l in col_links if l not in frame_of]
    link_names = active + static
    L = len(link_names)
    link_frame = np.array([frame_of.get(l, -1) for l in link_names], np.int32)
    link_pose = np.stack([T12(world[l]) for l in link_names])

    s_type, s_link, s_id, s_params, s_pose, s_off, s_cnt = [], [], [], [], [], [], []
    verts = []
    nv = 0
    for li, l in enumerate(link_names):
        for k, (T, g) in enumerate(u.links[l]):
            params = np.zeros(4)
            off, cnt = 0, 0
            if g.tag == "box":
                sz = np.array([float(v) for v in g.get("size").split()])
                h = sz / 2
                pts = np.array([[sx * h[0], sy * h[1], sz_ * h[2]] for sx in (-1, 1) for sy in (-1, 1) for sz_ in (-1, 1)])
                t = SHAPE_HULL
                params[:3] = h
            elif g.tag == "sphere":
                t = SHAPE_SPHERE
                params[0] = float(g.get("radius"))
                pts = None
            elif g.tag == "cylinder":
                t = SHAPE_CYLINDER
                params[0] = float(g.get("radius"))
                params[1] = float(g.get("length")) / 2
                pts = None
            elif g.tag == "mesh":
                t = SHAPE_HULL
                scale = np.array([float(v) for v in g.get("scale", "1 1 1").split()])
                pts = hull_vertices(read_stl(resolve(g.get("filename"))) * scale)
            else:
                raise ValueError(g.tag)
            if pts is not None:
                off, cnt = nv, len(pts)
                verts.append(pts)
                nv += cnt
            s_type.append(t)
            s_link.append(li)
            s_id.append(k)
            s_params.append(params)
            s_pose.append(T12(T))
            s_off.append(off)
            s_cnt.append(cnt)

    acm = np.zeros((L, L), np.uint8)
    idx = {l: i for i, l in enumerate(link_names)}
    for a, b in acm_pairs:
        if a in idx and b in idx:
            acm[idx[a], idx[b]] = acm[idx[b], idx[a]] = 1

    return dict(
        n_dof=np.int32(n_dof),
        joint_names=np.array([j["name"] for j in dof_joints]),
        joint_limits=np.array([j["limits"] for j in dof_joints], float),
        frame_parent=np.array(f_parent, np.int32), frame_joint_type=np.array(f_type, np.int32),
        frame_dof=np.array(f_dof, np.int32), frame_axis=np.array(f_axis, float).reshape(-1, 3),
        frame_origin=np.array(f_origin, float).reshape(-1, 12), base_pose=T12(world[base]),
        link_names=np.array(link_names), link_frame=link_frame, link_pose=link_pose,
        n_active_links=np.int32(len(active)),
        shape_type=np.array(s_type, np.int32), shape_link=np.array(s_link, np.int32),
        shape_id=np.array(s_id, np.int32), shape_params=np.array(s_params, float).reshape(-1, 4),
        shape_pose=np.array(s_pose, float).reshape(-1, 12), shape_vert_offset=np.array(s_off, np.int32),
        shape_vert_count=np.array(s_cnt, np.int32),
        verts=(np.concatenate(verts) if verts else np.zeros((0, 3))), acm=acm)


def main():
    os.makedirs(OUT, exist_ok=True)
    box = build_scene(f"{REF}/boxbot.urdf", f"{REF}/boxbot.srdf", "manipulator")
    np.savez_compressed(f"{OUT}/boxbot.npz", **box)
    sph = build_scene(f"{REF}/spherebot.urdf", f"{REF}/spherebot.srdf", "manipulator")
    np.savez_compressed(f"{OUT}/spherebot.npz", **sph)
    # trajopt/test/planning_unit.cpp:70-78: torso_lift_joint = 0, everything else default 0
    pr2 = build_scene(f"{REF}/arm_around_table.urdf", f"{REF}/pr2.srdf", "right_arm", {"torso_lift_joint": 0.0})
    cfg = json.load(open(f"{REF}/config/arm_around_table.json"))
    pr2["seed_traj"] = np.array(cfg["init_info"]["data"], float)
    np.savez_compressed(f"{OUT}/arm_around_table.npz", **pr2)
    for name, s in (("boxbot", box), ("spherebot", sph), ("arm_around_table", pr2)):
        na = int(s["n_active_links"])
        L = len(s["link_names"])
        P = sum(1 for a in range(L) for b in range(a + 1, L) if (a < na or b < na) and not s["acm"][a, b])
        print(f"{name}: n_dof={int(s['n_dof'])} frames={len(s['frame_parent'])} links={L} active={na} "
              f"shapes={len(s['shape_type'])} verts={len(s['verts'])} candidate_link_pairs={P}")


if __name__ == "__main__":
    sys.exit(main())
