"""Shared test helpers: scene fixtures, synthetic scenes, the host-compiled device numerics."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import tcb200  # noqa: E402

pkg = tcb200.load()
capi = pkg.capi
SCENES = os.path.join(ROOT, "trajopt-1_b200", "data", "scenes")
_pd = C.POINTER(C.c_double)


def load_scene(name):
    return dict(np.load(os.path.join(SCENES, name + ".npz")))


def rand_pose(rng, scale=1.0):
    """random rigid pose as 12 doubles"""
    a = rng.normal(size=3)
    a /= np.linalg.norm(a)
    th = rng.uniform(-np.pi, np.pi)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    R = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)
    T = np.zeros((3, 4))
    T[:, :3] = R
    T[:, 3] = rng.uniform(-scale, scale, 3)
    return T.reshape(12)


def perturb_pose(rng, pose, dth, dt):
    T = pose.reshape(3, 4).copy()
    a = rng.normal(size=3)
    a /= np.linalg.norm(a)
    th = rng.uniform(-dth, dth)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    R = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)
    T[:, :3] = R @ T[:, :3]
    T[:, 3] += rng.uniform(-dt, dt, 3)
    return T.reshape(12)


class HostCheck(capi.Lib):
    """tests/hostcheck/libhostcheck.so: the kernels' per-thread numerics compiled for the host."""

    def __init__(self):
        here = os.path.join(ROOT, "tests", "hostcheck")
        subprocess.check_call(["make", "-C", here, "-s"])
        super().__init__(os.path.join(here, "libhostcheck.so"), "hc_")

    def shape_distance(self, scene, sa, pose_a0, sb, pose_b0, threshold, pose_a1=None, pose_b1=None,
                       precision=capi.PRECISION_F64, max_verts=259, max_faces=520):
        def ptr(a):
            if a is None:
                return None, _pd()
            a = np.ascontiguousarray(a, dtype=np.float64).reshape(12)
            return a, capi._p(a, _pd)

        k0, a0 = ptr(pose_a0)
        k1, a1 = ptr(pose_a1)
        k2, b0 = ptr(pose_b0)
        k3, b1 = ptr(pose_b1)
        status, epa, gi, ei, sc = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_int()
        dist = C.c_double()
        n, pa, pb = np.zeros(3), np.zeros(3), np.zeros(3)
        self.check(self.dll.hc_shape_distance(scene.handle, C.c_int(precision), C.c_int(sa), a0, a1, C.c_int(sb), b0,
                                              b1, C.c_double(threshold), C.c_int(max_verts), C.c_int(max_faces),
                                              C.byref(status), C.byref(dist), capi._p(n, _pd), capi._p(pa, _pd),
                                              capi._p(pb, _pd), C.byref(epa), C.byref(gi), C.byref(ei), C.byref(sc)))
        return dict(status=status.value, contact=status.value == 1, distance=dist.value, normal=n, pa=pa, pb=pb,
                    epa=bool(epa.value), iters=gi.value, epa_iters=ei.value, scanned=sc.value)

    def support_lut_check(self, scene, dirs):
        """support look-up table vs full vertex scan on every hull that has a table"""
        dirs = np.ascontiguousarray(dirs, dtype=np.float64).reshape(-1, 3)
        mism, cand, vert = C.c_longlong(), C.c_longlong(), C.c_longlong()
        self.dll.hc_support_lut_check.restype = C.c_int
        hulls = self.dll.hc_support_lut_check(scene.handle, capi._p(dirs, _pd), C.c_int(len(dirs)), C.byref(mism),
                                              C.byref(cand), C.byref(vert))
        return dict(hulls=hulls, mismatches=mism.value, candidates=cand.value, vertices=vert.value)


def contact_keys(c):
    """unique key of a contact inside one evaluation"""
    return list(zip(c["traj"].tolist(), c["term"].tolist(), c["pair"].tolist(), c["substep"].tolist(),
                    c["shape_id"][:, 0].tolist(), c["shape_id"][:, 1].tolist()))


def compare(ro, rg, band=1e-9, want_rows=True, scene=None):
    """Parity report oracle result `ro` vs CUDA result `rg` (same scene, same inputs).

    Contacts are matched by (traj, term, link pair, substep, shape ids).  Contacts whose oracle distance
    lies within `band` of margin+buffer may be present on one side only (SURVEY 8c parity protocol)."""
    co, cg = ro.contacts(), rg.contacts()
    ko, kg = contact_keys(co), contact_keys(cg)
    io = {k: i for i, k in enumerate(ko)}
    ig = {k: i for i, k in enumerate(kg)}
    assert len(io) == len(ko) and len(ig) == len(kg), "duplicate contact keys"
    _, _, margin, _ = ro.scene.pairs()
    thr = margin + ro.scene.margin_buffer
    only_o = [k for k in ko if k not in ig]
    only_g = [k for k in kg if k not in io]
    bad_o = [k for k in only_o if abs(co["distance"][io[k]] - thr[k[2]]) > band]
    bad_g = [k for k in only_g if abs(cg["distance"][ig[k]] - thr[k[2]]) > band]
    common = [k for k in ko if k in ig]
    a = np.array([io[k] for k in common], dtype=np.int64)
    b = np.array([ig[k] for k in common], dtype=np.int64)
    rep = dict(n_contacts=len(ko), n_contacts_gpu=len(kg), n_common=len(common), unmatched_outside_band=len(bad_o) + len(bad_g),
               same_order=(ko == kg))
    cto, ctg = ro.counts(), rg.counts()
    rep["counts_equal"] = all(cto[f] == ctg[f] for f in ("n_states", "n_units", "n_dense", "n_candidates", "n_terms"))
    rep["n_narrowphase"] = (cto["n_narrowphase"], ctg["n_narrowphase"])
    rep["n_epa"] = (cto["n_epa"], ctg["n_epa"])
    rep["gjk_iters"] = (cto["gjk_iters"], ctg["gjk_iters"])
    rep["n_unconverged"] = (cto["n_unconverged"], ctg["n_unconverged"])
    if len(common) == 0:
        rep.update(max_dist_err=0.0, max_normal_err=0.0, max_point_err=0.0, max_cc_time_err=0.0, cc_type_mismatch=0,
                   max_grad_rel_err=0.0, max_row_rel_err=0.0, max_row_value_err=0.0, max_cost_err=0.0)
        return rep
    A, G = co[a], cg[b]
    rep["max_dist_err"] = float(np.abs(A["distance"] - G["distance"]).max())
    rep["max_normal_err"] = float(np.abs(A["normal"] - G["normal"]).max())
    rep["max_point_err"] = float(max(np.abs(A["nearest_points"] - G["nearest_points"]).max(),
                                     np.abs(A["nearest_points_local"] - G["nearest_points_local"]).max()))
    rep["max_cc_time_err"] = float(np.abs(A["cc_time"] - G["cc_time"]).max())
    rep["cc_type_mismatch"] = int((A["cc_type"] != G["cc_type"]).any(axis=1).sum())
    rep["ids_equal"] = bool((A["link_ids"] == G["link_ids"]).all() and (A["state"] == G["state"]).all())
    xo, xg = ro.contact_transforms()[a], rg.contact_transforms()[b]
    rep["max_transform_err"] = float(np.abs(xo - xg).max())
    same_cc = ~(A["cc_type"] != G["cc_type"]).any(axis=1)
    if scene is not None:
        # contacts inside the stated band of the continuous-contact decisions may differ in cc_type / cc_time, and
        # with them in everything scaled by cc_time; outside the band nothing may (cc_outside_band must be 0)
        out, ins, keys_in = cc_band_check(ro, rg, scene)
        rep["cc_outside_band"], rep["cc_inside_band"] = out, ins
        same_cc &= ~np.isin(packed_keys(A), keys_in)
    if want_rows:
        reco, go = ro.gradients()
        recg, gg = rg.gradients()
        go, gg = go[a], gg[b]
        den = np.maximum(np.abs(go).reshape(len(a), -1).max(axis=1), 1e-12)
        rel = np.abs(go - gg).reshape(len(a), -1).max(axis=1) / den
        rep["max_grad_rel_err"] = float(rel[same_cc].max()) if same_cc.any() else 0.0
        rowo, rowg = ro.rows(), rg.rows()
        worst, worst_v = 0.0, 0.0
        for i, j in zip(a, b):
            vo = rowo["val"][rowo["row_ptr"][i]:rowo["row_ptr"][i + 1]]
            vg = rowg["val"][rowg["row_ptr"][j]:rowg["row_ptr"][j + 1]]
            assert len(vo) == len(vg)
            assert (rowo["col_idx"][rowo["row_ptr"][i]:rowo["row_ptr"][i + 1]] ==
                    rowg["col_idx"][rowg["row_ptr"][j]:rowg["row_ptr"][j + 1]]).all()
            if len(vo) and same_cc[np.where(a == i)[0][0]]:
                # rows scaled by (1 - cc_time) ~ 1e-9 are numerically zero: floor the denominator well above rounding
                worst = max(worst, float(np.abs(vo - vg).max() / max(np.abs(vo).max(), 1e-9)))
            worst_v = max(worst_v, abs(float(rowo["value"][i] - rowg["value"][j])))
        rep["max_row_rel_err"] = worst
        rep["max_row_value_err"] = worst_v
        rep["rec_equal"] = bool((reco["has_gradient"][a] == recg["has_gradient"][b]).all())
    cto_term, cto_traj = ro.cost()
    ctg_term, ctg_traj = rg.cost()
    rep["max_cost_err"] = float(np.abs(cto_traj - ctg_traj).max()) if len(cto_traj) else 0.0
    if ko == kg:
        rep["term_ptr_equal"] = bool((ro.term_ptr() == rg.term_ptr()).all())
    return rep


def save_scene_file(arrays, path):
    """flat named-array container read by trajopt-1_b200/host/tc_scene_file.hpp"""
    import struct
    items = []
    for k, v in arrays.items():
        v = np.asarray(v)
        if v.dtype.kind == "f":
            items.append((k, 0, np.ascontiguousarray(v, dtype="<f8")))
        elif v.dtype == np.uint8:
            items.append((k, 2, np.ascontiguousarray(v)))
        elif v.dtype.kind in "iu":
            items.append((k, 1, np.ascontiguousarray(v, dtype="<i4")))
    with open(path, "wb") as f:
        f.write(b"TCSCENE1" + struct.pack("<I", len(items)))
        for k, code, v in items:
            name = k.encode()
            f.write(struct.pack("<I", len(name)) + name + struct.pack("<IQ", code, v.size) + v.tobytes())


class GoldenResult:
    """a frozen oracle result (tests/golden/*.npz, tools/make_golden.py) with the accessors compare() uses"""

    class _Scene:
        def __init__(self, pairs, margin_buffer):
            self._pairs, self.margin_buffer = pairs, margin_buffer

        def pairs(self):
            return self._pairs

    def __init__(self, name, live_scene):
        self.d = dict(np.load(os.path.join(ROOT, "tests", "golden", name + ".npz")))
        self.scene = GoldenResult._Scene(live_scene.pairs(), float(self.d["params"][2]))

    def contacts(self):
        return self.d["contacts"]

    def contact_transforms(self):
        return self.d["transforms"]

    def term_ptr(self):
        return self.d["term_ptr"]

    def gradients(self):
        return self.d["grad_rec"], self.d["grad"]

    def rows(self):
        d = self.d
        return dict(row_ptr=d["row_ptr"], col_idx=d["col_idx"], val=d["val"], value=d["row_value"], coeff=d["row_coeff"],
                    margin=d["row_margin"])

    def cost(self):
        return self.d["cost_term"], self.d["cost_traj"]

    def counts(self):
        c = dict(zip(("n_states", "n_units", "n_dense", "n_candidates", "n_terms"), self.d["counts"].tolist()))
        c.update(n_narrowphase=-1, n_epa=-1, gjk_iters=-1, n_unconverged=-1)
        return c


def packed_keys(c):
    """int64 key per contact, ordered like the result (traj, term, link pair, substep, shape ids)"""
    k = c["traj"].astype(np.int64)
    k = k * 1024 + c["term"]
    k = k * 16384 + c["pair"]
    k = k * 16384 + c["substep"]
    return k * 64 + c["shape_id"][:, 0].astype(np.int64) * 8 + c["shape_id"][:, 1]


def compare_fast(ro, rg, band=1e-9):
    """vectorised parity report for large batches (same protocol as compare(); rows are compared through the
    dense per-contact gradient blocks and, when both sides hold the same contact sequence, the CSR arrays)"""
    co, cg = ro.contacts(), rg.contacts()
    ko, kg = packed_keys(co), packed_keys(cg)
    assert (np.diff(ko) > 0).all() and (np.diff(kg) > 0).all(), "contacts not strictly sorted by key"
    common, a, b = np.intersect1d(ko, kg, assume_unique=True, return_indices=True)
    _, _, margin, _ = ro.scene.pairs()
    thr = margin + ro.scene.margin_buffer
    mo, mg = np.ones(len(ko), bool), np.ones(len(kg), bool)
    mo[a], mg[b] = False, False
    bad = int((np.abs(co["distance"][mo] - thr[co["pair"][mo]]) > band).sum() +
              (np.abs(cg["distance"][mg] - thr[cg["pair"][mg]]) > band).sum())
    A, G = co[a], cg[b]
    rep = dict(n_contacts=len(ko), n_contacts_gpu=len(kg), n_common=len(common), unmatched_outside_band=bad,
               same_order=bool(len(ko) == len(kg) and (ko == kg).all()))
    rep["max_dist_err"] = float(np.abs(A["distance"] - G["distance"]).max())
    rep["max_normal_err"] = float(np.abs(A["normal"] - G["normal"]).max())
    rep["max_point_err"] = float(np.abs(A["nearest_points"] - G["nearest_points"]).max())
    rep["max_cc_time_err"] = float(np.abs(A["cc_time"] - G["cc_time"]).max())
    same_cc = (A["cc_type"] == G["cc_type"]).all(axis=1)
    rep["cc_type_mismatch"] = int((~same_cc).sum())
    reco, go = ro.gradients()
    recg, gg = rg.gradients()
    go, gg = go[a].reshape(len(a), -1), gg[b].reshape(len(b), -1)
    den = np.maximum(np.abs(go).max(axis=1), 1e-12)
    rel = np.abs(go - gg).max(axis=1) / den
    rep["max_grad_rel_err"] = float(rel[same_cc].max()) if same_cc.any() else 0.0
    rep["rec_equal"] = bool((reco["has_gradient"][a] == recg["has_gradient"][b]).all())
    if rep["same_order"]:
        rowo, rowg = ro.rows(), rg.rows()
        rep["rows_same_structure"] = bool((rowo["row_ptr"] == rowg["row_ptr"]).all() and
                                          (rowo["col_idx"] == rowg["col_idx"]).all())
        if rep["rows_same_structure"] and len(rowo["val"]):
            starts = rowo["row_ptr"][:-1]
            nz = np.diff(rowo["row_ptr"]) > 0
            dmax = np.maximum.reduceat(np.abs(rowo["val"] - rowg["val"]), starts[nz])
            vmax = np.maximum.reduceat(np.abs(rowo["val"]), starts[nz])
            ok = same_cc[nz]
            # rows scaled by (1 - cc_time) ~ 1e-16 are numerically zero: floor the denominator well above rounding
            rep["max_row_rel_err"] = float((dmax / np.maximum(vmax, 1e-9))[ok].max())
            rep["max_row_value_err"] = float(np.abs(rowo["value"] - rowg["value"]).max())
    _, to = ro.cost()
    _, tg = rg.cost()
    rep["max_cost_rel_err"] = float((np.abs(to - tg) / np.maximum(np.abs(to), 1.0)).max())
    return rep


def check_invariants(r, sg, q, evaluator, lvs, fixed, margin_buffer, stack=capi.STACK_IFOPT):
    """size-independent properties of one evaluation (any library with the capi.Result accessors): strictly sorted
    contact keys, term offsets partitioning the contacts, the dense-eval count from the LVS rule, the cost as a
    checksum of the contacts, one row per contact with value = margin - distance.  Returns the contacts."""
    B, T, _ = q.shape
    c = r.contacts()
    cnt = r.counts()
    keys = packed_keys(c)
    assert cnt["overflow"] == 0 and cnt["n_unconverged"] == 0 and len(c) == cnt["n_contacts"]
    assert (np.diff(keys) > 0).all()  # sorted, no duplicates
    free = ~np.asarray(fixed, bool)
    if evaluator == capi.EVAL_DISCRETE:
        n_terms, units = B * int(free.sum()), B * int(free.sum())
        terms_per_traj = int(free.sum())
    else:
        seg_free = free[:-1] | free[1:]
        terms_per_traj = int(seg_free.sum())
        n_terms = B * terms_per_traj
        seg = np.linalg.norm(np.diff(q, axis=1), axis=2)[:, seg_free]
        sub = np.where(seg > lvs, np.ceil(seg / lvs) + 1, 2).astype(np.int64)
        if evaluator == capi.EVAL_CONTINUOUS:
            sub[:] = 2
        units = int((sub - 1).sum()) if evaluator in (capi.EVAL_CONTINUOUS, capi.EVAL_LVS_CONTINUOUS) else int(sub.sum())
    assert cnt["n_terms"] == n_terms and cnt["n_units"] == units and cnt["n_dense"] == units * sg.num_pairs()
    tp = r.term_ptr()
    assert tp[0] == 0 and tp[-1] == len(c) and (np.diff(tp) >= 0).all()
    if evaluator != capi.EVAL_DISCRETE and free.all() or (evaluator != capi.EVAL_DISCRETE and seg_free.all()):
        term_of = c["traj"].astype(np.int64) * terms_per_traj + c["term"]
        assert (np.searchsorted(term_of, np.arange(n_terms + 1)) == tp).all()
    # cost = checksum of the contacts: sum coeff * max(margin - d, 0), margin WITHOUT the buffer
    _, _, margin, coeff = sg.pairs()
    per = coeff[c["pair"]] * np.maximum(margin[c["pair"]] - c["distance"], 0.0)
    want = np.bincount(c["traj"], weights=per, minlength=B)
    _, got = r.cost()
    assert np.abs(got - want).max() <= 1e-9 * max(1.0, want.max())
    assert (c["distance"] <= margin[c["pair"]] + margin_buffer).all()
    # rows: one per contact, value = margin - distance, two timestep blocks of n_dof columns at most
    rows = r.rows()
    assert len(rows["value"]) == len(c) and rows["row_ptr"][-1] == len(rows["val"])
    if stack == capi.STACK_IFOPT:
        assert np.abs(rows["value"] - (margin[c["pair"]] - c["distance"])).max() <= 1e-12
    else:
        # legacy affine expressions d(x) = constant + coeffs . x reproduce the distance at the current x
        x = q.reshape(B, -1)
        lin = np.zeros(len(c))
        nzr = np.diff(rows["row_ptr"]) > 0
        row_of = np.repeat(np.arange(len(c)), np.diff(rows["row_ptr"]))
        np.add.at(lin, row_of, rows["val"] * x[c["traj"][row_of], rows["col_idx"]])
        assert np.abs(rows["value"] + lin - c["distance"]).max() <= 1e-9
    assert (np.diff(rows["row_ptr"]) <= 2 * sg.n_dof).all()
    first_col = c["term"].astype(np.int64) * sg.n_dof
    nz = np.diff(rows["row_ptr"]) > 0
    if nz.any():
        assert (rows["col_idx"][rows["row_ptr"][:-1][nz]] >= first_col[nz]).all()
        assert (rows["col_idx"][rows["row_ptr"][1:][nz] - 1] < first_col[nz] + 2 * sg.n_dof).all()
    return c


# ---- the stated band of the continuous-contact decisions ----
# calculateContinuousData is a chain of DISCRETE decisions on continuous quantities: GetAverageSupport averages the hull
# vertices whose support value lies within BULLET_EPSILON = 1e-3 of the running maximum (a vertex is in or out), the
# classes TIME0 / TIME1 / BETWEEN follow from comparing the two support values against BULLET_SUPPORT_FUNC_TOLERANCE =
# 0.01, and cc_time switches to 0.5 below BULLET_LENGTH_TOLERANCE = 0.001.  Across any of those thresholds cc_type /
# cc_time jump, so two implementations whose contact normals agree only to the stated 1e-5 can legitimately land on
# different sides.  Stated band: a contact whose decisive quantity (a support value in metres, or the length sum) lies
# within CC_BAND of its threshold may differ in cc_type / cc_time (and in the rows scaled by them); outside the band
# they must agree.  CC_BAND = 1e-5 x 0.1 m: the normal tolerance times a typical vertex distance.
CC_BAND = 1e-6


def _avg_support_margin(verts, nl):
    """GetAverageSupport scan (tc_contact.cuh averageSupport): returns (support, point, smallest distance of any of its
    comparisons to the +-BULLET_EPSILON thresholds)"""
    eps, mx, cnt, acc, closest = 1e-3, -1000.0, 0.0, np.zeros(3), np.inf
    for p in verts:
        sup = float(p @ nl)
        closest = min(closest, abs(sup - (mx + eps)), abs(sup - (mx - eps)))
        if sup > mx + eps:
            cnt, acc, mx = 1.0, p.copy(), sup
        elif sup < mx - eps:
            pass
        else:
            cnt += 1.0
            acc = acc + p
    return mx, acc / cnt, closest


def cc_decision_margin(c, xf, scene):
    """smallest distance [m] of the continuous-contact decisions of contact record `c` (transforms xf[2][2][12]) to
    their thresholds, over its swept hull links; inf when no decision is involved (cc_type NONE / primitives)"""
    shape_pose = np.asarray(scene["shape_pose"], dtype=np.float64).reshape(-1, 3, 4)
    allv = np.asarray(scene["verts"], dtype=np.float64).reshape(-1, 3)
    margin = np.inf
    for i in range(2):
        if c["cc_type"][i] == capi.CC_NONE:
            continue
        sel = np.nonzero((scene["shape_link"] == c["link_ids"][i]) & (scene["shape_id"] == c["shape_id"][i]))[0]
        if len(sel) != 1 or scene["shape_type"][sel[0]] != capi.SHAPE_HULL:
            continue
        k = sel[0]
        v = allv[scene["shape_vert_offset"][k]:scene["shape_vert_offset"][k] + scene["shape_vert_count"][k]]
        v = v @ shape_pose[k][:, :3].T + shape_pose[k][:, 3]
        n = np.array(c["normal"]) * (1.0 if i == 0 else -1.0)
        T0, T1 = xf[i, 0].reshape(3, 4), xf[i, 1].reshape(3, 4)
        s0, p0, m0 = _avg_support_margin(v, T0[:, :3].T @ n)
        s1, p1, m1 = _avg_support_margin(v, T1[:, :3].T @ n)
        pw0, pw1 = T0[:, :3] @ p0 + T0[:, 3], T1[:, :3] @ p1 + T1[:, 3]
        sup0, sup1 = float(n @ pw0), float(n @ pw1)
        pt = np.array(c["nearest_points"][i])
        lsum = np.linalg.norm(pt - pw0) + np.linalg.norm(pt - pw1)
        margin = min(margin, m0, m1, abs(abs(sup0 - sup1) - 0.01), abs(lsum - 0.001))
    return margin


def cc_band_check(ro, rg, scene, band=CC_BAND, cc_time_tol=1e-6):
    """Common contacts whose cc_type or cc_time differ between the two results: returns (number outside the stated
    band — must be 0 —, number inside it, keys of the in-band ones)"""
    co, cg = ro.contacts(), rg.contacts()
    _, a, b = np.intersect1d(packed_keys(co), packed_keys(cg), assume_unique=True, return_indices=True)
    A, G = co[a], cg[b]
    differ = (A["cc_type"] != G["cc_type"]).any(axis=1) | (np.abs(A["cc_time"] - G["cc_time"]).max(axis=1) > cc_time_tol)
    rows = np.nonzero(differ)[0]
    if len(rows) == 0:
        return 0, 0, np.zeros(0, np.int64)
    xf = ro.contact_transforms()[a]
    inside = [r for r in rows if cc_decision_margin(A[r], xf[r], scene) <= band]
    return len(rows) - len(inside), len(inside), packed_keys(A[np.array(inside, dtype=np.int64)]) if inside else np.zeros(0, np.int64)


def cc_type_outside_band(ro, rg, scene, band=CC_BAND):
    """(mismatches of cc_type / cc_time outside the stated band — must be 0 —, mismatches inside it)"""
    out, ins, _ = cc_band_check(ro, rg, scene, band)
    return out, ins
