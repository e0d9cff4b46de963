"""CPU suite: the kernels' per-thread numerics (csrc/*.cuh compiled for the host, -ffp-contract=off) against
the oracle on random shape pairs — same contact decision, same iteration counts, distances to 1e-12."""
import numpy as np
import pytest

from tests import helpers

capi = helpers.capi


@pytest.mark.parametrize("scene_name,n,seed", [("arm_around_table", 500, 1), ("boxbot", 150, 2), ("spherebot", 150, 3)])
def test_signed_distance_matches_oracle(oracle, hostcheck, scene_name, n, seed):
    s = helpers.load_scene(scene_name)
    so, sh = oracle.scene(s, 0.025, 20, 0.01), hostcheck.scene(s, 0.025, 20, 0.01)
    rng = np.random.default_rng(seed)
    nshape = len(s["shape_type"])
    n_contact = n_epa = 0
    for it in range(n):
        sa, sb = rng.integers(0, nshape, 2)
        scale = 0.3 if scene_name == "arm_around_table" else 1.0
        pa0, pb0 = helpers.rand_pose(rng, scale), helpers.rand_pose(rng, scale)
        pa1 = helpers.perturb_pose(rng, pa0, 0.2, 0.05) if it % 2 == 0 else None
        ro = oracle.shape_distance(so, sa, pa0, sb, pb0, 0.05, pose_a1=pa1)
        rh = hostcheck.shape_distance(sh, sa, pa0, sb, pb0, 0.05, pose_a1=pa1)
        assert ro["contact"] == rh["contact"], (it, ro["distance"], rh["distance"])
        if not ro["contact"]:
            continue
        n_contact += 1
        n_epa += int(ro["epa"])
        assert ro["epa"] == rh["epa"] and ro["iters"] == rh["iters"], (it, ro, rh)
        assert abs(ro["distance"] - rh["distance"]) <= 1e-12, (it, ro["distance"], rh["distance"])
        assert np.abs(ro["normal"] - rh["normal"]).max() <= 1e-9, it
        assert max(np.abs(ro["pa"] - rh["pa"]).max(), np.abs(ro["pb"] - rh["pb"]).max()) <= 1e-9, it
    assert n_contact > n // 20
    if scene_name != "spherebot":  # sphere cores never penetrate: pure GJK
        assert n_epa > 0


def test_fast_and_deep_polytope_agree(oracle, hostcheck):
    """an EPA that overflows the thread-local polytope is re-run with the large store: same answer as the oracle"""
    s = helpers.load_scene("arm_around_table")
    so, sh = oracle.scene(s, 0.025, 20, 0.01), hostcheck.scene(s, 0.025, 20, 0.01)
    rng = np.random.default_rng(7)
    nshape = len(s["shape_type"])
    seen_overflow = 0
    for it in range(400):
        sa, sb = rng.integers(0, nshape, 2)
        pa0, pb0 = helpers.rand_pose(rng, 0.05), helpers.rand_pose(rng, 0.05)  # deep overlaps
        ro = oracle.shape_distance(so, sa, pa0, sb, pb0, 0.05)
        small = hostcheck.shape_distance(sh, sa, pa0, sb, pb0, 0.05, max_verts=12, max_faces=20)
        if small["status"] == 2:  # ND_EPA_OVERFLOW: the device hands these to the deep kernel
            seen_overflow += 1
            small = hostcheck.shape_distance(sh, sa, pa0, sb, pb0, 0.05)
        assert small["contact"] == ro["contact"]
        if ro["contact"]:
            assert abs(ro["distance"] - small["distance"]) <= 1e-12
    assert seen_overflow > 0


def _lut_directions(rng, n):
    """random directions plus the awkward ones: axes, cube-map face / cell boundaries, tiny and huge magnitudes"""
    d = [rng.normal(size=(n, 3))]
    ax = np.concatenate([np.eye(3), -np.eye(3)])
    d.append(ax)
    g = np.linspace(-1, 1, 9)  # cell boundaries of the 8 x 8 cube-map faces
    for m in range(3):
        for sg in (1.0, -1.0):
            for u in g:
                for v in (g[0], g[3], g[4], g[8], 0.123):
                    e = np.zeros(3)
                    e[m], e[(m + 1) % 3], e[(m + 2) % 3] = sg, u, v
                    d.append(e[None])
                    d.append((e * (1 + 1e-7 * rng.normal(size=3)))[None])
    out = np.concatenate(d)
    out = np.concatenate([out, 1e-9 * out[:200], 1e12 * out[:200], 1e-200 * out[:50]])
    return out


@pytest.mark.parametrize("scene_name", ["arm_around_table", "cluttered", "dual_arm"])
def test_support_lut_equals_full_scan(hostcheck, scene_name):
    """the cube-map support table returns bit-for-bit the vertex the full scan returns, and prunes most vertices"""
    from tests import synthetic
    if scene_name == "arm_around_table":
        s = helpers.load_scene(scene_name)
    else:
        s = synthetic.cluttered_scene() if scene_name == "cluttered" else synthetic.dual_arm_scene()
    sh = hostcheck.scene(s, 0.025, 20, 0.01)
    r = hostcheck.support_lut_check(sh, _lut_directions(np.random.default_rng(11), 4000))
    assert r["hulls"] > 0, r
    assert r["mismatches"] == 0, r
    assert r["candidates"] < 0.25 * r["vertices"], r


def test_primitive_shapes_match_oracle(oracle, hostcheck):
    """capsule, cylinder, cone, sphere and small hulls against each other (tests/synthetic.primitives_scene): the
    out-of-line primitive support of the kernels against the oracle, discrete and swept"""
    from tests import synthetic
    s = synthetic.primitives_scene()
    so, sh = oracle.scene(s, 0.1, 1.0, 0.05), hostcheck.scene(s, 0.1, 1.0, 0.05)
    rng = np.random.default_rng(5)
    n = len(s["shape_type"])
    seen, n_contact, n_epa = set(), 0, 0
    for it in range(600):
        sa, sb = rng.integers(0, n, 2)
        pa0, pb0 = helpers.rand_pose(rng, 0.6), helpers.rand_pose(rng, 0.6)
        pa1 = helpers.perturb_pose(rng, pa0, 0.3, 0.1) if it % 2 == 0 else None
        ro = oracle.shape_distance(so, sa, pa0, sb, pb0, 0.2, pose_a1=pa1)
        rh = hostcheck.shape_distance(sh, sa, pa0, sb, pb0, 0.2, pose_a1=pa1)
        assert ro["contact"] == rh["contact"], (it, sa, sb)
        if not ro["contact"]:
            continue
        n_contact += 1
        n_epa += int(ro["epa"])
        seen.add((int(s["shape_type"][sa]), int(s["shape_type"][sb])))
        assert ro["epa"] == rh["epa"] and ro["iters"] == rh["iters"], (it, sa, sb, ro, rh)
        assert abs(ro["distance"] - rh["distance"]) <= 1e-12
        assert np.abs(ro["normal"] - rh["normal"]).max() <= 1e-9
    assert n_contact > 100 and n_epa > 10 and len(seen) >= 12, (n_contact, n_epa, sorted(seen))


def test_large_hull_without_table_matches_oracle(oracle, hostcheck):
    """a 400-vertex hull is beyond the look-up table's 8-bit vertex indices: the kernels fall back to the full scan
    (hullScanFull, both end poses of a sweep in one pass) — same answers as the oracle"""
    from tests import synthetic
    s = synthetic.primitives_scene()
    rng = np.random.default_rng(9)
    p = rng.normal(size=(400, 3))
    p = 0.5 * p / np.linalg.norm(p, axis=1, keepdims=True) * np.array([1.0, 0.7, 0.5])
    s["verts"] = np.concatenate([s["verts"], p])
    s["shape_vert_offset"][1], s["shape_vert_count"][1] = 16, 400  # the robot's hull shape
    so, sh = oracle.scene(s, 0.1, 1.0, 0.05), hostcheck.scene(s, 0.1, 1.0, 0.05)
    r = hostcheck.support_lut_check(sh, rng.normal(size=(10, 3)))
    assert r["hulls"] == 0  # the 8-vertex crate is below the table's minimum size, the 400-vertex hull above its maximum
    n_contact = 0
    for it in range(150):
        sb = int(rng.integers(2, 7))
        pa0, pb0 = helpers.rand_pose(rng, 0.6), helpers.rand_pose(rng, 0.6)
        pa1 = helpers.perturb_pose(rng, pa0, 0.3, 0.1) if it % 2 == 0 else None
        ro = oracle.shape_distance(so, 1, pa0, sb, pb0, 0.2, pose_a1=pa1)
        rh = hostcheck.shape_distance(sh, 1, pa0, sb, pb0, 0.2, pose_a1=pa1)
        assert ro["contact"] == rh["contact"], it
        if ro["contact"]:
            n_contact += 1
            assert ro["iters"] == rh["iters"] and abs(ro["distance"] - rh["distance"]) <= 1e-12, (it, ro, rh)
    assert n_contact > 30


def test_shared_sincos_and_fk_are_bit_identical_and_accurate(oracle, hostcheck):
    """The device FK (host-compiled), and the oracle's FK evaluate ONE sin/cos operation sequence: same bits out.
    That is what lets the contraction-free GPU kernels (tc_strict.cu) start GJK on curved cores from the oracle's own
    inputs.  Accuracy of the shared sin/cos against libm: <= 2.5e-16 absolute (it replaces libm's in both)."""
    import ctypes as C
    rng = np.random.default_rng(11)
    xs = np.concatenate([rng.uniform(-7, 7, 20000), rng.uniform(-1e5, 1e5, 5000), [0.0, -0.0, np.pi / 2, np.pi, 1e5, 1e-300],
                         np.arange(-8, 9) * (np.pi / 4)])
    so, co, sh, ch = C.c_double(), C.c_double(), C.c_double(), C.c_double()
    oracle.dll.orc_sincos.restype = None
    hostcheck.dll.hc_sincos.restype = None
    worst = 0.0
    for x in xs:
        oracle.dll.orc_sincos(C.c_double(x), C.byref(so), C.byref(co))
        hostcheck.dll.hc_sincos(C.c_double(x), C.byref(sh), C.byref(ch))
        assert so.value == sh.value and co.value == ch.value, x
        worst = max(worst, abs(so.value - np.sin(x)), abs(co.value - np.cos(x)))
    assert worst <= 2.5e-16, worst
    s = helpers.load_scene("arm_around_table")
    sco, sch = oracle.scene(s, 0.025, 20.0, 0.01), hostcheck.scene(s, 0.025, 20.0, 0.01)
    lim = s["joint_limits"]
    pd = C.POINTER(C.c_double)
    for _ in range(200):
        q = np.ascontiguousarray(rng.uniform(lim[:, 0], lim[:, 1]))
        a = oracle.fk(sco, q)
        b = np.zeros_like(a)
        assert hostcheck.dll.hc_fk(sch.handle, q.ctypes.data_as(pd), b.ctypes.data_as(pd)) == 0
        assert np.array_equal(a, b)
