"""GPU parity tests proper: CUDA path (through the C ABI) vs the CPU oracle on the same seeded inputs.
Tolerances are the north_star's: identical contact sets outside a 1e-9 m band around margin+buffer,
distances within 1e-6 m, gradient rows within 1e-5 relative (FP64 mode)."""
import numpy as np
import pytest

from tests import helpers

capi = helpers.capi
pytestmark = pytest.mark.gpu

DIST_TOL, ROW_TOL = 1e-6, 1e-5


def check(rep, rows=True):
    assert rep["counts_equal"], rep
    assert rep["unmatched_outside_band"] == 0, rep
    assert rep["max_dist_err"] <= DIST_TOL, rep
    assert rep["max_normal_err"] <= 1e-5, rep
    assert rep["max_cost_err"] <= 1e-6 * max(1.0, rep["n_contacts"]), rep
    if rows:
        assert rep["max_grad_rel_err"] <= ROW_TOL, rep
        assert rep["max_row_rel_err"] <= ROW_TOL, rep
        assert rep["max_row_value_err"] <= 1e-6, rep


def pair(cuda, oracle, name, margin, coeff, buf):
    s = helpers.load_scene(name)
    return s, cuda.scene(s, margin, coeff, buf), oracle.scene(s, margin, coeff, buf)


def test_boxbot_known_answers(cuda):
    """trajopt_ifopt/test/collision_unit.cpp:97-159: DiscreteCollisionConstraint(max_num_cnt=1) on boxbot"""
    s = helpers.load_scene("boxbot")
    sg = cuda.scene(s, 0.1, 1.0, 0.01)
    cases = [((-1.9, 0.01), -0.01, (0, 0)), ((-1.0, 0.01), 0.1, (1, 0)), ((0.01, 1.0), 0.1, (0, -1)),
             ((-0.9, 0.01), 0.2, (1, 0))]
    for q, val, jac in cases:
        r = cuda.eval(sg, np.array(q).reshape(1, 1, 2), capi.EVAL_DISCRETE)
        st = r.sets(1)
        assert abs(st["values"][0, 0] - val) <= 1e-6, (q, st["values"])
        assert np.abs(st["jac"][0, 0, 0] - np.array(jac)).max() <= 1e-6, (q, st["jac"])


@pytest.mark.parametrize("evaluator", [capi.EVAL_DISCRETE, capi.EVAL_LVS_DISCRETE, capi.EVAL_CONTINUOUS,
                                       capi.EVAL_LVS_CONTINUOUS])
@pytest.mark.parametrize("stack", [capi.STACK_IFOPT, capi.STACK_LEGACY])
def test_arm_around_table_seed(cuda, oracle, evaluator, stack):
    s, sg, so = pair(cuda, oracle, "arm_around_table", 0.025, 20.0, 0.01)
    q = s["seed_traj"][None]
    fixed = np.zeros(q.shape[1], np.uint8)
    fixed[0] = fixed[-1] = 1
    rg = cuda.eval(sg, q, evaluator, stack=stack, lvs=0.02, fixed=fixed)
    ro = oracle.eval(so, q, evaluator, stack=stack, lvs=0.02, fixed=fixed)
    rep = helpers.compare(ro, rg)
    assert rep["n_contacts"] > 0
    check(rep)
    assert rep["same_order"] and rep["term_ptr_equal"]


@pytest.mark.parametrize("evaluator", [capi.EVAL_LVS_DISCRETE, capi.EVAL_LVS_CONTINUOUS])
def test_arm_batch_random_seeds(cuda, oracle, evaluator):
    s, sg, so = pair(cuda, oracle, "arm_around_table", 0.025, 20.0, 0.01)
    rng = np.random.default_rng(7)
    B = 24
    q = np.repeat(s["seed_traj"][None], B, 0)
    q[:, 1:-1] += rng.normal(0, 0.3, (B, q.shape[1] - 2, 7))
    lim = s["joint_limits"]
    q = np.clip(q, lim[:, 0], lim[:, 1])
    fixed = np.zeros(q.shape[1], np.uint8)
    fixed[0] = fixed[-1] = 1
    rg = cuda.eval(sg, q, evaluator, lvs=0.05, fixed=fixed)
    ro = oracle.eval(so, q, evaluator, lvs=0.05, fixed=fixed, n_threads=8)
    check(helpers.compare(ro, rg))


@pytest.mark.parametrize("name", ["boxbot", "spherebot"])
@pytest.mark.parametrize("evaluator", [capi.EVAL_DISCRETE, capi.EVAL_LVS_DISCRETE, capi.EVAL_LVS_CONTINUOUS])
def test_small_bots(cuda, oracle, name, evaluator):
    s, sg, so = pair(cuda, oracle, name, 0.2, 1.0, 0.05)
    q = np.array([[[-1.9, 0.0], [0.0, 1.9], [1.9, 3.8]], [[-0.75, 0.75], [-0.2, 0.3], [0.4, 0.1]]])
    rg = cuda.eval(sg, q, evaluator, lvs=0.05)
    ro = oracle.eval(so, q, evaluator, lvs=0.05)
    rep = helpers.compare(ro, rg)
    assert rep["n_contacts"] > 0
    # face-face box contacts have non-unique witness points; the robots are prismatic so rows are immune
    check(rep)


@pytest.mark.parametrize("evaluator,max_cnt", [(capi.EVAL_DISCRETE, 3), (capi.EVAL_LVS_CONTINUOUS, 3),
                                               (capi.EVAL_LVS_DISCRETE, 2)])
def test_fixed_size_sets(cuda, oracle, evaluator, max_cnt):
    s, sg, so = pair(cuda, oracle, "arm_around_table", 0.025, 20.0, 0.01)
    q = s["seed_traj"][None]
    fixed = np.zeros(q.shape[1], np.uint8)
    fixed[0] = fixed[-1] = 1
    rg = cuda.eval(sg, q, evaluator, lvs=0.02, fixed=fixed)
    ro = oracle.eval(so, q, evaluator, lvs=0.02, fixed=fixed)
    a, b = ro.sets(max_cnt), rg.sets(max_cnt)
    assert (a["set_ptr"] == b["set_ptr"]).all()
    assert (a["sets"]["pair"] == b["sets"]["pair"]).all()
    assert np.abs(a["values"] - b["values"]).max() <= 1e-6
    assert np.abs(a["coeffs"] - b["coeffs"]).max() == 0
    den = np.maximum(np.abs(a["jac"]).max(axis=-1, keepdims=True), 1e-12)
    assert (np.abs(a["jac"] - b["jac"]) / den).max() <= ROW_TOL


def test_empty_and_errors(cuda):
    s = helpers.load_scene("boxbot")
    sg = cuda.scene(s, 0.1, 1.0, 0.01)
    r = cuda.eval(sg, np.array([[[5.0, 5.0]]]), capi.EVAL_DISCRETE)  # far away: no contacts
    assert r.counts()["n_contacts"] == 0 and len(r.contacts()) == 0
    assert r.rows()["row_ptr"].tolist() == [0]
    assert r.sets(2)["values"].tolist() == [[-0.01, -0.01]]
    r = cuda.eval(sg, np.zeros((1, 2, 2)), capi.EVAL_LVS_CONTINUOUS, fixed=[1, 1])  # all fixed: no terms
    assert r.counts()["n_terms"] == 0
    with pytest.raises(capi.TcError):
        cuda.eval(sg, np.zeros((1, 1, 2)), capi.EVAL_LVS_CONTINUOUS)  # segment evaluator needs T >= 2
    with pytest.raises(capi.TcError):
        cuda.eval(sg, np.zeros((1, 2, 2)), 7)


@pytest.mark.parametrize("evaluator,stack", [(capi.EVAL_LVS_CONTINUOUS, capi.STACK_IFOPT),
                                             (capi.EVAL_LVS_CONTINUOUS, capi.STACK_LEGACY),
                                             (capi.EVAL_DISCRETE, capi.STACK_IFOPT)])
def test_rows_folded_inside_the_pass_equal_rows_on_demand(cuda, evaluator, stack):
    """TC_WANT_ROWS without TC_WANT_GRADIENTS (what bench.py times) gives bit-identical rows"""
    s = helpers.load_scene("arm_around_table")
    sg = cuda.scene(s, 0.025, 20.0, 0.01)
    q = helpers.pkg.multistart.make_batch(s, 50, 8, t_steps=6)
    fixed = np.zeros(q.shape[1], np.uint8)
    fixed[0] = fixed[-1] = 1
    a = cuda.eval(sg, q, evaluator, stack=stack, lvs=0.02, fixed=fixed).rows()
    rb = cuda.eval(sg, q, evaluator, stack=stack, lvs=0.02, fixed=fixed, want=capi.WANT_LINEARISE)
    b = rb.rows()
    assert rb.counts()["nnz"] == len(b["val"]) > 0
    for k in ("row_ptr", "col_idx", "val", "value", "coeff", "margin"):
        assert (a[k] == b[k]).all(), k
    with pytest.raises(capi.TcError):
        rb.gradients()


def test_block_rows_equal_csr_rows(cuda):
    """tc_result_rows_blocks carries the same rows as tc_result_rows; the column indices follow from the block form"""
    s = helpers.load_scene("arm_around_table")
    sg = cuda.scene(s, 0.025, 20.0, 0.01)
    q = s["seed_traj"][None]
    fixed = np.zeros(q.shape[1], np.uint8)
    fixed[0] = fixed[-1] = 1
    for ev in (capi.EVAL_LVS_CONTINUOUS, capi.EVAL_DISCRETE):
        for want in (capi.WANT_LINEARISE, capi.WANT_ALL):
            r = cuda.eval(sg, q, ev, lvs=0.02, fixed=fixed, want=want)
            a, b = r.rows(), r.rows_blocks()
            assert len(a["val"]) > 0
            assert np.array_equal(a["row_ptr"], b["row_ptr"]) and np.array_equal(a["val"], b["val"])
            assert np.array_equal(a["value"], b["value"]) and np.array_equal(a["coeff"], b["coeff"])
            assert np.array_equal(a["col_idx"], capi.expand_columns(b["row_ptr"], b["first_col"], sg.n_dof))


@pytest.mark.parametrize("stack", [capi.STACK_LEGACY, capi.STACK_IFOPT])
@pytest.mark.parametrize("evaluator", [capi.EVAL_LVS_CONTINUOUS, capi.EVAL_LVS_DISCRETE, capi.EVAL_DISCRETE])
def test_qp_csc_block_matches_rows(cuda, oracle, stack, evaluator):
    """device-assembled hinge block [A_x | -I] (tc_result_qp_csc) against the same matrix built on the
    host with scipy from the rows (which the other tests hold against the oracle): identical structure and values"""
    import scipy.sparse as sp
    s, sg, so = pair(cuda, oracle, "arm_around_table", 0.025, 20.0, 0.01)
    q = helpers.pkg.multistart.make_batch(s, 7, 3, t_steps=6)
    B, T, n = q.shape
    fixed = np.zeros(T, np.uint8)
    fixed[0] = fixed[-1] = 1
    rg = cuda.eval(sg, q, evaluator, stack=stack, lvs=0.02, fixed=fixed, want=capi.WANT_ALL)
    m = rg.qp_csc()
    rows, co = rg.rows(), rg.contacts()
    check(helpers.compare(oracle.eval(so, q, evaluator, stack=stack, lvs=0.02, fixed=fixed), rg))
    nr = len(co)
    assert m["shape"] == (nr, B * T * n + nr) and nr > 0
    # host construction: global column = traj * T * n + local column
    counts = np.diff(rows["row_ptr"])
    ri = np.repeat(np.arange(nr), counts)
    gc = rows["col_idx"].astype(np.int64) + np.repeat(co["traj"].astype(np.int64), counts) * T * n
    sign = -1.0 if stack == capi.STACK_LEGACY else 1.0
    ax = sp.csr_matrix((sign * rows["val"], (ri, gc)), shape=(nr, B * T * n))
    want = sp.hstack([ax, -sp.identity(nr)]).tocsc()
    want.sort_indices()
    assert np.array_equal(m["col_ptr"], want.indptr)
    assert np.array_equal(m["row_idx"], want.indices)
    assert np.array_equal(m["val"], want.data)
    if stack == capi.STACK_LEGACY:
        upper = rows["value"] - rows["margin"]
    else:
        x0 = q.reshape(B, T * n)
        jx0 = np.array([rows["val"][rows["row_ptr"][i]:rows["row_ptr"][i + 1]] @
                        x0[co["traj"][i]][rows["col_idx"][rows["row_ptr"][i]:rows["row_ptr"][i + 1]]] for i in range(nr)])
        upper = jx0 - rows["value"]
    assert np.abs(m["upper"] - upper).max() <= 1e-12 * max(1.0, np.abs(upper).max())
    assert np.array_equal(m["slack_cost"], rows["coeff"])


def test_dynamic_environment_pose_update(cuda, oracle):
    """links outside the planning group moved between evaluations (tc_scene_set_link_poses): same results as a scene
    created with the new poses (collision_terms.cpp:399-411,659-660: dynamic_environment)"""
    s = helpers.load_scene("arm_around_table")
    sg = cuda.scene(s, 0.025, 20.0, 0.01)
    q = s["seed_traj"][None]
    fixed = np.zeros(q.shape[1], np.uint8)
    fixed[0] = fixed[-1] = 1
    before = cuda.eval(sg, q, capi.EVAL_LVS_CONTINUOUS, lvs=0.02, fixed=fixed).counts()["n_contacts"]
    moved = dict(s)
    lp = np.array(s["link_pose"], dtype=np.float64).reshape(-1, 12).copy()
    table = [i for i, nm in enumerate(s["link_names"]) if str(nm) == "table_link"][0]
    lp[table, [3, 7, 11]] += np.array([-0.12, 0.05, 0.08])  # the table comes closer to the arm
    moved["link_pose"] = lp
    sg.set_link_poses(lp)
    rg = cuda.eval(sg, q, capi.EVAL_LVS_CONTINUOUS, lvs=0.02, fixed=fixed)
    ro = oracle.eval(oracle.scene(moved, 0.025, 20.0, 0.01), q, capi.EVAL_LVS_CONTINUOUS, lvs=0.02, fixed=fixed)
    assert ro.counts()["n_contacts"] != before
    check(helpers.compare(ro, rg))
    fresh = cuda.eval(cuda.scene(moved, 0.025, 20.0, 0.01), q, capi.EVAL_LVS_CONTINUOUS, lvs=0.02, fixed=fixed)
    assert np.array_equal(fresh.contacts()["distance"], rg.contacts()["distance"])


@pytest.mark.parametrize("evaluator", [capi.EVAL_DISCRETE, capi.EVAL_LVS_DISCRETE, capi.EVAL_LVS_CONTINUOUS])
def test_primitive_shapes(cuda, oracle, evaluator):
    """every shape type of the C ABI (sphere, capsule, cylinder, cone, hull; compound link) through the whole pass"""
    from tests import synthetic
    s = synthetic.primitives_scene()
    sg, so = cuda.scene(s, 0.1, 1.0, 0.05), oracle.scene(s, 0.1, 1.0, 0.05)
    q = np.random.default_rng(3).uniform(-1.6, 1.6, (8, 5, 2))
    rg = cuda.eval(sg, q, evaluator, lvs=0.05)
    ro = oracle.eval(so, q, evaluator, lvs=0.05)
    rep = helpers.compare(ro, rg)
    assert rep["n_contacts"] > 50, rep
    # Cylinder and cone pairs run through the contraction-free kernels of tc_strict.cu (FK with the shared sin/cos,
    # GJK/EPA operation for operation as the oracle), so the north_star tolerances hold for every shape type: 1e-6 m,
    # normals / gradients / rows 1e-5 — the same `check` as every other scene.  (Round 1 needed 1e-4 / 1e-2 here.)
    check(rep)
    assert rep["cc_type_mismatch"] == 0, rep


@pytest.mark.parametrize("evaluator", [capi.EVAL_LVS_DISCRETE, capi.EVAL_LVS_CONTINUOUS])
def test_curved_obstacles_on_a_revolute_arm(cuda, oracle, evaluator):
    """cylinders and cones against hull links whose poses come out of a revolute (sin/cos) FK chain: the contraction-free
    FK + narrow phase (tc_strict.cu) keep the north_star tolerances there too"""
    from tests import synthetic
    s = synthetic.arm_with_curved_obstacles_scene()
    sg, so = cuda.scene(s, 0.05, 20.0, 0.02), oracle.scene(s, 0.05, 20.0, 0.02)
    q = helpers.pkg.multistart.make_batch(s, 300, 6, t_steps=6)
    fixed = np.zeros(q.shape[1], np.uint8)
    fixed[0] = fixed[-1] = 1
    rg = cuda.eval(sg, q, evaluator, lvs=0.05, fixed=fixed)
    ro = oracle.eval(so, q, evaluator, lvs=0.05, fixed=fixed, n_threads=8)
    rep = helpers.compare(ro, rg)
    curved = np.isin(ro.contacts()["link_ids"][:, 1], np.nonzero(np.char.startswith(s["link_names"].astype(str), "curved"))[0])
    assert curved.sum() > 30, (rep, int(curved.sum()))
    check(rep)


def test_speculative_pass_equals_exact_pass(cuda):
    """A result's second pass of the same shape is sized from its first (no read-back inside the call).  Whatever the
    new iterate does to the counts — more contacts than predicted (device flag, exact repeat inside tc_result_sync) or
    fewer — contacts and rows are bit-identical to those of a fresh result evaluated exactly."""
    s = helpers.load_scene("arm_around_table")
    sg = cuda.scene(s, 0.025, 20.0, 0.01)
    ms = helpers.pkg.multistart
    calm = np.repeat(s["seed_traj"][None], 16, 0)              # the fixture path: few contacts
    wild = ms.make_batch(s, 900, 16, t_steps=calm.shape[1])    # sigma 0.3 rad around it: many more
    fixed = np.zeros(calm.shape[1], np.uint8)
    fixed[0] = fixed[-1] = 1
    res = capi.Result(cuda, sg, cuda.new_result(sg), 16, calm.shape[1])
    seq = [calm, wild, wild, calm, calm]
    want_spec = [False, True, True, True, True]
    want_rep = [0, 1, 1, 1, 1]
    for q, ws, wr in zip(seq, want_spec, want_rep):
        r = cuda.eval(sg, q, capi.EVAL_LVS_CONTINUOUS, lvs=0.02, fixed=fixed, want=capi.WANT_LINEARISE, result=res)
        spec, rep = cuda.pass_info(res)
        assert rep == wr and (spec == ws or rep == wr), (spec, rep, ws, wr)
        fresh = cuda.eval(sg, q, capi.EVAL_LVS_CONTINUOUS, lvs=0.02, fixed=fixed, want=capi.WANT_LINEARISE)
        a, b = r.contacts(), fresh.contacts()
        assert len(a) == len(b) and a.tobytes() == b.tobytes()
        ra, rb = r.rows(), fresh.rows()
        for k in ("row_ptr", "col_idx", "val", "value"):
            assert np.array_equal(ra[k], rb[k]), k
        ca, cb = r.counts(), fresh.counts()
        assert all(ca[k] == cb[k] for k in ca if k != "overflow"), (ca, cb)
    assert cuda.pass_info(res)[1] >= 1  # the wilder iterate outgrew the prediction at least once


def test_eval_async_returns_before_the_stream_is_idle(cuda):
    """tc_eval_async on a warmed-up result only enqueues: right after it returns the stream still has work"""
    import torch
    s = helpers.load_scene("arm_around_table")
    sg = cuda.scene(s, 0.025, 20.0, 0.01)
    ms = helpers.pkg.multistart
    B = 1024
    q = torch.from_numpy(ms.make_batch(s, 0, B)).cuda()
    res = capi.Result(cuda, sg, cuda.new_result(sg), B, ms.T_STEPS)
    stream = torch.cuda.Stream()
    kw = dict(lvs=ms.LVS, fixed=ms.fixed_steps(), stream=stream.cuda_stream, want=capi.WANT_LINEARISE)
    torch.cuda.synchronize()
    cuda.eval_device(sg, q.data_ptr(), B, ms.T_STEPS, capi.EVAL_LVS_CONTINUOUS, res, **kw)  # exact pass, synchronised
    busy = 0
    for _ in range(3):
        cuda.eval_device(sg, q.data_ptr(), B, ms.T_STEPS, capi.EVAL_LVS_CONTINUOUS, res, sync=False, **kw)
        busy += 0 if stream.query() else 1
        cuda.sync(res)
        assert cuda.pass_info(res) == (True, 0)
    assert busy == 3, busy
    assert res.counts()["n_contacts"] > 100000


@pytest.mark.parametrize("n_slots,chunk", [(3, 16), (1, 16), (2, 64)])
def test_pipelined_host_pass_equals_one_pass(cuda, n_slots, chunk):
    """tc_pipeline_eval_rows (sub-batches on a ring of streams, compact rows) delivers bit for bit the rows, costs
    and term offsets of ONE tc_eval over the whole batch"""
    s = helpers.load_scene("arm_around_table")
    sg = cuda.scene(s, 0.025, 20.0, 0.01)
    ms = helpers.pkg.multistart
    B, T = 50, 6
    q = ms.make_batch(s, 200, B, t_steps=T)
    fixed = np.zeros(T, np.uint8)
    fixed[0] = fixed[-1] = 1
    ref = cuda.eval(sg, q, capi.EVAL_LVS_CONTINUOUS, lvs=0.02, fixed=fixed, want=capi.WANT_LINEARISE)
    rows, co = ref.rows(), ref.contacts()
    pipe = capi.Pipeline(cuda, sg, n_slots)
    tpt = T - 1
    pipe.reserve(B, tpt, len(co) + 10, len(rows["val"]) + 10)
    for rep in range(3):  # the later rounds run speculatively on warmed-up slots
        n_rows, nnz = pipe.eval_rows(q, capi.EVAL_LVS_CONTINUOUS, 0.02, fixed, chunk)
        o = pipe.out
        assert n_rows == len(co) and nnz == len(rows["val"])
        assert np.array_equal(o["term_ptr"], ref.term_ptr())
        assert np.array_equal(o["val"][:nnz], rows["val"]) and np.array_equal(o["value"][:n_rows], rows["value"])
        assert np.array_equal(o["pair"][:n_rows], co["pair"])
        assert np.array_equal(o["cost"], ref.cost()[1])
        blocks = np.tile([1] + [2] * (tpt - 2) + [1], B)  # first / last segment have a fixed end
        assert np.array_equal(capi.expand_compact_rows(o["term_ptr"], o["val"][:nnz], sg.n_dof, blocks), rows["row_ptr"])
    _, _, margin, coeff = sg.pairs()
    assert np.array_equal(margin[o["pair"][:n_rows]], rows["margin"]) and np.array_equal(coeff[o["pair"][:n_rows]], rows["coeff"])
    pipe.close()
    with pytest.raises(capi.TcError):
        small = capi.Pipeline(cuda, sg, 2)
        small.reserve(B, tpt, 5, 5)
        small.eval_rows(q, capi.EVAL_LVS_CONTINUOUS, 0.02, fixed, chunk)


@pytest.mark.parametrize("scene_name,evaluator,stack", [("arm_around_table", capi.EVAL_LVS_CONTINUOUS, capi.STACK_IFOPT),
                                                        ("arm_around_table", capi.EVAL_LVS_CONTINUOUS, capi.STACK_LEGACY),
                                                        ("arm_around_table", capi.EVAL_DISCRETE, capi.STACK_IFOPT),
                                                        ("dual_arm", capi.EVAL_LVS_CONTINUOUS, capi.STACK_IFOPT)])
def test_packed_rows_unpack_to_the_dense_rows(cuda, scene_name, evaluator, stack):
    """tc_pipeline_eval_rows_packed leaves out exactly the joints that move neither link of a row: unpacked with
    tc_scene_pair_dof_mask the values are those of ONE dense pass, every one of them, and what was left out is zero"""
    ms = helpers.pkg.multistart
    if scene_name == "dual_arm":
        from tests import synthetic
        s = synthetic.dual_arm_scene()  # 16 joints: the general row kernel
        B, T = 6, 8
        rng = np.random.default_rng(5)
        q = s["traj"][::12][:T][None] + rng.normal(0, 0.05, (B, T, int(s["n_dof"])))
    else:
        s = helpers.load_scene(scene_name)
        B, T = 40, 6
        q = ms.make_batch(s, 300, B, t_steps=T)
    sg = cuda.scene(s, 0.025, 20.0, 0.01)
    fixed = np.zeros(T, np.uint8)
    fixed[0] = 1
    ref = cuda.eval(sg, q, evaluator, stack=stack, lvs=0.02, fixed=fixed, want=capi.WANT_LINEARISE)
    rows, co = ref.rows(), ref.contacts()
    assert len(co) > 0
    single = evaluator == capi.EVAL_DISCRETE
    tpt = (T - 1) if single else (T - 1)  # one fixed waypoint: T - 1 states, or T - 1 segments (the first with one block)
    blocks_traj = [1] * tpt if single else [1] + [2] * (tpt - 1)
    pipe = capi.Pipeline(cuda, sg, 3)
    pipe.reserve(B, tpt, len(co) + 10, len(rows["val"]) + 10)
    masks = pipe.pair_masks(len(sg.pairs()[0]))
    for rep in range(3):
        n_rows, nnz = pipe.eval_rows(q, evaluator, 0.02, fixed, 16, stack=stack, packed=True)
        o = pipe.out
        assert n_rows == len(co) and 0 < nnz < len(rows["val"])
        assert np.array_equal(o["term_ptr"], ref.term_ptr()) and np.array_equal(o["pair"][:n_rows], co["pair"])
        row_ptr, dense = capi.unpack_rows(o["term_ptr"], o["val"][:nnz], o["pair"][:n_rows], masks, sg.n_dof,
                                          np.tile(blocks_traj, B))
        assert np.array_equal(row_ptr, rows["row_ptr"])
        assert np.array_equal(dense, rows["val"])  # -0.0 == 0.0: the entries left out are zeros of either sign
        assert np.array_equal(o["value"][:n_rows], rows["value"]) and np.array_equal(o["cost"], ref.cost()[1])
    pipe.close()


@pytest.mark.parametrize("evaluator", [capi.EVAL_DISCRETE, capi.EVAL_LVS_CONTINUOUS])
def test_dynamic_environment_moving_links_outside_the_group(cuda, oracle, evaluator):
    """collision_terms.cpp:399-411,828-836: with links that move outside the planning group the manager is active on
    every moving link, so those links are tested against the arm, the static scene AND each other; pairs without a group
    link have no gradient (constant rows) but count in the cost.  Then the links move (tc_scene_set_link_poses)."""
    from tests import synthetic
    s = synthetic.dynamic_env_scene()
    sg, so = cuda.scene(s, 0.05, 20.0, 0.02), oracle.scene(s, 0.05, 20.0, 0.02)
    assert sg.num_pairs() == so.num_pairs() > helpers.load_scene("arm_around_table")["acm"].shape[0]
    q = helpers.pkg.multistart.make_batch(helpers.load_scene("arm_around_table"), 70, 4, t_steps=5)
    fixed = np.zeros(q.shape[1], np.uint8)
    fixed[0] = 1
    na = int(s["n_active_links"])
    for step in range(2):
        rg = cuda.eval(sg, q, evaluator, lvs=0.05, fixed=fixed)
        ro = oracle.eval(so, q, evaluator, lvs=0.05, fixed=fixed, n_threads=8)
        rep = helpers.compare(ro, rg)
        check(rep)
        assert rep["same_order"] and rep["term_ptr_equal"]
        c = rg.contacts()
        env_env = (c["link_ids"][:, 0] >= na) & (c["link_ids"][:, 1] >= na)
        env_arm = (c["link_ids"][:, 0] < na) & (c["link_ids"][:, 1] >= na) & (c["link_ids"][:, 1] < na + 2)
        assert env_env.sum() > 0 and env_arm.sum() > 0, (int(env_env.sum()), int(env_arm.sum()))
        rows = rg.rows()
        for i in np.nonzero(env_env)[0][:20]:
            assert not rows["val"][rows["row_ptr"][i]:rows["row_ptr"][i + 1]].any()  # no gradient for this group
        # the cart rolls 4 cm: new scene poses on both sides
        lp = np.array(s["link_pose"], dtype=np.float64).reshape(-1, 12).copy()
        lp[na:na + 2, 3] += 0.04
        s = dict(s)
        s["link_pose"] = lp
        sg.set_link_poses(lp)
        so = oracle.scene(s, 0.05, 20.0, 0.02)
