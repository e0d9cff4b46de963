"""Numerical-gradient collision constraints and batched trajectory validation (8f N2):
host logic in trajopt-1_b200/constraints.py, driven by the oracle on the CPU and by the CUDA path on the GPU."""
import numpy as np
import pytest

from tests import helpers

capi = helpers.capi
cons = helpers.pkg.constraints

# the boxbot known answers of trajopt_ifopt/test/collision_unit.cpp:103-158 (margin 0.1, coeff 1, buffer 0.01)
KATS = [((-1.9, 0.01), -0.01, (0, 0)), ((-1.0, 0.01), 0.1, (1, 0)), ((0.01, 1.0), 0.1, (0, -1)),
        ((-0.9, 0.01), 0.2, (1, 0))]


def _boxbot_numerical(lib):
    s = helpers.load_scene("boxbot")
    sc = lib.scene(s, 0.1, 1.0, 0.01)
    c = cons.DiscreteCollisionNumericalConstraint(lib, sc, max_num_cnt=1)
    for q, val, jac in KATS:
        values, coeffs, J = c.evaluate(np.array(q))
        assert abs(values[0] - val) <= 1e-6, (q, values)
        # forward differences with delta = 1e-8 of a distance good to ~1e-10: 1e-2 absolute is what the scheme gives
        assert np.abs(J[0] - np.array(jac)).max() <= 2e-2, (q, J)
        assert coeffs[0] == 1.0
    assert c.passes == len(KATS)  # one batched pass per evaluate(), not n_dof + 1


def test_numerical_constraint_reproduces_boxbot_kats(oracle):
    _boxbot_numerical(oracle)


def test_numerical_discrete_matches_analytic_rows(oracle):
    """single-timestep sets hold one contact per shape pair, so d(max error)/dq is the analytic row"""
    s = helpers.load_scene("arm_around_table")
    sc = oracle.scene(s, 0.025, 20.0, 0.01)
    q = s["seed_traj"][2]
    values, coeffs, J = cons.DiscreteCollisionNumericalConstraint(oracle, sc, max_num_cnt=3).evaluate(q)
    st = oracle.eval(sc, q[None, None], capi.EVAL_DISCRETE).sets(3)
    n = min(3, int(st["set_ptr"][1]))
    assert n > 0
    assert np.abs(values[:n] - st["values"][0, :n]).max() <= 1e-12
    assert np.abs(coeffs[:n] - st["coeffs"][0, :n]).max() == 0
    assert np.abs(J[:n] - st["jac"][0, :n, 0]).max() <= 2e-2, (J[:n], st["jac"][0, :n, 0])
    assert np.all(values[n:] == -0.01) and np.all(J[n:] == 0)


def test_numerical_continuous_argument_checks(oracle):
    s = helpers.load_scene("boxbot")
    sc = oracle.scene(s, 0.1, 1.0, 0.01)
    with pytest.raises(RuntimeError, match="both fixed"):
        cons.ContinuousCollisionNumericalConstraint(oracle, sc, capi.EVAL_LVS_CONTINUOUS, 0.05, True, True)
    with pytest.raises(RuntimeError, match="max_num_cnt"):
        cons.ContinuousCollisionNumericalConstraint(oracle, sc, capi.EVAL_LVS_CONTINUOUS, 0.05, max_num_cnt=0)
    with pytest.raises(RuntimeError, match="max_num_cnt"):
        cons.DiscreteCollisionNumericalConstraint(oracle, sc, max_num_cnt=0)


def test_numerical_continuous_boxbot_cast(oracle):
    """box_cast path (trajopt_common/data/config/box_cast_test.json:41): the swept box hits the obstacle between the
    first two waypoints; the free end state gets a Jacobian block, a far-away segment returns -buffer and zeros"""
    s = helpers.load_scene("boxbot")
    sc = oracle.scene(s, 0.1, 1.0, 0.01)
    c = cons.ContinuousCollisionNumericalConstraint(oracle, sc, capi.EVAL_LVS_CONTINUOUS, 0.05, max_num_cnt=1)
    values, coeffs, J = c.evaluate(np.array([-1.9, 0.0]), np.array([0.0, 1.9]))
    assert values[0] > 0 and np.abs(J).max() > 0.1
    values, coeffs, J = c.evaluate(np.array([-5.0, 5.0]), np.array([-5.0, 6.0]))
    assert values[0] == -0.01 and np.all(J == 0)
    assert c.passes == 2


def test_check_trajectories_oracle(oracle):
    """the reference's planning tests: the seed trajectory collides (planning_unit.cpp:99-100)"""
    s = helpers.load_scene("arm_around_table")
    sc = oracle.scene(s, 0.025, 20.0, 0.01)
    ms = helpers.pkg.multistart
    q = np.concatenate([s["seed_traj"][None], np.repeat(s["seed_traj"][:1][None], len(s["seed_traj"]), axis=1)])
    out = cons.check_trajectories(oracle, sc, q, capi.EVAL_LVS_CONTINUOUS, lvs=0.02)
    assert out["in_collision"][0] and out["min_distance"][0] < 0
    # the second "trajectory" stands still at the start state
    still = oracle.eval(sc, q[1:2, :2], capi.EVAL_CONTINUOUS, want=capi.WANT_CONTACTS).contacts()
    want = still["distance"].min() if len(still) else np.inf
    assert out["min_distance"][1] == pytest.approx(want, abs=1e-9)
    assert ms is not None


@pytest.mark.gpu
def test_numerical_constraints_gpu_vs_oracle(cuda, oracle):
    _boxbot_numerical(cuda)
    s = helpers.load_scene("arm_around_table")
    sg, so = cuda.scene(s, 0.025, 20.0, 0.01), oracle.scene(s, 0.025, 20.0, 0.01)
    traj = s["seed_traj"]
    for t in (1, 2, 3):
        vg, cg, Jg = cons.DiscreteCollisionNumericalConstraint(cuda, sg, 3).evaluate(traj[t])
        vo, co, Jo = cons.DiscreteCollisionNumericalConstraint(oracle, so, 3).evaluate(traj[t])
        assert np.abs(vg - vo).max() <= 1e-6 and np.abs(cg - co).max() == 0
        # delta = 1e-8 turns FMA-level distance differences (~1e-12) into 1e-4; EPA's 1e-10 stop into 1e-2
        assert np.abs(Jg - Jo).max() <= 2e-2, (t, Jg, Jo)
    for f0, f1 in ((False, False), (True, False), (False, True)):
        a = cons.ContinuousCollisionNumericalConstraint(cuda, sg, capi.EVAL_LVS_CONTINUOUS, 0.02, f0, f1, 3)
        b = cons.ContinuousCollisionNumericalConstraint(oracle, so, capi.EVAL_LVS_CONTINUOUS, 0.02, f0, f1, 3)
        vg, cg, Jg = a.evaluate(traj[1], traj[2])
        vo, co, Jo = b.evaluate(traj[1], traj[2])
        assert np.abs(vg - vo).max() <= 1e-6 and np.abs(cg - co).max() == 0
        assert np.abs(Jg - Jo).max() <= 2e-2, (f0, f1, Jg, Jo)
        assert a.passes == 1


@pytest.mark.gpu
def test_check_trajectories_gpu_vs_oracle(cuda, oracle):
    s = helpers.load_scene("arm_around_table")
    sg, so = cuda.scene(s, 0.025, 20.0, 0.01), oracle.scene(s, 0.025, 20.0, 0.01)
    q = helpers.pkg.multistart.make_batch(s, 0, 24)
    for ev in (capi.EVAL_LVS_CONTINUOUS, capi.EVAL_LVS_DISCRETE):
        g = cons.check_trajectories(cuda, sg, q, ev, lvs=0.02)
        o = cons.check_trajectories(oracle, so, q, ev, lvs=0.02)
        assert np.array_equal(g["in_collision"], o["in_collision"])
        fin = np.isfinite(o["min_distance"])
        assert np.array_equal(fin, np.isfinite(g["min_distance"]))
        assert np.abs(g["min_distance"][fin] - o["min_distance"][fin]).max() <= 1e-6
        assert np.array_equal(g["n_contacts"], o["n_contacts"])


def test_expand_columns_rebuilds_csr_indices():
    """block form of the rows (tc_result_rows_blocks) -> CSR column indices, incl. absent blocks and an empty result"""
    first = np.array([[0, 7], [-1, 14], [7, -1], [21, 28]], np.int32)
    row_ptr = np.array([0, 14, 21, 28, 42], np.int64)
    col = capi.expand_columns(row_ptr, first, 7)
    want = np.concatenate([np.arange(0, 14), np.arange(14, 21), np.arange(7, 14), np.arange(21, 35)])
    assert np.array_equal(col, want)
    assert len(capi.expand_columns(np.array([0], np.int64), np.zeros((0, 2), np.int32), 7)) == 0
