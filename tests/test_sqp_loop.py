"""Closing the loop: the same trust-region SQP (tests/sqp.py, a restatement of trajopt's BasicTrustRegionSQP with an
in-repo OSQP-style QP) driven by the oracle and by the CUDA library.  Mirrors the reference's outcome tests
(trajopt/test/planning_unit.cpp:99-148, cast_cost_unit.cpp): the initial trajectory collides, the optimised one
keeps the safety margin; and the north_star's end-to-end parity: identical converged trajectories within 1e-4 rad."""
import numpy as np
import pytest

from tests import helpers, sqp

capi = helpers.capi


def _arm(lib):
    s = helpers.load_scene("arm_around_table")
    return sqp.CollisionProblem(lib, s, s["seed_traj"], capi.EVAL_LVS_CONTINUOUS, 0.025, 20.0, 0.01, 0.02)


def _boxbot_cast(lib):
    # trajopt_common/data/config/box_cast_test.json: straight path through the obstacle, 3 waypoints
    s = helpers.load_scene("boxbot")
    return sqp.CollisionProblem(lib, s, np.array([[-1.9, 0.0], [0.0, 1.9], [1.9, 3.8]]), capi.EVAL_LVS_CONTINUOUS, 0.2,
                                10.0, 0.05, 0.05)


def _min_distance(lib, prob, traj):
    r = lib.eval(prob.sc, traj[None], prob.evaluator, stack=capi.STACK_LEGACY, lvs=prob.lvs, fixed=prob.fixed)
    c = r.contacts()
    return c["distance"].min() if len(c) else np.inf


@pytest.mark.parametrize("make", [_boxbot_cast, _arm])
def test_oracle_driven_sqp_clears_the_collision(oracle, make):
    prob = make(oracle)
    assert _min_distance(oracle, prob, prob.q0) < 0  # the seed trajectory is in collision
    c0 = prob.cost(prob.q0[1:-1].reshape(-1))
    traj, cost, status = sqp.optimize(prob)
    assert status.startswith("converged") and cost < c0, (status, c0, cost)
    assert _min_distance(oracle, prob, traj) > 0  # collision free, as checkTrajectory asserts in the reference tests
    assert (traj[0] == prob.q0[0]).all() and (traj[-1] == prob.q0[-1]).all()


@pytest.mark.gpu
@pytest.mark.parametrize("make", [_boxbot_cast, _arm])
def test_cuda_and_oracle_converge_to_the_same_trajectory(cuda, oracle, make):
    po, pg = make(oracle), make(cuda)
    to, co, so = sqp.optimize(po)
    tg, cg, sg = sqp.optimize(pg)
    assert so == sg and po.passes == pg.passes, (so, sg, po.passes, pg.passes)
    assert np.abs(to - tg).max() <= 1e-4, np.abs(to - tg).max()  # rad (m for the prismatic boxbot)
    assert abs(co - cg) <= 1e-6 * max(1.0, abs(co))
    assert _min_distance(cuda, pg, tg) > 0
