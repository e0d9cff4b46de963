"""Seam B3: the contact manager driven directly — link poses in, contacts out (tc_contact_test) —
and the ContactTestType the evaluators pass through (ALL / CLOSEST / FIRST, tc_eval_desc.contact_test_type).
Reference call sites: trajopt/src/collision_terms.cpp:616-625,665-667,887-889,1132-1134 (setCollisionObjectsTransform +
contactTest), trajopt/src/problem_description.cpp:1669-1673 (contact_test_type)."""
import numpy as np
import pytest

from tests import helpers

capi = helpers.capi
ms = helpers.pkg.multistart


def _arm_states(oracle, so, s, n, seed=5, sigma=0.4):
    rng = np.random.default_rng(seed)
    lim = s["joint_limits"]
    base = s["seed_traj"] if "seed_traj" in s else s["traj"]
    q = np.clip(base[rng.integers(0, len(base), n)] + rng.normal(0, sigma, (n, int(s["n_dof"]))), lim[:, 0], lim[:, 1])
    return q, np.stack([oracle.fk(so, qi) for qi in q])


def _scene_for(ctype):
    """CLOSEST only differs from ALL on compound links (several shape pairs per link pair): the dual-arm scene of C5"""
    if ctype == capi.CONTACT_CLOSEST:
        from tests import synthetic
        return synthetic.dual_arm_scene(), 0.15
    return helpers.load_scene("arm_around_table"), 0.5


def _same_contacts(a, b, tol=1e-12):
    ca, cb = a.contacts(), b.contacts()
    assert len(ca) == len(cb)
    for f in ("traj", "pair", "link_ids", "shape_id", "cc_type"):
        assert np.array_equal(ca[f], cb[f]), f
    assert np.abs(ca["distance"] - cb["distance"]).max(initial=0) <= tol
    assert np.abs(ca["normal"] - cb["normal"]).max(initial=0) <= 1e-9
    assert np.abs(ca["cc_time"] - cb["cc_time"]).max(initial=0) <= 1e-9


def test_oracle_contact_test_equals_its_evaluators():
    """CPU: the manager driven with FK poses reports what the DISCRETE / CONTINUOUS evaluators report from joint values"""
    from oracle import pyoracle
    oracle = pyoracle.OracleLib()
    s = helpers.load_scene("arm_around_table")
    so = oracle.scene(s, 0.025, 20.0, 0.01)
    q, poses = _arm_states(oracle, so, s, 40)
    a = oracle.contact_test(so, poses)
    b = oracle.eval(so, q[:, None, :], capi.EVAL_DISCRETE)
    assert len(a.contacts()) > 20
    _same_contacts(a, b)
    assert np.array_equal(a.term_ptr(), b.term_ptr())
    a = oracle.contact_test(so, poses[:-1], poses[1:])
    b = oracle.eval(so, np.stack([q[:-1], q[1:]], axis=1), capi.EVAL_CONTINUOUS)
    assert len(a.contacts()) > 20 and (a.contacts()["cc_type"] != 0).any()
    _same_contacts(a, b)


@pytest.mark.parametrize("ctype", [capi.CONTACT_CLOSEST, capi.CONTACT_FIRST])
def test_oracle_contact_types_are_subsets_with_the_stated_choice(oracle, ctype):
    """CPU: CLOSEST keeps the closest result of every (state, link pair), FIRST the first result of every state"""
    s, sigma = _scene_for(ctype)
    so = oracle.scene(s, 0.05, 20.0, 0.02)
    q, poses = _arm_states(oracle, so, s, 10, sigma=sigma)
    call, sub = oracle.contact_test(so, poses).contacts(), oracle.contact_test(so, poses, contact_type=ctype).contacts()
    assert 0 < len(sub) < len(call)
    for st in range(len(q)):
        a, b = call[call["traj"] == st], sub[sub["traj"] == st]
        if ctype == capi.CONTACT_FIRST:
            assert len(b) == min(1, len(a))
            if len(a):
                assert b[0]["pair"] == a[0]["pair"] and b[0]["distance"] == a[0]["distance"]
        else:
            assert sorted(set(b["pair"])) == sorted(set(a["pair"])) and len(b) == len(set(a["pair"]))
            for c in b:
                assert c["distance"] <= a[a["pair"] == c["pair"]]["distance"].min() + 1e-9  # CLOSEST_TIE band


@pytest.mark.gpu
@pytest.mark.parametrize("cast", [False, True])
@pytest.mark.parametrize("ctype", [capi.CONTACT_ALL, capi.CONTACT_CLOSEST, capi.CONTACT_FIRST])
def test_contact_test_gpu_vs_oracle(cuda, oracle, cast, ctype):
    """tc_contact_test against the oracle's manager on the same poses: same contact sets, 1e-6 m, normals 1e-5"""
    s, sigma = _scene_for(ctype)
    sg, so = cuda.scene(s, 0.05, 20.0, 0.02), oracle.scene(s, 0.05, 20.0, 0.02)
    q, poses = _arm_states(oracle, so, s, 64, sigma=sigma)
    args = (poses[:-1], poses[1:]) if cast else (poses,)
    rg = cuda.contact_test(sg, *args, contact_type=ctype)
    ro = oracle.contact_test(so, *args, contact_type=ctype)
    rep = helpers.compare(ro, rg, want_rows=False)
    assert rep["n_contacts"] > 30 and rep["unmatched_outside_band"] == 0 and rep["same_order"], rep
    assert rep["max_dist_err"] <= 1e-6 and rep["max_normal_err"] <= 1e-5 and rep["max_cc_time_err"] <= 1e-5, rep
    assert rep["cc_type_mismatch"] == 0 and rep["max_transform_err"] <= 1e-12 and rep["term_ptr_equal"], rep
    with pytest.raises(capi.TcError):
        rg.rows()


@pytest.mark.gpu
@pytest.mark.parametrize("evaluator", [capi.EVAL_DISCRETE, capi.EVAL_LVS_DISCRETE, capi.EVAL_LVS_CONTINUOUS])
@pytest.mark.parametrize("ctype", [capi.CONTACT_CLOSEST, capi.CONTACT_FIRST])
def test_contact_test_type_through_the_evaluators(cuda, oracle, evaluator, ctype):
    """contact_request.type CLOSEST / FIRST through tc_eval: the manager's choice comes before the evaluator's filters
    (fixed first / last timestep here), rows and cost follow the reduced contact set"""
    s, sigma = _scene_for(ctype)
    sg, so = cuda.scene(s, 0.05, 20.0, 0.02), oracle.scene(s, 0.05, 20.0, 0.02)
    if "seed_traj" in s:
        q = ms.make_batch(s, 40, 6, t_steps=6)
    else:
        q = np.stack([_arm_states(oracle, so, s, 6, seed=40 + b, sigma=0.05)[0] for b in range(3)])
    fixed = np.zeros(q.shape[1], np.uint8)
    fixed[0] = fixed[-1] = 1
    rg = cuda.eval(sg, q, evaluator, lvs=0.05, fixed=fixed, contact_type=ctype)
    ro = oracle.eval(so, q, evaluator, lvs=0.05, fixed=fixed, contact_type=ctype, n_threads=8)
    n_all = oracle.eval(so, q, evaluator, lvs=0.05, fixed=fixed, n_threads=8).counts()["n_contacts"]
    rep = helpers.compare(ro, rg, scene=s)
    assert 0 < rep["n_contacts"] < n_all, (rep, n_all)
    assert rep["unmatched_outside_band"] == 0 and rep["same_order"] and rep["term_ptr_equal"], rep
    assert rep["cc_outside_band"] == 0, rep
    assert rep["max_dist_err"] <= 1e-6 and rep["max_normal_err"] <= 1e-5, rep
    assert rep["max_grad_rel_err"] <= 1e-5 and rep["max_row_rel_err"] <= 1e-5 and rep["max_cost_err"] <= 1e-6 * rep["n_contacts"], rep
