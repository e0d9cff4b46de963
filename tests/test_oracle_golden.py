"""CPU suite: pins the oracle against the reference's own known-answer vectors and closed forms."""
import numpy as np
import pytest

from tests import helpers

capi = helpers.capi


def test_boxbot_known_answers(oracle):
    """trajopt_ifopt/test/collision_unit.cpp:97-159 (the only numeric KAT on this path): margin 0.1, coeff 1,
    buffer 0.01, DiscreteCollisionConstraint(max_num_cnt=1); all EXPECT_NEAR(..., 1e-6)."""
    s = helpers.load_scene("boxbot")
    so = oracle.scene(s, 0.1, 1.0, 0.01)
    cases = [((-1.9, 0.01), -0.01, (0, 0)), ((-1.0, 0.01), 0.1, (1, 0)), ((0.01, 1.0), 0.1, (0, -1)),
             ((-0.9, 0.01), 0.2, (1, 0))]
    for q, val, jac in cases:
        r = oracle.eval(so, np.array(q).reshape(1, 1, 2), capi.EVAL_DISCRETE)
        st = r.sets(1)
        assert abs(st["values"][0, 0] - val) <= 1e-6, (q, st["values"])
        assert np.abs(st["jac"][0, 0, 0] - np.array(jac)).max() <= 1e-6, (q, st["jac"])


def test_spherebot_closed_form(oracle):
    """spherebot.urdf: r=0.5 sphere at (x, y) vs r=0.5 spheres at (0,0), (-0.75,0), (0,0.75):
    d = |ca - cb| - 1, n = (cb - ca)/|.|, rows = -+n_xy (SURVEY 8c golden vector 2)."""
    s = helpers.load_scene("spherebot")
    so = oracle.scene(s, 0.2, 1.0, 0.05)
    q = np.array([-0.75, 0.75])
    r = oracle.eval(so, q.reshape(1, 1, 2), capi.EVAL_DISCRETE)
    c = r.contacts()
    centers = {1: (0.0, 0.0), 2: (-0.75, 0.0), 3: (0.0, 0.75)}
    assert sorted(c["link_ids"][:, 1].tolist()) == [1, 2, 3]
    rec, g = r.gradients()
    for i, ct in enumerate(c):
        cb = np.array(centers[int(ct["link_ids"][1])])
        d = np.linalg.norm(q - cb) - 1.0
        n = (cb - q) / np.linalg.norm(cb - q)
        assert abs(ct["distance"] - d) <= 1e-9
        assert np.abs(ct["normal"][:2] - n).max() <= 1e-9
        assert np.abs(g[i, 0, 0] + n).max() <= 1e-9  # link A gets -n^T J, J = I for the two prismatic joints


def test_swept_sphere_closed_form(oracle):
    """sphere swept along a segment = capsule: distance to a static sphere is the point-segment distance"""
    s = helpers.load_scene("spherebot")
    so = oracle.scene(s, 0.5, 1.0, 0.5)
    q = np.array([[[-1.5, 1.2], [1.5, 1.2]]])
    r = oracle.eval(so, q, capi.EVAL_CONTINUOUS)
    c = r.contacts()
    assert len(c) == 3
    for ct in c:
        cb = {1: (0.0, 0.0), 2: (-0.75, 0.0), 3: (0.0, 0.75)}[int(ct["link_ids"][1])]
        d = abs(1.2 - cb[1]) - 1.0
        assert abs(ct["distance"] - d) <= 1e-9
        assert ct["cc_type"][0] == capi.CC_BETWEEN and ct["cc_type"][1] == capi.CC_NONE
        assert abs(ct["cc_time"][0] - (cb[0] + 1.5) / 3.0) <= 1e-6


@pytest.mark.parametrize("evaluator", [capi.EVAL_DISCRETE, capi.EVAL_LVS_DISCRETE])
def test_gradient_rows_finite_difference(oracle, evaluator):
    """executable form of the reference's Debug check (collision_terms.cpp:295-309): analytic rows vs central
    differences of the signed distance of the same contact"""
    s = helpers.load_scene("arm_around_table")
    so = oracle.scene(s, 0.025, 20.0, 0.01)
    q = s["seed_traj"][2:4].copy()[None]
    if evaluator == capi.EVAL_DISCRETE:
        q = q[:, :1]
    r = oracle.eval(so, q, evaluator, lvs=0.25)
    c = r.contacts()
    rows = r.rows()
    keys = helpers.contact_keys(c)
    # separated (GJK) contacts first, then penetrating ones: the signed distance is smooth in both regimes
    pick = ([i for i in range(len(c)) if not (c["flags"][i] & capi.FLAG_EPA)] + list(range(len(c))))[:4]
    assert pick
    h = 1e-6
    for i in pick:
        val = rows["val"][rows["row_ptr"][i]:rows["row_ptr"][i + 1]]
        cols = rows["col_idx"][rows["row_ptr"][i]:rows["row_ptr"][i + 1]]
        num = np.zeros(len(val))
        for k, col in enumerate(cols):
            d = []
            for sgn in (1, -1):
                qq = q.copy().reshape(-1)
                qq[col] += sgn * h
                rr = oracle.eval(so, qq.reshape(q.shape), evaluator, lvs=0.25, want=capi.WANT_CONTACTS)
                cc = rr.contacts()
                kk = helpers.contact_keys(cc)
                d.append(cc["distance"][kk.index(keys[i])])
            num[k] = -(d[0] - d[1]) / (2 * h)  # row = -d(distance)/dq
        # (the swept-shape rows of the cast evaluators are trajopt's cc_time-weighted approximation, not an
        # exact derivative, so they are not checked this way)
        tol = 2e-3
        assert np.abs(num - val).max() <= tol * max(1.0, np.abs(val).max()), (i, num, val)
