"""Host-side mirrors of the callers right above the evaluators, batched for the GPU.

  * numerical-gradient collision constraints
      trajopt_ifopt/src/constraints/collision/discrete_collision_numerical_constraint.cpp:65-200
      trajopt_ifopt/src/constraints/collision/continuous_collision_numerical_constraint.cpp:91-232
    The reference calls its evaluator once at x and once more per perturbed joint (delta = 1e-8, forward
    differences) and matches the perturbed gradient-result sets to the baseline's by (link pair, shape key).  Here
    the n_dof + 1 (discrete) or 2 n_dof + 1 (continuous) states go through ONE batched evaluation as independent
    "trajectories"; the matching and the difference quotients follow the reference line by line.
  * post-solve trajectory validation: the `checkTrajectory` call of the reference's outcome
    tests (trajopt/test/planning_unit.cpp:99-100,145-148) with a CONTINUOUS / LVS_* contact request, as one batched
    pass over many trajectories: is any contact closer than the given distance.

Everything here is written against the `eval` / `Result` interface of capi.Lib, so the same code drives
libtrajopt_cuda.so (the product) and, in the tests only, the CPU oracle.  No collision arithmetic happens here.
"""
import numpy as np

from . import capi

DELTA = 1e-8  # discrete_collision_numerical_constraint.cpp:158, continuous_collision_numerical_constraint.cpp:177


def _max_error(s, which):
    """GradientResultsSet::getMaxErrorT0 / T1 / getMaxError (trajopt_common/src/collision_types.cpp:203-263):
    which = 0: slot "excluding T1" of both links, 1: slot "excluding T0", 2: both slots.  None if no slot is set."""
    best = None
    for i in range(2):
        for k in range(2):
            if which != 2 and which != k:
                continue
            if s["has_error"][i][k]:
                v = float(s["error"][i][k])
                best = v if best is None else max(best, v)
    return best


def _term_sets(st, term):
    lo, hi = int(st["set_ptr"][term]), int(st["set_ptr"][term + 1])
    return st["sets"][lo:hi]


def _find(sets, base):
    """std::find_if on (key, shape_key) (discrete_collision_numerical_constraint.cpp:165-171)"""
    for s in sets:
        if s["pair"] == base["pair"] and tuple(s["shape_key"]) == tuple(base["shape_key"]):
            return s
    return None


class DiscreteCollisionNumericalConstraint:
    """One timestep, rows = max_num_cnt, Jacobian by forward differences."""

    def __init__(self, lib, scene, max_num_cnt=1):
        if max_num_cnt < 1:
            raise RuntimeError("max_num_cnt must be greater than zero!")  # ctor check of the reference (:55-56)
        self.lib, self.scene, self.max_num_cnt = lib, scene, int(max_num_cnt)
        self.passes = 0

    def evaluate(self, joint_vals):
        """-> values[max_num_cnt], coeffs[max_num_cnt], jac[max_num_cnt, n_dof] from ONE batched pass"""
        q = np.asarray(joint_vals, dtype=np.float64)
        n = self.scene.n_dof
        assert q.shape == (n,)
        states = np.repeat(q[None], n + 1, axis=0)
        for j in range(n):
            states[1 + j, j] = q[j] + DELTA
        r = self.lib.eval(self.scene, states[:, None, :], capi.EVAL_DISCRETE,
                          want=capi.WANT_CONTACTS | capi.WANT_GRADIENTS)
        self.passes += 1
        st = r.sets(self.max_num_cnt)
        mb = self.scene.margin_buffer
        values = np.full(self.max_num_cnt, -mb)
        coeffs = np.ones(self.max_num_cnt)
        jac = np.zeros((self.max_num_cnt, n))
        base = _term_sets(st, 0)
        cnt = min(self.max_num_cnt, len(base))
        for i in range(cnt):
            values[i] = _max_error(base[i], 0)  # getMaxErrorT0 (:128)
            coeffs[i] = base[i]["coeff"]
        for j in range(n):
            pert = _term_sets(st, 1 + j)
            for i in range(cnt):
                b0 = _max_error(base[i], 0)
                hit = _find(pert, base[i])
                e = _max_error(hit, 0) if hit is not None else -mb  # :173-182
                jac[i, j] = (e - b0) / DELTA
        return values, coeffs, jac


class ContinuousCollisionNumericalConstraint:
    """One segment (two states), rows = max_num_cnt, Jacobian blocks for both states by forward differences."""

    def __init__(self, lib, scene, evaluator, lvs, vars0_fixed=False, vars1_fixed=False, max_num_cnt=1):
        if vars0_fixed and vars1_fixed:
            raise RuntimeError("position_vars are both fixed!")  # :66-67
        if max_num_cnt < 1:
            raise RuntimeError("max_num_cnt must be greater than zero!")  # :69-70
        self.lib, self.scene, self.evaluator, self.lvs = lib, scene, evaluator, lvs
        self.f0, self.f1, self.max_num_cnt = bool(vars0_fixed), bool(vars1_fixed), int(max_num_cnt)
        self.passes = 0

    def _which(self):
        return 2 if (not self.f0 and not self.f1) else (0 if not self.f0 else 1)

    def evaluate(self, joint_vals0, joint_vals1):
        """-> values[max_num_cnt], coeffs[max_num_cnt], jac[max_num_cnt, 2, n_dof] (block p = state p)"""
        q0 = np.asarray(joint_vals0, dtype=np.float64)
        q1 = np.asarray(joint_vals1, dtype=np.float64)
        n = self.scene.n_dof
        pairs = np.empty((2 * n + 1, 2, n))
        pairs[0, 0], pairs[0, 1] = q0, q1
        for p, qp in enumerate((q0, q1)):
            for j in range(n):
                jv = qp.copy()
                jv[j] += DELTA
                # the reference passes the perturbed vector of block p as dof_vals0 and the unperturbed state 1 as
                # dof_vals1 for BOTH blocks (continuous_collision_numerical_constraint.cpp:182-183); kept as is
                pairs[1 + p * n + j, 0], pairs[1 + p * n + j, 1] = jv, q1
        fixed = np.array([self.f0, self.f1], dtype=np.uint8)
        r = self.lib.eval(self.scene, pairs, self.evaluator, lvs=self.lvs, fixed=fixed,
                          want=capi.WANT_CONTACTS | capi.WANT_GRADIENTS)
        self.passes += 1
        st = r.sets(self.max_num_cnt)
        mb = self.scene.margin_buffer
        which = self._which()
        values = np.full(self.max_num_cnt, -mb)
        coeffs = np.ones(self.max_num_cnt)
        jac = np.zeros((self.max_num_cnt, 2, n))
        base = _term_sets(st, 0)
        cnt = min(self.max_num_cnt, len(base))
        for i in range(cnt):
            values[i] = _max_error(base[i], which)
            coeffs[i] = base[i]["coeff"]
        for p in range(2):
            for j in range(n):
                pert = _term_sets(st, 1 + p * n + j)
                for i in range(cnt):
                    b0 = _max_error(base[i], which)
                    hit = _find(pert, base[i])
                    e = _max_error(hit, which) if hit is not None else -mb
                    jac[i, p, j] = (e - b0) / DELTA
        return values, coeffs, jac


def check_trajectories(lib, scene, trajs, evaluator=capi.EVAL_LVS_CONTINUOUS, lvs=0.02, contact_distance=0.0,
                       result=None):
    """Batched trajectory validation: trajs[B, T, n_dof] -> dict(in_collision[B] bool, min_distance[B] (inf when no
    contact is within margin + buffer), n_contacts[B]).  A trajectory is "in collision" when some contact of some
    state / swept sub-segment has distance < contact_distance — what the reference's tests assert on the result of
    checkTrajectory (trajopt/test/planning_unit.cpp:99-100,145-148).  All steps count (no fixed-step filtering).
    The scene's margin + buffer is the detection range, so contact_distance must not exceed it."""
    q = np.ascontiguousarray(trajs, dtype=np.float64)
    B = q.shape[0]
    r = lib.eval(scene, q, evaluator, lvs=lvs, fixed=None, want=capi.WANT_CONTACTS, result=result) \
        if result is not None else lib.eval(scene, q, evaluator, lvs=lvs, fixed=None, want=capi.WANT_CONTACTS)
    c = r.contacts()
    min_d = np.full(B, np.inf)
    n_c = np.zeros(B, dtype=np.int64)
    if len(c):
        np.minimum.at(min_d, c["traj"], c["distance"])
        np.add.at(n_c, c["traj"], 1)
    return dict(in_collision=min_d < contact_distance, min_distance=min_d, n_contacts=n_c)
