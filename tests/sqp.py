"""Test harness that closes the optimisation loop around the collision terms: a small
restatement of trajopt's BasicTrustRegionSQP (trajopt_sco/src/optimizers.cpp:560-960, defaults of
trajopt_sco/include/trajopt_sco/optimizers.hpp:92-125) for problems whose terms are a joint-velocity cost and a
hinge collision cost (trajopt/src/collision_terms.cpp:1267-1306), with the QP solved by an in-repo ADMM solver
that follows the published OSQP algorithm (OSQP / OsqpEigen are not installed here).

The evaluator is a parameter: anything with the capi `eval(...) -> Result` shape, i.e. the CPU oracle or the CUDA
library.  Driving the same loop with both shows the drop-in property end to end: identical converged trajectories
within 1e-4 rad (BASELINE.json north_star)."""
import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from tests import helpers

capi = helpers.capi


def solve_qp(P, q, A, l, u, eps=1e-9, max_iter=200000):
    """min 0.5 x'Px + q'x  s.t.  l <= Ax <= u   (OSQP's ADMM iteration with adaptive rho)"""
    n, m = P.shape[0], A.shape[0]
    sigma, alpha, rho = 1e-6, 1.6, 0.1
    eq = l == u

    def factor(rho):
        rho_vec = np.where(eq, 1e3 * rho, rho)
        K = sp.bmat([[P + sigma * sp.identity(n), A.T], [A, -sp.diags(1.0 / rho_vec)]], format="csc")
        return rho_vec, spla.splu(K)

    rho_vec, lu = factor(rho)
    x, z, y = np.zeros(n), np.zeros(m), np.zeros(m)
    for k in range(1, max_iter + 1):
        sol = lu.solve(np.concatenate([sigma * x - q, z - y / rho_vec]))
        xt, nu = sol[:n], sol[n:]
        zt = z + (nu - y) / rho_vec
        x = alpha * xt + (1 - alpha) * x
        zr = alpha * zt + (1 - alpha) * z
        zn = np.clip(zr + y / rho_vec, l, u)
        y = y + rho_vec * (zr - zn)
        z = zn
        if k % 25 == 0:
            Ax = A @ x
            r_prim = np.abs(Ax - z).max()
            r_dual = np.abs(P @ x + q + A.T @ y).max()
            if r_prim < eps and r_dual < eps:
                break
            if k % 200 == 0:
                num = r_prim / max(np.abs(Ax).max(), np.abs(z).max(), 1e-12)
                den = r_dual / max(np.abs(P @ x).max(), np.abs(A.T @ y).max(), np.abs(q).max(), 1e-12)
                ratio = np.sqrt(num / max(den, 1e-30))
                if ratio > 5 or ratio < 0.2:
                    rho = float(np.clip(rho * ratio, 1e-6, 1e6))
                    rho_vec, lu = factor(rho)
    return x, k


class CollisionProblem:
    """trajectory optimisation with a joint-velocity cost and a hinge collision cost; the first and last
    waypoints are fixed (the fixtures pin the start state and constrain the goal)"""

    def __init__(self, lib, scene, q_init, evaluator, margin, coeff, buffer, lvs, vel_coeff=1.0):
        self.lib, self.q0 = lib, np.array(q_init, dtype=np.float64)
        self.T, self.n = self.q0.shape
        self.sc = lib.scene(scene, margin, coeff, buffer)
        self.evaluator, self.lvs, self.vel_coeff = evaluator, lvs, vel_coeff
        self.fixed = np.zeros(self.T, np.uint8)
        self.fixed[0] = self.fixed[-1] = 1
        lim = scene["joint_limits"]
        self.lo, self.hi = np.tile(lim[:, 0], self.T - 2), np.tile(lim[:, 1], self.T - 2)
        self.passes = 0
        # velocity cost as a quadratic in the free variables: sum_t |x_{t+1} - x_t|^2
        nf, n = (self.T - 2) * self.n, self.n
        D = sp.lil_matrix(((self.T - 1) * n, self.T * n))
        for t in range(self.T - 1):
            for j in range(n):
                D[t * n + j, t * n + j] = -1.0
                D[t * n + j, (t + 1) * n + j] = 1.0
        self.D = D.tocsr()
        self.free = np.arange(n, (self.T - 1) * n)
        assert len(self.free) == nf

    def traj(self, xf):
        q = self.q0.copy()
        q[1:-1] = xf.reshape(self.T - 2, self.n)
        return q

    def _eval(self, q, want):
        self.passes += 1
        return self.lib.eval(self.sc, q[None], self.evaluator, stack=capi.STACK_LEGACY, lvs=self.lvs, fixed=self.fixed,
                             want=want)

    def cost(self, xf):
        """CollisionCost::value + JointVelCost::value at x (one geometry pass)"""
        q = self.traj(xf)
        _, traj_cost = self._eval(q, capi.WANT_COST | capi.WANT_CONTACTS).cost()
        v = self.D @ q.reshape(-1)
        return self.vel_coeff * float(v @ v) + float(traj_cost[0])

    def convexify(self, xf):
        """linearisation pass: affine signed distances d_i(x) = c_i + a_i . x for every contact"""
        q = self.traj(xf)
        r = self._eval(q, capi.WANT_ALL)
        rows = r.rows()
        m = len(rows["value"])
        Afull = sp.csr_matrix((rows["val"], rows["col_idx"], rows["row_ptr"]), shape=(m, self.T * self.n))
        const = rows["value"] + Afull[:, np.setdiff1d(np.arange(self.T * self.n), self.free)] @ \
            q.reshape(-1)[np.setdiff1d(np.arange(self.T * self.n), self.free)]
        return Afull[:, self.free].tocsr(), const, rows["margin"], rows["coeff"]

    def solve_model(self, xf, lin, box):
        """QP over [x_free, t]: min vel(x) + sum t_i,  t_i >= 0,  t_i >= coeff_i (margin_i - d_i(x)),  |x - x0| <= box"""
        Ad, const, margin, coeff = lin
        nf, m = len(xf), len(const)
        Df, Dfix = self.D[:, self.free], self.D
        qfull = self.traj(xf).reshape(-1)
        qfix = qfull.copy()
        qfix[self.free] = 0.0
        b = Dfix @ qfix  # velocity = Df x + b
        P = sp.block_diag([2 * self.vel_coeff * (Df.T @ Df), sp.csr_matrix((m, m))], format="csc")
        qv = np.concatenate([2 * self.vel_coeff * (Df.T @ b), np.ones(m)])
        # rows: hinge (coeff a_i . x + t_i >= coeff (margin - c_i)), t >= 0, box on x
        A = sp.vstack([sp.hstack([sp.diags(coeff) @ Ad, sp.identity(m)]),
                       sp.hstack([sp.csr_matrix((m, nf)), sp.identity(m)]),
                       sp.hstack([sp.identity(nf), sp.csr_matrix((nf, m))])], format="csc")
        l = np.concatenate([coeff * (margin - const), np.zeros(m), np.maximum(xf - box, self.lo)])
        u = np.concatenate([np.full(2 * m, np.inf), np.minimum(xf + box, self.hi)])
        z, iters = solve_qp(P, qv, A, l, u)
        x_new = z[:nf]
        v = Df @ x_new + b
        hinge = np.maximum(coeff * (margin - (const + Ad @ x_new)), 0.0)
        return x_new, self.vel_coeff * float(v @ v) + float(hinge.sum()), iters


def optimize(prob, max_iter=50, trust_box=0.1, improve_ratio_threshold=0.25, min_trust_box=1e-4, min_approx_improve=1e-4,
             trust_shrink=0.1, trust_expand=1.5, log=None):
    """BasicTrustRegionSQP::optimize for a cost-only problem (no penalty iterations)"""
    x = prob.x0() if hasattr(prob, "x0") else prob.q0[1:-1].reshape(-1).copy()
    old = prob.cost(x)
    status = "iteration limit"
    for it in range(max_iter):
        lin = prob.convexify(x)
        accepted = False
        while trust_box >= min_trust_box:
            x_new, model, qp_iters = prob.solve_model(x, lin, trust_box)
            approx = old - model
            if approx < -1e-5:
                raise RuntimeError("model merit increased: %g" % approx)
            if approx < min_approx_improve:
                status = "converged: model improvement below threshold"
                break
            new = prob.cost(x_new)
            exact = old - new
            if log is not None:
                log.append(dict(iter=it, box=trust_box, old=old, model=model, new=new, qp_iters=qp_iters, rows=len(lin[1])))
            if exact < 0 or exact / approx < improve_ratio_threshold:
                trust_box *= trust_shrink
            else:
                x, old, accepted = x_new, new, True
                trust_box *= trust_expand
                break
        if status.startswith("converged"):
            break
        if not accepted:
            status = "converged: trust region below minimum"
            break
    return prob.traj(x), old, status


class SingleStateProblem:
    """trajopt_common/data/config/simple_collision_test.json (trajopt/test/simple_collision_unit.cpp:62-124): ONE
    timestep of spherebot, a joint-position cost towards `target`, a DISCRETE collision cost (dist_pen 0.3, coeff 1) and a
    DISCRETE collision constraint (dist_pen 0.2, coeff 1).  BasicTrustRegionSQP turns the constraint into an l1 penalty
    with merit coefficient mu (merit_error_coeff, 10 at the start: trajopt_sco/include/trajopt_sco/optimizers.hpp:92-125);
    here mu is fixed, which is exact once it exceeds the multiplier.  Same interface as CollisionProblem."""

    def __init__(self, lib, scene, q_init, target=(0.0, 0.0), mu=10.0, buffer=0.05):
        self.lib, self.q0 = lib, np.array(q_init, dtype=np.float64).reshape(1, -1)
        self.n = self.q0.shape[1]
        self.target = np.asarray(target, dtype=np.float64)
        self.terms = [(lib.scene(scene, 0.3, 1.0, buffer), 1.0), (lib.scene(scene, 0.2, 1.0, buffer), mu)]
        lim = scene["joint_limits"]
        self.lo, self.hi = lim[:, 0].copy(), lim[:, 1].copy()
        self.passes = 0

    def x0(self):
        return self.q0.reshape(-1).copy()

    def traj(self, x):
        return x.reshape(1, self.n).copy()

    def _eval(self, sc, x, want):
        self.passes += 1
        return self.lib.eval(sc, x.reshape(1, 1, self.n), capi.EVAL_DISCRETE, stack=capi.STACK_LEGACY, want=want)

    def cost(self, x):
        c = float(((x - self.target) ** 2).sum())
        for sc, w in self.terms:
            c += w * float(self._eval(sc, x, capi.WANT_COST | capi.WANT_CONTACTS).cost()[1][0])
        return c

    def convexify(self, x):
        A, const, margin, coeff = [], [], [], []
        for sc, w in self.terms:
            rows = self._eval(sc, x, capi.WANT_ALL).rows()
            m = len(rows["value"])
            A.append(sp.csr_matrix((rows["val"], rows["col_idx"], rows["row_ptr"]), shape=(m, self.n)))
            const.append(rows["value"]), margin.append(rows["margin"]), coeff.append(w * rows["coeff"])
        return sp.vstack(A).tocsr(), np.concatenate(const), np.concatenate(margin), np.concatenate(coeff)

    def solve_model(self, x, lin, box):
        Ad, const, margin, coeff = lin
        n, m = self.n, len(const)
        P = sp.block_diag([2 * sp.identity(n), sp.csr_matrix((m, m))], format="csc")
        qv = np.concatenate([-2 * self.target, np.ones(m)])
        A = sp.vstack([sp.hstack([sp.diags(coeff) @ Ad, sp.identity(m)]),
                       sp.hstack([sp.csr_matrix((m, n)), sp.identity(m)]),
                       sp.hstack([sp.identity(n), sp.csr_matrix((n, m))])], format="csc")
        l = np.concatenate([coeff * (margin - const), np.zeros(m), np.maximum(x - box, self.lo)])
        u = np.concatenate([np.full(2 * m, np.inf), np.minimum(x + box, self.hi)])
        z, iters = solve_qp(P, qv, A, l, u)
        x_new = z[:n]
        hinge = np.maximum(coeff * (margin - (const + Ad @ x_new)), 0.0)
        return x_new, float(((x_new - self.target) ** 2).sum()) + float(hinge.sum()), iters
