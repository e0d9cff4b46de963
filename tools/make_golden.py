"""Freeze ORACLE outputs for the reference's fixtures as golden vectors (tests/golden/*.npz).

These are oracle-generated, NOT reference-generated: tesseract/Bullet cannot run in this environment
(DESIGN.md 2).  What they pin is (a) the oracle against itself over time and (b) the CUDA path against a
committed file rather than against a checker built on the GPU box.  The reference's own numeric vectors
(boxbot KATs) are asserted directly in tests/test_oracle_golden.py.

usage: python tools/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from tests import helpers  # noqa: E402

capi = helpers.capi
CASES = [
    # name, scene, margin, coeff, buffer, evaluator, stack, lvs, trajectory
    ("arm_seed_lvs_continuous", "arm_around_table", 0.025, 20.0, 0.01, capi.EVAL_LVS_CONTINUOUS, capi.STACK_IFOPT, 0.02, "seed"),
    ("arm_seed_lvs_discrete", "arm_around_table", 0.025, 20.0, 0.01, capi.EVAL_LVS_DISCRETE, capi.STACK_IFOPT, 0.02, "seed"),
    ("arm_seed_cast_legacy", "arm_around_table", 0.025, 20.0, 0.01, capi.EVAL_LVS_CONTINUOUS, capi.STACK_LEGACY, 0.02, "seed"),
    ("arm_seed_discrete", "arm_around_table", 0.025, 20.0, 0.01, capi.EVAL_DISCRETE, capi.STACK_IFOPT, 0.02, "seed"),
    ("boxbot_cast", "boxbot", 0.2, 10.0, 0.05, capi.EVAL_LVS_CONTINUOUS, capi.STACK_IFOPT, 0.05, "box_cast"),
]


def trajectory(scene, kind):
    if kind == "seed":
        return scene["seed_traj"][None]
    # trajopt_common/data/config/box_cast_test.json:41
    return np.array([[[-1.9, 0.0], [0.0, 1.9], [1.9, 3.8]]])


def main():
    orc = pyoracle.OracleLib()
    for name, scene_name, margin, coeff, buf, ev, stack, lvs, kind in CASES:
        s = helpers.load_scene(scene_name)
        so = orc.scene(s, margin, coeff, buf)
        q = trajectory(s, kind)
        fixed = np.zeros(q.shape[1], np.uint8)
        fixed[0] = fixed[-1] = 1
        r = orc.eval(so, q, ev, stack=stack, lvs=lvs, fixed=fixed)
        c = r.contacts()
        rec, grad = r.gradients()
        rows = r.rows()
        sets = r.sets(3)
        cost_term, cost_traj = r.cost()
        counts = r.counts()
        out = dict(q=q, fixed=fixed, params=np.array([margin, coeff, buf, lvs]), evaluator=np.int32(ev),
                   stack=np.int32(stack), contacts=c, transforms=r.contact_transforms(), term_ptr=r.term_ptr(),
                   grad_rec=rec, grad=grad, row_ptr=rows["row_ptr"], col_idx=rows["col_idx"], val=rows["val"],
                   row_value=rows["value"], row_coeff=rows["coeff"], row_margin=rows["margin"],
                   set_values=sets["values"], set_coeffs=sets["coeffs"], set_jac=sets["jac"], set_ptr=sets["set_ptr"],
                   cost_term=cost_term, cost_traj=cost_traj,
                   counts=np.array([counts[k] for k in ("n_states", "n_units", "n_dense", "n_candidates", "n_terms")]))
        path = os.path.join(ROOT, "tests", "golden", name + ".npz")
        np.savez_compressed(path, **out)
        print(name, "contacts", len(c), "rows", len(rows["value"]), "bytes", os.path.getsize(path))


if __name__ == "__main__":
    main()
