#!/usr/bin/env python3
"""Parity sweep beyond the test suite's seeds (GPU box; test infrastructure: imports the CPU checker).  For every
evaluator and both stacks, N further random seeds of the C4 workload are compared contact by contact with the CPU
checker at the north_star tolerances; prints one line per case and exits non-zero on the first violation.
usage: python tools/parity_sweep.py [seeds_per_case] [first_seed]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import helpers  # noqa: E402
from oracle import pyoracle  # noqa: E402

capi = helpers.capi
ms = helpers.pkg.multistart
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
first = int(sys.argv[2]) if len(sys.argv) > 2 else 10000
cuda, oracle = helpers.pkg.CudaLib(), pyoracle.OracleLib()
s = ms.load_scene()
sg, so = cuda.scene(s, ms.MARGIN, ms.COEFF, ms.BUFFER), oracle.scene(s, ms.MARGIN, ms.COEFF, ms.BUFFER)
names = {capi.EVAL_DISCRETE: "DISCRETE", capi.EVAL_LVS_DISCRETE: "LVS_DISCRETE", capi.EVAL_CONTINUOUS: "CONTINUOUS",
         capi.EVAL_LVS_CONTINUOUS: "LVS_CONTINUOUS"}
bad = 0
case = 0
for ev, name in names.items():
    for stack, sname in ((capi.STACK_IFOPT, "ifopt"), (capi.STACK_LEGACY, "legacy")):
        q = ms.make_batch(s, first + case * n, n)
        case += 1
        rg = cuda.eval(sg, q, ev, stack=stack, lvs=ms.LVS, fixed=ms.fixed_steps())
        ro = oracle.eval(so, q, ev, stack=stack, lvs=ms.LVS, fixed=ms.fixed_steps(), n_threads=oracle.num_threads())
        rep = helpers.compare_fast(ro, rg)
        cc_out = helpers.cc_type_outside_band(ro, rg, s)[0] if ev in (capi.EVAL_CONTINUOUS, capi.EVAL_LVS_CONTINUOUS) else 0
        co, cg = ro.counts(), rg.counts()
        counts_ok = all(co[k] == cg[k] for k in ("n_states", "n_units", "n_dense", "n_candidates", "n_terms", "n_narrowphase"))
        ok = (rep["unmatched_outside_band"] == 0 and cc_out == 0 and rep["max_dist_err"] <= 1e-6 and
              rep["max_normal_err"] <= 1e-5 and rep["max_grad_rel_err"] <= 1e-5 and rep["max_cost_rel_err"] <= 1e-6 and
              counts_ok and cg["n_unconverged"] == 0 and cg["overflow"] == 0)
        if rep.get("same_order"):
            ok = ok and rep["max_row_rel_err"] <= 1e-5 and rep["max_row_value_err"] <= 1e-6
        print(f"{name:15s} {sname:6s} seeds {n} contacts {rep['n_contacts']:8d} unmatched_outside_band {rep['unmatched_outside_band']} "
              f"cc_type_outside_band {cc_out} dist {rep['max_dist_err']:.2e} normal {rep['max_normal_err']:.2e} "
              f"grad {rep['max_grad_rel_err']:.2e} row {rep.get('max_row_rel_err', float('nan')):.2e} cost {rep['max_cost_rel_err']:.2e} "
              f"same_order {rep.get('same_order')} counts_equal {counts_ok} {'OK' if ok else 'VIOLATION'}", flush=True)
        bad += 0 if ok else 1
sys.exit(1 if bad else 0)
