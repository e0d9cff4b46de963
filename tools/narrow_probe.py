#!/usr/bin/env python3
"""Prints the stage times of one C4 linearisation pass (tuning probe).  usage: narrow_probe.py [seeds] [reps]; env
TC_LIB_PATH selects a build of tools/build_variant.sh, TC_PROBE_PREC=f32 / TC_PROBE_EVAL=discrete the mode, TAG a label."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tcb200  # noqa: E402

pkg = tcb200.load()
ms_, capi = pkg.multistart, pkg.capi
import torch  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
scene = ms_.load_scene("arm_around_table")
cu = pkg.CudaLib()
sg = cu.scene(scene, ms_.MARGIN, ms_.COEFF, ms_.BUFFER)
q = torch.from_numpy(ms_.make_batch(scene, 0, B)).cuda()
res = capi.Result(cu, sg, cu.new_result(sg), B, ms_.T_STEPS)
st = torch.cuda.current_stream().cuda_stream
prec = capi.PRECISION_F32 if os.environ.get("TC_PROBE_PREC") == "f32" else capi.PRECISION_F64
ev = capi.EVAL_LVS_DISCRETE if os.environ.get("TC_PROBE_EVAL") == "discrete" else capi.EVAL_LVS_CONTINUOUS
acc, wall = [], []
import time
for i in range(3 + reps):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    cu.eval_device(sg, q.data_ptr(), B, ms_.T_STEPS, ev, res, lvs=ms_.LVS, fixed=ms_.fixed_steps(),
                   stream=st, want=capi.WANT_LINEARISE, precision=prec)
    cu.sync(res)
    if i >= 3:
        wall.append(1e3 * (time.perf_counter() - t0))
        acc.append(cu.stage_ms(res))
m = np.mean(np.array(acc), axis=0)
c = res.counts()
print("tag=%s gjk=%.3f epa=%.3f narrow=%.3f fk=%.3f broad=%.3f deep=%.3f tail=%.3f total=%.3f wall=%.3f contacts=%d" % (
    os.environ.get("TAG", ""), m[8], m[9], m[5], m[0], m[6], m[2], m[3], m[4], float(np.mean(wall)), c["n_contacts"]))
