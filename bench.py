#!/usr/bin/env python3
"""bench.py — collision-term evaluations/s on the batched arm_around_table multi-start workload (config C4).

One "step" = one linearisation pass (FK -> broadphase -> GJK/EPA -> contacts -> gradient rows -> CSR rows
-> cost) over `--seeds` trajectories per GPU.  Units = N_dense = sum_b S(b) * P (swept sub-segments after
LVS subdivision x candidate link pairs), the count SURVEY.md 8(d) defines for CPU and GPU alike.

  python bench.py [--gpus N --steps K --warmup W]           our arm (CUDA path through the C ABI)
  python bench.py --impl reference ...                      CPU restatement (oracle) on the host cores
N > 1 is launched by torchrun, one rank per GPU; trajectories are sharded, no data-path collective;
one NCCL all-gather of (best cost, index) per step picks the multi-start winner.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "collision evals/sec (timestep x link-pair)"
UNIT = "evals/s"
import tcb200  # noqa: E402

pkg = tcb200.load()
ms_ = pkg.multistart
T_STEPS, LVS, MARGIN, COEFF, BUFFER = ms_.T_STEPS, ms_.LVS, ms_.MARGIN, ms_.COEFF, ms_.BUFFER
make_batch, fixed_steps = ms_.make_batch, ms_.fixed_steps


class ClockSampler(threading.Thread):
    """samples nvidia-smi clocks / throttle reasons during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                self.rows.append([x.strip() for x in out.stdout.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        rows = [r for r in self.rows if len(r) == 6 and r[0].isdigit()]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": float(np.median([int(r[0]) for r in rows])), "sm_max_mhz": float(rows[0][1]),
                "reasons": reasons, "samples": len(rows)}


def cpu_baseline(scene, seconds=12.0, n_seeds=None):
    """oracle (CPU restatement of the reference's evaluators, one evaluator per term, terms spread over all host
    threads like BasicTrustRegionSQPMultiThreaded) on a bounded sample of the same workload; also on ONE thread, so that
    the port's own thread scaling is on the record"""
    from oracle import pyoracle
    capi = pyoracle.capi
    orc = pyoracle.OracleLib()
    cores = orc.num_threads()
    so = orc.scene(scene, MARGIN, COEFF, BUFFER)
    n = n_seeds or max(8, 2 * cores)
    q = make_batch(scene, 0, n)

    def best_of(threads, budget, qq):
        best, dense, t_end = None, 0, time.time() + budget
        while True:
            t0 = time.perf_counter()
            r = orc.eval(so, qq, capi.EVAL_LVS_CONTINUOUS, lvs=LVS, fixed=fixed_steps(), n_threads=threads)
            r.rows()
            dt = time.perf_counter() - t0
            dense = r.counts()["n_dense"]
            best = dt if best is None else min(best, dt)
            if time.time() > t_end:
                return best, dense
    best, dense = best_of(cores, seconds, q)
    b1, d1 = best_of(1, min(6.0, seconds), q[:4])
    return {"value": dense / best, "unit": UNIT, "cores": cores, "kind": "port", "value_1thread": d1 / b1,
            "thread_scaling": (dense / best) / (d1 / b1),
            "sample": f"{n} seeds of the C4 workload (seeds 4000..{4000 + n - 1}), linearisation pass incl. rows, "
                      f"best wall time in {seconds:.0f} s, std::thread over terms; value_1thread: 4 of those seeds on "
                      "one thread"}, best, dense


def sqp_iteration_latency(cu, with_cpu):
    """BASELINE.json's second metric, ms per SQP iteration = one linearisation pass at x + one geometry (value-only)
    pass at the trial x, for ONE problem of the single-problem configs — the latency-bound regime a
    10-100-timestep problem lives in.  Host buffers, wall clock around the plain tc_eval calls (B = 1), median of 20.
    With `with_cpu` the same two passes through the CPU port (single thread, as one SQP thread would run them)."""
    capi = pkg.capi
    from tests import synthetic  # scene generators shared with the parity tests (no oracle involved)
    arm = ms_.load_scene("arm_around_table")
    c3, c5 = synthetic.cluttered_scene(), synthetic.dual_arm_scene()
    cases = [
        ("C2 arm_around_table T=10 LVS_CONTINUOUS", arm, make_batch(arm, 0, 1)[0], dict(margin=MARGIN, coeff=COEFF, buffer=BUFFER, lvs=LVS)),
        ("C3 64-hull clutter T=50 LVS_CONTINUOUS", c3, c3["traj"], dict(margin=0.025, coeff=20.0, buffer=0.05, lvs=0.02)),
        ("C5 dual arm 16 dof T=100 LVS_CONTINUOUS", c5, c5["traj"], dict(margin=0.025, coeff=20.0, buffer=0.01, lvs=0.05)),
    ]
    orc = None
    if with_cpu:
        from oracle import pyoracle
        orc = pyoracle.OracleLib()
    out = {}
    for name, sc, traj, prm in cases:
        q = np.ascontiguousarray(traj[None], dtype=np.float64)
        fixed = np.zeros(q.shape[1], np.uint8)
        fixed[0] = fixed[-1] = 1
        sg = cu.scene(sc, prm["margin"], prm["coeff"], prm["buffer"])
        # one result per kind of pass, as an SQP driver would hold them: each keeps seeing passes of ONE shape, so all
        # but the first are sized speculatively (no read-back inside the pass)
        res = (capi.Result(cu, sg, cu.new_result(sg), 1, q.shape[1]), capi.Result(cu, sg, cu.new_result(sg), 1, q.shape[1]))

        def one(lib, scene, result=None):
            kw = dict(result=result[0]) if result is not None else {}
            kw2 = dict(result=result[1]) if result is not None else {}
            t0 = time.perf_counter()
            # the CPU port assembles its rows from stored gradients; the device folds them into the rows on the fly
            lin = capi.WANT_LINEARISE if result is not None else capi.WANT_ALL
            r = lib.eval(scene, q, capi.EVAL_LVS_CONTINUOUS, lvs=prm["lvs"], fixed=fixed, want=lin, **kw)
            r.rows()
            r2 = lib.eval(scene, q, capi.EVAL_LVS_CONTINUOUS, lvs=prm["lvs"], fixed=fixed,
                          want=capi.WANT_COST | capi.WANT_CONTACTS, **kw2)
            r2.cost()
            return 1e3 * (time.perf_counter() - t0), r2.counts()["n_dense"]
        for _ in range(3):
            one(cu, sg, res)
        ts = [one(cu, sg, res) for _ in range(20)]
        entry = {"gpu_ms": float(np.median([t for t, _ in ts])), "dense_evals_per_pass": int(ts[0][1])}
        if orc is not None:
            so = orc.scene(sc, prm["margin"], prm["coeff"], prm["buffer"])
            entry["cpu_port_1thread_ms"] = float(np.median([one(orc, so)[0] for _ in range(3)]))
        out[name] = entry
    return out


def other_configs(cu):
    """The remaining SURVEY.md 8(d) lines, device-resident unless said otherwise (CUDA events around the passes, median of
    5 after 3 warm-up passes; every pass after the first is speculative): C1 boxbot DISCRETE at B = 2^20 states (evals/s)
    and B = 1 (us per pass, host buffers, wall clock), the LVS_DISCRETE term set of C4 (the "discrete" half of the
    reference's discrete + cast costs), and one full SQP iteration of C2 with the host QP included (the in-repo ADMM
    solver of tests/sqp.py standing in for OSQP, which is not installed here)."""
    import torch
    capi = pkg.capi
    out = {}
    box = ms_.load_scene("boxbot")
    sb = cu.scene(box, 0.1, 1.0, 0.01)
    rng = np.random.default_rng(1)

    def timed(scene_h, q_np, evaluator, want, reps=5, **kw):
        B, T = q_np.shape[0], q_np.shape[1]
        qd = torch.from_numpy(q_np).cuda()
        res = capi.Result(cu, scene_h, cu.new_result(scene_h), B, T)
        st = torch.cuda.current_stream().cuda_stream
        ts = []
        for i in range(3 + reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            cu.eval_device(scene_h, qd.data_ptr(), B, T, evaluator, res, stream=st, want=want, sync=False, **kw)
            b.record()
            cu.sync(res)
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(a.elapsed_time(b))
        return float(np.median(ts)), res.counts()
    q = rng.uniform(-2.2, 2.2, (1 << 20, 1, 2))
    ms1, c1 = timed(sb, q, capi.EVAL_DISCRETE, capi.WANT_LINEARISE)
    out["C1 boxbot DISCRETE B=2^20 (simple_collision_unit, 1 timestep per problem)"] = {
        "ms_per_pass": ms1, "evals_per_s": c1["n_dense"] / (ms1 * 1e-3), "n_contacts": c1["n_contacts"]}
    q1 = np.array([[[-0.9, 0.01]]])
    r1 = capi.Result(cu, sb, cu.new_result(sb), 1, 1)
    lat = []
    for i in range(30):
        t0 = time.perf_counter()
        r = cu.eval(sb, q1, capi.EVAL_DISCRETE, want=capi.WANT_LINEARISE, result=r1)
        r.rows()
        lat.append(1e6 * (time.perf_counter() - t0))
    out["C1 boxbot DISCRETE B=1"] = {"us_per_pass_host_buffers": float(np.median(lat[5:]))}
    arm = ms_.load_scene("arm_around_table")
    sa = cu.scene(arm, MARGIN, COEFF, BUFFER)
    qa = make_batch(arm, 0, 4096)
    ms2, c2 = timed(sa, qa, capi.EVAL_LVS_DISCRETE, capi.WANT_LINEARISE, lvs=LVS, fixed=fixed_steps())
    out["C4 seeds, LVS_DISCRETE term set (discrete half of discrete + cast)"] = {
        "ms_per_pass": ms2, "evals_per_s": c2["n_dense"] / (ms2 * 1e-3), "n_contacts": c2["n_contacts"]}
    ms3, c3 = timed(sa, qa, capi.EVAL_LVS_CONTINUOUS, capi.WANT_LINEARISE, lvs=LVS, fixed=fixed_steps(),
                    precision=capi.PRECISION_F32)
    out["C4 seeds, LVS_CONTINUOUS in FP32 (TC_PRECISION_F32; tolerance: tests/test_gpu_fullsize.py::test_fp32_mode_tolerance)"] = {
        "ms_per_pass": ms3, "evals_per_s": c3["n_dense"] / (ms3 * 1e-3), "n_contacts": c3["n_contacts"]}
    # one SQP iteration of C2 with the QP: linearisation pass + QP build and solve + one geometry pass at the trial point
    from tests import sqp
    prob = sqp.CollisionProblem(cu, arm, arm["seed_traj"], capi.EVAL_LVS_CONTINUOUS, MARGIN, COEFF, BUFFER, LVS)
    x = prob.q0[1:-1].reshape(-1).copy()
    prob.cost(x)
    t_lin, t_qp, t_val = [], [], []
    for _ in range(5):
        t0 = time.perf_counter()
        lin = prob.convexify(x)
        t1 = time.perf_counter()
        x_new, _, _ = prob.solve_model(x, lin, 0.1)
        t2 = time.perf_counter()
        prob.cost(x_new)
        t3 = time.perf_counter()
        t_lin.append(1e3 * (t1 - t0)), t_qp.append(1e3 * (t2 - t1)), t_val.append(1e3 * (t3 - t2))
    out["C2 one SQP iteration with the host QP (in-repo ADMM, scipy; OSQP not installed)"] = {
        "linearise_ms": float(np.median(t_lin)), "qp_ms": float(np.median(t_qp)), "value_pass_ms": float(np.median(t_val)),
        "total_ms": float(np.median(t_lin) + np.median(t_qp) + np.median(t_val)), "rows": int(len(lin[1]))}
    return out


def run_reference(args, scene, rank, world):
    """--impl reference: the reference's CPU path.  tesseract/Bullet/KDL cannot be built in this image
    (DESIGN.md), so this times the CPU restatement (oracle/) with all host threads, each step a bounded sample."""
    if rank != 0:
        return
    from oracle import pyoracle
    capi = pyoracle.capi
    orc = pyoracle.OracleLib()
    cores = orc.num_threads()
    so = orc.scene(scene, MARGIN, COEFF, BUFFER)
    n = max(8, 2 * cores)
    q = make_batch(scene, 0, n)
    times, dense = [], 0
    for it in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        r = orc.eval(so, q, capi.EVAL_LVS_CONTINUOUS, lvs=LVS, fixed=fixed_steps(), n_threads=cores)
        r.rows()
        r.cost()
        dt = time.perf_counter() - t0
        dense = r.counts()["n_dense"]
        if it >= args.warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    value = dense / (ms * 1e-3)
    sample = f"{n} seeds per step of the C4 workload, linearisation pass incl. rows and cost, {cores} host threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C4 batched multi-start arm_around_table (bounded sample)", "seeds_per_step": n,
                   "timesteps": T_STEPS, "evaluator": "LVS_CONTINUOUS", "lvs": LVS, "link_pairs": so.num_pairs()},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def run_ours(args, scene, rank, world, local_rank):
    import ctypes as C
    import torch
    capi = pkg.capi
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    cu = pkg.CudaLib(device=local_rank)
    sg = cu.scene(scene, MARGIN, COEFF, BUFFER)
    B = args.seeds
    q_host = make_batch(scene, rank * B, B)
    q_pin = torch.from_numpy(q_host).pin_memory()
    q_dev = q_pin.cuda()
    fixed = fixed_steps()
    res = capi.Result(cu, sg, cu.new_result(sg), B, T_STEPS)
    stream = torch.cuda.current_stream().cuda_stream
    comm = None
    if world > 1:
        idbuf = [None]
        if rank == 0:
            raw = C.create_string_buffer(128)
            cu.check(cu.dll.tc_nccl_unique_id(raw))
            idbuf = [raw.raw]
        dist.broadcast_object_list(idbuf, src=0)
        comm = C.c_void_p()
        cu.check(cu.dll.tc_nccl_comm_create(C.c_int(world), C.c_int(rank), idbuf[0], C.byref(comm)))
    cost_traj = np.zeros(B)
    win_traj = np.zeros(q_host.shape[1:])
    pd = C.POINTER(C.c_double)

    def pick_best():
        # arg-min over the per-trajectory costs on the device, record (cost, index, trajectory) exchanged with ONE
        # all-gather on the evaluation's stream (a single rank just reads its record back): tc_result_allgather_best
        bc, bi, br = C.c_double(), C.c_int64(), C.c_int32()
        cu.check(cu.dll.tc_result_allgather_best(res.handle, C.c_int64(rank * B), comm, C.c_int(world), C.byref(bc),
                                                 C.byref(bi), C.byref(br), win_traj.ctypes.data_as(pd)))
        return bc.value, bi.value, br.value

    def step_device():
        # want = cost + contacts + rows: the gradient vectors are folded into the CSR rows inside the pass.  The pass
        # is only enqueued (speculative sizing after the first one); pick_best waits for it and for the all-gather
        cu.eval_device(sg, q_dev.data_ptr(), B, T_STEPS, capi.EVAL_LVS_CONTINUOUS, res, lvs=LVS, fixed=fixed,
                       stream=stream, want=capi.WANT_LINEARISE, sync=False)
        return pick_best()

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step_device()
    counts = res.counts()
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    narrow_ms, launches, best = [], 0, None
    rank_ms = []
    e0.record()
    for _ in range(args.steps):
        best = step_device()
        st_ms = cu.stage_ms(res)
        narrow_ms.append(st_ms[5])
        rank_ms.append(st_ms[4])
        launches += cu.launches(res)[0]
    e1.record()
    barrier()
    sampler.stop_flag = True
    ms = e0.elapsed_time(e1) / args.steps
    counts = res.counts()
    stage = cu.stage_ms(res)

    # ---- device-side assembly of the hinge-cost QP block in CSC form, timed on its own: not part of
    # the step, reported next to it
    csc_ms = []
    for _ in range(3):
        step_device()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        cu.check(cu.dll.tc_result_qp_csc_device(res.handle, None, None, None, None, None, None, None, None))
        b.record()
        torch.cuda.synchronize()
        csc_ms.append(a.elapsed_time(b))

    # ---- end to end through the C ABI with HOST buffers: H2D of q, D2H of rows + costs, every step.
    # tc_pipeline_eval_rows: sub-batches of --e2e-chunk seeds on a ring of --e2e-workers streams inside the library,
    # compact rows (values + row value + link pair + term offsets) into pinned host arrays, one host thread.
    nd = sg.n_dof
    tpt = T_STEPS - 1
    pipe = capi.Pipeline(cu, sg, args.e2e_workers)
    n_rows0, nnz0 = counts["n_contacts"], counts["nnz"]

    def pinned(n, dt):
        return torch.empty(int(n), dtype={np.int64: torch.int64, np.float64: torch.float64, np.int32: torch.int32}[dt]).pin_memory().numpy()
    pipe.reserve(B, tpt, int(n_rows0 * 1.05) + 4096, int(nnz0 * 1.05) + 65536, alloc=pinned)
    q_np = q_pin.numpy()

    def step_e2e():
        n_rows, nnz_ = pipe.eval_rows(q_np, capi.EVAL_LVS_CONTINUOUS, LVS, fixed, args.e2e_chunk, packed=not args.e2e_dense)
        costs = pipe.out["cost"]
        i = int(np.argmin(costs))
        if comm is None:
            return (float(costs[i]), rank * B + i, 0), n_rows, nnz_
        bc, bi, br = C.c_double(), C.c_int64(), C.c_int32()
        cu.check(cu.dll.tc_allgather_best(C.c_double(costs[i]), C.c_int64(rank * B + i), comm, C.c_int(world),
                                          C.c_void_p(stream), C.byref(bc), C.byref(bi), C.byref(br)))
        return (bc.value, bi.value, br.value), n_rows, nnz_

    for _ in range(3):
        best_e2e, nc, nnz = step_e2e()
    assert best_e2e[1] == best[1] or best_e2e[0] == best[0], (best_e2e, best)  # same winner as the resident path
    # packed rows leave out the joints that move neither link of a row (exact zeros, a third of the values)
    assert nc == counts["n_contacts"] and (nnz == counts["nnz"] if args.e2e_dense else 0 < nnz < counts["nnz"])
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    barrier()
    e2e_ms = 1e3 * (time.perf_counter() - t0) / args.steps
    pipe.close()
    d2h = 8 * nnz + 12 * nc + 8 * B * tpt + 8 * B
    h2d = q_host.nbytes

    t = torch.tensor([ms, e2e_ms, float(np.mean(rank_ms))], dtype=torch.float64, device="cuda")
    tmin = torch.tensor([ms, e2e_ms, float(np.mean(rank_ms))], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(counts["n_dense"])], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    # per-rank spread: step = pass + winner exchange (max over ranks is the reported time); pass_ms = the kernels of the
    # pass alone.  step_max - pass_max is what the exchange and the waiting for the slowest rank cost
    per_rank = {"step_ms_min": float(tmin[0]), "step_ms_max": float(t[0]), "pass_ms_min": float(tmin[2]),
                "pass_ms_max": float(t[2]), "e2e_ms_min": float(tmin[1]), "e2e_ms_max": float(t[1])}
    ms, e2e_ms = float(t[0]), float(t[1])
    dense_total = float(tot[0])
    if rank == 0:
        peak64 = cu.measure_fma_peak(capi.PRECISION_F64)
        flops = 5.0 * counts["support_verts"] + 280.0 * counts["gjk_iters"] + 400.0 * counts["epa_iters"]
        nms = float(np.mean(narrow_ms))
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        flops_exec = 5.0 * counts["support_compared"] + 280.0 * counts["gjk_iters"] + 400.0 * counts["epa_iters"]
        traffic, traffic_src = None, None
        for name in ("r2_traffic.json", "r1_traffic.json"):
            try:
                tr = json.load(open(os.path.join(ROOT, "profiles", name)))
                traffic = sum(v["dram_bytes_per_launch"] for k, v in tr.items() if k.startswith(("k_gjk", "k_epa")))
                traffic_src = f"ncu --set full capture committed as profiles/{name} (" + str(tr.get("_commit", "round 1 build")) + \
                    "); not measured in this run"
                break
            except Exception:
                pass
        base, _, _ = cpu_baseline(scene) if (world == 1 and not args.no_cpu_baseline and not args.skip_extras) else ({"value": None, "unit": UNIT, "cores": 0,
                                                                            "kind": "port", "sample": "skipped"}, 0, 0)
        # HBM-bound rest of the pass (FK, broadphase, finalise, sort, gradients + rows, cost): algorithmic bytes of
        # SURVEY.md 8(d) over the time of everything that is not the narrow phase
        n_act = int(scene["n_active_links"])
        r_row = 2 * nd * 12 + 24
        hbm_bytes = (counts["n_states"] * (8 * nd + (96 + 96 + 48) * n_act) + 16 * counts["n_narrowphase"] +
                     counts["n_contacts"] * (176 + r_row))
        rest_ms = ms - nms
        extras = world == 1 and not args.skip_extras
        latency = sqp_iteration_latency(cu, with_cpu=not args.no_cpu_baseline) if extras else None
        others = other_configs(cu) if extras else None
        out = {
            "metric": METRIC, "value": dense_total / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C4 batched multi-start arm_around_table (BASELINE.json configs[3])",
                       "seeds_per_gpu": B, "global_seeds": B * world, "timesteps": T_STEPS, "n_dof": nd,
                       "evaluator": "LVS_CONTINUOUS", "lvs": LVS, "margin": MARGIN, "buffer": BUFFER,
                       "link_pairs": sg.num_pairs(), "parallelism": f"seeds sharded over {world} GPU(s)",
                       "l2": "inputs larger than L2 (poses+AABB arenas %.2f GB per pass)" %
                             (counts["n_states"] * (14 * 96 + 14 * 48) / 1e9)},
            "e2e": {"value": dense_total / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "rows_form": "dense blocks" if args.e2e_dense else "packed (structural non-zeros only, lossless)",
                    "row_values_per_step": int(nnz), "row_values_dense": int(counts["nnz"]),
                    "how": f"tc_pipeline_eval_rows{'' if args.e2e_dense else '_packed'}: sub-batches of {args.e2e_chunk} seeds "
                           f"(last one tapered) on a ring of {args.e2e_workers} streams inside the library, one host thread, "
                           "compact rows (values, row value, link pair, term offsets) into pinned host arrays, wall clock "
                           "around the whole batch"},
            "per_rank": per_rank,
            "gpu_launches": launches, "library_calls_per_step": cu.launches(res)[1],
            "roofline": {"bound": "fp64", "kernel": "narrow phase: k_gjk<double> + k_epa_wp<double> (2 launches)", "achieved": flops / (nms * 1e-3) / 1e12,
                         "peak": peak64, "unit": "TFLOP/s", "frac": flops / (nms * 1e-3) / 1e12 / peak64,
                         "frac_executed": flops_exec / (nms * 1e-3) / 1e12 / peak64,
                         "frac_note": "frac counts 5 flop per vertex of every queried hull (the reference's plain scan, "
                                      "SURVEY 8(d)); frac_executed counts the vertices the look-up tables left to compare",
                         "executed_flops_per_launch": flops_exec,
                         "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": "DFMA loop measured in this run (tc_measure_fma_peak)",
                         "kernel_ms": nms, "algorithmic_flops_per_launch": flops,
                         "hbm_peak_gbs": peaks.get("hbm_gbs")},
            "roofline_hbm": {"bound": "hbm", "kernel": "everything but the narrow phase (FK, broadphase, finalise, sort, "
                             "gradient rows, cost)", "achieved": hbm_bytes / (rest_ms * 1e-3) / 1e9,
                             "peak": peaks.get("hbm_gbs"), "unit": "GB/s",
                             "frac": (hbm_bytes / (rest_ms * 1e-3) / 1e9 / peaks["hbm_gbs"]) if peaks.get("hbm_gbs") else None,
                             "algorithmic_bytes": int(hbm_bytes), "ms": rest_ms,
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs (burst copy)"},
            "qp_csc": {"ms": float(np.mean(csc_ms[1:])), "what": "tc_result_qp_csc_device: hinge-cost QP block [A_x | -I] of "
                       "all rows assembled on the device in CSC form (3 kernels), outside the timed step",
                       "nnz": int(counts["nnz"] + counts["n_contacts"])},
            "sqp_iteration_ms": latency,
            "other_configs": others,
            "cpu_baseline": base,
            "counters": {k: counts[k] for k in ("n_dense", "n_units", "n_narrowphase", "n_contacts", "gjk_iters",
                                                "epa_iters", "n_epa", "support_verts", "support_compared", "n_unconverged")},
            "stage_ms": {"plan_fk": stage[0], "broadphase": stage[6], "narrow_kernels": stage[5], "gjk_kernel": stage[8],
                         "epa_kernel": stage[9], "deep_epa": stage[2],
                         "finalize_sort_rows": stage[3], "eval_total": stage[4]},
            "best": {"cost": best[0], "seed": best[1], "rank": best[2]},
            "clocks": sampler.summary(),
        }
        print(json.dumps(out))
    if comm is not None:
        cu.dll.tc_nccl_comm_destroy(comm)
    if dist is not None:
        dist.destroy_process_group()


def main():
    # Exactly ONE line may reach stdout (the JSON): libraries print there too (NCCL announces its version on the first
    # communicator), so file descriptor 1 is pointed at stderr for the run and the JSON goes to the saved descriptor.
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--seeds", type=int, default=4096, help="trajectories per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--skip-extras", action="store_true", help="tuning runs: no CPU baseline, single-problem latencies or "
                    "other configs (the default run reports them all)")
    ap.add_argument("--e2e-chunk", type=int, default=1024, help="seeds per sub-batch of the host-buffer (e2e) leg")
    ap.add_argument("--e2e-workers", type=int, default=4, help="streams (ring slots) of the e2e leg")
    ap.add_argument("--e2e-dense", action="store_true", help="e2e leg downloads dense row blocks (n_dof values per block) "
                                                             "instead of the packed form")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    scene = ms_.load_scene("arm_around_table")
    if args.impl == "reference":
        run_reference(args, scene, rank, world)
    else:
        run_ours(args, scene, rank, world, local_rank)
    real_stdout.flush()


if __name__ == "__main__":
    main()
