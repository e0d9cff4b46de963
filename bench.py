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
    threads like BasicTrustRegionSQPMultiThreaded) on a bounded sample of the same workload"""
    from oracle import pyoracle
    capi = pyoracle.capi
    orc = pyoracle.OracleLib()
    cores = orc.num_threads()
    so = orc.scene(scene, MARGIN, COEFF, BUFFER)
    n = n_seeds or max(8, 2 * cores)
    q = make_batch(scene, 0, n)
    best, dense, t_end = None, 0, time.time() + seconds
    while True:
        t0 = time.perf_counter()
        r = orc.eval(so, q, capi.EVAL_LVS_CONTINUOUS, lvs=LVS, fixed=fixed_steps(), n_threads=cores)
        r.rows()
        dt = time.perf_counter() - t0
        dense = r.counts()["n_dense"]
        best = dt if best is None else min(best, dt)
        if time.time() > t_end:
            break
    return {"value": dense / best, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n} seeds of the C4 workload (seeds 4000..{4000 + n - 1}), linearisation pass incl. rows, "
                      f"best wall time in {seconds:.0f} s, std::thread over terms"}, best, dense


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
        res = capi.Result(cu, sg, cu.new_result(sg), 1, q.shape[1])

        def one(lib, scene, result=None):
            kw = dict(result=result) if result is not None else {}
            t0 = time.perf_counter()
            # the CPU port assembles its rows from stored gradients; the device folds them into the rows on the fly
            lin = capi.WANT_LINEARISE if result is not None else capi.WANT_ALL
            r = lib.eval(scene, q, capi.EVAL_LVS_CONTINUOUS, lvs=prm["lvs"], fixed=fixed, want=lin, **kw)
            r.rows()
            r2 = lib.eval(scene, q, capi.EVAL_LVS_CONTINUOUS, lvs=prm["lvs"], fixed=fixed,
                          want=capi.WANT_COST | capi.WANT_CONTACTS, **kw)
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
        cu.check(cu.dll.tc_result_cost(res.handle, None, cost_traj.ctypes.data_as(pd)))
        i = int(np.argmin(cost_traj))
        if comm is None:
            return float(cost_traj[i]), rank * B + i, 0
        bc, bi, br = C.c_double(), C.c_int64(), C.c_int32()
        cu.check(cu.dll.tc_allgather_best(C.c_double(cost_traj[i]), C.c_int64(rank * B + i), comm, C.c_int(world),
                                          C.c_void_p(stream), C.byref(bc), C.byref(bi), C.byref(br)))
        # the owner ships the winning trajectory to every rank (T x n_dof doubles)
        win_traj[:] = q_host[i]
        cu.check(cu.dll.tc_broadcast_trajectory(win_traj.ctypes.data_as(pd), C.c_size_t(win_traj.size), C.c_int(br.value),
                                                comm, C.c_void_p(stream)))
        return bc.value, bi.value, br.value

    def step_device():
        # want = cost + contacts + rows: the gradient vectors are folded into the CSR rows inside the pass
        cu.eval_device(sg, q_dev.data_ptr(), B, T_STEPS, capi.EVAL_LVS_CONTINUOUS, res, lvs=LVS, fixed=fixed,
                       stream=stream, want=capi.WANT_LINEARISE)
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
    e0.record()
    for _ in range(args.steps):
        best = step_device()
        narrow_ms.append(cu.stage_ms(res)[5])
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
    # Sub-batches of 1024 seeds alternate between two streams so the D2H of one overlaps the kernels of the next
    # (multistart.PipelinedEvaluator); each sub-batch is a plain tc_eval + tc_result_rows + tc_result_cost.
    nd = sg.n_dof
    pipe = ms_.PipelinedEvaluator(cu, sg, capi.EVAL_LVS_CONTINUOUS, LVS, fixed, nd, chunk=args.e2e_chunk,
                                  n_workers=args.e2e_workers, concurrent=args.e2e_concurrent,
                                  taper=bool(args.e2e_taper), compact=args.e2e_rows == "blocks")
    q_np = q_pin.numpy()

    def step_e2e():
        costs, n_rows, nnz_, _ = pipe.run(q_np)
        i = int(np.argmin(costs))
        if comm is None:
            return (float(costs[i]), rank * B + i, 0), n_rows, nnz_
        bc, bi, br = C.c_double(), C.c_int64(), C.c_int32()
        cu.check(cu.dll.tc_allgather_best(C.c_double(costs[i]), C.c_int64(rank * B + i), comm, C.c_int(world),
                                          C.c_void_p(stream), C.byref(bc), C.byref(bi), C.byref(br)))
        return (bc.value, bi.value, br.value), n_rows, nnz_

    for _ in range(2):
        best_e2e, nc, nnz = step_e2e()
    assert best_e2e[1] == best[1] or best_e2e[0] == best[0], (best_e2e, best)  # same winner as the resident path
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    barrier()
    e2e_ms = 1e3 * (time.perf_counter() - t0) / args.steps
    pipe.close()
    n_sub = -(-B // args.e2e_chunk) + (2 if args.e2e_taper else 0)
    d2h = 8 * (nc + n_sub) + (8 * nnz + 8 * nc if args.e2e_rows == "blocks" else 12 * nnz) + 24 * nc + 8 * B
    h2d = q_host.nbytes

    t = torch.tensor([ms, e2e_ms], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(counts["n_dense"])], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
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
        traffic = None
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "r1_traffic.json")))
            traffic = sum(v["dram_bytes_per_launch"] for k, v in tr.items() if k.startswith(("k_gjk", "k_epa")))
        except Exception:
            pass
        base, _, _ = cpu_baseline(scene) if (world == 1 and not args.no_cpu_baseline) else ({"value": None, "unit": UNIT, "cores": 0,
                                                                            "kind": "port", "sample": "skipped"}, 0, 0)
        # HBM-bound rest of the pass (FK, broadphase, finalise, sort, gradients + rows, cost): algorithmic bytes of
        # SURVEY.md 8(d) over the time of everything that is not the narrow phase
        n_act = int(scene["n_active_links"])
        r_row = 2 * nd * 12 + 24
        hbm_bytes = (counts["n_states"] * (8 * nd + (96 + 96 + 48) * n_act) + 16 * counts["n_narrowphase"] +
                     counts["n_contacts"] * (176 + r_row))
        rest_ms = ms - nms
        latency = sqp_iteration_latency(cu, with_cpu=not args.no_cpu_baseline) if world == 1 else None
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
                    "how": f"{args.e2e_workers} streams x sub-batches of {args.e2e_chunk} seeds ({args.e2e_concurrent} computing at a "
                           f"time{', last one tapered' if args.e2e_taper else ''}), rows as {args.e2e_rows}, pinned host buffers, "
                           "wall clock around the whole batch"},
            "gpu_launches": launches, "library_calls_per_step": cu.launches(res)[1],
            "roofline": {"bound": "fp64", "kernel": "narrow phase: k_gjk<double> + k_epa_wp<double> (2 launches)", "achieved": flops / (nms * 1e-3) / 1e12,
                         "peak": peak64, "unit": "TFLOP/s", "frac": flops / (nms * 1e-3) / 1e12 / peak64,
                         "traffic": traffic, "peak_source": "DFMA loop measured in this run (tc_measure_fma_peak)",
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
            "cpu_baseline": base,
            "counters": {k: counts[k] for k in ("n_dense", "n_units", "n_narrowphase", "n_contacts", "gjk_iters",
                                                "epa_iters", "n_epa", "support_verts", "n_unconverged")},
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
    ap.add_argument("--e2e-chunk", type=int, default=1024, help="seeds per sub-batch of the host-buffer (e2e) leg")
    ap.add_argument("--e2e-workers", type=int, default=3, help="streams / host threads of the e2e leg")
    ap.add_argument("--e2e-concurrent", type=int, default=2, help="sub-batches allowed to compute at the same time")
    ap.add_argument("--e2e-rows", default="blocks", choices=["blocks", "csr"],
                    help="form in which the rows reach the host: block form (tc_result_rows_blocks: first column of the "
                         "T0 / T1 run per row) or full CSR with one column index per non-zero (tc_result_rows)")
    ap.add_argument("--e2e-taper", type=int, default=1, help="cut the last sub-batch into 1/2 + 1/4 + 1/4 (its download "
                    "is the part nothing overlaps)")
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
