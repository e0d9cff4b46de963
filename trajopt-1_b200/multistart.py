"""Host-side logic of the batched multi-start workload (BASELINE.json configs[3]):
seed generation, contiguous sharding of seeds over ranks, and winner selection.

The path shards over independent trajectories only; there is no data-path collective.  The single exchange
is the all-gather of every rank's (best cost, global seed index), followed by a broadcast of the winning
trajectory from its owner: `tc_allgather_best` / `tc_broadcast_trajectory` (NCCL, inside libtrajopt_cuda.so) on GPUs; `allgather_best` below is the same reduction over any torch.distributed group and
is what the world_size-2 gloo tests exercise on CPU."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCENES = os.path.join(ROOT, "tests", "golden", "scenes")

T_STEPS, LVS, MARGIN, COEFF, BUFFER, SIGMA = 10, 0.02, 0.025, 20.0, 0.01, 0.3


def load_scene(name="arm_around_table"):
    """scene fixture generated from the reference's URDF/SRDF/STL data by tools/make_fixtures.py"""
    return dict(np.load(os.path.join(SCENES, name + ".npz")))


def make_batch(scene, first_seed, n, t_steps=T_STEPS, sigma=SIGMA):
    """Seed k: start/goal of the fixture path resampled to `t_steps` waypoints, interior waypoints perturbed by
    N(0, sigma^2) rad per joint with rng seed 4000 + k, clamped to the URDF limits.  Depends only on k, so any
    sharding produces the same global batch."""
    w = np.linspace(0.0, 1.0, t_steps)[:, None]
    base = scene["seed_traj"][0] * (1 - w) + scene["seed_traj"][-1] * w
    q = np.repeat(base[None], n, 0)
    for i in range(n):
        rng = np.random.default_rng(4000 + first_seed + i)
        q[i, 1:-1] += rng.normal(0.0, sigma, (t_steps - 2, base.shape[1]))
    lim = scene["joint_limits"]
    return np.ascontiguousarray(np.clip(q, lim[:, 0], lim[:, 1]))


def fixed_steps(t_steps=T_STEPS):
    f = np.zeros(t_steps, np.uint8)
    f[0] = f[-1] = 1
    return f


def shard(n_global, rank, world):
    """contiguous shard [first, first + count) of rank: B_r = ceil(B / R), the tail rank may get fewer"""
    per = -(-n_global // world)
    first = min(rank * per, n_global)
    return first, max(0, min(per, n_global - first))


def select_best(costs, first_index=0):
    """(cost, global index) of the cheapest trajectory of a shard; ties go to the lowest index"""
    costs = np.asarray(costs, dtype=np.float64)
    if costs.size == 0:
        return float("inf"), -1
    i = int(np.argmin(costs))
    return float(costs[i]), first_index + i


def reduce_best(pairs):
    """winner over per-rank (cost, index) pairs -> (cost, index, rank); ties: lowest index, empty shards ignored"""
    best = (float("inf"), -1, -1)
    for r, (c, i) in enumerate(pairs):
        if i < 0:
            continue
        if c < best[0] or (c == best[0] and (best[1] < 0 or i < best[1])):
            best = (float(c), int(i), r)
    return best


def allgather_best(local_cost, local_index, group=None):
    """same exchange as tc_allgather_best over a torch.distributed group (16 bytes per rank)"""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    mine = torch.tensor([float(local_cost), float(local_index)], dtype=torch.float64)
    out = [torch.zeros(2, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    return reduce_best([(float(t[0]), int(t[1])) for t in out])


def broadcast_trajectory(traj, root, group=None):
    """same exchange as tc_broadcast_trajectory over a torch.distributed group: the owner of the winning seed
    ships its [T, n_dof] trajectory to every rank"""
    import torch
    import torch.distributed as dist
    t = torch.from_numpy(np.ascontiguousarray(traj, dtype=np.float64).copy())
    dist.broadcast(t, src=root, group=group)
    return t.numpy()


class PipelinedEvaluator:
    """End-to-end multi-start pass with HOST buffers: the batch is cut into sub-batches that go round-robin to a
    few worker threads, each with its own CUDA stream, tc_result and pinned output buffers.  While one worker's
    rows travel device -> host, the other workers' kernels run, so the PCIe time of the rows (the bulk of the
    bytes) hides behind the compute of the next sub-batch.  Every sub-batch is an ordinary
    tc_eval + tc_result_rows + tc_result_cost through the C ABI; seeds are independent, so the host QP of a
    sub-batch can start as soon as its rows have landed.

    torch is used for pinned host memory and stream objects only."""

    def __init__(self, cu, scene, evaluator, lvs, fixed, n_dof, chunk=1024, n_workers=2, rows_per_seed=1400,
                 concurrent=1, taper=True, compact=True):
        import queue
        import threading
        import torch
        self.cu, self.scene, self.evaluator, self.lvs, self.fixed = cu, scene, evaluator, lvs, fixed
        self.chunk, self.n_dof, self.taper, self.compact = chunk, n_dof, taper, compact
        self.torch = torch
        self.capi = cu.capi if hasattr(cu, "capi") else None
        self.workers = []
        self.jobs = [queue.Queue() for _ in range(n_workers)]
        self.done = queue.Queue()
        self.cap = int(chunk * rows_per_seed)
        # at most `concurrent` sub-batches compute at a time; the slot is given up once the device -> host copies
        # are enqueued, so a worker's copies run under other workers' kernels instead of all workers moving in
        # lock step
        self.compute = threading.Semaphore(concurrent)
        for w in range(n_workers):
            st = torch.cuda.Stream()
            bufs = self._alloc(self.cap)
            t = threading.Thread(target=self._work, args=(w, st, bufs), daemon=True)
            t.start()
            self.workers.append((t, st, bufs))

    def _alloc(self, cap):
        torch, nd = self.torch, self.n_dof
        pin = lambda n, dt: torch.empty(n, dtype=dt).pin_memory().numpy()
        return dict(cap=cap, row_ptr=pin(cap + 1, torch.int64), col=pin(2 * cap * nd, torch.int32),
                    val=pin(2 * cap * nd, torch.float64), value=pin(cap, torch.float64), coeff=pin(cap, torch.float64),
                    margin=pin(cap, torch.float64), cost=pin(self.chunk, torch.float64),
                    first=pin(2 * cap, torch.int32))

    def _work(self, w, stream, bufs):
        import ctypes as C
        from . import capi
        cu = self.cu
        torch = self.torch
        res = None
        pd, pi, pl = C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_int64)
        while True:
            job = self.jobs[w].get()
            if job is None:
                return
            idx, q = job
            try:
                if res is None:
                    res = capi.Result(cu, self.scene, cu.new_result(self.scene), len(q), q.shape[1])
                with self.compute:
                    r = cu.eval(self.scene, q, self.evaluator, lvs=self.lvs, fixed=self.fixed, result=res,
                                stream=stream.cuda_stream, want=capi.WANT_LINEARISE)
                    n = r.counts()["n_contacts"]
                    if n > bufs["cap"]:
                        bufs.update(self._alloc(int(n * 1.25)))
                    # the copies are enqueued before the next sub-batch may start (still under the lock) and
                    # awaited after it has started
                    # compact: the rows travel in block form (first column of the T0 / T1 run of every row instead
                    # of one column index per non-zero, tc_result_rows_blocks_async) — the device -> host path is what
                    # bounds this pass
                    fn = cu.dll.tc_result_rows_blocks_async if self.compact else cu.dll.tc_result_rows_async
                    cols = bufs["first"] if self.compact else bufs["col"]
                    cu.check(fn(res.handle, bufs["row_ptr"].ctypes.data_as(pl), cols.ctypes.data_as(pi),
                                bufs["val"].ctypes.data_as(pd), bufs["value"].ctypes.data_as(pd),
                                bufs["coeff"].ctypes.data_as(pd), bufs["margin"].ctypes.data_as(pd)))
                cu.check(cu.dll.tc_result_wait(res.handle))
                cu.check(cu.dll.tc_result_cost(res.handle, None, bufs["cost"].ctypes.data_as(pd)))
                c = r.counts()
                self.done.put((idx, w, n, c["nnz"], c["n_dense"], bufs["cost"][:len(q)].copy(), None))
            except Exception as e:  # surfaced by run()
                self.done.put((idx, w, 0, 0, 0, None, e))

    def run(self, q_pinned):
        """q_pinned: numpy view of pinned host memory [B, T, n_dof].  Returns (costs[B], n_rows, nnz, n_dense);
        the rows of sub-batch i are in worker (i % n_workers)'s pinned buffers when its entry arrives."""
        B = len(q_pinned)
        # sub-batch bounds: whole chunks, the last one cut in halves and quarters — the run ends with the download of
        # the last sub-batch, which nothing overlaps, so that one is kept small
        bounds, lo = [], 0
        while lo < B:
            hi = min(B, lo + self.chunk)
            if hi == B and self.taper and hi - lo >= 4:
                h, q = lo + (hi - lo) // 2, lo + 3 * (hi - lo) // 4
                bounds += [(lo, h), (h, q), (q, hi)]
            else:
                bounds.append((lo, hi))
            lo = hi
        for i, (lo, hi) in enumerate(bounds):
            self.jobs[i % len(self.jobs)].put((i, q_pinned[lo:hi]))
        costs = np.zeros(B)
        rows = nnz = dense = 0
        for _ in range(len(bounds)):
            idx, w, n, z, d, cost, err = self.done.get()
            if err is not None:
                raise err
            costs[bounds[idx][0]: bounds[idx][0] + len(cost)] = cost
            rows, nnz, dense = rows + n, nnz + z, dense + d
        return costs, rows, nnz, dense

    def close(self):
        for q in self.jobs:
            q.put(None)
        for t, _, _ in self.workers:
            t.join(timeout=10)
