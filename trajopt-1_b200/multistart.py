"""Host-side logic of the batched multi-start workload (BASELINE.json configs[3]):
seed generation, contiguous sharding of seeds over ranks, and winner selection.

The path shards over independent trajectories only; there is no data-path collective.  The single exchange
is the all-gather of every rank's best (cost, global seed index, trajectory): `tc_result_allgather_best` (NCCL, inside
libtrajopt_cuda.so, fed from the device) on GPUs; `allgather_best` / `broadcast_trajectory` below are the same
reduction over any torch.distributed group and are what the world_size-2 gloo tests exercise on CPU.
The host-buffer pass over a large batch (sub-batches on a ring of streams, compact rows) lives inside the library:
`tc_pipeline_eval_rows` (capi.Pipeline)."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCENES = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "scenes")

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
