"""CPU suite: host logic of the sharded multi-start — contiguous sharding, winner selection, and the
one exchange (all-gather of best cost/index) run with two gloo ranks on CPU."""
import os
import socket

import numpy as np
import pytest

from tests import helpers

ms = helpers.pkg.multistart
capi = helpers.capi


def test_shards_are_contiguous_and_cover_everything():
    for n in (0, 1, 7, 8, 4096, 4097):
        for world in (1, 2, 3, 4, 8):
            spans = [ms.shard(n, r, world) for r in range(world)]
            assert sum(c for _, c in spans) == n
            pos = 0
            for first, count in spans:
                assert count == 0 or first == pos
                pos += count


def test_batches_do_not_depend_on_the_sharding():
    s = helpers.load_scene("arm_around_table")
    whole = ms.make_batch(s, 0, 6)
    parts = np.concatenate([ms.make_batch(s, *ms.shard(6, r, 3)) for r in range(3)])
    assert (whole == parts).all()
    lim = s["joint_limits"]
    assert (whole >= lim[:, 0]).all() and (whole <= lim[:, 1]).all()
    assert (whole[:, 0] == whole[0, 0]).all() and (whole[:, -1] == whole[0, -1]).all()  # endpoints shared


def test_winner_selection_rules():
    assert ms.select_best([3.0, 1.0, 1.0], 10) == (1.0, 11)
    assert ms.select_best([], 0) == (float("inf"), -1)
    assert ms.reduce_best([(2.0, 5), (1.0, 9), (1.0, 3)]) == (1.0, 3, 2)
    assert ms.reduce_best([(float("inf"), -1), (4.0, 8)]) == (4.0, 8, 1)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, n_seeds, out):
    import torch.distributed as dist
    from oracle import pyoracle
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        scene = helpers.load_scene("arm_around_table")
        first, count = ms.shard(n_seeds, rank, world)
        orc = pyoracle.OracleLib()
        so = orc.scene(scene, ms.MARGIN, ms.COEFF, ms.BUFFER)
        q = ms.make_batch(scene, first, count)
        r = orc.eval(so, q, capi.EVAL_LVS_CONTINUOUS, lvs=ms.LVS, fixed=ms.fixed_steps(), n_threads=2)
        _, cost = r.cost()
        local = ms.select_best(cost, first)
        best = ms.allgather_best(*local)
        # the owner ships the winning trajectory; the other ranks pass a buffer of the same shape
        mine = q[local[1] - first] if count else np.zeros(q.shape[1:])
        win = ms.broadcast_trajectory(mine, best[2])
        dense = r.counts()["n_dense"]
        out.put((rank, best, dense, win))
    finally:
        dist.destroy_process_group()


def test_two_ranks_agree_with_one(oracle):
    """world_size 2 over gloo: both ranks see the same winner, equal to the single-process answer; the unit
    count the bench reports (N_dense) is additive over shards"""
    import torch.multiprocessing as mp
    n_seeds, world = 6, 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, world, port, n_seeds, out)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted((out.get(timeout=240) for _ in range(world)), key=lambda g: g[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    scene = helpers.load_scene("arm_around_table")
    so = oracle.scene(scene, ms.MARGIN, ms.COEFF, ms.BUFFER)
    r = oracle.eval(so, ms.make_batch(scene, 0, n_seeds), capi.EVAL_LVS_CONTINUOUS, lvs=ms.LVS, fixed=ms.fixed_steps())
    _, cost = r.cost()
    want_cost, want_idx = ms.select_best(cost, 0)
    assert got[0][1] == got[1][1]
    c, i, owner = got[0][1]
    assert i == want_idx and abs(c - want_cost) <= 1e-12 * max(1.0, abs(want_cost))
    assert owner == (0 if i < ms.shard(n_seeds, 0, world)[1] else 1)
    assert got[0][2] + got[1][2] == r.counts()["n_dense"]
    # every rank holds the winner's trajectory after the broadcast
    want_traj = ms.make_batch(scene, want_idx, 1)[0]
    assert np.array_equal(got[0][3], want_traj) and np.array_equal(got[1][3], want_traj)


@pytest.mark.gpu
def test_result_allgather_best_single_rank_reads_the_device_winner(cuda):
    """tc_result_allgather_best with one rank: arg-min over the per-trajectory costs and the winner's joint values come
    back from the device record (lowest index on ties), no communicator needed"""
    import ctypes as C
    from tests import helpers
    capi = helpers.capi
    s = ms.load_scene()
    sg = cuda.scene(s, ms.MARGIN, ms.COEFF, ms.BUFFER)
    q = ms.make_batch(s, 11, 24)
    r = cuda.eval(sg, q, capi.EVAL_LVS_CONTINUOUS, lvs=ms.LVS, fixed=ms.fixed_steps(), want=capi.WANT_COST | capi.WANT_CONTACTS)
    _, costs = r.cost()
    bc, bi, br = C.c_double(), C.c_int64(), C.c_int32()
    traj = np.zeros(q.shape[1:])
    cuda.check(cuda.dll.tc_result_allgather_best(r.handle, C.c_int64(1000), None, C.c_int(1), C.byref(bc), C.byref(bi), C.byref(br),
                                                 traj.ctypes.data_as(C.POINTER(C.c_double))))
    i = int(np.argmin(costs))
    assert bc.value == costs[i] and bi.value == 1000 + i and br.value == 0
    assert np.array_equal(traj, q[i])
