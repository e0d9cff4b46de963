"""tc_cost_noalloc (the allocation-free entry point, static device scratch) against tc_result_cost, through the C ABI.
Collected last on purpose (file name).  Two calls in a row share the static array; the
second uses a different batch so stale contents of the first would show."""
import numpy as np
import pytest

from tests import helpers

capi = helpers.capi
pytestmark = pytest.mark.gpu

COST_TOL = 1e-12  # relative; same arithmetic and summation order as tc_result_cost, so the difference should be 0


def _batch(s, B, seed):
    rng = np.random.default_rng(seed)
    q = np.repeat(s["seed_traj"][None], B, 0)
    q[:, 1:-1] += rng.normal(0, 0.3, (B, q.shape[1] - 2, q.shape[2]))
    lim = s["joint_limits"]
    return np.clip(q, lim[:, 0], lim[:, 1])


def _noalloc_cost(cuda, sg, r, torch, stream=None):
    """per-trajectory cost of result r through tc_cost_noalloc; the test owns every device array"""
    cnt = r.counts()
    n_terms = cnt["n_terms"]
    assert n_terms % r.B == 0
    c = r.contacts()
    _, _, margin, coeff = sg.pairs()
    dev = torch.device("cuda:0")
    term_ptr = torch.from_numpy(r.term_ptr()).to(dev)
    dist = torch.from_numpy(np.ascontiguousarray(c["distance"])).to(dev)
    mar = torch.from_numpy(np.ascontiguousarray(margin[c["pair"]])).to(dev)
    coe = torch.from_numpy(np.ascontiguousarray(coeff[c["pair"]])).to(dev)
    out = torch.full((r.B,), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    status = cuda.cost_noalloc(term_ptr.data_ptr(), n_terms, n_terms // r.B, dist.data_ptr(), mar.data_ptr(),
                               coe.data_ptr(), out.data_ptr(), stream=stream.cuda_stream if stream else None)
    cuda.check(status)
    if stream is not None:
        stream.synchronize()  # the call only enqueues; the caller's stream orders it
    else:
        torch.cuda.synchronize()
    return out.cpu().numpy()


def test_noalloc_cost_matches_result_cost_twice(cuda):
    import torch
    s = helpers.load_scene("arm_around_table")
    sg = cuda.scene(s, 0.025, 20.0, 0.01)
    fixed = np.zeros(s["seed_traj"].shape[0], np.uint8)
    fixed[0] = fixed[-1] = 1
    seen = 0
    for B, seed, evaluator in ((24, 7, capi.EVAL_LVS_CONTINUOUS), (9, 11, capi.EVAL_LVS_DISCRETE)):
        r = cuda.eval(sg, _batch(s, B, seed), evaluator, lvs=0.05, fixed=fixed)
        _, want = r.cost()
        got = _noalloc_cost(cuda, sg, r, torch)
        assert np.all(np.isfinite(got))
        assert np.allclose(got, want, rtol=COST_TOL, atol=0.0), np.abs(got - want).max()
        seen += int(np.count_nonzero(want))
    assert seen > 0  # the batches do produce hinge cost


def test_noalloc_on_a_caller_stream(cuda):
    import torch
    s = helpers.load_scene("arm_around_table")
    sg = cuda.scene(s, 0.025, 20.0, 0.01)
    r = cuda.eval(sg, _batch(s, 16, 3), capi.EVAL_LVS_CONTINUOUS, lvs=0.05)
    _, want = r.cost()
    side = torch.cuda.Stream()
    for _ in range(2):  # back to back on one stream: ordered, so the shared scratch is safe
        got = _noalloc_cost(cuda, sg, r, torch, stream=side)
        assert np.allclose(got, want, rtol=COST_TOL, atol=0.0), np.abs(got - want).max()
    assert np.count_nonzero(want) > 0


def test_noalloc_inside_a_cuda_graph(cuda):
    """the use the entry point exists for: captured into a CUDA graph (no allocation, no synchronisation, the scratch
    address resolved by tc_init beforehand) and replayed on new inputs written into the same device arrays"""
    import torch
    s = helpers.load_scene("arm_around_table")
    sg = cuda.scene(s, 0.025, 20.0, 0.01)
    fixed = np.zeros(s["seed_traj"].shape[0], np.uint8)
    fixed[0] = fixed[-1] = 1
    B = 12
    runs = [cuda.eval(sg, _batch(s, B, seed), capi.EVAL_LVS_CONTINUOUS, lvs=0.05, fixed=fixed,
                      result=None) for seed in (21, 22)]
    _, _, margin, coeff = sg.pairs()
    dev = torch.device("cuda:0")
    n_terms = runs[0].counts()["n_terms"]
    cap = max(len(r.contacts()) for r in runs)
    term_ptr = torch.zeros(n_terms + 1, dtype=torch.int64, device=dev)
    dist, mar, coe = (torch.zeros(cap, dtype=torch.float64, device=dev) for _ in range(3))
    out = torch.zeros(B, dtype=torch.float64, device=dev)

    def load(r):
        c = r.contacts()
        term_ptr.copy_(torch.from_numpy(r.term_ptr()))
        dist[:len(c)].copy_(torch.from_numpy(np.ascontiguousarray(c["distance"])))
        mar[:len(c)].copy_(torch.from_numpy(np.ascontiguousarray(margin[c["pair"]])))
        coe[:len(c)].copy_(torch.from_numpy(np.ascontiguousarray(coeff[c["pair"]])))

    load(runs[0])
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        cuda.check(cuda.cost_noalloc(term_ptr.data_ptr(), n_terms, n_terms // B, dist.data_ptr(), mar.data_ptr(),
                                     coe.data_ptr(), out.data_ptr(), stream=torch.cuda.current_stream().cuda_stream))
    for r in (runs[0], runs[1], runs[0]):
        load(r)
        out.fill_(float("nan"))
        g.replay()
        torch.cuda.synchronize()
        _, want = r.cost()
        assert np.allclose(out.cpu().numpy(), want, rtol=COST_TOL, atol=0.0)
    assert np.count_nonzero(runs[1].cost()[1]) > 0


def test_noalloc_rejects_sizes_beyond_the_static_scratch(cuda):
    import torch
    dev = torch.device("cuda:0")
    z = torch.zeros(8, dtype=torch.float64, device=dev)
    tp = torch.zeros(8, dtype=torch.int64, device=dev)
    n = capi.NOALLOC_MAX_TERMS + 1
    status = cuda.cost_noalloc(tp.data_ptr(), n, 1, z.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr())
    assert status == capi.TC_ERR_OVERFLOW
    assert "TC_NOALLOC_MAX_TERMS" in cuda.last_error()
    status = cuda.cost_noalloc(tp.data_ptr(), 7, 2, z.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr())
    assert status == capi.TC_ERR_INVALID
