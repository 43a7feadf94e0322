"""GPU: avr_graph_propagate_noalloc (scratch in the library's static __device__ array, no allocation inside
the call) against avr_graph_propagate_device, through the C ABI, on two calls in a row.

Collected last on purpose: it must pass alone in a fresh process and
after every other test.  Bar: bit-exact, like the other SpMV tests (same kernels, same summation order).
"""
import ctypes as C

import numpy as np
import pytest

from acceleratedvolrenderer_b200 import scenes
from acceleratedvolrenderer_b200.api import AvrError

pytestmark = pytest.mark.gpu


def _bits(t):
    return t.cpu().numpy().view(np.uint32)


def test_noalloc_matches_the_allocating_entry_twice_in_a_row(ctx):
    import torch
    dev = torch.device("cuda", 0)
    # two graphs of different sizes, the larger one FIRST: the second call then runs over scratch that still
    # holds the first call's longer vectors and its flags, so stale contents would show
    cases = [(300_000, 5, 7, 11), (70_001, 4, 3, 12)]
    graphs = []
    for n, deg, bounces, seed in cases:
        rowptr, col, val = scenes.random_graph(n, avg_deg=deg, seed=seed)
        L0 = torch.from_numpy(np.random.default_rng(seed).random(n, dtype=np.float32)).to(dev)
        graphs.append((ctx.create_graph(rowptr, col, val), L0, bounces, n))
    want = []
    for g, L0, bounces, n in graphs:
        ref = torch.full((n,), float("nan"), device=dev)
        it = g.propagate_device(L0.data_ptr(), bounces, ref.data_ptr(), want_iters=True)
        want.append((ref, it))
    torch.cuda.synchronize()
    # everything the calls below need exists now: from here to the comparison nothing may allocate on the device
    outs = [torch.full((n,), float("nan"), device=dev) for _, _, _, n in graphs]
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
    got_iters, free = [], []
    for (g, L0, bounces, n), out in zip(graphs, outs):
        got_iters.append(g.propagate_noalloc(L0.data_ptr(), bounces, out.data_ptr(), want_iters=True))
        torch.cuda.synchronize()
        free.append(torch.cuda.mem_get_info(dev)[0])
    assert torch.cuda.memory_stats(dev)["allocation.all.allocated"] == before
    # the first call may pay the driver's lazy load of the init kernel; after it the device's free memory stands still
    assert free[1] >= free[0], "device memory shrank across a no-alloc call"
    for (ref, it), out, git in zip(want, outs, got_iters):
        assert git == it
        assert np.array_equal(_bits(out), _bits(ref))
    # third and fourth call: NaN/Inf early-out (flags raised), then a clean call right after it — the flags
    # of the previous call must not leak
    g, L0, bounces, n = graphs[1]
    rowptr, col, val = scenes.random_graph(n, avg_deg=4, seed=12)
    big = ctx.create_graph(rowptr, col, val * 1e7)
    ref = torch.empty(n, device=dev); out = torch.empty(n, device=dev)
    it_ref = big.propagate_device(L0.data_ptr(), 20, ref.data_ptr(), want_iters=True)
    it_got = big.propagate_noalloc(L0.data_ptr(), 20, out.data_ptr(), want_iters=True)
    assert it_got == it_ref < 20
    assert np.array_equal(_bits(out), _bits(ref))
    out2 = torch.empty(n, device=dev)
    assert g.propagate_noalloc(L0.data_ptr(), bounces, out2.data_ptr(), want_iters=True) == want[1][1]
    assert np.array_equal(_bits(out2), _bits(want[1][0]))
    # asynchronous form (iters_done = NULL) on the caller's stream
    out3 = torch.empty(n, device=dev)
    s = torch.cuda.Stream(dev)
    s.wait_stream(torch.cuda.current_stream(dev))
    g.propagate_noalloc(L0.data_ptr(), bounces, out3.data_ptr(), stream=s.cuda_stream)
    s.synchronize()
    assert np.array_equal(_bits(out3), _bits(want[1][0]))
    big.close()
    for g, *_ in graphs:
        g.close()


def test_noalloc_refuses_graphs_beyond_the_static_scratch(ctx):
    import torch
    from acceleratedvolrenderer_b200.api import Graph
    n = Graph.NOALLOC_MAX_VERTICES + 1
    rowptr = np.arange(n + 1, dtype=np.int64)           # one self-edge per vertex
    col = np.arange(n, dtype=np.int32); val = np.full(n, 0.5, dtype=np.float32)
    g = ctx.create_graph(rowptr, col, val)
    L0 = torch.ones(n, device="cuda:0"); out = torch.zeros(n, device="cuda:0")
    with pytest.raises(AvrError) as e:
        g.propagate_noalloc(L0.data_ptr(), 2, out.data_ptr())
    assert e.value.status == 4 and "1048576" in str(e.value) and "AVR_NOALLOC_MAX_VERTICES" in str(e.value)
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0                 # refused before touching the output
    # the limit itself is accepted
    g.close()
    n = Graph.NOALLOC_MAX_VERTICES
    g = ctx.create_graph(rowptr[:n + 1], col[:n], val[:n])
    ref = torch.empty(n, device="cuda:0")
    g.propagate_device(L0.data_ptr(), 3, ref.data_ptr(), want_iters=True)
    g.propagate_noalloc(L0.data_ptr(), 3, out.data_ptr(), want_iters=True)
    assert np.array_equal(_bits(out[:n]), _bits(ref))
    assert float(ref[0]) == 1.0 + 0.5 + 0.25 + 0.125
    g.close()
