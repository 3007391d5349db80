"""nes_rank_weights_static: the K6 entry point that allocates nothing (rank scratch in a fixed device array of the library; no counterpart
in the reference -- include/nes_b200.h) against the existing nes_rank_weights, through the C
ABI, on calls in a row that share the scratch.  Ranks are integers and both entry points run the same float64 operations, so the outputs
must be bit-identical.  This file sorts last on purpose (pytest -x at the end of a round must not hide other tests behind it)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ctx():
    import nes_b200
    return nes_b200, nes_b200.NesContext(counts=[4096], seed=5, table_len=1 << 16)


def _fitness(F, seed, ties):
    rs = np.random.RandomState(seed)
    fit = rs.randn(F, 2)
    if ties:                                   # equal fitnesses take the stable order: a third of the entries repeat
        flat = fit.reshape(-1)
        idx = rs.randint(0, flat.size, flat.size // 3)
        flat[idx] = flat[rs.randint(0, flat.size, idx.size)]
    return np.ascontiguousarray(fit, dtype=np.float64)


def test_static_rank_weights_equal_the_allocating_entry_point_on_calls_in_a_row():
    nes_b200, ctx = _ctx()
    lib, h, ptr = ctx.lib, ctx._h, nes_b200._lib.ptr
    dev = ctx.dev
    # the second call is smaller than the first (the scratch still holds the first call's ranks past its own 2F entries), the third
    # has ties, the last is the largest the entry point takes
    cases = [(5000, 1, False), (37, 2, False), (1203, 3, True), (1, 4, False), (nes_b200.NesContext.STATIC_RANK_MAX_F, 5, True)]
    bufs = []
    for F, seed, ties in cases:                # every tensor exists before the calls under test
        fit = _fitness(F, seed, ties)
        bufs.append((fit, torch.from_numpy(fit).to(dev), torch.full((F, 2), -7.0, dtype=torch.float64, device=dev),
                     torch.full((F,), -7.0, dtype=torch.float32, device=dev)))
    torch.cuda.synchronize()
    free_before = None
    for k, (fit, d_fit, d_cen, d_w) in enumerate(bufs):
        F = fit.shape[0]
        if k == 1:
            free_before = torch.cuda.mem_get_info(dev)[0]          # after the first call: module loading is behind us
        nes_b200._lib.check(lib.nes_rank_weights_static(h, ptr(d_fit), ctypes.c_int64(F), ptr(d_cen), ptr(d_w)))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] == free_before, 'device memory changed across calls that must not allocate'
    for fit, d_fit, d_cen, d_w in bufs:
        want_cen, want_w = ctx.rank_weights(fit)                    # the existing entry point (host buffers, context scratch)
        assert np.array_equal(d_cen.cpu().numpy(), want_cen.numpy())
        assert np.array_equal(d_w.cpu().numpy(), want_w.numpy())
    # either output may be left out
    fit, d_fit, d_cen, d_w = bufs[2]
    d_w2 = torch.zeros_like(d_w)
    ctx.rank_weights_static(d_fit, None, d_w2)
    d_c2 = torch.zeros_like(d_cen)
    ctx.rank_weights_static(d_fit, d_c2, None)
    assert torch.equal(d_w2, d_w) and torch.equal(d_c2, d_cen)
    ctx.close()


def test_static_rank_weights_refuses_what_it_cannot_serve():
    nes_b200, ctx = _ctx()
    Fmax = nes_b200.NesContext.STATIC_RANK_MAX_F
    big = torch.zeros((Fmax + 1, 2), dtype=torch.float64, device=ctx.dev)
    w = torch.zeros(Fmax + 1, dtype=torch.float32, device=ctx.dev)
    with pytest.raises(nes_b200.NesError, match='NES_STATIC_RANK_MAX_F'):
        ctx.rank_weights_static(big, None, w)
    host = torch.zeros((8, 2), dtype=torch.float64)
    with pytest.raises(nes_b200.NesError, match='device pointers'):
        ctx.rank_weights_static(host, None, w[:8])
    with pytest.raises(nes_b200.NesError):
        ctx.rank_weights_static(big[:8].contiguous(), None, None)
    ctx.rank_weights_static(big[:0].contiguous(), None, None)       # F = 0: nothing to do
    ctx.close()


def test_static_rank_weights_runs_under_stream_capture():
    """the reason the entry point exists: under a global-mode stream capture any cudaMalloc / cudaFree / synchronising call from this
    thread invalidates the capture; the static entry point is captured and the replayed graph gives the same weights"""
    nes_b200, ctx = _ctx()
    dev = ctx.dev
    F = 2500
    fit = _fitness(F, 9, True)
    d_fit = torch.from_numpy(fit).to(dev)
    d_cen = torch.zeros((F, 2), dtype=torch.float64, device=dev)
    d_w = torch.zeros(F, dtype=torch.float32, device=dev)
    ctx.rank_weights_static(d_fit, d_cen, d_w)                      # warm: symbol address and kernels loaded outside the capture
    want_cen, want_w = ctx.rank_weights(fit)
    d_cen.zero_(); d_w.zero_()
    torch.cuda.synchronize()
    side = torch.cuda.Stream(device=dev)
    graph = torch.cuda.CUDAGraph()
    old = ctx._stream
    try:
        with torch.cuda.graph(graph, stream=side, capture_error_mode='global'):
            ctx.use_stream(torch.cuda.current_stream(dev))
            ctx.rank_weights_static(d_fit, d_cen, d_w)
    finally:
        ctx.use_stream(old)
    torch.cuda.synchronize()
    assert float(d_w.abs().sum()) == 0.0                            # captured, not executed
    graph.replay()
    torch.cuda.synchronize()
    assert np.array_equal(d_cen.cpu().numpy(), want_cen.numpy()) and np.array_equal(d_w.cpu().numpy(), want_w.numpy())
    del graph
    ctx.close()
