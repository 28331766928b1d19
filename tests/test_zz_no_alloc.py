"""GPU: bk_batcher_static_scratch (csrc/batcher_static.cu) -- the on-device batcher with its scratch in a fixed __device__ array
inside the library, for callers that cannot call cudaMalloc.  No counterpart in the reference;
the checker is the existing entry point bk_test_batcher (same kernels, cudaMalloc'd scratch) on the same
inputs.  Integer work: bit-exact.  Named test_zz_* so that pytest collects it last.

The caller owns every buffer: device memory comes from torch here (plumbing), passed as raw pointers through the C ABI."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

_i32p = ctypes.POINTER(ctypes.c_int32)
MAX_SEQ = 65536          # BK_STATIC_BATCHER_MAX_SEQ (include/bk_api.h)
E_ARG, E_TOO_LARGE = -1, -2


def _inputs(n_seq, seed):
    rng = np.random.default_rng(seed)
    lens = rng.integers(1, 513, n_seq).astype(np.int32)
    if n_seq >= 1000:
        lens[rng.integers(0, n_seq, n_seq // 3)] = 128      # many ties: the sort must be stable
    toks = rng.integers(0, 30522, int(lens.sum())).astype(np.int32)
    return lens, toks


def _existing(lib, lens, toks):
    n_seq, T = len(lens), len(toks)
    order, cu = np.zeros(n_seq, np.int32), np.zeros(n_seq + 1, np.int32)
    tok, pos = np.zeros(T, np.int32), np.zeros(T, np.int32)
    p = lambda a: a.ctypes.data_as(_i32p)
    assert lib.bk_test_batcher(0, n_seq, p(lens), p(toks), p(order), p(cu), p(tok), p(pos)) == 0
    return order, cu, tok, pos


def _static(lib, torch, lens, toks):
    n_seq, T = len(lens), len(toks)
    dev = torch.device("cuda:0")
    d_lens, d_toks = torch.from_numpy(lens).to(dev), torch.from_numpy(toks).to(dev)
    # outputs pre-filled with a pattern: every element must be written by the call
    outs = [torch.full((n,), -7, dtype=torch.int32, device=dev) for n in (n_seq, n_seq + 1, T, T)]
    torch.cuda.synchronize()
    rc = lib.bk_batcher_static_scratch(0, n_seq, T, d_lens.data_ptr(), d_toks.data_ptr(), *[o.data_ptr() for o in outs], None)
    assert rc == 0
    torch.cuda.synchronize()
    return [o.cpu().numpy() for o in outs]


def test_static_scratch_batcher_equals_existing_entry_point(lib):
    """two calls in a row (a large batch, then a smaller one that finds the first call's leftovers in the shared scratch, then the
    single-sentence path), each compared with bk_test_batcher: order, cu_seqlens, packed ids and positions bit-identical."""
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    for n_seq, seed in ((5000, 1), (1025, 2), (33, 3), (1, 4), (8192, 5)):
        lens, toks = _inputs(n_seq, seed)
        want = _existing(lib, lens, toks)
        got = _static(lib, torch, lens, toks)
        for name, g, w in zip(("order", "cu", "tok", "pos"), got, want):
            assert np.array_equal(g, w), (n_seq, name)


def test_static_scratch_batcher_rejects_oversize(lib):
    """beyond the stated maximum the call fails with its own error code (and launches nothing); bad arguments likewise."""
    torch = pytest.importorskip("torch")
    buf = torch.zeros(16, dtype=torch.int32, device="cuda:0")
    p = buf.data_ptr()
    assert lib.bk_batcher_static_scratch(0, MAX_SEQ + 1, MAX_SEQ + 1, p, p, p, p, p, p, None) == E_TOO_LARGE
    assert lib.bk_batcher_static_scratch(0, 0, 0, p, p, p, p, p, p, None) == E_ARG
    assert lib.bk_batcher_static_scratch(0, 4, 3, p, p, p, p, p, p, None) == E_ARG           # fewer tokens than sentences
    assert lib.bk_batcher_static_scratch(0, 4, 4, p, None, p, p, p, p, None) == E_ARG
    torch.cuda.synchronize()


def test_static_scratch_batcher_at_the_limit(lib):
    """exactly BK_STATIC_BATCHER_MAX_SEQ sentences (short ones, to keep the test small) use the whole scratch array."""
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(9)
    lens = rng.integers(1, 9, MAX_SEQ).astype(np.int32)
    toks = rng.integers(0, 30522, int(lens.sum())).astype(np.int32)
    want = _existing(lib, lens, toks)
    got = _static(lib, torch, lens, toks)
    for name, g, w in zip(("order", "cu", "tok", "pos"), got, want):
        assert np.array_equal(g, w), name
