"""GPU: each sm_100a kernel alone through its C-ABI test entry (bk_test_*), against NumPy / oracle/."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

_u16p = ctypes.POINTER(ctypes.c_uint16)
_f32p = ctypes.POINTER(ctypes.c_float)
_i32p = ctypes.POINTER(ctypes.c_int32)


def h(a):
    return np.ascontiguousarray(a, dtype=np.float16)


def u16(a):
    return a.view(np.uint16).ctypes.data_as(_u16p)


def gelu16(orc_lib, x):
    x = np.ascontiguousarray(x, np.float32)
    y = np.zeros_like(x)
    orc_lib.oracle_gelu(x.size, x.ctypes.data_as(_f32p), y.ctypes.data_as(_f32p))
    return y


def run_gemm(lib, A, Wraw, wtype, N, bias, resid, epi, want_ms=False):
    M, K = A.shape
    out = np.zeros((M, N), np.float16)
    ms = ctypes.c_float(0)
    rc = lib.bk_test_gemm(0, M, N, K, u16(A), Wraw.ctypes.data, wtype, bias.ctypes.data_as(_f32p),
                          u16(resid) if resid is not None else None, epi, u16(out), ctypes.byref(ms) if want_ms else None)
    assert rc == 0
    return out, ms.value


GEMM_SHAPES = [(128, 256, 64), (128, 128, 128), (77, 64, 128), (300, 384, 384), (1000, 2304, 768), (515, 768, 3072),
               (4096, 3072, 768), (129, 1152, 384)]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
@pytest.mark.parametrize("epi", [0, 1, 2])
def test_gemm_f16_weights(lib, binding, orc, make_model, M, N, K, epi):
    rng = np.random.default_rng(M * 7 + N * 3 + K + epi)
    A = h(rng.standard_normal((M, K)) * 0.5)
    W = h(rng.standard_normal((N, K)) * 0.05)
    bias = (rng.standard_normal(N) * 0.1).astype(np.float32)
    resid = h(rng.standard_normal((M, N))) if epi == 1 else None
    out, _ = run_gemm(lib, A, W, binding.BK_F16, N, bias, resid, epi)
    ref = A.astype(np.float32) @ W.astype(np.float32).T + bias
    if epi == 1:
        ref = ref + resid.astype(np.float32)
    if epi == 2:
        o = orc.Oracle(make_model("tiny", "f32", 13, 300))
        ref = gelu16(o.lib, ref)
    err = np.abs(out.astype(np.float32) - ref)
    tol = 2e-3 * np.abs(ref) + 2e-3
    if epi == 2:  # a 1-ulp f16 difference of the GELU input can move the output by one f16 ulp
        assert (err <= tol).mean() > 0.999 and err.max() < 2e-2
    else:
        assert (err <= tol).all(), (float(err.max()), np.unravel_index(err.argmax(), err.shape))


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("bn", [64, 128, 192, 256])
@pytest.mark.parametrize("epi", [0, 1])
def test_gemm_forced_tile_width(lib, binding, bn, epi, cg):
    """every tile shape of the persistent kernel (the engine picks one per problem): width 64..256, one CTA
    per 128-row tile or a CTA pair (cta_group::2) per 256-row tile; several tiles per CTA"""
    M, N, K = 128 * 150 + 37, 768, 256      # > 148 tiles for every width: exercises staging-box reuse across tiles
    rng = np.random.default_rng(bn + epi + cg)
    A = h(rng.standard_normal((M, K)) * 0.5)
    W = h(rng.standard_normal((N, K)) * 0.05)
    bias = (rng.standard_normal(N) * 0.1).astype(np.float32)
    resid = h(rng.standard_normal((M, N))) if epi == 1 else None
    out, _ = run_gemm(lib, A, W, binding.BK_F16, N, bias, resid, epi | (bn << 8) | (cg << 20))
    ref = A.astype(np.float32) @ W.astype(np.float32).T + bias
    if epi == 1:
        ref = ref + resid.astype(np.float32)
    err = np.abs(out.astype(np.float32) - ref)
    assert (err <= 2e-3 * np.abs(ref) + 2e-3).all(), (float(err.max()), np.unravel_index(err.argmax(), err.shape))


def test_gelu_all_f16_inputs(lib, orc, make_model):
    """the epilogue's exp2/rcp GELU against the reference's f16 table (ggml.c:1433-1441) on every finite f16"""
    x = np.arange(65536, dtype=np.uint32).astype(np.uint16).view(np.float16)
    x = np.ascontiguousarray(x[np.isfinite(x)].astype(np.float32))
    y = np.zeros_like(x)
    assert lib.bk_test_gelu(0, x.size, x.ctypes.data_as(_f32p), y.ctypes.data_as(_f32p)) == 0
    o = orc.Oracle(make_model("tiny", "f32", 13, 300))
    want = gelu16(o.lib, x)
    bad = y != want
    # differences are 1-ulp f16 near-ties only, and rare
    assert bad.sum() <= 32, int(bad.sum())
    ulp = np.abs(y[bad].astype(np.float16).view(np.int16).astype(np.int32) - want[bad].astype(np.float16).view(np.int16).astype(np.int32))
    assert (ulp <= 1).all()


@pytest.mark.parametrize("wname,wtype", [("q4_0", 2), ("q4_1", 3), ("f32", 0)])
def test_gemm_other_weight_storage(lib, gio, wname, wtype):
    """Q4_0/Q4_1/F32 weight storage: the product decodes GGUF blocks itself; compare with the NumPy decode."""
    M, N, K = 200, 384, 768
    rng = np.random.default_rng(3)
    A = h(rng.standard_normal((M, K)) * 0.5)
    Wf = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    if wtype == 0:
        raw, Wd = Wf, Wf.astype(np.float16).astype(np.float32)  # F32 weights are used as f16 operands
    else:
        raw = (gio.quantize_q4_0 if wtype == 2 else gio.quantize_q4_1)(Wf)
        Wd = gio.dequantize(raw, wtype, (N, K)).astype(np.float16).astype(np.float32)
    bias = np.zeros(N, np.float32)
    out, _ = run_gemm(lib, A, np.ascontiguousarray(raw), wtype, N, bias, None, 0)
    ref = A.astype(np.float32) @ Wd.T
    err = np.abs(out.astype(np.float32) - ref)
    assert (err <= 2e-3 * np.abs(ref) + 2e-3).all(), float(err.max())


@pytest.mark.parametrize("wname,wtype", [("q4_0", 2), ("q4_1", 3)])
@pytest.mark.parametrize("M,N,K,epi", [(200, 384, 768, 0), (1000, 2304, 768, 0), (333, 768, 3072, 1), (130, 96, 128, 2), (5, 3072, 768, 2)])
def test_gemm_q4_decoded_in_shared_memory(lib, gio, wname, wtype, M, N, K, epi):
    """W kept as GGUF Q4 blocks in device memory and decoded by the GEMM's decode warps: only packed bytes cross
    HBM.  Must equal the path that expands W to f16 at load bit for bit (same values, same MMA K order)."""
    rng = np.random.default_rng(M + N + K + wtype)
    A = h(rng.standard_normal((M, K)) * 0.5)
    Wf = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    raw = np.ascontiguousarray((gio.quantize_q4_0 if wtype == 2 else gio.quantize_q4_1)(Wf))
    Wd = gio.dequantize(raw, wtype, (N, K)).astype(np.float16).astype(np.float32)
    bias = (rng.standard_normal(N) * 0.1).astype(np.float32)
    resid = h(rng.standard_normal((M, N))) if epi == 1 else None
    at_load, _ = run_gemm(lib, A, raw, wtype, N, bias, resid, epi)
    in_smem, _ = run_gemm(lib, A, raw, wtype, N, bias, resid, epi | (1 << 28))
    assert np.array_equal(at_load.view(np.uint16), in_smem.view(np.uint16))
    if epi != 2:
        ref = A.astype(np.float32) @ Wd.T + bias + (resid.astype(np.float32) if epi == 1 else 0)
        err = np.abs(in_smem.astype(np.float32) - ref)
        assert (err <= 2e-3 * np.abs(ref) + 2e-3).all(), float(err.max())


@pytest.mark.parametrize("wname,wtype", [("q4_0", 2), ("q4_1", 3)])
@pytest.mark.parametrize("M,N,K,epi", [(1000, 2304, 768, 0), (300, 3072, 768, 2), (5000, 1152, 384, 0), (777, 3072, 1024, 3), (129, 128, 64, 0),
                                       (1, 256, 128, 2), (20000, 2304, 768, 0), (4096, 4096, 1024, 4), (257, 1536, 384, 3)])
def test_gemm_q4_weight_stationary(lib, gio, wname, wtype, M, N, K, epi):
    """gemm_ws.cu: the packed GGUF blocks are the only copy of W; every CTA pair decodes its 128-column tile once into shared
    memory and streams all row tiles past it (QKV and FFN-up of Q4 models above the latency regime).  Must equal, bit for bit,
    the kernel fed from the f16 copy expanded at load (same decoder values, same MMA K order, same epilogue arithmetic) --
    every epilogue these projections use: bias, bias + GELU, bias + GELU + Q8_0 / Q8_1 rounding."""
    rng = np.random.default_rng(M + N + K + wtype + epi)
    A = h(rng.standard_normal((M, K)) * 0.5)
    Wf = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    raw = np.ascontiguousarray((gio.quantize_q4_0 if wtype == 2 else gio.quantize_q4_1)(Wf))
    bias = (rng.standard_normal(N) * 0.1).astype(np.float32)
    at_load, _ = run_gemm(lib, A, raw, wtype, N, bias, None, epi)
    # the automatic tile width and every width whose resident slice fits: (227 KB - 34.5 KB of boxes / barriers - 3 A stages) of smem
    widths = [0] + [bn for bn in (128, 192, 256) if N % bn == 0 and (K // 64) * (bn // 2) * 128 <= 227 * 1024 - 34304 - 3 * 16384 and 74 // (N // bn) >= 1]
    assert len(widths) >= 2
    for bn in widths:
        ws, _ = run_gemm(lib, A, raw, wtype, N, bias, None, epi | (1 << 29) | (bn << 8))
        assert np.array_equal(at_load.view(np.uint16), ws.view(np.uint16)), (bn, int((at_load.view(np.uint16) != ws.view(np.uint16)).sum()))
    if epi == 0:
        Wd = gio.dequantize(raw, wtype, (N, K)).astype(np.float16).astype(np.float32)
        ref = A.astype(np.float32) @ Wd.T + bias
        err = np.abs(ws.astype(np.float32) - ref)
        assert (err <= 2e-3 * np.abs(ref) + 2e-3).all(), float(err.max())


@pytest.mark.parametrize("qepi,fn", [(3, "oracle_quant_q8_0_roundtrip"), (4, "oracle_quant_q8_1_roundtrip")])
def test_gemm_gelu_q8_epilogue(lib, binding, orc, make_model, qepi, fn):
    M, N, K = 256, 512, 128
    rng = np.random.default_rng(9)
    A = h(rng.standard_normal((M, K)) * 0.5)
    W = h(rng.standard_normal((N, K)) * 0.1)
    bias = (rng.standard_normal(N) * 0.1).astype(np.float32)
    plain, _ = run_gemm(lib, A, W, binding.BK_F16, N, bias, None, 2)
    q, _ = run_gemm(lib, A, W, binding.BK_F16, N, bias, None, qepi)
    o = orc.Oracle(make_model("tiny", "f32", 13, 300))
    src = np.ascontiguousarray(plain.astype(np.float32))
    want = np.zeros_like(src)
    getattr(o.lib, fn)(src.ctypes.data_as(_f32p), want.ctypes.data_as(_f32p), src.size)
    # the kernel quantises the same f16-valued GELU outputs -> same integers; only the final f16 store rounds
    assert np.allclose(q.astype(np.float32), want, rtol=1e-3, atol=1e-6)


@pytest.mark.parametrize("T,d", [(1, 64), (37, 128), (1000, 384), (513, 768), (300, 1024)])
@pytest.mark.parametrize("qmode", [0, 1, 2])
def test_add_ln(lib, orc, make_model, T, d, qmode):
    rng = np.random.default_rng(T + d)
    pre = h(rng.standard_normal((T, d)) * 2 + 0.3)
    gamma = (1 + 0.1 * rng.standard_normal(d)).astype(np.float32)
    beta = (0.1 * rng.standard_normal(d)).astype(np.float32)
    y = np.zeros((T, d), np.float16)
    yq = np.zeros((T, d), np.float16)
    assert lib.bk_test_add_ln(0, T, d, 1e-12, u16(pre), gamma.ctypes.data_as(_f32p), beta.ctypes.data_as(_f32p), qmode, u16(y), u16(yq)) == 0
    o = orc.Oracle(make_model("tiny", "f32", 13, 300))
    o.hp.eps = 1e-12
    ref = o.layernorm(pre.astype(np.float32), gamma, beta)
    assert np.allclose(y.astype(np.float32), ref, rtol=1e-3, atol=1e-3)
    if qmode:
        want = np.zeros_like(ref)
        fn = o.lib.oracle_quant_q8_0_roundtrip if qmode == 1 else o.lib.oracle_quant_q8_1_roundtrip
        fn(ref.ctypes.data_as(_f32p), want.ctypes.data_as(_f32p), ref.size)
        err = np.abs(yq.astype(np.float32) - want)
        step = np.abs(ref).reshape(T, d // 32, 32).max(-1, keepdims=True) / 127.0
        step = np.broadcast_to(step, (T, d // 32, 32)).reshape(T, d)
        # fp32-vs-double LN sums can flip a rounding decision for values sitting on a .5 boundary: count them, never more than one step
        flips = int((err > 1e-3 * np.abs(want) + 1e-6).sum())
        print(f"\nadd_ln T={T} d={d} qmode={qmode}: {flips} of {err.size} Q8 roundings flipped by one step ({flips / err.size:.2e})")
        assert flips <= 0.005 * err.size
        assert (err <= step * 1.01 + 1e-3).all()


@pytest.mark.parametrize("M,N,K", [(256, 768, 768), (1000, 768, 768), (37, 768, 3072), (5000, 384, 384), (300, 384, 1536),
                                   (777, 1024, 1024), (256, 1024, 4096), (20000, 768, 768)])
@pytest.mark.parametrize("qmode", [0, 1, 2])
def test_gemm_ln_fused(lib, orc, make_model, M, N, K, qmode):
    """projection + bias + residual + LayerNorm (+ Q8 rounding) in ONE kernel (gemm_ln.cu: clusters of CTA pairs hold
    whole rows in TMEM, statistics exchanged through distributed shared memory) against fp32 NumPy + the oracle's
    LayerNorm (ggml.c:8976-9004) and Q8 round trip (ggml-quants.c:702-747, 944-985)."""
    if qmode and M > 5000:
        pytest.skip("large case once")
    rng = np.random.default_rng(M + N + K + qmode)
    A = h(rng.standard_normal((M, K)) * 0.5)
    W = h(rng.standard_normal((N, K)) * 0.05)
    bias = (rng.standard_normal(N) * 0.1).astype(np.float32)
    resid = h(rng.standard_normal((M, N)) + 0.3)
    gamma = (1 + 0.1 * rng.standard_normal(N)).astype(np.float32)
    beta = (0.1 * rng.standard_normal(N)).astype(np.float32)
    y = np.zeros((M, N), np.float16)
    yq = np.zeros((M, N), np.float16)
    assert lib.bk_test_gemm_ln(0, M, N, K, u16(A), u16(W), bias.ctypes.data_as(_f32p), u16(resid), gamma.ctypes.data_as(_f32p),
                               beta.ctypes.data_as(_f32p), 1e-12, qmode, u16(y), u16(yq), None) == 0
    pre = A.astype(np.float32) @ W.astype(np.float32).T + bias + resid.astype(np.float32)
    o = orc.Oracle(make_model("tiny", "f32", 13, 300))
    o.hp.eps = 1e-12
    ref = o.layernorm(np.ascontiguousarray(pre), gamma, beta)
    err = np.abs(y.astype(np.float32) - ref)
    assert (err <= 2e-3 * np.abs(ref) + 2e-3).all(), (float(err.max()), np.argwhere(err > 2e-3 * np.abs(ref) + 2e-3)[:5])
    if qmode:
        got = y.astype(np.float32)   # the kernel quantises its own fp32 LN value; compare against the round trip of y
        want = np.zeros_like(got)
        fn = o.lib.oracle_quant_q8_0_roundtrip if qmode == 1 else o.lib.oracle_quant_q8_1_roundtrip
        fn(np.ascontiguousarray(ref).ctypes.data_as(_f32p), want.ctypes.data_as(_f32p), ref.size)
        step = np.abs(ref).reshape(M, N // 32, 32).max(-1, keepdims=True) / 127.0
        step = np.broadcast_to(step, (M, N // 32, 32)).reshape(M, N)
        errq = np.abs(yq.astype(np.float32) - want)
        # fp32 GEMM-order differences can flip a .5 rounding: count them, never more than one step
        flips = int((errq > 1e-3 * np.abs(want) + 1e-5).sum())
        print(f"\ngemm_ln M={M} N={N} K={K} qmode={qmode}: {flips} of {errq.size} Q8 roundings flipped by one step ({flips / errq.size:.2e})")
        assert flips <= 0.01 * errq.size
        assert (errq <= step * 1.01 + 2e-3).all()


@pytest.mark.parametrize("d,H,lens", [(128, 4, [5, 128, 1, 33]), (256, 4, [64, 17, 200]), (768, 12, [128, 128, 16]),
                                      (384, 12, [512, 2]), (256, 4, [512, 300, 129])])
def test_attention(lib, orc, make_model, d, H, lens):
    rng = np.random.default_rng(d + sum(lens))
    T = sum(lens)
    qkv = h(rng.standard_normal((T, 3 * d)) * 1.2)
    out = np.zeros((T, d), np.float16)
    L = np.array(lens, np.int32)
    assert lib.bk_test_attention(0, len(lens), L.ctypes.data_as(_i32p), d, H, u16(qkv), 0, u16(out)) == 0
    o = orc.Oracle(make_model("tiny", "f32", 13, 300))
    off = 0
    for n in lens:
        blk = qkv[off:off + n].astype(np.float32)
        ref = o.attention(blk[:, :d], blk[:, d:2 * d], blk[:, 2 * d:], H)
        got = out[off:off + n].astype(np.float32)
        assert np.allclose(got, ref, rtol=2e-3, atol=2e-3), (n, float(np.abs(got - ref).max()))
        off += n


@pytest.mark.parametrize("d,H,n_seq,qmode", [(768, 12, 90, 0), (384, 12, 100, 0), (768, 12, 61, 1), (128, 2, 700, 2)])
def test_attention_persistent_kernel(lib, orc, make_model, monkeypatch, d, H, n_seq, qmode):
    """launches whose longest sentence has <= 128 tokens run persistent CTAs that walk over several (sentence, head)
    items with two of them in flight (attention_persist.cuh): more items than CTAs, every key-chunk / slice boundary,
    against the oracle and bit for bit against the one-CTA-per-item kernel"""
    rng = np.random.default_rng(d + n_seq)
    lens = [1, 15, 16, 17, 32, 33, 63, 64, 65, 96, 97, 127, 128] + [int(v) for v in rng.integers(1, 129, n_seq - 13)]
    lens = [lens[i] for i in rng.permutation(len(lens))]
    T = sum(lens)
    qkv = h(rng.standard_normal((T, 3 * d)) * 1.2)
    L = np.array(lens, np.int32)
    outs = []
    for persist, loop in (("128", "1"), ("0", "0"), ("0", "1")):   # persistent two-stage | one CTA per item | looping CTAs (default)
        monkeypatch.setenv("BERT_B200_ATT_PERSIST", persist)
        monkeypatch.setenv("BERT_B200_ATT_LOOP", loop)
        out = np.zeros((T, d), np.float16)
        assert lib.bk_test_attention(0, len(lens), L.ctypes.data_as(_i32p), d, H, u16(qkv), qmode, u16(out)) == 0
        outs.append(out)
    assert np.array_equal(outs[0].view(np.uint16), outs[1].view(np.uint16))
    assert np.array_equal(outs[2].view(np.uint16), outs[1].view(np.uint16))
    if qmode == 0:
        o = orc.Oracle(make_model("tiny", "f32", 13, 300))
        off = 0
        for i, n in enumerate(lens):
            if i < 40 or n in (1, 16, 17, 64, 65, 128):
                blk = qkv[off:off + n].astype(np.float32)
                ref = o.attention(blk[:, :d], blk[:, d:2 * d], blk[:, 2 * d:], H)
                got = outs[0][off:off + n].astype(np.float32)
                assert np.allclose(got, ref, rtol=2e-3, atol=2e-3), (i, n, float(np.abs(got - ref).max()))
            off += n


@pytest.mark.parametrize("d,H,qmode,reps", [(768, 12, 0, 1), (384, 12, 0, 1), (256, 4, 1, 1), (128, 4, 2, 1),
                                             (256, 4, 1, 5), (128, 4, 2, 7)])   # reps > 1: several items per persistent CTA with the Q8 epilogue
def test_attention_streaming_kernel(lib, orc, make_model, monkeypatch, d, H, qmode, reps):
    """launches with sentences of more than 128 tokens stream the scores through a two-slot TMEM ring and compute them
    twice (attention_stream.cuh): every chunk / query-block boundary, short sentences inside a long launch, against the
    oracle and bit for bit against the kernel that keeps the whole score block resident"""
    rng = np.random.default_rng(d + qmode)
    lens = [512, 511, 385, 384, 383, 300, 257, 256, 255, 200, 129, 128, 127, 100, 64, 33, 17, 16, 5, 1] + [int(v) for v in rng.integers(1, 513, 12)]
    for _ in range(reps - 1):
        lens += [int(v) for v in rng.integers(1, 513, 32)]
    T = sum(lens)
    qkv = h(rng.standard_normal((T, 3 * d)) * 1.2)
    L = np.array(lens, np.int32)
    outs = []
    # exp / service warps (attention_xs.cuh) / P in tensor memory (attention_pt.cuh) / two blocks in flight (attention_duo.cuh) /
    # streaming through shared memory / resident scores
    for stream, xs, pt, duo in (("1", "1", "0", "0"), ("1", "0", "1", "0"), ("1", "0", "0", "1"), ("1", "0", "0", "0"), ("0", "0", "0", "0")):
        monkeypatch.setenv("BERT_B200_ATT_STREAM", stream)
        monkeypatch.setenv("BERT_B200_ATT_XS", xs)
        monkeypatch.setenv("BERT_B200_ATT_PT", pt)
        monkeypatch.setenv("BERT_B200_ATT_DUO", duo)
        out = np.zeros((T, d), np.float16)
        assert lib.bk_test_attention(0, len(lens), L.ctypes.data_as(_i32p), d, H, u16(qkv), qmode, u16(out)) == 0
        outs.append(out)
    for k in range(4):
        assert np.array_equal(outs[k].view(np.uint16), outs[4].view(np.uint16)), k
    if qmode == 0:
        o = orc.Oracle(make_model("tiny", "f32", 13, 300))
        off = 0
        for i, n in enumerate(lens):
            blk = qkv[off:off + n].astype(np.float32)
            ref = o.attention(blk[:, :d], blk[:, d:2 * d], blk[:, 2 * d:], H)
            got = outs[0][off:off + n].astype(np.float32)
            assert np.allclose(got, ref, rtol=2e-3, atol=2e-3), (i, n, float(np.abs(got - ref).max()))
            off += n


@pytest.mark.parametrize("d,lens", [(64, [1, 2, 64]), (384, [128, 7]), (768, [128] * 5 + [512]), (1024, [33])])
def test_pool_norm(lib, orc, make_model, d, lens):
    rng = np.random.default_rng(d)
    T = sum(lens)
    x = h(rng.standard_normal((T, d)))
    out = np.zeros((len(lens), d), np.float32)
    L = np.array(lens, np.int32)
    assert lib.bk_test_pool(0, len(lens), L.ctypes.data_as(_i32p), d, u16(x), out.ctypes.data_as(_f32p)) == 0
    o = orc.Oracle(make_model("tiny", "f32", 13, 300))
    off = 0
    for i, n in enumerate(lens):
        ref = o.pool(x[off:off + n].astype(np.float32))
        assert np.allclose(out[i], ref, rtol=1e-5, atol=1e-6)
        off += n


@pytest.mark.parametrize("n_seq", [1, 2, 31, 32, 33, 1000, 1024, 1025, 5000, 8192])
def test_on_device_batcher(lib, n_seq):
    """batch_sort + batch_pack (rowwise.cu): stable sort by length, longest first; packed ids, positions, cu_seqlens and the
    sorted-position -> caller-index map must equal NumPy's stable argsort.  Integer work: bit-exact."""
    rng = np.random.default_rng(n_seq)
    lens = rng.integers(1, 513, n_seq).astype(np.int32)
    if n_seq >= 1000:
        lens[rng.integers(0, n_seq, n_seq // 3)] = 128          # many ties: stability matters
    toks = rng.integers(0, 30522, int(lens.sum())).astype(np.int32)
    T = int(lens.sum())
    order, cu = np.zeros(n_seq, np.int32), np.zeros(n_seq + 1, np.int32)
    tok, pos = np.zeros(T, np.int32), np.zeros(T, np.int32)
    p = lambda a: a.ctypes.data_as(_i32p)
    assert lib.bk_test_batcher(0, n_seq, p(lens), p(toks), p(order), p(cu), p(tok), p(pos)) == 0
    want_order = np.argsort(-lens.astype(np.int64), kind="stable").astype(np.int32)
    assert np.array_equal(order, want_order)
    want_cu = np.concatenate([[0], np.cumsum(lens[want_order])]).astype(np.int32)
    assert np.array_equal(cu, want_cu)
    src = np.concatenate([[0], np.cumsum(lens)])[:-1]
    want_tok = np.concatenate([toks[src[i]:src[i] + lens[i]] for i in want_order])
    want_pos = np.concatenate([np.arange(lens[i]) for i in want_order]).astype(np.int32)
    assert np.array_equal(tok, want_tok) and np.array_equal(pos, want_pos)


@pytest.mark.parametrize("ftype", ["f32", "f16", "q4_0", "q4_1"])
@pytest.mark.parametrize("shape", ["mini", "mini384", "bert-base"])
def test_embed_ln(lib, binding, gio, orc, make_model, shape, ftype):
    """embeddings gather (three tables in their GGUF storage type, decoded in registers) + LayerNorm alone against the oracle's
    embed (bert.cpp:868-898: (type[0] + word[tok]) + pos[i], ggml_get_rows decode, LayerNorm with double sums); for Q4
    files also the Q8-rounded copy against the oracle's round trip of ITS OWN LayerNorm output (flipped roundings counted)."""
    path = make_model(shape, ftype, seed=51, n_vocab=600)
    f = gio.read_gguf(path)
    hp = f.hparams()
    o = orc.Oracle(path)
    rng = np.random.default_rng(7)
    tokens = rng.integers(0, hp.n_vocab, 300).astype(np.int32)
    want = o.embed(tokens)                                    # positions 0..T-1
    pos = np.arange(len(tokens), dtype=np.int32)
    d, T = hp.n_embd, len(tokens)
    t = {k: np.ascontiguousarray(f.tensors[k][2]) for k in ("embeddings.word_embeddings.weight", "embeddings.token_type_embeddings.weight",
                                                              "embeddings.position_embeddings.weight")}
    gamma, beta = f.f32("embeddings.LayerNorm.weight"), f.f32("embeddings.LayerNorm.bias")
    qmode = {"q4_0": 1, "q4_1": 2}.get(ftype, 0)
    x, xq = np.zeros((T, d), np.float16), np.zeros((T, d), np.float16)
    wtype = f.tensors["embeddings.word_embeddings.weight"][0]
    assert lib.bk_test_embed_ln(0, T, d, hp.n_vocab, hp.n_max_tokens, wtype, t["embeddings.word_embeddings.weight"].ctypes.data,
                                t["embeddings.token_type_embeddings.weight"].ctypes.data, t["embeddings.position_embeddings.weight"].ctypes.data,
                                gamma.ctypes.data_as(_f32p), beta.ctypes.data_as(_f32p), hp.eps, tokens.ctypes.data_as(_i32p),
                                pos.ctypes.data_as(_i32p), qmode, u16(x), u16(xq)) == 0
    # x is the f16 rounding of the fp32 LayerNorm output: half an f16 ulp + fp32-vs-double statistics
    err = np.abs(x.astype(np.float32) - want)
    assert (err <= 6e-4 * np.abs(want) + 1e-6).all(), float((err / (np.abs(want) + 1e-6)).max())
    if qmode:
        rt = np.zeros_like(want)
        fn = o.lib.oracle_quant_q8_0_roundtrip if qmode == 1 else o.lib.oracle_quant_q8_1_roundtrip
        fn(np.ascontiguousarray(want).ctypes.data_as(_f32p), rt.ctypes.data_as(_f32p), want.size)
        errq = np.abs(xq.astype(np.float32) - rt)
        step = np.broadcast_to(np.abs(want).reshape(T, d // 32, 32).max(-1, keepdims=True) / 127.0, (T, d // 32, 32)).reshape(T, d)
        flips = int((errq > 1e-3 * np.abs(rt) + 1e-6).sum())
        print(f"\nembed_ln {shape} {ftype}: {flips} of {errq.size} Q8 roundings flipped by one step")
        assert flips <= 0.002 * errq.size and (errq <= step * 1.01 + 1e-3).all()


@pytest.mark.parametrize("d,H,n_seq,qmode", [(768, 12, 700, 0), (768, 12, 100, 1), (128, 4, 3000, 0), (1024, 16, 64, 2), (256, 4, 13, 0)])
def test_attention_loop_kernel(lib, monkeypatch, d, H, n_seq, qmode):
    """attention_loop.cuh: the per-item kernel with the CTA kept on its SM (four per SM, each walking over several
    (sentence, head) items; the next item's loads requested when the current P.V MMAs retire).  More items than CTAs, fewer
    items than CTAs, every key-chunk / slice boundary: bit for bit the one-CTA-per-item kernel."""
    rng = np.random.default_rng(d + n_seq + qmode)
    lens = [1, 15, 16, 17, 32, 33, 63, 64, 65, 96, 97, 127, 128] + [int(v) for v in rng.integers(1, 129, n_seq - 13)]
    lens = [lens[i] for i in rng.permutation(len(lens))]
    T = sum(lens)
    qkv = h(rng.standard_normal((T, 3 * d)) * 1.2)
    L = np.array(lens, np.int32)
    monkeypatch.setenv("BERT_B200_ATT_PERSIST", "0")
    outs = []
    for loop in ("1", "0"):
        monkeypatch.setenv("BERT_B200_ATT_LOOP", loop)
        out = np.zeros((T, d), np.float16)
        assert lib.bk_test_attention(0, len(lens), L.ctypes.data_as(_i32p), d, H, u16(qkv), qmode, u16(out)) == 0
        outs.append(out)
    assert np.array_equal(outs[0].view(np.uint16), outs[1].view(np.uint16)), int((outs[0].view(np.uint16) != outs[1].view(np.uint16)).sum())


@pytest.mark.parametrize("d,H,n64,n32,n16,qmode", [(768, 12, 37, 45, 203, 0), (768, 12, 1, 0, 0, 1), (128, 4, 0, 3, 1000, 0), (1024, 16, 64, 64, 64, 2),
                                                   (256, 4, 2, 4, 8, 0), (384, 12, 0, 0, 17, 1), (768, 12, 300, 0, 0, 0)])
def test_attention_packed_short_sentences(lib, monkeypatch, d, H, n64, n32, n16, qmode):
    """attention_pack.cuh: 2 / 4 / 8 sentences of 33..64 / 17..32 / 1..16 tokens share one 128-row tile (slots on 16-key
    boundaries, the other slots' P entries exact zeros).  Every class boundary, partly filled last tiles, more tiles than CTAs:
    equal to the one-sentence-per-tile kernel bit for bit (a signed zero apart)."""
    rng = np.random.default_rng(d + n64 + 3 * n32 + 7 * n16 + qmode)
    lens = ([64, 33][:n64] + [int(v) for v in rng.integers(33, 65, max(0, n64 - 2))] + [32, 17][:n32] + [int(v) for v in rng.integers(17, 33, max(0, n32 - 2))] +
            [16, 1][:n16] + [int(v) for v in rng.integers(1, 17, max(0, n16 - 2))])
    assert len(lens) == n64 + n32 + n16
    T = sum(lens)
    qkv = h(rng.standard_normal((T, 3 * d)) * 1.2)
    L = np.array(lens, np.int32)
    packed = np.zeros((T, d), np.float16)
    assert lib.bk_test_attention_packed(0, n64, n32, n16, L.ctypes.data_as(_i32p), d, H, u16(qkv), qmode, u16(packed)) == 0
    monkeypatch.setenv("BERT_B200_ATT_PERSIST", "0")
    monkeypatch.setenv("BERT_B200_ATT_LOOP", "1")
    alone = np.zeros((T, d), np.float16)
    assert lib.bk_test_attention(0, len(lens), L.ctypes.data_as(_i32p), d, H, u16(qkv), qmode, u16(alone)) == 0
    assert np.isfinite(packed.astype(np.float32)).all()
    assert np.array_equal(packed, alone), int((packed != alone).sum())     # float compare: -0 == +0
