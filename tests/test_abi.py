"""CPU: build/libbert.so loads and exports every symbol include/*.h declares; the loader fails loudly
(NULL + stderr) without a GPU instead of falling back to anything."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header, macro):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = "\n".join(l for l in src.splitlines() if not l.lstrip().startswith("#"))
    return sorted(set(re.findall(macro + r"\s+[\w\s\*]*?\b(\w+)\s*\(", src)))


def test_exports_every_declared_symbol(binding):
    lib = binding.load_library()
    bert = _declared("bert.h", "BERT_API")
    bk = _declared("bk_api.h", "BK_API")
    assert len(bert) >= 12 and len(bk) >= 18, (bert, bk)
    for s in bert + bk:
        assert hasattr(lib, s), s
    assert sorted(bert) == sorted(binding.BERT_SYMBOLS)
    assert sorted(bk) == sorted(binding.BK_SYMBOLS)


def test_no_torch_no_oracle_linked():
    import subprocess
    out = subprocess.run(["ldd", os.path.join(ROOT, "build", "libbert.so")], capture_output=True, text=True).stdout
    assert "torch" not in out and "oracle" not in out and "bert_ref" not in out


def test_quantize_matches_reference_codec(binding, gio, make_model, tmp_path):
    """bert_model_quantize (host tool) vs the NumPy restatement of ggml-quants.c:428-504; both were
    checked byte-identical to the reference's own bert_model_quantize output in the authoring container."""
    lib = binding.load_library()
    src = make_model("mini", "f16", seed=31)
    for ft, name in ((2, "q4_0"), (3, "q4_1")):
        dst = str(tmp_path / f"q_{name}.gguf")
        assert lib.bert_model_quantize(src.encode(), dst.encode(), ft)
        a, b = gio.read_gguf(dst), gio.read_gguf(make_model("mini", name, seed=31))
        assert set(a.tensors) == set(b.tensors)
        for k in b.tensors:
            assert a.tensors[k][0] == b.tensors[k][0], k
            assert np.array_equal(a.tensors[k][2], b.tensors[k][2]), k
        assert a.kv["general.file_type"] == ft and a.kv["general.quantization_version"] == 2
    assert not lib.bert_model_quantize(src.encode(), str(tmp_path / "x.gguf").encode(), 7)


@pytest.mark.parametrize("shape", ["mini", "mini384"])
def test_quantize_pinned_to_the_reference_quantizer(shape, binding, gio, make_model, tmp_path):
    """row f2 pin: the REFERENCE'S OWN bert_model_quantize (bert.cpp:1313-1599, exported by oracle/_ref/libbert_ref.so) run on
    the same F16 file must write the same tensors (types, shapes, bytes, order) and the same two added KVs
    (bert.cpp:1377-1378) as build/libbert.so's bert_model_quantize and as gguf_io's NumPy writer (which makes every Q4 test
    and bench model).  All other KVs must be carried over unchanged."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    if not O.available_reference():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    ref = ctypes.CDLL(O.LIB_REF)
    O.bind_bert_api(ref)
    lib = binding.load_library()
    src = make_model(shape, "f16", seed=33)
    f16 = gio.read_gguf(src)
    for ft, name in ((2, "q4_0"), (3, "q4_1")):
        mine, theirs = str(tmp_path / f"mine_{name}.gguf"), str(tmp_path / f"ref_{name}.gguf")
        assert lib.bert_model_quantize(src.encode(), mine.encode(), ft)
        sys.stdout.flush()
        saved, devnull = os.dup(1), os.open(os.devnull, os.O_WRONLY)       # the reference printf()s a line per tensor
        os.dup2(devnull, 1)
        try:
            assert ref.bert_model_quantize(src.encode(), theirs.encode(), ft)
        finally:
            os.dup2(saved, 1)
            os.close(saved)
            os.close(devnull)
        a, b, c = gio.read_gguf(mine), gio.read_gguf(theirs), gio.read_gguf(make_model(shape, name, seed=33))
        assert list(a.tensors) == list(b.tensors) == list(c.tensors)        # same order in the file
        n_q = 0
        for k in b.tensors:
            for other in (a, c):
                assert other.tensors[k][0] == b.tensors[k][0] and other.tensors[k][1] == b.tensors[k][1], k
                assert np.array_equal(other.tensors[k][2], b.tensors[k][2]), k
            n_q += b.tensors[k][0] == ft
        assert n_q == 3 + 6 * f16.hparams().n_layer                           # 2-D ".weight" tensors only (bert.cpp:1431-1436)
        assert a.kv["general.file_type"] == b.kv["general.file_type"] == ft
        assert a.kv["general.quantization_version"] == b.kv["general.quantization_version"] == 2
        for k, v in f16.kv.items():
            for other in (a, b):
                w = other.kv[k]
                assert (np.array_equal(v, w) if isinstance(v, np.ndarray) else v == w), k
        assert set(a.kv) == set(b.kv)


def test_load_fails_loudly_on_bad_file(binding, tmp_path):
    lib = binding.load_library()
    bad = tmp_path / "bad.gguf"
    bad.write_bytes(b"GGUF" + b"\0" * 40)
    assert not lib.bert_load_from_file(str(bad).encode())
    assert not lib.bert_load_from_file(b"/nonexistent/file.gguf")
