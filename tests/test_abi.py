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


def test_load_fails_loudly_on_bad_file(binding, tmp_path):
    lib = binding.load_library()
    bad = tmp_path / "bad.gguf"
    bad.write_bytes(b"GGUF" + b"\0" * 40)
    assert not lib.bert_load_from_file(str(bad).encode())
    assert not lib.bert_load_from_file(b"/nonexistent/file.gguf")
