"""CPU: the GGUF producer side -- embedding.cpp_b200/convert_to_gguf.py against the REFERENCE'S OWN
converter (models/convert-to-gguf.py:145-346) run here on the same random-init HuggingFace BertModel directory, then the
converted file through the reference's own loader + bert_eval (oracle/_ref).

The reference's script is executed unmodified (runpy) with one shim for its third-party dependency: gguf-py dropped
GGUFWriter.add_source_hf_repo after the version the reference was written against (the wheel here is 0.19)."""
import json
import os
import runpy
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SCRIPT = "/root/reference/models/convert-to-gguf.py"

RUNNER = r'''
import runpy, sys, gguf
if not hasattr(gguf.GGUFWriter, "add_source_hf_repo"):
    gguf.GGUFWriter.add_source_hf_repo = lambda self, v: self.add_string("general.source_hf_repo", v)
sys.argv = ["convert-to-gguf.py"] + sys.argv[1:]
runpy.run_path(%r, run_name="__main__")
''' % REF_SCRIPT


@pytest.fixture(scope="module")
def hf_dir(tmp_path_factory, gio):
    transformers = pytest.importorskip("transformers")
    torch = pytest.importorskip("torch")
    tokenizers = pytest.importorskip("tokenizers")
    d = str(tmp_path_factory.mktemp("hf") / "tinybert")
    os.makedirs(d)
    vocab = gio.synthetic_vocab(300)
    seen = set()
    for i, t in enumerate(vocab):            # HF tokenizers need distinct entries
        if t in seen:
            vocab[i] = f"[dup{i}]"
        seen.add(vocab[i])
    tok = tokenizers.Tokenizer.from_str(gio.stub_tokenizer_json(vocab))
    fast = transformers.PreTrainedTokenizerFast(tokenizer_object=tok, unk_token="[UNK]", sep_token="[SEP]", pad_token="[PAD]",
                                                cls_token="[CLS]", mask_token="[MASK]")
    fast.save_pretrained(d)
    torch.manual_seed(0)
    cfg = transformers.BertConfig(vocab_size=300, hidden_size=64, num_hidden_layers=2, num_attention_heads=2, intermediate_size=128,
                                  max_position_embeddings=64)
    transformers.BertModel(cfg).save_pretrained(d)
    return d


@pytest.mark.parametrize("ftype", [0, 1])
def test_converter_matches_the_reference_converter(ftype, hf_dir, gio, tmp_path):
    if not os.path.exists(REF_SCRIPT):
        pytest.skip("needs /root/reference")
    pytest.importorskip("gguf")
    pytest.importorskip("sentencepiece")     # imported (not used) by the reference script
    conv = __import__("importlib").util.spec_from_file_location("convert_to_gguf", os.path.join(ROOT, "embedding.cpp_b200", "convert_to_gguf.py"))
    mod = __import__("importlib").util.module_from_spec(conv)
    conv.loader.exec_module(mod)
    mine, theirs = str(tmp_path / "mine.gguf"), str(tmp_path / "ref.gguf")
    runner = tmp_path / "run_ref.py"
    runner.write_text(RUNNER)
    r = subprocess.run([sys.executable, str(runner), hf_dir, str(ftype), "--outfile", theirs], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 0 and os.path.exists(theirs), r.stderr[-2000:]
    mod.convert(hf_dir, ftype, mine)
    a, b = gio.read_gguf(mine), gio.read_gguf(theirs)
    # every KV the reference writes, with the same value (tokens, scores, types, blob, hyper-parameters, special ids) ...
    assert set(a.kv) == set(b.kv), (sorted(set(a.kv) ^ set(b.kv)))
    for k, v in b.kv.items():
        w = a.kv[k]
        assert (np.array_equal(v, w) if isinstance(v, np.ndarray) else v == w), k
    # ... and the same tensors: names, order, ggml types (F16 only for 2-D ".weight" when ftype = 1), shapes, bytes
    assert list(a.tensors) == list(b.tensors)
    assert len(b.tensors) == 5 + 16 * 2 and not any("pooler" in k or "position_ids" in k for k in b.tensors)
    for k in b.tensors:
        assert a.tensors[k][0] == b.tensors[k][0] and a.tensors[k][1] == b.tensors[k][1], k
        assert np.array_equal(a.tensors[k][2], b.tensors[k][2]), k
        want_f16 = ftype == 1 and k.endswith(".weight") and len(b.tensors[k][1]) == 2
        assert b.tensors[k][0] == (gio.GGML_F16 if want_f16 else gio.GGML_F32), k


def test_converted_file_loads_in_the_reference_and_the_oracle(hf_dir, gio, orc, tmp_path):
    """the converted file through the consumers: the reference's own loader + bert_eval (oracle/_ref) and the pinned C
    restatement give the same bits; the embedding equals what the HF model itself computes (mean pooling, L2)"""
    import importlib.util
    spec = importlib.util.spec_from_file_location("convert_to_gguf", os.path.join(ROOT, "embedding.cpp_b200", "convert_to_gguf.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    path = mod.convert(hf_dir, 0, str(tmp_path / "m.gguf"))
    toks = np.array([101, 110, 120, 130, 140, 102], np.int32)
    o = orc.Oracle(path).eval(toks)
    if orc.available_reference():
        r = orc.Reference(path, n_threads=2)
        assert np.array_equal(r.eval(toks), o)
        r.close()
    import torch
    import transformers
    m = transformers.BertModel.from_pretrained(hf_dir).eval()
    with torch.no_grad():
        h = m(input_ids=torch.tensor(toks[None].astype(np.int64))).last_hidden_state[0].mean(0).numpy()
    h = h / np.linalg.norm(h)
    # HF uses erf-GELU and fp32 exp, the reference tanh-GELU through f16 tables: close, not equal
    assert float(h @ o) > 0.9999, float(h @ o)
