"""CPU: the C++ WordPiece tokenizer (csrc/tokenizer_wp.cpp) against the Python `tokenizers` wheel -- the
same Rust library the reference links (crate tokenizers, pinned 0.13.4 in tokenizers-cpp/rust/Cargo.toml:11;
the wheel here is newer) -- on synthetic WordPiece vocabularies.  Token ids must match exactly.
The reference's own 4 known answers (examples/test_tokenizer.cpp:70-73) are checked in full on the part of the
bert-base-uncased vocabulary that the answers themselves pin (tests/golden/tokenizer_known_answers.json)."""
import json
import os

import numpy as np
import pytest

tokenizers = pytest.importorskip("tokenizers")


def build_vocab():
    toks = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"]
    toks += list("abcdefghijklmnopqrstuvwxyz0123456789")
    toks += ["##" + c for c in "abcdefghijklmnopqrstuvwxyz0123456789"]
    toks += list("!\"#$%&'()*+,-./:;<=>?@[\\]^_`{|}~") + ["¿", "¡", "—", "。", "、", "«", "»", "…"]
    toks += ["the", "quick", "brown", "fox", "jump", "##s", "##ed", "##ing", "over", "lazy", "dog", "hello", "world",
             "quebec", "paris", "embed", "##ding", "##dings", "token", "##izer", "##ization", "don", "t", "un", "##believ",
             "##able", "2023", "20", "##23", "cafe", "naive", "uber", "strasse", "ss", "中", "文", "日", "本", "語",
             "한", "ᄒ", "ᅡ", "ᆫ", "##ᅡ", "##ᆫ", "привет", "мир",
             "##ет", "при", "##в", "ελληνικα", "ς", "σ", "i", "##\u0307",
             "a", "##b", "ab", "##c", "abc", "##abc", "x" * 100, "fi", "ﬁ", "①", "1", "km", "㎞"]
    seen, out = set(), []
    for t in toks:
        if t not in seen:
            seen.add(t)
            out.append(t)
    return out


def make_json(vocab, lowercase=True, strip_accents=None, chinese=True, clean=True, truncation=None):
    added = [{"id": i, "content": t, "single_word": False, "lstrip": False, "rstrip": False, "normalized": False, "special": True}
             for i, t in enumerate(vocab) if t in ("[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]")]
    return json.dumps({
        "version": "1.0", "truncation": truncation, "padding": None, "added_tokens": added,
        "normalizer": {"type": "BertNormalizer", "clean_text": clean, "handle_chinese_chars": chinese,
                       "strip_accents": strip_accents, "lowercase": lowercase},
        "pre_tokenizer": {"type": "BertPreTokenizer"},
        "post_processor": None,
        "decoder": {"type": "WordPiece", "prefix": "##", "cleanup": True},
        "model": {"type": "WordPiece", "unk_token": "[UNK]", "continuing_subword_prefix": "##",
                  "max_input_chars_per_word": 100, "vocab": {t: i for i, t in enumerate(vocab)}},
    }, ensure_ascii=False)


TEXTS = [
    "", " ", "hello world", "Hello, World!", "The quick brown fox jumps over the lazy dog.", "don't", "Québec",
    "Café naïve Über Straße", "embeddings tokenizer tokenization", "unbelievable", "2023 20232023",
    "中文ab日本語", "한", "Привет мир", "ΕΛΛΗΝΙΚΆ Σ",
    "İstanbul", "a\tb\nc\r\nd", "x\u0001y\ufffdz\u200bw\u00adq", "hello[SEP]world [CLS] [MASK]", "[sep] [UNK]",
    "¿qué? ¡sí! — «ok»… 。、", "x" * 100, "x" * 101, "abcabc ab abc abcab zzz",
    "ﬁ ① ㎞", "e\u0301 a\u0323\u0308 s\u0323\u0307", "tab\u00a0nbsp\u3000ideographic\u2028ls",
    "snake_case __init__ a_b", "1+1=2; #hash @user $5 50% ^caret ~tilde `tick`", "  leading and trailing  ",
    "\U0001F600 emoji \U00020000", "for (int i = 0; i < n; i++) { sum += a[i] * b[i]; }",
]


def random_texts(n, seed):
    rng = np.random.default_rng(seed)
    alphabet = list("abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789    .,!?'-_()[]") + \
        ["é", "ü", "ß", "Ç", "ñ", "中", "文", "日", "한", "П", "р", "и", "в",
         "е", "т", "Σ", "ς", "İ", "\u0301", "\u00a0", "\t", "—", "…", "[SEP]", "[UNK]", "##"]
    return ["".join(rng.choice(alphabet, size=int(rng.integers(0, 60)))) for _ in range(n)]


CONFIGS = [dict(), dict(lowercase=False), dict(lowercase=True, strip_accents=False), dict(lowercase=False, strip_accents=True),
           dict(chinese=False), dict(clean=False),
           dict(truncation={"direction": "Right", "max_length": 8, "strategy": "LongestFirst", "stride": 0})]


@pytest.mark.parametrize("cfg", CONFIGS, ids=[str(i) for i in range(len(CONFIGS))])
def test_token_ids_match_hf_tokenizers(cfg, binding):
    vocab = build_vocab()
    js = make_json(vocab, **cfg)
    hf = tokenizers.Tokenizer.from_str(js)
    mine = binding.Tokenizer(js)
    bad = []
    for text in TEXTS + random_texts(400, 7):
        want = hf.encode(text, add_special_tokens=False).ids   # huggingface_tokenizer.cc:28-35
        got = mine.encode(text)
        if got != want:
            bad.append((text, got, want))
    mine.close()
    assert not bad, bad[:5]


def test_gguf_stub_tokenizer_json(binding, gio):
    """the tokenizer.json the fixture writer stores in every synthetic GGUF file"""
    vocab = gio.synthetic_vocab(1000)
    js = gio.stub_tokenizer_json(vocab)
    hf = tokenizers.Tokenizer.from_str(js)
    mine = binding.Tokenizer(js)
    for text in ["the model embeds tokens", "Quebec, Paris!", "embedding embeddings", "zzzz qqq [SEP] the"]:
        assert mine.encode(text) == hf.encode(text, add_special_tokens=False).ids
    mine.close()


def test_reference_known_answers(binding):
    """all four known answers of the reference (examples/test_tokenizer.cpp:70-73), [CLS] / [SEP] wrapping of bert_tokenize
    (bert.cpp:756-779) included.  bert-base-uncased's vocabulary is not available offline; the 94 entries the answers
    themselves pin (tests/golden/gen_tokenizer_known_answers.py) sit at their true ids in a 30522-entry vocabulary."""
    ka = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tokenizer_known_answers.json"), encoding="utf-8"))
    vocab = [f"[unused{i}]" for i in range(30522)]
    for i, t in {0: "[PAD]", 100: "[UNK]", 101: "[CLS]", 102: "[SEP]", 103: "[MASK]"}.items():
        vocab[i] = t
    for piece, i in ka["vocab"].items():
        vocab[i] = piece
    js = make_json(vocab)                                 # bert-base-uncased's normalizer / pre-tokenizer / model settings
    mine = binding.Tokenizer(js)
    hf = tokenizers.Tokenizer.from_str(js)
    assert len(ka["cases"]) == 4
    for c in ka["cases"]:
        got = [101] + mine.encode(c["text"]) + [102]
        assert got == c["ids"], (c["text"][:40], got[:12], c["ids"][:12])
        assert hf.encode(c["text"], add_special_tokens=False).ids == c["ids"][1:-1]   # the stand-in oracle agrees with the reference's answers
    mine.close()
    p = os.environ.get("BERT_BASE_UNCASED_TOKENIZER_JSON")       # the real file, when somebody has it
    if p and os.path.exists(p):
        real = binding.Tokenizer(open(p, encoding="utf-8").read())
        for c in ka["cases"]:
            assert [101] + real.encode(c["text"]) + [102] == c["ids"]
        real.close()
