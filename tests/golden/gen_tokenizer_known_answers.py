"""Transcribes the reference's four tokenizer known answers (examples/test_tokenizer.cpp:70-73: text -> ids of
bert-base-uncased, [CLS] = 101 and [SEP] = 102 included) into tests/golden/tokenizer_known_answers.json, together with
the vocabulary entries those answers imply.

The bert-base-uncased vocabulary is not available offline, but the answers pin 94 of its entries: the text is normalised
and pre-tokenised (BertNormalizer + BertPreTokenizer, Python `tokenizers` wheel), every word is matched with its id;
the five words that the reference splits into several pieces (7 ids for 3 words in vector 2, 205 ids for 197 words in
vector 4) are split as bert-base-uncased does: sy ##om ##me / ta ##alla / tan ##aan / evaluate ##post ##fi ##x /
is ##di ##git.  The script checks that piece <-> id is a bijection over all four vectors (e.g. "is" = 2003 both as a
word in vector 3 and as the first piece of "isdigit" in vector 4), which is what makes the reconstruction trustworthy.

    python tests/golden/gen_tokenizer_known_answers.py        (authoring container only: reads /root/reference)"""
import json
import os
import re

from tokenizers import normalizers, pre_tokenizers

HERE = os.path.dirname(os.path.abspath(__file__))
SPLITS = {"syomme": ["sy", "##om", "##me"], "taalla": ["ta", "##alla"], "tanaan": ["tan", "##aan"],
          "evaluatepostfix": ["evaluate", "##post", "##fi", "##x"], "isdigit": ["is", "##di", "##git"]}


def main():
    src = open("/root/reference/examples/test_tokenizer.cpp", encoding="utf-8").read()
    pat = re.compile(r'tokenizer_test\(bctx,\s*"((?:[^"\\]|\\.)*)",\s*\{([^}]*)\}\);')
    cases = []
    for m in pat.finditer(src):
        text = m.group(1).encode("utf-8").decode("unicode_escape").encode("latin-1").decode("utf-8")   # C escapes (\t, \")
        cases.append({"text": text, "ids": [int(x) for x in m.group(2).split(",")]})
    assert len(cases) == 4
    norm = normalizers.BertNormalizer(clean_text=True, handle_chinese_chars=True, strip_accents=None, lowercase=True)
    pre = pre_tokenizers.BertPreTokenizer()
    vocab = {}
    for c in cases:
        pieces = []
        for w, _ in pre.pre_tokenize_str(norm.normalize_str(c["text"])):
            pieces += SPLITS.get(w, [w])
        body = c["ids"][1:-1]
        assert c["ids"][0] == 101 and c["ids"][-1] == 102 and len(pieces) == len(body), c["text"][:30]
        for p, i in zip(pieces, body):
            assert vocab.setdefault(p, i) == i, (p, vocab[p], i)
    assert len(set(vocab.values())) == len(vocab)          # bijection
    json.dump({"source": "examples/test_tokenizer.cpp:70-73", "cases": cases, "vocab": vocab},
              open(os.path.join(HERE, "tokenizer_known_answers.json"), "w", encoding="utf-8"), ensure_ascii=False, indent=0)
    print(len(cases), "vectors,", len(vocab), "vocabulary entries")


if __name__ == "__main__":
    main()
