// tokenizer_wp.cpp -- WordPiece tokenisation from an HF tokenizer.json, in C++.
//
// The reference delegates to the Rust crate `tokenizers` 0.13.4 (tokenizers-cpp/rust/Cargo.toml:11) through
// tokenizers_encode(..., add_special_tokens = false) (tokenizers-cpp/src/huggingface_tokenizer.cc:28-35).
// That crate is not vendored and there is no Rust toolchain here, so this file restates the published
// algorithm of the pipeline BERT-family files use:
//   added_tokens split -> BertNormalizer{clean_text, handle_chinese_chars, strip_accents, lowercase}
//   -> BertPreTokenizer (whitespace split, every punctuation char isolated) -> WordPiece (greedy
//   longest-match-first, "##" continuation prefix, whole word -> unk_token on failure,
//   max_input_chars_per_word) -> truncation (right, max_length).
// Parity is checked against the Python `tokenizers` wheel (same Rust library) in tests/test_tokenizer.py.
#include "tokenizer_wp.h"

#include <algorithm>
#include <cstring>
#include <map>
#include <memory>

namespace bk {
namespace {

#include "unicode_tables.inc"

// ------------------------------------------------------------------------------- mini JSON
struct JValue {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    bool b = false;
    double num = 0;
    std::string str;
    std::vector<JValue> arr;
    std::vector<std::pair<std::string, JValue>> obj;
    const JValue *get(const char *key) const {
        if (kind != Obj) return nullptr;
        for (const auto &kv : obj) if (kv.first == key) return &kv.second;
        return nullptr;
    }
    bool is_null() const { return kind == Null; }
};

void append_utf8(std::string &s, uint32_t cp) {
    if (cp < 0x80) s.push_back((char)cp);
    else if (cp < 0x800) { s.push_back((char)(0xC0 | (cp >> 6))); s.push_back((char)(0x80 | (cp & 0x3F))); }
    else if (cp < 0x10000) { s.push_back((char)(0xE0 | (cp >> 12))); s.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); s.push_back((char)(0x80 | (cp & 0x3F))); }
    else { s.push_back((char)(0xF0 | (cp >> 18))); s.push_back((char)(0x80 | ((cp >> 12) & 0x3F))); s.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); s.push_back((char)(0x80 | (cp & 0x3F))); }
}

struct JParser {
    const char *p, *end;
    std::string err;
    void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) p++; }
    bool fail(const char *m) { if (err.empty()) err = m; return false; }
    bool hex4(uint32_t &v) {
        if (end - p < 4) return fail("bad \\u escape");
        v = 0;
        for (int i = 0; i < 4; i++) {
            const char c = *p++;
            v <<= 4;
            if (c >= '0' && c <= '9') v |= c - '0';
            else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
            else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
            else return fail("bad \\u escape");
        }
        return true;
    }
    bool string(std::string &out) {
        if (p >= end || *p != '"') return fail("expected string");
        p++;
        out.clear();
        while (p < end && *p != '"') {
            if (*p == '\\') {
                if (++p >= end) return fail("bad escape");
                const char c = *p++;
                switch (c) {
                    case 'n': out.push_back('\n'); break;
                    case 't': out.push_back('\t'); break;
                    case 'r': out.push_back('\r'); break;
                    case 'b': out.push_back('\b'); break;
                    case 'f': out.push_back('\f'); break;
                    case 'u': {
                        uint32_t cp;
                        if (!hex4(cp)) return false;
                        if (cp >= 0xD800 && cp <= 0xDBFF && end - p >= 6 && p[0] == '\\' && p[1] == 'u') {
                            p += 2;
                            uint32_t lo;
                            if (!hex4(lo)) return false;
                            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                        }
                        append_utf8(out, cp);
                        break;
                    }
                    default: out.push_back(c);
                }
            } else out.push_back(*p++);
        }
        if (p >= end) return fail("unterminated string");
        p++;
        return true;
    }
    bool value(JValue &v, int depth = 0) {
        if (depth > 64) return fail("nesting too deep");
        ws();
        if (p >= end) return fail("unexpected end");
        if (*p == '{') {
            p++;
            v.kind = JValue::Obj;
            ws();
            if (p < end && *p == '}') { p++; return true; }
            for (;;) {
                ws();
                std::string k;
                if (!string(k)) return false;
                ws();
                if (p >= end || *p != ':') return fail("expected ':'");
                p++;
                v.obj.emplace_back(std::move(k), JValue());
                if (!value(v.obj.back().second, depth + 1)) return false;
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == '}') { p++; return true; }
                return fail("expected ',' or '}'");
            }
        }
        if (*p == '[') {
            p++;
            v.kind = JValue::Arr;
            ws();
            if (p < end && *p == ']') { p++; return true; }
            for (;;) {
                v.arr.emplace_back();
                if (!value(v.arr.back(), depth + 1)) return false;
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == ']') { p++; return true; }
                return fail("expected ',' or ']'");
            }
        }
        if (*p == '"') { v.kind = JValue::Str; return string(v.str); }
        if (end - p >= 4 && !strncmp(p, "true", 4)) { v.kind = JValue::Bool; v.b = true; p += 4; return true; }
        if (end - p >= 5 && !strncmp(p, "false", 5)) { v.kind = JValue::Bool; v.b = false; p += 5; return true; }
        if (end - p >= 4 && !strncmp(p, "null", 4)) { v.kind = JValue::Null; p += 4; return true; }
        char *e = nullptr;
        v.num = strtod(p, &e);
        if (e == p) return fail("unexpected character");
        v.kind = JValue::Num;
        p = e;
        return true;
    }
};

// -------------------------------------------------------------------------------- unicode
template <size_t N> bool in_ranges(const uint32_t (&tab)[N][2], uint32_t cp) {
    size_t lo = 0, hi = N;
    while (lo < hi) {
        const size_t mid = (lo + hi) / 2;
        if (cp < tab[mid][0]) hi = mid;
        else if (cp > tab[mid][1]) lo = mid + 1;
        else return true;
    }
    return false;
}
bool is_whitespace(uint32_t c) {  // Unicode White_Space (Rust char::is_whitespace)
    return (c >= 9 && c <= 13) || c == 0x20 || c == 0x85 || c == 0xA0 || c == 0x1680 || (c >= 0x2000 && c <= 0x200A) ||
           c == 0x2028 || c == 0x2029 || c == 0x202F || c == 0x205F || c == 0x3000;
}
bool is_control(uint32_t c) {  // tokenizers/src/normalizers/bert.rs: \t \n \r are whitespace, otherwise category C*
    if (c == '\t' || c == '\n' || c == '\r') return false;
    return in_ranges(kOther, c);
}
bool is_chinese(uint32_t c) {  // bert.rs is_chinese_char
    return (c >= 0x4E00 && c <= 0x9FFF) || (c >= 0x3400 && c <= 0x4DBF) || (c >= 0x20000 && c <= 0x2A6DF) ||
           (c >= 0x2A700 && c <= 0x2B73F) || (c >= 0x2B740 && c <= 0x2B81F) || (c >= 0x2B920 && c <= 0x2CEAF) ||
           (c >= 0xF900 && c <= 0xFAFF) || (c >= 0x2F800 && c <= 0x2FA1F);
}
bool is_punct(uint32_t c) {  // pre_tokenizers/bert.rs is_bert_punc: ASCII punctuation or Unicode P*
    if ((c >= 33 && c <= 47) || (c >= 58 && c <= 64) || (c >= 91 && c <= 96) || (c >= 123 && c <= 126)) return true;
    return in_ranges(kPunct, c);
}
uint8_t ccc_of(uint32_t cp) {
    size_t lo = 0, hi = sizeof(kCcc) / sizeof(kCcc[0]);
    while (lo < hi) {
        const size_t mid = (lo + hi) / 2;
        if (kCcc[mid].cp < cp) lo = mid + 1; else hi = mid;
    }
    return (lo < sizeof(kCcc) / sizeof(kCcc[0]) && kCcc[lo].cp == cp) ? kCcc[lo].ccc : 0;
}
void nfd_append(uint32_t cp, std::vector<uint32_t> &out) {
    if (cp >= 0xAC00 && cp <= 0xD7A3) {  // Hangul syllable: algorithmic decomposition
        const uint32_t s = cp - 0xAC00, l = s / 588, v = (s % 588) / 28, t = s % 28;
        out.push_back(0x1100 + l);
        out.push_back(0x1161 + v);
        if (t) out.push_back(0x11A7 + t);
        return;
    }
    size_t lo = 0, hi = sizeof(kDecomp) / sizeof(kDecomp[0]);
    while (lo < hi) {
        const size_t mid = (lo + hi) / 2;
        if (kDecomp[mid].cp < cp) lo = mid + 1; else hi = mid;
    }
    if (lo < sizeof(kDecomp) / sizeof(kDecomp[0]) && kDecomp[lo].cp == cp)
        for (uint32_t i = 0; i < kDecomp[lo].n; i++) out.push_back(kDecomp[lo].d[i]);
    else out.push_back(cp);
}
std::vector<uint32_t> nfd(const std::vector<uint32_t> &in) {
    std::vector<uint32_t> out;
    out.reserve(in.size() + 8);
    for (uint32_t c : in) nfd_append(c, out);
    // canonical ordering: stable sort of each run of non-starters by combining class
    size_t i = 0;
    while (i < out.size()) {
        if (ccc_of(out[i]) == 0) { i++; continue; }
        size_t j = i;
        while (j < out.size() && ccc_of(out[j]) != 0) j++;
        std::stable_sort(out.begin() + i, out.begin() + j, [](uint32_t a, uint32_t b) { return ccc_of(a) < ccc_of(b); });
        i = j;
    }
    return out;
}
void lower_append(uint32_t cp, std::vector<uint32_t> &out) {
    if (cp < 0x80) { out.push_back((cp >= 'A' && cp <= 'Z') ? cp + 32 : cp); return; }
    size_t lo = 0, hi = sizeof(kLower) / sizeof(kLower[0]);
    while (lo < hi) {
        const size_t mid = (lo + hi) / 2;
        if (kLower[mid].cp < cp) lo = mid + 1; else hi = mid;
    }
    if (lo < sizeof(kLower) / sizeof(kLower[0]) && kLower[lo].cp == cp)
        for (uint32_t i = 0; i < kLower[lo].n; i++) out.push_back(kLower[lo].d[i]);
    else out.push_back(cp);
}
std::vector<uint32_t> decode_utf8(const std::string &s) {
    std::vector<uint32_t> out;
    out.reserve(s.size());
    const unsigned char *p = (const unsigned char *)s.data(), *e = p + s.size();
    while (p < e) {
        uint32_t c = *p;
        int n = 0;
        if (c < 0x80) n = 0;
        else if ((c & 0xE0) == 0xC0) { c &= 0x1F; n = 1; }
        else if ((c & 0xF0) == 0xE0) { c &= 0x0F; n = 2; }
        else if ((c & 0xF8) == 0xF0) { c &= 0x07; n = 3; }
        else { out.push_back(0xFFFD); p++; continue; }
        if (e - p <= n) { out.push_back(0xFFFD); break; }
        bool ok = true;
        for (int i = 1; i <= n; i++) {
            if ((p[i] & 0xC0) != 0x80) { ok = false; break; }
            c = (c << 6) | (p[i] & 0x3F);
        }
        if (!ok) { out.push_back(0xFFFD); p++; continue; }
        out.push_back(c);
        p += n + 1;
    }
    return out;
}
std::string encode_utf8(const uint32_t *b, const uint32_t *e) {
    std::string s;
    for (; b < e; b++) append_utf8(s, *b);
    return s;
}

}  // namespace

// =============================================================================== tokenizer
bool WordPieceTokenizer::load(const std::string &json, std::string &err) {
    JParser jp{json.data(), json.data() + json.size(), {}};
    JValue root;
    if (!jp.value(root) || root.kind != JValue::Obj) { err = "tokenizer.json: " + (jp.err.empty() ? std::string("not an object") : jp.err); return false; }
    const JValue *model = root.get("model");
    if (!model || model->kind != JValue::Obj) { err = "tokenizer.json: no model section"; return false; }
    const JValue *type = model->get("type");
    const JValue *vocab = model->get("vocab");
    if (type && type->kind == JValue::Str && type->str != "WordPiece") { err = "tokenizer model '" + type->str + "' is not WordPiece"; return false; }
    if (!vocab || vocab->kind != JValue::Obj) { err = "tokenizer.json: model.vocab missing"; return false; }
    vocab_.reserve(vocab->obj.size() * 2);
    for (const auto &kv : vocab->obj) vocab_[kv.first] = (int32_t)kv.second.num;
    if (const JValue *v = model->get("unk_token")) if (v->kind == JValue::Str) unk_token_ = v->str;
    if (const JValue *v = model->get("continuing_subword_prefix")) if (v->kind == JValue::Str) prefix_ = v->str;
    if (const JValue *v = model->get("max_input_chars_per_word")) if (v->kind == JValue::Num) max_chars_ = (int)v->num;
    auto it = vocab_.find(unk_token_);
    if (it == vocab_.end()) { err = "unk_token '" + unk_token_ + "' not in vocab"; return false; }
    unk_id_ = it->second;

    const JValue *norm = root.get("normalizer");
    has_normalizer_ = norm && norm->kind == JValue::Obj;
    if (has_normalizer_) {
        const JValue *t = norm->get("type");
        if (!t || t->kind != JValue::Str || t->str != "BertNormalizer") { err = "normalizer type not supported (only BertNormalizer)"; return false; }
        auto flag = [&](const char *k, bool dflt) { const JValue *v = norm->get(k); return (v && v->kind == JValue::Bool) ? v->b : dflt; };
        clean_text_ = flag("clean_text", true);
        chinese_ = flag("handle_chinese_chars", true);
        lowercase_ = flag("lowercase", true);
        const JValue *sa = norm->get("strip_accents");
        strip_accents_ = (sa && sa->kind == JValue::Bool) ? sa->b : lowercase_;  // null follows lowercase (bert.rs)
    }
    const JValue *pre = root.get("pre_tokenizer");
    has_pretok_ = pre && pre->kind == JValue::Obj;
    if (has_pretok_) {
        const JValue *t = pre->get("type");
        if (!t || t->kind != JValue::Str || t->str != "BertPreTokenizer") { err = "pre_tokenizer type not supported (only BertPreTokenizer)"; return false; }
    }
    if (const JValue *at = root.get("added_tokens")) {
        if (at->kind == JValue::Arr)
            for (const JValue &a : at->arr) {
                const JValue *c = a.get("content"), *id = a.get("id");
                if (!c || c->kind != JValue::Str || !id || c->str.empty()) continue;
                auto fl = [&](const char *k, bool dflt) { const JValue *v = a.get(k); return (v && v->kind == JValue::Bool) ? v->b : dflt; };
                added_.push_back({c->str, (int32_t)id->num, fl("single_word", false), fl("lstrip", false), fl("rstrip", false), fl("normalized", false)});
            }
    }
    if (const JValue *tr = root.get("truncation"))
        if (tr->kind == JValue::Obj)
            if (const JValue *ml = tr->get("max_length")) if (ml->kind == JValue::Num) trunc_max_ = (int)ml->num;
    return true;
}

std::vector<uint32_t> WordPieceTokenizer::normalize(const std::vector<uint32_t> &in) const {
    if (!has_normalizer_) return in;
    std::vector<uint32_t> a;
    a.reserve(in.size() + 16);
    if (clean_text_) {
        for (uint32_t c : in) {
            if (c == 0 || c == 0xFFFD || is_control(c)) continue;
            a.push_back(is_whitespace(c) ? ' ' : c);
        }
    } else a = in;
    if (chinese_) {
        std::vector<uint32_t> b;
        b.reserve(a.size() + 16);
        for (uint32_t c : a) {
            if (is_chinese(c)) { b.push_back(' '); b.push_back(c); b.push_back(' '); }
            else b.push_back(c);
        }
        a.swap(b);
    }
    if (strip_accents_) {
        std::vector<uint32_t> d = nfd(a);
        a.clear();
        for (uint32_t c : d) if (!in_ranges(kMn, c)) a.push_back(c);
    }
    if (lowercase_) {
        std::vector<uint32_t> b;
        b.reserve(a.size() + 8);
        for (uint32_t c : a) lower_append(c, b);
        a.swap(b);
    }
    return a;
}

void WordPieceTokenizer::wordpiece(const std::vector<uint32_t> &word, std::vector<int32_t> &out) const {
    const int n = (int)word.size();
    if (n == 0) return;
    if (n > max_chars_) { out.push_back(unk_id_); return; }
    std::vector<int32_t> pieces;
    int start = 0;
    std::string cand;
    while (start < n) {
        int end = n;
        int32_t found = -1;
        while (start < end) {
            cand = start > 0 ? prefix_ : std::string();
            cand += encode_utf8(word.data() + start, word.data() + end);
            auto it = vocab_.find(cand);
            if (it != vocab_.end()) { found = it->second; break; }
            end--;
        }
        if (found < 0) { out.push_back(unk_id_); return; }  // the whole word becomes [UNK]
        pieces.push_back(found);
        start = end;
    }
    out.insert(out.end(), pieces.begin(), pieces.end());
}

void WordPieceTokenizer::encode_piece(const std::string &piece, std::vector<int32_t> &out) const {
    const std::vector<uint32_t> cps = normalize(decode_utf8(piece));
    std::vector<uint32_t> word;
    auto flush = [&]() { if (!word.empty()) { wordpiece(word, out); word.clear(); } };
    for (uint32_t c : cps) {
        if (has_pretok_) {
            if (is_whitespace(c)) { flush(); continue; }
            if (is_punct(c)) { flush(); word.push_back(c); flush(); continue; }
        }
        word.push_back(c);
    }
    flush();
}

std::vector<int32_t> WordPieceTokenizer::encode(const std::string &text) const {
    std::vector<int32_t> out;
    // added (special) tokens are cut out of the raw text first: leftmost match, longest content on ties
    size_t pos = 0, seg = 0;
    const size_t n = text.size();
    auto is_word_byte = [&](size_t i) { const unsigned char c = (unsigned char)text[i]; return c >= 0x80 || (c >= '0' && c <= '9') || (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || c == '_'; };
    while (pos < n && !added_.empty()) {
        const Added *best = nullptr;
        for (const Added &a : added_) {
            if (a.content.size() > n - pos || text.compare(pos, a.content.size(), a.content) != 0) continue;
            if (a.single_word) {
                const size_t e = pos + a.content.size();
                if ((pos > 0 && is_word_byte(pos - 1)) || (e < n && is_word_byte(e))) continue;
            }
            if (!best || a.content.size() > best->content.size()) best = &a;
        }
        if (!best) { pos++; continue; }
        size_t left_end = pos, right_begin = pos + best->content.size();
        if (best->lstrip) while (left_end > seg && (text[left_end - 1] == ' ' || text[left_end - 1] == '\t' || text[left_end - 1] == '\n')) left_end--;
        if (best->rstrip) while (right_begin < n && (text[right_begin] == ' ' || text[right_begin] == '\t' || text[right_begin] == '\n')) right_begin++;
        if (left_end > seg) encode_piece(text.substr(seg, left_end - seg), out);
        out.push_back(best->id);
        pos = seg = right_begin;
    }
    if (seg < n) encode_piece(text.substr(seg), out);
    if (trunc_max_ >= 0 && (int)out.size() > trunc_max_) out.resize(trunc_max_);
    return out;
}

}  // namespace bk
