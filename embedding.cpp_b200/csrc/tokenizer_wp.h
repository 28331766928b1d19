// tokenizer_wp.h -- C++ WordPiece tokenizer driven by the HF tokenizer.json stored in the GGUF file
// (key blob.tokenizer.json).  Replaces tokenizer.cpp + tokenizers-cpp + the Rust `tokenizers` crate
// (reference tokenizer.cpp:30-53, tokenizers-cpp/src/huggingface_tokenizer.cc:28-35): encode() is
// `Tokenizer::encode(text, add_special_tokens=false)` for BertNormalizer + BertPreTokenizer + WordPiece.
#pragma once
#include <cstdint>
#include <string>
#include <unordered_map>
#include <vector>

namespace bk {

class WordPieceTokenizer {
  public:
    bool load(const std::string &tokenizer_json, std::string &err);
    std::vector<int32_t> encode(const std::string &text) const;

  private:
    struct Added { std::string content; int32_t id; bool single_word, lstrip, rstrip, normalized; };
    std::unordered_map<std::string, int32_t> vocab_;
    std::vector<Added> added_;
    std::string unk_token_ = "[UNK]", prefix_ = "##";
    int32_t unk_id_ = 0;
    int max_chars_ = 100;
    bool clean_text_ = true, chinese_ = true, strip_accents_ = true, lowercase_ = true;
    bool has_normalizer_ = false, has_pretok_ = false;
    int trunc_max_ = -1;
    void encode_piece(const std::string &piece, std::vector<int32_t> &out) const;
    std::vector<uint32_t> normalize(const std::vector<uint32_t> &cps) const;
    void wordpiece(const std::vector<uint32_t> &word, std::vector<int32_t> &out) const;
};

}  // namespace bk
