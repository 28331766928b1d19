// Stand-in for tokenizers-cpp's public header, used ONLY when compiling the
// reference's bert.cpp into oracle/_ref/libbert_ref.so (test infrastructure).
// The real header pulls in a Rust static library (crate `tokenizers` 0.13.4)
// that cannot be built in this image (no cargo/rustc).  The oracle is driven
// with token ids through bert_eval, so only the type has to exist.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace tokenizers {
class Tokenizer {
 public:
  virtual ~Tokenizer() {}
  virtual std::vector<int32_t> Encode(const std::string& text) = 0;
  virtual std::string Decode(const std::vector<int32_t>& ids) = 0;
};
}  // namespace tokenizers
