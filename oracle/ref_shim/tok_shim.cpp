// Replaces the reference's tokenizer.cpp when building oracle/_ref (test
// infrastructure, never linked into the product).  tokenizer.cpp cannot be
// used: it needs the Rust tokenizers crate, and bert_tokenizer::load() falls
// off the end of a non-void function (tokenizer.cpp:30-37), which gcc 13
// turns into a trap.  The oracle is only ever driven through bert_eval with
// token ids, so encode() returning nothing is sufficient: bert_tokenize then
// yields [CLS][SEP].
#include "tokenizer.h"

bert_tokenizer::bert_tokenizer() {}
bert_tokenizer::~bert_tokenizer() {}
bool bert_tokenizer::load(const std::string&) { return true; }
std::string bert_tokenizer::decode(const std::vector<int>&) { return std::string(); }
std::string bert_tokenizer::decode(const int32_t) { return std::string(); }
std::vector<int> bert_tokenizer::encode(const std::string&) { return std::vector<int>(); }
std::string bert_tokenizer::load_bytes_from_file(const std::string&) { return std::string(); }
void bert_tokenizer::print_encode_result(const std::vector<int>&) {}
