// gguf.h -- minimal GGUF v2/v3 reader/writer for BERT encoder files (host C++).
// Container layout: reference ggml/src/ggml.c:18017-18306 and ggml/docs/gguf.md; key names gguf.h:256-302.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace bk {

enum GgufType : uint32_t {  // ggml.c:17878-17909
    GT_U8 = 0, GT_I8, GT_U16, GT_I16, GT_U32, GT_I32, GT_F32, GT_BOOL, GT_STR, GT_ARR, GT_U64, GT_I64, GT_F64
};

struct GgufValue {
    uint32_t type = GT_U32;
    uint32_t elem_type = GT_U32;      // arrays
    uint64_t u = 0;                   // any integer / bool scalar
    double f = 0;                     // float scalars
    std::string s;                    // strings
    std::vector<std::string> strs;    // array<string>
    uint64_t arr_n = 0;               // arrays: element count
    size_t raw_off = 0, raw_len = 0;  // byte range of the whole value inside the file (for verbatim copy)
};

struct GgufTensor {
    std::string name;
    uint32_t n_dims = 0;
    uint64_t ne[4] = {1, 1, 1, 1};  // ne[0] = contiguous dim
    uint32_t type = 0;              // ggml_type
    uint64_t offset = 0;            // relative to data section
    size_t nbytes = 0;
};

struct GgufFile {
    uint32_t version = 0;
    uint32_t alignment = 32;
    std::vector<std::pair<std::string, GgufValue>> kv;  // file order
    std::map<std::string, size_t> kv_index;
    std::vector<GgufTensor> tensors;
    std::map<std::string, size_t> tensor_index;
    size_t data_off = 0;
    // mapping
    const uint8_t *base = nullptr;
    size_t size = 0;
    int fd = -1;

    ~GgufFile();
    bool open(const char *path, std::string &err);
    const GgufValue *find(const std::string &key) const;
    bool get_u32(const std::string &key, uint32_t &out) const;
    bool get_f32(const std::string &key, float &out) const;
    const GgufTensor *tensor(const std::string &name) const;
    const uint8_t *tensor_data(const GgufTensor &t) const { return base + data_off + t.offset; }
};

size_t ggml_tensor_bytes(uint32_t type, const uint64_t ne[4]);  // 0 for unsupported types

}  // namespace bk
