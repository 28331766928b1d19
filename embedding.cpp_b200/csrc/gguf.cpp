// gguf.cpp -- GGUF v2/v3 parser over an mmap of the model file (replaces gguf_init_from_file,
// reference ggml/src/ggml.c:18017-18306, and the gguf_file wrapper gguf.h:51-141).
#include "gguf.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstring>

namespace bk {

// sizes come from an untrusted file: every product is overflow-checked (0 = unsupported / implausible, rejected by the loader)
size_t ggml_tensor_bytes(uint32_t type, const uint64_t ne[4]) {
    uint64_t n = 1;
    for (int i = 0; i < 4; i++)
        if (__builtin_mul_overflow(n, ne[i], &n)) return 0;
    uint64_t bytes = 0;
    switch (type) {
        case 0: if (__builtin_mul_overflow(n, (uint64_t)4, &bytes)) return 0; return (size_t)bytes;   // F32
        case 1: if (__builtin_mul_overflow(n, (uint64_t)2, &bytes)) return 0; return (size_t)bytes;   // F16
        case 2: if (ne[0] % 32 || __builtin_mul_overflow(n / 32, (uint64_t)18, &bytes)) return 0; return (size_t)bytes;  // Q4_0: f16 d + 16 B
        case 3: if (ne[0] % 32 || __builtin_mul_overflow(n / 32, (uint64_t)20, &bytes)) return 0; return (size_t)bytes;  // Q4_1: f16 d + f16 m + 16 B
    }
    return 0;
}

namespace {
struct Cursor {
    const uint8_t *p;
    size_t n, pos = 0;
    bool ok = true;
    template <class T> T rd() {
        T v{};
        if (pos + sizeof(T) > n) { ok = false; return v; }
        memcpy(&v, p + pos, sizeof(T));
        pos += sizeof(T);
        return v;
    }
    std::string str() {
        const uint64_t len = rd<uint64_t>();
        if (!ok || len > n - pos) { ok = false; return {}; }
        std::string s(reinterpret_cast<const char *>(p + pos), (size_t)len);
        pos += (size_t)len;
        return s;
    }
};
size_t scalar_size(uint32_t t) {
    switch (t) {
        case GT_U8: case GT_I8: case GT_BOOL: return 1;
        case GT_U16: case GT_I16: return 2;
        case GT_U32: case GT_I32: case GT_F32: return 4;
        case GT_U64: case GT_I64: case GT_F64: return 8;
    }
    return 0;
}
bool read_scalar(Cursor &c, uint32_t t, GgufValue &v) {
    switch (t) {
        case GT_U8: v.u = c.rd<uint8_t>(); break;
        case GT_I8: v.u = (uint64_t)(int64_t)c.rd<int8_t>(); break;
        case GT_U16: v.u = c.rd<uint16_t>(); break;
        case GT_I16: v.u = (uint64_t)(int64_t)c.rd<int16_t>(); break;
        case GT_U32: v.u = c.rd<uint32_t>(); break;
        case GT_I32: v.u = (uint64_t)(int64_t)c.rd<int32_t>(); break;
        case GT_U64: v.u = c.rd<uint64_t>(); break;
        case GT_I64: v.u = (uint64_t)c.rd<int64_t>(); break;
        case GT_BOOL: v.u = c.rd<uint8_t>() != 0; break;
        case GT_F32: v.f = c.rd<float>(); break;
        case GT_F64: v.f = c.rd<double>(); break;
        default: return false;
    }
    if (t != GT_F32 && t != GT_F64) v.f = (double)(int64_t)v.u;
    return c.ok;
}
}  // namespace

GgufFile::~GgufFile() {
    if (base) munmap(const_cast<uint8_t *>(base), size);
    if (fd >= 0) close(fd);
}

bool GgufFile::open(const char *path, std::string &err) {
    fd = ::open(path, O_RDONLY);
    if (fd < 0) { err = std::string("cannot open ") + path; return false; }
    struct stat sb;
    if (fstat(fd, &sb) != 0 || sb.st_size < 24) { err = "file too small to be GGUF"; return false; }
    size = (size_t)sb.st_size;
    void *m = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
    if (m == MAP_FAILED) { err = "mmap failed"; base = nullptr; return false; }
    base = static_cast<const uint8_t *>(m);
    Cursor c{base, size};
    if (memcmp(base, "GGUF", 4) != 0) { err = "bad magic (not a GGUF file)"; return false; }
    c.pos = 4;
    version = c.rd<uint32_t>();
    if (version < 2 || version > 3) { err = "unsupported GGUF version " + std::to_string(version); return false; }  // v1 rejected, ggml.c:18057
    const uint64_t n_tensors = c.rd<uint64_t>(), n_kv = c.rd<uint64_t>();
    if (n_tensors > (1u << 20) || n_kv > (1u << 20)) { err = "implausible tensor / kv count"; return false; }
    for (uint64_t i = 0; i < n_kv && c.ok; i++) {
        std::string key = c.str();
        GgufValue v;
        v.type = c.rd<uint32_t>();
        v.raw_off = c.pos;
        if (v.type == GT_STR) v.s = c.str();
        else if (v.type == GT_ARR) {
            v.elem_type = c.rd<uint32_t>();
            v.arr_n = c.rd<uint64_t>();
            if (v.elem_type == GT_STR) {
                if (v.arr_n > size) { c.ok = false; break; }
                v.strs.reserve((size_t)v.arr_n);
                for (uint64_t j = 0; j < v.arr_n && c.ok; j++) v.strs.push_back(c.str());
            } else {
                const size_t es = scalar_size(v.elem_type);
                if (!es || v.arr_n > (size - c.pos) / es) { c.ok = false; break; }
                c.pos += (size_t)v.arr_n * es;
            }
        } else if (!read_scalar(c, v.type, v)) { err = "bad value type for key " + key; return false; }
        v.raw_len = c.pos - v.raw_off;
        kv_index[key] = kv.size();
        kv.emplace_back(std::move(key), std::move(v));
    }
    if (!c.ok) { err = "truncated GGUF key/value section"; return false; }
    uint32_t al = 0;
    if (get_u32("general.alignment", al) && al) alignment = al;
    for (uint64_t i = 0; i < n_tensors && c.ok; i++) {
        GgufTensor t;
        t.name = c.str();
        t.n_dims = c.rd<uint32_t>();
        if (t.n_dims < 1 || t.n_dims > 4) { err = "tensor " + t.name + ": bad n_dims"; return false; }
        for (uint32_t k = 0; k < t.n_dims; k++) t.ne[k] = c.rd<uint64_t>();
        t.type = c.rd<uint32_t>();
        t.offset = c.rd<uint64_t>();
        t.nbytes = ggml_tensor_bytes(t.type, t.ne);
        tensor_index[t.name] = tensors.size();
        tensors.push_back(std::move(t));
    }
    if (!c.ok) { err = "truncated GGUF tensor-info section"; return false; }
    if (alignment > (1u << 20) || (alignment & (alignment - 1))) { err = "implausible general.alignment"; return false; }
    data_off = (c.pos + alignment - 1) / alignment * alignment;
    if (data_off > size) { err = "truncated GGUF file (no data section)"; return false; }
    const uint64_t room = size - data_off;
    for (const auto &t : tensors) {
        if (t.nbytes == 0) continue;  // unsupported type: reported by the model loader if the tensor is needed
        if (t.offset > room || t.nbytes > room - t.offset) { err = "tensor " + t.name + " extends past end of file"; return false; }  // no wrap-around
    }
    return true;
}

const GgufValue *GgufFile::find(const std::string &key) const {
    auto it = kv_index.find(key);
    return it == kv_index.end() ? nullptr : &kv[it->second].second;
}
bool GgufFile::get_u32(const std::string &key, uint32_t &out) const {
    const GgufValue *v = find(key);
    if (!v || v->type == GT_STR || v->type == GT_ARR || v->type == GT_F32 || v->type == GT_F64) return false;
    out = (uint32_t)v->u;
    return true;
}
bool GgufFile::get_f32(const std::string &key, float &out) const {
    const GgufValue *v = find(key);
    if (!v || v->type == GT_STR || v->type == GT_ARR) return false;
    out = (float)v->f;
    return true;
}
const GgufTensor *GgufFile::tensor(const std::string &name) const {
    auto it = tensor_index.find(name);
    return it == tensor_index.end() ? nullptr : &tensors[it->second];
}

}  // namespace bk
