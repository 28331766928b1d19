// quantize.cpp -- bert_model_quantize: F32/F16 GGUF -> Q4_0 / Q4_1 GGUF (host-side file tool).
// Replaces reference bert.cpp:1313-1599; same selection rule (2-D tensors whose name ends in "weight",
// bert.cpp:1431-1436), same block codec (quantize_row_q4_0_reference / quantize_row_q4_1_reference,
// ggml-quants.c:428-510), same 32-byte tensor padding (bert.cpp:1561-1562) and the two KVs the
// reference adds (general.quantization_version = 2, general.file_type; bert.cpp:1377-1378).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/bert.h"
#include "gguf.h"

using namespace bk;

namespace {

float h2f(uint16_t h) {
    const uint32_t s = (uint32_t)(h & 0x8000) << 16, e = (h >> 10) & 0x1F, m = h & 0x3FF;
    uint32_t u;
    if (e == 0) {
        if (m == 0) u = s;
        else {
            int sh = 0;
            uint32_t mm = m;
            while (!(mm & 0x400)) { mm <<= 1; sh++; }
            u = s | ((uint32_t)(127 - 15 - sh + 1) << 23) | ((mm & 0x3FF) << 13);
        }
    } else if (e == 31) u = s | 0x7F800000u | (m << 13);
    else u = s | ((e + 112) << 23) | (m << 13);
    float f;
    memcpy(&f, &u, 4);
    return f;
}
uint16_t f2h(float f) {  // round-to-nearest-even, like _cvtss_sh(x, 0)
    uint32_t u;
    memcpy(&u, &f, 4);
    const uint32_t s = (u >> 16) & 0x8000;
    u &= 0x7FFFFFFF;
    if (u >= 0x7F800000u) return (uint16_t)(s | 0x7C00 | ((u > 0x7F800000u) ? 0x200 : 0));
    if (u >= 0x477FF000u) return (uint16_t)(s | 0x7C00);  // rounds to inf
    if (u < 0x33000001u) return (uint16_t)s;              // rounds to zero (< half of the min subnormal)
    const int e = (int)(u >> 23) - 127;
    uint32_t m = (u & 0x7FFFFF) | 0x800000;
    int shift = (e < -14) ? (13 + (-14 - e)) : 13;
    uint32_t q = m >> shift;
    const uint32_t rem = m & ((1u << shift) - 1), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (q & 1))) q++;
    uint32_t r = (e < -14) ? q : (((uint32_t)(e + 15) << 10) + (q - 0x400u));
    return (uint16_t)(s | r);
}

void quant_q4_0(const float *x, uint8_t *y, int64_t k) {  // ggml-quants.c:428-463
    for (int64_t b = 0; b < k / 32; b++, x += 32, y += 18) {
        float amax = 0.f, mx = 0.f;
        for (int j = 0; j < 32; j++) { const float v = x[j]; if (amax < fabsf(v)) { amax = fabsf(v); mx = v; } }
        const float d = mx / -8.f, id = d ? 1.0f / d : 0.0f;
        const uint16_t dh = f2h(d);
        memcpy(y, &dh, 2);
        for (int j = 0; j < 16; j++) {
            const float x0 = x[j] * id, x1 = x[16 + j] * id;
            int q0 = (int8_t)(x0 + 8.5f), q1 = (int8_t)(x1 + 8.5f);
            if (q0 > 15) q0 = 15;
            if (q1 > 15) q1 = 15;
            y[2 + j] = (uint8_t)(q0 | (q1 << 4));
        }
    }
}
void quant_q4_1(const float *x, uint8_t *y, int64_t k) {  // ggml-quants.c:469-504
    for (int64_t b = 0; b < k / 32; b++, x += 32, y += 20) {
        float mn = INFINITY, mx = -INFINITY;
        for (int j = 0; j < 32; j++) { const float v = x[j]; if (v < mn) mn = v; if (v > mx) mx = v; }
        const float d = (mx - mn) / 15.f, id = d ? 1.0f / d : 0.0f;
        const uint16_t dh = f2h(d), mh = f2h(mn);
        memcpy(y, &dh, 2);
        memcpy(y + 2, &mh, 2);
        for (int j = 0; j < 16; j++) {
            const float x0 = (x[j] - mn) * id, x1 = (x[16 + j] - mn) * id;
            int q0 = (int8_t)(x0 + 0.5f), q1 = (int8_t)(x1 + 0.5f);
            if (q0 > 15) q0 = 15;
            if (q1 > 15) q1 = 15;
            y[4 + j] = (uint8_t)(q0 | (q1 << 4));
        }
    }
}

void put_str(std::vector<uint8_t> &o, const std::string &s) {
    const uint64_t n = s.size();
    o.insert(o.end(), (const uint8_t *)&n, (const uint8_t *)&n + 8);
    o.insert(o.end(), s.begin(), s.end());
}
template <class T> void put(std::vector<uint8_t> &o, T v) { o.insert(o.end(), (const uint8_t *)&v, (const uint8_t *)&v + sizeof(T)); }
bool ends_with(const std::string &s, const char *suf) {
    const size_t n = strlen(suf);
    return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

}  // namespace

extern "C" BERT_API bool bert_model_quantize(const char *fname_inp, const char *fname_out, int ftype) {
    if (ftype != 2 && ftype != 3) { fprintf(stderr, "%s: invalid quantization type %d (2 = q4_0, 3 = q4_1)\n", __func__, ftype); return false; }
    GgufFile g;
    std::string err;
    if (!g.open(fname_inp, err)) { fprintf(stderr, "%s: failed to open '%s': %s\n", __func__, fname_inp, err.c_str()); return false; }
    const uint32_t AL = g.alignment;  // the input's general.alignment KV is carried over verbatim, so the output must honour it (default 32)
    std::vector<uint8_t> head;
    head.insert(head.end(), {'G', 'G', 'U', 'F'});
    put<uint32_t>(head, 3);
    put<uint64_t>(head, g.tensors.size());
    uint64_t n_kv = 0;
    std::vector<uint8_t> kvs;
    for (const auto &kvp : g.kv) {
        if (kvp.first == "general.quantization_version" || kvp.first == "general.file_type") continue;
        put_str(kvs, kvp.first);
        put<uint32_t>(kvs, kvp.second.type);
        kvs.insert(kvs.end(), g.base + kvp.second.raw_off, g.base + kvp.second.raw_off + kvp.second.raw_len);
        n_kv++;
    }
    put_str(kvs, "general.quantization_version"); put<uint32_t>(kvs, GT_U32); put<uint32_t>(kvs, 2); n_kv++;
    put_str(kvs, "general.file_type"); put<uint32_t>(kvs, GT_U32); put<uint32_t>(kvs, (uint32_t)ftype); n_kv++;
    put<uint64_t>(head, n_kv);
    head.insert(head.end(), kvs.begin(), kvs.end());

    // new tensor types / sizes / offsets
    struct Out { uint32_t type; size_t bytes; uint64_t off; bool quant; };
    std::vector<Out> outs(g.tensors.size());
    uint64_t off = 0;
    for (size_t i = 0; i < g.tensors.size(); i++) {
        const GgufTensor &t = g.tensors[i];
        if (t.nbytes == 0) { fprintf(stderr, "%s: tensor '%s' has unsupported type %u\n", __func__, t.name.c_str(), t.type); return false; }
        const bool q = ends_with(t.name, "weight") && t.n_dims == 2 && (t.type == 0 || t.type == 1) && t.ne[0] % 32 == 0;
        outs[i].quant = q;
        outs[i].type = q ? (uint32_t)ftype : t.type;
        outs[i].bytes = q ? ggml_tensor_bytes((uint32_t)ftype, t.ne) : t.nbytes;
        outs[i].off = off;
        off += (outs[i].bytes + AL - 1) / AL * AL;
    }
    for (size_t i = 0; i < g.tensors.size(); i++) {
        const GgufTensor &t = g.tensors[i];
        put_str(head, t.name);
        put<uint32_t>(head, t.n_dims);
        for (uint32_t k = 0; k < t.n_dims; k++) put<uint64_t>(head, t.ne[k]);
        put<uint32_t>(head, outs[i].type);
        put<uint64_t>(head, outs[i].off);
    }
    while (head.size() % AL) head.push_back(0);

    FILE *fo = fopen(fname_out, "wb");
    if (!fo) { fprintf(stderr, "%s: cannot create '%s'\n", __func__, fname_out); return false; }
    bool ok = fwrite(head.data(), 1, head.size(), fo) == head.size();
    std::vector<float> f32;
    std::vector<uint8_t> qbuf;
    for (size_t i = 0; i < g.tensors.size() && ok; i++) {
        const GgufTensor &t = g.tensors[i];
        const uint8_t *src = g.tensor_data(t);
        const uint8_t *data = src;
        if (outs[i].quant) {
            const int64_t K = (int64_t)t.ne[0], rows = (int64_t)t.ne[1];
            f32.resize((size_t)K * rows);
            if (t.type == 0) memcpy(f32.data(), src, f32.size() * 4);
            else { const uint16_t *hp = (const uint16_t *)src; for (size_t k = 0; k < f32.size(); k++) f32[k] = h2f(hp[k]); }
            qbuf.resize(outs[i].bytes);
            const size_t rb = outs[i].bytes / (size_t)rows;
            const int nt = (int)std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
            std::vector<std::thread> th;
            for (int w = 0; w < nt; w++)
                th.emplace_back([&, w]() {
                    for (int64_t r = w; r < rows; r += nt)
                        (ftype == 2 ? quant_q4_0 : quant_q4_1)(f32.data() + (size_t)r * K, qbuf.data() + (size_t)r * rb, K);
                });
            for (auto &x : th) x.join();
            data = qbuf.data();
        }
        ok = fwrite(data, 1, outs[i].bytes, fo) == outs[i].bytes;
        static const uint8_t zeros[32] = {0};
        const size_t pad = (AL - outs[i].bytes % AL) % AL;
        if (ok && pad) ok = fwrite(zeros, 1, pad, fo) == pad;
    }
    ok = (fclose(fo) == 0) && ok;
    if (!ok) fprintf(stderr, "%s: write to '%s' failed\n", __func__, fname_out);
    return ok;
}
