// batcher_static.cu -- bk_batcher_static_scratch: the on-device batcher (rowwise.cu: batch_sort_kernel + batch_pack_kernel)
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
// No counterpart in the reference (include/bk_api.h).
//
// The batcher needs two per-sentence scratch arrays between its two kernels: caller_off[n_seq] (where each sentence starts in
// the caller's concatenated ids) and src_off[n_seq] (the same, in sorted order).  The engine takes them from its cudaMalloc'd
// workspace; here they live in ONE file-scope __device__ array of fixed size, so the entry point allocates nothing: no
// cudaMalloc / cudaFree / cudaMallocAsync / host allocation, only two kernel launches on the caller's stream.  The kernels are
// the product's own (launch_batcher), unchanged.
//
//   * limit: BK_STATIC_BATCHER_MAX_SEQ = 65536 sentences per call (2 x 65536 x 4 B = 512 KiB, held by every GPU that loads
//     the library until the process exits); more is rejected with BK_STATIC_E_TOO_LARGE.
//   * nothing in the scratch has to start at zero: both arrays are fully written for [0, n_seq) by batch_sort_kernel before
//     batch_sort_kernel / batch_pack_kernel read them (rows beyond n_seq keep the previous call's contents and are never read).
//   * EVERY call on a device shares the array: two calls whose kernels overlap on the GPU (different streams, or graphs replayed
//     concurrently) race on it.  Calls on one stream are ordered and safe.
//   * the array's name is a host-side stand-in in host code, not a device address; the address the kernels get comes from
//     cudaGetSymbolAddress, resolved once per device with that device current (first call on that device; make that call outside
//     stream capture -- afterwards the entry point only launches).
#include <cstdio>
#include <mutex>

#include "kernels.h"
#include "../../include/bk_api.h"

using namespace bk;

namespace {
constexpr int kMaxDevices = 64;
// [0, MAX) caller_off, [MAX, 2 MAX) src_off
static __device__ __align__(256) int32_t g_batcher_scratch[2 * BK_STATIC_BATCHER_MAX_SEQ];

std::mutex g_mu;
int32_t *g_scratch_addr[kMaxDevices];  // per device: the array's device address, nullptr until resolved

int fail(int code, const char *what) {
    fprintf(stderr, "bk_batcher_static_scratch: %s\n", what);
    return code;
}

// device address of the scratch array on `device` (which must be current)
int32_t *scratch_on(int device, cudaError_t *err) {
    std::lock_guard<std::mutex> lk(g_mu);
    *err = cudaSuccess;
    if (!g_scratch_addr[device]) {
        void *p = nullptr;
        *err = cudaGetSymbolAddress(&p, g_batcher_scratch);
        if (*err != cudaSuccess) return nullptr;
        g_scratch_addr[device] = static_cast<int32_t *>(p);
    }
    return g_scratch_addr[device];
}
}  // namespace

BK_API int bk_batcher_static_scratch(int device, int32_t n_seq, int32_t n_tokens, const int32_t *d_lens, const int32_t *d_tokens,
                                     int32_t *d_order, int32_t *d_cu, int32_t *d_tok, int32_t *d_pos, void *stream) {
    char msg[256];
    if (device < 0 || device >= kMaxDevices) {
        snprintf(msg, sizeof msg, "device %d outside [0, %d)", device, kMaxDevices);
        return fail(BK_STATIC_E_ARG, msg);
    }
    if (n_seq < 1 || !d_lens || !d_tokens || !d_order || !d_cu || !d_tok || !d_pos) return fail(BK_STATIC_E_ARG, "n_seq < 1 or a null pointer");
    if (n_seq > BK_STATIC_BATCHER_MAX_SEQ) {
        snprintf(msg, sizeof msg, "n_seq = %d exceeds the static scratch (BK_STATIC_BATCHER_MAX_SEQ = %d sentences per call); split the batch",
                 n_seq, BK_STATIC_BATCHER_MAX_SEQ);
        return fail(BK_STATIC_E_TOO_LARGE, msg);
    }
    if (n_tokens < n_seq || (int64_t)n_tokens > (int64_t)n_seq * 512) {
        snprintf(msg, sizeof msg, "n_tokens = %d impossible for %d sentences of 1..512 tokens", n_tokens, n_seq);
        return fail(BK_STATIC_E_ARG, msg);
    }
    cudaError_t err = cudaSetDevice(device);
    if (err != cudaSuccess) return fail(BK_STATIC_E_CUDA, cudaGetErrorString(err));
    int32_t *scratch = scratch_on(device, &err);
    if (!scratch) {
        snprintf(msg, sizeof msg, "cudaGetSymbolAddress on device %d: %s", device, cudaGetErrorString(err));
        return fail(BK_STATIC_E_CUDA, msg);
    }
    BatcherParams b{d_tokens, d_lens, n_seq, n_tokens, d_tok, d_pos, d_cu, d_order,
                    /*src_off=*/scratch + BK_STATIC_BATCHER_MAX_SEQ, /*caller_off=*/scratch};
    err = launch_batcher(b, static_cast<cudaStream_t>(stream));
    if (err != cudaSuccess) {
        snprintf(msg, sizeof msg, "launch: %s", cudaGetErrorString(err));
        return fail(BK_STATIC_E_CUDA, msg);
    }
    return 0;
}
