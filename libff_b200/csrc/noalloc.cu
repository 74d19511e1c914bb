// b200msm_batch_to_special_noalloc: batch normalisation of DEVICE-resident group elements for
// callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc).  No counterpart in the reference.
// The arithmetic is batch_to_special's (batch_kernels.cuh: k_extract_z, the k_inv_*
// tree, k_apply_zinv) and the result equals b200msm_batch_to_special's limb for limb; the only
// difference is where the tree's scratch lives: the file-scope device array below instead of the
// engine's grow-only cudaMalloc slots (SL_INV_VAL / SL_INV_PRE / SL_INV_PROD).
//
// This entry point calls no cudaMalloc / cudaFree and no allocating library routine: it resolves
// the array's device address once per device (cudaGetSymbolAddress, with that device current) and
// launches kernels on the caller's stream.  It needs no engine context (b200msm_init is not
// required), only a current device.
//
// Every call on a device shares the one array, which keeps the previous call's contents.  Nothing
// in the inversion tree needs to start at zero -- each scratch word a call reads was written by
// an earlier kernel of the SAME call (z by k_extract_z, prefix products and chunk products by
// k_inv_up / k_inv_root) -- so nothing is cleared.  Calls that overlap on the GPU (two streams,
// two host threads on one device) race on the array: serialise them.
#include <mutex>

#include "batch_kernels.cuh"
#include "engine.hpp"
#include "../../include/b200msm.h"

namespace b200msm {
namespace {

constexpr size_t NOALLOC_MAX_POINTS = B200MSM_NOALLOC_MAX_POINTS;
constexpr size_t NOALLOC_MAX_ELEM_BYTES = 96;  // widest coordinate: bw6_761 Fq, bls12_377/381 Fq2

struct TreeSizes {
    size_t pre, prod;  // field elements in the prefix-product and chunk-product levels
    int levels;
};
constexpr TreeSizes tree_sizes(size_t n) {
    TreeSizes t{n, 0, 1};
    size_t s = n;
    while (s > (size_t)INV_CHUNK) {
        s = (s + INV_CHUNK - 1) / INV_CHUNK;
        t.pre += s;
        t.prod += s;
        t.levels++;
    }
    return t;
}
constexpr size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }
constexpr size_t scratch_bytes(size_t n, size_t eb) {
    return align256(n * eb) + align256(tree_sizes(n).pre * eb) +
           align256((tree_sizes(n).prod ? tree_sizes(n).prod : 1) * eb);
}
constexpr size_t NOALLOC_SCRATCH_BYTES = scratch_bytes(NOALLOC_MAX_POINTS, NOALLOC_MAX_ELEM_BYTES);
static_assert(NOALLOC_SCRATCH_BYTES <= (8u << 20), "the static scratch is meant to stay modest");

// Held by every GPU that loads the library until the process exits (6.4 MiB).  In host code this
// name is only a handle for cudaGetSymbolAddress, never a device address.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_noalloc_mutex;
unsigned char *g_scratch_ptr[MAX_DEVICES];  // device address of the array, per device

// once per device, with that device current
bool scratch_for_current_device(unsigned char **out) {
    int dev = -1;
    B200_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) return set_error("batch_to_special_noalloc: device %d", dev), false;
    std::lock_guard<std::mutex> lk(g_noalloc_mutex);
    if (!g_scratch_ptr[dev]) {
        void *p = nullptr;
        B200_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return set_error("batch_to_special_noalloc: no device address for the scratch array"), false;
        g_scratch_ptr[dev] = (unsigned char *)p;
    }
    *out = g_scratch_ptr[dev];
    return true;
}

#define NOALLOC_LAUNCH_CHECK()                                                                  \
    do {                                                                                        \
        cudaError_t e_ = cudaPeekAtLastError();                                                 \
        if (e_ != cudaSuccess) {                                                                \
            set_error("batch_to_special_noalloc: kernel launch failed: %s (%s:%d)",             \
                      cudaGetErrorString(e_), __FILE__, __LINE__);                              \
            return false;                                                                       \
        }                                                                                       \
    } while (0)

template <class C>
bool to_special_static(unsigned char *scratch, const uint32_t *in, uint32_t *out, size_t n,
                       cudaStream_t stream) {
    using F = typename C::Fq;
    constexpr size_t eb = sizeof(uint32_t) * F::N;
    static_assert(eb <= NOALLOC_MAX_ELEM_BYTES, "scratch sized for 96-byte coordinates");
    // level sizes n0 = n, n1 = ceil(n0 / CHUNK), ... until <= CHUNK (at most 4 levels for 2^15)
    size_t sizes[8];
    int L = 0;
    sizes[L++] = n;
    while (sizes[L - 1] > (size_t)INV_CHUNK) {
        sizes[L] = (sizes[L - 1] + INV_CHUNK - 1) / INV_CHUNK;
        L++;
    }
    uint32_t *z = (uint32_t *)scratch;
    uint32_t *pre = (uint32_t *)(scratch + align256(n * eb));
    uint32_t *prod = (uint32_t *)(scratch + align256(n * eb) + align256(tree_sizes(n).pre * eb));
    uint32_t *lv_val[8], *lv_pre[8];
    lv_val[0] = z;
    lv_pre[0] = pre;
    {
        uint32_t *pp = pre + sizes[0] * F::N, *pr = prod;
        for (int l = 1; l < L; l++) {
            lv_val[l] = pr;
            lv_pre[l] = pp;
            pr += sizes[l] * F::N;
            pp += sizes[l] * F::N;
        }
    }
    auto blocks = [](size_t t) { return (unsigned)((t + 127) / 128); };
    k_extract_z<C><<<blocks(n), 128, 0, stream>>>(in, z, (uint32_t)n);
    NOALLOC_LAUNCH_CHECK();
    for (int l = 0; l + 1 < L; l++) {
        k_inv_up<F><<<blocks(sizes[l + 1]), 128, 0, stream>>>(lv_val[l], lv_pre[l], lv_val[l + 1],
                                                             (uint32_t)sizes[l]);
        NOALLOC_LAUNCH_CHECK();
    }
    k_inv_root<F><<<1, 32, 0, stream>>>(lv_val[L - 1], lv_pre[L - 1], (uint32_t)sizes[L - 1]);
    NOALLOC_LAUNCH_CHECK();
    for (int l = L - 1; l-- > 0;) {
        k_inv_down<F><<<blocks(sizes[l + 1]), 128, 0, stream>>>(lv_val[l], lv_pre[l], lv_val[l + 1],
                                                               (uint32_t)sizes[l]);
        NOALLOC_LAUNCH_CHECK();
    }
    k_apply_zinv<C><<<blocks(n), 128, 0, stream>>>(in, z, out, nullptr, nullptr, (uint32_t)n);
    NOALLOC_LAUNCH_CHECK();
    return true;
}

}  // namespace
}  // namespace b200msm

using namespace b200msm;

extern "C" bool b200msm_batch_to_special_noalloc(uint32_t group, const void *d_in, void *d_out,
                                                 size_t bytes, size_t n, void *cuda_stream) {
    const size_t gb = b200msm_group_bytes(group);
    if (gb == 0) return set_error("batch_to_special_noalloc: unknown group id %u", group), false;
    if (n > NOALLOC_MAX_POINTS)
        return set_error("batch_to_special_noalloc: n = %zu exceeds the static scratch's maximum of "
                         "%zu points per call (B200MSM_NOALLOC_MAX_POINTS); split the batch or use "
                         "b200msm_batch_to_special",
                         n, NOALLOC_MAX_POINTS),
               false;
    if (bytes != n * gb)
        return set_error("batch_to_special_noalloc: buffer size does not match n = %zu", n), false;
    if (n == 0) return true;
    if (!d_in || !d_out) return set_error("batch_to_special_noalloc: null buffer"), false;
    unsigned char *scratch = nullptr;
    if (!scratch_for_current_device(&scratch)) return false;
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    const uint32_t *in = (const uint32_t *)d_in;
    uint32_t *out = (uint32_t *)d_out;
    switch (group) {
        case 0: return to_special_static<AltBn128G1>(scratch, in, out, n, stream);
        case 1: return to_special_static<AltBn128G2>(scratch, in, out, n, stream);
        case 2: return to_special_static<Bls12377G1>(scratch, in, out, n, stream);
        case 3: return to_special_static<Bw6761G1>(scratch, in, out, n, stream);
        case 4: return to_special_static<Bls12377G2>(scratch, in, out, n, stream);
        case 5: return to_special_static<Bw6761G2>(scratch, in, out, n, stream);
        case 6: return to_special_static<Bls12381G1>(scratch, in, out, n, stream);
        case 7: return to_special_static<Bls12381G2>(scratch, in, out, n, stream);
    }
    return set_error("batch_to_special_noalloc: unknown group id %u", group), false;
}
