// Integer-pipe microbenchmarks for the MSM roofline: MEASURED_PEAKS.json has no
// integer peak, so the denominators used by bench.py are measured here.
//   imad32     independent 32-bit IMAD chains (mad.lo.u32)
//   imadwide   independent IMAD.WIDE.U32 chains (mad.wide.u32, 64-bit accumulate)
//   dfma       independent FP64 FMA chains; dfma_plus_imad: the two interleaved 1:1
//   mont<N>    back-to-back Montgomery products of the field code itself (fp.cuh)
// Prints one JSON object.  Usage: microbench [iters]
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <cstdlib>

#include "fp.cuh"
#include "../../tools/fp64_product/fp_dfma.cuh"  // the rejected FP64-pipe product: measured here only

using namespace b200msm;

#define CK(x)                                                                      \
    do {                                                                           \
        cudaError_t e = (x);                                                       \
        if (e != cudaSuccess) {                                                    \
            fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e));                \
            exit(1);                                                               \
        }                                                                          \
    } while (0)

constexpr int ILP = 8;

__global__ void k_imad32(uint32_t *out, uint32_t seed, int iters) {
    uint32_t a[ILP], m = seed | 1u, x = threadIdx.x + 3;
#pragma unroll
    for (int i = 0; i < ILP; i++) a[i] = seed + i + threadIdx.x;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++)
#pragma unroll
            for (int i = 0; i < ILP; i++)
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(m), "r"(x));
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_imadwide(uint64_t *out, uint32_t seed, int iters) {
    uint64_t a[ILP];
    uint32_t m = seed | 1u, x = threadIdx.x + 3;
#pragma unroll
    for (int i = 0; i < ILP; i++) a[i] = seed + i + threadIdx.x;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++)
#pragma unroll
            for (int i = 0; i < ILP; i++)
                asm volatile("{ .reg .u32 lo; cvt.u32.u64 lo, %0; mad.wide.u32 %0, lo, %1, %0; }"
                             : "+l"(a[i])
                             : "r"(m));
    }
    uint64_t s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// FP64 pipe: independent DFMA chains; and DFMA interleaved 1:1 with IMAD (do the two pipes overlap?)
__global__ void k_dfma(double *out, double seed, int iters) {
    double a[ILP], m = seed, x = 1e-9 * threadIdx.x;
#pragma unroll
    for (int i = 0; i < ILP; i++) a[i] = seed + i + threadIdx.x;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++)
#pragma unroll
            for (int i = 0; i < ILP; i++)
                asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a[i]) : "d"(m), "d"(x));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_dfma_imad(double *out, double seed, int iters) {
    double a[ILP], m = seed, x = 1e-9 * threadIdx.x;
    uint32_t b[ILP], mi = (uint32_t)seed | 1u, xi = threadIdx.x + 3;
#pragma unroll
    for (int i = 0; i < ILP; i++) a[i] = seed + i + threadIdx.x, b[i] = i + threadIdx.x;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++)
#pragma unroll
            for (int i = 0; i < ILP; i++) {
                asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a[i]) : "d"(m), "d"(x));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(b[i]) : "r"(mi), "r"(xi));
            }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += a[i] + b[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// DFMA with three distinct register operands per chain (the shape of the products in
// fp_dfma.cuh), alone and interleaved 1:1 with IMAD.WIDE: does the FP64 pipe run beside the
// 64-bit integer multiply-adds of fp.cuh?
template <bool WIDE, bool FP>
__global__ void k_dfma3_wide(double *out, double seed, int iters) {
    double a[ILP], b[ILP], c[ILP];
    uint64_t w[ILP];
    uint32_t m = (uint32_t)seed | 1u;
#pragma unroll
    for (int i = 0; i < ILP; i++) {
        a[i] = seed + i + threadIdx.x, b[i] = 1.0 + 1e-9 * (i + threadIdx.x), c[i] = 1e-7 * i;
        w[i] = i + threadIdx.x;
    }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++)
#pragma unroll
            for (int i = 0; i < ILP; i++) {
                if (FP) asm volatile("fma.rz.f64 %0, %1, %2, %0;" : "+d"(a[i]) : "d"(b[i]), "d"(c[(i + u) % ILP]));
                if (WIDE)
                    asm volatile("{ .reg .u32 lo; cvt.u32.u64 lo, %0; mad.wide.u32 %0, lo, %1, %0; }"
                                 : "+l"(w[i])
                                 : "r"(m));
            }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += a[i] + (double)w[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// Random gather of 64-byte records (two uint4 per record), the access pattern of the bucket
// accumulation; idx = hash of the record number.  Reports useful GB/s.
__global__ void k_gather64(const uint4 *__restrict__ src, uint4 *__restrict__ dst, uint32_t mask,
                           uint32_t count) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    uint32_t h = i * 2654435761u;
    h ^= h >> 15;
    h *= 2246822519u;
    h ^= h >> 13;
    const uint4 *p = src + (size_t)(h & mask) * 4;
    uint4 a = p[0], b = p[1], c = p[2], d = p[3];
    a.x ^= b.x ^ c.x ^ d.x;
    a.y ^= b.y ^ c.y ^ d.y;
    a.z ^= b.z ^ c.z ^ d.z;
    a.w ^= b.w ^ c.w ^ d.w;
    dst[i] = a;
}

template <class P>
__global__ void k_mont(uint32_t *out, uint32_t seed, int iters) {
    using F = Fp<P>;
    F a = F::one(), b = F::one();
    a.v[0] ^= seed + threadIdx.x;
    b.v[1] ^= seed * 3 + blockIdx.x;
    a.v[P::N - 1] &= 0x0fffffff;
    b.v[P::N - 1] &= 0x0fffffff;
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
        a = F::mul_inline(a, b);
        b = F::mul_inline(b, a);
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < P::N; i++) s ^= a.v[i] ^ b.v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// the Montgomery product on the FP64 pipe: same canonical result as Fp<P>::mul
template <class P>
__device__ __forceinline__ Fp<P> mul_fp64(const Fp<P> &a, const Fp<P> &b) {
    Fp<P> t;
    FpDfma<P>::mul_words(t.v, a.v, b.v);
    return Fp<P>::reduce_once(t);
}

// NI integer-pipe products and ND FP64-pipe products (fp_dfma.cuh) per iteration on independent
// chains: do the two multipliers of the SM add up?
template <class P, int NI, int ND>
__global__ void k_mont_mix(uint32_t *out, uint32_t seed, int iters) {
    using F = Fp<P>;
    F x[NI + ND], b = F::one();
    b.v[1] ^= seed * 3 + blockIdx.x;
    b.v[P::N - 1] &= 0x0fffffff;
#pragma unroll
    for (int k = 0; k < NI + ND; k++) {
        x[k] = F::one();
        x[k].v[0] ^= seed + threadIdx.x + k;
        x[k].v[P::N - 1] &= 0x0fffffff;
    }
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < NI; k++) x[k] = F::mul_inline(x[k], b);
#pragma unroll
        for (int k = 0; k < ND; k++) x[NI + k] = mul_fp64<P>(x[NI + k], b);
    }
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < NI + ND; k++)
#pragma unroll
        for (int i = 0; i < P::N; i++) s ^= x[k].v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class K>
static float time_ms(K launch) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    launch();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 3; r++) {
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    return best;
}

int main(int argc, char **argv) {
    int iters = argc > 1 ? atoi(argv[1]) : 2000;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    void *out;
    CK(cudaMalloc(&out, (size_t)sms * 8 * 1024 * 8));
    printf("{\"gpu\": \"%s\", \"sms\": %d", prop.name, sms);
    for (int warps : {4, 8, 16, 32}) {
        int threads = warps * 32 > 1024 ? 1024 : warps * 32;
        int blocks = sms * (warps * 32 / threads);
        double ops = (double)blocks * threads * iters * 8.0 * ILP;
        float ms = time_ms([&] { k_imad32<<<blocks, threads>>>((uint32_t *)out, 12345, iters); });
        printf(", \"imad32_w%d_Tops\": %.3f", warps, ops / ms * 1e-9);
        ms = time_ms([&] { k_imadwide<<<blocks, threads>>>((uint64_t *)out, 12345, iters); });
        printf(", \"imadwide_w%d_Tops\": %.3f", warps, ops / ms * 1e-9);
        ms = time_ms([&] { k_dfma<<<blocks, threads>>>((double *)out, 1.0000001, iters); });
        printf(", \"dfma_w%d_Tops\": %.3f", warps, ops / ms * 1e-9);
        ms = time_ms([&] { k_dfma_imad<<<blocks, threads>>>((double *)out, 1.0000001, iters); });
        printf(", \"dfma_plus_imad_w%d_Tpairs\": %.3f", warps, ops / ms * 1e-9);
    }
    for (int wps : {4, 8, 12, 16}) {  // resident warps per SM
        int blocks = sms * wps / 4;
        double muls = (double)blocks * 128 * iters * 2.0;
        float ms = time_ms([&] { k_mont<AltBn128FqParams><<<blocks, 128>>>((uint32_t *)out, 7, iters); });
        printf(", \"mont8_w%d_Gmul\": %.2f", wps, muls / ms * 1e-6);
        ms = time_ms([&] { k_mont<Bls12377FqParams><<<blocks, 128>>>((uint32_t *)out, 7, iters / 2); });
        printf(", \"mont12_w%d_Gmul\": %.2f", wps, muls / 2 / ms * 1e-6);
        if (wps <= 8) {
            ms = time_ms([&] { k_mont<Bw6761FqParams><<<blocks, 128>>>((uint32_t *)out, 7, iters / 8); });
            printf(", \"mont24_w%d_Gmul\": %.2f", wps, muls / 8 / ms * 1e-6);
        }
    }
    if (argc > 2 && !strcmp(argv[2], "prof")) {  // one launch per mix, for ncu
#define PROF(P, NI, ND) k_mont_mix<P, NI, ND><<<sms * 4, 128>>>((uint32_t *)out, 7, iters);
        PROF(AltBn128FqParams, 2, 0)
        PROF(AltBn128FqParams, 0, 2)
        PROF(AltBn128FqParams, 1, 1)
        PROF(AltBn128FqParams, 2, 1)
        CK(cudaDeviceSynchronize());
        return 0;
    }
    if (argc > 2) {
        for (int warps : {8, 16, 32}) {
            int threads = warps * 32 > 1024 ? 1024 : warps * 32;
            int blocks = sms * (warps * 32 / threads);
            double ops = (double)blocks * threads * iters * 8.0 * ILP;
            float ms = time_ms([&] { k_dfma3_wide<false, true><<<blocks, threads>>>((double *)out, 1.0000001, iters); });
            printf(", \"dfma3_w%d_Tops\": %.3f", warps, ops / ms * 1e-9);
            ms = time_ms([&] { k_dfma3_wide<true, false><<<blocks, threads>>>((double *)out, 1.0000001, iters); });
            printf(", \"wide_w%d_Tops\": %.3f", warps, ops / ms * 1e-9);
            ms = time_ms([&] { k_dfma3_wide<true, true><<<blocks, threads>>>((double *)out, 1.0000001, iters); });
            printf(", \"dfma3_plus_wide_w%d_Tpairs\": %.3f", warps, ops / ms * 1e-9);
        }
    }
    if (argc > 2) {  // FP64-pipe products and integer/FP64 mixes (products/s, all chains counted)
#define MIX(P, NI, ND, tag, div)                                                                   \
    for (int wps : {12, 16, 24, 32}) {                                                             \
        int blocks = sms * wps / 4;                                                                \
        int it = iters / div;                                                                      \
        double muls = (double)blocks * 128 * it * (NI + ND);                                       \
        float ms = time_ms([&] { k_mont_mix<P, NI, ND><<<blocks, 128>>>((uint32_t *)out, 7, it); }); \
        printf(", \"" tag "_i%dd%d_w%d_Gmul\": %.2f", NI, ND, wps, muls / ms * 1e-6);              \
    }
        MIX(AltBn128FqParams, 2, 0, "mont8", 1)
        MIX(AltBn128FqParams, 0, 2, "mont8", 1)
        MIX(AltBn128FqParams, 1, 1, "mont8", 1)
        MIX(AltBn128FqParams, 2, 1, "mont8", 1)
        MIX(AltBn128FqParams, 3, 2, "mont8", 1)
        MIX(AltBn128FqParams, 2, 2, "mont8", 1)
        MIX(Bls12377FqParams, 2, 0, "mont12", 2)
        MIX(Bls12377FqParams, 0, 2, "mont12", 2)
        MIX(Bls12377FqParams, 1, 1, "mont12", 2)
        MIX(Bls12377FqParams, 2, 1, "mont12", 2)
        printf("}\n");
        return 0;
    }
    {
        const uint32_t recs = 1u << 24, count = 1u << 26;  // 1 GiB of 64-byte records
        uint4 *src, *dst;
        CK(cudaMalloc(&src, (size_t)recs * 64));
        CK(cudaMalloc(&dst, (size_t)count * 16));
        CK(cudaMemset(src, 1, (size_t)recs * 64));
        for (int gran : {128, 64, 32}) {
            cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, gran);
            float ms = time_ms([&] { k_gather64<<<count / 256, 256>>>(src, dst, recs - 1, count); });
            printf(", \"gather64_l2gran%d_GBps\": %.1f", gran, (double)count * 64 / ms * 1e-6);
        }
        cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, 128);
    }
    printf("}\n");
    return 0;
}
