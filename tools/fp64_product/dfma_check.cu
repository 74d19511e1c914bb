// Checks the FP64-pipe Montgomery product (fp_dfma.cuh) against the engine's integer product
// Fp<P>::mul (itself pinned to the reference by tests/test_leaves_gpu.py) on the device, bit for
// bit, for the seven moduli of the path.  The FP64 product was measured and rejected (DESIGN.md §4,
// profiles/microbench_r01_fp64.json); it is not part of libb200msm.so and lives here with its test.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr \
//        -I libff_b200/csrc -o tools/fp64_product/dfma_check tools/fp64_product/dfma_check.cu
// Prints one line per field and exits non-zero on the first mismatch.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "fp.cuh"
#include "fp_dfma.cuh"

using namespace b200msm;

#define CK(x)                                                                          \
    do {                                                                               \
        cudaError_t e_ = (x);                                                          \
        if (e_ != cudaSuccess) {                                                       \
            fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_));                   \
            exit(2);                                                                   \
        }                                                                              \
    } while (0)

template <class P>
__global__ void k_both(const uint32_t *a, const uint32_t *b, uint32_t *r_int, uint32_t *r_fp64, uint32_t n) {
    using F = Fp<P>;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    F x = F::load(a + (size_t)i * F::N), y = F::load(b + (size_t)i * F::N);
    F::mul(x, y).store(r_int + (size_t)i * F::N);
    F t;
    FpDfma<P>::mul_words(t.v, x.v, y.v);
    F::reduce_once(t).store(r_fp64 + (size_t)i * F::N);
}

// operands below p: random words with the top limb masked under the modulus' top limb, plus the
// corner values 0, 1 (as an integer), p - 1
template <class P>
static std::vector<uint32_t> operands(uint32_t n, uint64_t seed) {
    constexpr int N = P::N;
    std::vector<uint32_t> v((size_t)n * N);
    uint64_t s = seed * 0x9E3779B97F4A7C15ull + 1;
    auto next = [&]() {
        s ^= s << 13, s ^= s >> 7, s ^= s << 17;
        return (uint32_t)(s >> 16);
    };
    uint32_t top = P::mod(N - 1);
    uint32_t mask = 1;
    while (mask <= top / 2) mask <<= 1;  // highest power of two <= top
    mask -= 1;                           // values below it are below p in the top limb
    for (uint32_t i = 0; i < n; i++) {
        for (int k = 0; k < N; k++) v[(size_t)i * N + k] = next();
        v[(size_t)i * N + N - 1] &= mask;
    }
    auto set = [&](uint32_t i, int kind) {
        for (int k = 0; k < N; k++)
            v[(size_t)i * N + k] = kind == 0 ? 0u : kind == 1 ? (k == 0) : P::mod(k) - (k == 0);
    };
    if (n > 8) {
        set(seed & 1 ? 0 : 1, 0);
        set(2, 1);
        set(3 + (seed & 1), 2);
        set(6, 2);
    }
    return v;
}

template <class P>
static bool check(const char *name) {
    constexpr int N = P::N;
    const uint32_t n = 1 << 16;
    auto a = operands<P>(n, 11), b = operands<P>(n, 12);
    uint32_t *da, *db, *di, *dd;
    const size_t bytes = (size_t)n * N * 4;
    CK(cudaMalloc(&da, bytes));
    CK(cudaMalloc(&db, bytes));
    CK(cudaMalloc(&di, bytes));
    CK(cudaMalloc(&dd, bytes));
    CK(cudaMemcpy(da, a.data(), bytes, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(db, b.data(), bytes, cudaMemcpyHostToDevice));
    k_both<P><<<(n + 127) / 128, 128>>>(da, db, di, dd, n);
    CK(cudaDeviceSynchronize());
    std::vector<uint32_t> ri((size_t)n * N), rd((size_t)n * N);
    CK(cudaMemcpy(ri.data(), di, bytes, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(rd.data(), dd, bytes, cudaMemcpyDeviceToHost));
    cudaFree(da), cudaFree(db), cudaFree(di), cudaFree(dd);
    size_t nonzero = 0;
    for (size_t k = 0; k < ri.size(); k++) {
        nonzero += ri[k] != 0;
        if (ri[k] != rd[k]) {
            printf("%s: MISMATCH at element %zu word %zu\n", name, k / N, k % N);
            return false;
        }
    }
    if (nonzero < ri.size() / 2) {
        printf("%s: suspicious all-zero output\n", name);
        return false;
    }
    printf("%s: %u products equal\n", name, n);
    return true;
}

int main() {
    bool ok = true;
    ok &= check<AltBn128FqParams>("alt_bn128_fq");
    ok &= check<AltBn128FrParams>("alt_bn128_fr");
    ok &= check<Bls12377FqParams>("bls12_377_fq");
    ok &= check<Bls12377FrParams>("bls12_377_fr");
    ok &= check<Bls12381FqParams>("bls12_381_fq");
    ok &= check<Bls12381FrParams>("bls12_381_fr");
    ok &= check<Bw6761FqParams>("bw6_761_fq");
    printf(ok ? "dfma_check ok\n" : "dfma_check FAILED\n");
    return ok ? 0 : 1;
}
