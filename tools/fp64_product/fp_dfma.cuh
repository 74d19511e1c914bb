// Montgomery product on the FP64 pipe (sm_100a): the second multiplier of the SM.
//
// B200 issues DFMA at the same 64 lanes/clk/SM as IMAD and on a DIFFERENT pipe (microbench:
// dfma 16.4 T/s, dfma interleaved 1:1 with imad 16.9 T pairs/s), so a field multiplication done
// with DFMAs runs beside the integer ones of fp.cuh instead of queueing behind them.  This file
// computes exactly the same function as Fp<P>::mul -- libff's Fp_model::mul_reduce,
// libff/algebra/fields/fp.tcc:50-228: a*b*R^-1 mod p, canonical, R = 2^(32 N) -- with 52-bit limbs
// held in doubles.
//
// Exact 52x52 -> 104-bit products from two round-toward-zero FMAs (N. Emmart's construction):
//   hi = fma_rz(a, b, 2^104)          -> 2^104 + floor(a b / 2^52) 2^52   (ulp of the sum is 2^52)
//   lo = fma_rz(a, b, 2^104 + 2^52 - hi) -> 2^52 + (a b mod 2^52)          (exact)
// so the mantissa fields of hi and lo ARE the two halves.  Their bit patterns are added into
// 64-bit integer columns; the exponent fields are multiples of 2^52 whose total per column is
// known at compile time and pre-subtracted.
//
// R differs from the integer representation's: with M limbs the reduction divides by 2^(52 M),
// not 2^(32 N).  Operand a is therefore split at a bit offset of S = 52 M - 32 N (i.e. a 2^S is
// multiplied), which costs nothing, and a 2^S b < p 2^(52 M) keeps the result below 2p.
#pragma once
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define B200_HD __host__ __device__ __forceinline__
#else
#define B200_HD inline
#include <cmath>
#endif

namespace b200msm {
namespace dfma {

constexpr uint64_t MASK52 = (1ull << 52) - 1;
constexpr uint64_t BIAS_LO = 0x4330000000000000ull;  // bits of 2^52
constexpr uint64_t BIAS_HI = 0x4670000000000000ull;  // bits of 2^104

B200_HD double fma_rz(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
    return __fma_rz(a, b, c);
#else
    return std::fma(a, b, c);  // host checker: caller sets FE_TOWARDZERO
#endif
}
B200_HD uint64_t bits_of(double d) {
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(d);
#else
    uint64_t u;
    std::memcpy(&u, &d, 8);
    return u;
#endif
}
B200_HD double double_of(uint32_t hi, uint32_t lo) {
#if defined(__CUDA_ARCH__)
    return __hiloint2double((int)hi, (int)lo);
#else
    uint64_t u = ((uint64_t)hi << 32) | lo;
    double d;
    std::memcpy(&d, &u, 8);
    return d;
#endif
}
B200_HD uint32_t shr_pair(uint32_t lo, uint32_t hi, int sh) {  // (hi:lo >> sh) low word, 0 < sh < 32
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, sh);
#else
    return (uint32_t)((((uint64_t)hi << 32) | lo) >> sh);
#endif
}

// an integer < 2^52 given as (hi 20 bits, lo 32 bits) -> the double of that value
B200_HD double limb_to_double(uint32_t hi20, uint32_t lo32) {
    return double_of(hi20 | 0x43300000u, lo32) - 4503599627370496.0;
}

}  // namespace dfma

template <class P>
struct FpDfma {
    static constexpr int N = P::N;
    static constexpr int M = (32 * N + 51) / 52;  // 5 / 8 / 15 limbs for 8 / 12 / 24 words
    static constexpr int S = 52 * M - 32 * N;     // 4 / 32 / 12

    // 52-bit limb k of p
    static constexpr uint64_t mod52(int k) {
        uint64_t r = 0;
        for (int bit = 0; bit < 52; bit++) {
            int pos = 52 * k + bit;
            if (pos < 32 * N && ((P::mod(pos >> 5) >> (pos & 31)) & 1)) r |= 1ull << bit;
        }
        return r;
    }
    // -p^-1 mod 2^52
    static constexpr uint64_t pinv52() {
        uint64_t p0 = (uint64_t)P::mod(0) | ((uint64_t)P::mod(1) << 32);
        uint64_t x = p0;  // p0 * x == 1 mod 2^3 for odd p0
        for (int i = 0; i < 6; i++) x *= 2 - p0 * x;
        return (0 - x) & dfma::MASK52;
    }
    // #products (i, j), 0 <= i, j < M, with i + j == k
    static constexpr int cnt(int k) {
        if (k < 0 || k > 2 * M - 2) return 0;
        return (k < M ? k : 2 * M - 2 - k) + 1;
    }
    // exponent-field total that column k receives over the whole product (a b and q p terms)
    static constexpr uint64_t col_bias(int k) {
        return 2ull * (uint64_t)cnt(k) * dfma::BIAS_LO + 2ull * (uint64_t)cnt(k - 1) * dfma::BIAS_HI;
    }

    B200_HD static uint32_t word_at(const uint32_t *a, int i) { return (i >= 0 && i < N) ? a[i] : 0u; }
    // bits [o, o + 32) of the N-word integer a; o is a compile-time constant after unrolling
    B200_HD static uint32_t bits32(const uint32_t *a, int o) {
        int w = o >= 0 ? o / 32 : -((31 - o) / 32);
        int sh = o - 32 * w;
        if (sh == 0) return word_at(a, w);
        return dfma::shr_pair(word_at(a, w), word_at(a, w + 1), sh);
    }
    // 52-bit limb k of (a << shift) as a double
    B200_HD static double limb(const uint32_t *a, int k, int shift) {
        int o = 52 * k - shift;
        return dfma::limb_to_double(bits32(a, o + 32) & 0xFFFFFu, bits32(a, o));
    }

    B200_HD static void mac(uint64_t &c0, uint64_t &c1, double x, double y) {
        double hi = dfma::fma_rz(x, y, 20282409603651670423947251286016.0);            // 2^104
        double lo = dfma::fma_rz(x, y, 20282409603651674927546878656512.0 - hi);       // 2^104 + 2^52
        c0 += dfma::bits_of(lo);
        c1 += dfma::bits_of(hi);
    }

    // r = a * b * 2^(-32 N) mod p in [0, 2p); r, a, b are N 32-bit words (a, b < p)
    B200_HD static void mul_words(uint32_t *r, const uint32_t *a, const uint32_t *b) {
        double ad[M], bd[M], pd[M];
#pragma unroll
        for (int k = 0; k < M; k++) {
            ad[k] = limb(a, k, S);
            bd[k] = limb(b, k, 0);
            pd[k] = (double)mod52(k);
        }
        uint64_t col[2 * M + 1];
#pragma unroll
        for (int k = 0; k < 2 * M; k++) col[k] = 0 - col_bias(k);
#pragma unroll
        for (int i = 0; i < M; i++) {
#pragma unroll
            for (int j = 0; j < M; j++) mac(col[i + j], col[i + j + 1], ad[j], bd[i]);
            uint64_t q = ((col[i] & dfma::MASK52) * pinv52()) & dfma::MASK52;
            double qd = dfma::limb_to_double((uint32_t)(q >> 32), (uint32_t)q);
#pragma unroll
            for (int j = 0; j < M; j++) mac(col[i + j], col[i + j + 1], qd, pd[j]);
            col[i + 1] += col[i] >> 52;
        }
        // T = sum col[M + k] 2^(52 k): normalise to 52-bit limbs, then repack into 32-bit words
        uint64_t L[M + 1], c = 0;
#pragma unroll
        for (int k = 0; k < M; k++) {
            uint64_t v = col[M + k] + c;
            L[k] = k == M - 1 ? v : (v & dfma::MASK52);
            c = v >> 52;
        }
        L[M] = 0;
#pragma unroll
        for (int w = 0; w < N; w++) {
            int k = (32 * w) / 52, off = (32 * w) % 52;
            uint64_t x = L[k] >> off;
            if (off > 20) x |= L[k + 1] << (52 - off);
            r[w] = (uint32_t)x;
        }
    }
};

}  // namespace b200msm
