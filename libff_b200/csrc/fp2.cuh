// Quadratic extension Fq2 = Fq[u]/(u^2 + NR) for G2 coordinates, NR a small positive integer:
// alt_bn128 and bls12_381 u^2 = -1 (alt_bn128_init.cpp:138-140, bls12_381_init.cpp:137-139),
// bls12_377 u^2 = -5 (bls12_377_init.cpp:170-176).
//
// Replaces Fp2_model (libff/algebra/fields/fp2.tcc): mul :101-114 (Karatsuba), squared_complex
// :141-151, inverse :153-168, add/sub/neg :79-120.  The reference multiplies by `non_residue`
// with a full Fq multiplication; here -NR x is a few additions and a subtraction -- same value,
// canonical result.
// Memory layout = the reference's: c0 limbs then c1 limbs (fp2.hpp `Fp_model coeffs[2]`).
#pragma once
#include "fp.cuh"

namespace b200msm {

#ifndef B200_FP2_LAZY
#define B200_FP2_LAZY 1  // 0: every base-field product of an Fq2 product is reduced on its own
#endif
template <class P, int NR = 1, int CALL_MIN = B200_CALL_MIN_LIMBS>
struct Fp2 {
    using F = Fp<P, CALL_MIN>;
    static constexpr int N = 2 * P::N;  // 32-bit words per element
    using Params = P;
    F c0, c1;

    B200_DEV static Fp2 zero() { return Fp2{F::zero(), F::zero()}; }
    B200_DEV static Fp2 one() { return Fp2{F::one(), F::zero()}; }
    B200_DEV bool is_zero() const { return c0.is_zero() && c1.is_zero(); }
    B200_DEV bool operator==(const Fp2 &b) const { return c0 == b.c0 && c1 == b.c1; }
    B200_DEV static Fp2 add(const Fp2 &a, const Fp2 &b) {
        return Fp2{F::add(a.c0, b.c0), F::add(a.c1, b.c1)};
    }
    B200_DEV static Fp2 dbl(const Fp2 &a) { return add(a, a); }
    B200_DEV static Fp2 sub(const Fp2 &a, const Fp2 &b) {
        return Fp2{F::sub(a.c0, b.c0), F::sub(a.c1, b.c1)};
    }
    B200_DEV static Fp2 neg(const Fp2 &a) { return Fp2{F::neg(a.c0), F::neg(a.c1)}; }
    B200_DEV static Fp2 cneg(const Fp2 &a, bool f) { return Fp2{F::cneg(a.c0, f), F::cneg(a.c1, f)}; }
    B200_DEV static Fp2 select(bool f, const Fp2 &a, const Fp2 &b) {
        return Fp2{F::select(f, a.c0, b.c0), F::select(f, a.c1, b.c1)};
    }
    // NR * x by additions (NR is 1 or 5)
    B200_DEV static F times_nr(const F &x) {
        if constexpr (NR == 1) {
            return x;
        } else {
            static_assert(NR == 5, "non-residue -1 or -5");
            return F::add(F::dbl(F::dbl(x)), x);
        }
    }
    // (a0 + a1 u)(b0 + b1 u) = (a0 b0 - NR a1 b1) + ((a0 + a1)(b0 + b1) - a0 b0 - a1 b1) u
    // (out of line, operands BY VALUE: they travel in registers; by reference they would sit in
    // local memory -- 224 local loads / stores around every call in a product chain)
    // Karatsuba with LAZY reduction: the three base-field products stay unreduced 2N-limb
    // integers, the two coordinates are formed there and reduced once each -- 3 N^2 + 2 (N^2 + N)
    // limb products instead of 3 (2 N^2 + N).  Bounds: sums of two canonical elements are < 2p,
    // so v2 < 4 p^2 < p R (two spare bits); v0 - NR v1 > -NR p^2 > -p R (NR = 5 only for
    // bls12_377 with seven spare bits), brought back into [0, p R) by adding p R once when the
    // difference is negative.  Measured (profiles/policy_sweep_r02b.log): alt_bn128 G2 2^20
    // 12.88 -> 11.47 ms.
    B200_DEV static Fp2 mul_lazy(const Fp2 &a, const Fp2 &b) {
        constexpr int BN = P::N;
        uint32_t v0[2 * BN], v1[2 * BN], v2[2 * BN], sa[BN], sb[BN];
        F::mul_wide(v0, a.c0.v, b.c0.v);
        F::mul_wide(v1, a.c1.v, b.c1.v);
        F::add_plain(sa, a.c0.v, a.c1.v);
        F::add_plain(sb, b.c0.v, b.c1.v);
        F::mul_wide(v2, sa, sb);
        F::wide_sub(v2, v0);
        F::wide_sub(v2, v1);  // a0 b1 + a1 b0 >= 0
        const F c1 = F::redc_wide(v2);
        uint32_t borrow = F::wide_sub(v0, v1);
        if constexpr (NR == 5) {
            // v0 - 5 v1 = (v0 - v1) - 4 v1 > -p R: at most one of the two subtractions wraps
            // (once the value is negative, 2^(64 N) - 5 p^2 > 0 keeps the second from borrowing)
#pragma unroll
            for (int i = 2 * BN - 1; i > 0; i--) v1[i] = __funnelshift_l(v1[i - 1], v1[i], 2);
            v1[0] <<= 2;
            borrow |= F::wide_sub(v0, v1);
        }
        F::wide_add_pR(v0, borrow);
        return Fp2{F::redc_wide(v0), c1};
    }
    __device__ __noinline__ static Fp2 mul(Fp2 a, Fp2 b) {
        if constexpr (B200_FP2_LAZY != 0) return mul_lazy(a, b);
        F v0 = F::mul(a.c0, b.c0);
        F v1 = F::mul(a.c1, b.c1);
        F s = F::mul(F::add(a.c0, a.c1), F::add(b.c0, b.c1));
        return Fp2{F::sub(v0, times_nr(v1)), F::sub(F::sub(s, v0), v1)};
    }
    // a b + c d.  Sharing the two reductions between the products as well (six unreduced
    // products, two reductions, one out-of-line routine with four operands by value) measured
    // slower than two lazy products: alt_bn128 G2 2^20 11.43 -> 11.69 ms, bls12_381 G2 27.35 -> 28.35.
    B200_DEV static Fp2 mul2(const Fp2 &a, const Fp2 &b, const Fp2 &c, const Fp2 &d) {
        return add(mul(a, b), mul(c, d));
    }
    // (a0 + a1 u)^2 = (a0 + a1)(a0 - NR a1) + (NR - 1) a0 a1 + 2 a0 a1 u   (fp2.tcc:141-151)
    __device__ __noinline__ static Fp2 sqr(Fp2 a) {
        F t = F::mul(a.c0, a.c1);
        F c0 = F::mul(F::add(a.c0, a.c1), F::sub(a.c0, times_nr(a.c1)));
        if constexpr (NR == 5) c0 = F::add(c0, F::dbl(F::dbl(t)));
        return Fp2{c0, F::dbl(t)};
    }
    // 1 / (a0 + a1 u) = (a0 - a1 u) / (a0^2 + NR a1^2)
    __device__ __noinline__ static Fp2 inv(Fp2 a) {
        F d = F::inv(F::add(F::sqr(a.c0), times_nr(F::sqr(a.c1))));
        return Fp2{F::mul(a.c0, d), F::neg(F::mul(a.c1, d))};
    }
    B200_DEV static Fp2 load(const uint32_t *p) { return Fp2{F::load(p), F::load(p + P::N)}; }
    B200_DEV void store(uint32_t *p) const {
        c0.store(p);
        c1.store(p + P::N);
    }
};

// The same field with another inline / call policy for its base-field products (CALL_MIN limbs).
template <class F, int CM>
struct Rebind;
template <class P, int X, int CM>
struct Rebind<Fp<P, X>, CM> {
    using type = Fp<P, CM>;
};
template <class P, int NR, int X, int CM>
struct Rebind<Fp2<P, NR, X>, CM> {
    using type = Fp2<P, NR, CM>;
};
template <class F>
struct BaseField;
template <class P, int X>
struct BaseField<Fp<P, X>> {
    using type = Fp<P, X>;
};
template <class P, int NR, int X>
struct BaseField<Fp2<P, NR, X>> {
    using type = Fp<P, X>;
};
template <class F>
struct BaseLimbs;
template <class P, int X>
struct BaseLimbs<Fp<P, X>> {
    static constexpr int N = P::N;
    static constexpr bool EXT = false;
};
template <class P, int NR, int X>
struct BaseLimbs<Fp2<P, NR, X>> {
    static constexpr int N = P::N;
    static constexpr bool EXT = true;
};
// Tail kernels (carry rounds, bucket reduction, window totals) are latency- and instruction-cache
// bound: with 12-limb products inlined they are 270 KB of code; calling one shared copy is
// faster there (bls12_377 G1 2^22: carry rounds 2.65 -> 1.82 ms, bucket reduction 2.35 -> 1.73).
#ifndef B200_TAIL_CALL_MIN_LIMBS
#define B200_TAIL_CALL_MIN_LIMBS 12
#endif
#ifndef B200_TAIL_ROLLED_MIN_LIMBS
// prime fields from this many limbs inline the ROLLED product instead.  That won (r01) while
// called products took their operands through local memory; with operands by value the call is
// faster for both wide classes (profiles/policy_sweep_r02b.log: tail kernels of bls12_377 G1 2^22
// 4.68 -> 4.25 ms, bw6_761 G1 2^19 7.41 -> 6.94 ms), so no field rolls its tail products now.
#define B200_TAIL_ROLLED_MIN_LIMBS 99
#endif
template <class F>
using TailField = typename Rebind<F, ((!BaseLimbs<F>::EXT && BaseLimbs<F>::N >= B200_TAIL_ROLLED_MIN_LIMBS)
                                          ? -BaseLimbs<F>::N
                                      : BaseLimbs<F>::N >= B200_CALL_MIN_LIMBS      ? B200_CALL_MIN_LIMBS
                                      : BaseLimbs<F>::N >= B200_TAIL_CALL_MIN_LIMBS ? B200_TAIL_CALL_MIN_LIMBS
                                                                                    : B200_CALL_MIN_LIMBS)>::type;
// Product policy of the ACCUMULATION kernels, per field class -- every entry measured, and
// re-measured in r02 after out-of-line routines took their operands by value (before that a call
// moved its operands through local memory, which is what made inlined and rolled forms win in r01):
//   8-limb prime fields        unrolled inline
//   12-limb prime fields       called (B200_ACCUM_PRIME12_POLICY); r01 rolled bls12_377 Fq because
//                              its unrolled kernel ran at 75 % where bls12_381's ran at 85 %: the
//                              cause was the reduction's lowering for moduli = 1 mod 2^32 (fp.cuh,
//                              ModTab), not instruction fetch
//   24-limb prime field        the shared-memory slot kernel (msm_kernels.cuh); AccumRolled only
//                              selects the product of its register fallback / merging variant
//   quadratic extensions       Fq2 product called, lazy reduction inside (mul_lazy above)
template <class P>
struct AccumRolled {
    static constexpr bool value = false;
};
#ifndef B200_ACCUM_ROLLED_377
#define B200_ACCUM_ROLLED_377 0
#endif
template <>
struct AccumRolled<Bls12377FqParams> {
    static constexpr bool value = B200_ACCUM_ROLLED_377 != 0;
};
// bw6_761 Fq (24 limbs), register kernels only: rolled inline (a by-value call of two 24-limb
// operands still goes through the stack: 2^19 accumulation 30.5 -> 25.5 ms in r01)
template <>
struct AccumRolled<Bw6761FqParams> {
    static constexpr bool value = true;
};
#ifndef B200_ACCUM_EXT12_POLICY
// base-field products of the 12-limb quadratic extensions outside the lazy product (squarings,
// inversions): called
#define B200_ACCUM_EXT12_POLICY 12
#endif
#ifndef B200_ACCUM_PRIME12_POLICY
// 12-limb prime fields call their products in the accumulation too, now that operands travel by
// value: the kernel shrinks to a third and no longer misses the instruction cache (bls12_377 G1
// 2^22: accumulation 22.85 -> 22.21 ms; bls12_381 G1 unchanged at 23.7)
#define B200_ACCUM_PRIME12_POLICY 12
#endif
template <class F>
using AccumField = typename Rebind<F, (BaseLimbs<F>::EXT && BaseLimbs<F>::N >= 12
                                           ? (AccumRolled<typename F::Params>::value ? -12 : B200_ACCUM_EXT12_POLICY)
                                       : (!BaseLimbs<F>::EXT && AccumRolled<typename F::Params>::value)
                                           ? -BaseLimbs<F>::N
                                       : (!BaseLimbs<F>::EXT && BaseLimbs<F>::N == 12) ? B200_ACCUM_PRIME12_POLICY
                                           : B200_CALL_MIN_LIMBS)>::type;

}  // namespace b200msm
