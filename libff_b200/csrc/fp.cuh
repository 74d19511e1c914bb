// Prime-field arithmetic on 32-bit limbs held in registers (sm_100a).
//
// Replaces the reference's Fp_model<n, modulus> hot leaves:
//   mul_reduce      libff/algebra/fields/fp.tcc:50-228   (x86 CIOS asm / GMP)  -> Fp::mul / Fp::sqr
//   operator+= / -= libff/algebra/fields/fp.tcc:350-547                          -> Fp::add / Fp::sub
//   operator-()     libff/algebra/fields/fp.tcc:616-630                          -> Fp::neg
//   invert          libff/algebra/fields/fp.tcc:679-727  (mpn_gcdext)            -> Fp::inv (Fermat)
//   as_bigint       libff/algebra/fields/fp.tcc:270-281                          -> Fp::from_mont
//
// Representation contract: an element is the reference's `mont_repr`, x*R mod p with
// R = 2^(64 n) = 2^(32 N); reinterpreting the host's little-endian 64-bit limbs as 2n 32-bit limbs
// is a no-op, so uploads/downloads need no conversion.  Every routine returns a canonical
// representative in [0, p), exactly like the reference, so outputs compare limb for limb.
//
// Multiplication is a word-serial Montgomery product on two interleaved accumulators ("even"
// columns and "odd" columns).  The a[j]*b_i partial products of even j land on register pairs
// (j, j+1) of one accumulator and those of odd j on pairs of the other, so every
// mad.lo.cc/madc.hi.cc couple is a single carry chain step on an aligned 64-bit register pair;
// ptxas turns each couple into one IMAD.WIDE.U32.X.  The one-limb right shift of the reduction
// step is free: the two accumulators swap roles each row.  All moduli on this path leave at least
// two spare bits in the top limb, which is what keeps the accumulators from overflowing.
#pragma once
#include <cstdint>
#include "field_params.cuh"

namespace b200msm {

#define B200_DEV __device__ __forceinline__
// Fields of at least this many 32-bit limbs call one shared copy of the Montgomery product
// instead of inlining it at every use (instruction-cache footprint against call overhead).
// Squaring of a 24-limb field under the rolled policy: 0 = dedicated squaring inlined (unrolled),
// 1 = the same as one shared called copy, 2 = the rolled product a * a
#ifndef B200_ROLLED_SQR_POLICY
#define B200_ROLLED_SQR_POLICY 0
#endif
#ifndef B200_CALL_MIN_LIMBS
#define B200_CALL_MIN_LIMBS 24
#endif

// Moduli with p = 1 mod 2^32 (so INV = 2^32 - 1): bls12_377 Fq and the scalar fields of both BLS12
// curves.  When ptxas sees those two constants it lowers EVERY mad.lo.cc / madc.hi.cc couple of the
// reduction to an IMAD.X + IMAD.HI.U32.X pair instead of one IMAD.WIDE.U32.X: 361 instead of 241
// multiply-pipe instructions per 12-limb product (cuobjdump -sass; the other moduli fuse with
// immediates).  For these moduli the reduction therefore reads p and INV from constant memory,
// which ptxas cannot fold.
template <class P>
struct ModTab {
    uint32_t v[P::N];
    uint32_t inv;
};
template <class P>
constexpr ModTab<P> make_modtab() {
    ModTab<P> t{};
    for (int i = 0; i < P::N; i++) t.v[i] = P::mod(i);
    t.inv = P::INV;
    return t;
}
template <class P>
__constant__ ModTab<P> c_modtab = make_modtab<P>();

namespace limb {

// acc[j], acc[j+1] = a[j] * bi for even j < N
template <int N>
B200_DEV void row_set(uint32_t *acc, const uint32_t *a, uint32_t bi) {
#pragma unroll
    for (int j = 0; j < N; j += 2) {
        asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(acc[j]) : "r"(a[j]), "r"(bi));
        asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(acc[j + 1]) : "r"(a[j]), "r"(bi));
    }
}

// acc[j], acc[j+1] += a[j] * bi for even j < N, one carry chain; carry-out stays in CC
template <int N>
B200_DEV void row_acc(uint32_t *acc, const uint32_t *a, uint32_t bi) {
    asm volatile("mad.lo.cc.u32 %0, %1, %2, %0;" : "+r"(acc[0]) : "r"(a[0]), "r"(bi));
    asm volatile("madc.hi.cc.u32 %0, %1, %2, %0;" : "+r"(acc[1]) : "r"(a[0]), "r"(bi));
#pragma unroll
    for (int j = 2; j < N; j += 2) {
        asm volatile("madc.lo.cc.u32 %0, %1, %2, %0;" : "+r"(acc[j]) : "r"(a[j]), "r"(bi));
        asm volatile("madc.hi.cc.u32 %0, %1, %2, %0;" : "+r"(acc[j + 1]) : "r"(a[j]), "r"(bi));
    }
}

// acc[j], acc[j+1] = a[j] * bi + acc[j+2], acc[j+3] (+ carry-in from CC): accumulate while
// dropping the two consumed low limbs
template <int N>
B200_DEV void row_acc_shift(uint32_t *acc, const uint32_t *a, uint32_t bi) {
#pragma unroll
    for (int j = 0; j < N - 2; j += 2) {
        asm volatile("madc.lo.cc.u32 %0, %1, %2, %3;" : "=r"(acc[j]) : "r"(a[j]), "r"(bi), "r"(acc[j + 2]));
        asm volatile("madc.hi.cc.u32 %0, %1, %2, %3;"
            : "=r"(acc[j + 1])
            : "r"(a[j]), "r"(bi), "r"(acc[j + 3]));
    }
    asm volatile("madc.lo.cc.u32 %0, %1, %2, 0;" : "=r"(acc[N - 2]) : "r"(a[N - 2]), "r"(bi));
    asm volatile("madc.hi.u32 %0, %1, %2, 0;" : "=r"(acc[N - 1]) : "r"(a[N - 2]), "r"(bi));
}

// row_acc restricted to the even columns j >= J0 (J0 even, J0 < N): the rows of a squaring only
// reach the columns at or above their own index
template <int N, int J0>
B200_DEV void row_acc_from(uint32_t *acc, const uint32_t *a, uint32_t bi) {
    asm volatile("mad.lo.cc.u32 %0, %1, %2, %0;" : "+r"(acc[J0]) : "r"(a[J0]), "r"(bi));
    asm volatile("madc.hi.cc.u32 %0, %1, %2, %0;" : "+r"(acc[J0 + 1]) : "r"(a[J0]), "r"(bi));
#pragma unroll
    for (int j = J0 + 2; j < N; j += 2) {
        asm volatile("madc.lo.cc.u32 %0, %1, %2, %0;" : "+r"(acc[j]) : "r"(a[j]), "r"(bi));
        asm volatile("madc.hi.cc.u32 %0, %1, %2, %0;" : "+r"(acc[j + 1]) : "r"(a[j]), "r"(bi));
    }
}

// row_acc_shift whose products start at column J0: the columns below only move down two limbs
// (with the carry-in from CC rippling through them)
template <int N, int J0>
B200_DEV void row_acc_shift_from(uint32_t *acc, const uint32_t *a, uint32_t bi) {
#pragma unroll
    for (int j = 0; j < N - 2; j += 2) {
        if (j < J0) {
            asm volatile("addc.cc.u32 %0, %1, 0;" : "=r"(acc[j]) : "r"(acc[j + 2]));
            asm volatile("addc.cc.u32 %0, %1, 0;" : "=r"(acc[j + 1]) : "r"(acc[j + 3]));
        } else {
            asm volatile("madc.lo.cc.u32 %0, %1, %2, %3;" : "=r"(acc[j]) : "r"(a[j]), "r"(bi), "r"(acc[j + 2]));
            asm volatile("madc.hi.cc.u32 %0, %1, %2, %3;"
                : "=r"(acc[j + 1])
                : "r"(a[j]), "r"(bi), "r"(acc[j + 3]));
        }
    }
    if (N - 2 < J0) {
        asm volatile("addc.cc.u32 %0, 0, 0;" : "=r"(acc[N - 2]));
        asm volatile("addc.u32 %0, 0, 0;" : "=r"(acc[N - 1]));
    } else {
        asm volatile("madc.lo.cc.u32 %0, %1, %2, 0;" : "=r"(acc[N - 2]) : "r"(a[N - 2]), "r"(bi));
        asm volatile("madc.hi.u32 %0, %1, %2, 0;" : "=r"(acc[N - 1]) : "r"(a[N - 2]), "r"(bi));
    }
}

}  // namespace limb

// CALL_MIN > 0: fields of at least this many limbs call one shared copy of the product (see
// B200_CALL_MIN_LIMBS).  CALL_MIN < 0: fields of at least -CALL_MIN limbs inline a ROLLED product
// (row pairs in a loop) -- small code without the call.  Kernels pick the policy per use through
// Rebind (fp2.cuh): the layout of an element does not depend on it.
template <class P, int CALL_MIN = B200_CALL_MIN_LIMBS>
struct Fp {
    static constexpr int N = P::N;
    static constexpr bool ROLLED = CALL_MIN < 0 && N >= -CALL_MIN;
    static constexpr bool CALLS = CALL_MIN > 0 && N >= CALL_MIN;
    using Params = P;
    // The squaring rows and the shared-reduction product accumulate up to 3p before the shift,
    // which needs p < 2^(32 N - 2); every coordinate field here has the two spare bits, the
    // 255-bit bls12_381 Fr (scalars only) has one and takes the plain product instead.
    static constexpr bool SPARE2 = (P::mod(N - 1) >> 30) == 0;
    static constexpr bool HIDE_MOD = P::mod(0) == 1u;  // see ModTab
    uint32_t v[N];

    B200_DEV static Fp zero() {
        Fp r;
#pragma unroll
        for (int i = 0; i < N; i++) r.v[i] = 0;
        return r;
    }
    B200_DEV static Fp one() {
        Fp r;
#pragma unroll
        for (int i = 0; i < N; i++) r.v[i] = P::one(i);
        return r;
    }
    B200_DEV bool is_zero() const {
        uint32_t o = v[0];
#pragma unroll
        for (int i = 1; i < N; i++) o |= v[i];
        return o == 0;
    }
    B200_DEV bool operator==(const Fp &b) const {
        uint32_t o = v[0] ^ b.v[0];
#pragma unroll
        for (int i = 1; i < N; i++) o |= v[i] ^ b.v[i];
        return o == 0;
    }

    // r = a - p if a >= p else a     (a < 2p)
    B200_DEV static Fp reduce_once(const Fp &a) {
        Fp s;
        asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(s.v[0]) : "r"(a.v[0]), "r"(P::mod(0)));
#pragma unroll
        for (int i = 1; i < N; i++)
            asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(s.v[i]) : "r"(a.v[i]), "r"(P::mod(i)));
        uint32_t borrow;
        asm volatile("subc.u32 %0, 0, 0;" : "=r"(borrow));
        Fp r;
#pragma unroll
        for (int i = 0; i < N; i++) r.v[i] = borrow ? a.v[i] : s.v[i];
        return r;
    }

    // fp.tcc:350-465 -- top limb has spare bits, so a + b never carries out of N limbs
    B200_DEV static Fp add(const Fp &a, const Fp &b) {
        Fp t;
        asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(t.v[0]) : "r"(a.v[0]), "r"(b.v[0]));
#pragma unroll
        for (int i = 1; i < N - 1; i++)
            asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(t.v[i]) : "r"(a.v[i]), "r"(b.v[i]));
        asm volatile("addc.u32 %0, %1, %2;" : "=r"(t.v[N - 1]) : "r"(a.v[N - 1]), "r"(b.v[N - 1]));
        return reduce_once(t);
    }
    B200_DEV static Fp dbl(const Fp &a) { return add(a, a); }

    // fp.tcc:467-547
    B200_DEV static Fp sub(const Fp &a, const Fp &b) {
        Fp t;
        asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(t.v[0]) : "r"(a.v[0]), "r"(b.v[0]));
#pragma unroll
        for (int i = 1; i < N; i++)
            asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(t.v[i]) : "r"(a.v[i]), "r"(b.v[i]));
        uint32_t borrow;
        asm volatile("subc.u32 %0, 0, 0;" : "=r"(borrow));  // 0 or 0xffffffff
        Fp r;
        asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r.v[0]) : "r"(t.v[0]), "r"(P::mod(0) & borrow));
#pragma unroll
        for (int i = 1; i < N - 1; i++)
            asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r.v[i]) : "r"(t.v[i]), "r"(P::mod(i) & borrow));
        asm volatile("addc.u32 %0, %1, %2;"
            : "=r"(r.v[N - 1])
            : "r"(t.v[N - 1]), "r"(P::mod(N - 1) & borrow));
        return r;
    }

    // fp.tcc:616-630 -- the negative of zero is zero
    B200_DEV static Fp neg(const Fp &a) {
        uint32_t nz = a.is_zero() ? 0u : 0xffffffffu;
        Fp r;
        asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(r.v[0]) : "r"(P::mod(0) & nz), "r"(a.v[0]));
#pragma unroll
        for (int i = 1; i < N - 1; i++)
            asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(r.v[i]) : "r"(P::mod(i) & nz), "r"(a.v[i]));
        asm volatile("subc.u32 %0, %1, %2;" : "=r"(r.v[N - 1]) : "r"(P::mod(N - 1) & nz), "r"(a.v[N - 1]));
        return r;
    }
    B200_DEV static Fp cneg(const Fp &a, bool flag) {
        Fp n = neg(a);
        Fp r;
#pragma unroll
        for (int i = 0; i < N; i++) r.v[i] = flag ? n.v[i] : a.v[i];
        return r;
    }
    B200_DEV static Fp select(bool flag, const Fp &a, const Fp &b) {  // flag ? a : b
        Fp r;
#pragma unroll
        for (int i = 0; i < N; i++) r.v[i] = flag ? a.v[i] : b.v[i];
        return r;
    }

    // Second half of every row: m = T[0]*INV; T += m*p; T >>= 32, with T = E + (O << 32).  On
    // return the roles of E and O are swapped for the caller.
    B200_DEV static void reduce_row(uint32_t *E, uint32_t *O) {
        uint32_t mod[N];
#pragma unroll
        for (int i = 0; i < N; i++) mod[i] = HIDE_MOD ? c_modtab<P>.v[i] : P::mod(i);
        uint32_t m = E[0] * (HIDE_MOD ? c_modtab<P>.inv : P::INV);
        limb::row_acc<N>(O, mod + 1, m);
        // O never carries out of N limbs (T < 2^(32(N+1))); this keeps the flag dead
        asm volatile("addc.u32 %0, %0, 0;" : "+r"(O[N - 1]));
        limb::row_acc<N>(E, mod, m);
        asm volatile("addc.u32 %0, %0, 0;" : "+r"(O[N - 1]));
    }

    // One row of the Montgomery product: T += a*bi, then reduce_row.
    B200_DEV static void mont_row(uint32_t *E, uint32_t *O, const uint32_t *a, uint32_t bi,
                                  bool first) {
        if (first) {
            limb::row_set<N>(O, a + 1, bi);
            limb::row_set<N>(E, a, bi);
        } else {
            asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(E[0]) : "r"(O[1]));
            limb::row_acc_shift<N>(O, a + 1, bi);
            limb::row_acc<N>(E, a, bi);
            asm volatile("addc.u32 %0, %0, 0;" : "+r"(O[N - 1]));
        }
        reduce_row(E, O);
    }

    // One row of a sum of two products: T += a*bi + c*di, then reduce_row.  One reduction serves
    // both products; a*b + c*d < 2 p^2 < p R / 2 keeps the result below 2p.
    B200_DEV static void mont_row2(uint32_t *E, uint32_t *O, const uint32_t *a, uint32_t bi,
                                   const uint32_t *c, uint32_t di, bool first) {
        if (first) {
            limb::row_set<N>(O, a + 1, bi);
            limb::row_set<N>(E, a, bi);
        } else {
            asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(E[0]) : "r"(O[1]));
            limb::row_acc_shift<N>(O, a + 1, bi);
            limb::row_acc<N>(E, a, bi);
            asm volatile("addc.u32 %0, %0, 0;" : "+r"(O[N - 1]));
        }
        limb::row_acc<N>(O, c + 1, di);
        asm volatile("addc.u32 %0, %0, 0;" : "+r"(O[N - 1]));  // dead flag, as in reduce_row
        limb::row_acc<N>(E, c, di);
        asm volatile("addc.u32 %0, %0, 0;" : "+r"(O[N - 1]));
        reduce_row(E, O);
    }

    // Row I of a squaring.  a^2 = sum_I 2^(32 I) a_I w_I with w_I = a_I 2^(32 I) + sum_{j>I}
    // 2 a_j 2^(32 j): limb I of w_I is a_I, limb I+1 is limb I+1 of 2a without its lowest bit
    // (that bit is the top bit of a_I), the limbs above are those of 2a.  Row I therefore only
    // reaches columns >= I: N (N + 1) / 2 products instead of N^2 (fp.tcc:632-677 is the
    // reference's squaring; only its 3-limb asm path is specialised there).
    template <int I>
    B200_DEV static void sqr_row(uint32_t *E, uint32_t *O, const uint32_t *a, const uint32_t *a2) {
        uint32_t w[N];
#pragma unroll
        for (int j = 0; j < N; j++) w[j] = a2[j];
        w[I] = a[I];
        if constexpr (I + 1 < N) w[I + 1] = a2[I + 1] & ~1u;
        constexpr int J0E = (I % 2 == 0) ? I : I + 1;  // even columns >= I
        constexpr int J0O = (I % 2 == 1) ? I - 1 : I;  // odd columns >= I, as indices into w + 1
        if constexpr (I == 0) {
            limb::row_set<N>(O, w + 1, a[0]);
            limb::row_set<N>(E, w, a[0]);
        } else {
            asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(E[0]) : "r"(O[1]));
            limb::row_acc_shift_from<N, J0O>(O, w + 1, a[I]);
            if constexpr (J0E < N) {
                limb::row_acc_from<N, J0E>(E, w, a[I]);
                asm volatile("addc.u32 %0, %0, 0;" : "+r"(O[N - 1]));
            }
        }
        reduce_row(E, O);
    }
    template <int I>
    B200_DEV static void sqr_rows(uint32_t *E, uint32_t *O, const uint32_t *a, const uint32_t *a2) {
        if constexpr (I < N) {
            sqr_row<I>(E, O, a, a2);
            sqr_rows<I + 1>(O, E, a, a2);  // the accumulators swap roles every row
        }
    }
    // T = E + (O << 32) after the last row -> canonical element
    B200_DEV static Fp finish(const uint32_t *E, const uint32_t *O) {
        Fp t;
        asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(t.v[0]) : "r"(E[0]), "r"(O[1]));
#pragma unroll
        for (int i = 1; i < N - 1; i++)
            asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(t.v[i]) : "r"(E[i]), "r"(O[i + 1]));
        asm volatile("addc.u32 %0, %1, 0;" : "=r"(t.v[N - 1]) : "r"(E[N - 1]));
        return reduce_once(t);
    }

    // ---- separated product and reduction (lazy reduction in the quadratic extension, fp2.cuh) ----
    // One row of the plain integer product: T += a*bi, no reduction.  The caller takes the finished
    // low limb E[0] after every row; the one-limb shift is the same role swap as in mont_row.
    B200_DEV static void wide_row(uint32_t *E, uint32_t *O, const uint32_t *a, uint32_t bi, bool first) {
        if (first) {
            limb::row_set<N>(O, a + 1, bi);
            limb::row_set<N>(E, a, bi);
        } else {
            asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(E[0]) : "r"(O[1]));
            limb::row_acc_shift<N>(O, a + 1, bi);
            limb::row_acc<N>(E, a, bi);
            asm volatile("addc.u32 %0, %0, 0;" : "+r"(O[N - 1]));
        }
    }
    // T[0 .. 2N) = a * b as integers, any a, b < 2^(32 N) (e.g. unreduced sums of two elements)
    B200_DEV static void mul_wide(uint32_t *T, const uint32_t *a, const uint32_t *b) {
        uint32_t E[N], O[N];
#pragma unroll
        for (int i = 0; i < N; i += 2) {
            wide_row(E, O, a, b[i], i == 0);
            T[i] = E[0];
            wide_row(O, E, a, b[i + 1], false);
            T[i + 1] = O[0];
        }
        asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(T[N]) : "r"(E[0]), "r"(O[1]));
#pragma unroll
        for (int i = 1; i < N - 1; i++)
            asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(T[N + i]) : "r"(E[i]), "r"(O[i + 1]));
        asm volatile("addc.u32 %0, %1, 0;" : "=r"(T[2 * N - 1]) : "r"(E[N - 1]));
    }
    // One row of a reduction without a product: m = T[0]*INV; T = (T + m*p) >> 32.  The shift of
    // the previous row is folded into this row's multiply-adds (acc[j] = p*m + acc[j+2]).
    B200_DEV static void redc_row(uint32_t *E, uint32_t *O, bool first) {
        uint32_t mod[N];
#pragma unroll
        for (int i = 0; i < N; i++) mod[i] = HIDE_MOD ? c_modtab<P>.v[i] : P::mod(i);
        const uint32_t inv = HIDE_MOD ? c_modtab<P>.inv : P::INV;
        if (first) {
            const uint32_t m = E[0] * inv;
            limb::row_set<N>(O, mod + 1, m);
            limb::row_acc<N>(E, mod, m);
            asm volatile("addc.u32 %0, %0, 0;" : "+r"(O[N - 1]));
        } else {
            asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(E[0]) : "r"(O[1]));
            const uint32_t m = E[0] * inv;  // mul.lo leaves the carry flag alone
            limb::row_acc_shift<N>(O, mod + 1, m);
            limb::row_acc<N>(E, mod, m);
            asm volatile("addc.u32 %0, %0, 0;" : "+r"(O[N - 1]));
        }
    }
    // T / R mod p, canonical, for 0 <= T < p R given as 2N limbs: N reduction rows on the low
    // half (their value ends <= p), plus the high half (< p), one conditional subtraction.
    B200_DEV static Fp redc_wide(const uint32_t *T) {
        uint32_t E[N], O[N];
#pragma unroll
        for (int i = 0; i < N; i++) E[i] = T[i];
#pragma unroll
        for (int i = 0; i < N; i += 2) {
            redc_row(E, O, i == 0);
            redc_row(O, E, false);
        }
        Fp t;
        asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(t.v[0]) : "r"(E[0]), "r"(O[1]));
#pragma unroll
        for (int i = 1; i < N - 1; i++)
            asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(t.v[i]) : "r"(E[i]), "r"(O[i + 1]));
        asm volatile("addc.u32 %0, %1, 0;" : "=r"(t.v[N - 1]) : "r"(E[N - 1]));
        asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(t.v[0]) : "r"(T[N]));
#pragma unroll
        for (int i = 1; i < N - 1; i++) asm volatile("addc.cc.u32 %0, %0, %1;" : "+r"(t.v[i]) : "r"(T[N + i]));
        asm volatile("addc.u32 %0, %0, %1;" : "+r"(t.v[N - 1]) : "r"(T[2 * N - 1]));
        return reduce_once(t);
    }
    // a + b as integers (no reduction; canonical a, b: the sum fits, the top limb has spare bits)
    B200_DEV static void add_plain(uint32_t *r, const uint32_t *a, const uint32_t *b) {
        asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r[0]) : "r"(a[0]), "r"(b[0]));
#pragma unroll
        for (int i = 1; i < N - 1; i++) asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r[i]) : "r"(a[i]), "r"(b[i]));
        asm volatile("addc.u32 %0, %1, %2;" : "=r"(r[N - 1]) : "r"(a[N - 1]), "r"(b[N - 1]));
    }
    // X -= Y over 2N limbs; returns the borrow as a mask (0 or 0xffffffff)
    B200_DEV static uint32_t wide_sub(uint32_t *X, const uint32_t *Y) {
        asm volatile("sub.cc.u32 %0, %0, %1;" : "+r"(X[0]) : "r"(Y[0]));
#pragma unroll
        for (int i = 1; i < 2 * N; i++) asm volatile("subc.cc.u32 %0, %0, %1;" : "+r"(X[i]) : "r"(Y[i]));
        uint32_t borrow;
        asm volatile("subc.u32 %0, 0, 0;" : "=r"(borrow));
        return borrow;
    }
    // X += (p & mask) * R: brings a difference of two products in (-p R, 0) back into [0, p R)
    B200_DEV static void wide_add_pR(uint32_t *X, uint32_t mask) {
        asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(X[N]) : "r"(P::mod(0) & mask));
#pragma unroll
        for (int i = 1; i < N - 1; i++)
            asm volatile("addc.cc.u32 %0, %0, %1;" : "+r"(X[N + i]) : "r"(P::mod(i) & mask));
        asm volatile("addc.u32 %0, %0, %1;" : "+r"(X[2 * N - 1]) : "r"(P::mod(N - 1) & mask));
    }
    // the same function as mul through the separated form (parity tests of the two halves)
    B200_DEV static Fp mul_separated(const Fp &a, const Fp &b) {
        uint32_t T[2 * N];
        mul_wide(T, a.v, b.v);
        return redc_wide(T);
    }

    // fp.tcc:50-228: a*b*R^-1 mod p, canonical.  Wide fields call one shared copy of the
    // product (a 24-limb product is ~1300 instructions; inlining a dozen per group operation
    // only thrashes the instruction cache and the compiler).
    B200_DEV static Fp mul(const Fp &a, const Fp &b) {
        if constexpr (ROLLED)
            return mul_rolled(a, b);
        else if constexpr (CALLS)
            return mul_call(a, b);
        else
            return mul_inline(a, b);
    }
    // The same product with the row pairs in a loop (first pair peeled): a third of the code of
    // the fully unrolled form.  The multiplier's limbs rotate through registers so that every
    // index stays static.
    B200_DEV static Fp mul_rolled(const Fp &a, const Fp &b) {
        uint32_t E[N], O[N], bb[N];
#pragma unroll
        for (int i = 0; i < N; i++) bb[i] = b.v[i];
        mont_row(E, O, a.v, bb[0], true);
        mont_row(O, E, a.v, bb[1], false);
#pragma unroll 1
        for (int it = 1; it < N / 2; it++) {
#pragma unroll
            for (int k = 0; k + 2 < N; k++) bb[k] = bb[k + 2];
            mont_row(E, O, a.v, bb[0], false);
            mont_row(O, E, a.v, bb[1], false);
        }
        return finish(E, O);
    }
    __device__ __noinline__ static Fp mul_call(Fp a, Fp b) { return mul_inline(a, b); }  // by value: registers
    B200_DEV static Fp mul_inline(const Fp &a, const Fp &b) {
        uint32_t E[N], O[N];
#pragma unroll
        for (int i = 0; i < N; i += 2) {
            mont_row(E, O, a.v, b.v[i], i == 0);
            mont_row(O, E, a.v, b.v[i + 1], false);
        }
        return finish(E, O);
    }

    // fp.tcc:632-677: a*a*R^-1 mod p with N (N + 1) / 2 + N^2 limb products instead of 2 N^2
    B200_DEV static Fp sqr(const Fp &a) {
        if constexpr (!SPARE2)
            return mul(a, a);
        else if constexpr (ROLLED && N >= 24 && B200_ROLLED_SQR_POLICY == 2)
            return mul_rolled(a, a);
        else if constexpr (CALLS || (ROLLED && N >= 24 && B200_ROLLED_SQR_POLICY == 1))
            return sqr_call(a);
        else
            return sqr_inline(a);
    }
    __device__ __noinline__ static Fp sqr_call(Fp a) { return sqr_inline(a); }
    B200_DEV static Fp sqr_inline(const Fp &a) {
        uint32_t a2[N];  // 2a fits: the top limb has spare bits
        asm volatile("add.cc.u32 %0, %1, %1;" : "=r"(a2[0]) : "r"(a.v[0]));
#pragma unroll
        for (int i = 1; i < N - 1; i++) asm volatile("addc.cc.u32 %0, %1, %1;" : "=r"(a2[i]) : "r"(a.v[i]));
        asm volatile("addc.u32 %0, %1, %1;" : "=r"(a2[N - 1]) : "r"(a.v[N - 1]));
        uint32_t E[N], O[N];
        sqr_rows<0>(E, O, a.v, a2);
        return finish(E, O);
    }

    // (a*b + c*d)*R^-1 mod p, canonical, with ONE reduction: 3 N^2 limb products instead of
    // 4 N^2.  The group law's Y3 = R (Q - X3) - Y1 PPP is such a sum (with -Y1).
    B200_DEV static Fp mul2(const Fp &a, const Fp &b, const Fp &c, const Fp &d) {
        if constexpr (!SPARE2)
            return add(mul(a, b), mul(c, d));
        else if constexpr (ROLLED)
            return mul2_rolled(a, b, c, d);
        else if constexpr (CALLS)
            return mul2_call(a, b, c, d);
        else
            return mul2_inline(a, b, c, d);
    }
    B200_DEV static Fp mul2_rolled(const Fp &a, const Fp &b, const Fp &c, const Fp &d) {
        uint32_t E[N], O[N], bb[N], dd[N];
#pragma unroll
        for (int i = 0; i < N; i++) bb[i] = b.v[i], dd[i] = d.v[i];
        mont_row2(E, O, a.v, bb[0], c.v, dd[0], true);
        mont_row2(O, E, a.v, bb[1], c.v, dd[1], false);
#pragma unroll 1
        for (int it = 1; it < N / 2; it++) {
#pragma unroll
            for (int k = 0; k + 2 < N; k++) bb[k] = bb[k + 2], dd[k] = dd[k + 2];
            mont_row2(E, O, a.v, bb[0], c.v, dd[0], false);
            mont_row2(O, E, a.v, bb[1], c.v, dd[1], false);
        }
        return finish(E, O);
    }
    __device__ __noinline__ static Fp mul2_call(Fp a, Fp b, Fp c, Fp d) {
        return mul2_inline(a, b, c, d);
    }
    B200_DEV static Fp mul2_inline(const Fp &a, const Fp &b, const Fp &c, const Fp &d) {
        uint32_t E[N], O[N];
#pragma unroll
        for (int i = 0; i < N; i += 2) {
            mont_row2(E, O, a.v, b.v[i], c.v, d.v[i], i == 0);
            mont_row2(O, E, a.v, b.v[i + 1], c.v, d.v[i + 1], false);
        }
        return finish(E, O);
    }

    // ---- products whose second operand streams from shared memory ---------------------------
    // Limb k of the streamed operand lies at sb[k * stride] (word-interleaved per-thread slots:
    // conflict-free).  Only `a`, the two accumulators and one limb of b are in registers -- the
    // register-resident product also holds all of b (and rotates it to keep its indices static).
    // Row pairs in a loop, first pair peeled, as mul_rolled.
    B200_DEV static Fp mul_sb_inline(const Fp &a, const uint32_t *sb, int stride) {
        uint32_t E[N], O[N];
        uint32_t b0 = sb[0], b1 = sb[stride];
        uint32_t n0 = sb[2 * (size_t)stride], n1 = sb[3 * (size_t)stride];  // one row pair ahead
        mont_row(E, O, a.v, b0, true);
        mont_row(O, E, a.v, b1, false);
#pragma unroll 1
        for (int it = 1; it < N / 2; it++) {
            b0 = n0, b1 = n1;
            if (it + 1 < N / 2) {
                const uint32_t *q = sb + (size_t)(2 * it + 2) * stride;
                n0 = q[0], n1 = q[stride];
            }
            mont_row(E, O, a.v, b0, false);
            mont_row(O, E, a.v, b1, false);
        }
        return finish(E, O);
    }
    B200_DEV static Fp mul_sb(const Fp &a, const uint32_t *sb, int stride) {
        return mul_sb_inline(a, sb, stride);
    }
    // a * b + c * d with one reduction (see mul2), b and d streamed
    B200_DEV static Fp mul2_sb(const Fp &a, const uint32_t *sb, const Fp &c, const uint32_t *sd,
                               int stride) {
        static_assert(SPARE2, "shared-reduction product needs two spare bits");
        uint32_t E[N], O[N];
        mont_row2(E, O, a.v, sb[0], c.v, sd[0], true);
        mont_row2(O, E, a.v, sb[stride], c.v, sd[stride], false);
#pragma unroll 1
        for (int it = 1; it < N / 2; it++) {
            const uint32_t *q = sb + (size_t)(2 * it) * stride, *r = sd + (size_t)(2 * it) * stride;
            mont_row2(E, O, a.v, q[0], c.v, r[0], false);
            mont_row2(O, E, a.v, q[stride], c.v, r[stride], false);
        }
        return finish(E, O);
    }
    // rows [0, N) of a product whose accumulators E, O the caller owns: lets a fused routine start
    // the next product's first row without re-materialising anything
    B200_DEV static Fp lds(const uint32_t *s, int stride) {
        Fp r;
#pragma unroll
        for (int i = 0; i < N; i++) r.v[i] = s[(size_t)i * stride];
        return r;
    }
    B200_DEV void sts(uint32_t *s, int stride) const {
#pragma unroll
        for (int i = 0; i < N; i++) s[(size_t)i * stride] = v[i];
    }

    // Montgomery -> plain integer (fp.tcc:270-281): one reduction of (a, 0, ..., 0)
    B200_DEV static Fp from_mont(const Fp &a) {
        Fp o = zero();
        o.v[0] = 1;
        return mul(a, o);
    }
    // plain integer (< p) -> Montgomery (fp.tcc:230-235)
    B200_DEV static Fp to_mont(const Fp &a) {
        Fp r2;
#pragma unroll
        for (int i = 0; i < N; i++) r2.v[i] = P::r2(i);
        return mul(a, r2);
    }

    // a^(p-2): the canonical inverse, equal to the reference's invert() output (the inverse in
    // [0, p) is unique).  inv(0) = 0.
    __device__ __noinline__ static Fp inv(Fp a) {
        Fp r = one();
#pragma unroll 1
        for (int i = P::BITS - 1; i >= 0; i--) {
            r = sqr(r);
            uint32_t w = 0;
            // p-2 limb lookup without a dynamically indexed local array
#pragma unroll
            for (int k = 0; k < N; k++)
                if ((i >> 5) == k) w = P::pm2(k);
            if ((w >> (i & 31)) & 1) r = mul(r, a);
        }
        return r;
    }

    B200_DEV static Fp load(const uint32_t *p) {
        Fp r;
        if constexpr (N % 4 == 0) {
#pragma unroll
            for (int i = 0; i < N; i += 4) {
                uint4 q = *reinterpret_cast<const uint4 *>(p + i);
                r.v[i] = q.x, r.v[i + 1] = q.y, r.v[i + 2] = q.z, r.v[i + 3] = q.w;
            }
        } else {
#pragma unroll
            for (int i = 0; i < N; i += 2) {
                uint2 q = *reinterpret_cast<const uint2 *>(p + i);
                r.v[i] = q.x, r.v[i + 1] = q.y;
            }
        }
        return r;
    }
    B200_DEV void store(uint32_t *p) const {
        if constexpr (N % 4 == 0) {
#pragma unroll
            for (int i = 0; i < N; i += 4)
                *reinterpret_cast<uint4 *>(p + i) = make_uint4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        } else {
#pragma unroll
            for (int i = 0; i < N; i += 2) *reinterpret_cast<uint2 *>(p + i) = make_uint2(v[i], v[i + 1]);
        }
    }
};

}  // namespace b200msm
