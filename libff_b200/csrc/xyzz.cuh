// Group arithmetic for short-Weierstrass curves with a = 0 in extended Jacobian ("XYZZ")
// coordinates: x = X/ZZ, y = Y/ZZZ with ZZ^3 = ZZZ^2; infinity <=> ZZ = 0.
//
// This is the device replacement for the reference's per-curve group law
//   alt_bn128_G1::mixed_add / add / dbl   libff/algebra/curves/alt_bn128/alt_bn128_g1.cpp:149-326
//   alt_bn128_G2::mixed_add / add / dbl   libff/algebra/curves/alt_bn128/alt_bn128_g2.cpp:166-349
//   bls12_377_G1::mixed_add / add / dbl   libff/algebra/curves/bls12_377/bls12_377_g1.cpp:121-352
//   bw6_761_G1::mixed_add / add / dbl     libff/algebra/curves/bw6_761/bw6_761_g1.cpp:111-358
// The reference works in Jacobian (or, for bw6_761, homogeneous projective) coordinates; the
// group element computed is the same, only its representative differs, and every result leaves
// the device converted back to the reference's coordinate system (to_jacobian / to_projective)
// so that libff's operator== and to_affine_coordinates() see the identical point.
//
// Exceptional cases handled exactly as the reference does (alt_bn128_g1.cpp:149-283): either
// operand at infinity, P + P (falls to doubling), P + (-P) (infinity).
// Formulas: EFD madd-2008-s (8M+2S), add-2008-s (12M+2S), dbl-2008-s-1 / mdbl-2008-s-1 with a = 0.
#pragma once
#include "fp.cuh"
#include "fp2.cuh"

namespace b200msm {

template <class F>
struct Affine {
    F x, y;  // (0, 0) encodes infinity on the device (not on any curve here: b != 0)
    static constexpr int WORDS = 2 * F::N;
    B200_DEV bool is_inf() const { return x.is_zero() && y.is_zero(); }
    B200_DEV static Affine load(const uint32_t *p) { return Affine{F::load(p), F::load(p + F::N)}; }
    B200_DEV void store(uint32_t *p) const {
        x.store(p);
        y.store(p + F::N);
    }
};

template <class F>
struct XYZZ {
    F X, Y, ZZ, ZZZ;
    static constexpr int WORDS = 4 * F::N;

    B200_DEV static XYZZ inf() { return XYZZ{F::zero(), F::one(), F::zero(), F::zero()}; }
    B200_DEV bool is_inf() const { return ZZ.is_zero(); }
    B200_DEV static XYZZ from_affine(const Affine<F> &p) {
        if (p.is_inf()) return inf();
        return XYZZ{p.x, p.y, F::one(), F::one()};
    }
    B200_DEV static XYZZ load(const uint32_t *p) {
        return XYZZ{F::load(p), F::load(p + F::N), F::load(p + 2 * F::N), F::load(p + 3 * F::N)};
    }
    B200_DEV void store(uint32_t *p) const {
        X.store(p);
        Y.store(p + F::N);
        ZZ.store(p + 2 * F::N);
        ZZZ.store(p + 3 * F::N);
    }
    B200_DEV static XYZZ neg(const XYZZ &p) { return XYZZ{p.X, F::neg(p.Y), p.ZZ, p.ZZZ}; }

    // 2 * (x, y), affine input, not infinity
    __device__ __noinline__ static XYZZ dbl_affine(F x, F y) {
        F U = F::dbl(y);
        F V = F::sqr(U);
        F W = F::mul(U, V);
        F S = F::mul(x, V);
        F xx = F::sqr(x);
        F M = F::add(F::dbl(xx), xx);
        F X3 = F::sub(F::sqr(M), F::dbl(S));
        F Y3 = F::mul2(M, F::sub(S, X3), F::neg(W), y);
        return XYZZ{X3, Y3, V, W};
    }

    __device__ __noinline__ static XYZZ dbl(XYZZ p) {
        if (p.is_inf()) return p;
        F U = F::dbl(p.Y);
        F V = F::sqr(U);
        F W = F::mul(U, V);
        F S = F::mul(p.X, V);
        F xx = F::sqr(p.X);
        F M = F::add(F::dbl(xx), xx);
        F X3 = F::sub(F::sqr(M), F::dbl(S));
        F Y3 = F::mul2(M, F::sub(S, X3), F::neg(W), p.Y);
        return XYZZ{X3, Y3, F::mul(V, p.ZZ), F::mul(W, p.ZZZ)};
    }

    // acc += (x2, y2); the affine operand must not be infinity (callers filter it)
    B200_DEV static void madd(XYZZ &acc, const F &x2, const F &y2) {
        if (acc.is_inf()) {
            acc = XYZZ{x2, y2, F::one(), F::one()};
            return;
        }
        F Pd = F::sub(F::mul(x2, acc.ZZ), acc.X);
        F R = F::sub(F::mul(y2, acc.ZZZ), acc.Y);
        if (Pd.is_zero()) {
            if (R.is_zero())
                acc = dbl_affine(x2, y2);
            else
                acc = inf();
            return;
        }
        F PP = F::sqr(Pd);
        F PPP = F::mul(Pd, PP);
        F Q = F::mul(acc.X, PP);
        F X3 = F::sub(F::sub(F::sqr(R), PPP), F::dbl(Q));
        F Y3 = F::mul2(R, F::sub(Q, X3), F::neg(acc.Y), PPP);  // one reduction for both products
        acc.X = X3;
        acc.Y = Y3;
        acc.ZZ = F::mul(acc.ZZ, PP);
        acc.ZZZ = F::mul(acc.ZZZ, PPP);
    }

    // acc += q
    // (out of line with the points BY VALUE: what fits travels in registers instead of through
    // local memory)
    __device__ __noinline__ static XYZZ add_value(XYZZ acc, XYZZ q) {
        add_inline(acc, q);
        return acc;
    }
    B200_DEV static void add(XYZZ &acc, const XYZZ &q) { acc = add_value(acc, q); }
    // the same inlined at the call site: no operands through the stack (the out-of-line routine
    // takes its points by reference, i.e. from local memory); for kernels that have ONE addition
    // site and a field short enough for the code to stay cached
    B200_DEV static void add_inline(XYZZ &acc, const XYZZ &q) {
        if (q.is_inf()) return;
        if (acc.is_inf()) {
            acc = q;
            return;
        }
        F U1 = F::mul(acc.X, q.ZZ);
        F S1 = F::mul(acc.Y, q.ZZZ);
        F Pd = F::sub(F::mul(q.X, acc.ZZ), U1);
        F R = F::sub(F::mul(q.Y, acc.ZZZ), S1);
        if (Pd.is_zero()) {
            if (R.is_zero())
                acc = dbl(acc);
            else
                acc = inf();
            return;
        }
        F PP = F::sqr(Pd);
        F PPP = F::mul(Pd, PP);
        F Q = F::mul(U1, PP);
        F X3 = F::sub(F::sub(F::sqr(R), PPP), F::dbl(Q));
        F Y3 = F::mul2(R, F::sub(Q, X3), F::neg(S1), PPP);
        acc.X = X3;
        acc.Y = Y3;
        acc.ZZ = F::mul(F::mul(acc.ZZ, q.ZZ), PP);
        acc.ZZZ = F::mul(F::mul(acc.ZZZ, q.ZZZ), PPP);
    }
    B200_DEV static XYZZ select(bool flag, const XYZZ &a, const XYZZ &b) {  // flag ? a : b
        return XYZZ{F::select(flag, a.X, b.X), F::select(flag, a.Y, b.Y), F::select(flag, a.ZZ, b.ZZ),
                    F::select(flag, a.ZZZ, b.ZZZ)};
    }

    // Representatives in the reference's coordinate systems (no inversion needed):
    //   Jacobian   (X ZZ, Y ZZZ, ZZ):         x = X ZZ / ZZ^2,  y = Y ZZZ / ZZ^3
    //   projective (X ZZZ, Y ZZ, ZZ ZZZ):     x = X ZZZ / (ZZ ZZZ), y = Y ZZ / (ZZ ZZZ)
    // Infinity is written as the reference's canonical zero (0, 1, 0)
    // (alt_bn128_init.cpp:287-288, bw6_761_init.cpp:285-286).
    B200_DEV void to_jacobian(F &X_, F &Y_, F &Z_) const {
        if (is_inf()) {
            X_ = F::zero(), Y_ = F::one(), Z_ = F::zero();
            return;
        }
        X_ = F::mul(X, ZZ);
        Y_ = F::mul(Y, ZZZ);
        Z_ = ZZ;
    }
    B200_DEV void to_projective(F &X_, F &Y_, F &Z_) const {
        if (is_inf()) {
            X_ = F::zero(), Y_ = F::one(), Z_ = F::zero();
            return;
        }
        X_ = F::mul(X, ZZZ);
        Y_ = F::mul(Y, ZZ);
        Z_ = F::mul(ZZ, ZZZ);
    }
};

// k * p for a small non-negative k (double-and-add, MSB first)
template <class F>
__device__ __noinline__ XYZZ<F> small_mul(XYZZ<F> p, uint32_t k) {
    XYZZ<F> r = XYZZ<F>::inf();
    if (k == 0 || p.is_inf()) return r;
    for (int bit = 31 - __clz(k); bit >= 0; bit--) {
        r = XYZZ<F>::dbl(r);
        if ((k >> bit) & 1) XYZZ<F>::add(r, p);
    }
    return r;
}

}  // namespace b200msm
