// Batched normalisation and fixed-base exponentiation kernels.
//
// batch_to_special (libff/algebra/scalar_multiplication/multiexp.tcc:949-974) ->
//   G::batch_to_special_all_non_zeros (alt_bn128_g1.cpp:456-477, alt_bn128_g2.cpp:486-507,
//   bls12_377_g1.cpp:512-533, bw6_761_g1.cpp:489-506) -> batch_invert (field_utils.tcc:419-439).
// The reference's Montgomery trick is one serial chain over the whole vector; here it is a
// tree of chunks: every thread multiplies up CHUNK consecutive values (k_inv_up), the chunk
// products are inverted recursively, and the way down (k_inv_down) turns prefix products into
// individual inverses.  One field inversion (Fermat) happens at the root.  Affine coordinates are
// canonical, so the output equals the reference's limb for limb.
//
// batch_exp / windowed_exp (multiexp.tcc:848-915): one thread per scalar adds one table entry per
// window row.
#pragma once
#include "curves.cuh"

namespace b200msm {

constexpr int INV_CHUNK = 16;

// z[i] = Z coordinate of point i, or 1 where the point is infinity (its inverse is never used)
template <class C>
__global__ void k_extract_z(const uint32_t *__restrict__ raw, uint32_t *__restrict__ z, uint32_t n) {
    using R = Raw<C>;
    using F = typename C::Fq;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    F X, Y, Z;
    R::load(raw + (size_t)i * R::PW, X, Y, Z);
    if (R::is_inf(X, Z)) Z = F::one();
    Z.store(z + (size_t)i * F::N);
}

// pre[i] = product of the values before i inside its chunk; prod[t] = product of chunk t
template <class F>
__global__ void k_inv_up(const uint32_t *__restrict__ val, uint32_t *__restrict__ pre,
                         uint32_t *__restrict__ prod, uint32_t n) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t lo = (uint64_t)t * INV_CHUNK;
    if (lo >= n) return;
    uint32_t hi = (uint32_t)min((uint64_t)n, lo + INV_CHUNK);
    F run = F::one();
    for (uint32_t i = (uint32_t)lo; i < hi; i++) {
        run.store(pre + (size_t)i * F::N);
        run = F::mul(run, F::load(val + (size_t)i * F::N));
    }
    run.store(prod + (size_t)t * F::N);
}

// prod[t] now holds the inverse of the chunk product: val[i] <- 1 / val[i]
template <class F>
__global__ void k_inv_down(uint32_t *__restrict__ val, const uint32_t *__restrict__ pre,
                           const uint32_t *__restrict__ prod, uint32_t n) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t lo = (uint64_t)t * INV_CHUNK;
    if (lo >= n) return;
    uint32_t hi = (uint32_t)min((uint64_t)n, lo + INV_CHUNK);
    F inv = F::load(prod + (size_t)t * F::N);
    for (uint32_t i = hi; i-- > (uint32_t)lo;) {
        F v = F::load(val + (size_t)i * F::N);
        F::mul(inv, F::load(pre + (size_t)i * F::N)).store(val + (size_t)i * F::N);
        inv = F::mul(inv, v);
    }
}

// root of the tree: n <= INV_CHUNK values, one thread, one field inversion
template <class F>
__global__ void k_inv_root(uint32_t *__restrict__ val, uint32_t *__restrict__ pre, uint32_t n) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    F run = F::one();
    for (uint32_t i = 0; i < n; i++) {
        run.store(pre + (size_t)i * F::N);
        run = F::mul(run, F::load(val + (size_t)i * F::N));
    }
    F inv = F::inv(run);
    for (uint32_t i = n; i-- > 0;) {
        F v = F::load(val + (size_t)i * F::N);
        F::mul(inv, F::load(pre + (size_t)i * F::N)).store(val + (size_t)i * F::N);
        inv = F::mul(inv, v);
    }
}

// Applies zinv[i] = 1/Z_i.  Jacobian: (X/Z^2, Y/Z^3, 1); projective: (X/Z, Y/Z, 1); infinity ->
// (0, 1, 0).  Writes reference layout to out_raw (may alias raw) and/or device affine to aff.
template <class C>
__global__ void k_apply_zinv(const uint32_t *raw, const uint32_t *__restrict__ zinv,
                             uint32_t *out_raw, uint32_t *__restrict__ aff,
                             uint8_t *__restrict__ inf_flag, uint32_t n) {
    using R = Raw<C>;
    using F = typename C::Fq;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    F X, Y, Z;
    R::load(raw + (size_t)i * R::PW, X, Y, Z);
    bool inf = R::is_inf(X, Z);
    F x, y;
    if (inf) {
        x = F::zero(), y = F::one();
    } else {
        F zi = F::load(zinv + (size_t)i * F::N);
        if constexpr (C::PROJECTIVE) {
            x = F::mul(X, zi);
            y = F::mul(Y, zi);
        } else {
            F zi2 = F::sqr(zi);
            x = F::mul(X, zi2);
            y = F::mul(Y, F::mul(zi2, zi));
        }
    }
    if (out_raw) R::store(out_raw + (size_t)i * R::PW, x, y, inf ? F::zero() : F::one());
    if (aff) {
        Affine<F> a{x, inf ? F::zero() : y};
        a.store(aff + (size_t)i * R::AW);
        inf_flag[i] = inf;
    }
}

// windowed_exp (multiexp.tcc:848-873): out[i] = sum_r table[r][bits(s_i, r*w, w)], table already
// normalised to device affine, row-major rows x row_len.  With `coeff` the exponent is
// coeff * v[i] (batch_exp_with_coeff, multiexp.tcc:917-947: one Fr product ahead of the digits).
template <class C>
__global__ void __launch_bounds__(128)
    k_batch_exp(const uint32_t *__restrict__ scalars, const uint32_t *__restrict__ coeff,
                const uint32_t *__restrict__ table_aff, uint32_t rows, uint32_t row_len, int window,
                uint32_t *__restrict__ out_raw, uint32_t n) {
    using R = Raw<C>;
    using F = typename C::Fq;
    using Fr = typename C::Fr;
    using X = XYZZ<F>;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr s = Fr::load(scalars + (size_t)i * Fr::N);
    if (coeff != nullptr) s = Fr::mul(Fr::load(coeff), s);  // Montgomery forms: stays one
    s = Fr::from_mont(s);
    const uint32_t mask = (1u << window) - 1;
    X acc = X::inf();
    // the next row's entry is fetched while the current addition runs
    uint32_t d = s.v[0] & mask;
    Affine<F> p = Affine<F>::load(table_aff + (size_t)(d < row_len ? d : 0) * R::AW);
    for (uint32_t r = 0; r < rows; r++) {
        const Affine<F> pcur = p;
#pragma unroll
        for (int k = 0; k < Fr::N - 1; k++) s.v[k] = __funnelshift_r(s.v[k], s.v[k + 1], window);
        s.v[Fr::N - 1] >>= window;
        if (r + 1 < rows) {
            d = s.v[0] & mask;
            p = Affine<F>::load(table_aff + ((size_t)(r + 1) * row_len + (d < row_len ? d : 0)) * R::AW);
        }
        if (!pcur.is_inf()) X::madd(acc, pcur.x, pcur.y);
    }
    F Xo, Yo, Zo;
    R::from_xyzz(acc, Xo, Yo, Zo);
    R::store(out_raw + (size_t)i * R::PW, Xo, Yo, Zo);
}

// the same point over the same field with another inline / call policy (identical word layout)
template <class XB, class XA>
B200_DEV XB xyzz_cast(const XA &p) {
    uint32_t tmp[XA::WORDS];
    p.store(tmp);
    return XB::load(tmp);
}

// ---- get_window_table on the device (multiexp.tcc:809-846) --------------------------------------
// row_base[r] = 2^(r w) g: thread r walks its own r w doublings (the longest chain, scalar_bits
// doublings, is what the reference's serial loop walks too, :838-840).
template <class C>
__global__ void k_table_row_bases(const uint32_t *__restrict__ g_raw, int window, uint32_t rows,
                                  uint32_t *__restrict__ row_base) {
    using R = Raw<C>;
    using F = typename C::Fq;
    using X = XYZZ<TailField<F>>;
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    F Xi, Yi, Zi;
    R::load(g_raw, Xi, Yi, Zi);
    X p = xyzz_cast<X>(R::to_xyzz(Xi, Yi, Zi));
    for (uint32_t k = 0; k < r * (uint32_t)window; k++) p = X::dbl(p);
    p.store(row_base + (size_t)r * R::XW);
}

// entries [r][j0 .. j0 + TABLE_RUN): the first by double-and-add (j0 * row_base[r]), the rest by
// running additions of row_base[r] -- the reference's inner loop (:830-833) cut into runs.
constexpr uint32_t TABLE_RUN = 64;
template <class C>
__global__ void __launch_bounds__(64)
    k_table_fill(const uint32_t *__restrict__ row_base, uint32_t rows, uint32_t row_len,
                 uint32_t last_len, uint32_t *__restrict__ table_raw) {
    using R = Raw<C>;
    using F = TailField<typename C::Fq>;
    using X = XYZZ<F>;
    const uint32_t runs = (row_len + TABLE_RUN - 1) / TABLE_RUN;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= rows * runs) return;
    const uint32_t r = t / runs, j0 = (t % runs) * TABLE_RUN;
    const uint32_t len = r + 1 == rows ? last_len : row_len;  // filled entries of this row
    const X base = X::load(row_base + (size_t)r * R::XW);
    X cur = small_mul<F>(base, j0);
    for (uint32_t j = j0; j < j0 + TABLE_RUN && j < row_len; j++) {
        uint32_t *dst = table_raw + ((size_t)r * row_len + j) * R::PW;
        typename C::Fq Xo, Yo, Zo;
        if (j < len) {
            R::from_xyzz(xyzz_cast<XYZZ<typename C::Fq>>(cur), Xo, Yo, Zo);
            X::add(cur, base);
        } else {  // T::zero() tail of the last row
            Xo = C::Fq::zero(), Yo = C::Fq::one(), Zo = C::Fq::zero();
        }
        R::store(dst, Xo, Yo, Zo);
    }
}

// device affine ((0, 0) = zero) -> the reference's special form (x, y, 1) / (0, 1, 0)
template <class C>
__global__ void k_affine_to_special(const uint32_t *__restrict__ aff, uint32_t *__restrict__ raw,
                                    uint32_t n) {
    using R = Raw<C>;
    using F = typename C::Fq;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Affine<F> p = Affine<F>::load(aff + (size_t)i * R::AW);
    const bool inf = p.is_inf();
    R::store(raw + (size_t)i * R::PW, p.x, inf ? F::one() : p.y, inf ? F::zero() : F::one());
}

}  // namespace b200msm
