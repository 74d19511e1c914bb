// Wire-format MSM: device-side decode / validate / encode of the byte format of the reference's
// C ABI (ffi/ffi_serialization.hpp:12-16, ffi/ffi_serialization.tcc:98-187, sizes ffi/ffi.h:13-17,
// 55-59): every base-field value is big-endian, left-padded to the in-memory width, PLAIN (not
// Montgomery); an extension-field element lists its coefficients highest order first (c1 | c0);
// a point is affine X | Y with (0, 1) standing for infinity (.tcc:150-168); a scalar is one Fr
// value.  Reading fails when a value is >= its modulus (.tcc:64-75) and, for points, when
// is_well_formed() or is_in_safe_subgroup() fails (.tcc:165).
#pragma once
#include "curves.cuh"

namespace b200msm {

// Curve coefficient b in Montgomery form (all curves here have a = 0): alt_bn128 y^2 = x^3 + 3
// (alt_bn128_init.cpp:266), its twist b' = 3 / (9 + u) (alt_bn128_init.cpp:267-268), bls12_377
// y^2 = x^3 + 1 (bls12_377_init.cpp:301), bw6_761 y^2 = x^3 - 1 (bw6_761_init.cpp:267-271).
template <class C>
struct CurveB;
template <>
struct CurveB<AltBn128G1> {
    using F = AltBn128G1::Fq;
    B200_DEV static F get() {
        F one = F::one(), two = F::dbl(one);
        return F::add(two, one);
    }
};
template <>
struct CurveB<AltBn128G2> {
    using F = AltBn128G2::Fq;
    B200_DEV static F get() {
        constexpr uint32_t c0[8] = {0x77b802a8u, 0x3bf938e3u, 0x3633535du, 0x020b1b27u,
                                    0x49755260u, 0x26b7edf0u, 0x4384a86du, 0x2514c632u};
        constexpr uint32_t c1[8] = {0xd1dcff67u, 0x38e7ecccu, 0x93ce0d3eu, 0x65f0b37du,
                                    0x22ac00aau, 0xd749d0ddu, 0x4a688d4du, 0x0141b9ceu};
        F r;
#pragma unroll
        for (int i = 0; i < 8; i++) r.c0.v[i] = c0[i], r.c1.v[i] = c1[i];
        return r;
    }
};
template <>
struct CurveB<Bls12377G1> {
    using F = Bls12377G1::Fq;
    B200_DEV static F get() { return F::one(); }
};
template <>
struct CurveB<Bw6761G1> {
    using F = Bw6761G1::Fq;
    B200_DEV static F get() { return F::neg(F::one()); }
};

// bls12_377 twist y^2 = x^3 + 1/u (bls12_377_init.cpp:302-304): 1/u = -(1/5) u
template <>
struct CurveB<Bls12377G2> {
    using F = Bls12377G2::Fq;
    B200_DEV static F get() {
        constexpr uint32_t c1[12] = {0x66666685u, 0x80722666u, 0x899999a9u, 0x8df55926u,
                                     0xd64f34cfu, 0x7fe4561au, 0xb6e4f01bu, 0xb95da6d8u,
                                     0xfc142743u, 0x4b747cccu, 0x70f49f43u, 0x0039c3fau};
        F r = F::zero();
#pragma unroll
        for (int i = 0; i < 12; i++) r.c1.v[i] = c1[i];
        return r;
    }
};
// bw6_761 twist y^2 = x^3 + 4 (bw6_761_init.cpp:272-282)
template <>
struct CurveB<Bw6761G2> {
    using F = Bw6761G2::Fq;
    B200_DEV static F get() { return F::dbl(F::dbl(F::one())); }
};
// bls12_381 y^2 = x^3 + 4 and its twist y^2 = x^3 + 4 (1 + u) (bls12_381_init.cpp:418-420)
template <>
struct CurveB<Bls12381G1> {
    using F = Bls12381G1::Fq;
    B200_DEV static F get() { return F::dbl(F::dbl(F::one())); }
};
template <>
struct CurveB<Bls12381G2> {
    using F = Bls12381G2::Fq;
    B200_DEV static F get() {
        auto four = Bls12381G1::Fq::dbl(Bls12381G1::Fq::dbl(Bls12381G1::Fq::one()));
        return F{four, four};
    }
};

// error bits accumulated by the decode kernels
enum WireError : uint32_t {
    WIRE_ERR_RANGE = 1,     // a coordinate or scalar >= its modulus
    WIRE_ERR_CURVE = 2,     // point not on the curve
    WIRE_ERR_SUBGROUP = 4,  // point outside the prime-order subgroup
};

// big-endian bytes -> little-endian 32-bit limbs of one prime-field value; false if >= modulus
template <class P>
B200_DEV bool wire_read_fp(const uint8_t *src, Fp<P> &out) {
    constexpr int N = P::N;
#pragma unroll
    for (int k = 0; k < N; k++) {
        const uint8_t *b = src + 4 * (N - 1 - k);
        out.v[k] = ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3];
    }
    bool lt = false, decided = false;  // out < modulus ?
#pragma unroll
    for (int k = N - 1; k >= 0; k--) {
        if (!decided && out.v[k] != P::mod(k)) {
            lt = out.v[k] < P::mod(k);
            decided = true;
        }
    }
    return lt;
}
template <class P>
B200_DEV void wire_write_fp(uint8_t *dst, const Fp<P> &plain) {
    constexpr int N = P::N;
#pragma unroll
    for (int k = 0; k < N; k++) {
        uint8_t *b = dst + 4 * (N - 1 - k);
        const uint32_t v = plain.v[k];
        b[0] = (uint8_t)(v >> 24), b[1] = (uint8_t)(v >> 16), b[2] = (uint8_t)(v >> 8), b[3] = (uint8_t)v;
    }
}

// coordinate-field element (Fp or Fp2) from the wire, converted to Montgomery form
template <class P>
B200_DEV bool wire_read_coord(const uint8_t *src, Fp<P> &out, bool &is_zero, bool &is_one) {
    Fp<P> plain;
    bool ok = wire_read_fp<P>(src, plain);
    is_zero = plain.is_zero();
    uint32_t o = plain.v[0] ^ 1u;
#pragma unroll
    for (int k = 1; k < P::N; k++) o |= plain.v[k];
    is_one = o == 0;
    out = Fp<P>::to_mont(plain);
    return ok;
}
template <class P, int NR>
B200_DEV bool wire_read_coord(const uint8_t *src, Fp2<P, NR> &out, bool &is_zero, bool &is_one) {
    bool z0, o0, z1, o1;
    bool ok1 = wire_read_coord<P>(src, out.c1, z1, o1);               // highest order first
    bool ok0 = wire_read_coord<P>(src + 4 * P::N, out.c0, z0, o0);
    is_zero = z0 && z1;
    is_one = o0 && z1;
    return ok0 && ok1;
}
template <class P>
B200_DEV void wire_write_coord(uint8_t *dst, const Fp<P> &mont) {
    wire_write_fp<P>(dst, Fp<P>::from_mont(mont));
}
template <class P, int NR>
B200_DEV void wire_write_coord(uint8_t *dst, const Fp2<P, NR> &mont) {
    wire_write_fp<P>(dst, Fp<P>::from_mont(mont.c1));
    wire_write_fp<P>(dst + 4 * P::N, Fp<P>::from_mont(mont.c0));
}

// points: wire -> device affine (Montgomery) + infinity flags; validation into *err
template <class C>
__global__ void __launch_bounds__(128)
    k_wire_decode_points(const uint8_t *__restrict__ wire, uint32_t *__restrict__ aff,
                         uint8_t *__restrict__ inf_flag, uint32_t *__restrict__ err, uint32_t n,
                         uint32_t flags) {
    using R = Raw<C>;
    using F = typename C::Fq;
    using Fr = typename C::Fr;
    using X = XYZZ<F>;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    constexpr size_t CB = (size_t)R::CW * 4;
    F x, y;
    bool xz, xo, yz, yo;
    uint32_t e = 0;
    if (!wire_read_coord(wire + (size_t)i * 2 * CB, x, xz, xo)) e |= WIRE_ERR_RANGE;
    if (!wire_read_coord(wire + (size_t)i * 2 * CB + CB, y, yz, yo)) e |= WIRE_ERR_RANGE;
    const bool inf = xz && yo;  // (0, 1), ffi_serialization.tcc:157-162
    if (!inf && e == 0) {
        if (flags & 1u) {  // is_well_formed: y^2 = x^3 + b
            F lhs = F::sqr(y);
            F rhs = F::add(F::mul(F::sqr(x), x), CurveB<C>::get());
            if (!(lhs == rhs)) e |= WIRE_ERR_CURVE;
        }
        if ((flags & 2u) && e == 0 && C::ID != 0) {  // is_in_safe_subgroup: r P = 0 (trivial for alt_bn128 G1)
            X acc = X::inf();
            for (int bit = Fr::Params::BITS - 1; bit >= 0; bit--) {
                acc = X::dbl(acc);
                uint32_t w = 0;
#pragma unroll
                for (int k = 0; k < Fr::N; k++)
                    if ((bit >> 5) == k) w = Fr::Params::mod(k);
                if ((w >> (bit & 31)) & 1) X::madd(acc, x, y);
            }
            if (!acc.is_inf()) e |= WIRE_ERR_SUBGROUP;
        }
    }
    Affine<F> a{inf ? F::zero() : x, inf ? F::zero() : y};
    a.store(aff + (size_t)i * R::AW);
    inf_flag[i] = inf;
    if (e) atomicOr(err, e);
}

// ---- libff's binary stream encoding (what multi_exp_stream reads) -------------------------------
// group_element_codec<encoding_binary, Form, compression_off> (curve_serialization.tcc:79-101):
// X then Y; an extension-field coordinate is coeffs[0] then coeffs[1] (field_serialization.tcc:
// 127-147); every prime-field element is its bigint with ALL bytes reversed, i.e. big-endian
// (:197-222), holding the plain value (form_plain) or the Montgomery representation
// (form_montgomery).  X = 0 and Y = 1 encodes the zero element.  Like the reference's reader, no
// range or curve check is made.
template <class P>
B200_DEV void stream_read_coord(const uint8_t *src, bool mont, Fp<P> &out) {
    Fp<P> v;
    (void)wire_read_fp<P>(src, v);
    out = mont ? v : Fp<P>::to_mont(v);
}
template <class P, int NR>
B200_DEV void stream_read_coord(const uint8_t *src, bool mont, Fp2<P, NR> &out) {
    stream_read_coord<P>(src, mont, out.c0);  // lowest order first
    stream_read_coord<P>(src + 4 * P::N, mont, out.c1);
}

template <class C>
__global__ void __launch_bounds__(128)
    k_stream_decode_points(const uint8_t *__restrict__ bytes, uint32_t *__restrict__ aff,
                           uint8_t *__restrict__ inf_flag, uint32_t n, uint32_t flags) {
    using R = Raw<C>;
    using F = typename C::Fq;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    constexpr size_t CB = (size_t)R::CW * 4;
    F x, y;
    stream_read_coord(bytes + (size_t)i * 2 * CB, (flags & 1u) != 0, x);
    stream_read_coord(bytes + (size_t)i * 2 * CB + CB, (flags & 1u) != 0, y);
    const bool inf = x.is_zero() && y == F::one();
    Affine<F> a{inf ? F::zero() : x, inf ? F::zero() : y};
    a.store(aff + (size_t)i * R::AW);
    inf_flag[i] = inf;
}

// scalars: wire -> Montgomery limbs (what the MSM pipeline takes)
template <class C>
__global__ void k_wire_decode_scalars(const uint8_t *__restrict__ wire, uint32_t *__restrict__ scalars,
                                      uint32_t *__restrict__ err, uint32_t n) {
    using Fr = typename C::Fr;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr s;
    if (!wire_read_fp<typename Fr::Params>(wire + (size_t)i * Fr::N * 4, s)) atomicOr(err, (uint32_t)WIRE_ERR_RANGE);
    Fr::to_mont(s).store(scalars + (size_t)i * Fr::N);
}

// result: reference-layout element -> affine wire bytes (group_element_write, .tcc:170-187)
template <class C>
__global__ void k_wire_encode_point(const uint32_t *__restrict__ raw, uint8_t *__restrict__ wire) {
    using R = Raw<C>;
    using F = typename C::Fq;
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    constexpr size_t CB = (size_t)R::CW * 4;
    F Xc, Yc, Zc;
    R::load(raw, Xc, Yc, Zc);
    F x = F::zero(), y = F::one();  // infinity is written as (0, 1)
    if (!R::is_inf(Xc, Zc)) {
        F zi = F::inv(Zc);
        if constexpr (C::PROJECTIVE) {
            x = F::mul(Xc, zi);
            y = F::mul(Yc, zi);
        } else {
            F zi2 = F::sqr(zi);
            x = F::mul(Xc, zi2);
            y = F::mul(Yc, F::mul(zi2, zi));
        }
    }
    wire_write_coord(wire, x);
    wire_write_coord(wire + CB, y);
}

}  // namespace b200msm
