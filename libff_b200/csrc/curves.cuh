// The four groups of the MSM path and how their elements lie in the reference's memory.
//
// In-memory layout: a libff group element is three coordinate-field elements
// X | Y | Z, each `Fq::N` little-endian 32-bit words of Montgomery representation, contiguous,
// stride = sizeof(GroupT):
//   alt_bn128_G1  Jacobian, 96 B      libff/algebra/curves/alt_bn128/alt_bn128_g1.hpp:44-46
//   alt_bn128_G2  Jacobian/Fq2, 192 B libff/algebra/curves/alt_bn128/alt_bn128_g2.hpp:45
//   bls12_377_G1  Jacobian, 144 B     libff/algebra/curves/bls12_377/bls12_377_g1.hpp:44
//   bw6_761_G1    homogeneous projective, 288 B   libff/algebra/curves/bw6_761/bw6_761_g1.hpp:37-39
// and the remaining groups of the reference's pairing-friendly families:
//   bls12_377_G2  Jacobian/Fq2, 288 B libff/algebra/curves/bls12_377/bls12_377_g2.hpp:45
//   bw6_761_G2    homogeneous projective over Fq, 288 B  bw6_761/bw6_761_g2.hpp:39
//   bls12_381_G1  Jacobian, 144 B     libff/algebra/curves/bls12_381/bls12_381_g1.hpp:44
//   bls12_381_G2  Jacobian/Fq2, 288 B libff/algebra/curves/bls12_381/bls12_381_g2.hpp:45
// Infinity: Jacobian Z = 0 (alt_bn128_g1.cpp:91); bw6_761 X = 0 and Z = 0 (bw6_761_g1.cpp:79-82).
#pragma once
#include "xyzz.cuh"

namespace b200msm {

struct AltBn128G1 {
    static constexpr int ID = 0;
    using Fq = Fp<AltBn128FqParams>;
    using Fr = Fp<AltBn128FrParams>;
    static constexpr bool PROJECTIVE = false;
};
struct AltBn128G2 {
    static constexpr int ID = 1;
    using Fq = Fp2<AltBn128FqParams>;
    using Fr = Fp<AltBn128FrParams>;
    static constexpr bool PROJECTIVE = false;
};
struct Bls12377G1 {
    static constexpr int ID = 2;
    using Fq = Fp<Bls12377FqParams>;
    using Fr = Fp<Bls12377FrParams>;
    static constexpr bool PROJECTIVE = false;
};
struct Bw6761G1 {
    static constexpr int ID = 3;
    using Fq = Fp<Bw6761FqParams>;
    using Fr = Fp<Bls12377FqParams>;  // bw6_761 Fr is bls12_377 Fq (bw6_761_init.cpp:35-59)
    static constexpr bool PROJECTIVE = true;
};

struct Bls12377G2 {
    static constexpr int ID = 4;
    using Fq = Fp2<Bls12377FqParams, 5>;  // u^2 = -5 (bls12_377_init.cpp:170-176)
    using Fr = Fp<Bls12377FrParams>;
    static constexpr bool PROJECTIVE = false;
};
struct Bw6761G2 {
    static constexpr int ID = 5;
    using Fq = Fp<Bw6761FqParams>;        // the M-twist is defined over Fq itself
    using Fr = Fp<Bls12377FqParams>;
    static constexpr bool PROJECTIVE = true;
};
struct Bls12381G1 {
    static constexpr int ID = 6;
    using Fq = Fp<Bls12381FqParams>;
    using Fr = Fp<Bls12381FrParams>;
    static constexpr bool PROJECTIVE = false;
};
struct Bls12381G2 {
    static constexpr int ID = 7;
    using Fq = Fp2<Bls12381FqParams, 1>;  // u^2 = -1 (bls12_381_init.cpp:137-139)
    using Fr = Fp<Bls12381FrParams>;
    static constexpr bool PROJECTIVE = false;
};

template <class C>
struct Raw {
    using F = typename C::Fq;
    static constexpr int CW = F::N;        // words per coordinate
    static constexpr int PW = 3 * F::N;    // words per group element
    static constexpr int AW = 2 * F::N;    // words per device affine point
    static constexpr int XW = 4 * F::N;    // words per XYZZ point
    static constexpr int SW = C::Fr::N;    // words per scalar

    B200_DEV static void load(const uint32_t *p, F &X, F &Y, F &Z) {
        X = F::load(p);
        Y = F::load(p + CW);
        Z = F::load(p + 2 * CW);
    }
    B200_DEV static void store(uint32_t *p, const F &X, const F &Y, const F &Z) {
        X.store(p);
        Y.store(p + CW);
        Z.store(p + 2 * CW);
    }
    B200_DEV static bool is_inf(const F &X, const F &Z) {
        if constexpr (C::PROJECTIVE)
            return X.is_zero() && Z.is_zero();
        else
            return Z.is_zero();
    }
    // XYZZ of a reference-layout element (no inversion: Jacobian (X,Y,Z) = XYZZ (X,Y,Z^2,Z^3);
    // projective (X,Y,Z) = XYZZ (X Z, Y Z^2, Z^2, Z^3))
    B200_DEV static XYZZ<F> to_xyzz(const F &X, const F &Y, const F &Z) {
        if (is_inf(X, Z)) return XYZZ<F>::inf();
        F zz = F::sqr(Z);
        F zzz = F::mul(zz, Z);
        if constexpr (C::PROJECTIVE)
            return XYZZ<F>{F::mul(X, Z), F::mul(Y, zz), zz, zzz};
        else
            return XYZZ<F>{X, Y, zz, zzz};
    }
    B200_DEV static void from_xyzz(const XYZZ<F> &p, F &X, F &Y, F &Z) {
        if constexpr (C::PROJECTIVE)
            p.to_projective(X, Y, Z);
        else
            p.to_jacobian(X, Y, Z);
    }
};

}  // namespace b200msm
