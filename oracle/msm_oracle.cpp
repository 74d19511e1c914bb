// TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's MSM path ("port" oracle).
//
// Plain C++ (no GMP, no reference headers): 64-bit-limb Montgomery fields, the reference's own
// Jacobian / homogeneous-projective group formulas, and its bucket-method multi_exp, batch_exp
// and batch_to_special, each function citing the clearmatics/libff file:line it follows.  It is
// the checker used when the compiled reference (oracle/_ref/libffref.so) is not available, and it
// is itself pinned against that library and against tests/golden/ (tests/test_oracle.py).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load it; the product (libff_b200/) never does.
//
// Build: make -C oracle oracle   ->  oracle/libmsmoracle.so
#include <openssl/sha.h>

#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;

// ------------------------------------------------------------------------------------------------
// Multi-precision helpers on little-endian 64-bit limbs (libff/algebra/fields/bigint.hpp:28-65)

template <int N>
struct Params {
    uint64_t mod[N];
    uint64_t inv;     // -p^-1 mod 2^64        (e.g. alt_bn128_init.cpp:62,103)
    uint64_t r1[N];   // R mod p = one()       (fp.tcc:230-235)
    uint64_t r2[N];   // R^2 mod p (Rsquared)  (alt_bn128_init.cpp:88-93)
    int bits;
};

template <int N>
static int cmp_n(const uint64_t *a, const uint64_t *b) {
    for (int i = N - 1; i >= 0; i--)
        if (a[i] != b[i]) return a[i] < b[i] ? -1 : 1;
    return 0;
}
template <int N>
static uint64_t add_n(uint64_t *r, const uint64_t *a, const uint64_t *b) {
    u128 c = 0;
    for (int i = 0; i < N; i++) {
        c += (u128)a[i] + b[i];
        r[i] = (uint64_t)c;
        c >>= 64;
    }
    return (uint64_t)c;
}
template <int N>
static uint64_t sub_n(uint64_t *r, const uint64_t *a, const uint64_t *b) {
    uint64_t borrow = 0;
    for (int i = 0; i < N; i++) {
        u128 d = (u128)a[i] - b[i] - borrow;
        r[i] = (uint64_t)d;
        borrow = (uint64_t)(d >> 64) & 1;
    }
    return borrow;
}

static int hexval(char c) { return c <= '9' ? c - '0' : (c | 32) - 'a' + 10; }

template <int N>
static void parse_hex(const char *s, uint64_t *out) {
    memset(out, 0, 8 * N);
    size_t len = strlen(s);
    for (size_t k = 0; k < len; k++) {
        size_t bit = 4 * (len - 1 - k);
        out[bit / 64] |= (uint64_t)hexval(s[k]) << (bit % 64);
    }
}

template <int N>
static void setup_params(Params<N> &P, const char *mod_hex) {
    parse_hex<N>(mod_hex, P.mod);
    int top = N - 1;
    while (P.mod[top] == 0) top--;
    P.bits = 64 * top + (64 - __builtin_clzll(P.mod[top]));
    uint64_t x = 1;  // Newton iteration for p^-1 mod 2^64
    for (int i = 0; i < 6; i++) x *= 2 - P.mod[0] * x;
    P.inv = (uint64_t)0 - x;
    // 2^(64 N) mod p and 2^(128 N) mod p by repeated doubling of 1
    uint64_t v[N] = {1};
    auto dbl_mod = [&]() {
        uint64_t carry = add_n<N>(v, v, v);
        if (carry || cmp_n<N>(v, P.mod) >= 0) sub_n<N>(v, v, P.mod);
    };
    for (int i = 0; i < 64 * N; i++) dbl_mod();
    memcpy(P.r1, v, sizeof v);
    for (int i = 0; i < 64 * N; i++) dbl_mod();
    memcpy(P.r2, v, sizeof v);
}

// ------------------------------------------------------------------------------------------------
// Fp_model<n, modulus> (libff/algebra/fields/fp.hpp:43, fp.tcc).  Elements are mont_repr limbs.

template <int N_, int ID>
struct Fp {
    static constexpr int N = N_;
    static constexpr int WORDS = N_;
    static Params<N_> P;
    uint64_t v[N_];

    static Fp zero() {
        Fp r;
        memset(r.v, 0, sizeof r.v);
        return r;
    }
    static Fp one() {
        Fp r;
        memcpy(r.v, P.r1, sizeof r.v);
        return r;
    }
    bool is_zero() const {  // fp.tcc:301-306
        uint64_t o = 0;
        for (int i = 0; i < N; i++) o |= v[i];
        return o == 0;
    }
    bool operator==(const Fp &b) const { return memcmp(v, b.v, sizeof v) == 0; }  // fp.tcc:289-293
    bool operator!=(const Fp &b) const { return !(*this == b); }

    // fp.tcc:350-465 (generic branch :449-462): add, then subtract p if carry or >= p
    Fp operator+(const Fp &b) const {
        Fp r;
        uint64_t carry = add_n<N>(r.v, v, b.v);
        if (carry || cmp_n<N>(r.v, P.mod) >= 0) sub_n<N>(r.v, r.v, P.mod);
        return r;
    }
    // fp.tcc:467-547 (generic branch :529-545): if a < b add p first, then subtract
    Fp operator-(const Fp &b) const {
        Fp r;
        if (sub_n<N>(r.v, v, b.v)) add_n<N>(r.v, r.v, P.mod);
        return r;
    }
    // fp.tcc:616-630: -0 = 0, else p - a
    Fp operator-() const {
        if (is_zero()) return *this;
        Fp r;
        sub_n<N>(r.v, P.mod, v);
        return r;
    }
    // mul_reduce, generic branch fp.tcc:204-227: full 2n-limb product, then n rounds of
    // k = inv * res[i]; res += k * p << (64 i); result = upper half, minus p if >= p.
    static Fp mul_reduce(const uint64_t *a, const uint64_t *b) {
        uint64_t res[2 * N + 1] = {0};
        for (int i = 0; i < N; i++) {
            u128 c = 0;
            for (int j = 0; j < N; j++) {
                c += (u128)a[j] * b[i] + res[i + j];
                res[i + j] = (uint64_t)c;
                c >>= 64;
            }
            res[i + N] = (uint64_t)c;
        }
        for (int i = 0; i < N; i++) {
            uint64_t k = P.inv * res[i];
            u128 c = 0;
            for (int j = 0; j < N; j++) {
                c += (u128)k * P.mod[j] + res[i + j];
                res[i + j] = (uint64_t)c;
                c >>= 64;
            }
            for (int j = i + N; c != 0 && j <= 2 * N; j++) {  // mpn_add_1 carry propagation
                c += res[j];
                res[j] = (uint64_t)c;
                c >>= 64;
            }
        }
        Fp r;
        memcpy(r.v, res + N, sizeof r.v);
        if (res[2 * N] || cmp_n<N>(r.v, P.mod) >= 0) sub_n<N>(r.v, r.v, P.mod);
        return r;
    }
    Fp operator*(const Fp &b) const { return mul_reduce(v, b.v); }  // fp.tcc:549-559
    Fp squared() const { return mul_reduce(v, v); }                 // fp.tcc:632-677
    // as_bigint (fp.tcc:270-281): one reduction of (mont_repr, 1)
    void as_bigint(uint64_t *out) const {
        uint64_t o[N] = {1};
        Fp r = mul_reduce(v, o);
        memcpy(out, r.v, sizeof r.v);
    }
    // Fp_model(bigint) (fp.tcc:230-235): multiply by Rsquared
    static Fp from_bigint(const uint64_t *plain) { return mul_reduce(plain, P.r2); }
    // invert (fp.tcc:679-727) uses mpn_gcdext; the inverse in [0, p) is unique, so a^(p-2) is the
    // same element.  Only reached once per normalisation.
    Fp inverse() const {
        uint64_t e[N], two[N] = {2};
        sub_n<N>(e, P.mod, two);
        Fp r = one();
        for (int i = P.bits - 1; i >= 0; i--) {
            r = r.squared();
            if ((e[i / 64] >> (i % 64)) & 1) r = r * *this;
        }
        return r;
    }
    void load(const uint64_t *p) { memcpy(v, p, sizeof v); }
    void store(uint64_t *p) const { memcpy(p, v, sizeof v); }
};
template <int N_, int ID>
Params<N_> Fp<N_, ID>::P;

// Fp2_model (libff/algebra/fields/fp2.hpp, fp2.tcc:79-168): c0 + c1 U, U^2 = non_residue
template <class F>
struct Fp2 {
    static constexpr int WORDS = 2 * F::WORDS;
    static F non_residue;  // alt_bn128_init.cpp:138-140
    F c0, c1;
    static Fp2 zero() { return Fp2{F::zero(), F::zero()}; }
    static Fp2 one() { return Fp2{F::one(), F::zero()}; }
    bool is_zero() const { return c0.is_zero() && c1.is_zero(); }
    bool operator==(const Fp2 &b) const { return c0 == b.c0 && c1 == b.c1; }
    bool operator!=(const Fp2 &b) const { return !(*this == b); }
    Fp2 operator+(const Fp2 &b) const { return Fp2{c0 + b.c0, c1 + b.c1}; }  // fp2.tcc:79-85
    Fp2 operator-(const Fp2 &b) const { return Fp2{c0 - b.c0, c1 - b.c1}; }  // fp2.tcc:87-93
    Fp2 operator-() const { return Fp2{-c0, -c1}; }                          // fp2.tcc:116-120
    Fp2 operator*(const Fp2 &b) const {  // fp2.tcc:101-114 (Karatsuba)
        const F aA = c0 * b.c0, bB = c1 * b.c1;
        return Fp2{aA + non_residue * bB, (c0 + c1) * (b.c0 + b.c1) - aA - bB};
    }
    Fp2 squared() const {  // squared_complex, fp2.tcc:141-151
        const F ab = c0 * c1;
        return Fp2{(c0 + c1) * (c0 + non_residue * c1) - ab - non_residue * ab, ab + ab};
    }
    Fp2 inverse() const {  // fp2.tcc:153-168
        const F t0 = c0.squared(), t1 = c1.squared();
        const F t3 = (t0 - non_residue * t1).inverse();
        return Fp2{c0 * t3, -(c1 * t3)};
    }
    void load(const uint64_t *p) {
        c0.load(p);
        c1.load(p + F::WORDS);
    }
    void store(uint64_t *p) const {
        c0.store(p);
        c1.store(p + F::WORDS);
    }
};
template <class F>
F Fp2<F>::non_residue;

// ------------------------------------------------------------------------------------------------
// Groups.  Jacobian (alt_bn128 G1/G2, bls12_377 G1) and homogeneous projective (bw6_761 G1), all
// with curve coefficient a = 0 (alt_bn128_init.cpp:293, bw6_761_init.cpp:299).

template <class F, int GID>
struct Jac {
    using Fq = F;
    static constexpr bool FOLDED_EQ = GID != 0;  // see operator+
    F X, Y, Z;
    static Jac zero() { return Jac{F::zero(), F::one(), F::zero()}; }  // alt_bn128_init.cpp:287-288
    bool is_zero() const { return Z.is_zero(); }                        // alt_bn128_g1.cpp:91
    bool is_special() const { return is_zero() || Z == F::one(); }
    Jac operator-() const { return Jac{X, -Y, Z}; }                     // alt_bn128_g1.cpp:139-142

    bool operator==(const Jac &o) const {  // alt_bn128_g1.cpp:93-127
        if (is_zero()) return o.is_zero();
        if (o.is_zero()) return false;
        const F Z1s = Z.squared(), Z2s = o.Z.squared();
        if ((X * Z2s) != (o.X * Z1s)) return false;
        const F Z1c = Z * Z1s, Z2c = o.Z * Z2s;
        return !((Y * Z2c) != (o.Y * Z1c));
    }
    // dbl-2009-l, alt_bn128_g1.cpp:285-326 (same in alt_bn128_g2.cpp:309-349, bls12_377_g1.cpp:312-352)
    Jac dbl() const {
        if (is_zero()) return *this;
        const F A = X.squared(), B = Y.squared(), C = B.squared();
        F D = (X + B).squared() - A - C;
        D = D + D;
        const F E = A + A + A, Fv = E.squared();
        const F X3 = Fv - (D + D);
        F eightC = C + C;
        eightC = eightC + eightC;
        eightC = eightC + eightC;
        const F Y3 = E * (D - X3) - eightC;
        const F Y1Z1 = Y * Z;
        return Jac{X3, Y3, Y1Z1 + Y1Z1};
    }
    // add-2007-bl.  alt_bn128_G1::operator+ is add() with the operator== pre-check
    // (alt_bn128_g1.cpp:134-137,149-206); alt_bn128_G2 and bls12_377_G1 fold the check into the
    // formula (alt_bn128_g2.cpp:166-219, bls12_377_g1.cpp:121-160).  Same X3, Y3, Z3 either way.
    Jac operator+(const Jac &o) const {
        if (is_zero()) return o;
        if (o.is_zero()) return *this;
        if (!FOLDED_EQ && *this == o) return dbl();
        const F Z1Z1 = Z.squared(), Z2Z2 = o.Z.squared();
        const F U1 = X * Z2Z2, U2 = o.X * Z1Z1;
        const F S1 = FOLDED_EQ ? Y * (o.Z * Z2Z2) : Y * o.Z * Z2Z2;
        const F S2 = FOLDED_EQ ? o.Y * (Z * Z1Z1) : o.Y * Z * Z1Z1;
        if (FOLDED_EQ && U1 == U2 && S1 == S2) return dbl();
        const F H = U2 - U1, S2mS1 = S2 - S1;
        const F I = (H + H).squared(), J = H * I;
        const F r = S2mS1 + S2mS1, V = U1 * I;
        const F X3 = r.squared() - J - (V + V);
        const F S1J = S1 * J;
        const F Y3 = r * (V - X3) - (S1J + S1J);
        const F Z3 = ((Z + o.Z).squared() - Z1Z1 - Z2Z2) * H;
        return Jac{X3, Y3, Z3};
    }
    // madd-2007-bl, alt_bn128_g1.cpp:208-283 (alt_bn128_g2.cpp:232-307, bls12_377_g1.cpp:236-310)
    Jac mixed_add(const Jac &o) const {
        if (is_zero()) return o;
        if (o.is_zero()) return *this;
        const F Z1Z1 = Z.squared();
        const F U2 = o.X * Z1Z1;
        const F Z1c = Z * Z1Z1;
        const F S2 = o.Y * Z1c;
        if (X == U2 && Y == S2) return dbl();
        const F H = U2 - X, HH = H.squared();
        F I = HH + HH;
        I = I + I;
        const F J = H * I;
        F r = S2 - Y;
        r = r + r;
        const F V = X * I;
        const F X3 = r.squared() - J - V - V;
        F Y3 = Y * J;
        Y3 = r * (V - X3) - Y3 - Y3;
        const F Z3 = (Z + H).squared() - Z1Z1 - HH;
        return Jac{X3, Y3, Z3};
    }
    // to_affine_coordinates / to_special, alt_bn128_g1.cpp:68-84
    void to_special() {
        if (is_zero()) {
            *this = zero();
            return;
        }
        const F Zi = Z.inverse(), Z2i = Zi.squared(), Z3i = Z2i * Zi;
        X = X * Z2i, Y = Y * Z3i, Z = F::one();
    }
    // batch_to_special_all_non_zeros tail, alt_bn128_g1.cpp:468-476 (zinv = 1/Z)
    void apply_zinv(const F &zinv) {
        const F Z2 = zinv.squared(), Z3 = zinv * Z2;
        X = X * Z2, Y = Y * Z3, Z = F::one();
    }
};

template <class F, int GID>
struct Proj {
    using Fq = F;
    F X, Y, Z;
    static Proj zero() { return Proj{F::zero(), F::one(), F::zero()}; }  // bw6_761_init.cpp:285-286
    bool is_zero() const { return X.is_zero() && Z.is_zero(); }           // bw6_761_g1.cpp:79-82
    bool is_special() const { return is_zero() || Z == F::one(); }
    Proj operator-() const { return Proj{X, -Y, Z}; }                     // bw6_761_g1.cpp:196-199
    bool operator==(const Proj &o) const {                                // bw6_761_g1.cpp:84-101
        if (is_zero()) return o.is_zero();
        if (o.is_zero()) return false;
        return !(((X * o.Z) != (o.X * Z)) || ((Y * o.Z) != (o.Y * Z)));
    }
    Proj dbl() const {  // dbl-2007-bl, bw6_761_g1.cpp:318-358
        if (is_zero()) return *this;
        const F XX = X.squared(), w = XX + XX + XX;
        const F Y1Z1 = Y * Z, s = Y1Z1 + Y1Z1;
        const F ss = s.squared(), sss = s * ss;
        const F R = Y * s, RR = R.squared();
        const F B = (X + R).squared() - XX - RR;
        const F h = w.squared() - (B + B);
        return Proj{h * s, w * (B - h) - (RR + RR), sss};
    }
    Proj operator+(const Proj &o) const {  // add-1998-cmo-2, bw6_761_g1.cpp:111-194
        if (is_zero()) return o;
        if (o.is_zero()) return *this;
        const F X1Z2 = X * o.Z, X2Z1 = Z * o.X, Y1Z2 = Y * o.Z, Y2Z1 = Z * o.Y;
        if (X1Z2 == X2Z1 && Y1Z2 == Y2Z1) return dbl();
        const F Z1Z2 = Z * o.Z;
        const F u = Y2Z1 - Y1Z2, uu = u.squared();
        const F v = X2Z1 - X1Z2, vv = v.squared(), vvv = v * vv;
        const F R = vv * X1Z2;
        const F A = uu * Z1Z2 - (vvv + R + R);
        return Proj{v * A, u * (R - A) - vvv * Y1Z2, vvv * Z1Z2};
    }
    Proj mixed_add(const Proj &o) const {  // madd-1998-cmo, bw6_761_g1.cpp:263-316
        if (is_zero()) return o;
        if (o.is_zero()) return *this;
        const F X2Z1 = Z * o.X, Y2Z1 = Z * o.Y;
        if (X == X2Z1 && Y == Y2Z1) return dbl();
        const F u = Y2Z1 - Y, uu = u.squared();
        const F v = X2Z1 - X, vv = v.squared(), vvv = v * vv;
        const F R = vv * X;
        const F A = uu * Z - vvv - R - R;
        return Proj{v * A, u * (R - A) - vvv * Y, vvv * Z};
    }
    void to_special() {  // bw6_761_g1.cpp:58-72
        if (is_zero()) {
            *this = zero();
            return;
        }
        const F Zi = Z.inverse();
        X = X * Zi, Y = Y * Zi, Z = F::one();
    }
    void apply_zinv(const F &zinv) { X = X * zinv, Y = Y * zinv, Z = F::one(); }  // bw6_761_g1.cpp:501-505
};

template <class G>
static G load_point(const uint64_t *p) {
    G g;
    constexpr int W = G::Fq::WORDS;
    g.X.load(p), g.Y.load(p + W), g.Z.load(p + 2 * W);
    return g;
}
template <class G>
static void store_point(uint64_t *p, const G &g) {
    constexpr int W = G::Fq::WORDS;
    g.X.store(p), g.Y.store(p + W), g.Z.store(p + 2 * W);
}

// ------------------------------------------------------------------------------------------------
// Scalars as plain integers: bigint<n>::num_bits / test_bit (bigint.tcc:89-136), digit extraction
// (field_utils.tcc:50-100) and signed recoding (field_utils.tcc:102-239).

template <int N>
struct Big {
    uint64_t d[N];
    size_t num_bits() const {  // bigint.tcc:89-109
        for (int i = N - 1; i >= 0; i--)
            if (d[i]) return 64 * (size_t)i + (64 - __builtin_clzll(d[i]));
        return 0;
    }
    bool test_bit(size_t bit) const {  // bigint.tcc:126-136: out-of-range bits read as 0
        if (bit >= 64 * (size_t)N) return false;
        return (d[bit / 64] >> (bit % 64)) & 1;
    }
};

// field_get_digit (field_utils.tcc:50-100): the c-bit digit at index idx, spanning <= 2 limbs
template <int N>
static size_t get_digit(const Big<N> &v, size_t c, size_t idx) {
    const size_t start = c * idx, end = start + c;
    const size_t lo = start / 64, hi = end / 64;
    size_t value = 0;
    if (lo < (size_t)N) value = v.d[lo] >> (start - lo * 64);
    if (hi < (size_t)N && hi != lo) {
        const size_t hi_bits = end - hi * 64;
        if (hi_bits) value |= v.d[hi] << (c - hi_bits);
    }
    return value & (((size_t)1 << c) - 1);
}

// field_get_signed_digit (field_utils.tcc:167-203): re-walks digits 0..idx with the carry
template <int N>
static long get_signed_digit(const Big<N> &v, size_t c, size_t idx) {
    const size_t carry_mask = (size_t)1 << (c - 1), overflow_mask = (size_t)1 << c;
    size_t carry = 0, overflow = 0, digit, i = 0;
    do {
        carry = overflow | carry;
        digit = get_digit(v, c, i) + carry;
        overflow = (digit & overflow_mask) >> c;
        carry = (digit & carry_mask) >> (c - 1);
        ++i;
    } while (i <= idx);
    return (long)((1 - overflow) * (digit - carry * overflow_mask));
}

// libff::log2 (common/utils.cpp:32-44) and pippenger_optimal_c (multiexp.tcc:35-40), size_t math
static size_t ceil_log2(size_t n) {
    size_t r = (n & (n - 1)) == 0 ? 0 : 1;
    while (n > 1) {
        n >>= 1;
        r++;
    }
    return r;
}
static size_t pippenger_optimal_c(size_t n) {
    const size_t l = ceil_log2(n);
    return l - (l / 3 - 2);
}

// ------------------------------------------------------------------------------------------------
// multi_exp_method_BDLO12 (multiexp.tcc:284-380): unsigned c-bit windows, MSB window first.
template <class G, class Fr, bool SPECIAL>
static G bdlo12_inner(const G *bases, const Fr *exps, size_t length) {
    const size_t c = pippenger_optimal_c(length);
    std::vector<Big<Fr::N>> bi(length);
    size_t num_bits = 0;
    for (size_t i = 0; i < length; i++) {
        exps[i].as_bigint(bi[i].d);
        num_bits = std::max(num_bits, bi[i].num_bits());
    }
    const size_t num_groups = (num_bits + c - 1) / c;
    G result = G::zero();
    bool result_nonzero = false;
    for (size_t k = num_groups - 1; k <= num_groups; k--) {
        if (result_nonzero)
            for (size_t i = 0; i < c; i++) result = result.dbl();
        std::vector<G> buckets((size_t)1 << c, G::zero());
        std::vector<bool> bucket_nonzero((size_t)1 << c);
        for (size_t i = 0; i < length; i++) {
            size_t id = 0;
            for (size_t j = 0; j < c; j++)
                if (bi[i].test_bit(k * c + j)) id |= (size_t)1 << j;
            if (id == 0) continue;
            if (bucket_nonzero[id]) {
                buckets[id] = SPECIAL ? buckets[id].mixed_add(bases[i]) : buckets[id] + bases[i];
            } else {
                buckets[id] = bases[i];
                bucket_nonzero[id] = true;
            }
        }
        G running_sum = G::zero();
        bool running_sum_nonzero = false;
        for (size_t i = ((size_t)1 << c) - 1; i > 0; i--) {
            if (bucket_nonzero[i]) {
                if (running_sum_nonzero) {
                    running_sum = running_sum + buckets[i];
                } else {
                    running_sum = buckets[i];
                    running_sum_nonzero = true;
                }
            }
            if (running_sum_nonzero) {
                if (result_nonzero) {
                    result = result + running_sum;
                } else {
                    result = running_sum;
                    result_nonzero = true;
                }
            }
        }
    }
    return result;
}

// multi_exp_method_BDLO12_signed (multiexp.tcc:45-125, 507-641): digits in [-2^(c-1), 2^(c-1)),
// 2^(c-1) buckets, c = pippenger_optimal_c + 1, (num_bits + 2 + c - 1) / c rounds.
template <class G, class Fr, bool SPECIAL>
static G bdlo12_signed_inner(const G *bases, const Fr *exps, size_t n) {
    const size_t c = pippenger_optimal_c(n) + 1;
    std::vector<Big<Fr::N>> bi(n);
    size_t num_bits = 0;
    for (size_t i = 0; i < n; i++) {
        exps[i].as_bigint(bi[i].d);
        num_bits = std::max(num_bits, bi[i].num_bits());
    }
    const size_t num_rounds = (num_bits + 2 + c - 1) / c;
    const size_t num_buckets = (size_t)1 << (c - 1);
    std::vector<G> buckets(num_buckets, G::zero());
    std::vector<bool> hit(num_buckets);
    auto round = [&](size_t digit_idx) -> G {
        hit.assign(num_buckets, false);
        size_t non_zero = 0;
        for (size_t i = 0; i < n; i++) {
            const long digit = get_signed_digit(bi[i], c, digit_idx);
            if (digit == 0) continue;
            const size_t b = (size_t)(digit < 0 ? -digit : digit) - 1;
            const G el = digit < 0 ? -bases[i] : bases[i];
            if (hit[b]) {
                buckets[b] = SPECIAL ? buckets[b].mixed_add(el) : buckets[b] + el;
            } else {
                buckets[b] = el;
                hit[b] = true;
            }
            non_zero++;
        }
        if (non_zero == 0) return G::zero();
        // multiexp_accumulate_buckets<normal> (multiexp.tcc:90-125)
        size_t i = num_buckets - 1;
        while (!hit[i]) --i;
        G sum = buckets[i], acc = buckets[i];
        while (i > 0) {
            --i;
            if (hit[i]) acc = acc + buckets[i];
            sum = sum + acc;
        }
        return sum;
    };
    if (n == 0) return G::zero();
    G result = round(num_rounds - 1);
    for (size_t r = 1; r < num_rounds; r++) {
        for (size_t i = 0; i < c; i++) result = result.dbl();
        result = result + round(num_rounds - 1 - r);
    }
    return result;
}

// ------------------------------------------------------------------------------------------------
// Streamed MSM (multiexp_stream.tcc) and the binary stream encoding it reads.

template <class F>
static size_t field_num_bits() {  // Fp_model::num_bits = bit length of the modulus
    Big<F::N> m;
    memcpy(m.d, F::P.mod, sizeof m.d);
    return m.num_bits();
}

// field_get_signed_digits (field_utils.tcc:205-239): one pass, num_digits digits; a carry out of
// the last digit is dropped, exactly as the reference does
template <int N>
static void get_signed_digits(std::vector<long> &digits, const Big<N> &v, size_t c, size_t num_digits) {
    const size_t carry_mask = (size_t)1 << (c - 1), overflow_mask = (size_t)1 << c;
    size_t carry = 0, overflow = 0;
    for (size_t k = 0; k < num_digits; k++) {
        carry = overflow | carry;
        const size_t digit = get_digit(v, c, k) + carry;
        overflow = (digit & overflow_mask) >> c;
        carry = (digit & carry_mask) >> (c - 1);
        digits[k] = (long)((1 - overflow) * (digit - carry * overflow_mask));
    }
}

// multi_exp_add_element_to_bucket_with_signed_digit<special> (multiexp.tcc:45-81)
template <class G>
static void add_to_bucket(std::vector<G> &buckets, std::vector<bool> &hit, const G &el, long digit) {
    if (digit == 0) return;
    const size_t b = (size_t)(digit < 0 ? -digit : digit) - 1;
    const G e = digit < 0 ? -el : el;
    if (hit[b]) {
        buckets[b] = buckets[b].mixed_add(e);
    } else {
        buckets[b] = e;
        hit[b] = true;
    }
}

// multiexp_accumulate_buckets<normal> (multiexp.tcc:90-125).  *defined = false when no bucket is
// set: the reference requires one (":88-89 Caller must ensure ...") and reads out of bounds otherwise.
template <class G>
static G accumulate_buckets(const std::vector<G> &buckets, const std::vector<bool> &hit, bool *defined) {
    size_t i = buckets.size();
    while (i > 0 && !hit[i - 1]) --i;
    if (i == 0) {
        *defined = false;
        return G::zero();
    }
    --i;
    G sum = buckets[i], acc = buckets[i];
    while (i > 0) {
        --i;
        if (hit[i]) acc = acc + buckets[i];
        sum = sum + acc;
    }
    return sum;
}

// multi_exp_base_elements_from_fifo_all_rounds (multiexp_stream.tcc:51-122) with
// c = bdlo12_signed_optimal_c(n) (:170, multiexp.tcc:636-640)
template <class G, class Fr>
static G stream_all_rounds(const G *bases, const Fr *exps, size_t n, bool *defined) {
    const size_t c = pippenger_optimal_c(n) + 1;
    const size_t num_digits = (field_num_bits<Fr>() + 2 + c - 1) / c;
    const size_t num_buckets = (size_t)1 << (c - 1);
    std::vector<std::vector<G>> buckets(num_digits, std::vector<G>(num_buckets, G::zero()));
    std::vector<std::vector<bool>> hit(num_digits, std::vector<bool>(num_buckets, false));
    std::vector<long> digits(num_digits);
    for (size_t i = 0; i < n; i++) {
        Big<Fr::N> bi;
        exps[i].as_bigint(bi.d);
        get_signed_digits(digits, bi, c, num_digits);
        for (size_t k = 0; k < num_digits; k++) add_to_bucket(buckets[k], hit[k], bases[i], digits[k]);
    }
    G result = accumulate_buckets(buckets[num_digits - 1], hit[num_digits - 1], defined);
    for (size_t k = num_digits - 1; k-- > 0;) {
        for (size_t i = 0; i < c; i++) result = result.dbl();
        result = result + accumulate_buckets(buckets[k], hit[k], defined);
    }
    return result;
}

// multi_exp_precompute_from_fifo (multiexp_stream.tcc:124-162): table holds, per base, its
// num_digits multiples by 2^(k c); ONE bucket set
template <class G, class Fr>
static G stream_precompute(const G *table, const Fr *exps, size_t n, size_t c, bool *defined) {
    const size_t num_digits = (field_num_bits<Fr>() + c - 1) / c;  // :205
    const size_t num_buckets = (size_t)1 << (c - 1);
    std::vector<G> buckets(num_buckets, G::zero());
    std::vector<bool> hit(num_buckets, false);
    std::vector<long> digits(num_digits);
    for (size_t i = 0; i < n; i++) {
        Big<Fr::N> bi;
        exps[i].as_bigint(bi.d);
        get_signed_digits(digits, bi, c, num_digits);
        for (size_t k = 0; k < num_digits; k++)
            add_to_bucket(buckets, hit, table[i * num_digits + k], digits[k]);
    }
    return accumulate_buckets(buckets, hit, defined);
}

// field_element_codec<encoding_binary, Form, Fp_model> (field_serialization.tcc:197-222): the
// bigint (plain value or mont_repr) with all its bytes reversed; extension fields coefficient by
// coefficient, lowest first (:127-147)
template <int N, int ID>
static void stream_put(std::vector<uint8_t> &out, const Fp<N, ID> &x, bool mont) {
    uint64_t w[N];
    if (mont) {
        memcpy(w, x.v, sizeof w);
    } else {
        x.as_bigint(w);
    }
    const uint8_t *b = (const uint8_t *)w;
    for (size_t k = sizeof w; k-- > 0;) out.push_back(b[k]);
}
template <class F>
static void stream_put(std::vector<uint8_t> &out, const Fp2<F> &x, bool mont) {
    stream_put(out, x.c0, mont);
    stream_put(out, x.c1, mont);
}
template <int N, int ID>
static const uint8_t *stream_get(const uint8_t *in, Fp<N, ID> &x, bool mont) {
    uint64_t w[N];
    uint8_t *b = (uint8_t *)w;
    for (size_t k = 0; k < sizeof w; k++) b[sizeof w - 1 - k] = in[k];
    x = mont ? Fp<N, ID>::zero() : Fp<N, ID>::from_bigint(w);
    if (mont) memcpy(x.v, w, sizeof w);
    return in + sizeof w;
}
template <class F>
static const uint8_t *stream_get(const uint8_t *in, Fp2<F> &x, bool mont) {
    in = stream_get(in, x.c0, mont);
    return stream_get(in, x.c1, mont);
}
// group_element_codec<encoding_binary, Form, compression_off> (curve_serialization.tcc:79-101)
template <class G>
static void stream_put_point(std::vector<uint8_t> &out, G p, bool mont) {
    p.to_special();  // to_affine_coordinates: zero becomes (0, 1, 0)
    stream_put(out, p.X, mont);
    stream_put(out, p.Y, mont);
}
template <class G>
static const uint8_t *stream_get_point(const uint8_t *in, G &p, bool mont) {
    in = stream_get(in, p.X, mont);
    in = stream_get(in, p.Y, mont);
    typedef decltype(p.X) F;
    p.Z = (p.X.is_zero() && p.Y == F::one()) ? F::zero() : F::one();
    return in;
}

// libff::multi_exp (multiexp.tcc:643-688): `chunks` ranges (last takes the remainder), OpenMP over
// ranges, serial sum of the partial results.
template <class G, class Fr>
static G multi_exp(const G *bases, const Fr *exps, size_t total, size_t chunks, bool special,
                   bool signed_digits) {
    auto inner = [&](const G *b, const Fr *e, size_t len) {
        if (signed_digits)
            return special ? bdlo12_signed_inner<G, Fr, true>(b, e, len)
                           : bdlo12_signed_inner<G, Fr, false>(b, e, len);
        return special ? bdlo12_inner<G, Fr, true>(b, e, len) : bdlo12_inner<G, Fr, false>(b, e, len);
    };
    if (total < chunks || chunks == 1) return inner(bases, exps, total);
    const size_t one = total / chunks;
    std::vector<G> partial(chunks, G::zero());
#pragma omp parallel for
    for (size_t i = 0; i < chunks; i++) {
        const size_t lo = i * one, hi = i == chunks - 1 ? total : (i + 1) * one;
        partial[i] = inner(bases + lo, exps + lo, hi - lo);
    }
    G fin = G::zero();
    for (size_t i = 0; i < chunks; i++) fin = fin + partial[i];
    return fin;
}

// batch_invert (field_utils.tcc:419-439): prefix products, ONE inversion, back-substitution
template <class F>
static void batch_invert(std::vector<F> &vec) {
    std::vector<F> prod;
    prod.reserve(vec.size());
    F acc = F::one();
    for (const F &el : vec) {
        prod.push_back(acc);
        acc = acc * el;
    }
    F acc_inverse = acc.inverse();
    for (long i = (long)vec.size() - 1; i >= 0; --i) {
        const F old_el = vec[i];
        vec[i] = acc_inverse * prod[i];
        acc_inverse = acc_inverse * old_el;
    }
}

// batch_to_special (multiexp.tcc:949-974) + G::batch_to_special_all_non_zeros
// (alt_bn128_g1.cpp:456-477, bw6_761_g1.cpp:489-506)
template <class G>
static void batch_to_special(std::vector<G> &vec) {
    std::vector<typename G::Fq> zs;
    for (const G &g : vec)
        if (!g.is_zero()) zs.push_back(g.Z);
    batch_invert(zs);
    size_t k = 0;
    for (G &g : vec) {
        if (!g.is_zero())
            g.apply_zinv(zs[k++]);
        else
            g = G::zero();
    }
}

// get_window_table (multiexp.tcc:809-846): table[outer][inner] = inner * 2^(outer*window) * g,
// last row shorter, unused tail left at zero
template <class G>
static void window_table(size_t scalar_size, size_t window, const G &g, G *out) {
    const size_t in_window = (size_t)1 << window;
    const size_t outerc = (scalar_size + window - 1) / window;
    const size_t last_in_window = (size_t)1 << (scalar_size - (outerc - 1) * window);
    for (size_t i = 0; i < outerc * in_window; i++) out[i] = G::zero();
    G gouter = g;
    for (size_t outer = 0; outer < outerc; outer++) {
        G ginner = G::zero();
        const size_t cur = outer == outerc - 1 ? last_in_window : in_window;
        for (size_t inner = 0; inner < cur; inner++) {
            out[outer * in_window + inner] = ginner;
            ginner = ginner + gouter;
        }
        for (size_t i = 0; i < window; i++) gouter = gouter + gouter;
    }
}

// windowed_exp (multiexp.tcc:848-873)
template <class G, class Fr>
static G windowed_exp(size_t scalar_size, size_t window, const G *table, const Fr &pow) {
    const size_t in_window = (size_t)1 << window;
    const size_t outerc = (scalar_size + window - 1) / window;
    Big<Fr::N> v;
    pow.as_bigint(v.d);
    G res = table[0];
    for (size_t outer = 0; outer < outerc; outer++) {
        size_t inner = 0;
        for (size_t i = 0; i < window; i++)
            if (v.test_bit(outer * window + i)) inner |= (size_t)1 << i;
        res = res + table[outer * in_window + inner];
    }
    return res;
}

// scalar * base, plain double-and-add over as_bigint (the reference's operator* goes through
// wNAF, scalar_multiplication/wnaf.tcc; the group element is the same)
template <class G, class Fr>
static G scalar_mul(const G &base, const Fr &s) {
    Big<Fr::N> v;
    s.as_bigint(v.d);
    G r = G::zero();
    for (long i = (long)v.num_bits() - 1; i >= 0; i--) {
        r = r.dbl();
        if (v.test_bit((size_t)i)) r = r + base;
    }
    return r;
}

// SHA512_rng<FieldT>(idx) (common/rng.tcc:26-71): SHA-512 of idx || iter (two little-endian u64),
// low num_limbs words, bits above the modulus MSB cleared, rejection-sampled; then FieldT(bigint).
template <class Fr>
static Fr sha512_rng(uint64_t idx) {
    constexpr int N = Fr::N;
    for (uint64_t iter = 0;; iter++) {
        uint64_t in[2] = {idx, iter};
        unsigned char hash[SHA512_DIGEST_LENGTH];
        SHA512((const unsigned char *)in, sizeof in, hash);
        uint64_t v[N];
        memcpy(v, hash, sizeof v);
        const int bits = Fr::P.bits;
        for (int b = bits; b < 64 * N; b++) v[b / 64] &= ~((uint64_t)1 << (b % 64));
        if (cmp_n<N>(v, Fr::P.mod) < 0) return Fr::from_bigint(v);
    }
}

// ------------------------------------------------------------------------------------------------
// The four (group, scalar field) pairs of the path (SURVEY.md §8 sizes table)

typedef Fp<4, 0> BnFq;    // alt_bn128_init.cpp:84-104
typedef Fp<4, 1> BnFr;    // alt_bn128_init.cpp:43-63
typedef Fp<6, 2> BlsFq;   // bls12_377_init.cpp:101-128
typedef Fp<4, 3> BlsFr;   // bls12_377_init.cpp:58-82
typedef Fp<12, 4> Bw6Fq;  // bw6_761_init.cpp:81-116
typedef Fp<6, 5> Bw6Fr;   // bw6_761_init.cpp:35-59 (= bls12_377 Fq modulus)
typedef Fp2<BnFq> BnFq2;
// the remaining groups of the same families
typedef Fp<6, 6> Bls381Fq;  // bls12_381_init.cpp:76-116
typedef Fp<4, 7> Bls381Fr;  // bls12_381_init.cpp:36-74
typedef Fp2<BlsFq> BlsFq2;  // u^2 = -5, bls12_377_init.cpp:170-176
typedef Fp2<Bls381Fq> Bls381Fq2;  // u^2 = -1, bls12_381_init.cpp:137-139

typedef Jac<BnFq, 0> BnG1;
typedef Jac<BnFq2, 1> BnG2;
typedef Jac<BlsFq, 2> BlsG1;
typedef Proj<Bw6Fq, 3> Bw6G1;
typedef Jac<BlsFq2, 4> BlsG2;       // bls12_377_g2.cpp (same formulas as alt_bn128_g2.cpp)
typedef Proj<Bw6Fq, 5> Bw6G2;       // bw6_761_g2.cpp (same formulas as bw6_761_g1.cpp, over Fq)
typedef Jac<Bls381Fq, 6> Bls381G1;  // bls12_381_g1.cpp
typedef Jac<Bls381Fq2, 7> Bls381G2; // bls12_381_g2.cpp

template <class F>
static F fq_from_hex(const char *h) {
    uint64_t v[F::N];
    parse_hex<F::N>(h, v);
    return F::from_bigint(v);
}

static BnG1 g_bn_g1;
static BnG2 g_bn_g2;
static BlsG1 g_bls_g1;
static Bw6G1 g_bw6_g1;
static BlsG2 g_bls_g2;
static Bw6G2 g_bw6_g2;
static Bls381G1 g_bls381_g1;
static Bls381G2 g_bls381_g2;
static bool g_ready = false;

static void init_all() {
    if (g_ready) return;
    setup_params(BnFq::P, "30644e72e131a029b85045b68181585d97816a916871ca8d3c208c16d87cfd47");
    setup_params(BnFr::P, "30644e72e131a029b85045b68181585d2833e84879b9709143e1f593f0000001");
    setup_params(BlsFq::P,
                 "1ae3a4617c510eac63b05c06ca1493b1a22d9f300f5138f1ef3622fba094800170b5d44300000008"
                 "508c00000000001");
    setup_params(BlsFr::P, "12ab655e9a2ca55660b44d1e5c37b00159aa76fed00000010a11800000000001");
    setup_params(Bw6Fq::P,
                 "122e824fb83ce0ad187c94004faff3eb926186a81d14688528275ef8087be41707ba638e584e9190"
                 "3cebaff25b423048689c8ed12f9fd9071dcd3dc73ebff2e98a116c25667a8f8160cf8aeeaf0a437e"
                 "6913e6870000082f49d00000000008b");
    setup_params(Bw6Fr::P,
                 "1ae3a4617c510eac63b05c06ca1493b1a22d9f300f5138f1ef3622fba094800170b5d44300000008"
                 "508c00000000001");
    setup_params(Bls381Fq::P,
                 "1a0111ea397fe69a4b1ba7b6434bacd764774b84f38512bf6730d2a0f6b0f6241eabfffeb153ffff"
                 "b9feffffffffaaab");
    setup_params(Bls381Fr::P, "73eda753299d7d483339d80809a1d80553bda402fffe5bfeffffffff00000001");
    BnFq2::non_residue = -BnFq::one();  // q - 1, alt_bn128_init.cpp:138-140
    BlsFq2::non_residue = -fq_from_hex<BlsFq>("5");        // bls12_377_init.cpp:170-176
    Bls381Fq2::non_residue = -Bls381Fq::one();             // bls12_381_init.cpp:137-139
    // generators: bls12_377_init.cpp:392-420, bw6_761_init.cpp:340-360, bls12_381_init.cpp:441-450,
    // 517-534 (decimal literals there; read back from the compiled reference in hexadecimal)
    g_bls_g2 = BlsG2{
        BlsFq2{fq_from_hex<BlsFq>("b997fef930828fe1b9e6a1707b8aa508a3dbfd7fe2246499c709226a0a6fef49"
                                  "f85b3a375363f4f8f6ea3fbd159f8a"),
               fq_from_hex<BlsFq>("d6ac33b84947d9845f81a57a136bfa326e915fabc8cd6a57ff133b42d00f62e4"
                                  "e1af460228cd5184deae976fa62596")},
        BlsFq2{fq_from_hex<BlsFq>("118dd509b2e9a13744a507d515a595dbb7e3b63df568866473790184bdf83636"
                                  "c94df2b7a962cb2af4337f07cb7e622"),
               fq_from_hex<BlsFq>("185067c6ca76d992f064a432bd9f9be832b0cac2d824d0518f77d39e76c3e146"
                                  "afb825f2092218d038867d7f337a010")},
        BlsFq2::one()};
    g_bw6_g2 = Bw6G2{
        fq_from_hex<Bw6Fq>("110133241d9b816c852a82e69d660f9d61053aac5a7115f4c06201013890f6d26b41c5da"
                           "b3da268734ec3f1f09feb58c5bbcae9ac70e7c7963317a300e1b6bace6948cb3cd208d70"
                           "0e96efbc2ad54b06410cf4fe1bf995ba830c194cd025f1c"),
        fq_from_hex<Bw6Fq>("17c3357761369f8179eb10e4b6d2dc26b7cf9acec2181c81a78e2753ffe3160a1d86c80b"
                           "95a59c94c97eb733293fef64f293dbd2c712b88906c170ffa823003ea96fcd504affc758"
                           "aa2d3a3c5a02a591ec0594f9eac689eb70a16728c73b61"),
        Bw6Fq::one()};
    g_bls381_g1 = Bls381G1{
        fq_from_hex<Bls381Fq>("17f1d3a73197d7942695638c4fa9ac0fc3688c4f9774b905a14e3a3f171bac586c55"
                              "e83ff97a1aeffb3af00adb22c6bb"),
        fq_from_hex<Bls381Fq>("8b3f481e3aaa0f1a09e30ed741d8ae4fcf5e095d5d00af600db18cb2c04b3edd03cc"
                              "744a2888ae40caa232946c5e7e1"),
        Bls381Fq::one()};
    g_bls381_g2 = Bls381G2{
        Bls381Fq2{fq_from_hex<Bls381Fq>("24aa2b2f08f0a91260805272dc51051c6e47ad4fa403b02b4510b647ae"
                                        "3d1770bac0326a805bbefd48056c8c121bdb8"),
                  fq_from_hex<Bls381Fq>("13e02b6052719f607dacd3a088274f65596bd0d09920b61ab5da61bbdc"
                                        "7f5049334cf11213945d57e5ac7d055d042b7e")},
        Bls381Fq2{fq_from_hex<Bls381Fq>("ce5d527727d6e118cc9cdc6da2e351aadfd9baa8cbdd3a76d429a6951"
                                        "60d12c923ac9cc3baca289e193548608b82801"),
                  fq_from_hex<Bls381Fq>("606c4a02ea734cc32acd2b02bc28b99cb3e287e85a763af267492ab57"
                                        "2e99ab3f370d275cec1da1aaa9075ff05f79be")},
        Bls381Fq2::one()};
    // generators: alt_bn128_init.cpp:289-290,358-369; bls12_377_init.cpp:324-331;
    // bw6_761_init.cpp:287-296 (decimal literals there, hexadecimal here)
    g_bn_g1 = BnG1{fq_from_hex<BnFq>("1"), fq_from_hex<BnFq>("2"), BnFq::one()};
    g_bn_g2 = BnG2{
        BnFq2{fq_from_hex<BnFq>("1800deef121f1e76426a00665e5c4479674322d4f75edadd46debd5cd992f6ed"),
              fq_from_hex<BnFq>("198e9393920d483a7260bfb731fb5d25f1aa493335a9e71297e485b7aef312c2")},
        BnFq2{fq_from_hex<BnFq>("12c85ea5db8c6deb4aab71808dcb408fe3d1e7690c43d37b4ce6cc0166fa7daa"),
              fq_from_hex<BnFq>("90689d0585ff075ec9e99ad690c3395bc4b313370b38ef355acdadcd122975b")},
        BnFq2::one()};
    g_bls_g1 = BlsG1{
        fq_from_hex<BlsFq>("8848defe740a67c8fc6225bf87ff5485951e2caa9d41bb188282c8bd37cb5cd5481512ff"
                           "cd394eeab9b16eb21be9ef"),
        fq_from_hex<BlsFq>("1914a69c5102eff1f674f5d30afeec4bd7fb348ca3e52d96d182ad44fb82305c2fe3d363"
                           "4a9591afd82de55559c8ea6"),
        BlsFq::one()};
    g_bw6_g1 = Bw6G1{
        fq_from_hex<Bw6Fq>("1075b020ea190c8b277ce98a477beaee6a0cfb7551b27f0ee05c54b85f56fc779017ffac"
                           "15520ac11dbfcd294c2e746a17a54ce47729b905bd71fa0c9ea097103758f9a280ca27f6"
                           "750dd0356133e82055928aca6af603f4088f3af66e5b43d"),
        fq_from_hex<Bw6Fq>("58b84e0a6fc574e6fd637b45cc2a420f952589884c9ec61a7348d2a2e573a3265909f1af"
                           "7e0dbac5b8fa1771b5b806cc685d31717a4c55be3fb90b6fc2cdd49f9df141b3053253b2"
                           "b08119cad0fb93ad1cb2be0b20d2a1bafc8f2db4e95363"),
        Bw6Fq::one()};
    g_ready = true;
}

template <class G> static const G &generator();
template <> const BnG1 &generator<BnG1>() { return g_bn_g1; }
template <> const BnG2 &generator<BnG2>() { return g_bn_g2; }
template <> const BlsG1 &generator<BlsG1>() { return g_bls_g1; }
template <> const Bw6G1 &generator<Bw6G1>() { return g_bw6_g1; }
template <> const BlsG2 &generator<BlsG2>() { return g_bls_g2; }
template <> const Bw6G2 &generator<Bw6G2>() { return g_bw6_g2; }
template <> const Bls381G1 &generator<Bls381G1>() { return g_bls381_g1; }
template <> const Bls381G2 &generator<Bls381G2>() { return g_bls381_g2; }

template <class G, class Fr>
struct Ops {
    static constexpr size_t GW = 3 * G::Fq::WORDS;
    static std::vector<G> load_points(const void *p, size_t n) {
        std::vector<G> v(n);
        for (size_t i = 0; i < n; i++) v[i] = load_point<G>((const uint64_t *)p + i * GW);
        return v;
    }
    static void store_points(void *p, const std::vector<G> &v) {
        for (size_t i = 0; i < v.size(); i++) store_point<G>((uint64_t *)p + i * GW, v[i]);
    }
    static std::vector<Fr> load_scalars(const void *p, size_t n) {
        std::vector<Fr> v(n);
        for (size_t i = 0; i < n; i++) v[i].load((const uint64_t *)p + i * Fr::N);
        return v;
    }
    static void gen_scalars(uint64_t start, size_t n, void *out) {
#pragma omp parallel for
        for (size_t i = 0; i < n; i++) sha512_rng<Fr>(start + i).store((uint64_t *)out + i * Fr::N);
    }
    // arithmetic walk P_i = (a + i d) G, normalised
    static void gen_bases_walk(size_t start, size_t n, void *out) {
        const Fr a = sha512_rng<Fr>(1ull << 63), d = sha512_rng<Fr>((1ull << 63) + 1);
        const G D = scalar_mul(generator<G>(), d);
        std::vector<G> v(n);
        const size_t seg = 4096, nseg = (n + seg - 1) / seg;
#pragma omp parallel for
        for (size_t s = 0; s < nseg; s++) {
            const size_t lo = s * seg, hi = std::min(n, lo + seg);
            uint64_t k[Fr::N] = {start + lo};
            G cur = scalar_mul(generator<G>(), a + Fr::from_bigint(k) * d);
            for (size_t i = lo; i < hi; i++) {
                v[i] = cur;
                cur = cur + D;
            }
        }
        batch_to_special(v);
        store_points(out, v);
    }
    static double multi_exp_(int method, int form, const void *bases, const void *scalars, size_t n,
                             size_t chunks, void *out, int normalize) {
        std::vector<G> b = load_points(bases, n);
        std::vector<Fr> s = load_scalars(scalars, n);
        auto t0 = std::chrono::steady_clock::now();
        G r = multi_exp<G, Fr>(b.data(), s.data(), n, chunks, form == 1, method == 1);
        double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (normalize) r.to_special();
        store_point<G>((uint64_t *)out, r);
        return sec;
    }
    // c != 0: every element followed by its multiples by 2^c (profile_multiexp.cpp:120-151)
    static size_t stream_write(int form, const void *in, size_t n, size_t c, void *out) {
        const size_t per = c ? (field_num_bits<Fr>() + c - 1) / c : 1;
        std::vector<uint8_t> bytes;
        for (size_t i = 0; i < n; i++) {
            G el = load_point<G>((const uint64_t *)in + i * GW);
            stream_put_point(bytes, el, form != 0);
            for (size_t k = 1; k < per; k++) {
                for (size_t j = 0; j < c; j++) el = el.dbl();
                stream_put_point(bytes, el, form != 0);
            }
        }
        if (out) memcpy(out, bytes.data(), bytes.size());
        return bytes.size();
    }
    // multi_exp_stream (c == 0) / multi_exp_stream_with_precompute; 1 computed, 0 undefined in the
    // reference (an empty bucket set is summed)
    static int multi_exp_stream_(int form, const void *bytes, size_t nbytes, const void *scalars,
                                 size_t n, size_t c, void *out, int normalize) {
        const size_t per = c ? (field_num_bits<Fr>() + c - 1) / c : 1;
        std::vector<G> el(n * per);
        const uint8_t *in = (const uint8_t *)bytes;
        if (nbytes != n * per * 2 * sizeof(G) / 3) return 0;
        for (size_t i = 0; i < n * per; i++) in = stream_get_point(in, el[i], form != 0);
        std::vector<Fr> s = load_scalars(scalars, n);
        bool defined = true;
        G r = c ? stream_precompute<G, Fr>(el.data(), s.data(), n, c, &defined)
                : stream_all_rounds<G, Fr>(el.data(), s.data(), n, &defined);
        if (!defined) return 0;
        if (normalize) r.to_special();
        store_point<G>((uint64_t *)out, r);
        return 1;
    }
    static void group_op(int op, const void *a, const void *b, void *out, size_t n) {
        for (size_t i = 0; i < n; i++) {
            G x = load_point<G>((const uint64_t *)a + i * GW), y = G::zero(), r;
            if (b) y = load_point<G>((const uint64_t *)b + i * GW);
            switch (op) {
                case 0: r = x + y; break;
                case 1: r = x + y; break;  // add(): same element as operator+
                case 2: r = x.mixed_add(y); break;
                case 3: r = x.dbl(); break;
                case 4: r = -x; break;
                default: r = x; r.to_special(); break;
            }
            store_point<G>((uint64_t *)out + i * GW, r);
        }
    }
};

#define DISPATCH(gid, call)                                     \
    switch (gid) {                                              \
        case 0: return Ops<BnG1, BnFr>::call;                   \
        case 1: return Ops<BnG2, BnFr>::call;                   \
        case 2: return Ops<BlsG1, BlsFr>::call;                 \
        case 3: return Ops<Bw6G1, Bw6Fr>::call;                 \
        case 4: return Ops<BlsG2, BlsFr>::call;                 \
        case 5: return Ops<Bw6G2, Bw6Fr>::call;                 \
        case 6: return Ops<Bls381G1, Bls381Fr>::call;           \
        case 7: return Ops<Bls381G2, Bls381Fr>::call;           \
    }

template <class F>
static void field_op(int op, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t n) {
    constexpr int W = F::WORDS;
    for (size_t i = 0; i < n; i++) {
        F x, y = F::zero(), r;
        x.load(a + i * W);
        if (b) y.load(b + i * W);
        switch (op) {
            case 0: r = x * y; break;
            case 1: r = x.squared(); break;
            case 2: r = x + y; break;
            case 3: r = x - y; break;
            case 4: r = -x; break;
            default: r = x.inverse(); break;
        }
        r.store(out + i * W);
    }
}
template <class F>
static void field_convert(int op, const uint64_t *a, uint64_t *out, size_t n) {
    for (size_t i = 0; i < n; i++) {
        if (op == 6) {
            F x;
            x.load(a + i * F::N);
            x.as_bigint(out + i * F::N);
        } else {
            F::from_bigint(a + i * F::N).store(out + i * F::N);
        }
    }
}
template <class F>
static long signed_digit_of(const uint64_t *mont, size_t c, size_t idx) {
    F x;
    x.load(mont);
    Big<F::N> v;
    x.as_bigint(v.d);
    return get_signed_digit(v, c, idx);
}
// field_get_num_signed_digits (field_utils.tcc:102-165)
template <class F>
static size_t num_signed_digits(size_t c) {
    Big<F::N> m1;
    (-F::one()).as_bigint(m1.d);
    const size_t naive = (F::P.bits + 1 + c - 1) / c;
    const size_t sign_bit = (size_t)1 << (c - 1), max_signed = sign_bit - 1;
    for (size_t i = naive - 1; i < naive; --i) {
        const size_t d = get_digit(m1, c, i);
        if (d & sign_bit) return naive + 1;
        if (d != max_signed) break;
    }
    return naive;
}

#define FIELD_SWITCH(fid, expr)                    \
    switch (fid) {                                 \
        case 0: { using F = BnFq; expr; } break;   \
        case 1: { using F = BnFr; expr; } break;   \
        case 2: { using F = BlsFq; expr; } break;  \
        case 3: { using F = BlsFr; expr; } break;  \
        case 4: { using F = Bw6Fq; expr; } break;  \
        case 5: { using F = Bw6Fr; expr; } break;  \
        case 8: { using F = Bls381Fq; expr; } break; \
        case 9: { using F = Bls381Fr; expr; } break; \
    }

// ------------------------------------------------------------------------------------------------
// C entry points; same shapes as oracle/ref_driver.cpp's ref_* so that oracle/port.py and
// oracle/ref.py are interchangeable checkers.
template <class G, class Fr>
static void scalar_mul_c(const void *base, const void *scalar, void *out, int normalize) {
    G b = load_point<G>((const uint64_t *)base);
    Fr s;
    s.load((const uint64_t *)scalar);
    G r = scalar_mul(b, s);
    if (normalize) r.to_special();
    store_point<G>((uint64_t *)out, r);
}

template <class G>
static void to_special_c(void *inout, size_t n) {
    std::vector<G> v = Ops<G, BnFr>::load_points(inout, n);
    batch_to_special(v);
    Ops<G, BnFr>::store_points(inout, v);
}

template <class G>
static void window_table_c(size_t scalar_bits, size_t window, const void *g, void *out) {
    const size_t rows = (scalar_bits + window - 1) / window;
    std::vector<G> t(rows << window);
    window_table<G>(scalar_bits, window, load_point<G>((const uint64_t *)g), t.data());
    Ops<G, BnFr>::store_points(out, t);
}

// batch_exp (multiexp.tcc:875-915): OpenMP over scalars
template <class G, class Fr>
static double batch_exp_c(size_t scalar_bits, size_t window, const void *table, size_t rows,
                          size_t row_len, const void *scalars, size_t n, void *out, int normalize) {
    std::vector<G> t = Ops<G, Fr>::load_points(table, rows * row_len);
    std::vector<Fr> s = Ops<G, Fr>::load_scalars(scalars, n);
    std::vector<G> res(n, G::zero());
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for
    for (size_t i = 0; i < n; i++) res[i] = windowed_exp<G, Fr>(scalar_bits, window, t.data(), s[i]);
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (normalize) batch_to_special(res);
    Ops<G, Fr>::store_points(out, res);
    return sec;
}

template <class G>
static void group_eq_c(const void *a, const void *b, uint8_t *out, size_t n) {
    constexpr size_t GW = 3 * G::Fq::WORDS;
    for (size_t i = 0; i < n; i++)
        out[i] = load_point<G>((const uint64_t *)a + i * GW) == load_point<G>((const uint64_t *)b + i * GW);
}

extern "C" {

int orc_init() {
    init_all();
    return 1;
}
int orc_num_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void orc_set_num_threads(int t) {
#ifdef _OPENMP
    omp_set_num_threads(t);
#endif
}
size_t orc_group_bytes(int gid) {
    static const size_t b[8] = {96, 192, 144, 288, 288, 288, 144, 288};
    return gid >= 0 && gid < 8 ? b[gid] : 0;
}
size_t orc_scalar_bytes(int gid) {
    static const size_t b[8] = {32, 32, 32, 48, 32, 48, 32, 32};
    return gid >= 0 && gid < 8 ? b[gid] : 0;
}
size_t orc_scalar_bits(int gid) {
    init_all();
    switch (gid) {
        case 0: case 1: return BnFr::P.bits;
        case 2: case 4: return BlsFr::P.bits;
        case 3: case 5: return Bw6Fr::P.bits;
        case 6: case 7: return Bls381Fr::P.bits;
    }
    return 0;
}
void orc_gen_scalars(int gid, uint64_t start, size_t n, void *out) {
    init_all();
    DISPATCH(gid, gen_scalars(start, n, out));
}
void orc_gen_bases_walk(int gid, size_t start, size_t n, void *out) {
    init_all();
    DISPATCH(gid, gen_bases_walk(start, n, out));
}
void orc_scalar_mul(int gid, const void *base, const void *scalar, void *out, int normalize) {
    init_all();
    switch (gid) {
        case 0: return scalar_mul_c<BnG1, BnFr>(base, scalar, out, normalize);
        case 1: return scalar_mul_c<BnG2, BnFr>(base, scalar, out, normalize);
        case 2: return scalar_mul_c<BlsG1, BlsFr>(base, scalar, out, normalize);
        case 3: return scalar_mul_c<Bw6G1, Bw6Fr>(base, scalar, out, normalize);
        case 4: return scalar_mul_c<BlsG2, BlsFr>(base, scalar, out, normalize);
        case 5: return scalar_mul_c<Bw6G2, Bw6Fr>(base, scalar, out, normalize);
        case 6: return scalar_mul_c<Bls381G1, Bls381Fr>(base, scalar, out, normalize);
        case 7: return scalar_mul_c<Bls381G2, Bls381Fr>(base, scalar, out, normalize);
    }
}
// method: 0 BDLO12, 1 BDLO12_signed; form: 0 normal, 1 special.  Returns seconds inside multi_exp.
double orc_multi_exp(int gid, int method, int form, const void *bases, const void *scalars,
                     size_t n, size_t chunks, void *out, int normalize) {
    init_all();
    DISPATCH(gid, multi_exp_(method, form, bases, scalars, n, chunks, out, normalize));
    return -1.0;
}
size_t orc_stream_write(int gid, int form, const void *in, size_t n, size_t c, void *out) {
    init_all();
    DISPATCH(gid, stream_write(form, in, n, c, out));
    return 0;
}
int orc_multi_exp_stream(int gid, int form, const void *bytes, size_t nbytes, const void *scalars,
                         size_t n, size_t c, void *out, int normalize) {
    init_all();
    DISPATCH(gid, multi_exp_stream_(form, bytes, nbytes, scalars, n, c, out, normalize));
    return -1;
}
size_t orc_bdlo12_signed_optimal_c(size_t n) { return pippenger_optimal_c(n) + 1; }
void orc_batch_to_special(int gid, void *inout, size_t n) {
    init_all();
    switch (gid) {
        case 0: return to_special_c<BnG1>(inout, n);
        case 1: return to_special_c<BnG2>(inout, n);
        case 2: return to_special_c<BlsG1>(inout, n);
        case 3: return to_special_c<Bw6G1>(inout, n);
        case 4: return to_special_c<BlsG2>(inout, n);
        case 5: return to_special_c<Bw6G2>(inout, n);
        case 6: return to_special_c<Bls381G1>(inout, n);
        case 7: return to_special_c<Bls381G2>(inout, n);
    }
}
size_t orc_window_table_rows(int gid, size_t scalar_bits, size_t window) {
    (void)gid;
    return (scalar_bits + window - 1) / window;
}
void orc_window_table(int gid, size_t scalar_bits, size_t window, const void *g, void *out) {
    init_all();
    switch (gid) {
        case 0: return window_table_c<BnG1>(scalar_bits, window, g, out);
        case 1: return window_table_c<BnG2>(scalar_bits, window, g, out);
        case 2: return window_table_c<BlsG1>(scalar_bits, window, g, out);
        case 3: return window_table_c<Bw6G1>(scalar_bits, window, g, out);
        case 4: return window_table_c<BlsG2>(scalar_bits, window, g, out);
        case 5: return window_table_c<Bw6G2>(scalar_bits, window, g, out);
        case 6: return window_table_c<Bls381G1>(scalar_bits, window, g, out);
        case 7: return window_table_c<Bls381G2>(scalar_bits, window, g, out);
    }
}
double orc_batch_exp(int gid, size_t scalar_bits, size_t window, const void *table, size_t rows,
                     size_t row_len, const void *scalars, size_t n, void *out, int normalize) {
    init_all();
    switch (gid) {
        case 0: return batch_exp_c<BnG1, BnFr>(scalar_bits, window, table, rows, row_len, scalars, n, out, normalize);
        case 1: return batch_exp_c<BnG2, BnFr>(scalar_bits, window, table, rows, row_len, scalars, n, out, normalize);
        case 2: return batch_exp_c<BlsG1, BlsFr>(scalar_bits, window, table, rows, row_len, scalars, n, out, normalize);
        case 3: return batch_exp_c<Bw6G1, Bw6Fr>(scalar_bits, window, table, rows, row_len, scalars, n, out, normalize);
        case 4: return batch_exp_c<BlsG2, BlsFr>(scalar_bits, window, table, rows, row_len, scalars, n, out, normalize);
        case 5: return batch_exp_c<Bw6G2, Bw6Fr>(scalar_bits, window, table, rows, row_len, scalars, n, out, normalize);
        case 6: return batch_exp_c<Bls381G1, Bls381Fr>(scalar_bits, window, table, rows, row_len, scalars, n, out, normalize);
        case 7: return batch_exp_c<Bls381G2, Bls381Fr>(scalar_bits, window, table, rows, row_len, scalars, n, out, normalize);
    }
    return -1.0;
}
void orc_group_op(int gid, int op, const void *a, const void *b, void *out, size_t n) {
    init_all();
    DISPATCH(gid, group_op(op, a, b, out, n));
}
void orc_group_eq(int gid, const void *a, const void *b, uint8_t *out, size_t n) {
    init_all();
    switch (gid) {
        case 0: return group_eq_c<BnG1>(a, b, out, n);
        case 1: return group_eq_c<BnG2>(a, b, out, n);
        case 2: return group_eq_c<BlsG1>(a, b, out, n);
        case 3: return group_eq_c<Bw6G1>(a, b, out, n);
        case 4: return group_eq_c<BlsG2>(a, b, out, n);
        case 5: return group_eq_c<Bw6G2>(a, b, out, n);
        case 6: return group_eq_c<Bls381G1>(a, b, out, n);
        case 7: return group_eq_c<Bls381G2>(a, b, out, n);
    }
}
void orc_group_constants(int gid, void *zero, void *one) {
    init_all();
    switch (gid) {
        case 0: store_point((uint64_t *)zero, BnG1::zero()); store_point((uint64_t *)one, g_bn_g1); break;
        case 1: store_point((uint64_t *)zero, BnG2::zero()); store_point((uint64_t *)one, g_bn_g2); break;
        case 2: store_point((uint64_t *)zero, BlsG1::zero()); store_point((uint64_t *)one, g_bls_g1); break;
        case 3: store_point((uint64_t *)zero, Bw6G1::zero()); store_point((uint64_t *)one, g_bw6_g1); break;
        case 4: store_point((uint64_t *)zero, BlsG2::zero()); store_point((uint64_t *)one, g_bls_g2); break;
        case 5: store_point((uint64_t *)zero, Bw6G2::zero()); store_point((uint64_t *)one, g_bw6_g2); break;
        case 6: store_point((uint64_t *)zero, Bls381G1::zero()); store_point((uint64_t *)one, g_bls381_g1); break;
        case 7: store_point((uint64_t *)zero, Bls381G2::zero()); store_point((uint64_t *)one, g_bls381_g2); break;
    }
}
size_t orc_field_bytes(int fid) {
    static const size_t b[11] = {32, 32, 48, 32, 96, 48, 64, 96, 48, 32, 96};
    return fid >= 0 && fid < 11 ? b[fid] : 0;
}
void orc_field_op(int fid, int op, const void *a, const void *b, void *out, size_t n) {
    init_all();
    const uint64_t *pa = (const uint64_t *)a, *pb = (const uint64_t *)b;
    uint64_t *po = (uint64_t *)out;
    if (fid == 6) return field_op<BnFq2>(op, pa, pb, po, n);
    if (fid == 7) return field_op<BlsFq2>(op, pa, pb, po, n);
    if (fid == 10) return field_op<Bls381Fq2>(op, pa, pb, po, n);
    if (op >= 6) {
        FIELD_SWITCH(fid, field_convert<F>(op, pa, po, n));
    } else {
        FIELD_SWITCH(fid, field_op<F>(op, pa, pb, po, n));
    }
}
long orc_signed_digit(int fid, const void *scalar_mont, size_t c, size_t idx) {
    init_all();
    FIELD_SWITCH(fid, return signed_digit_of<F>((const uint64_t *)scalar_mont, c, idx));
    return 0;
}
// field_get_digit (field_utils.tcc:50-100) of the plain value of a Montgomery-form scalar
size_t orc_get_digit(int fid, const void *scalar_mont, size_t c, size_t idx) {
    init_all();
    FIELD_SWITCH(fid, {
        F x;
        x.load((const uint64_t *)scalar_mont);
        Big<F::N> v;
        x.as_bigint(v.d);
        return get_digit(v, c, idx);
    });
    return 0;
}
size_t orc_num_signed_digits(int fid, size_t c) {
    init_all();
    FIELD_SWITCH(fid, return num_signed_digits<F>(c));
    return 0;
}
size_t orc_pippenger_optimal_c(size_t n) { return pippenger_optimal_c(n); }

}  // extern "C"
