// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// Thin extern "C" driver around the UNMODIFIED reference (clearmatics/libff,
// compiled from the sources where they lie under /root/reference by
// oracle/Makefile into oracle/_ref/libffref.so).  It exposes the reference's
// own implementation of the MSM hot path over raw in-memory
// buffers so that tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// `--impl reference` legs can (a) pin the restatement in oracle/msm_oracle.cpp,
// (b) check the CUDA path, (c) time libff's CPU multi_exp on the host cores.
//
// Every buffer is the reference's in-memory layout: 64-bit little-endian
// limbs, Montgomery form, coordinates X|Y|Z contiguous, stride sizeof(GroupT).
//
// Reference entry points exercised (file:line in /root/reference):
//   multi_exp                     libff/algebra/scalar_multiplication/multiexp.tcc:643-688
//   multi_exp_inner (BDLO12)      multiexp.tcc:276-381
//   multi_exp_inner (BDLO12_signed) multiexp.tcc:507-633
//   batch_exp / get_window_table  multiexp.tcc:809-915
//   batch_to_special              multiexp.tcc:949-974
//   field_get_signed_digit        libff/algebra/fields/field_utils.tcc:167-203
//   SHA512_rng                    libff/common/rng.tcc:26-71
//   ffi::group_element_read/write, field_element_read/write   ffi/ffi_serialization.tcc:98-187
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <libff/algebra/curves/alt_bn128/alt_bn128_pp.hpp>
#include <libff/algebra/curves/bls12_377/bls12_377_pp.hpp>
#include <libff/algebra/curves/bls12_381/bls12_381_pp.hpp>
#include <libff/algebra/curves/bw6_761/bw6_761_pp.hpp>
#include <libff/algebra/fields/field_utils.hpp>
#include <libff/algebra/scalar_multiplication/multiexp.hpp>
#include <libff/algebra/scalar_multiplication/multiexp_stream.hpp>
#include <sstream>
#include <libff/common/profiling.hpp>
#include <libff/common/rng.hpp>

// the reference's own byte format (header-only templates; ffi.cpp itself is not compiled)
#include <ffi/ffi_serialization.hpp>
#include <ffi/ffi_serialization.tcc>

#ifdef _OPENMP
#include <omp.h>
#endif

using namespace libff;

namespace
{

enum group_id {
    G_ALT_BN128_G1 = 0,
    G_ALT_BN128_G2 = 1,
    G_BLS12_377_G1 = 2,
    G_BW6_761_G1 = 3,
    G_BLS12_377_G2 = 4,
    G_BW6_761_G2 = 5,
    G_BLS12_381_G1 = 6,
    G_BLS12_381_G2 = 7
};
enum field_id {
    F_ALT_BN128_FQ = 0,
    F_ALT_BN128_FR = 1,
    F_BLS12_377_FQ = 2,
    F_BLS12_377_FR = 3,
    F_BW6_761_FQ = 4,
    F_BW6_761_FR = 5,
    F_ALT_BN128_FQ2 = 6,
    F_BLS12_377_FQ2 = 7,
    F_BLS12_381_FQ = 8,
    F_BLS12_381_FR = 9,
    F_BLS12_381_FQ2 = 10
};

template<typename T> std::vector<T> load_vec(const void *p, size_t n)
{
    // Group/field objects are trivially copyable aggregates of limbs.
    std::vector<T> v(n);
    if (n) {
        std::memcpy((void *)v.data(), p, n * sizeof(T));
    }
    return v;
}

template<typename T> void store_vec(void *p, const std::vector<T> &v)
{
    if (!v.empty()) {
        std::memcpy(p, (const void *)v.data(), v.size() * sizeof(T));
    }
}

template<typename G, typename Fr> struct ops {
    static void gen_scalars(uint64_t start, size_t n, void *out)
    {
        std::vector<Fr> v(n);
#pragma omp parallel for
        for (size_t i = 0; i < n; ++i) {
            v[i] = SHA512_rng<Fr>(start + i);
        }
        store_vec(out, v);
    }

    // Arithmetic walk P_i = P_0 + i*D, P_0 = a*G, D = d*G with
    // (a, d) = (SHA512_rng(2^63), SHA512_rng(2^63 + 1)); SURVEY.md §8d.
    static void gen_bases_walk(size_t start, size_t n, void *out)
    {
        const Fr a = SHA512_rng<Fr>(1ull << 63);
        const Fr d = SHA512_rng<Fr>((1ull << 63) + 1);
        const G D = d * G::one();
        std::vector<G> v(n);
        const size_t seg = 4096;
        const size_t nseg = (n + seg - 1) / seg;
#pragma omp parallel for
        for (size_t s = 0; s < nseg; ++s) {
            const size_t lo = s * seg;
            const size_t hi = std::min(n, lo + seg);
            Fr idx = Fr((unsigned long)(start + lo));
            G cur = (a + idx * d) * G::one();
            for (size_t i = lo; i < hi; ++i) {
                v[i] = cur;
                cur = cur + D;
            }
        }
        batch_to_special<G>(v);
        store_vec(out, v);
    }

    static void scalar_mul(const void *base, const void *scalar, void *out, int normalize)
    {
        G b;
        Fr s;
        std::memcpy((void *)&b, base, sizeof(G));
        std::memcpy((void *)&s, scalar, sizeof(Fr));
        G r = s * b;
        if (normalize) {
            r.to_affine_coordinates();
        }
        std::memcpy(out, (const void *)&r, sizeof(G));
    }

    template<multi_exp_method M, multi_exp_base_form F>
    static G run(const std::vector<G> &b, const std::vector<Fr> &s, size_t chunks)
    {
        return multi_exp<G, Fr, M, F>(b.begin(), b.end(), s.begin(), s.end(), chunks);
    }

    static int multi_exp_(int method, int form, const void *bases, const void *scalars, size_t n,
                          size_t chunks, void *out, int normalize)
    {
        const std::vector<G> b = load_vec<G>(bases, n);
        const std::vector<Fr> s = load_vec<Fr>(scalars, n);
        G r;
        switch (method * 2 + form) {
        case 0: r = run<multi_exp_method_BDLO12, multi_exp_base_form_normal>(b, s, chunks); break;
        case 1: r = run<multi_exp_method_BDLO12, multi_exp_base_form_special>(b, s, chunks); break;
        case 2: r = run<multi_exp_method_BDLO12_signed, multi_exp_base_form_normal>(b, s, chunks); break;
        case 3: r = run<multi_exp_method_BDLO12_signed, multi_exp_base_form_special>(b, s, chunks); break;
        case 4: r = run<multi_exp_method_naive_plain, multi_exp_base_form_normal>(b, s, chunks); break;
        case 5: r = run<multi_exp_method_naive_plain, multi_exp_base_form_special>(b, s, chunks); break;
        case 6: r = run<multi_exp_method_naive, multi_exp_base_form_normal>(b, s, chunks); break;
        case 7: r = run<multi_exp_method_naive, multi_exp_base_form_special>(b, s, chunks); break;
        default: return 0;
        }
        if (normalize) {
            r.to_affine_coordinates();
        }
        std::memcpy(out, (const void *)&r, sizeof(G));
        return 1;
    }

    // Same as multi_exp_ but the vectors are views the caller already owns as
    // reference objects: avoids the copy so the timed region is the MSM alone.
    static double multi_exp_timed(int form, const void *bases, const void *scalars, size_t n,
                                  size_t chunks, void *out)
    {
        const std::vector<G> b = load_vec<G>(bases, n);
        const std::vector<Fr> s = load_vec<Fr>(scalars, n);
        const long long t0 = get_nsec_time();
        G r = form ? run<multi_exp_method_BDLO12, multi_exp_base_form_special>(b, s, chunks)
                   : run<multi_exp_method_BDLO12, multi_exp_base_form_normal>(b, s, chunks);
        const long long t1 = get_nsec_time();
        r.to_affine_coordinates();
        std::memcpy(out, (const void *)&r, sizeof(G));
        return (double)(t1 - t0) * 1e-9;
    }

    static void batch_to_special_(void *inout, size_t n)
    {
        std::vector<G> v = load_vec<G>(inout, n);
        batch_to_special<G>(v);
        store_vec(inout, v);
    }

    static size_t window_table_rows(size_t scalar_bits, size_t window)
    {
        return (scalar_bits + window - 1) / window;
    }

    static void window_table_(size_t scalar_bits, size_t window, const void *g, void *out)
    {
        G base;
        std::memcpy((void *)&base, g, sizeof(G));
        const window_table<G> t = get_window_table<G>(scalar_bits, window, base);
        char *p = (char *)out;
        for (size_t r = 0; r < t.size(); ++r) {
            std::memcpy(p, (const void *)t[r].data(), t[r].size() * sizeof(G));
            p += t[r].size() * sizeof(G);
        }
    }

    static double batch_exp_(size_t scalar_bits, size_t window, const void *table, size_t rows,
                             size_t row_len, const void *scalars, size_t n, void *out,
                             int normalize)
    {
        window_table<G> t(rows);
        const char *p = (const char *)table;
        for (size_t r = 0; r < rows; ++r) {
            t[r] = load_vec<G>(p + r * row_len * sizeof(G), row_len);
        }
        const std::vector<Fr> s = load_vec<Fr>(scalars, n);
        const long long t0 = get_nsec_time();
        std::vector<G> res = batch_exp<G, Fr>(scalar_bits, window, t, s);
        const long long t1 = get_nsec_time();
        if (normalize) {
            batch_to_special<G>(res);
        }
        store_vec(out, res);
        return (double)(t1 - t0) * 1e-9;
    }

    // op: 0 add (operator+), 1 add(), 2 mixed_add, 3 dbl, 4 neg, 5 to_affine
    static void group_op(int op, const void *a, const void *b, void *out, size_t n)
    {
        std::vector<G> va = load_vec<G>(a, n);
        std::vector<G> vb = b ? load_vec<G>(b, n) : std::vector<G>();
        std::vector<G> r(n);
        for (size_t i = 0; i < n; ++i) {
            switch (op) {
            case 0: r[i] = va[i] + vb[i]; break;
            case 1: r[i] = va[i].add(vb[i]); break;
            case 2: r[i] = va[i].mixed_add(vb[i]); break;
            case 3: r[i] = va[i].dbl(); break;
            case 4: r[i] = -va[i]; break;
            case 5:
                r[i] = va[i];
                r[i].to_affine_coordinates();
                break;
            }
        }
        store_vec(out, r);
    }

    static void group_eq(const void *a, const void *b, uint8_t *out, size_t n)
    {
        std::vector<G> va = load_vec<G>(a, n);
        std::vector<G> vb = load_vec<G>(b, n);
        for (size_t i = 0; i < n; ++i) {
            out[i] = (va[i] == vb[i]) ? 1 : 0;
        }
    }

    static void constants(void *zero, void *one)
    {
        G z = G::zero();
        G o = G::one();
        std::memcpy(zero, (const void *)&z, sizeof(G));
        std::memcpy(one, (const void *)&o, sizeof(G));
    }
};

typedef ops<alt_bn128_G1, alt_bn128_Fr> O0;
typedef ops<alt_bn128_G2, alt_bn128_Fr> O1;
typedef ops<bls12_377_G1, bls12_377_Fr> O2;
typedef ops<bw6_761_G1, bw6_761_Fr> O3;
typedef ops<bls12_377_G2, bls12_377_Fr> O4;
typedef ops<bw6_761_G2, bw6_761_Fr> O5;
typedef ops<bls12_381_G1, bls12_381_Fr> O6;
typedef ops<bls12_381_G2, bls12_381_Fr> O7;

// op: 0 mul, 1 sqr, 2 add, 3 sub, 4 neg, 5 inverse, 6 as_bigint (mont->plain),
//     7 from bigint (plain->mont)
template<typename F> void field_op(int op, const void *a, const void *b, void *out, size_t n)
{
    std::vector<F> va = load_vec<F>(a, n);
    std::vector<F> vb = b ? load_vec<F>(b, n) : std::vector<F>();
    std::vector<F> r(n);
    for (size_t i = 0; i < n; ++i) {
        switch (op) {
        case 0: r[i] = va[i] * vb[i]; break;
        case 1: r[i] = va[i].squared(); break;
        case 2: r[i] = va[i] + vb[i]; break;
        case 3: r[i] = va[i] - vb[i]; break;
        case 4: r[i] = -va[i]; break;
        case 5: r[i] = va[i].is_zero() ? va[i] : va[i].inverse(); break;
        default: break;
        }
    }
    store_vec(out, r);
}

template<typename F> void field_convert(int op, const void *a, void *out, size_t n)
{
    // Prime fields only: op 6 Montgomery -> plain limbs, op 7 plain -> Montgomery.
    const size_t w = sizeof(F);
    for (size_t i = 0; i < n; ++i) {
        if (op == 6) {
            F x;
            std::memcpy((void *)&x, (const char *)a + i * w, w);
            auto bi = x.as_bigint();
            std::memcpy((char *)out + i * w, (const void *)bi.data, w);
        } else {
            bigint<F::num_limbs> bi;
            std::memcpy((void *)bi.data, (const char *)a + i * w, w);
            F x(bi);
            std::memcpy((char *)out + i * w, (const void *)&x, w);
        }
    }
}

template<typename F> long signed_digit(const void *scalar_mont, size_t c, size_t idx)
{
    F x;
    std::memcpy((void *)&x, scalar_mont, sizeof(F));
    return (long)field_get_signed_digit(x.as_bigint(), c, idx);
}

bool g_inited = false;

} // namespace

extern "C" {

int ref_init()
{
    if (!g_inited) {
        inhibit_profiling_info = true;
        inhibit_profiling_counters = true;
        alt_bn128_pp::init_public_params();
        bls12_377_pp::init_public_params();
        bw6_761_pp::init_public_params();
        bls12_381_pp::init_public_params();
        g_inited = true;
    }
    return 1;
}

int ref_num_threads()
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void ref_set_num_threads(int t)
{
#ifdef _OPENMP
    omp_set_num_threads(t);
#else
    (void)t;
#endif
}

size_t ref_group_bytes(int gid)
{
    switch (gid) {
    case 0: return sizeof(alt_bn128_G1);
    case 1: return sizeof(alt_bn128_G2);
    case 2: return sizeof(bls12_377_G1);
    case 3: return sizeof(bw6_761_G1);
    case 4: return sizeof(bls12_377_G2);
    case 5: return sizeof(bw6_761_G2);
    case 6: return sizeof(bls12_381_G1);
    case 7: return sizeof(bls12_381_G2);
    }
    return 0;
}

size_t ref_scalar_bytes(int gid)
{
    switch (gid) {
    case 0: return sizeof(alt_bn128_Fr);
    case 1: return sizeof(alt_bn128_Fr);
    case 2: return sizeof(bls12_377_Fr);
    case 3: return sizeof(bw6_761_Fr);
    case 4: return sizeof(bls12_377_Fr);
    case 5: return sizeof(bw6_761_Fr);
    case 6: return sizeof(bls12_381_Fr);
    case 7: return sizeof(bls12_381_Fr);
    }
    return 0;
}

size_t ref_scalar_bits(int gid)
{
    switch (gid) {
    case 0: return alt_bn128_Fr::size_in_bits();
    case 1: return alt_bn128_Fr::size_in_bits();
    case 2: return bls12_377_Fr::size_in_bits();
    case 3: return bw6_761_Fr::size_in_bits();
    case 4: return bls12_377_Fr::size_in_bits();
    case 5: return bw6_761_Fr::size_in_bits();
    case 6: return bls12_381_Fr::size_in_bits();
    case 7: return bls12_381_Fr::size_in_bits();
    }
    return 0;
}

#define DISPATCH(gid, CALL)                                                                        \
    switch (gid) {                                                                                 \
    case 0: return O0::CALL;                                                                       \
    case 1: return O1::CALL;                                                                       \
    case 2: return O2::CALL;                                                                       \
    case 3: return O3::CALL;                                                                       \
    case 4: return O4::CALL;                                                                       \
    case 5: return O5::CALL;                                                                       \
    case 6: return O6::CALL;                                                                       \
    case 7: return O7::CALL;                                                                       \
    }

void ref_gen_scalars(int gid, uint64_t start, size_t n, void *out)
{
    DISPATCH(gid, gen_scalars(start, n, out));
}

void ref_gen_bases_walk(int gid, size_t start, size_t n, void *out)
{
    DISPATCH(gid, gen_bases_walk(start, n, out));
}

void ref_scalar_mul(int gid, const void *base, const void *scalar, void *out, int normalize)
{
    DISPATCH(gid, scalar_mul(base, scalar, out, normalize));
}

int ref_multi_exp(int gid, int method, int form, const void *bases, const void *scalars, size_t n,
                  size_t chunks, void *out, int normalize)
{
    DISPATCH(gid, multi_exp_(method, form, bases, scalars, n, chunks, out, normalize));
    return 0;
}

double ref_multi_exp_timed(int gid, int form, const void *bases, const void *scalars, size_t n,
                           size_t chunks, void *out)
{
    DISPATCH(gid, multi_exp_timed(form, bases, scalars, n, chunks, out));
    return -1.0;
}

void ref_batch_to_special(int gid, void *inout, size_t n)
{
    DISPATCH(gid, batch_to_special_(inout, n));
}

size_t ref_window_table_rows(int gid, size_t scalar_bits, size_t window)
{
    DISPATCH(gid, window_table_rows(scalar_bits, window));
    return 0;
}

void ref_window_table(int gid, size_t scalar_bits, size_t window, const void *g, void *out)
{
    DISPATCH(gid, window_table_(scalar_bits, window, g, out));
}

double ref_batch_exp(int gid, size_t scalar_bits, size_t window, const void *table, size_t rows,
                     size_t row_len, const void *scalars, size_t n, void *out, int normalize)
{
    DISPATCH(gid,
             batch_exp_(scalar_bits, window, table, rows, row_len, scalars, n, out, normalize));
    return -1.0;
}

size_t ref_exp_window_size(int gid, size_t n)
{
    switch (gid) {
    case 0: return get_exp_window_size<alt_bn128_G1>(n);
    case 1: return get_exp_window_size<alt_bn128_G2>(n);
    case 2: return get_exp_window_size<bls12_377_G1>(n);
    case 3: return get_exp_window_size<bw6_761_G1>(n);
    case 4: return get_exp_window_size<bls12_377_G2>(n);
    case 5: return get_exp_window_size<bw6_761_G2>(n);
    case 6: return get_exp_window_size<bls12_381_G1>(n);
    case 7: return get_exp_window_size<bls12_381_G2>(n);
    }
    return 0;
}

void ref_group_op(int gid, int op, const void *a, const void *b, void *out, size_t n)
{
    DISPATCH(gid, group_op(op, a, b, out, n));
}

void ref_group_eq(int gid, const void *a, const void *b, uint8_t *out, size_t n)
{
    DISPATCH(gid, group_eq(a, b, out, n));
}

void ref_group_constants(int gid, void *zero, void *one)
{
    DISPATCH(gid, constants(zero, one));
}

size_t ref_field_bytes(int fid)
{
    switch (fid) {
    case F_ALT_BN128_FQ: return sizeof(alt_bn128_Fq);
    case F_ALT_BN128_FR: return sizeof(alt_bn128_Fr);
    case F_BLS12_377_FQ: return sizeof(bls12_377_Fq);
    case F_BLS12_377_FR: return sizeof(bls12_377_Fr);
    case F_BW6_761_FQ: return sizeof(bw6_761_Fq);
    case F_BW6_761_FR: return sizeof(bw6_761_Fr);
    case F_ALT_BN128_FQ2: return sizeof(alt_bn128_Fq2);
    case F_BLS12_377_FQ2: return sizeof(bls12_377_Fq2);
    case F_BLS12_381_FQ: return sizeof(bls12_381_Fq);
    case F_BLS12_381_FR: return sizeof(bls12_381_Fr);
    case F_BLS12_381_FQ2: return sizeof(bls12_381_Fq2);
    }
    return 0;
}

void ref_field_op(int fid, int op, const void *a, const void *b, void *out, size_t n)
{
    if (op >= 6) {
        switch (fid) {
        case F_ALT_BN128_FQ: return field_convert<alt_bn128_Fq>(op, a, out, n);
        case F_ALT_BN128_FR: return field_convert<alt_bn128_Fr>(op, a, out, n);
        case F_BLS12_377_FQ: return field_convert<bls12_377_Fq>(op, a, out, n);
        case F_BLS12_377_FR: return field_convert<bls12_377_Fr>(op, a, out, n);
        case F_BW6_761_FQ: return field_convert<bw6_761_Fq>(op, a, out, n);
        case F_BW6_761_FR: return field_convert<bw6_761_Fr>(op, a, out, n);
        case F_BLS12_381_FQ: return field_convert<bls12_381_Fq>(op, a, out, n);
        case F_BLS12_381_FR: return field_convert<bls12_381_Fr>(op, a, out, n);
        }
        return;
    }
    switch (fid) {
    case F_ALT_BN128_FQ: return field_op<alt_bn128_Fq>(op, a, b, out, n);
    case F_ALT_BN128_FR: return field_op<alt_bn128_Fr>(op, a, b, out, n);
    case F_BLS12_377_FQ: return field_op<bls12_377_Fq>(op, a, b, out, n);
    case F_BLS12_377_FR: return field_op<bls12_377_Fr>(op, a, b, out, n);
    case F_BW6_761_FQ: return field_op<bw6_761_Fq>(op, a, b, out, n);
    case F_BW6_761_FR: return field_op<bw6_761_Fr>(op, a, b, out, n);
    case F_ALT_BN128_FQ2: return field_op<alt_bn128_Fq2>(op, a, b, out, n);
    case F_BLS12_377_FQ2: return field_op<bls12_377_Fq2>(op, a, b, out, n);
    case F_BLS12_381_FQ: return field_op<bls12_381_Fq>(op, a, b, out, n);
    case F_BLS12_381_FR: return field_op<bls12_381_Fr>(op, a, b, out, n);
    case F_BLS12_381_FQ2: return field_op<bls12_381_Fq2>(op, a, b, out, n);
    }
}

long ref_signed_digit(int fid, const void *scalar_mont, size_t c, size_t idx)
{
    switch (fid) {
    case F_ALT_BN128_FR: return signed_digit<alt_bn128_Fr>(scalar_mont, c, idx);
    case F_BLS12_377_FR: return signed_digit<bls12_377_Fr>(scalar_mont, c, idx);
    case F_BW6_761_FR: return signed_digit<bw6_761_Fr>(scalar_mont, c, idx);
    case F_BLS12_381_FR: return signed_digit<bls12_381_Fr>(scalar_mont, c, idx);
    }
    return 0;
}

size_t ref_num_signed_digits(int fid, size_t c)
{
    switch (fid) {
    case F_ALT_BN128_FR: return field_get_num_signed_digits<alt_bn128_Fr>(c);
    case F_BLS12_377_FR: return field_get_num_signed_digits<bls12_377_Fr>(c);
    case F_BW6_761_FR: return field_get_num_signed_digits<bw6_761_Fr>(c);
    case F_BLS12_381_FR: return field_get_num_signed_digits<bls12_381_Fr>(c);
    }
    return 0;
}

size_t ref_pippenger_optimal_c(size_t n) { return internal::pippenger_optimal_c(n); }

} // extern "C"

// ---- wire format of the reference's ffi layer (ffi/ffi_serialization.tcc:98-187) ----
// n elements; returns the number of elements the reference's reader/writer accepted.
template<typename G> static size_t wire_group_write(const void *in, void *out, size_t n)
{
    const size_t wb = 2 * sizeof(((G *)nullptr)->X);
    size_t ok = 0;
    for (size_t i = 0; i < n; ++i) {
        G g;
        std::memcpy((void *)&g, (const char *)in + i * sizeof(G), sizeof(G));
        ok += ffi::group_element_write(g, (char *)out + i * wb, wb) ? 1 : 0;
    }
    return ok;
}
template<typename G> static size_t wire_group_read(const void *in, void *out, uint8_t *okv, size_t n)
{
    const size_t wb = 2 * sizeof(((G *)nullptr)->X);
    size_t ok = 0;
    for (size_t i = 0; i < n; ++i) {
        G g;
        const bool r = ffi::group_element_read(g, (const char *)in + i * wb, wb);
        okv[i] = r;
        ok += r ? 1 : 0;
        std::memcpy((char *)out + i * sizeof(G), (const void *)&g, sizeof(G));
    }
    return ok;
}
template<typename F> static size_t wire_field_write(const void *in, void *out, size_t n)
{
    size_t ok = 0;
    for (size_t i = 0; i < n; ++i) {
        F f;
        std::memcpy((void *)&f, (const char *)in + i * sizeof(F), sizeof(F));
        ok += ffi::field_element_write(f, (char *)out + i * sizeof(F), sizeof(F)) ? 1 : 0;
    }
    return ok;
}
template<typename F> static size_t wire_field_read(const void *in, void *out, uint8_t *okv, size_t n)
{
    size_t ok = 0;
    for (size_t i = 0; i < n; ++i) {
        F f = F::zero();
        const bool r = ffi::field_element_read(f, (const char *)in + i * sizeof(F), sizeof(F));
        okv[i] = r;
        ok += r ? 1 : 0;
        std::memcpy((char *)out + i * sizeof(F), (const void *)&f, sizeof(F));
    }
    return ok;
}

extern "C" {
size_t ref_wire_group_write(int gid, const void *in, void *out, size_t n)
{
    switch (gid) {
    case 0: return wire_group_write<alt_bn128_G1>(in, out, n);
    case 1: return wire_group_write<alt_bn128_G2>(in, out, n);
    case 2: return wire_group_write<bls12_377_G1>(in, out, n);
    case 3: return wire_group_write<bw6_761_G1>(in, out, n);
    case 4: return wire_group_write<bls12_377_G2>(in, out, n);
    case 5: return wire_group_write<bw6_761_G2>(in, out, n);
    case 6: return wire_group_write<bls12_381_G1>(in, out, n);
    case 7: return wire_group_write<bls12_381_G2>(in, out, n);
    }
    return 0;
}
size_t ref_wire_group_read(int gid, const void *in, void *out, uint8_t *okv, size_t n)
{
    switch (gid) {
    case 0: return wire_group_read<alt_bn128_G1>(in, out, okv, n);
    case 1: return wire_group_read<alt_bn128_G2>(in, out, okv, n);
    case 2: return wire_group_read<bls12_377_G1>(in, out, okv, n);
    case 3: return wire_group_read<bw6_761_G1>(in, out, okv, n);
    case 4: return wire_group_read<bls12_377_G2>(in, out, okv, n);
    case 5: return wire_group_read<bw6_761_G2>(in, out, okv, n);
    case 6: return wire_group_read<bls12_381_G1>(in, out, okv, n);
    case 7: return wire_group_read<bls12_381_G2>(in, out, okv, n);
    }
    return 0;
}
size_t ref_wire_scalar_write(int gid, const void *in, void *out, size_t n)
{
    switch (gid) {
    case 0: case 1: return wire_field_write<alt_bn128_Fr>(in, out, n);
    case 2: return wire_field_write<bls12_377_Fr>(in, out, n);
    case 3: return wire_field_write<bw6_761_Fr>(in, out, n);
    case 4: return wire_field_write<bls12_377_Fr>(in, out, n);
    case 5: return wire_field_write<bw6_761_Fr>(in, out, n);
    case 6: case 7: return wire_field_write<bls12_381_Fr>(in, out, n);
    }
    return 0;
}
size_t ref_wire_scalar_read(int gid, const void *in, void *out, uint8_t *okv, size_t n)
{
    switch (gid) {
    case 0: case 1: return wire_field_read<alt_bn128_Fr>(in, out, okv, n);
    case 2: return wire_field_read<bls12_377_Fr>(in, out, okv, n);
    case 3: return wire_field_read<bw6_761_Fr>(in, out, okv, n);
    case 4: return wire_field_read<bls12_377_Fr>(in, out, okv, n);
    case 5: return wire_field_read<bw6_761_Fr>(in, out, okv, n);
    case 6: case 7: return wire_field_read<bls12_381_Fr>(in, out, okv, n);
    }
    return 0;
}
}

// ---- streamed / precomputed MSM (multiexp_stream.tcc:164-223) and the binary stream encoding it
// reads (curve_serialization.tcc:79-101), uncompressed, form 0 = plain, 1 = montgomery ----
template<typename G, form_t Form> static size_t stream_write_t(const void *in, size_t n, size_t c, void *out)
{
    using Fr = typename G::scalar_field;
    std::ostringstream os(std::ios_base::out | std::ios_base::binary);
    const size_t per = c ? (Fr::num_bits + c - 1) / c : 1;
    for (size_t i = 0; i < n; ++i) {
        G el;
        std::memcpy((void *)&el, (const char *)in + i * sizeof(G), sizeof(G));
        // profile_multiexp.cpp:137-147: the element, then each further multiple by 2^c
        group_write<encoding_binary, Form, compression_off>(el, os);
        for (size_t k = 1; k < per; ++k) {
            for (size_t j = 0; j < c; ++j) {
                el = el.dbl();
            }
            group_write<encoding_binary, Form, compression_off>(el, os);
        }
    }
    const std::string b = os.str();
    if (out) {
        std::memcpy(out, b.data(), b.size());
    }
    return b.size();
}
// multiexp_accumulate_buckets requires a non-empty bucket ("Caller must ensure that at least one
// bucket is not empty", multiexp.tcc:88-89, an assert that NDEBUG removes); the stream functions
// call it for EVERY digit position (multiexp_stream.tcc:96-118), also the top ones that hold
// nothing but a possible carry, and for the single bucket set of the precompute variant.  Where
// that precondition fails the reference reads bucket_hit[-1...]: undefined, no value to compare to.
template<typename Fr> static bool stream_msm_defined(const std::vector<Fr> &ex, size_t c, size_t num_digits, bool per_round)
{
    std::vector<ssize_t> digits(num_digits);
    std::vector<bool> hit(num_digits, false);
    bool any = false;
    for (const Fr &e : ex) {
        field_get_signed_digits(digits, e, c, num_digits);
        for (size_t k = 0; k < num_digits; ++k) {
            if (digits[k] != 0) {
                hit[k] = true;
                any = true;
            }
        }
    }
    if (!per_round) {
        return any;
    }
    for (size_t k = 0; k < num_digits; ++k) {
        if (!hit[k]) {
            return false;
        }
    }
    return true;
}

template<typename G, form_t Form>
static int stream_msm_t(const void *bytes, size_t nbytes, const void *scalars, size_t n, size_t c,
                        void *out, int normalize)
{
    using Fr = typename G::scalar_field;
    std::istringstream is(std::string((const char *)bytes, nbytes), std::ios_base::in | std::ios_base::binary);
    const std::vector<Fr> ex = load_vec<Fr>(scalars, n);
    // multi_exp_stream / multi_exp_stream_with_precompute (multiexp_stream.tcc:164-223) run their
    // reader in a second thread over an SPSC FIFO whose counters are relaxed atomics: the element
    // data is not published to the consumer and the call crashes now and then (seen here on
    // bw6_761).  A checker must be deterministic, so the SAME reference reader and the SAME
    // reference consumer are called one after the other on a FIFO that holds every element;
    // set REF_STREAM_THREADED=1 to go through the two-thread entry points instead.
    G r;
    const bool threaded = std::getenv("REF_STREAM_THREADED") != nullptr;
    if (c ? !stream_msm_defined(ex, c, (Fr::num_bits + c - 1) / c, false)
          : !stream_msm_defined(
                ex, bdlo12_signed_optimal_c(n), (Fr::num_bits + 2 + bdlo12_signed_optimal_c(n) - 1) / bdlo12_signed_optimal_c(n), true)) {
        return 0;
    }
    if (threaded) {
        r = c ? multi_exp_stream_with_precompute<Form, compression_off, G, Fr>(is, ex, c)
              : multi_exp_stream<Form, compression_off, G, Fr>(is, ex);
    } else if (c) {
        const size_t num_digits = (Fr::num_bits + c - 1) / c;  // multiexp_stream.tcc:205
        concurrent_buffer_fifo_spsc<G> fifo(std::max<size_t>(n, 1), num_digits);
        element_buffers_from_stream_producer<Form, compression_off, G>(is, fifo, n, num_digits);
        r = multi_exp_precompute_from_fifo<G, Fr>(fifo, ex, c, num_digits);
    } else {
        const size_t cc = bdlo12_signed_optimal_c(n);  // multiexp_stream.tcc:170
        concurrent_fifo_spsc<G> fifo(std::max<size_t>(n, 1));
        elements_from_stream_producer<Form, compression_off, G>(is, fifo, n);
        r = multi_exp_base_elements_from_fifo_all_rounds<G, Fr>(fifo, ex, cc);
    }
    if (normalize) {
        r.to_affine_coordinates();
    }
    std::memcpy(out, (const void *)&r, sizeof(G));
    return 1;
}
template<typename G> static size_t stream_write_g(int form, const void *in, size_t n, size_t c, void *out)
{
    return form ? stream_write_t<G, form_montgomery>(in, n, c, out) : stream_write_t<G, form_plain>(in, n, c, out);
}
template<typename G>
static int stream_msm_g(int form, const void *bytes, size_t nbytes, const void *scalars, size_t n, size_t c,
                        void *out, int normalize)
{
    return form ? stream_msm_t<G, form_montgomery>(bytes, nbytes, scalars, n, c, out, normalize)
                : stream_msm_t<G, form_plain>(bytes, nbytes, scalars, n, c, out, normalize);
}

extern "C" {
// bytes written (out may be NULL to size the buffer); c != 0: every base followed by its multiples
size_t ref_stream_write(int gid, int form, const void *in, size_t n, size_t c, void *out)
{
    switch (gid) {
    case 0: return stream_write_g<alt_bn128_G1>(form, in, n, c, out);
    case 1: return stream_write_g<alt_bn128_G2>(form, in, n, c, out);
    case 2: return stream_write_g<bls12_377_G1>(form, in, n, c, out);
    case 3: return stream_write_g<bw6_761_G1>(form, in, n, c, out);
    case 4: return stream_write_g<bls12_377_G2>(form, in, n, c, out);
    case 5: return stream_write_g<bw6_761_G2>(form, in, n, c, out);
    case 6: return stream_write_g<bls12_381_G1>(form, in, n, c, out);
    case 7: return stream_write_g<bls12_381_G2>(form, in, n, c, out);
    }
    return 0;
}
// multi_exp_stream (c == 0) / multi_exp_stream_with_precompute (c != 0) on an in-memory stream;
// 1 = computed, 0 = the reference's behaviour is undefined for this input (see stream_msm_defined),
// -1 = unknown group
int ref_multi_exp_stream(int gid, int form, const void *bytes, size_t nbytes, const void *scalars, size_t n,
                         size_t c, void *out, int normalize)
{
    switch (gid) {
    case 0: return stream_msm_g<alt_bn128_G1>(form, bytes, nbytes, scalars, n, c, out, normalize);
    case 1: return stream_msm_g<alt_bn128_G2>(form, bytes, nbytes, scalars, n, c, out, normalize);
    case 2: return stream_msm_g<bls12_377_G1>(form, bytes, nbytes, scalars, n, c, out, normalize);
    case 3: return stream_msm_g<bw6_761_G1>(form, bytes, nbytes, scalars, n, c, out, normalize);
    case 4: return stream_msm_g<bls12_377_G2>(form, bytes, nbytes, scalars, n, c, out, normalize);
    case 5: return stream_msm_g<bw6_761_G2>(form, bytes, nbytes, scalars, n, c, out, normalize);
    case 6: return stream_msm_g<bls12_381_G1>(form, bytes, nbytes, scalars, n, c, out, normalize);
    case 7: return stream_msm_g<bls12_381_G2>(form, bytes, nbytes, scalars, n, c, out, normalize);
    }
    return -1;
}
size_t ref_bdlo12_signed_optimal_c(size_t n) { return bdlo12_signed_optimal_c(n); }
}
