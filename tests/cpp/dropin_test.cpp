// Drop-in test: UNMODIFIED libff headers + include/libff_b200.hpp, call sites written exactly as
// in the reference's own tests (libff/algebra/scalar_multiplication/tests/test_multiexp.cpp:112-256,
// 344-390).  With the header included, libff::multi_exp<G, Fr, BDLO12[_signed], Form>,
// multi_exp_implementation<...>::multi_exp_inner, batch_exp and batch_to_special run on the GPU
// through the b200msm C ABI; the naive methods stay on the reference's CPU code and serve as the
// expected values.  Built by `make -C oracle dropin` into oracle/_ref/ (it links reference objects,
// so the binary is not tracked); run on the GPU box by tests/test_dropin_gpu.py.
#include <libff/algebra/curves/alt_bn128/alt_bn128_pp.hpp>
#include <libff/algebra/curves/bls12_377/bls12_377_pp.hpp>
#include <libff/algebra/curves/bls12_381/bls12_381_pp.hpp>
#include <libff/algebra/curves/bw6_761/bw6_761_pp.hpp>
#include <libff/algebra/scalar_multiplication/multiexp.hpp>
#include <libff/algebra/scalar_multiplication/multiexp_stream.hpp>
#include <libff/common/rng.hpp>
#include <sstream>

#include "libff_b200.hpp"  // <- the only line a libff user adds

#include <cstdio>
#include <cstring>

using namespace libff;

static int g_checks = 0, g_failures = 0;
#define CHECK(cond)                                                          \
    do {                                                                     \
        ++g_checks;                                                          \
        if (!(cond)) {                                                       \
            ++g_failures;                                                    \
            printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond);           \
        }                                                                    \
    } while (0)

template<typename GroupT> bool same_affine(GroupT a, GroupT b)
{
    a.to_affine_coordinates();
    b.to_affine_coordinates();
    return std::memcmp((const void *)&a, (const void *)&b, sizeof(GroupT)) == 0;
}

// test_multiexp.cpp:112-158
template<typename GroupT, multi_exp_method Method, multi_exp_base_form BaseForm>
void test_multiexp_inner()
{
    using FieldT = typename GroupT::scalar_field;
    const size_t c = 4;
    std::vector<GroupT> bases{
        FieldT(1) * GroupT::one(), FieldT(2) * GroupT::one(), FieldT(3) * GroupT::one(),
        FieldT(4) * GroupT::one(), FieldT(1) * GroupT::one(), FieldT(2) * GroupT::one(),
        FieldT(3) * GroupT::one(), FieldT(4) * GroupT::one(),
    };
    for (auto &b : bases) {
        b.to_affine_coordinates();
    }
    std::vector<FieldT> exponents{
        FieldT(4 << (c * 0)), FieldT(3 << (c * 1)), FieldT(2 << (c * 2)), FieldT(1 << (c * 3)),
        FieldT(4 << (c * 0)), FieldT(3 << (c * 1)), FieldT(2 << (c * 2)), FieldT(1 << (c * 3)),
    };
    const long expected_unencoded =
        2 * ((4 << 0) + ((3 << c) * 2) + ((2 << (2 * c)) * 3) + ((1 << (3 * c)) * 4));
    const GroupT expected = FieldT(expected_unencoded) * GroupT::one();
    const GroupT actual =
        internal::multi_exp_implementation<GroupT, FieldT, Method, BaseForm>::multi_exp_inner(
            bases.begin(), bases.end(), exponents.begin(), exponents.end());
    CHECK(expected == actual);
    CHECK(same_affine(expected, actual));
}

// test_multiexp.cpp:205-256
template<typename GroupT, multi_exp_method Method, multi_exp_base_form BaseForm>
void test_multi_exp_config(size_t num_elements)
{
    using Field = typename GroupT::scalar_field;
    std::vector<GroupT> base_elements;
    std::vector<Field> scalars;
    GroupT g = GroupT::one();
    size_t expect_scalar = 0;
    for (size_t i = 0; i < num_elements; ++i) {
        base_elements.push_back(g);
        g = g + GroupT::one();
        scalars.push_back(num_elements - i);
        expect_scalar += (i + 1) * (num_elements - i);
    }
    const GroupT expect = Field(expect_scalar) * GroupT::one();
    if (BaseForm == multi_exp_base_form_special) {
        batch_to_special(base_elements);
    }
    for (size_t chunks : {1, 2, 4}) {
        const GroupT result = multi_exp<GroupT, Field, Method, BaseForm>(
            base_elements.begin(), base_elements.end(), scalars.begin(), scalars.end(), chunks);
        CHECK(expect == result);
        CHECK(same_affine(expect, result));
    }
}

template<typename GroupT, multi_exp_method Method> void test_multi_exp_group_method()
{
    for (size_t n : {1, 2, 3, 4, 5, 256, 257}) {
        test_multi_exp_config<GroupT, Method, multi_exp_base_form_normal>(n);
        test_multi_exp_config<GroupT, Method, multi_exp_base_form_special>(n);
    }
    test_multiexp_inner<GroupT, Method, multi_exp_base_form_normal>();
    test_multiexp_inner<GroupT, Method, multi_exp_base_form_special>();
}

// random full-width scalars (SHA512_rng, rng.tcc:26-71) against the reference's CPU methods
template<typename GroupT> void test_random(size_t n)
{
    using Field = typename GroupT::scalar_field;
    std::vector<GroupT> bases;
    std::vector<Field> scalars;
    GroupT g = SHA512_rng<Field>(77) * GroupT::one();
    const GroupT d = SHA512_rng<Field>(78) * GroupT::one();
    for (size_t i = 0; i < n; ++i) {
        bases.push_back(g);
        g = g + d;
        scalars.push_back(SHA512_rng<Field>(i));
    }
    // CPU (reference code, not specialised): naive_plain
    const GroupT expect = multi_exp<GroupT, Field, multi_exp_method_naive_plain>(
        bases.begin(), bases.end(), scalars.begin(), scalars.end(), 1);
    // GPU, Jacobian bases
    const GroupT r_normal = multi_exp<GroupT, Field, multi_exp_method_BDLO12>(
        bases.begin(), bases.end(), scalars.begin(), scalars.end(), 1);
    CHECK(expect == r_normal);
    // GPU batch_to_special against the reference's scalar to_special()
    std::vector<GroupT> special(bases);
    batch_to_special(special);
    for (size_t i = 0; i < n; ++i) {
        GroupT t = bases[i];
        t.to_special();
        CHECK(std::memcmp((const void *)&t, (const void *)&special[i], sizeof(GroupT)) == 0);
    }
    const GroupT r_special =
        multi_exp<GroupT, Field, multi_exp_method_BDLO12_signed, multi_exp_base_form_special>(
            special.begin(), special.end(), scalars.begin(), scalars.end(), 8);
    CHECK(expect == r_special);
    CHECK(same_affine(expect, r_special));
    // the wrapper that production callers use (multiexp.tcc:690-757) picks up the specialisation
    scalars[0] = Field::zero();
    scalars[1] = Field::one();
    inhibit_profiling_info = true;
    const GroupT expect2 = multi_exp<GroupT, Field, multi_exp_method_naive_plain>(
        bases.begin(), bases.end(), scalars.begin(), scalars.end(), 1);
    const GroupT r_filter = multi_exp_filter_one_zero<GroupT, Field, multi_exp_method_BDLO12>(
        bases.begin(), bases.end(), scalars.begin(), scalars.end(), 1);
    CHECK(expect2 == r_filter);
    // empty input (multiexp.tcc:291-309)
    const GroupT r_empty = multi_exp<GroupT, Field, multi_exp_method_BDLO12>(
        bases.begin(), bases.begin(), scalars.begin(), scalars.begin(), 1);
    CHECK(r_empty.is_zero());
}

// batch_exp (multiexp.tcc:875-915) against windowed_exp on the CPU (multiexp.tcc:848-873)
template<typename GroupT> void test_batch_exp(size_t n, size_t window)
{
    using Field = typename GroupT::scalar_field;
    const size_t scalar_size = Field::size_in_bits();
    const window_table<GroupT> table = get_window_table(scalar_size, window, GroupT::one());
    std::vector<Field> v;
    for (size_t i = 0; i < n; ++i) {
        v.push_back(SHA512_rng<Field>(1000 + i));
    }
    v[0] = Field::zero();
    v[1] = -Field::one();
    const std::vector<GroupT> got = batch_exp(scalar_size, window, table, v);
    CHECK(got.size() == n);
    for (size_t i = 0; i < n && i < got.size(); ++i) {
        const GroupT want = windowed_exp(scalar_size, window, table, v[i]);
        CHECK(want == got[i]);
    }
    const std::vector<GroupT> part = batch_exp(scalar_size, window, table, v, n / 2);
    CHECK(part.size() == n / 2);
    // batch_exp_with_coeff (multiexp.tcc:917-947): on the device too (one Fr product ahead of the
    // digits); expected values from the reference's windowed_exp on the CPU
    const std::vector<GroupT> unit = batch_exp_with_coeff(scalar_size, window, table, Field::one(), v);
    const Field c = SHA512_rng<Field>(77);
    const std::vector<GroupT> scaled = batch_exp_with_coeff(scalar_size, window, table, c, v);
    CHECK(unit.size() == n && scaled.size() == n);
    for (size_t i = 0; i < n && i < scaled.size(); ++i) {
        CHECK(unit[i] == got[i]);
        CHECK(windowed_exp(scalar_size, window, table, c * v[i]) == scaled[i]);
    }
    // get_window_table (multiexp.tcc:809-846) is built on the device: every entry j 2^(r w) g,
    // the last row's tail zero -- checked against the group law on the CPU
    const size_t in_window = 1ul << window;
    const size_t outerc = (scalar_size + window - 1) / window;
    const size_t last_in_window = 1ul << (scalar_size - (outerc - 1) * window);
    CHECK(table.size() == outerc);
    GroupT gouter = GroupT::one();
    for (size_t outer = 0; outer < outerc && outer < table.size(); ++outer) {
        CHECK(table[outer].size() == in_window);
        GroupT ginner = GroupT::zero();
        const size_t cur = outer == outerc - 1 ? last_in_window : in_window;
        for (size_t inner = 0; inner < in_window; ++inner) {
            CHECK(table[outer][inner] == (inner < cur ? ginner : GroupT::zero()));
            ginner = ginner + gouter;
        }
        for (size_t i = 0; i < window; ++i) {
            gouter = gouter + gouter;
        }
    }
    // a table the caller changed afterwards is no longer the resident one: it is uploaded
    window_table<GroupT> other = table;
    other[0][1] = other[0][2];
    const std::vector<GroupT> got2 = batch_exp(scalar_size, window, other, v);
    for (size_t i = 0; i < n && i < got2.size(); ++i) {
        CHECK(windowed_exp(scalar_size, window, other, v[i]) == got2[i]);
    }
}

// multi_exp_stream / multi_exp_stream_with_precompute (multiexp_stream.tcc:164-223) on streams
// written by the reference's own group_write, as profile_multiexp.cpp:100-151 does; expected
// value from the CPU naive method
template<typename GroupT, form_t Form, compression_t Comp> void test_stream(size_t n)
{
    using Field = typename GroupT::scalar_field;
    std::vector<GroupT> bases;
    std::vector<Field> scalars;
    GroupT g = SHA512_rng<Field>(91) * GroupT::one();
    const GroupT d = SHA512_rng<Field>(92) * GroupT::one();
    for (size_t i = 0; i < n; ++i) {
        bases.push_back(g);
        g = g + d;
        scalars.push_back(SHA512_rng<Field>(500 + i));
    }
    bases[n / 2] = GroupT::zero();
    const GroupT expect = multi_exp<GroupT, Field, multi_exp_method_naive_plain>(
        bases.begin(), bases.end(), scalars.begin(), scalars.end(), 1);
    {
        std::ostringstream os(std::ios_base::out | std::ios_base::binary);
        for (const GroupT &el : bases) {
            group_write<encoding_binary, Form, Comp>(el, os);
        }
        std::istringstream is(os.str(), std::ios_base::in | std::ios_base::binary);
        const GroupT r = multi_exp_stream<Form, Comp, GroupT, Field>(is, scalars);
        CHECK(expect == r);
    }
    {
        const size_t c = bdlo12_signed_optimal_c(n);
        const size_t per = (Field::num_bits + c - 1) / c;
        std::ostringstream os(std::ios_base::out | std::ios_base::binary);
        for (GroupT el : bases) {
            group_write<encoding_binary, Form, Comp>(el, os);
            for (size_t i = 0; i + 1 < per; ++i) {
                for (size_t j = 0; j < c; ++j) {
                    el = el.dbl();
                }
                group_write<encoding_binary, Form, Comp>(el, os);
            }
        }
        std::istringstream is(os.str(), std::ios_base::in | std::ios_base::binary);
        const GroupT r = multi_exp_stream_with_precompute<Form, Comp, GroupT, Field>(is, scalars, c);
        CHECK(expect == r);
    }
}

// TestEdgeCase1 (test_multiexp.cpp:344-390), literal vectors
void test_edge_case1()
{
    auto fq = [](const char *dec) { return alt_bn128_Fq(bigint<alt_bn128_q_limbs>(dec)); };
    const alt_bn128_G1 p0(
        fq("9718951969731899135137883357393855980400393826120313062226815407726206576592"),
        fq("10028136500182002989802635832456611151885266158376934693401545671954333615444"),
        alt_bn128_Fq::one());
    const alt_bn128_G1 p1(
        fq("17855957492011111566052623837030895280947925668538672721363773163892284645508"),
        fq("8941137791600821643453541768300973235207650866386901989397206694130900646332"),
        alt_bn128_Fq::one());
    CHECK(p0.is_well_formed());
    CHECK(p1.is_well_formed());
    const std::vector<alt_bn128_G1> bases{p0, p1};
    const std::vector<alt_bn128_Fr> scalars{alt_bn128_Fr(1), alt_bn128_Fr(0x79D69)};
    const alt_bn128_G1 expect = p0 + alt_bn128_Fr(0x79D69) * p1;
    const alt_bn128_G1 r1 =
        multi_exp<alt_bn128_G1, alt_bn128_Fr, multi_exp_method_BDLO12_signed>(
            bases.begin(), bases.end(), scalars.begin(), scalars.end(), 1);
    const alt_bn128_G1 r2 =
        multi_exp<alt_bn128_G1, alt_bn128_Fr, multi_exp_method_BDLO12, multi_exp_base_form_special>(
            bases.begin(), bases.end(), scalars.begin(), scalars.end(), 1);
    CHECK(expect == r1);
    CHECK(expect == r2);
}

int main()
{
    inhibit_profiling_info = true;
    inhibit_profiling_counters = true;
    alt_bn128_pp::init_public_params();
    bls12_377_pp::init_public_params();
    bw6_761_pp::init_public_params();
    bls12_381_pp::init_public_params();
    try {
        test_multi_exp_group_method<alt_bn128_G1, multi_exp_method_BDLO12>();
        test_multi_exp_group_method<alt_bn128_G1, multi_exp_method_BDLO12_signed>();
        test_multi_exp_group_method<alt_bn128_G2, multi_exp_method_BDLO12>();
        test_multi_exp_group_method<alt_bn128_G2, multi_exp_method_BDLO12_signed>();
        test_multi_exp_group_method<bls12_377_G1, multi_exp_method_BDLO12>();
        test_multi_exp_group_method<bls12_377_G1, multi_exp_method_BDLO12_signed>();
        test_multi_exp_group_method<bw6_761_G1, multi_exp_method_BDLO12>();
        test_multi_exp_group_method<bw6_761_G1, multi_exp_method_BDLO12_signed>();
        test_random<alt_bn128_G1>(1000);
        test_random<alt_bn128_G2>(300);
        test_random<bls12_377_G1>(500);
        test_random<bw6_761_G1>(200);
        test_batch_exp<alt_bn128_G1>(200, 5);
        test_batch_exp<alt_bn128_G2>(60, 4);
        test_batch_exp<bls12_377_G1>(100, 5);
        test_batch_exp<bw6_761_G1>(40, 4);
        // the remaining groups (test_multiexp.cpp:322-326 runs bls12_381)
        test_multi_exp_group_method<bls12_381_G1, multi_exp_method_BDLO12>();
        test_multi_exp_group_method<bls12_381_G1, multi_exp_method_BDLO12_signed>();
        test_multi_exp_group_method<bls12_381_G2, multi_exp_method_BDLO12>();
        test_multi_exp_group_method<bls12_381_G2, multi_exp_method_BDLO12_signed>();
        test_multi_exp_group_method<bls12_377_G2, multi_exp_method_BDLO12_signed>();
        test_multi_exp_group_method<bw6_761_G2, multi_exp_method_BDLO12_signed>();
        test_random<bls12_381_G1>(500);
        test_random<bls12_381_G2>(200);
        test_random<bls12_377_G2>(200);
        test_random<bw6_761_G2>(100);
        test_batch_exp<bls12_381_G1>(100, 5);
        test_batch_exp<bls12_381_G2>(40, 4);
        test_batch_exp<bls12_377_G2>(40, 4);
        test_batch_exp<bw6_761_G2>(30, 4);
        test_stream<alt_bn128_G1, form_montgomery, compression_off>(300);
        test_stream<alt_bn128_G1, form_plain, compression_off>(100);
        test_stream<alt_bn128_G1, form_plain, compression_on>(100);
        test_stream<alt_bn128_G2, form_montgomery, compression_off>(60);
        test_stream<bls12_377_G1, form_plain, compression_off>(80);
        test_stream<bw6_761_G1, form_montgomery, compression_off>(40);
        test_stream<bls12_381_G1, form_montgomery, compression_on>(60);
        test_stream<bls12_381_G2, form_plain, compression_off>(40);
        test_edge_case1();
    } catch (const std::exception &e) {
        printf("EXCEPTION %s\n", e.what());
        return 2;
    }
    printf("engine contexts: %zu\n", b200msm_device_count());
    printf("dropin_test: %d checks, %d failures, %llu GPU kernel launches\n", g_checks, g_failures,
           (unsigned long long)b200msm_launch_count());
    return g_failures == 0 && b200msm_launch_count() > 0 ? 0 : 1;
}
