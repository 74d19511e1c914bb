/** libff_b200.hpp -- source-level drop-in that routes libff's MSM hot path to the B200 engine.
 *
 * Include this header immediately after
 *     #include <libff/algebra/scalar_multiplication/multiexp.hpp>
 * (and the curve headers) and before any call site, then link libb200msm.so.  It provides explicit
 * (full) specialisations of the reference templates for the groups of the path:
 *
 *   libff::multi_exp<G, Fr, multi_exp_method_BDLO12 | _BDLO12_signed, normal | special>
 *                                                   multiexp.hpp:63-73, multiexp.tcc:643-688
 *   libff::internal::multi_exp_implementation<G, Fr, Method, Form>::multi_exp_inner
 *                                                   multiexp.tcc:198-212, 276-381, 507-633
 *   libff::multi_exp_filter_one_zero<G, Fr, Method, Form>   multiexp.hpp:78-88, multiexp.tcc:690-757
 *   libff::batch_exp<G, Fr>(scalar_size, window, table, v[, num_entries])
 *                                                   multiexp.hpp:120-134, multiexp.tcc:875-915
 *   libff::batch_exp_with_coeff<G, Fr>(scalar_size, window, table, coeff, v)
 *                                                   multiexp.hpp:136-142, multiexp.tcc:917-947
 *   libff::get_window_table<G>(scalar_size, window, g)   multiexp.hpp:107-111, multiexp.tcc:809-846
 *   libff::batch_to_special<G>(vec)                 multiexp.hpp:144, multiexp.tcc:949-974
 *   libff::multi_exp_stream<Form, Comp, G, Fr>(istream, exponents)
 *   libff::multi_exp_stream_with_precompute<Form, Comp, G, Fr>(istream, exponents, c)
 *                                                   multiexp_stream.hpp:19-41, .tcc:164-223
 *
 * for G in { alt_bn128_G1, alt_bn128_G2, bls12_377_G1, bw6_761_G1 } (the path's four groups) and
 * { bls12_377_G2, bw6_761_G2, bls12_381_G1, bls12_381_G2 } (bls12_381 is in the
 * reference's own multiexp tests, test_multiexp.cpp:322-326) with their scalar fields.
 * Call sites stay source-identical (libff's own test, test_multiexp.cpp, compiles unchanged: see
 * tests/cpp/dropin_test.cpp).  Everything forwards to the C ABI of include/b200msm.h with
 * `&*vec.begin()`: no conversion, libff's in-memory layout is the wire format.
 *
 * Semantics kept from the reference: the result is the same group element (operator== true,
 * identical affine coordinates); batch_to_special output is limb-identical; n == 0 returns
 * G::zero().  Differences, all permitted by the reference's contract: the returned (X:Y:Z)
 * representative may differ (get_window_table returns special-form entries); `chunks` (an OpenMP
 * split hint, multiexp.tcc:655-687) is ignored -- the range is split over the bound GPUs instead,
 * by the same rule.  The profiling side effects of batch_exp / batch_exp_with_coeff /
 * batch_to_special (multiexp.tcc:887-913, 923-945, 951, 973) are kept and, as in the reference,
 * silenced by libff::inhibit_profiling_info.  There is no CPU fallback: if the engine cannot run
 * (no CUDA device, library error) a std::runtime_error carrying b200msm_last_error() is thrown.
 *
 * Devices.  On first use the process binds (libff::b200::init / init_devices, or lazily):
 *   $B200MSM_DEVICES = "0,1,2,3"   exactly these device ordinals, else
 *   $B200MSM_DEVICE or $LOCAL_RANK that one device (one process per GPU under a launcher), else
 *   $B200MSM_NUM_GPUS = k          the first k visible devices, else
 *   every visible device.
 * With several devices bound, one multi_exp call runs on all of them: the index range is cut into
 * one contiguous range per device exactly as the reference cuts it into `chunks` for its OpenMP
 * threads (multiexp.tcc:655-678), and the partial sums are added (multiexp.tcc:681-685).
 *
 * Window tables.  get_window_table builds the table on the device, keeps it resident and hands the
 * caller an ordinary window_table<G>; a later batch_exp / batch_exp_with_coeff with that table is
 * served from the resident copy (recognised by its shape and a fingerprint of sampled entries;
 * $B200MSM_TABLE_CACHE=0 turns this off, every call then uploads its table).
 */
#ifndef LIBFF_B200_HPP_
#define LIBFF_B200_HPP_

#include <cstdlib>
#include <libff/algebra/curves/alt_bn128/alt_bn128_pp.hpp>
#include <libff/algebra/curves/bls12_377/bls12_377_pp.hpp>
#include <libff/algebra/curves/bls12_381/bls12_381_pp.hpp>
#include <libff/algebra/curves/bw6_761/bw6_761_pp.hpp>
#include <libff/algebra/scalar_multiplication/multiexp.hpp>
#include <libff/algebra/scalar_multiplication/multiexp_stream.hpp>
#include <libff/common/profiling.hpp>
#include <cstdio>
#include <cstring>
#include <exception>
#include <istream>
#include <list>
#include <mutex>
#include <thread>
#include <stdexcept>
#include <string>
#include <vector>

#include "b200msm.h"

namespace libff
{
namespace b200
{

template<typename GroupT> struct group_id;
template<> struct group_id<alt_bn128_G1> {
    static constexpr uint32_t value = B200MSM_ALT_BN128_G1;
};
template<> struct group_id<alt_bn128_G2> {
    static constexpr uint32_t value = B200MSM_ALT_BN128_G2;
};
template<> struct group_id<bls12_377_G1> {
    static constexpr uint32_t value = B200MSM_BLS12_377_G1;
};
template<> struct group_id<bw6_761_G1> {
    static constexpr uint32_t value = B200MSM_BW6_761_G1;
};
template<> struct group_id<bls12_377_G2> {
    static constexpr uint32_t value = B200MSM_BLS12_377_G2;
};
template<> struct group_id<bw6_761_G2> {
    static constexpr uint32_t value = B200MSM_BW6_761_G2;
};
template<> struct group_id<bls12_381_G1> {
    static constexpr uint32_t value = B200MSM_BLS12_381_G1;
};
template<> struct group_id<bls12_381_G2> {
    static constexpr uint32_t value = B200MSM_BLS12_381_G2;
};

[[noreturn]] inline void fail(const char *what)
{
    throw std::runtime_error(
        std::string(what) + ": " + b200msm_last_error() + " (b200msm has no CPU fallback)");
}

/// Binds the process to a CUDA device (b200msm_init). Safe to call repeatedly.
inline void init(int device)
{
    if (!b200msm_init(device)) {
        fail("b200msm_init");
    }
}

/// Binds the process to several GPUs of the box; a call then spreads over all of them.
inline void init_devices(const std::vector<int> &devices)
{
    if (!b200msm_init_devices(devices.data(), devices.size())) {
        fail("b200msm_init_devices");
    }
}

inline void ensure_init()
{
    static std::once_flag once;
    std::call_once(once, [] {
        if (b200msm_device_count() > 0) {
            return; // the application bound its devices itself
        }
        if (const char *list = std::getenv("B200MSM_DEVICES")) {
            std::vector<int> ids;
            for (const char *p = list; *p;) {
                char *end = nullptr;
                const long v = std::strtol(p, &end, 10);
                if (end == p) {
                    break;
                }
                ids.push_back((int)v);
                p = (*end == ',') ? end + 1 : end;
            }
            init_devices(ids);
            return;
        }
        const char *d = std::getenv("B200MSM_DEVICE");
        if (!d) {
            d = std::getenv("LOCAL_RANK");
        }
        if (d) {
            init(std::atoi(d));
            return;
        }
        if (const char *k = std::getenv("B200MSM_NUM_GPUS")) {
            std::vector<int> ids;
            for (int i = 0; i < std::atoi(k); ++i) {
                ids.push_back(i);
            }
            if (!ids.empty()) {
                init_devices(ids);
                return;
            }
        }
        if (b200msm_init_all() <= 0) {
            fail("b200msm_init_all");
        }
    });
}

/// The layout contract of include/b200msm.h, checked at compile time against libff's types.
template<typename GroupT, typename FieldT> struct layout_check {
    static_assert(
        sizeof(FieldT) == FieldT::num_limbs * sizeof(mp_limb_t),
        "scalar must be exactly its Montgomery limbs");
    static_assert(sizeof(mp_limb_t) == 8, "64-bit limbs expected");
};

template<typename GroupT, typename FieldT>
GroupT multi_exp(
    typename std::vector<GroupT>::const_iterator vec_start,
    typename std::vector<GroupT>::const_iterator vec_end,
    typename std::vector<FieldT>::const_iterator scalar_start,
    typename std::vector<FieldT>::const_iterator scalar_end,
    multi_exp_base_form form)
{
    (void)sizeof(layout_check<GroupT, FieldT>);
    const size_t n = vec_end - vec_start;
    if ((size_t)(scalar_end - scalar_start) != n) {
        throw std::runtime_error("b200 multi_exp: bases and scalars differ in length");
    }
    if (n == 0) {
        // multiexp.tcc:291-309; &*vec_start would not be dereferenceable
        return GroupT::zero();
    }
    ensure_init();
    if (sizeof(GroupT) != b200msm_group_bytes(group_id<GroupT>::value) ||
        sizeof(FieldT) != b200msm_scalar_bytes(group_id<GroupT>::value)) {
        throw std::runtime_error("b200 multi_exp: sizeof(GroupT)/sizeof(FieldT) mismatch");
    }
    GroupT result;
    const bool ok = b200msm_multi_exp(
        group_id<GroupT>::value,
        (const void *)&*vec_start,
        n * sizeof(GroupT),
        (const void *)&*scalar_start,
        n * sizeof(FieldT),
        n,
        form == multi_exp_base_form_special ? B200MSM_FORM_SPECIAL : B200MSM_FORM_NORMAL,
        (void *)&result,
        sizeof(GroupT));
    if (!ok) {
        fail("b200msm_multi_exp");
    }
    return result;
}

// ---- resident window tables -------------------------------------------------------------------
// get_window_table leaves its table on the device; batch_exp recognises it again by shape and by a
// fingerprint over sampled entries (every row's entries 1, 2 and last, plus the table's corners).
struct table_cache_entry {
    uint32_t gid;
    size_t scalar_size, window, rows, row_len;
    uint64_t fingerprint, handle;
};

inline std::list<table_cache_entry> &table_cache()
{
    static std::list<table_cache_entry> cache;
    return cache;
}
inline std::mutex &table_cache_mutex()
{
    static std::mutex m;
    return m;
}
inline bool table_cache_enabled()
{
    const char *e = std::getenv("B200MSM_TABLE_CACHE");
    return !(e && e[0] == '0');
}

template<typename GroupT> uint64_t table_fingerprint(const window_table<GroupT> &table)
{
    uint64_t h = 1469598103934665603ull; // FNV-1a over the sampled entries' bytes
    auto mix = [&h](const GroupT &el) {
        const unsigned char *p = (const unsigned char *)&el;
        for (size_t i = 0; i < sizeof(GroupT); ++i) {
            h = (h ^ p[i]) * 1099511628211ull;
        }
    };
    for (const std::vector<GroupT> &row : table) {
        const size_t len = row.size();
        if (len == 0) {
            continue;
        }
        mix(row[0]);
        mix(row[len > 1 ? 1 : 0]);
        mix(row[len > 2 ? 2 : 0]);
        mix(row[len / 2]);
        mix(row[len - 1]);
    }
    return h;
}

template<typename GroupT>
uint64_t find_resident_table(const size_t scalar_size, const size_t window, const window_table<GroupT> &table)
{
    if (!table_cache_enabled() || table.empty()) {
        return 0;
    }
    const uint64_t fp = table_fingerprint(table);
    std::lock_guard<std::mutex> lk(table_cache_mutex());
    for (const table_cache_entry &e : table_cache()) {
        if (e.gid == group_id<GroupT>::value && e.scalar_size == scalar_size && e.window == window &&
            e.rows == table.size() && e.row_len == table[0].size() && e.fingerprint == fp) {
            return e.handle;
        }
    }
    return 0;
}

/// get_window_table (multiexp.tcc:809-846) built on the device; the table stays resident.
template<typename GroupT>
window_table<GroupT> get_window_table(const size_t scalar_size, const size_t window, const GroupT &g)
{
    ensure_init();
    const uint32_t gid = group_id<GroupT>::value;
    const size_t in_window = 1ul << window;
    const size_t outerc = (scalar_size + window - 1) / window;
    uint64_t handle = 0;
    if (!b200msm_table_build(gid, scalar_size, window, (const void *)&g, sizeof(GroupT), &handle)) {
        fail("b200msm_table_build");
    }
    window_table<GroupT> powers_of_g(outerc, std::vector<GroupT>(in_window, GroupT::zero()));
    std::vector<void *> rows(outerc);
    for (size_t r = 0; r < outerc; ++r) {
        rows[r] = (void *)powers_of_g[r].data();
    }
    if (!b200msm_table_download(handle, rows.data(), outerc, in_window)) {
        (void)b200msm_release(handle);
        fail("b200msm_table_download");
    }
    if (!table_cache_enabled()) {
        (void)b200msm_release(handle);
        return powers_of_g;
    }
    std::lock_guard<std::mutex> lk(table_cache_mutex());
    std::list<table_cache_entry> &cache = table_cache();
    cache.push_front(table_cache_entry{
        gid, scalar_size, window, outerc, in_window, table_fingerprint(powers_of_g), handle});
    while (cache.size() > 4) { // a prover holds a G1 and a G2 table at a time
        (void)b200msm_release(cache.back().handle);
        cache.pop_back();
    }
    return powers_of_g;
}

/// batch_exp (coeff == nullptr, multiexp.tcc:875-915) and batch_exp_with_coeff (:917-947)
template<typename GroupT, typename FieldT>
std::vector<GroupT> batch_exp(
    const size_t scalar_size,
    const size_t window,
    const window_table<GroupT> &table,
    const FieldT *coeff,
    const std::vector<FieldT> &v,
    size_t num_entries)
{
    if (!inhibit_profiling_info) {
        print_indent(); // multiexp.tcc:887-889
    }
    std::vector<GroupT> res(num_entries, table[0][0]); // multiexp.tcc:890
    if (num_entries > v.size()) {
        throw std::runtime_error("b200 batch_exp: num_entries exceeds the scalar vector");
    }
    if (num_entries != 0) {
        ensure_init();
        const uint32_t gid = group_id<GroupT>::value;
        const uint64_t resident = find_resident_table(scalar_size, window, table);
        bool ok;
        if (resident) {
            ok = b200msm_batch_exp_resident(
                gid,
                resident,
                (const void *)v.data(),
                num_entries * sizeof(FieldT),
                num_entries,
                (const void *)coeff,
                coeff ? sizeof(FieldT) : 0,
                (void *)res.data(),
                num_entries * sizeof(GroupT));
        } else {
            std::vector<const void *> rows(table.size());
            for (size_t r = 0; r < table.size(); ++r) {
                if (table[r].size() != table[0].size()) {
                    // get_window_table gives every row 1 << window entries (multiexp.tcc:828-836)
                    throw std::runtime_error("b200 batch_exp: ragged window table");
                }
                rows[r] = (const void *)table[r].data();
            }
            ok = b200msm_batch_exp_with_coeff(
                gid,
                scalar_size,
                window,
                rows.data(),
                rows.size(),
                table[0].size(),
                (const void *)coeff,
                coeff ? sizeof(FieldT) : 0,
                (const void *)v.data(),
                num_entries * sizeof(FieldT),
                num_entries,
                (void *)res.data(),
                num_entries * sizeof(GroupT));
        }
        if (!ok) {
            fail("b200msm_batch_exp");
        }
    }
    if (!inhibit_profiling_info) {
        // the reference prints one dot per 10000 entries and " DONE!" (multiexp.tcc:899-912)
        for (size_t i = 0; i < num_entries; i += 10000) {
            printf(".");
        }
        printf(" DONE!\n");
        fflush(stdout);
    }
    return res;
}

template<typename GroupT> void batch_to_special(std::vector<GroupT> &vec)
{
    enter_block("Batch-convert elements to special form"); // multiexp.tcc:951
    if (!vec.empty()) {
        ensure_init();
        if (!b200msm_batch_to_special(
                group_id<GroupT>::value,
                (void *)vec.data(),
                vec.size() * sizeof(GroupT),
                vec.size())) {
            fail("b200msm_batch_to_special");
        }
    }
    leave_block("Batch-convert elements to special form"); // multiexp.tcc:973
}

/// multi_exp_stream / multi_exp_stream_with_precompute (multiexp_stream.tcc:164-223): the base
/// elements are read from the stream in chunks by a helper thread (the reference's producer
/// thread, :176-180) while the device reduces the previous chunk.  Uncompressed streams are
/// uploaded as bytes and decoded on the device; compressed ones are decoded by libff's own
/// group_read (a square root per element) and pushed as in-memory elements.
/// precompute_c == 0: plain stream.
template<form_t Form, compression_t Comp, typename GroupT, typename FieldT>
GroupT multi_exp_stream(
    std::istream &in_s, const std::vector<FieldT> &exponents, const size_t precompute_c)
{
    (void)sizeof(layout_check<GroupT, FieldT>);
    const size_t n = exponents.size();
    if (n == 0) {
        return GroupT::zero();
    }
    ensure_init();
    const uint32_t gid = group_id<GroupT>::value;
    const size_t per = precompute_c ? b200msm_precompute_digits(gid, (uint32_t)precompute_c) : 1;
    if (per == 0) {
        throw std::runtime_error("b200 multi_exp_stream: unsupported precompute_c");
    }
    const bool raw_bytes = Comp == compression_off;
    const uint32_t encoding =
        !raw_bytes ? B200MSM_STREAM_MEMORY_SPECIAL
                   : (Form == form_montgomery ? B200MSM_STREAM_BINARY_MONTGOMERY
                                              : B200MSM_STREAM_BINARY_PLAIN);
    const size_t elem_bytes = raw_bytes ? b200msm_wire_point_bytes(gid) : sizeof(GroupT);
    uint64_t stream = 0;
    if (!b200msm_stream_begin(
            gid,
            (const void *)exponents.data(),
            n * sizeof(FieldT),
            n,
            encoding,
            (uint32_t)precompute_c,
            &stream)) {
        fail("b200msm_stream_begin");
    }
    // releases the stream (its device buffers) on every way out, exceptions included
    struct stream_guard {
        uint64_t handle;
        ~stream_guard()
        {
            if (handle) {
                GroupT scratch;
                (void)b200msm_stream_finish(handle, (void *)&scratch, sizeof(GroupT));
            }
        }
    } guard{stream};
    // bases per chunk: large enough to amortise the fixed cost of one device MSM
    const size_t chunk = std::max<size_t>(1, std::min<size_t>(n, ((size_t)1 << 20) / per));
    std::vector<char> buf[2];
    buf[0].resize(chunk * per * elem_bytes);
    buf[1].resize(chunk * per * elem_bytes);
    bool read_ok = true;
    std::exception_ptr read_error; // an exception of the reader (it may run on a helper thread)
    auto read_chunk = [&](std::vector<char> &dst, size_t count) {
        try {
            if (raw_bytes) {
                in_s.read(dst.data(), (std::streamsize)(count * per * elem_bytes));
                if (!in_s) {
                    read_ok = false;
                }
            } else {
                // the buffer is raw storage; GroupT is a trivially destructible aggregate of
                // limbs, so the elements built here need no destructor call
                GroupT *g = (GroupT *)dst.data();
                for (size_t i = 0; i < count * per; ++i) {
                    new (g + i) GroupT();
                    group_read<encoding_binary, Form, Comp>(g[i], in_s);
                }
                if (!in_s) {
                    read_ok = false;
                }
            }
        } catch (...) {
            read_error = std::current_exception();
            read_ok = false;
        }
    };
    size_t done = 0, cur = 0;
    size_t count = std::min(chunk, n);
    read_chunk(buf[cur], count);
    bool ok = true;
    while (ok && read_ok && done < n) {
        const size_t next_count = std::min(chunk, n - done - count);
        std::thread reader;
        if (next_count) {
            try {
                reader = std::thread([&, next_count] { read_chunk(buf[cur ^ 1], next_count); });
            } catch (...) { // no thread to be had: read on this one after the push
            }
        }
        ok = b200msm_stream_push(stream, buf[cur].data(), count * per * elem_bytes, count);
        if (reader.joinable()) {
            reader.join();
        } else if (next_count) {
            read_chunk(buf[cur ^ 1], next_count);
        }
        done += count;
        count = next_count;
        cur ^= 1;
    }
    GroupT result;
    if (read_error) {
        std::rethrow_exception(read_error); // e.g. a malformed compressed point; guard releases
    }
    if (!ok || !read_ok) {
        const std::string why = ok ? "stream ended early" : b200msm_last_error();
        throw std::runtime_error("b200 multi_exp_stream: " + why);
    }
    guard.handle = 0; // finish releases the stream itself, whatever it returns
    if (!b200msm_stream_finish(stream, (void *)&result, sizeof(GroupT))) {
        fail("b200msm_stream_finish");
    }
    return result;
}

} // namespace b200
} // namespace libff

// ---- explicit specialisations of the reference templates ------------------------------------

#define LIBFF_B200_SPECIALISE_MULTI_EXP(G, F, METHOD, FORM)                                   \
    namespace libff                                                                           \
    {                                                                                         \
    template<>                                                                                \
    inline G multi_exp<G, F, METHOD, FORM>(                                                   \
        std::vector<G>::const_iterator vec_start,                                             \
        std::vector<G>::const_iterator vec_end,                                               \
        std::vector<F>::const_iterator scalar_start,                                          \
        std::vector<F>::const_iterator scalar_end,                                            \
        const size_t chunks)                                                                  \
    {                                                                                         \
        (void)chunks; /* OpenMP split hint of the CPU path, multiexp.tcc:655-687 */           \
        return b200::multi_exp<G, F>(vec_start, vec_end, scalar_start, scalar_end, FORM);     \
    }                                                                                         \
    /* multiexp.tcc:690-757: the CPU pre-pass (skip zeros, add the bases of unit scalars one  \
       by one) is pointless on the device, where a zero scalar yields no bucket item and a    \
       unit scalar exactly one */                                                             \
    template<>                                                                                \
    inline G multi_exp_filter_one_zero<G, F, METHOD, FORM>(                                   \
        std::vector<G>::const_iterator vec_start,                                             \
        std::vector<G>::const_iterator vec_end,                                               \
        std::vector<F>::const_iterator scalar_start,                                          \
        std::vector<F>::const_iterator scalar_end,                                            \
        const size_t chunks)                                                                  \
    {                                                                                         \
        (void)chunks;                                                                         \
        return b200::multi_exp<G, F>(vec_start, vec_end, scalar_start, scalar_end, FORM);     \
    }                                                                                         \
    namespace internal                                                                        \
    {                                                                                         \
    template<> class multi_exp_implementation<G, F, METHOD, FORM>                             \
    {                                                                                         \
    public:                                                                                   \
        static G multi_exp_inner(                                                             \
            std::vector<G>::const_iterator bases,                                             \
            std::vector<G>::const_iterator bases_end,                                         \
            std::vector<F>::const_iterator exponents,                                         \
            std::vector<F>::const_iterator exponents_end)                                     \
        {                                                                                     \
            return b200::multi_exp<G, F>(bases, bases_end, exponents, exponents_end, FORM);   \
        }                                                                                     \
    };                                                                                        \
    }                                                                                         \
    }

#define LIBFF_B200_SPECIALISE_STREAM(G, F, FORM, COMP)                                        \
    namespace libff                                                                           \
    {                                                                                         \
    template<>                                                                                \
    inline G multi_exp_stream<FORM, COMP, G, F>(                                              \
        std::istream & base_elements_in, const std::vector<F> &exponents)                     \
    {                                                                                         \
        return b200::multi_exp_stream<FORM, COMP, G, F>(base_elements_in, exponents, 0);      \
    }                                                                                         \
    template<>                                                                                \
    inline G multi_exp_stream_with_precompute<FORM, COMP, G, F>(                              \
        std::istream & precomputed_elements_in,                                               \
        const std::vector<F> &exponents,                                                      \
        const size_t precompute_c)                                                            \
    {                                                                                         \
        return b200::multi_exp_stream<FORM, COMP, G, F>(                                      \
            precomputed_elements_in, exponents, precompute_c);                                \
    }                                                                                         \
    }

#define LIBFF_B200_SPECIALISE_GROUP(G, F)                                                     \
    LIBFF_B200_SPECIALISE_STREAM(G, F, form_plain, compression_off)                           \
    LIBFF_B200_SPECIALISE_STREAM(G, F, form_plain, compression_on)                            \
    LIBFF_B200_SPECIALISE_STREAM(G, F, form_montgomery, compression_off)                      \
    LIBFF_B200_SPECIALISE_STREAM(G, F, form_montgomery, compression_on)                       \
    LIBFF_B200_SPECIALISE_MULTI_EXP(G, F, multi_exp_method_BDLO12, multi_exp_base_form_normal) \
    LIBFF_B200_SPECIALISE_MULTI_EXP(G, F, multi_exp_method_BDLO12, multi_exp_base_form_special) \
    LIBFF_B200_SPECIALISE_MULTI_EXP(                                                          \
        G, F, multi_exp_method_BDLO12_signed, multi_exp_base_form_normal)                     \
    LIBFF_B200_SPECIALISE_MULTI_EXP(                                                          \
        G, F, multi_exp_method_BDLO12_signed, multi_exp_base_form_special)                    \
    namespace libff                                                                           \
    {                                                                                         \
    template<>                                                                                \
    inline std::vector<G> batch_exp<G, F>(                                                    \
        const size_t scalar_size,                                                             \
        const size_t window,                                                                  \
        const window_table<G> &table,                                                         \
        const std::vector<F> &v,                                                              \
        size_t num_entries)                                                                   \
    {                                                                                         \
        return b200::batch_exp<G, F>(scalar_size, window, table, nullptr, v, num_entries);    \
    }                                                                                         \
    template<>                                                                                \
    inline std::vector<G> batch_exp<G, F>(                                                    \
        const size_t scalar_size,                                                             \
        const size_t window,                                                                  \
        const window_table<G> &table,                                                         \
        const std::vector<F> &v)                                                              \
    {                                                                                         \
        return b200::batch_exp<G, F>(scalar_size, window, table, nullptr, v, v.size());       \
    }                                                                                         \
    /* the reference calls windowed_exp directly here (multiexp.tcc:934), so it needs its own \
       specialisation: one Fr product by coeff ahead of the digit extraction on the device */ \
    template<>                                                                                \
    inline std::vector<G> batch_exp_with_coeff<G, F>(                                         \
        const size_t scalar_size,                                                             \
        const size_t window,                                                                  \
        const window_table<G> &table,                                                         \
        const F &coeff,                                                                       \
        const std::vector<F> &v)                                                              \
    {                                                                                         \
        return b200::batch_exp<G, F>(scalar_size, window, table, &coeff, v, v.size());        \
    }                                                                                         \
    template<>                                                                                \
    inline window_table<G> get_window_table<G>(                                               \
        const size_t scalar_size, const size_t window, const G &g)                            \
    {                                                                                         \
        return b200::get_window_table<G>(scalar_size, window, g);                             \
    }                                                                                         \
    template<> inline void batch_to_special<G>(std::vector<G> & vec)                          \
    {                                                                                         \
        b200::batch_to_special<G>(vec);                                                       \
    }                                                                                         \
    }

LIBFF_B200_SPECIALISE_GROUP(alt_bn128_G1, alt_bn128_Fr)
LIBFF_B200_SPECIALISE_GROUP(alt_bn128_G2, alt_bn128_Fr)
LIBFF_B200_SPECIALISE_GROUP(bls12_377_G1, bls12_377_Fr)
LIBFF_B200_SPECIALISE_GROUP(bw6_761_G1, bw6_761_Fr)
LIBFF_B200_SPECIALISE_GROUP(bls12_377_G2, bls12_377_Fr)
LIBFF_B200_SPECIALISE_GROUP(bw6_761_G2, bw6_761_Fr)
LIBFF_B200_SPECIALISE_GROUP(bls12_381_G1, bls12_381_Fr)
LIBFF_B200_SPECIALISE_GROUP(bls12_381_G2, bls12_381_Fr)

#undef LIBFF_B200_SPECIALISE_GROUP
#undef LIBFF_B200_SPECIALISE_STREAM
#undef LIBFF_B200_SPECIALISE_MULTI_EXP

#endif // LIBFF_B200_HPP_
