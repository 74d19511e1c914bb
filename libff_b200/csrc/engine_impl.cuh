// Templated host orchestration of the device pipelines; instantiated once per group.
#pragma once
#include <algorithm>

#include "batch_kernels.cuh"
#include "debug_kernels.cuh"
#include "engine.hpp"
#include "msm_kernels.cuh"
#include "wire_kernels.cuh"

namespace b200msm {

#define B200_LAUNCH_CHECK(ctx)            \
    do {                                  \
        (ctx).launches++;                 \
        B200_CUDA(cudaPeekAtLastError()); \
    } while (0)

inline unsigned cdiv(size_t a, size_t b) { return (unsigned)((a + b - 1) / b); }

// Window size: minimises  n*W*10 + W*2^(c-1)*53, W = ceil((bits + 1) / c): 10 field
// multiplications per accumulated entry (8M+2S mixed addition) against the measured cost of one
// bucket in the reduction (two 12M+2S additions + chunk offset, at lower occupancy; B200,
// alt_bn128 G1, 2^24 points: 0.157 ns per entry vs 0.84 ns per bucket).  The reference's own choice is pippenger_optimal_c
// (multiexp.tcc:35-40); any c gives the same group element.
template <class C>
uint32_t pick_window_bits(const Ctx &ctx, size_t n) {
    if (ctx.opt_window_bits > 0) return (uint32_t)std::min(std::max(ctx.opt_window_bits, 2), 24);
    const int bits = C::Fr::Params::BITS;
    double best = 0;
    uint32_t best_c = 2;
    for (uint32_t c = 2; c <= 22; c++) {
        double W = (double)((bits + 1 + c - 1) / c);
        // (quadratic extensions: 40 -- measured at 2^20, profiles/segment_window_sweep_r01e.log)
        const double bucket_cost = BaseLimbs<typename C::Fq>::EXT ? 40.0 : 53.0;
        double cost = (double)n * W * 10.0 + W * (double)(1u << (c - 1)) * bucket_cost;
        if (c == 2 || cost < best) best = cost, best_c = c;
    }
    return best_c;
}

template <class C>
bool ingest_impl(Ctx &ctx, const void *d_raw, size_t n, uint32_t form, void *d_aff, uint8_t *d_inf);
template <class C>
bool to_special_impl(Ctx &ctx, void *d_raw, size_t n);

// ------------------------------------------------------------------------------------------
// 1 / val[i] for n field elements (in place)
template <class F>
bool batch_invert_dev(Ctx &ctx, uint32_t *val, size_t n) {
    // level sizes n0 = n, n1 = ceil(n0 / CHUNK), ... until <= CHUNK
    std::vector<size_t> sizes{n};
    while (sizes.back() > (size_t)INV_CHUNK) sizes.push_back((sizes.back() + INV_CHUNK - 1) / INV_CHUNK);
    size_t pre_total = 0, prod_total = 0;
    for (size_t l = 0; l < sizes.size(); l++) {
        pre_total += sizes[l];
        if (l > 0) prod_total += sizes[l];
    }
    uint32_t *pre, *prod;
    const size_t eb = sizeof(uint32_t) * F::N;
    if (!reserve_as(ctx, SL_INV_PRE, pre_total * eb, &pre)) return false;
    if (!reserve_as(ctx, SL_INV_PROD, std::max<size_t>(prod_total, 1) * eb, &prod)) return false;
    std::vector<uint32_t *> lv_val{val}, lv_pre{pre};
    {
        uint32_t *pp = pre + sizes[0] * F::N, *pr = prod;
        for (size_t l = 1; l < sizes.size(); l++) {
            lv_val.push_back(pr);
            lv_pre.push_back(pp);
            pr += sizes[l] * F::N;
            pp += sizes[l] * F::N;
        }
    }
    const size_t L = sizes.size();
    for (size_t l = 0; l + 1 < L; l++) {
        unsigned threads = (unsigned)sizes[l + 1];
        k_inv_up<F><<<cdiv(threads, 128), 128, 0, ctx.stream>>>(lv_val[l], lv_pre[l], lv_val[l + 1],
                                                              (uint32_t)sizes[l]);
        B200_LAUNCH_CHECK(ctx);
    }
    k_inv_root<F><<<1, 32, 0, ctx.stream>>>(lv_val[L - 1], lv_pre[L - 1], (uint32_t)sizes[L - 1]);
    B200_LAUNCH_CHECK(ctx);
    for (size_t l = L - 1; l-- > 0;) {
        unsigned threads = (unsigned)sizes[l + 1];
        k_inv_down<F><<<cdiv(threads, 128), 128, 0, ctx.stream>>>(lv_val[l], lv_pre[l], lv_val[l + 1],
                                                                (uint32_t)sizes[l]);
        B200_LAUNCH_CHECK(ctx);
    }
    return true;
}

// Normalises n reference-layout points; writes reference layout (out_raw, may alias raw) and/or
// device affine + flags.
template <class C>
bool normalise_dev(Ctx &ctx, const uint32_t *raw, size_t n, uint32_t *out_raw, uint32_t *aff,
                   uint8_t *inf) {
    using F = typename C::Fq;
    if (n == 0) return true;
    uint32_t *z;
    if (!reserve_as(ctx, SL_INV_VAL, n * sizeof(uint32_t) * F::N, &z)) return false;
    k_extract_z<C><<<cdiv(n, 128), 128, 0, ctx.stream>>>(raw, z, (uint32_t)n);
    B200_LAUNCH_CHECK(ctx);
    if (!batch_invert_dev<F>(ctx, z, n)) return false;
    k_apply_zinv<C><<<cdiv(n, 128), 128, 0, ctx.stream>>>(raw, z, out_raw, aff, inf, (uint32_t)n);
    B200_LAUNCH_CHECK(ctx);
    return true;
}

template <class C>
bool to_special_impl(Ctx &ctx, void *d_raw, size_t n) {
    return normalise_dev<C>(ctx, (const uint32_t *)d_raw, n, (uint32_t *)d_raw, nullptr, nullptr);
}

template <class C>
bool ingest_impl(Ctx &ctx, const void *d_raw, size_t n, uint32_t form, void *d_aff, uint8_t *d_inf) {
    if (n == 0) return true;
    if (form == 2) {  // packed upload: d_raw == d_aff holds device-affine records already
        k_flags_from_affine<C><<<cdiv(n, 256), 256, 0, ctx.stream>>>((const uint32_t *)d_aff, d_inf,
                                                                   (uint32_t)n);
        B200_LAUNCH_CHECK(ctx);
        return true;
    }
    if (form == 1) {
        k_ingest_special<C><<<cdiv(n, 128), 128, 0, ctx.stream>>>((const uint32_t *)d_raw,
                                                                (uint32_t *)d_aff, d_inf, (uint32_t)n);
        B200_LAUNCH_CHECK(ctx);
        return true;
    }
    return normalise_dev<C>(ctx, (const uint32_t *)d_raw, n, nullptr, (uint32_t *)d_aff, d_inf);
}

template <class C>
bool sum_impl(Ctx &ctx, const void *d_raw, size_t count, void *d_out) {
    k_sum_raw<C><<<1, 32, 0, ctx.stream>>>((const uint32_t *)d_raw, (uint32_t)count, (uint32_t *)d_out);
    B200_LAUNCH_CHECK(ctx);
    return true;
}

// ------------------------------------------------------------------------------------------
// Carries per thread in the later accumulation rounds.  Every thread walks its carries serially
// with full additions (14 field products each): long segments need fewer rounds, short ones keep
// the serial chain of the wide fields (microseconds per product) short.
template <class C>
inline uint32_t pick_carry_segment(const Ctx &ctx) {
    if (ctx.opt_carry_segment > 0) return (uint32_t)ctx.opt_carry_segment;
    constexpr int words = Raw<C>::CW;  // 32-bit words per coordinate
    (void)words;
    return 4u;  // measured best (or equal) for 8-, 12-, 16- and 24-word coordinates
}

// Sorted entries per thread in the first accumulation round: long segments amortise the carry
// rounds (one full addition per segment), short ones keep all SMs busy on small inputs.
// Measured (profiles/segment_window_sweep_r01e.log): prime fields gain from 128-entry segments
// when enough threads remain (alt_bn128 G1 2^22: 12.9 -> 12.5 ms, 2^24: 42.6 -> 42.3); the 24-limb
// field, whose carry rounds are the most expensive and whose kernel holds only 256 threads per SM,
// wants them even below one wave (bw6_761 2^19: 46.2 -> 42.3 ms, 2^17: 21.1 -> 19.0); quadratic
// extensions lose with more than 64.
template <class C>
inline uint32_t pick_segment(const Ctx &ctx, size_t total) {
    if (ctx.opt_segment > 0) return ((uint32_t)ctx.opt_segment + 3u) & ~3u;  // uint4 item loads
    constexpr bool ext = BaseLimbs<typename C::Fq>::EXT;
    constexpr int limbs = BaseLimbs<typename C::Fq>::N;
    const size_t want_threads = (!ext && limbs >= 24) ? 148 * 128 : 148 * 384 * 4;
    uint32_t seg = ext ? 64 : 128;
    while (seg > 8 && total / seg < want_threads) seg >>= 1;
    return seg;
}

// the shared-memory accumulation kernel exists for the wide prime fields only
template <class C>
bool attr_accum_sm() {
    if constexpr (SmemAccum<C>::value)
        B200_CUDA(cudaFuncSetAttribute(k_accum_first_sm<C>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)SmemAccum<C>::BYTES));
    return true;
}
template <class C, class... A>
void launch_accum_sm(unsigned blocks, cudaStream_t st, const Entry *sorted, const uint32_t *offsets,
                     const uint32_t *seg_bucket, uint32_t total, uint32_t seg, uint32_t t1,
                     const uint32_t *aff, uint32_t *buckets, uint32_t *carry_key, uint32_t *carry_pt,
                     uint32_t key_base) {
    if constexpr (SmemAccum<C>::value)
        k_accum_first_sm<C><<<blocks, SmemAccum<C>::THREADS, SmemAccum<C>::BYTES, st>>>(
            sorted, offsets, seg_bucket, total, seg, 0u, t1, aff, buckets, carry_key, carry_pt, key_base);
}

// ------------------------------------------------------------------------------------------
// Geometry of one MSM call
template <class C>
struct MsmGeom {
    bool pre, ranged;
    int c, W;
    uint32_t B, bstride;
    size_t M;
    int fold_bit;
};

template <class C>
MsmGeom<C> msm_geom(const Ctx &ctx, size_t n, const MsmShape *shape) {
    MsmGeom<C> g;
    g.pre = shape != nullptr;
    // shared-bucket point ranges of one call (Ctx::range_c): fixed digit size, buckets kept
    g.ranged = !g.pre && ctx.range_c > 0;
    g.c = g.pre ? shape->c : g.ranged ? ctx.range_c : (int)pick_window_bits<C>(ctx, n);
    g.W = g.pre ? shape->W : (C::Fr::Params::BITS + 1 + g.c - 1) / g.c;
    g.B = 1u << (g.c - 1);
    g.bstride = g.pre ? 0u : g.B;  // precomputed multiples share one bucket set
    g.M = g.pre ? (size_t)g.B : (size_t)g.W * g.B;
    // W c == scalar bits: the top digit may carry out.  Default: k_digits folds such scalars to
    // r - s (the true sum); drop_top_carry: the reference's recoding, carry dropped (-2).
    g.fold_bit = -1;
    if (g.pre && g.W * g.c - 1 < C::Fr::Params::BITS) g.fold_bit = g.W * g.c - 1;
    if (g.pre && shape->drop_top_carry) g.fold_bit = -2;
    return g;
}

// Measured (profiles/sweep_sort_r02.log): the overlap hides the sort -- at 2^24 the exposed digit
// pass falls from 2.0 to 0.5 ms and the scatter from 2.35 to 0.2 ms -- but the accumulation grows
// by the same 3.6 ms (42.04 against 42.14 ms in total; within 1 % at 2^20..2^22 too): its blocks
// hold every register of an SM, so the sort's blocks only run in slots an accumulation block gave
// up.  The groups therefore stay an option (sort_overlap > 1); the default is one sort, which also
// keeps the phase timings of a call disjoint.
#ifndef B200_SORT_OVERLAP_MIN_LOG2
#define B200_SORT_OVERLAP_MIN_LOG2 62
#endif
// Groups of windows (ascending).  The counting sort of group g + 1 -- digit pass, scan, scatter:
// L2-atomic / random-access bound -- runs on the sort stream under the accumulation of group g,
// which is bound by the integer multiply pipe: different resources, so the 10 % of the step the
// sort used to hold at 2^24 disappears behind the accumulation.  The first group is one window,
// so that the only exposed sort is short.
inline std::vector<int> pick_sort_groups(const Ctx &ctx, size_t n, int W, bool single) {
    std::vector<int> sizes;
    int want = ctx.opt_sort_overlap ? ctx.opt_sort_overlap : ctx.opt_window_groups;
    if (single || W <= 3 || want == 1) return {W};
    if (want == 0) {
        if ((uint64_t)n * W < (1ull << B200_SORT_OVERLAP_MIN_LOG2)) return {W};  // launch overheads dominate
        // Sorting a window costs ~13 % of accumulating it, so a group may be several times larger
        // than the one whose accumulation hides its sort: 1, 4, the rest.
        if (W <= 6) return {1, W - 1};
        return {1, 4, W - 5};
    }
    want = std::min(want, W);
    sizes.push_back(1);
    int left = W - 1;
    for (int g = 1; g < want; g++) {
        const int take = (left + (want - g) - 1) / (want - g);
        sizes.push_back(take);
        left -= take;
    }
    return sizes;
}

// Upper bound of the accumulation threads (= carry entries) of a group holding `items` sorted items
template <class C>
inline size_t segments_bound(const Ctx &ctx, size_t items) {
    if (ctx.opt_segment > 0) return items / (((size_t)ctx.opt_segment + 3u) & ~(size_t)3u) + 2;
    constexpr bool ext = BaseLimbs<typename C::Fq>::EXT;
    constexpr int limbs = BaseLimbs<typename C::Fq>::N;
    const size_t want_threads = (!ext && limbs >= 24) ? 148 * 128 : 148 * 384 * 4;
    return std::max<size_t>(2 * want_threads, items / (ext ? 64 : 128)) + 2;
}

// Scratch of one MSM, reserved once before any kernel is queued (nothing regrows mid-call)
template <class C>
struct MsmScratch {
    int32_t *digits[2];
    uint32_t *ranks[2], *seg_bucket[2], *offsets[2];
    Entry *sorted[2];
    uint32_t *hist, *buckets, *meta, *tile_sums, *live;
    uint32_t *keyA, *ptA, *keyB, *ptB, *keyC, *ptC;
    size_t carry_cap;
};

template <class C>
bool msm_reserve(Ctx &ctx, const MsmGeom<C> &g, size_t n, const std::vector<int> &groups,
                 MsmScratch<C> &sc) {
    using R = Raw<C>;
    int wg_max = 0;
    for (int w : groups) wg_max = std::max(wg_max, w);
    const size_t items = (size_t)wg_max * n;
    const bool two = groups.size() > 1 || g.ranged;  // a second buffer set only where sorts overlap
    static const Slot DIG[2] = {SL_DIGITS, SL_DIGITS_B}, RNK[2] = {SL_RANKS, SL_RANKS_B},
                      SRT[2] = {SL_SORTED, SL_SORTED_B}, SGB[2] = {SL_SEG_BUCKET, SL_SEG_BUCKET_B},
                      OFF[2] = {SL_OFFSETS, SL_OFFSETS_B};
    size_t carry_cap = 0;
    for (int w : groups) carry_cap += segments_bound<C>(ctx, (size_t)w * n);
    for (int b = 0; b < 2; b++) {
        if (b == 1 && !two) {
            sc.digits[1] = sc.digits[0], sc.ranks[1] = sc.ranks[0], sc.sorted[1] = sc.sorted[0];
            sc.seg_bucket[1] = sc.seg_bucket[0], sc.offsets[1] = sc.offsets[0];
            break;
        }
        if (!reserve_as(ctx, DIG[b], items * 4, &sc.digits[b])) return false;
        if (!reserve_as(ctx, RNK[b], items * 4, &sc.ranks[b])) return false;
        // the sorted array is read four items at a time
        if (!reserve_as(ctx, SRT[b], (items + 4) * sizeof(Entry), &sc.sorted[b])) return false;
        if (!reserve_as(ctx, SGB[b], (segments_bound<C>(ctx, items) + 1) * 4, &sc.seg_bucket[b]))
            return false;
        // every group's slice is followed by one extra entry (its item count)
        if (!reserve_as(ctx, OFF[b], (g.M + groups.size() + 2) * 4, &sc.offsets[b])) return false;
    }
    if (!reserve_as(ctx, SL_HIST, g.M * 4, &sc.hist)) return false;
    if (!reserve_as(ctx, SL_BUCKETS, g.M * R::XW * 4, &sc.buckets)) return false;
    if (!reserve_as(ctx, SL_META, 64, &sc.meta)) return false;
    const size_t tiles_max = (size_t)cdiv(g.M, SCAN_TILE) + 1;
    if (!reserve_as(ctx, SL_TILE_SUMS, (tiles_max + 1) * 4, &sc.tile_sums)) return false;
    sc.live = nullptr;
    if (g.ranged && !reserve_as(ctx, SL_LIVE, g.M * 4, &sc.live)) return false;
    const uint32_t SEG_CARRY = pick_carry_segment<C>(ctx);
    const size_t t2 = (carry_cap + 1 + SEG_CARRY - 1) / SEG_CARRY, t3 = (t2 + SEG_CARRY - 1) / SEG_CARRY;
    if (!reserve_as(ctx, SL_CARRY_KEY_A, carry_cap * 4, &sc.keyA)) return false;
    if (!reserve_as(ctx, SL_CARRY_PT_A, carry_cap * R::XW * 4, &sc.ptA)) return false;
    if (!reserve_as(ctx, SL_CARRY_KEY_B, t2 * 4, &sc.keyB)) return false;
    if (!reserve_as(ctx, SL_CARRY_PT_B, t2 * R::XW * 4, &sc.ptB)) return false;
    if (!reserve_as(ctx, SL_CARRY_KEY_C, t3 * 4, &sc.keyC)) return false;
    if (!reserve_as(ctx, SL_CARRY_PT_C, t3 * R::XW * 4, &sc.ptC)) return false;
    sc.carry_cap = carry_cap;
    return true;
}

template <class C>
bool msm_prepare_impl(Ctx &ctx, size_t n, const MsmShape *shape) {
    if (n == 0) return true;
    const MsmGeom<C> g = msm_geom<C>(ctx, n, shape);
    MsmScratch<C> sc;
    return msm_reserve<C>(ctx, g, n, pick_sort_groups(ctx, n, g.W, g.pre || g.ranged), sc);
}

template <class C>
bool msm_impl(Ctx &ctx, const void *d_aff, const uint8_t *d_inf, const void *d_scalars, size_t n,
              void *d_out, const MsmShape *shape) {
    using R = Raw<C>;
    cudaStream_t st = ctx.stream;
    if (n == 0) {  // multiexp.tcc:291-309: zero element
        k_sum_raw<C><<<1, 32, 0, st>>>(nullptr, 0, (uint32_t *)d_out);
        B200_LAUNCH_CHECK(ctx);
        return true;
    }
    if (n >= (1ull << 31)) {
        set_error("multi_exp: n = %zu exceeds the 2^31 - 1 points one device shard addresses", n);
        return false;
    }
    const MsmGeom<C> geo = msm_geom<C>(ctx, n, shape);
    const bool pre = geo.pre, ranged = geo.ranged;
    const bool range_first = !ranged || ctx.range_first, range_last = !ranged || ctx.range_last;
    const int c = geo.c, W = geo.W;
    const uint32_t B = geo.B, bstride = geo.bstride;
    const size_t M = geo.M;
    if ((uint64_t)n * W >= (pre ? (1ull << 31) : (1ull << 32))) {
        set_error("multi_exp: n * windows = %llu does not fit 32-bit positions",
                  (unsigned long long)n * W);
        return false;
    }
    std::vector<int> gsizes = pick_sort_groups(ctx, n, W, pre || ranged);
    MsmScratch<C> sc;
    if (!msm_reserve<C>(ctx, geo, n, gsizes, sc)) return false;
    uint32_t *const buckets = sc.buckets;
    uint32_t *const used_windows = sc.meta + 8;

    // --- the reduction's scratch (sizes do not depend on the data)
    const uint32_t V = pre ? (B >= 4096 ? 16u : 1u) : 1u;
    const uint32_t Br = B / V;
    // buckets per thread of the bucket reduction: 16, more only beyond 2^18 threads.  (While one
    // CTA per window walked the chunk totals the limit was 2^16 threads, K = 128 at 2^24 points;
    // with the window totals spread over a cluster, K = 32 is faster there: bucket reduction
    // 3.82 -> 3.04 ms, window totals + Horner 0.87 -> 1.13, profiles/sweep_reduce_cluster_r02b.log)
    uint32_t K = 16;
    while ((M / K) > 262144 && K < 128) K <<= 1;
    if (ctx.opt_reduce_chunk > 0) K = 1u << std::min(ctx.opt_reduce_chunk, 12);
    K = std::min(K, Br);
    const uint32_t nchunks = Br / K;
    uint32_t log2K = 0, log2Br = 0;
    while ((1u << log2K) < K) log2K++;
    while ((1u << log2Br) < Br) log2Br++;
    const int Wr_max = pre ? (int)V : W;
    uint32_t *chunks, *chunk_T, *windows, *window_T = nullptr, *horner;
    if (!reserve_as(ctx, SL_CHUNKS, (size_t)Wr_max * nchunks * R::XW * 4, &chunks)) return false;
    if (!reserve_as(ctx, SL_CHUNK_TOTALS, (size_t)Wr_max * nchunks * R::XW * 4, &chunk_T)) return false;
    if (!reserve_as(ctx, SL_WINDOWS, (size_t)2 * Wr_max * R::XW * 4, &windows)) return false;
    if (!reserve_as(ctx, SL_HORNER, (size_t)R::XW * 4, &horner)) return false;
    if (pre) window_T = windows + (size_t)Wr_max * R::XW;
    constexpr int WR_THREADS = 256;
    const size_t wr_smem = (size_t)WR_THREADS * R::XW * 4;
    // the same kernel spread over a cluster of CTAs (one warp per scheduler on each SM)
    constexpr int WR_CL_THREADS = 128;
    const size_t wr_cl_smem = (size_t)WR_CL_THREADS * R::XW * 4;
    if (!(ctx.attr_mask & (1u << C::ID))) {  // function attributes are per device
        B200_CUDA(cudaFuncSetAttribute(k_window_reduce<C, WR_THREADS>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wr_smem));
        B200_CUDA(cudaFuncSetAttribute(k_window_reduce<C, WR_CL_THREADS>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wr_cl_smem));
        if (!attr_accum_sm<C>()) return false;
        ctx.attr_mask |= 1u << C::ID;
    }
    // wide prime fields: the accumulator lives in shared memory (option coop_accum = 1 turns it off)
    const bool use_smem_accum = SmemAccum<C>::value && ctx.opt_coop_accum != 1;

    // --- stream choreography.  Sorts run on the sort stream, accumulations alternate between the
    // main and the second accumulation stream (the partly filled last wave of one launch overlaps
    // the first wave of the next).  Buffer set p = parity of the sort; ev_sort[p] = "sort into set
    // p done", ev_sort[2 + p] = "the accumulation that read set p is done".
    cudaStream_t ss = ctx.sort_stream;
    B200_CUDA(cudaEventRecord(ctx.ev_group_go, st));
    // inputs (uploads, ingest) were queued on the main stream -- or, for the ranges of a pipelined
    // host call that share one bucket set, on the sort stream: their sort then starts under the
    // previous range's accumulation (an independent MSM per range may not: its digit pass
    // rewrites the histogram the previous range's bucket reduction still reads)
    if (!(ctx.sort_inputs_ready && ranged)) B200_CUDA(cudaStreamWaitEvent(ss, ctx.ev_group_go, 0));
    if (range_first) B200_CUDA(cudaMemsetAsync(used_windows, 0, 4, ss));
    if (ranged && range_first) B200_CUDA(cudaMemsetAsync(sc.live, 0, M * 4, ss));

    const size_t G = gsizes.size();
    const uint32_t SEG_CARRY = pick_carry_segment<C>(ctx);
    size_t toff = 0;      // carry entries written so far (all groups share one carry list)
    int w_lo = 0;
    int Wu = W;           // windows that hold items (known after the first digit pass)
    size_t call_items = 0;  // estimate of the sorted items of all groups
    bool any_items = false;
    cudaStream_t last_as = st;
    // sort of group g: queued on the sort stream; returns the group's item count (host sync)
    struct Sorted {
        int p;
        uint32_t total, seg;
        size_t t1;
        int lo, hi;
    };
    // First half: histogram, digits, scan and the read-back of the item count, all queued.
    auto sort_begin = [&](int lo, int hi, bool first, Sorted &out) -> bool {
        const int p = ctx.sort_parity;
        ctx.sort_parity ^= 1;
        const size_t mg = pre ? M : (size_t)(hi - lo) * B;
        uint32_t *hist_g = sc.hist + (pre ? 0 : (size_t)lo * B);
        // (consecutive groups use different buffer sets, so the item count a scan writes behind
        // its slice never lands in a neighbour's)
        uint32_t *off_g = sc.offsets[p] + (pre ? 0 : (size_t)lo * B);
        // the accumulation that last read buffer set p must be done
        B200_CUDA(cudaStreamWaitEvent(ss, ctx.ev_sort[2 + p], 0));
        B200_CUDA(cudaMemsetAsync(hist_g, 0, mg * 4, ss));
        // the first pass walks every window once to find the highest one in use
        k_digits<C><<<cdiv(n, 256), 256, 0, ss>>>((const uint32_t *)d_scalars, d_inf, sc.digits[p],
                                                   sc.ranks[p], hist_g, used_windows, (uint32_t)n, c, lo,
                                                   hi, first ? W : hi, bstride, geo.fold_bit,
                                                   pre ? shape->pstride : 1u, pre ? shape->wstride : 0u);
        B200_LAUNCH_CHECK(ctx);
        if (first) B200_CUDA(cudaEventRecord(ctx.ev[PH_DIGITS + 1], ss));
        const uint32_t tiles = cdiv(mg, SCAN_TILE);
        k_scan_tiles<<<tiles, SCAN_THREADS, 0, ss>>>(hist_g, sc.tile_sums, (uint32_t)mg);
        B200_LAUNCH_CHECK(ctx);
        k_scan_tile_sums<<<1, 1024, 0, ss>>>(sc.tile_sums, tiles);
        B200_LAUNCH_CHECK(ctx);
        k_scan_apply<<<tiles, SCAN_THREADS, 0, ss>>>(hist_g, sc.tile_sums, tiles, off_g, (uint32_t)mg);
        B200_LAUNCH_CHECK(ctx);
        if (first) B200_CUDA(cudaEventRecord(ctx.ev[PH_SCAN + 1], ss));
        // the number of sorted items sizes the accumulation segments (pinned destination: the
        // copies are queued, the host waits for them in sort_finish)
        B200_CUDA(cudaMemcpyAsync(&ctx.host_meta[2 * p], off_g + mg, 4, cudaMemcpyDeviceToHost, ss));
        B200_CUDA(cudaMemcpyAsync(&ctx.host_meta[2 * p + 1], used_windows, 4, cudaMemcpyDeviceToHost, ss));
        out.p = p, out.lo = lo, out.hi = hi;
        return true;
    };
    // Second half: the host learns the item count, then the scatter is queued.
    auto sort_finish = [&](bool first, Sorted &out) -> bool {
        const int p = out.p;
        const int wg = out.hi - out.lo;
        uint32_t *off_g = sc.offsets[p] + (pre ? 0 : (size_t)out.lo * B);
        B200_CUDA(cudaStreamSynchronize(ss));
        out.total = ctx.host_meta[2 * p];
        if (first) Wu = pre ? 1 : std::max<int>(1, std::min<int>(W, (int)ctx.host_meta[2 * p + 1]));
        // One segment length for the whole call, the one a single sort of all windows would get
        // (estimated from the first group, a single window): short segments in the small early
        // groups would only multiply the carries, and those groups run beside the next sort anyway.
        if (first) call_items = G > 1 ? (size_t)out.total * (size_t)W / (size_t)(out.hi - out.lo) : (size_t)out.total;
        out.seg = pick_segment<C>(ctx, std::max(call_items, (size_t)out.total));
        out.t1 = ((size_t)out.total + out.seg - 1) / out.seg;
        // (an estimate far below this group's own load -- a sparse first window -- must not
        // overrun the carry entries reserved for the group)
        const size_t room = segments_bound<C>(ctx, (size_t)wg * n) - 2;
        while (out.t1 > room) {
            out.seg *= 2;
            out.t1 = ((size_t)out.total + out.seg - 1) / out.seg;
        }
        if (out.total > 0) {
            k_scatter<C><<<dim3(cdiv(n, 256), wg), 256, 0, ss>>>(
                sc.digits[p], sc.ranks[p], off_g, sc.sorted[p], sc.seg_bucket[p], (uint32_t)n, out.seg,
                bstride, pre ? shape->pstride : 1u, pre ? shape->wstride : 0u);
            B200_LAUNCH_CHECK(ctx);
        }
        if (first) B200_CUDA(cudaEventRecord(ctx.ev[PH_SCATTER + 1], ss));
        B200_CUDA(cudaEventRecord(ctx.ev_sort[p], ss));
        return true;
    };
    auto accumulate_group = [&](const Sorted &sg, size_t g) -> bool {
        cudaStream_t as = (g & 1) ? ctx.acc_stream : st;
        last_as = as;
        B200_CUDA(cudaStreamWaitEvent(as, ctx.ev_sort[sg.p], 0));
        if (g == 1) {  // the second stream joins: it must not run ahead of the call's start
            B200_CUDA(cudaStreamWaitEvent(as, ctx.ev_group_go, 0));
        }
        if (sg.total > 0) {
            any_items = true;
            if (toff + sg.t1 > sc.carry_cap) {
                set_error("multi_exp: carry list overflow (%zu + %zu > %zu)", toff, sg.t1, sc.carry_cap);
                return false;
            }
            const uint32_t *off_g = sc.offsets[sg.p] + (pre ? 0 : (size_t)sg.lo * B);
            uint32_t *bk = buckets + (pre ? 0 : (size_t)sg.lo * B * R::XW);
            const uint32_t key_base = pre ? 0u : (uint32_t)sg.lo * B;
            if (ranged)
                k_accum_first_merge<C><<<cdiv(sg.t1, 128), 128, 0, as>>>(
                    sc.sorted[sg.p], off_g, sc.seg_bucket[sg.p], sg.total, sg.seg, 0u, (uint32_t)sg.t1,
                    (const uint32_t *)d_aff, bk, sc.keyA + toff, sc.ptA + toff * R::XW, key_base, sc.live);
            else if (use_smem_accum)
                launch_accum_sm<C>(cdiv(sg.t1, SmemAccum<C>::THREADS), as, sc.sorted[sg.p], off_g,
                                   sc.seg_bucket[sg.p], sg.total, sg.seg, (uint32_t)sg.t1,
                                   (const uint32_t *)d_aff, bk, sc.keyA + toff, sc.ptA + toff * R::XW,
                                   key_base);
            else
                k_accum_first<C><<<cdiv(sg.t1, 128), 128, 0, as>>>(
                    sc.sorted[sg.p], off_g, sc.seg_bucket[sg.p], sg.total, sg.seg, 0u, (uint32_t)sg.t1,
                    (const uint32_t *)d_aff, bk, sc.keyA + toff, sc.ptA + toff * R::XW, key_base);
            B200_LAUNCH_CHECK(ctx);
            toff += sg.t1;
        }
        B200_CUDA(cudaEventRecord(ctx.ev_sort[2 + sg.p], as));
        return true;
    };

    if (ctx.tail_pending && !ranged) {  // the previous chunk's tail reads chunks / chunk_T / windows
        B200_CUDA(cudaStreamWaitEvent(st, ctx.ev_tail_done, 0));
        ctx.tail_pending = false;
    }
    // group 0 is sorted up front; from then on the first half of sort g + 1 is queued, then
    // accumulation g is launched, and only then the host waits for the item count of g + 1
    Sorted cur, nxt;
    if (!sort_begin(0, gsizes[0], true, cur) || !sort_finish(true, cur)) return false;
    if (ctx.tail_pending) {  // ranged: the previous range's carry rounds add into the same buckets
        B200_CUDA(cudaStreamWaitEvent(st, ctx.ev_tail_done, 0));
        ctx.tail_pending = false;
    }
    w_lo = gsizes[0];
    for (size_t g = 0; g < G; g++) {
        bool have_next = false;
        if (g + 1 < G && w_lo < Wu) {  // windows from Wu up hold no items
            const int hi = std::min(w_lo + gsizes[g + 1], Wu);
            if (!sort_begin(w_lo, hi, false, nxt)) return false;
            w_lo = hi;
            have_next = true;
        }
        if (!accumulate_group(cur, g)) return false;
        if (!have_next) break;
        if (!sort_finish(false, nxt)) return false;
        cur = nxt;
    }
    if (last_as != st) {  // the main stream continues after the last accumulation
        B200_CUDA(cudaEventRecord(ctx.ev_group[1], last_as));
        B200_CUDA(cudaStreamWaitEvent(st, ctx.ev_group[1], 0));
    }
    if (G > 1) {  // ... and after the other stream's
        B200_CUDA(cudaStreamWaitEvent(st, ctx.ev_sort[2], 0));
        B200_CUDA(cudaStreamWaitEvent(st, ctx.ev_sort[3], 0));
    } else {
        B200_CUDA(cudaStreamWaitEvent(st, ctx.ev_sort[2 + cur.p], 0));
    }
    B200_CUDA(cudaEventRecord(ctx.ev[PH_ACCUM + 1], st));

    // tails on the side stream: for a range that is not the last one the "tail" is its carry
    // rounds, which run under the digit pass and the sort of the next range; for a deferred chunk
    // (streamed MSM) everything after the bucket reduction
    const bool side = ctx.defer_tail || (ranged && !range_last);
    cudaStream_t ts = (ranged && !range_last) ? ctx.tail_stream : st;
    if (ts != st) {
        B200_CUDA(cudaEventRecord(ctx.ev_group[0], st));
        B200_CUDA(cudaStreamWaitEvent(ts, ctx.ev_group[0], 0));
    }
    // --- carry rounds over the whole carry list (keys ascending across the groups)
    if (any_items) {
        const uint32_t *kin = sc.keyA, *pin = sc.ptA;
        uint32_t *kout = sc.keyB, *pout = sc.ptB;
        uint32_t begin = 0, endp = (uint32_t)toff;
        while (endp > begin) {
            const size_t tn = ((size_t)(endp - begin) + SEG_CARRY - 1) / SEG_CARRY;
            k_accum_carry<C><<<cdiv(tn, 128), 128, 0, ts>>>(kin, pin, begin, endp, 0u, CARRY_DEAD,
                                                            SEG_CARRY, buckets, kout, pout);
            B200_LAUNCH_CHECK(ctx);
            if (tn == 1) break;  // a single thread owns every run it sees: no carries remain
            kin = kout, pin = pout;
            kout = (kout == sc.keyB) ? sc.keyC : sc.keyB;
            pout = (pout == sc.ptB) ? sc.ptC : sc.ptB;
            begin = 0, endp = (uint32_t)tn;
        }
    }
    B200_CUDA(cudaEventRecord(ctx.ev[PH_CARRY + 1], st));
    if (!range_last) {  // the buckets stay; the next range of the call adds to them
        B200_CUDA(cudaEventRecord(ctx.ev_tail_done, ts));
        ctx.tail_pending = true;
        B200_CUDA(cudaEventRecord(ctx.ev[PH_BUCKET_REDUCE + 1], st));
        B200_CUDA(cudaEventRecord(ctx.ev[PH_FINAL + 1], st));
        return true;
    }
    // --- bucket reduction, window totals, Horner
    const int Wr = pre ? (int)V : Wu;
    k_bucket_reduce<C><<<dim3(cdiv(nchunks, 64), Wr), 64, 0, st>>>(
        buckets, ranged ? sc.live : sc.hist, Br, K, 0u, chunks, chunk_T);
    B200_LAUNCH_CHECK(ctx);
    B200_CUDA(cudaEventRecord(ctx.ev[PH_BUCKET_REDUCE + 1], st));
    cudaStream_t fs = st;
    if (ctx.defer_tail) {
        // the next chunk's digit pass rewrites hist / buckets: it may start from here on, the
        // latency-bound window totals and the Horner chain move to the side stream
        fs = ctx.tail_stream;
        B200_CUDA(cudaEventRecord(ctx.ev_tail_go, st));
        B200_CUDA(cudaStreamWaitEvent(fs, ctx.ev_tail_go, 0));
    }
    // CTAs per window: as many as still give every thread a chunk, at most 8 (portable cluster size)
    unsigned wr_cl = ctx.opt_reduce_cluster > 0 ? (unsigned)ctx.opt_reduce_cluster : 8u;
    while (wr_cl > 1 && (size_t)wr_cl * WR_CL_THREADS > std::max<size_t>(nchunks, WR_THREADS)) wr_cl >>= 1;
    if (wr_cl > 1) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)Wr * wr_cl);
        cfg.blockDim = dim3(WR_CL_THREADS);
        cfg.dynamicSmemBytes = wr_cl_smem;
        cfg.stream = fs;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = wr_cl;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        B200_CUDA(cudaLaunchKernelEx(&cfg, k_window_reduce<C, WR_CL_THREADS>, (const uint32_t *)chunks,
                                     (const uint32_t *)chunk_T, (uint32_t)nchunks, (uint32_t)log2K, 0u,
                                     windows, window_T));
        ctx.launches++;
    } else {
        k_window_reduce<C, WR_THREADS><<<Wr, WR_THREADS, wr_smem, fs>>>(chunks, chunk_T, nchunks, log2K, 0u,
                                                                        windows, window_T);
        B200_LAUNCH_CHECK(ctx);
    }
    if (pre)
        k_final_slices<C><<<1, 32, 0, fs>>>(windows, window_T, (int)V, (int)log2Br, (uint32_t *)d_out);
    else
        k_final<C><<<1, 32, 0, fs>>>(windows, 0, Wr, c, horner, true, true, (uint32_t *)d_out);
    B200_LAUNCH_CHECK(ctx);
    if (ctx.defer_tail) {
        B200_CUDA(cudaEventRecord(ctx.ev_tail_done, fs));
        ctx.tail_pending = true;
    }
    (void)side;
    B200_CUDA(cudaEventRecord(ctx.ev[PH_FINAL + 1], st));
    return true;
}

// table[w * n + i] = 2^(w c) P_i in device affine form (the multiples the reference's
// multi_exp_stream_with_precompute reads from its stream, multiexp_stream.hpp:27-41), built on the
// device: c doublings per step in XYZZ, one batch normalisation per multiple.
template <class C>
bool precompute_impl(Ctx &ctx, const void *d_aff, const uint8_t *d_inf, size_t n, int c, int W,
                     void *d_table, uint8_t *d_table_inf) {
    using R = Raw<C>;
    if (n == 0 || W <= 0) return true;
    uint32_t *state, *raw;
    if (!reserve_as(ctx, SL_PRE_STATE, n * R::XW * 4, &state)) return false;
    if (!reserve_as(ctx, SL_PRE_RAW, n * R::PW * 4, &raw)) return false;
    uint32_t *table = (uint32_t *)d_table;
    B200_CUDA(cudaMemcpyAsync(table, d_aff, n * R::AW * 4, cudaMemcpyDeviceToDevice, ctx.stream));
    B200_CUDA(cudaMemcpyAsync(d_table_inf, d_inf, n, cudaMemcpyDeviceToDevice, ctx.stream));
    for (int w = 1; w < W; w++) {
        k_pre_step<C><<<cdiv(n, 128), 128, 0, ctx.stream>>>((const uint32_t *)d_aff, d_inf, state, raw,
                                                            (uint32_t)n, c, w == 1);
        B200_LAUNCH_CHECK(ctx);
        // a multiple of a point of small even order (unchecked input in a cofactor group) can be
        // the zero element although the base is not: every multiple keeps its own flag
        if (!normalise_dev<C>(ctx, raw, n, nullptr, table + (size_t)w * n * R::AW,
                              d_table_inf + (size_t)w * n))
            return false;
    }
    return true;
}

// ---- fixed-base exponentiation (multiexp.tcc:809-947) ----------------------------------------
template <class C>
bool table_normalise_impl(Ctx &ctx, const void *d_table_raw, size_t entries, void *d_table_aff) {
    uint8_t *tinf;
    if (!reserve_as(ctx, SL_TABLE_INF, entries, &tinf)) return false;
    // table entries are arbitrary Jacobian / projective representatives: normalised once per table
    return normalise_dev<C>(ctx, (const uint32_t *)d_table_raw, entries, nullptr,
                            (uint32_t *)d_table_aff, tinf);
}

// get_window_table (multiexp.tcc:809-846) on the device: entry [r][j] = j 2^(r w) g for j below
// the row's length (the last row is shorter, :815; its tail stays zero), reference layout.
template <class C>
bool table_build_impl(Ctx &ctx, const void *d_g_raw, int scalar_bits, int window, size_t rows,
                      size_t row_len, void *d_table_raw) {
    using R = Raw<C>;
    uint32_t *row_base;
    if (!reserve_as(ctx, SL_PRE_STATE, rows * R::XW * 4, &row_base)) return false;
    k_table_row_bases<C><<<cdiv(rows, 32), 32, 0, ctx.stream>>>((const uint32_t *)d_g_raw, window,
                                                                (uint32_t)rows, row_base);
    B200_LAUNCH_CHECK(ctx);
    const uint32_t last_len = 1u << (scalar_bits - (int)(rows - 1) * window);
    const uint32_t blocks_per_row = cdiv(row_len, TABLE_RUN);
    const size_t threads = rows * blocks_per_row;
    k_table_fill<C><<<cdiv(threads, 64), 64, 0, ctx.stream>>>(row_base, (uint32_t)rows,
                                                              (uint32_t)row_len, last_len,
                                                              (uint32_t *)d_table_raw);
    B200_LAUNCH_CHECK(ctx);
    return true;
}

template <class C>
bool table_export_impl(Ctx &ctx, const void *d_table_aff, size_t entries, void *d_table_raw) {
    if (entries == 0) return true;
    k_affine_to_special<C><<<cdiv(entries, 128), 128, 0, ctx.stream>>>(
        (const uint32_t *)d_table_aff, (uint32_t *)d_table_raw, (uint32_t)entries);
    B200_LAUNCH_CHECK(ctx);
    return true;
}

template <class C>
bool batch_exp_impl(Ctx &ctx, const void *d_table_aff, size_t rows, size_t row_len, int window,
                    const void *d_scalars, const void *d_coeff, size_t n, void *d_out) {
    if (n == 0) return true;
    k_batch_exp<C><<<cdiv(n, 128), 128, 0, ctx.stream>>>(
        (const uint32_t *)d_scalars, (const uint32_t *)d_coeff, (const uint32_t *)d_table_aff,
        (uint32_t)rows, (uint32_t)row_len, window, (uint32_t *)d_out, (uint32_t)n);
    B200_LAUNCH_CHECK(ctx);
    return true;
}

template <class C>
uint32_t window_bits_impl(const Ctx &ctx, size_t n) {
    return pick_window_bits<C>(ctx, n);
}

template <class C>
bool wire_decode_impl(Ctx &ctx, const void *d_wire_points, const void *d_wire_scalars, size_t n,
                      uint32_t flags, void *d_aff, uint8_t *d_inf, void *d_scalars, uint32_t *d_err) {
    B200_CUDA(cudaMemsetAsync(d_err, 0, 4, ctx.stream));
    if (n == 0) return true;
    k_wire_decode_points<C><<<cdiv(n, 128), 128, 0, ctx.stream>>>(
        (const uint8_t *)d_wire_points, (uint32_t *)d_aff, d_inf, d_err, (uint32_t)n, flags);
    B200_LAUNCH_CHECK(ctx);
    k_wire_decode_scalars<C><<<cdiv(n, 256), 256, 0, ctx.stream>>>(
        (const uint8_t *)d_wire_scalars, (uint32_t *)d_scalars, d_err, (uint32_t)n);
    B200_LAUNCH_CHECK(ctx);
    return true;
}

template <class C>
bool stream_decode_impl(Ctx &ctx, const void *d_bytes, size_t n, uint32_t flags, void *d_aff,
                        uint8_t *d_inf) {
    if (n == 0) return true;
    k_stream_decode_points<C><<<cdiv(n, 128), 128, 0, ctx.stream>>>(
        (const uint8_t *)d_bytes, (uint32_t *)d_aff, d_inf, (uint32_t)n, flags);
    B200_LAUNCH_CHECK(ctx);
    return true;
}

template <class C>
bool wire_encode_impl(Ctx &ctx, const void *d_raw, void *d_wire_out) {
    k_wire_encode_point<C><<<1, 32, 0, ctx.stream>>>((const uint32_t *)d_raw, (uint8_t *)d_wire_out);
    B200_LAUNCH_CHECK(ctx);
    return true;
}

template <class C>
size_t debug_elem_bytes_impl(int op) {
    using R = Raw<C>;
    if (op <= DBG_FQ_INV || op == DBG_FQ_MUL2 ||
        (op >= DBG_FQ_MUL_ROLLED && op <= DBG_FQ_MUL_SEPARATED))
        return (size_t)R::CW * 4;
    if (op == DBG_FR_FROM_MONT || op == DBG_FR_MUL) return (size_t)R::SW * 4;
    if (op >= DBG_G_ADD && op <= DBG_G_NEG) return (size_t)R::PW * 4;
    return 0;
}

template <class C>
bool debug_op_impl(Ctx &ctx, int op, const void *d_a, const void *d_b, void *d_out, size_t n) {
    if (n == 0) return true;
    k_debug_op<C><<<cdiv(n, 64), 64, 0, ctx.stream>>>(op, (const uint32_t *)d_a, (const uint32_t *)d_b,
                                                       (uint32_t *)d_out, (uint32_t)n);
    B200_LAUNCH_CHECK(ctx);
    return true;
}

template <class C>
const GroupOps *make_ops() {
    using R = Raw<C>;
    static const GroupOps ops = {(uint32_t)C::ID,
                                 (size_t)R::PW * 4,
                                 (size_t)R::SW * 4,
                                 (size_t)R::AW * 4,
                                 C::Fr::Params::BITS,
                                 &ingest_impl<C>,
                                 &msm_impl<C>,
                                 &msm_prepare_impl<C>,
                                 &precompute_impl<C>,
                                 &stream_decode_impl<C>,
                                 &sum_impl<C>,
                                 &to_special_impl<C>,
                                 &table_normalise_impl<C>,
                                 &table_build_impl<C>,
                                 &table_export_impl<C>,
                                 &batch_exp_impl<C>,
                                 &window_bits_impl<C>,
                                 &wire_decode_impl<C>,
                                 &wire_encode_impl<C>,
                                 &debug_op_impl<C>,
                                 &debug_elem_bytes_impl<C>};
    return &ops;
}

}  // namespace b200msm
