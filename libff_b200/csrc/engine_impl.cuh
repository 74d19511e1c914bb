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

template <class C>
bool msm_impl(Ctx &ctx, const void *d_aff, const uint8_t *d_inf, const void *d_scalars, size_t n,
              void *d_out, const MsmShape *shape) {
    using R = Raw<C>;
    using F = typename C::Fq;
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
    const bool pre = shape != nullptr;
    // shared-bucket point ranges of one call (Ctx::range_c): fixed digit size, buckets kept
    const bool ranged = !pre && ctx.range_c > 0;
    const bool range_first = !ranged || ctx.range_first, range_last = !ranged || ctx.range_last;
    const int c = pre ? shape->c : ranged ? ctx.range_c : (int)pick_window_bits<C>(ctx, n);
    const int W = pre ? shape->W : (C::Fr::Params::BITS + 1 + c - 1) / c;
    const uint32_t B = 1u << (c - 1);
    const uint32_t bstride = pre ? 0u : B;         // precomputed multiples share one bucket set
    const size_t M = pre ? (size_t)B : (size_t)W * B;
    // W c == scalar bits: the top digit may carry out; k_digits folds such scalars to r - s
    const int fold_bit = pre && W * c - 1 < C::Fr::Params::BITS ? W * c - 1 : -1;
    if ((uint64_t)n * W >= (pre ? (1ull << 31) : (1ull << 32))) {
        set_error("multi_exp: n * windows = %llu does not fit 32-bit positions",
                  (unsigned long long)n * W);
        return false;
    }
    int32_t *digits;
    uint32_t *hist, *offsets, *buckets, *chunks, *windows;
    Entry *sorted;
    if (!reserve_as(ctx, SL_DIGITS, (size_t)W * n * 4, &digits)) return false;
    if (!reserve_as(ctx, SL_HIST, M * 4, &hist)) return false;
    if (!reserve_as(ctx, SL_OFFSETS, (M + 2) * 4, &offsets)) return false;  // + total, used windows
    if (!reserve_as(ctx, SL_BUCKETS, M * R::XW * 4, &buckets)) return false;

    // --- digits + histogram
    B200_CUDA(cudaMemsetAsync(hist, 0, M * 4, st));
    // (the highest used window is a maximum over all ranges of a call)
    if (range_first) B200_CUDA(cudaMemsetAsync(offsets + M + 1, 0, 4, st));
    uint32_t *live = nullptr;
    if (ranged) {
        if (!reserve_as(ctx, SL_LIVE, M * 4, &live)) return false;
        if (range_first) B200_CUDA(cudaMemsetAsync(live, 0, M * 4, st));
    }
    uint32_t *ranks;
    if (!reserve_as(ctx, SL_RANKS, (size_t)W * n * 4, &ranks)) return false;
    k_digits<C><<<cdiv(n, 256), 256, 0, st>>>((const uint32_t *)d_scalars, d_inf, digits, ranks, hist,
                                               offsets + M + 1, (uint32_t)n, c, W, bstride, fold_bit);
    B200_LAUNCH_CHECK(ctx);
    B200_CUDA(cudaEventRecord(ctx.ev[PH_DIGITS + 1], st));
    // --- offsets
    const uint32_t tiles = cdiv(M, SCAN_TILE);
    uint32_t *tile_sums;
    if (!reserve_as(ctx, SL_TILE_SUMS, ((size_t)tiles + 1) * 4, &tile_sums)) return false;
    k_scan_tiles<<<tiles, SCAN_THREADS, 0, st>>>(hist, tile_sums, (uint32_t)M);
    B200_LAUNCH_CHECK(ctx);
    k_scan_tile_sums<<<1, 1024, 0, st>>>(tile_sums, tiles);
    B200_LAUNCH_CHECK(ctx);
    k_scan_apply<<<tiles, SCAN_THREADS, 0, st>>>(hist, tile_sums, tiles, offsets, (uint32_t)M);
    B200_LAUNCH_CHECK(ctx);
    B200_CUDA(cudaEventRecord(ctx.ev[PH_SCAN + 1], st));
    // the number of sorted items is known after the scan; it sizes the accumulation segments.
    // Also read back: the first sorted position of every window (offsets[w * B], one strided copy).
    uint32_t meta[2] = {0, 0};
    std::vector<uint32_t> wpos((size_t)W + 1, 0);
    B200_CUDA(cudaMemcpyAsync(meta, offsets + M, 8, cudaMemcpyDeviceToHost, st));
    if (!pre)
        B200_CUDA(cudaMemcpy2DAsync(wpos.data(), 4, offsets, (size_t)B * 4, 4, (size_t)W + 1,
                                    cudaMemcpyDeviceToHost, st));
    B200_CUDA(cudaStreamSynchronize(st));
    const uint32_t total = meta[0];
    // windows (bucket sets) that hold items
    // (an earlier range may have used higher windows than this one: meta[1] is cumulative)
    const int Wu = pre ? 1 : std::max<int>(1, std::min<int>(W, (int)meta[1]));
    const uint32_t seg = pick_segment<C>(ctx, total);
    const size_t t1 = ((size_t)total + seg - 1) / seg;
    uint32_t *seg_bucket;
    if (!reserve_as(ctx, SL_SEG_BUCKET, (t1 + 1) * 4, &seg_bucket)) return false;
    // the sorted array is read four items at a time
    if (!reserve_as(ctx, SL_SORTED, ((size_t)W * n + 4) * sizeof(Entry), &sorted)) return false;
    // --- scatter
    k_scatter<C><<<dim3(cdiv(n, 256), W), 256, 0, st>>>(digits, ranks, offsets, sorted, seg_bucket,
                                                       (uint32_t)n, seg, bstride,
                                                       pre ? shape->pstride : 1u,
                                                       pre ? shape->wstride : 0u);
    B200_LAUNCH_CHECK(ctx);
    B200_CUDA(cudaEventRecord(ctx.ev[PH_SCATTER + 1], st));

    // --- reduction geometry (Br and K are powers of two, K <= Br).  A single bucket set
    // (precomputed multiples) is cut into V slices that go through the reduction like V windows.
    const uint32_t V = pre ? (B >= 4096 ? 16u : 1u) : 1u;
    const uint32_t Br = B / V;
    const int Wr = pre ? (int)V : Wu;
    uint32_t K = 16;
    while ((M / K) > 65536 && K < 128) K <<= 1;
    if (ctx.opt_reduce_chunk > 0) K = 1u << std::min(ctx.opt_reduce_chunk, 12);
    K = std::min(K, Br);
    const uint32_t nchunks = Br / K;
    uint32_t log2K = 0, log2Br = 0;
    while ((1u << log2K) < K) log2K++;
    while ((1u << log2Br) < Br) log2Br++;
    uint32_t *chunk_T, *window_T = nullptr, *horner;
    if (!reserve_as(ctx, SL_CHUNKS, (size_t)Wr * nchunks * R::XW * 4, &chunks)) return false;
    if (!reserve_as(ctx, SL_CHUNK_TOTALS, (size_t)Wr * nchunks * R::XW * 4, &chunk_T)) return false;
    if (!reserve_as(ctx, SL_WINDOWS, (size_t)2 * Wr * R::XW * 4, &windows)) return false;
    if (!reserve_as(ctx, SL_HORNER, (size_t)R::XW * 4, &horner)) return false;
    if (pre) window_T = windows + (size_t)Wr * R::XW;
    constexpr int WR_THREADS = 256;
    const size_t wr_smem = (size_t)WR_THREADS * R::XW * 4;
    static bool attr_set = false;
    if (!attr_set) {
        B200_CUDA(cudaFuncSetAttribute(k_window_reduce<C, WR_THREADS>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wr_smem));
        attr_set = true;
    }
    if (ctx.tail_pending) {  // the previous range's tail reads chunks / chunk_T / windows
        B200_CUDA(cudaStreamWaitEvent(st, ctx.ev_tail_done, 0));
        ctx.tail_pending = false;
    }

    // --- groups of windows, walked from the top window down.  The accumulation of a group is one
    // launch of k_accum_first over the group's segments; everything after it -- carry rounds,
    // bucket reduction, window totals, the group's stretch of the Horner chain -- is a string of
    // small or latency-bound kernels that runs on the high-priority side stream UNDER the
    // accumulation of the next group.  Only the bottom group's string (one window) is exposed.
    // Consecutive accumulations alternate between two streams so that the last, partly filled
    // wave of one launch overlaps the first wave of the next.
    struct Group {
        int lo, hi;  // windows [lo, hi)
    };
    std::vector<Group> groups;  // top-down
    if (pre || total == 0 || ranged) {
        groups.push_back(Group{0, Wr});
    } else {
        int ng = ctx.opt_window_groups;
        std::vector<int> sizes;  // bottom-up
        // Measured (profiles/groups_sweep_r01d.log): more than one group does not pay.  The strings
        // of small kernels are chains of dependent field products; next to an accumulation that
        // saturates the multiply pipe of every SM their warps wait for issue slots and the chains
        // stretch more than the overlap hides (2^24 alt_bn128 G1: 42.5 ms with one group, 48.6 with
        // six; bw6_761 2^19: 50 against 95 ms).  The default is therefore ONE group; the option
        // stays for experiments (an SM partition reserved for the side stream would change this).
        if (ng == 0) ng = 1;
        if (ng == 1 || Wu == 1) {
            sizes.push_back(Wu);
        } else if (ng > 1 && ng < 64) {
            for (int g = 0; g < std::min(ng, Wu); g++) sizes.push_back(0);
            for (int w = 0; w < Wu; w++) sizes[(size_t)w * sizes.size() / Wu]++;
        } else {  // 64: 1, 1, 2, then threes (bottom-up)
            int left = Wu;
            for (int want : {1, 1, 2}) {
                if (left <= 0) break;
                sizes.push_back(std::min(want, left));
                left -= sizes.back();
            }
            while (left > 0) {
                sizes.push_back(std::min(3, left));
                left -= sizes.back();
            }
        }
        int hi = Wu;
        for (size_t g = sizes.size(); g-- > 0;) {
            groups.push_back(Group{hi - sizes[g], hi});
            hi -= sizes[g];
        }
    }
    const size_t G = groups.size();
    // tails on the side stream; for a range that is not the last one the "tail" is its carry
    // rounds, which run under the digit pass and the sort of the next range
    const bool side = G > 1 || ctx.defer_tail || (ranged && !range_last);
    cudaStream_t ts = side ? ctx.tail_stream : st;
    const uint32_t SEG_CARRY = pick_carry_segment<C>(ctx);
    uint32_t *keyA = nullptr, *ptA = nullptr, *keyB = nullptr, *ptB = nullptr, *keyC = nullptr, *ptC = nullptr;
    if (total > 0) {
        const size_t t2 = (t1 + 1 + SEG_CARRY - 1) / SEG_CARRY, t3 = (t2 + SEG_CARRY - 1) / SEG_CARRY;
        if (!reserve_as(ctx, SL_CARRY_KEY_A, t1 * 4, &keyA)) return false;
        if (!reserve_as(ctx, SL_CARRY_PT_A, t1 * R::XW * 4, &ptA)) return false;
        if (!reserve_as(ctx, SL_CARRY_KEY_B, t2 * 4, &keyB)) return false;
        if (!reserve_as(ctx, SL_CARRY_PT_B, t2 * R::XW * 4, &ptB)) return false;
        if (!reserve_as(ctx, SL_CARRY_KEY_C, t3 * 4, &keyC)) return false;
        if (!reserve_as(ctx, SL_CARRY_PT_C, t3 * R::XW * 4, &ptC)) return false;
    }
    if (G > 1) {
        B200_CUDA(cudaEventRecord(ctx.ev_group_go, st));
        B200_CUDA(cudaStreamWaitEvent(ctx.acc_stream, ctx.ev_group_go, 0));
    }
    // segment that holds the first item of window w (the segment straddling two windows belongs
    // to the upper one)
    auto seg_of_window = [&](int w) -> uint32_t { return w <= 0 ? 0u : wpos[(size_t)w] / seg; };
    for (size_t g = 0; g < G; g++) {
        const Group &gr = groups[g];
        const bool first = g == 0, last = g + 1 == G;
        cudaStream_t as = (g & 1) ? ctx.acc_stream : st;
        if (total > 0) {
            const uint32_t tb = (pre || last) ? 0u : seg_of_window(gr.lo);
            const uint32_t te = (pre || first) ? (uint32_t)t1 : seg_of_window(gr.hi);
            if (te > tb) {
                if (ranged)
                    k_accum_first_merge<C><<<cdiv(te - tb, 128), 128, 0, as>>>(
                        sorted, offsets, seg_bucket, total, seg, tb, te, (const uint32_t *)d_aff,
                        buckets, keyA, ptA, live);
                else
                    k_accum_first<C><<<cdiv(te - tb, 128), 128, 0, as>>>(
                        sorted, offsets, seg_bucket, total, seg, tb, te, (const uint32_t *)d_aff,
                        buckets, keyA, ptA);
                B200_LAUNCH_CHECK(ctx);
            }
            if (side) {
                B200_CUDA(cudaEventRecord(ctx.ev_group[g & 1], as));
                B200_CUDA(cudaStreamWaitEvent(ts, ctx.ev_group[g & 1], 0));
            }
            if (last) B200_CUDA(cudaEventRecord(ctx.ev[PH_ACCUM + 1], as));
            // carry rounds over entries [tb, te] (both boundary entries; each is live in one group)
            const uint32_t cb = tb, ce = std::min<uint32_t>((uint32_t)t1, te + 1);
            const uint32_t key_lo = pre ? 0u : (uint32_t)gr.lo * B;
            const uint32_t key_hi = pre ? CARRY_DEAD : (uint32_t)gr.hi * B;
            const uint32_t *kin = keyA, *pin = ptA;
            uint32_t *kout = keyB, *pout = ptB;
            uint32_t begin = cb, endp = ce;
            while (endp > begin) {
                const size_t tn = ((size_t)(endp - begin) + SEG_CARRY - 1) / SEG_CARRY;
                k_accum_carry<C><<<cdiv(tn, 128), 128, 0, ts>>>(kin, pin, begin, endp, key_lo, key_hi,
                                                                SEG_CARRY, buckets, kout, pout);
                B200_LAUNCH_CHECK(ctx);
                if (tn == 1) break;  // a single thread owns every run it sees: no carries remain
                kin = kout, pin = pout;
                kout = (kout == keyB) ? keyC : keyB;
                pout = (pout == ptB) ? ptC : ptB;
                begin = 0, endp = (uint32_t)tn;
            }
        } else if (last) {
            B200_CUDA(cudaEventRecord(ctx.ev[PH_ACCUM + 1], st));
        }
        if (!side) B200_CUDA(cudaEventRecord(ctx.ev[PH_CARRY + 1], st));
        if (!range_last) break;  // the buckets stay; the next range of the call adds to them
        // bucket reduction, window totals, Horner stretch of this group
        k_bucket_reduce<C><<<dim3(cdiv(nchunks, 64), gr.hi - gr.lo), 64, 0, ts>>>(
            buckets, ranged ? live : hist, Br, K, (uint32_t)gr.lo, chunks, chunk_T);
        B200_LAUNCH_CHECK(ctx);
        if (!side) B200_CUDA(cudaEventRecord(ctx.ev[PH_BUCKET_REDUCE + 1], st));
        if (last && ctx.defer_tail) {
            // the next range's digit pass rewrites hist / buckets: it may start from here on
            B200_CUDA(cudaEventRecord(ctx.ev_tail_go, ts));
            B200_CUDA(cudaStreamWaitEvent(st, ctx.ev_tail_go, 0));
        }
        k_window_reduce<C, WR_THREADS><<<gr.hi - gr.lo, WR_THREADS, wr_smem, ts>>>(
            chunks, chunk_T, nchunks, log2K, (uint32_t)gr.lo, windows, window_T);
        B200_LAUNCH_CHECK(ctx);
        if (pre)
            k_final_slices<C><<<1, 32, 0, ts>>>(windows, window_T, (int)V, (int)log2Br, (uint32_t *)d_out);
        else
            k_final<C><<<1, 32, 0, ts>>>(windows, gr.lo, gr.hi, c, horner, first, last, (uint32_t *)d_out);
        B200_LAUNCH_CHECK(ctx);
    }
    if (G > 1 && (G & 1) == 0) {
        // the last accumulation ran on the second stream: the caller's stream must see it
        // (ev_group[1] was recorded after it)
        B200_CUDA(cudaStreamWaitEvent(st, ctx.ev_group[1], 0));
    }
    if (side) {
        B200_CUDA(cudaEventRecord(ctx.ev_tail_done, ts));
        if (ctx.defer_tail || (ranged && !range_last)) {
            ctx.tail_pending = true;
        } else {
            B200_CUDA(cudaStreamWaitEvent(st, ctx.ev_tail_done, 0));
        }
    }
    if (side) {  // overlapped: what is left after the last accumulation counts as carry rounds
        B200_CUDA(cudaEventRecord(ctx.ev[PH_CARRY + 1], st));
        B200_CUDA(cudaEventRecord(ctx.ev[PH_BUCKET_REDUCE + 1], st));
    }
    B200_CUDA(cudaEventRecord(ctx.ev[PH_FINAL + 1], st));
    return true;
}

// table[w * n + i] = 2^(w c) P_i in device affine form (the multiples the reference's
// multi_exp_stream_with_precompute reads from its stream, multiexp_stream.hpp:27-41), built on the
// device: c doublings per step in XYZZ, one batch normalisation per multiple.
template <class C>
bool precompute_impl(Ctx &ctx, const void *d_aff, const uint8_t *d_inf, size_t n, int c, int W,
                     void *d_table) {
    using R = Raw<C>;
    if (n == 0 || W <= 0) return true;
    uint32_t *state, *raw;
    uint8_t *inf_scratch;
    if (!reserve_as(ctx, SL_PRE_STATE, n * R::XW * 4, &state)) return false;
    if (!reserve_as(ctx, SL_PRE_RAW, n * R::PW * 4, &raw)) return false;
    if (!reserve_as(ctx, SL_TABLE_INF, n, &inf_scratch)) return false;
    uint32_t *table = (uint32_t *)d_table;
    B200_CUDA(cudaMemcpyAsync(table, d_aff, n * R::AW * 4, cudaMemcpyDeviceToDevice, ctx.stream));
    for (int w = 1; w < W; w++) {
        k_pre_step<C><<<cdiv(n, 128), 128, 0, ctx.stream>>>((const uint32_t *)d_aff, d_inf, state, raw,
                                                            (uint32_t)n, c, w == 1);
        B200_LAUNCH_CHECK(ctx);
        if (!normalise_dev<C>(ctx, raw, n, nullptr, table + (size_t)w * n * R::AW, inf_scratch))
            return false;
    }
    return true;
}

template <class C>
bool batch_exp_impl(Ctx &ctx, const void *d_table_raw, size_t rows, size_t row_len, int window,
                    const void *d_scalars, size_t n, void *d_out) {
    using R = Raw<C>;
    const size_t entries = rows * row_len;
    uint32_t *taff;
    uint8_t *tinf;
    if (!reserve_as(ctx, SL_TABLE_AFF, entries * R::AW * 4, &taff)) return false;
    if (!reserve_as(ctx, SL_TABLE_INF, entries, &tinf)) return false;
    // table entries are arbitrary Jacobian/projective representatives: normalise once
    if (!normalise_dev<C>(ctx, (const uint32_t *)d_table_raw, entries, nullptr, taff, tinf)) return false;
    if (n == 0) return true;
    k_batch_exp<C><<<cdiv(n, 128), 128, 0, ctx.stream>>>((const uint32_t *)d_scalars, taff,
                                                         (uint32_t)rows, (uint32_t)row_len, window,
                                                         (uint32_t *)d_out, (uint32_t)n);
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
    if (op <= DBG_FQ_INV || op == DBG_FQ_MUL_D || op == DBG_FQ_MUL2 ||
        (op >= DBG_FQ_MUL_ROLLED && op <= DBG_FQ_SQR_CALLED))
        return (size_t)R::CW * 4;
    if (op == DBG_FR_FROM_MONT || op == DBG_FR_MUL || op == DBG_FR_MUL_D) return (size_t)R::SW * 4;
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
                                 &precompute_impl<C>,
                                 &stream_decode_impl<C>,
                                 &sum_impl<C>,
                                 &to_special_impl<C>,
                                 &batch_exp_impl<C>,
                                 &window_bits_impl<C>,
                                 &wire_decode_impl<C>,
                                 &wire_encode_impl<C>,
                                 &debug_op_impl<C>,
                                 &debug_elem_bytes_impl<C>};
    return &ops;
}

}  // namespace b200msm
