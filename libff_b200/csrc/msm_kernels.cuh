// MSM kernels (bucket method) -- device replacement for multi_exp_method_BDLO12 /
// BDLO12_signed (libff/algebra/scalar_multiplication/multiexp.tcc:276-381, 507-633).
//
// Pipeline (one launch each, all windows at once):
//   k_digits        scalars: Montgomery -> plain (fp.tcc:270-281) -> signed c-bit digits
//                   (field_utils.tcc:167-239 restated as a single carry pass) + bucket histogram
//   k_scan_*        exclusive prefix sum of the histogram -> bucket offsets (3 launches)
//   k_scatter       counting-sort scatter of (point index | sign) words to offsets[bucket] + the
//                   rank the histogram pass handed out (no atomics); also records the bucket
//                   that owns the first item of every accumulation segment
//   k_accum_first   balanced segmented accumulation: every thread sums SEG consecutive sorted
//                   entries with mixed additions (multiexp.tcc:319-345 / :45-81); sums of buckets
//                   that begin inside the segment are stored, the leading partial becomes a carry
//   k_accum_carry   same on the (already sorted) carry list with full additions, until no carries
//   k_bucket_reduce running sums over chunks of buckets (multiexp.tcc:90-125, 354-376)
//   k_window_reduce per-window total
//   k_final         Horner over windows (multiexp.tcc:309-314) -> reference coordinates
// Any bucket-method schedule yields the same group element; parity is on the affine result.
#pragma once
#include <cooperative_groups.h>

#include "curves.cuh"
#include "fp_coop.cuh"

namespace b200msm {

// One sorted item is a 32-bit word: point index with the digit's sign in bit 31.  Its bucket is
// implied by its position (offsets[b] <= pos < offsets[b + 1]); 4-byte items keep the scatter
// target of one window (n x 4 B) inside the 126 MB L2, so the random writes combine there.
typedef uint32_t Entry;

// ------------------------------------------------------------------------------------------
// bases: reference layout -> device affine (x, y); infinity -> (0, 0) + flag
template <class C>
__global__ void k_ingest_special(const uint32_t *__restrict__ raw, uint32_t *__restrict__ aff,
                                 uint8_t *__restrict__ inf_flag, uint32_t n) {
    using R = Raw<C>;
    using F = typename C::Fq;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    F X, Y, Z;
    R::load(raw + (size_t)i * R::PW, X, Y, Z);
    bool inf = R::is_inf(X, Z);
    Affine<F> a{inf ? F::zero() : X, inf ? F::zero() : Y};
    a.store(aff + (size_t)i * R::AW);
    inf_flag[i] = inf;
}

// packed upload: the array already holds device-affine records; only the flags are derived
template <class C>
__global__ void k_flags_from_affine(const uint32_t *__restrict__ aff, uint8_t *__restrict__ inf_flag,
                                    uint32_t n) {
    using R = Raw<C>;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 *p = reinterpret_cast<const uint4 *>(aff + (size_t)i * R::AW);
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k < R::AW / 4; k++) {
        const uint4 v = p[k];
        acc |= v.x | v.y | v.z | v.w;
    }
    inf_flag[i] = acc == 0;
}

// ------------------------------------------------------------------------------------------
// Warp-aggregated counter updates: lanes that hit the same bucket (witness-like inputs put a
// large share of all scalars into one bucket) issue ONE atomic between them.  Must be called by
// all 32 lanes.
__device__ __forceinline__ void warp_hist_add(uint32_t *hist, bool valid, uint32_t bucket) {
    const uint32_t lane = threadIdx.x & 31;
    if (!__any_sync(0xffffffffu, valid)) return;  // e.g. the windows above short scalars
    const uint32_t peers = __match_any_sync(0xffffffffu, valid ? bucket : 0xffffffe0u + lane);
    if (valid && lane == (uint32_t)__ffs(peers) - 1) atomicAdd(&hist[bucket], (uint32_t)__popc(peers));
}
// returns this lane's slot: base of the group's reservation + rank inside the group
__device__ __forceinline__ uint32_t warp_cursor_take(uint32_t *cursor, bool valid, uint32_t bucket) {
    const uint32_t lane = threadIdx.x & 31;
    if (!__any_sync(0xffffffffu, valid)) return 0;
    const uint32_t peers = __match_any_sync(0xffffffffu, valid ? bucket : 0xffffffe0u + lane);
    const uint32_t leader = (uint32_t)__ffs(peers) - 1;
    uint32_t base = 0;
    if (valid && lane == leader) base = atomicAdd(&cursor[bucket], (uint32_t)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + (uint32_t)__popc(peers & ((1u << lane) - 1));
}

// ------------------------------------------------------------------------------------------
// scalars -> signed digits + histogram.  digits[w * n + i] in [-2^(c-1), 2^(c-1)]
// bstride = buckets per window (2^(c-1)), or 0 when all windows share ONE bucket set (precomputed
// multiples 2^(w c) P_i, multiexp_stream.tcc:124-162); inf_flag then holds one flag per table
// entry (i * pstride + w * wstride) instead of one per base.
// Windows [w_lo, w_hi) are emitted (a group of windows of one MSM; the carry chain below w_lo is
// walked without output) into digits[(w - w_lo) * n + i]; their buckets are (w - w_lo) * bstride + ..
// The walk continues to w_top >= w_hi without output, only to find the highest window in use.
// fold_bit >= 0: W c equals the scalar width, so a scalar with bit W c - 1 set would carry out of
// the top digit (the reference's field_get_signed_digits drops that carry, field_utils.tcc:205-239);
// such a scalar is replaced by r - s, whose top digit is small, and all its digits change sign:
// s P = -(r - s) P.  fold_bit == -2: the reference's own recoding instead (a digit of exactly
// 2^(c-1) becomes -2^(c-1) with a carry, field_utils.tcc:224-236) and the carry out of the top
// digit is dropped as the reference drops it.
template <class C>
__global__ void k_digits(const uint32_t *__restrict__ scalars, const uint8_t *__restrict__ inf_flag,
                         int32_t *__restrict__ digits, uint32_t *__restrict__ ranks,
                         uint32_t *__restrict__ hist, uint32_t *__restrict__ used_windows,
                         uint32_t n, int c, int w_lo, int w_hi, int w_top, uint32_t bstride,
                         int fold_bit, uint32_t pstride, uint32_t wstride) {
    using Fr = typename C::Fr;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t top = 0;  // 1 + index of the highest window with a non-zero digit
    {
        const bool live = i < n;  // every lane runs the loop: the histogram update is warp-wide
        Fr s = live ? Fr::from_mont(Fr::load(scalars + (size_t)i * Fr::N)) : Fr::zero();
        const bool per_entry = bstride == 0;
        bool skip = !live || (!per_entry && inf_flag != nullptr && inf_flag[i]);
        bool flip = false;
        if (fold_bit >= 0) {
            uint32_t word = 0;
#pragma unroll
            for (int k = 0; k < Fr::N; k++)
                if (k == (fold_bit >> 5)) word = s.v[k];
            flip = (word >> (fold_bit & 31)) & 1u;
            if (flip) s = Fr::neg(s);
        }
        const uint32_t mask = (1u << c) - 1, half = 1u << (c - 1);
        const uint32_t tie = fold_bit == -2 ? 1u : 0u;  // reference recoding: 2^(c-1) carries
        uint32_t carry = 0;
        for (int w = 0; w < w_top; w++) {
            uint32_t d = (s.v[0] & mask) + carry;
            carry = d + tie > half;
            int32_t sd = (int32_t)d - (int32_t)(carry << c);
#pragma unroll
            for (int k = 0; k < Fr::N - 1; k++) s.v[k] = __funnelshift_r(s.v[k], s.v[k + 1], c);
            s.v[Fr::N - 1] >>= c;
            if (skip) sd = 0;
            if (w < w_lo || w >= w_hi) {  // uniform across the block
                if (sd != 0) top = w + 1;
                continue;
            }
            if (flip) sd = -sd;
            if (per_entry && sd != 0 && inf_flag != nullptr &&
                inf_flag[(size_t)i * pstride + (size_t)w * wstride])
                sd = 0;
            const size_t slot = (size_t)(w - w_lo) * n + i;
            if (live) digits[slot] = sd;
            if (sd != 0) top = w + 1;
            // the counter's value before this item = the item's rank inside its bucket: the
            // scatter then needs no second round of atomics
            const uint32_t rank = warp_cursor_take(
                hist, sd != 0, (uint32_t)(w - w_lo) * bstride + (uint32_t)(sd < 0 ? -sd : sd) - 1);
            if (live) ranks[slot] = rank;
        }
    }
    // the reference sizes its loop by the longest scalar (num_bits, multiexp.tcc:298-306); here
    // the windows above the highest non-zero digit are dropped from the reduction phases
    top = __reduce_max_sync(0xffffffffu, top);
    if ((threadIdx.x & 31) == 0 && top > 0) atomicMax(used_windows, top);
}

// ------------------------------------------------------------------------------------------
// Exclusive scan of the m bucket counters in three launches: per-tile sums, scan of the tile
// sums (one block), per-tile scan + tile offset.  offsets has m + 1 entries.
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;                          // counters per thread
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;   // counters per block

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t *warp_sums,
                                                         uint32_t &block_total) {
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) warp_sums[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        uint32_t w = lane < (blockDim.x >> 5) ? warp_sums[lane] : 0;
        uint32_t winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t o = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += o;
        }
        warp_sums[lane] = winc - w;                    // exclusive prefix of warp totals
        if (lane == 31) warp_sums[32] = winc;          // block total
    }
    __syncthreads();
    block_total = warp_sums[32];
    return warp_sums[wid] + inc - v;
}

static __global__ void __launch_bounds__(SCAN_THREADS)
    k_scan_tiles(const uint32_t *__restrict__ hist, uint32_t *__restrict__ tile_sums, uint32_t m) {
    __shared__ uint32_t ws[33];
    const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++)
        if (base + j < m) s += hist[base + j];
    uint32_t total;
    block_exclusive_scan(s, ws, total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// in-place exclusive scan of the tile sums by one block; tile_sums[tiles] = grand total
static __global__ void __launch_bounds__(1024)
    k_scan_tile_sums(uint32_t *__restrict__ tile_sums, uint32_t tiles) {
    __shared__ uint32_t ws[33];
    uint32_t carry = 0;
    for (uint32_t lo = 0; lo < tiles; lo += 1024) {
        const uint32_t k = lo + threadIdx.x;
        uint32_t v = k < tiles ? tile_sums[k] : 0, total;
        uint32_t ex = block_exclusive_scan(v, ws, total);
        if (k < tiles) tile_sums[k] = carry + ex;
        carry += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) tile_sums[tiles] = carry;
}

static __global__ void __launch_bounds__(SCAN_THREADS)
    k_scan_apply(const uint32_t *__restrict__ hist, const uint32_t *__restrict__ tile_sums,
                 uint32_t tiles, uint32_t *__restrict__ offsets, uint32_t m) {
    __shared__ uint32_t ws[33];
    const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS], s = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) {
        v[j] = base + j < m ? hist[base + j] : 0;
        s += v[j];
    }
    uint32_t total;
    uint32_t run = tile_sums[blockIdx.x] + block_exclusive_scan(s, ws, total);
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) {
        if (base + j < m) {
            offsets[base + j] = run;
        }
        run += v[j];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) offsets[m] = tile_sums[tiles];
}

// ------------------------------------------------------------------------------------------
// item = index of the point to add: i for the plain MSM (pstride 1, wstride 0); the position of
// 2^(w c) P_i in a precomputed table otherwise (i * pstride + w * wstride).
template <class C>
__global__ void k_scatter(const int32_t *__restrict__ digits, const uint32_t *__restrict__ ranks,
                          const uint32_t *__restrict__ offsets, Entry *__restrict__ sorted,
                          uint32_t *__restrict__ seg_bucket, uint32_t n, uint32_t SEG,
                          uint32_t bstride, uint32_t pstride, uint32_t wstride) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t w = blockIdx.y;
    const int32_t d = i < n ? digits[(size_t)w * n + i] : 0;
    if (d == 0) return;
    const uint32_t b = w * bstride + (uint32_t)(d < 0 ? -d : d) - 1;
    const uint32_t pos = offsets[b] + ranks[(size_t)w * n + i];
    sorted[pos] = (i * pstride + w * wstride) | (d < 0 ? 0x80000000u : 0u);
    // every position is written exactly once, so exactly one thread sees each segment start
    if (pos % SEG == 0) seg_bucket[pos / SEG] = b;
}

// ------------------------------------------------------------------------------------------
// First accumulation round: thread t owns sorted[t*SEG, (t+1)*SEG) (SEG chosen by the host, a
// multiple of 4).  Bucket boundaries come from the offsets array, walked from the segment's
// first bucket (recorded by the scatter).
// MERGE: the buckets already hold the sums of earlier point ranges of the same call (one bucket
// set shared by all upload ranges of a host-buffer multi_exp).  `live[b] != 0` says bucket b has
// been written before; the thread that owns the head of a run starts from that value instead of
// the neutral element and marks the bucket live.  Exactly one thread per launch owns a bucket.
// L2 prefetch of one bucket accumulator (XW words; a prefetch never faults, so running past the
// last bucket is harmless)
template <int XW>
__device__ __forceinline__ void prefetch_bucket(const uint32_t *bucket_acc, uint32_t b) {
    const char *p = reinterpret_cast<const char *>(bucket_acc + (size_t)b * XW);
#pragma unroll
    for (int o = 0; o < XW * 4; o += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + o));
}

template <class C, bool MERGE>
__device__ __forceinline__ void accum_first_body(
    const Entry *__restrict__ sorted, const uint32_t *__restrict__ offsets,
    const uint32_t *__restrict__ seg_bucket, uint32_t total, uint32_t SEG, uint32_t t_begin,
    uint32_t t_end, const uint32_t *__restrict__ aff, uint32_t *__restrict__ bucket_acc,
    uint32_t *__restrict__ carry_key, uint32_t *__restrict__ carry_pt, uint32_t key_base,
    uint32_t *live) {
    using R = Raw<C>;
    using F = AccumField<typename C::Fq>;
    using X = XYZZ<F>;
    // segments [t_begin, t_end): one group of windows (the host walks the windows top-down in
    // groups so that the latency-bound reduction of one group runs under the next accumulation)
    const uint32_t t = t_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= t_end) return;
    const uint64_t start64 = (uint64_t)t * SEG;
    if (start64 >= total) return;
    const uint32_t start = (uint32_t)start64;
    const uint32_t end = (uint32_t)min((uint64_t)total, start64 + SEG);

    uint32_t cur = seg_bucket[t];
    uint32_t next = offsets[cur + 1];            // first position of the following bucket
    bool head_carry = start > offsets[cur];      // the bucket began in an earlier segment
    carry_key[t] = key_base + cur;  // buckets are local to the launch's window group, keys global
    bool carried = false;
    X acc = X::inf();
    if (MERGE) {
        // the thread walks consecutive buckets: pull the next accumulators towards the SM early
        prefetch_bucket<R::XW>(bucket_acc, cur + 1);
        prefetch_bucket<R::XW>(bucket_acc, cur + 2);
        if (!head_carry) {
            if (live[cur] != 0) acc = X::load(bucket_acc + (size_t)cur * R::XW);
            live[cur] = 1;
        }
    }
    uint4 q = *reinterpret_cast<const uint4 *>(sorted + start);  // 4 items per load
    Entry e = q.x;
    Affine<F> p = Affine<F>::load(aff + (size_t)(e & 0x7fffffffu) * R::AW);
    for (uint32_t pos = start; pos < end; pos++) {
        const Entry ecur = e;
        const Affine<F> pcur = p;
        if (pos + 1 < end) {  // prefetch the next item and its point
            const uint32_t k = (pos + 1 - start) & 3u;
            if (k == 0) q = *reinterpret_cast<const uint4 *>(sorted + pos + 1);
            e = k == 0 ? q.x : k == 1 ? q.y : k == 2 ? q.z : q.w;
            p = Affine<F>::load(aff + (size_t)(e & 0x7fffffffu) * R::AW);
        }
        if (pos == next) {  // the run of `cur` is complete; `pos` starts a later, non-empty bucket
            if (head_carry && !carried) {
                acc.store(carry_pt + (size_t)t * R::XW);
                carried = true;
            } else {
                acc.store(bucket_acc + (size_t)cur * R::XW);
            }
            head_carry = false;
            do {
                cur++;
                next = offsets[cur + 1];
            } while (next <= pos);  // skip empty buckets
            acc = X::inf();
            if (MERGE) {
                prefetch_bucket<R::XW>(bucket_acc, cur + 2);
                if (live[cur] != 0) acc = X::load(bucket_acc + (size_t)cur * R::XW);
                live[cur] = 1;
            }
        }
        X::madd(acc, pcur.x, F::cneg(pcur.y, (ecur >> 31) != 0));
    }
    if (head_carry && !carried) {
        acc.store(carry_pt + (size_t)t * R::XW);
        carried = true;
    } else {
        acc.store(bucket_acc + (size_t)cur * R::XW);
    }
    if (!carried) X::inf().store(carry_pt + (size_t)t * R::XW);
}

template <class C>
__global__ void __launch_bounds__(128)
    k_accum_first(const Entry *__restrict__ sorted, const uint32_t *__restrict__ offsets,
                  const uint32_t *__restrict__ seg_bucket, uint32_t total, uint32_t SEG,
                  uint32_t t_begin, uint32_t t_end, const uint32_t *__restrict__ aff,
                  uint32_t *__restrict__ bucket_acc, uint32_t *__restrict__ carry_key,
                  uint32_t *__restrict__ carry_pt, uint32_t key_base) {
    accum_first_body<C, false>(sorted, offsets, seg_bucket, total, SEG, t_begin, t_end, aff,
                               bucket_acc, carry_key, carry_pt, key_base, nullptr);
}

template <class C>
__global__ void __launch_bounds__(128, Raw<C>::XW <= 32 ? 4 : 1)
    k_accum_first_merge(const Entry *__restrict__ sorted, const uint32_t *__restrict__ offsets,
                        const uint32_t *__restrict__ seg_bucket, uint32_t total, uint32_t SEG,
                        uint32_t t_begin, uint32_t t_end, const uint32_t *__restrict__ aff,
                        uint32_t *__restrict__ bucket_acc, uint32_t *__restrict__ carry_key,
                        uint32_t *__restrict__ carry_pt, uint32_t key_base,
                        uint32_t *__restrict__ live) {
    accum_first_body<C, true>(sorted, offsets, seg_bucket, total, SEG, t_begin, t_end, aff,
                              bucket_acc, carry_key, carry_pt, key_base, live);
}

// ------------------------------------------------------------------------------------------
// First accumulation round of the WIDE prime fields (24 limbs: bw6_761) with the accumulator in
// SHARED MEMORY.  One thread per segment as above, but a 24-limb XYZZ accumulator (96 words), the
// point being added (48) and the temporaries of a mixed addition do not fit 255 registers: the
// register-resident kernel spills (628 bytes of spill stores per thread, 8 warps per SM, 63 % of
// the multiply pipe).  Here every thread owns six word-interleaved slots of shared memory
// (X, Y, ZZ, ZZZ and two temporaries; limb k of slot e at smem[(e N + k) blockDim + tid]: no bank
// conflicts) and every product streams its second operand from there one limb per row
// (Fp::mul_sb): a product keeps only its first operand and the two accumulators in registers.
// The next point is pulled towards the SM with an L2 prefetch instead of an early register load.
#ifndef B200_SM_THREADS
#define B200_SM_THREADS 64
#endif
#ifndef B200_SM_MINBLOCKS
#define B200_SM_MINBLOCKS 4
#endif
#ifndef B200_SM_MIN_LIMBS
#define B200_SM_MIN_LIMBS 24
#endif
#ifndef B200_SM_MINBLOCKS_12
#define B200_SM_MINBLOCKS_12 8
#endif
template <class C>
struct SmemAccum {
    static constexpr bool value =
        !BaseLimbs<typename C::Fq>::EXT && BaseLimbs<typename C::Fq>::N >= B200_SM_MIN_LIMBS;
    static constexpr int SLOTS = 8;
    static constexpr int THREADS = B200_SM_THREADS;
    static constexpr int MINBLOCKS = BaseLimbs<typename C::Fq>::N >= 24 ? B200_SM_MINBLOCKS : B200_SM_MINBLOCKS_12;
    static constexpr size_t BYTES = (size_t)SLOTS * BaseLimbs<typename C::Fq>::N * THREADS * 4;
};

// The accumulator and the temporaries of a mixed addition in word-interleaved shared-memory slots;
// the arithmetic is a short program of slot-to-slot operations, and every product of it runs
// through ONE out-of-line routine (uprod).  Measured on the way here (profiles/sm_accum_r02.md):
// a 24-limb row loop inlined at every product is 60-140 KB of code and the kernel starves on
// instruction fetch (sm__icc hit rate 68 %, "no instruction" 4.7 stall cycles per issue at 12
// warps); a called register-to-register product passes its operands through the stack.  uprod
// loads its multiplicands into registers, streams the other operands one limb per row from the
// slots and writes the result back, so nothing is live across a call, the kernel needs the
// registers of one product, and all warps of an SM loop in the same 18 KB of code.
template <class F>
struct SmemXYZZ {
    static constexpr int N = F::N;
    uint32_t *base;  // this thread's column of the block's slots
    int stride;      // words between consecutive limbs = threads per block
    bool inf;        // the accumulator is the neutral element (its slots are then meaningless)
    enum { SX = 0, SY = 1, SZZ = 2, SZZZ = 3, ST0 = 4, ST1 = 5, ST2 = 6, ST3 = 7, NONE = -1 };

    B200_DEV SmemXYZZ(uint32_t *smem, int tid, int threads) : base(smem + tid), stride(threads), inf(true) {}
    B200_DEV static uint32_t *slot_of(uint32_t *base, int stride, int e) { return base + (size_t)e * N * stride; }
    B200_DEV uint32_t *slot(int e) const { return slot_of(base, stride, e); }

    // dst <- a b [- c d] [- s]   (slot numbers; c = NONE: one product; s = NONE: no subtraction).
    // The two-product form shares one Montgomery reduction (Fp::mul2).
    __device__ __noinline__ static void uprod(uint32_t *base, int stride, int dst, int a, int b, int c,
                                              int d, int s) {
        const uint32_t *sb = slot_of(base, stride, b);
        const bool two = c != NONE;
        const uint32_t *sd = slot_of(base, stride, two ? d : b);
        const F A = F::lds(slot_of(base, stride, a), stride);
        F Cn = F::zero();
        if (two) Cn = F::neg(F::lds(slot_of(base, stride, c), stride));
        uint32_t E[N], O[N];
        uint32_t b0 = sb[0], b1 = sb[stride], d0 = sd[0], d1 = sd[stride];
        if (two) {
            F::mont_row2(E, O, A.v, b0, Cn.v, d0, true);
            F::mont_row2(O, E, A.v, b1, Cn.v, d1, false);
        } else {
            F::mont_row(E, O, A.v, b0, true);
            F::mont_row(O, E, A.v, b1, false);
        }
#pragma unroll 1
        for (int it = 1; it < N / 2; it++) {
            const size_t o = (size_t)(2 * it) * stride;
            b0 = sb[o], b1 = sb[o + stride];
            if (two) {
                d0 = sd[o], d1 = sd[o + stride];
                F::mont_row2(E, O, A.v, b0, Cn.v, d0, false);
                F::mont_row2(O, E, A.v, b1, Cn.v, d1, false);
            } else {
                F::mont_row(E, O, A.v, b0, false);
                F::mont_row(O, E, A.v, b1, false);
            }
        }
        F r = F::finish(E, O);
        if (s != NONE) r = F::sub(r, F::lds(slot_of(base, stride, s), stride));
        r.sts(slot_of(base, stride, dst), stride);
    }
    B200_DEV void prod(int dst, int a, int b, int s = NONE) const { uprod(base, stride, dst, a, b, NONE, NONE, s); }
    // Two independent products at once: d1 <- a1 b1 [- s1], d2 <- a2 b2 [- s2].  With two warps per
    // scheduler a single product's carry chains leave the multiply pipe waiting on their own
    // latency (ncu: 3.3 "wait" stall cycles per issue); the rows of two products interleave into
    // twice the independent chains.
    __device__ __noinline__ static void uprod_pair(uint32_t *base, int stride, int d1, int a1, int b1,
                                                   int s1, int d2, int a2, int b2, int s2) {
        const uint32_t *p1 = slot_of(base, stride, b1), *p2 = slot_of(base, stride, b2);
        const F A1 = F::lds(slot_of(base, stride, a1), stride), A2 = F::lds(slot_of(base, stride, a2), stride);
        uint32_t E1[N], O1[N], E2[N], O2[N];
        F::mont_row(E1, O1, A1.v, p1[0], true);
        F::mont_row(E2, O2, A2.v, p2[0], true);
        F::mont_row(O1, E1, A1.v, p1[stride], false);
        F::mont_row(O2, E2, A2.v, p2[stride], false);
#pragma unroll 1
        for (int it = 1; it < N / 2; it++) {
            const size_t o = (size_t)(2 * it) * stride;
            const uint32_t x0 = p1[o], x1 = p1[o + stride], y0 = p2[o], y1 = p2[o + stride];
            F::mont_row(E1, O1, A1.v, x0, false);
            F::mont_row(E2, O2, A2.v, y0, false);
            F::mont_row(O1, E1, A1.v, x1, false);
            F::mont_row(O2, E2, A2.v, y1, false);
        }
        F r = F::finish(E1, O1);
        if (s1 != NONE) r = F::sub(r, F::lds(slot_of(base, stride, s1), stride));
        r.sts(slot_of(base, stride, d1), stride);
        r = F::finish(E2, O2);
        if (s2 != NONE) r = F::sub(r, F::lds(slot_of(base, stride, s2), stride));
        r.sts(slot_of(base, stride, d2), stride);
    }
    B200_DEV void pair(int d1, int a1, int b1, int s1, int d2, int a2, int b2, int s2) const {
        uprod_pair(base, stride, d1, a1, b1, s1, d2, a2, b2, s2);
    }
    B200_DEV bool is_zero(int a) const {
        const uint32_t *p = slot(a);
        uint32_t o = 0;
#pragma unroll
        for (int i = 0; i < N; i++) o |= p[(size_t)i * stride];
        return o == 0;
    }

    // (out of line like the products: the entry function stays the segment walk alone)
    __device__ __noinline__ static void store_slots(uint32_t *base, int stride, uint32_t *dst, bool inf) {
        if (inf) {
            XYZZ<F>::inf().store(dst);
            return;
        }
#pragma unroll 1
        for (int e = 0; e < 4; e++) F::lds(slot_of(base, stride, e), stride).store(dst + e * N);
    }
    B200_DEV void store_global(uint32_t *dst) const { store_slots(base, stride, dst, inf); }
    // slots (ex, ey) <- (x2, +-y2): vector loads from global memory (with most of the SM's L1
    // carved out as shared memory, limb-wise reads of global memory would all go to the L2)
    __device__ __noinline__ static void stage_point(uint32_t *base, int stride, const uint32_t *gp,
                                                    bool negate, int ex, int ey) {
        F::load(gp).sts(slot_of(base, stride, ex), stride);
        F::cneg(F::load(gp + N), negate).sts(slot_of(base, stride, ey), stride);
    }
    B200_DEV void set_one(int e) const { F::one().sts(slot(e), stride); }

    // acc += (x2, +-y2), the affine point at gp (not infinity); madd-2008-s as XYZZ::madd
    B200_DEV void madd(const uint32_t *gp, bool negate) {
        if (inf) {
            stage_point(base, stride, gp, negate, SX, SY);
            set_one(SZZ), set_one(SZZZ);
            inf = false;
            return;
        }
        stage_point(base, stride, gp, negate, ST0, ST1);
        pair(ST0, SZZ, ST0, SX, ST1, SZZZ, ST1, SY);          // Pd = x2 ZZ - X, R = +-y2 ZZZ - Y
        if (is_zero(ST0)) {                                   // same x: P + P or P - P
            if (is_zero(ST1))
                set_double(gp, negate);
            else
                inf = true;
            return;
        }
        pair(ST2, ST0, ST0, NONE, ST3, ST1, ST1, NONE);       // PP = Pd^2, R^2
        pair(ST0, ST2, ST0, NONE, SZZ, SZZ, ST2, NONE);       // PPP = PP Pd, ZZ PP
        pair(SX, SX, ST2, NONE, SZZZ, SZZZ, ST0, NONE);       // Q = X PP, ZZZ PPP
        {
            const F Q = F::lds(slot(SX), stride);
            const F X3 = F::sub(F::sub(F::lds(slot(ST3), stride), F::lds(slot(ST0), stride)), F::dbl(Q));
            X3.sts(slot(SX), stride);                         // R^2 - PPP - 2 Q
            F::sub(Q, X3).sts(slot(ST2), stride);             // Q - X3
        }
        uprod(base, stride, SY, ST2, ST1, SY, ST0, NONE);     // Y3 = (Q - X3) R - Y PPP
    }
    // slots <- 2 (x2, +-y2): mdbl-2008-s-1 as XYZZ::dbl_affine, as a slot program
    B200_DEV void set_double(const uint32_t *gp, bool negate) {
        stage_point(base, stride, gp, negate, SX, SY);
        {
            const F y = F::lds(slot(SY), stride);
            F::dbl(y).sts(slot(ST0), stride);            // U = 2 y
        }
        pair(SZZ, ST0, ST0, NONE, ST2, SX, SX, NONE);      // V = U^2 = ZZ, x^2
        pair(SZZZ, ST0, SZZ, NONE, ST1, SX, SZZ, NONE);    // W = U V = ZZZ, S = x V
        {
            const F t = F::lds(slot(ST2), stride);
            t.sts(slot(ST0), stride);
        }
        {
            const F xx = F::lds(slot(ST0), stride);
            F::add(F::dbl(xx), xx).sts(slot(ST0), stride);   // M = 3 x^2
        }
        prod(SX, ST0, ST0);                 // M^2
        {
            const F S = F::lds(slot(ST1), stride);
            const F X3 = F::sub(F::lds(slot(SX), stride), F::dbl(S));
            X3.sts(slot(SX), stride);
            F::sub(S, X3).sts(slot(ST1), stride);        // S - X3
        }
        uprod(base, stride, SY, ST1, ST0, SZZZ, SY, NONE);  // Y3 = (S - X3) M - W y
    }
};

template <class C>
__global__ void __launch_bounds__(SmemAccum<C>::THREADS, SmemAccum<C>::MINBLOCKS)
    k_accum_first_sm(const Entry *__restrict__ sorted, const uint32_t *__restrict__ offsets,
                     const uint32_t *__restrict__ seg_bucket, uint32_t total, uint32_t SEG,
                     uint32_t t_begin, uint32_t t_end, const uint32_t *__restrict__ aff,
                     uint32_t *__restrict__ bucket_acc, uint32_t *__restrict__ carry_key,
                     uint32_t *__restrict__ carry_pt, uint32_t key_base) {
    using R = Raw<C>;
    using F = AccumField<typename C::Fq>;
    using X = XYZZ<F>;
    extern __shared__ uint32_t smem_acc[];
    const uint32_t t = t_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= t_end) return;
    const uint64_t start64 = (uint64_t)t * SEG;
    if (start64 >= total) return;
    const uint32_t start = (uint32_t)start64;
    const uint32_t end = (uint32_t)min((uint64_t)total, start64 + SEG);

    uint32_t cur = seg_bucket[t];
    uint32_t next = offsets[cur + 1];            // first position of the following bucket
    bool head_carry = start > offsets[cur];      // the bucket began in an earlier segment
    carry_key[t] = key_base + cur;
    bool carried = false;
    SmemXYZZ<F> acc(smem_acc, (int)threadIdx.x, (int)blockDim.x);
    uint4 q = *reinterpret_cast<const uint4 *>(sorted + start);  // 4 items per load
    Entry e = q.x;
    for (uint32_t pos = start; pos < end; pos++) {
        const Entry ecur = e;
        if (pos + 1 < end) {  // the next item; its point only as an L2 prefetch
            const uint32_t k = (pos + 1 - start) & 3u;
            if (k == 0) q = *reinterpret_cast<const uint4 *>(sorted + pos + 1);
            e = k == 0 ? q.x : k == 1 ? q.y : k == 2 ? q.z : q.w;
            const char *np = reinterpret_cast<const char *>(aff + (size_t)(e & 0x7fffffffu) * R::AW);
#pragma unroll
            for (int o = 0; o < R::AW * 4; o += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(np + o));
        }
        if (pos == next) {  // the run of `cur` is complete; `pos` starts a later, non-empty bucket
            if (head_carry && !carried) {
                acc.store_global(carry_pt + (size_t)t * R::XW);
                carried = true;
            } else {
                acc.store_global(bucket_acc + (size_t)cur * R::XW);
            }
            head_carry = false;
            do {
                cur++;
                next = offsets[cur + 1];
            } while (next <= pos);  // skip empty buckets
            acc.inf = true;
        }
        acc.madd(aff + (size_t)(ecur & 0x7fffffffu) * R::AW, (ecur >> 31) != 0);
    }
    if (head_carry && !carried) {
        acc.store_global(carry_pt + (size_t)t * R::XW);
        carried = true;
    } else {
        acc.store_global(bucket_acc + (size_t)cur * R::XW);
    }
    if (!carried) X::inf().store(carry_pt + (size_t)t * R::XW);
}

// Later rounds: the carry list (keys ascending, one XYZZ point each) is reduced the same way with
// full additions; sums of runs that begin a bucket in this list are added into bucket_acc.
// The list is entries [begin, end) of keys / pts; only keys in [key_lo, key_hi) are live (a group
// of windows shares its two boundary entries with its neighbours: each is live in exactly one).
constexpr uint32_t CARRY_DEAD = 0xfffffffeu;
template <class C>
__global__ void __launch_bounds__(128)
    k_accum_carry(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ pts,
                  uint32_t begin, uint32_t end, uint32_t key_lo, uint32_t key_hi, uint32_t SEG,
                  uint32_t *__restrict__ bucket_acc, uint32_t *__restrict__ carry_key,
                  uint32_t *__restrict__ carry_pt) {
    using R = Raw<C>;
    using F = TailField<typename C::Fq>;
    using X = XYZZ<F>;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t start64 = (uint64_t)begin + (uint64_t)t * SEG;
    if (start64 >= end) return;
    const uint32_t start = (uint32_t)start64;
    const uint32_t stop = (uint32_t)min((uint64_t)end, start64 + SEG);
    auto key_at = [&](uint32_t pos) {
        const uint32_t k = keys[pos];
        return (k >= key_lo && k < key_hi) ? k : CARRY_DEAD;
    };
    uint32_t cur = key_at(start);
    bool head_carry = cur != CARRY_DEAD && start > begin && key_at(start - 1) == cur;
    bool carried = false;
    carry_key[t] = cur;
    X acc = X::inf();
    for (uint32_t pos = start; pos <= stop; pos++) {
        uint32_t k = pos < stop ? key_at(pos) : 0xffffffffu;
        if (k != cur) {
            if (head_carry && !carried) {
                acc.store(carry_pt + (size_t)t * R::XW);
                carried = true;
            } else if (cur != CARRY_DEAD && !acc.is_inf()) {
                X b = X::load(bucket_acc + (size_t)cur * R::XW);
                X::add(b, acc);
                b.store(bucket_acc + (size_t)cur * R::XW);
            }
            head_carry = false;
            cur = k;
            acc = X::inf();
        }
        if (pos < stop && k != CARRY_DEAD) {
            X q = X::load(pts + (size_t)pos * R::XW);
            X::add(acc, q);
        }
    }
    if (!carried) X::inf().store(carry_pt + (size_t)t * R::XW);
}

// ------------------------------------------------------------------------------------------
// Bucket reduction in two levels of running sums (multiexp.tcc:90-125, 354-376 is the serial
// form).  With chunks of K buckets and (b + 1) = (b - jK + 1) + jK:
//   window sum = sum_j L_j + K * sum_j j T_j,   L_j = sum_{b in chunk j} (b - jK + 1) bucket[b],
//                                               T_j = sum_{b in chunk j} bucket[b].
// Level 1 (k_bucket_reduce): thread (w, j) walks its K buckets from the top; `running` ends as
// T_j and `acc` as L_j -- two additions per bucket, no scalar multiplication.
// B = buckets per window (a power of two), K a power of two <= B; grid.y = windows.
template <class C>
__global__ void __launch_bounds__(64)
    k_bucket_reduce(const uint32_t *__restrict__ bucket_acc, const uint32_t *__restrict__ hist,
                    uint32_t B, uint32_t K, uint32_t w0, uint32_t *__restrict__ chunk_L,
                    uint32_t *__restrict__ chunk_T) {
    using R = Raw<C>;
    using F = TailField<typename C::Fq>;
    using X = XYZZ<F>;
    const uint32_t chunks = B / K;
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t w = w0 + blockIdx.y;
    if (j >= chunks) return;
    const uint32_t lo = j * K, hi = lo + K;
    X running = X::inf(), acc = X::inf();
    if constexpr (!BaseLimbs<typename C::Fq>::EXT && BaseLimbs<typename C::Fq>::N <= 8) {
        // Short prime fields: ONE inlined addition site serves both steps of a bucket (running +=
        // bucket, acc += running), its operands picked by selects.  The out-of-line addition takes
        // its points from local memory; here they stay in registers.
        for (uint32_t b = hi; b-- > lo;) {
            const size_t g = (size_t)w * B + b;
            X q = X::inf();
            if (hist[g] != 0) q = X::load(bucket_acc + g * R::XW);  // empty buckets were never written
#pragma unroll 1
            for (int step = 0; step < 2; step++) {
                X dst = X::select(step == 0, running, acc);
                const X src = X::select(step == 0, q, running);
                X::add_inline(dst, src);
                if (step == 0)
                    running = dst;
                else
                    acc = dst;
            }
        }
    } else {
        for (uint32_t b = hi; b-- > lo;) {
            const size_t g = (size_t)w * B + b;
            if (hist[g] != 0) {  // empty buckets were never written
                X q = X::load(bucket_acc + g * R::XW);
                X::add(running, q);
            }
            X::add(acc, running);
        }
    }
    acc.store(chunk_L + ((size_t)w * chunks + j) * R::XW);
    running.store(chunk_T + ((size_t)w * chunks + j) * R::XW);
}

// Level 2, one block per window: thread t owns Q consecutive chunks (Q a power of two) and the
// same trick gives  sum_j j T_j = sum_t A_t + Q * sum_t t R_t  with  R_t = sum_j T_j and
// A_t = sum_j (j - tQ) T_j over its chunks.  The remaining weights t < THREADS are applied with a
// short double-and-add in parallel; K and Q are powers of two, i.e. doublings.  A shared-memory
// tree then adds the THREADS partial results.
// A window may be spread over the CTAs of a thread-block CLUSTER (launch attribute; CL = 1 for a
// plain launch): 8 warps of dependent product chains on ONE SM queue for its multiply pipe while
// the other SMs idle, so CL x THREADS threads on CL SMs walk shorter chains without sharing a
// pipe, and CTA 0 collects the CL partial results through distributed shared memory.
template <class C, int THREADS>
__global__ void __launch_bounds__(THREADS)
    k_window_reduce(const uint32_t *__restrict__ chunk_L, const uint32_t *__restrict__ chunk_T,
                    uint32_t chunks, uint32_t log2K, uint32_t w0,
                    uint32_t *__restrict__ window_sums, uint32_t *__restrict__ window_totals) {
    using R = Raw<C>;
    using F = TailField<typename C::Fq>;
    using X = XYZZ<F>;
    extern __shared__ uint32_t smem[];
    cooperative_groups::cluster_group cluster = cooperative_groups::this_cluster();
    const uint32_t CL = cluster.num_blocks(), rank = cluster.block_rank();
    const uint32_t lt = threadIdx.x;                  // thread inside its CTA
    const uint32_t w = w0 + blockIdx.x / CL, t = rank * THREADS + lt;  // thread inside the window
    const uint32_t TT = (uint32_t)THREADS * CL;
    const uint32_t Q = chunks > TT ? chunks / TT : 1u;
    const uint32_t log2Q = 31 - __clz(Q);
    X total = X::inf(), plain = X::inf();
    if (t * Q < chunks) {
        X sumL = X::inf(), run = X::inf(), acc = X::inf();
        if constexpr (!BaseLimbs<typename C::Fq>::EXT && BaseLimbs<typename C::Fq>::N <= 8) {
            // one inlined addition site for the three steps of a chunk (see k_bucket_reduce)
            for (uint32_t j = (t + 1) * Q; j-- > t * Q;) {
                const size_t g = ((size_t)w * chunks + j) * R::XW;
                const X qT = X::load(chunk_T + g), qL = X::load(chunk_L + g);
#pragma unroll 1
                for (int step = 0; step < 3; step++) {
                    // acc += run;  run += T_j;  sumL += L_j
                    X dst = step == 0 ? acc : step == 1 ? run : sumL;
                    const X src = step == 0 ? run : step == 1 ? qT : qL;
                    X::add_inline(dst, src);
                    if (step == 0)
                        acc = dst;
                    else if (step == 1)
                        run = dst;
                    else
                        sumL = dst;
                }
            }
        } else {
            for (uint32_t j = (t + 1) * Q; j-- > t * Q;) {
                const size_t g = ((size_t)w * chunks + j) * R::XW;
                X::add(acc, run);  // T_j is counted (j - tQ) times: once per later step
                X q = X::load(chunk_T + g);
                X::add(run, q);
                q = X::load(chunk_L + g);
                X::add(sumL, q);
            }
        }
        plain = run;
        X y = small_mul<F>(run, t);
        for (uint32_t k = 0; k < log2Q; k++) y = X::dbl(y);
        X::add(y, acc);
        for (uint32_t k = 0; k < log2K; k++) y = X::dbl(y);
        total = sumL;
        X::add(total, y);
    }
    // sum over the threads of the CTA (shared-memory tree), then over the CTAs of the cluster
    auto reduce = [&](X &val) {
        val.store(smem + (size_t)lt * R::XW);
        __syncthreads();
        for (uint32_t d = THREADS / 2; d > 0; d >>= 1) {
            if (lt < d) {
                X q = X::load(smem + (size_t)(lt + d) * R::XW);
                X::add(val, q);
                val.store(smem + (size_t)lt * R::XW);
            }
            __syncthreads();
        }
        if (CL > 1) {
            cluster.sync();  // slot 0 of every CTA holds that CTA's partial result
            if (rank == 0 && lt > 0 && lt < CL) {
                val = X::load(cluster.map_shared_rank(smem, lt));
                val.store(smem + (size_t)lt * R::XW);
            }
            __syncthreads();
            for (uint32_t d = CL / 2; d > 0; d >>= 1) {
                if (rank == 0 && lt < d) {
                    X q = X::load(smem + (size_t)(lt + d) * R::XW);
                    X::add(val, q);
                    val.store(smem + (size_t)lt * R::XW);
                }
                __syncthreads();
            }
            cluster.sync();  // remote shared memory must outlive the reads above
        }
    };
    reduce(total);
    if (rank == 0 && lt == 0) total.store(window_sums + (size_t)w * R::XW);
    if (window_totals == nullptr) return;  // uniform across the cluster
    // the unweighted sum of the window's buckets (used when one bucket set is cut into slices)
    __syncthreads();
    reduce(plain);
    if (rank == 0 && lt == 0) plain.store(window_totals + (size_t)w * R::XW);
}

// One doubling shared by the lanes of a warp.  The c doublings between windows are a serial
// chain (254..377 in total) that one thread walks at one multiplication latency after the other;
// the nine products of dbl-2008-s-1 have depth three, so lanes 0..3 take one product each per
// level and exchange results through shared memory: 3 product latencies per doubling instead
// of 9.  Every lane holds the full point before and after.  `sh` has 4 * F::N words.
//
// Quadratic extensions go one step further (dbl_coop_ext / add_coop_ext): a product over Fp2 is
// three independent base-field products (Karatsuba), so lanes 3g .. 3g+2 take one each for
// product g and a level costs ONE base-field product latency instead of three.  `sh` then holds
// 12 base-field elements = 6 * F::N words.
template <class F2>
__device__ __forceinline__ void ext_products(const F2 &a, const F2 &b, uint32_t *sh, uint32_t lane) {
    using B = typename F2::F;
    constexpr int BN = F2::N / 2;
    const uint32_t part = lane % 3u;
    const B x = part == 0 ? a.c0 : part == 1 ? a.c1 : B::add(a.c0, a.c1);
    const B y = part == 0 ? b.c0 : part == 1 ? b.c1 : B::add(b.c0, b.c1);
    const B v = B::mul(x, y);
    if (lane < 12) v.store(sh + lane * BN);
    __syncwarp();
}
// product g of the last ext_products call: (v0 - NR v1) + (s - v0 - v1) u
template <class F2>
__device__ __forceinline__ F2 ext_result(const uint32_t *sh, int g) {
    using B = typename F2::F;
    constexpr int BN = F2::N / 2;
    const B v0 = B::load(sh + (3 * g) * BN), v1 = B::load(sh + (3 * g + 1) * BN),
            s = B::load(sh + (3 * g + 2) * BN);
    return F2{B::sub(v0, F2::times_nr(v1)), B::sub(B::sub(s, v0), v1)};
}

template <class F>
__device__ __forceinline__ XYZZ<F> dbl_coop_ext(const XYZZ<F> &p, uint32_t *sh, uint32_t lane) {
    const uint32_t g = (lane / 3u) & 3u;  // product this lane works on
    const F U = F::dbl(p.Y);
    // level 1: V = U^2, xx = X^2
    ext_products<F>(F::select(g == 1, p.X, U), F::select(g == 1, p.X, U), sh, lane);
    const F V = ext_result<F>(sh, 0), xx = ext_result<F>(sh, 1);
    __syncwarp();
    const F M = F::add(F::dbl(xx), xx);
    // level 2: W = U V, S = X V, M^2, ZZ' = V ZZ
    ext_products<F>(g == 0 ? U : g == 1 ? p.X : g == 2 ? M : V,
                    g == 0 ? V : g == 1 ? V : g == 2 ? M : p.ZZ, sh, lane);
    const F Wv = ext_result<F>(sh, 0), S = ext_result<F>(sh, 1), MM = ext_result<F>(sh, 2),
            ZZ3 = ext_result<F>(sh, 3);
    __syncwarp();
    const F X3 = F::sub(MM, F::dbl(S));
    // level 3: M (S - X3), W Y, ZZZ' = W ZZZ
    ext_products<F>(g == 0 ? M : Wv, g == 0 ? F::sub(S, X3) : g == 1 ? p.Y : p.ZZZ, sh, lane);
    const F Y3 = F::sub(ext_result<F>(sh, 0), ext_result<F>(sh, 1));
    const F ZZZ3 = ext_result<F>(sh, 2);
    __syncwarp();
    return XYZZ<F>{X3, Y3, ZZ3, ZZZ3};
}

// acc += q over a quadratic extension (add-2008-s in four levels, see add_coop_wide); every lane
// holds both points; exceptional cases are uniform across the warp and take the one-lane routine
template <class F>
__device__ __forceinline__ void add_coop_ext(XYZZ<F> &acc, const XYZZ<F> &q, uint32_t *sh,
                                             uint32_t lane) {
    if (q.is_inf()) return;
    if (acc.is_inf()) {
        acc = q;
        return;
    }
    const uint32_t g = (lane / 3u) & 3u;
    // level 1: U1 = X1 ZZ2, S1 = Y1 ZZZ2, U2 = X2 ZZ1, S2 = Y2 ZZZ1
    ext_products<F>(g == 0 ? acc.X : g == 1 ? acc.Y : g == 2 ? q.X : q.Y,
                    g == 0 ? q.ZZ : g == 1 ? q.ZZZ : g == 2 ? acc.ZZ : acc.ZZZ, sh, lane);
    const F U1 = ext_result<F>(sh, 0), S1 = ext_result<F>(sh, 1);
    const F Pd = F::sub(ext_result<F>(sh, 2), U1), R = F::sub(ext_result<F>(sh, 3), S1);
    __syncwarp();
    if (Pd.is_zero()) {  // uniform
        XYZZ<F>::add(acc, q);
        return;
    }
    // level 2: PP = P^2, RR = R^2, ZZ1 ZZ2, ZZZ1 ZZZ2
    ext_products<F>(g == 0 ? Pd : g == 1 ? R : g == 2 ? acc.ZZ : acc.ZZZ,
                    g == 0 ? Pd : g == 1 ? R : g == 2 ? q.ZZ : q.ZZZ, sh, lane);
    const F PP = ext_result<F>(sh, 0), RR = ext_result<F>(sh, 1), ZZ12 = ext_result<F>(sh, 2),
            ZZZ12 = ext_result<F>(sh, 3);
    __syncwarp();
    // level 3: PPP = P PP, Q = U1 PP, ZZ3 = ZZ12 PP
    ext_products<F>(g == 0 ? Pd : g == 1 ? U1 : ZZ12, PP, sh, lane);
    const F PPP = ext_result<F>(sh, 0), Q = ext_result<F>(sh, 1), ZZ3 = ext_result<F>(sh, 2);
    __syncwarp();
    const F X3 = F::sub(F::sub(RR, PPP), F::dbl(Q));
    // level 4: R (Q - X3), S1 PPP, ZZZ3 = ZZZ12 PPP
    ext_products<F>(g == 0 ? R : g == 1 ? S1 : ZZZ12, g == 0 ? F::sub(Q, X3) : PPP, sh, lane);
    acc.X = X3;
    acc.Y = F::sub(ext_result<F>(sh, 0), ext_result<F>(sh, 1));
    acc.ZZ = ZZ3;
    acc.ZZZ = ext_result<F>(sh, 2);
    __syncwarp();
}

template <class F>
__device__ __forceinline__ XYZZ<F> dbl_coop(const XYZZ<F> &p, uint32_t *sh, uint32_t lane) {
    if constexpr (BaseLimbs<F>::EXT) return dbl_coop_ext<F>(p, sh, lane);
    const uint32_t slot = (lane & 3u) * F::N;
    const F U = F::dbl(p.Y);
    // level 1: V = U^2, xx = X^2
    F t = F::mul(F::select(lane == 1, p.X, U), F::select(lane == 1, p.X, U));
    __syncwarp();
    if (lane < 4) t.store(sh + slot);
    __syncwarp();
    const F V = F::load(sh), xx = F::load(sh + F::N);
    const F M = F::add(F::dbl(xx), xx);
    // level 2: W = U V, S = X V, M^2, ZZ' = V ZZ
    t = F::mul(lane == 0 ? U : lane == 1 ? p.X : lane == 2 ? M : V,
               lane == 0 ? V : lane == 1 ? V : lane == 2 ? M : p.ZZ);
    __syncwarp();
    if (lane < 4) t.store(sh + slot);
    __syncwarp();
    const F Wv = F::load(sh), S = F::load(sh + F::N), MM = F::load(sh + 2 * F::N),
            ZZ3 = F::load(sh + 3 * F::N);
    const F X3 = F::sub(MM, F::dbl(S));
    // level 3: M (S - X3), W Y, ZZZ' = W ZZZ
    t = F::mul(lane == 0 ? M : Wv, lane == 0 ? F::sub(S, X3) : lane == 1 ? p.Y : p.ZZZ);
    __syncwarp();
    if (lane < 4) t.store(sh + slot);
    __syncwarp();
    const F Y3 = F::sub(F::load(sh), F::load(sh + F::N));
    return XYZZ<F>{X3, Y3, ZZ3, F::load(sh + 2 * F::N)};
}

// The same doubling for the wide prime fields, every product spread over the lanes of a group
// (fp_coop.cuh): groups 0..3 take the up to four products of a level.  `sh` has 4 * F::N words
// for the results, `win` CoopField<F>::WIN_WORDS for the products' accumulator windows.
template <class F>
__device__ __forceinline__ XYZZ<F> dbl_coop_wide(const XYZZ<F> &p, uint32_t *sh, uint32_t *win,
                                                 uint32_t lane) {
    using CF = CoopField<F>;
    const uint32_t g = (lane / CF::T) & 3u;            // product this lane works on
    const bool writer = lane % CF::T == 0 && lane < 4 * CF::T;
    const F U = F::dbl(p.Y);
    // level 1: V = U^2, xx = X^2
    F t = CF::mul(F::select(g == 1, p.X, U), F::select(g == 1, p.X, U), win, lane);
    if (writer) t.store(sh + g * F::N);
    __syncwarp();
    const F V = F::load(sh), xx = F::load(sh + F::N);
    __syncwarp();
    const F M = F::add(F::dbl(xx), xx);
    // level 2: W = U V, S = X V, M^2, ZZ' = V ZZ
    t = CF::mul(g == 0 ? U : g == 1 ? p.X : g == 2 ? M : V, g == 0 ? V : g == 1 ? V : g == 2 ? M : p.ZZ,
                win, lane);
    if (writer) t.store(sh + g * F::N);
    __syncwarp();
    const F Wv = F::load(sh), S = F::load(sh + F::N), MM = F::load(sh + 2 * F::N),
            ZZ3 = F::load(sh + 3 * F::N);
    __syncwarp();
    const F X3 = F::sub(MM, F::dbl(S));
    // level 3: M (S - X3), W Y, ZZZ' = W ZZZ
    t = CF::mul(g == 0 ? M : Wv, g == 0 ? F::sub(S, X3) : g == 1 ? p.Y : p.ZZZ, win, lane);
    if (writer) t.store(sh + g * F::N);
    __syncwarp();
    const F Y3 = F::sub(F::load(sh), F::load(sh + F::N));
    const F ZZZ3 = F::load(sh + 2 * F::N);
    __syncwarp();
    return XYZZ<F>{X3, Y3, ZZ3, ZZZ3};
}

// acc += q (add-2008-s, as XYZZ::add) with the 14 products in four levels of up to four
// lane-parallel products.  Every lane holds both points; exceptional cases (an operand at
// infinity, equal x) are uniform across the warp and take the one-lane routine.
template <class F>
__device__ __forceinline__ void add_coop_wide(XYZZ<F> &acc, const XYZZ<F> &q, uint32_t *sh,
                                              uint32_t *win, uint32_t lane) {
    using CF = CoopField<F>;
    if (q.is_inf()) return;
    if (acc.is_inf()) {
        acc = q;
        return;
    }
    const uint32_t g = (lane / CF::T) & 3u;
    const bool writer = lane % CF::T == 0 && lane < 4 * CF::T;
    // level 1: U1 = X1 ZZ2, S1 = Y1 ZZZ2, U2 = X2 ZZ1, S2 = Y2 ZZZ1
    F t = CF::mul(g == 0 ? acc.X : g == 1 ? acc.Y : g == 2 ? q.X : q.Y,
                  g == 0 ? q.ZZ : g == 1 ? q.ZZZ : g == 2 ? acc.ZZ : acc.ZZZ, win, lane);
    if (writer) t.store(sh + g * F::N);
    __syncwarp();
    const F U1 = F::load(sh), S1 = F::load(sh + F::N);
    const F Pd = F::sub(F::load(sh + 2 * F::N), U1), R = F::sub(F::load(sh + 3 * F::N), S1);
    __syncwarp();
    if (Pd.is_zero()) {  // uniform
        XYZZ<F>::add(acc, q);
        return;
    }
    // level 2: PP = P^2, RR = R^2, ZZ1 ZZ2, ZZZ1 ZZZ2
    t = CF::mul(g == 0 ? Pd : g == 1 ? R : g == 2 ? acc.ZZ : acc.ZZZ,
                g == 0 ? Pd : g == 1 ? R : g == 2 ? q.ZZ : q.ZZZ, win, lane);
    if (writer) t.store(sh + g * F::N);
    __syncwarp();
    const F PP = F::load(sh), RR = F::load(sh + F::N), ZZ12 = F::load(sh + 2 * F::N),
            ZZZ12 = F::load(sh + 3 * F::N);
    __syncwarp();
    // level 3: PPP = P PP, Q = U1 PP, ZZ3 = ZZ12 PP
    t = CF::mul(g == 0 ? Pd : g == 1 ? U1 : ZZ12, PP, win, lane);
    if (writer) t.store(sh + g * F::N);
    __syncwarp();
    const F PPP = F::load(sh), Q = F::load(sh + F::N), ZZ3 = F::load(sh + 2 * F::N);
    __syncwarp();
    const F X3 = F::sub(F::sub(RR, PPP), F::dbl(Q));
    // level 4: R (Q - X3), S1 PPP, ZZZ3 = ZZZ12 PPP
    t = CF::mul(g == 0 ? R : g == 1 ? S1 : ZZZ12, g == 0 ? F::sub(Q, X3) : PPP, win, lane);
    if (writer) t.store(sh + g * F::N);
    __syncwarp();
    acc.X = X3;
    acc.Y = F::sub(F::load(sh), F::load(sh + F::N));
    acc.ZZ = ZZ3;
    acc.ZZZ = F::load(sh + 2 * F::N);
    __syncwarp();
}

// result = sum_w 2^(w c) window_sums[w] (Horner from the top window, multiexp.tcc:309-314),
// written in the reference's coordinate system.  One warp; see dbl_coop.  The chain may be walked
// in steps, one per group of windows [w_lo, w_hi) from the top group down: `state` carries the
// running value between steps (read unless `first`), out_raw is written when `last`.
template <class C>
__global__ void __launch_bounds__(32) k_final(const uint32_t *__restrict__ window_sums, int w_lo, int w_hi,
                                              int c, uint32_t *__restrict__ state, bool first, bool last,
                                              uint32_t *__restrict__ out_raw) {
    using R = Raw<C>;
    using F = typename C::Fq;
    using X = XYZZ<F>;
    __shared__ __align__(16) uint32_t sh[6 * F::N];
    constexpr bool WIDE = CoopField<F>::OK;
    __shared__ uint32_t win[CoopField<F>::WIN_WORDS];
    if (blockIdx.x != 0) return;
    const uint32_t lane = threadIdx.x & 31;
    X r = first ? X::inf() : X::load(state);
    for (int w = w_hi - 1; w >= w_lo; w--) {
        if (!r.is_inf()) {  // uniform across the warp: every lane holds the same point
            for (int k = 0; k < c; k++) {
                if constexpr (WIDE)
                    r = dbl_coop_wide<F>(r, sh, win, lane);
                else
                    r = dbl_coop<F>(r, sh, lane);
            }
        }
        X q = X::load(window_sums + (size_t)w * R::XW);
        if constexpr (WIDE)
            add_coop_wide<F>(r, q, sh, win, lane);
        else if constexpr (BaseLimbs<F>::EXT)
            add_coop_ext<F>(r, q, sh, lane);
        else
            X::add(r, q);
    }
    if (lane == 0) {
        if (!last) r.store(state);
        if (last) {
            F Xo, Yo, Zo;
            R::from_xyzz(r, Xo, Yo, Zo);
            R::store(out_raw, Xo, Yo, Zo);
        }
    }
}

// Single bucket set (precomputed multiples) cut into V slices of 2^log2S consecutive buckets that
// went through the reduction like V windows: slice g contributed S_g = sum (b_local + 1) bucket and
// T_g = sum bucket, and the set's value is  sum_g S_g + 2^log2S * sum_g g T_g  (running sums from
// the top slice, then log2S doublings).
template <class C>
__global__ void __launch_bounds__(32)
    k_final_slices(const uint32_t *__restrict__ slice_sums, const uint32_t *__restrict__ slice_totals,
                   int V, int log2S, uint32_t *__restrict__ out_raw) {
    using R = Raw<C>;
    using F = typename C::Fq;
    using X = XYZZ<F>;
    __shared__ __align__(16) uint32_t sh[6 * F::N];
    if (blockIdx.x != 0) return;
    const uint32_t lane = threadIdx.x & 31;
    X run = X::inf(), acc = X::inf(), sumS = X::inf();
    for (int g = V - 1; g >= 0; g--) {
        X q = X::load(slice_sums + (size_t)g * R::XW);
        X::add(sumS, q);
        if (g > 0) {
            q = X::load(slice_totals + (size_t)g * R::XW);
            X::add(run, q);
            X::add(acc, run);
        }
    }
    if (!acc.is_inf())  // uniform across the warp
        for (int k = 0; k < log2S; k++) acc = dbl_coop<F>(acc, sh, lane);
    X::add(acc, sumS);
    if (lane == 0) {
        F Xo, Yo, Zo;
        R::from_xyzz(acc, Xo, Yo, Zo);
        R::store(out_raw, Xo, Yo, Zo);
    }
}

// One step of a table build: state[i] <- 2^c state[i] (state = the base itself when `first`),
// also written in the reference's layout for the batch normalisation that makes it affine.
template <class C>
__global__ void __launch_bounds__(128)
    k_pre_step(const uint32_t *__restrict__ aff, const uint8_t *__restrict__ inf_flag,
               uint32_t *__restrict__ state, uint32_t *__restrict__ raw, uint32_t n, int c,
               bool first) {
    using R = Raw<C>;
    using F = typename C::Fq;
    using X = XYZZ<F>;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    X p;
    if (first) {
        p = inf_flag[i] ? X::inf() : X::from_affine(Affine<F>::load(aff + (size_t)i * R::AW));
    } else {
        p = X::load(state + (size_t)i * R::XW);
    }
    for (int k = 0; k < c; k++) p = X::dbl(p);
    p.store(state + (size_t)i * R::XW);
    F Xo, Yo, Zo;
    R::from_xyzz(p, Xo, Yo, Zo);
    R::store(raw + (size_t)i * R::PW, Xo, Yo, Zo);
}

// Sum of `count` reference-layout group elements (partials of the GPU shards, multiexp.tcc:681-685)
template <class C>
__global__ void k_sum_raw(const uint32_t *__restrict__ raw, uint32_t count,
                          uint32_t *__restrict__ out_raw) {
    using R = Raw<C>;
    using F = typename C::Fq;
    using X = XYZZ<F>;
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    X r = X::inf();
    for (uint32_t i = 0; i < count; i++) {
        F Xi, Yi, Zi;
        R::load(raw + (size_t)i * R::PW, Xi, Yi, Zi);
        X q = R::to_xyzz(Xi, Yi, Zi);
        X::add(r, q);
    }
    F Xo, Yo, Zo;
    R::from_xyzz(r, Xo, Yo, Zo);
    R::store(out_raw, Xo, Yo, Zo);
}

}  // namespace b200msm
