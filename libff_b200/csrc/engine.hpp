// Host-side engine state shared by the per-group translation units (no templates here).
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

namespace b200msm {

enum Phase {
    PH_INGEST = 0,
    PH_DIGITS,
    PH_SCAN,
    PH_SCATTER,
    PH_ACCUM,
    PH_CARRY,
    PH_BUCKET_REDUCE,
    PH_FINAL,
    PH_TOTAL,
    PH_COUNT
};

void set_error(const char *fmt, ...);

#define B200_CUDA(call)                                                                     \
    do {                                                                                    \
        cudaError_t e_ = (call);                                                            \
        if (e_ != cudaSuccess) {                                                            \
            ::b200msm::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),    \
                                 __FILE__, __LINE__);                                       \
            return false;                                                                   \
        }                                                                                   \
    } while (0)

// Grow-only device scratch slots: the pipeline's buffers are reused across calls so that a
// steady-state multi_exp performs no cudaMalloc.
enum Slot {
    SL_RAW = 0,     // uploaded reference-layout bases
    SL_AFF,         // device affine bases
    SL_INF,         // infinity flags
    SL_SCALARS,     // uploaded scalars
    SL_DIGITS,
    SL_RANKS,       // rank of every item inside its bucket (the histogram counter it saw)
    SL_HIST,
    SL_OFFSETS,
    SL_CURSOR,
    SL_TILE_SUMS,
    SL_SORTED,
    SL_SEG_BUCKET,
    SL_BUCKETS,
    SL_CARRY_KEY_A,
    SL_CARRY_PT_A,
    SL_CARRY_KEY_B,
    SL_CARRY_PT_B,
    SL_CARRY_KEY_C,
    SL_CARRY_PT_C,
    SL_HORNER,      // running value of the Horner chain between window groups
    SL_CHUNKS,
    SL_CHUNK_TOTALS,
    SL_WINDOWS,
    SL_OUT,
    SL_PARTIALS,    // per-range partial results of a pipelined host call
    SL_INV_VAL,
    SL_INV_PRE,
    SL_INV_PROD,
    SL_TABLE_RAW,
    SL_TABLE_AFF,
    SL_TABLE_INF,
    SL_BEXP_OUT,
    SL_WIRE_IN,
    SL_WIRE_SC,
    SL_WIRE_MISC,
    SL_PRE_STATE,   // XYZZ running multiples while a table is built
    SL_PRE_RAW,     // the same in reference layout, input of the batch normalisation
    SL_STREAM_IN,   // encoded chunk of a streamed MSM
    SL_LIVE,        // per bucket: written by an earlier point range of the same call
    SL_SUM_IN,      // partial points of the device shards, summed on the first device
    SL_DIGITS_B,    // second digit / rank set: the sort of one window group under the
    SL_RANKS_B,     //   accumulation of the previous one
    SL_SORTED_B,
    SL_SEG_BUCKET_B,
    SL_OFFSETS_B,
    SL_META,        // item count of a sort, highest window in use
    SL_COUNT
};

constexpr int MAX_H2D_PARTS = 16;

// Pinned staging ring of one engine context.  Pageable sources -- what a std::vector<G> of a libff
// caller is -- would be staged by the driver at ~10 GB/s; instead STAGE_LANES host threads copy
// slices into these buffers and issue the DMA from there, which keeps the PCIe link busy.
constexpr int STAGE_LANES_MAX = 16;
constexpr size_t STAGE_BYTES = (size_t)8 << 20;
struct Stager {
    void *buf[STAGE_LANES_MAX][2] = {};
    cudaEvent_t ev[STAGE_LANES_MAX][2] = {};
    int lanes = 0;  // lanes whose buffers exist
    bool init(int want_lanes);
    void release();
};

// One engine context per bound device (b200msm_init_devices): its streams, events, scratch pool
// and options.  Nothing in the pipelines is process-global, so several contexts -- one per GPU of
// the box -- run concurrently from their own host threads.
struct Ctx {
    int device = -1;
    int index = 0;               // position in the list of bound devices
    uint32_t attr_mask = 0;      // groups whose kernels have had their function attributes set
    Stager stager;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;          // host -> device uploads of a pipelined call
    cudaEvent_t ev_copy[MAX_H2D_PARTS] = {};     // "range k is on the device"
    cudaEvent_t ev_sync = nullptr;
    // The last two kernels of an MSM (per-window totals, Horner) are one block per window and one
    // warp: latency-bound.  In a pipelined host call they run on a high-priority side stream
    // under the next range's digit/scatter/accumulation kernels.
    cudaStream_t tail_stream = nullptr;
    cudaEvent_t ev_tail_go = nullptr, ev_tail_done = nullptr;
    // window groups of one MSM: consecutive accumulations alternate between `stream` and
    // `acc_stream`, the reductions that follow each of them go to `tail_stream`
    cudaStream_t acc_stream = nullptr;
    cudaEvent_t ev_group_go = nullptr, ev_group[2] = {};
    // the counting sort of window group g + 1 (L2 / atomic bound) runs here under the
    // accumulation of group g (multiply-pipe bound)
    cudaStream_t sort_stream = nullptr;
    cudaEvent_t ev_sort[4] = {};  // [p]: sort into buffer set p done; [2 + p]: its reader done
    int sort_parity = 0;          // buffer set the next sort writes
    uint32_t *host_meta = nullptr;   // pinned: item counts read back between the halves of a sort
    bool sort_inputs_ready = false;  // the caller queued the inputs' upload / ingest on sort_stream
    bool defer_tail = false;    // set by the pipelined caller
    // Point ranges of ONE host-buffer call that share a bucket set (set by the pipelined caller):
    // every range runs digits / sort / accumulation with the digit size of the whole call and adds
    // into the same buckets; only the last range reduces the buckets.  range_c == 0: off.
    int range_c = 0;
    bool range_first = false, range_last = false;
    bool tail_pending = false;  // a deferred tail may still be reading the reduction scratch
    void *slot_ptr[SL_COUNT] = {};
    size_t slot_bytes[SL_COUNT] = {};
    cudaEvent_t ev[PH_COUNT + 1] = {};
    cudaEvent_t ev_timer[2] = {};                // b200msm_timer_start / _stop
    float last_ms[PH_COUNT] = {};
    bool have_timings = false;
    uint64_t launches = 0;
    int opt_window_bits = 0;  // 0 = automatic
    int opt_segment = 0;      // 0 = default
    int opt_h2d_parts = 0;    // 0 = automatic
    int opt_h2d_shared_buckets = 1; // upload ranges of one call share a bucket set (else: one MSM each)
    int opt_trace = 0;        // print per-range phase times of a pipelined call (development aid)
    int opt_h2d_growth = 0;   // size ratio of consecutive ranges in percent; 0 = default (130)
    int opt_window_groups = 0; // groups the windows are walked in (top-down); 0 = automatic, 1 = one
    int opt_carry_segment = 0; // carries per thread in the later accumulation rounds; 0 = automatic
    int opt_reduce_chunk = 0; // log2 of the buckets per thread in the reduction; 0 = automatic
    int opt_reduce_cluster = 0; // CTAs (SMs) one window's totals are spread over; 0 = automatic (up to 8), 1 = one
    // special-form bases leave Z behind on upload: 0 never (default: measured, the host-side packing
    // loses to the plain copy on a 16-core host, profiles/pack_sweep_r02.log), 1 pageable sources, 2 always
    int opt_h2d_pack = 0;
    int opt_h2d_taper = 0;    // trailing upload ranges shrink again (steps of growth per range from the end); 0 = off
    int opt_stage_lanes = 8;  // host threads that stage pageable buffers through the pinned ring
    int opt_sort_overlap = 0; // window groups whose sort overlaps the previous accumulation; 0 = automatic
    int opt_coop_accum = 0;   // lane-cooperative accumulation of the wide fields: 0 automatic, 1 off, 2 on

    bool reserve(Slot s, size_t bytes, void **out);
    void release_all();
};

template <class T>
inline bool reserve_as(Ctx &ctx, Slot s, size_t bytes, T **out) {
    void *p = nullptr;
    if (!ctx.reserve(s, bytes, &p)) return false;
    *out = static_cast<T *>(p);
    return true;
}

// Shape of an MSM over a table of precomputed multiples 2^(w c) P_i (multiexp_stream.tcc:124-162):
// the digit of window w of scalar i selects table[i * pstride + w * wstride], and every window
// feeds the same 2^(c-1) buckets.  nullptr = the plain MSM (one bucket set per window).
struct MsmShape {
    int c;              // digit bits the table was built for
    int W;              // digits per scalar = multiples per base = ceil(scalar bits / c)
    uint32_t pstride;   // table index stride between bases
    uint32_t wstride;   // table index stride between multiples of one base
    // Where W c equals the scalar width the reference's field_get_signed_digits drops the carry out
    // of the top digit (field_utils.tcc:205-239).  false: such scalars are folded to r - s and the
    // result is the true sum; true: the carry is dropped as the reference does, bit for bit.
    bool drop_top_carry = false;
};

// Type-erased per-group entry points, implemented once per group in group_*.cu.
struct GroupOps {
    uint32_t id;
    size_t group_bytes, scalar_bytes, affine_bytes;
    int scalar_bits;
    // d_raw (reference layout, device) -> d_aff, d_inf.  form 0: any Z, 1: special (Z = 1 or the
    // zero element), 2: d_raw == d_aff already holds device-affine records (packed upload): flags only
    bool (*ingest)(Ctx &, const void *d_raw, size_t n, uint32_t form, void *d_aff, uint8_t *d_inf);
    // the MSM proper; result (reference layout) into d_out.  With a shape, d_inf holds one flag
    // per table ENTRY (index i * pstride + w * wstride), else one per base.
    bool (*msm)(Ctx &, const void *d_aff, const uint8_t *d_inf, const void *d_scalars, size_t n,
                void *d_out, const MsmShape *shape);
    // reserves the scratch of an MSM over n points (upper bounds), so that a pipelined call does
    // not regrow slots between its ranges
    bool (*msm_prepare)(Ctx &, size_t n, const MsmShape *shape);
    // table[w * n + i] = 2^(w c) P_i (device affine), w < W, from n device-affine bases;
    // d_table_inf[w * n + i] = 1 where that multiple is the zero element
    bool (*precompute)(Ctx &, const void *d_aff, const uint8_t *d_inf, size_t n, int c, int W,
                       void *d_table, uint8_t *d_table_inf);
    // libff's binary stream encoding (curve_serialization.tcc:79-101, uncompressed) -> device
    // affine + infinity flags; flags bit 0: Montgomery form (else plain)
    bool (*stream_decode)(Ctx &, const void *d_bytes, size_t n, uint32_t flags, void *d_aff,
                          uint8_t *d_inf);
    bool (*sum)(Ctx &, const void *d_raw, size_t count, void *d_out);
    bool (*to_special)(Ctx &, void *d_raw, size_t n);
    // fixed-base tables (multiexp.tcc:809-846): raw reference-layout entries -> device affine
    // ((0, 0) = zero element); the device build of get_window_table; affine -> special-form raw
    bool (*table_normalise)(Ctx &, const void *d_table_raw, size_t entries, void *d_table_aff);
    bool (*table_build)(Ctx &, const void *d_g_raw, int scalar_bits, int window, size_t rows,
                        size_t row_len, void *d_table_raw);
    bool (*table_export)(Ctx &, const void *d_table_aff, size_t entries, void *d_table_raw);
    // out[i] = windowed_exp(table, [coeff *] scalars[i]) (multiexp.tcc:848-947); d_coeff may be null
    bool (*batch_exp)(Ctx &, const void *d_table_aff, size_t rows, size_t row_len, int window,
                      const void *d_scalars, const void *d_coeff, size_t n, void *d_out);
    uint32_t (*window_bits)(const Ctx &, size_t n);
    // wire format of the reference's ffi layer (wire_kernels.cuh): decode + validate, encode
    bool (*wire_decode)(Ctx &, const void *d_wire_points, const void *d_wire_scalars, size_t n,
                        uint32_t flags, void *d_aff, uint8_t *d_inf, void *d_scalars, uint32_t *d_err);
    bool (*wire_encode)(Ctx &, const void *d_raw, void *d_wire_out);
    // element-wise leaf operation (debug_kernels.cuh); returns bytes per element via elem_bytes
    bool (*debug_op)(Ctx &, int op, const void *d_a, const void *d_b, void *d_out, size_t n);
    size_t (*debug_elem_bytes)(int op);
};

const GroupOps *ops_alt_bn128_g1();
const GroupOps *ops_alt_bn128_g2();
const GroupOps *ops_bls12_377_g1();
const GroupOps *ops_bw6_761_g1();
const GroupOps *ops_bls12_377_g2();
const GroupOps *ops_bw6_761_g2();
const GroupOps *ops_bls12_381_g1();
const GroupOps *ops_bls12_381_g2();

}  // namespace b200msm
