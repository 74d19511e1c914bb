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
    SL_COUNT
};

constexpr int MAX_H2D_PARTS = 16;

struct Ctx {
    int device = -1;
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
};

// Type-erased per-group entry points, implemented once per group in group_*.cu.
struct GroupOps {
    uint32_t id;
    size_t group_bytes, scalar_bytes, affine_bytes;
    int scalar_bits;
    // d_raw (reference layout, device) -> d_aff, d_inf
    bool (*ingest)(Ctx &, const void *d_raw, size_t n, uint32_t form, void *d_aff, uint8_t *d_inf);
    // the MSM proper; result (reference layout) into d_out
    bool (*msm)(Ctx &, const void *d_aff, const uint8_t *d_inf, const void *d_scalars, size_t n,
                void *d_out, const MsmShape *shape);
    // table[w * n + i] = 2^(w c) P_i (device affine), w < W, from n device-affine bases
    bool (*precompute)(Ctx &, const void *d_aff, const uint8_t *d_inf, size_t n, int c, int W,
                       void *d_table);
    // libff's binary stream encoding (curve_serialization.tcc:79-101, uncompressed) -> device
    // affine + infinity flags; flags bit 0: Montgomery form (else plain)
    bool (*stream_decode)(Ctx &, const void *d_bytes, size_t n, uint32_t flags, void *d_aff,
                          uint8_t *d_inf);
    bool (*sum)(Ctx &, const void *d_raw, size_t count, void *d_out);
    bool (*to_special)(Ctx &, void *d_raw, size_t n);
    bool (*batch_exp)(Ctx &, const void *d_table_raw, size_t rows, size_t row_len, int window,
                      const void *d_scalars, size_t n, void *d_out);
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
