// extern "C" boundary of the engine (declared in include/b200msm.h).  Conventions follow the
// reference's ffi layer (ffi/ffi.cpp:16-54,124-260): bool status, exact-size checks, caller-owned
// buffers, outputs written only on success, no exceptions across the boundary.
//
// Multi-GPU: the engine binds any number of devices of the box (b200msm_init_devices).  A call
// cuts its index range into one contiguous range per device -- `one = n / D`, the last range takes
// the remainder: the reference's own `chunks` split, multiexp.tcc:663-678 -- runs the single-GPU
// pipeline for every range from its own host thread, and adds the D partial points on the first
// device (multiexp.tcc:681-685).  Everything below `fan_out` works on ONE context.
#include "../../include/b200msm.h"

#include <atomic>
#include <condition_variable>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <thread>

#include "engine.hpp"

namespace b200msm {

static thread_local std::string g_error;

void set_error(const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_error = buf;
}

bool Ctx::reserve(Slot s, size_t bytes, void **out) {
    if (bytes == 0) bytes = 16;
    if (slot_bytes[s] < bytes) {
        if (slot_ptr[s]) {
            // kernels on any of the context's streams may still use the slot (deferred tails,
            // carry rounds of an earlier point range, uploads): wait for all of them
            B200_CUDA(cudaStreamSynchronize(stream));
            B200_CUDA(cudaStreamSynchronize(tail_stream));
            B200_CUDA(cudaStreamSynchronize(acc_stream));
            B200_CUDA(cudaStreamSynchronize(copy_stream));
            if (sort_stream) B200_CUDA(cudaStreamSynchronize(sort_stream));
            B200_CUDA(cudaFree(slot_ptr[s]));
            slot_ptr[s] = nullptr;
            slot_bytes[s] = 0;
        }
        size_t want = bytes + bytes / 16;  // a little head-room against near-miss regrowth
        cudaError_t e = cudaMalloc(&slot_ptr[s], want);
        if (e != cudaSuccess) {
            want = bytes;
            (void)cudaGetLastError();
            e = cudaMalloc(&slot_ptr[s], want);
        }
        if (e != cudaSuccess) {
            set_error("cudaMalloc of %zu bytes (scratch slot %d, device %d) failed: %s", want, (int)s,
                      device, cudaGetErrorString(e));
            slot_ptr[s] = nullptr;
            return false;
        }
        slot_bytes[s] = want;
    }
    *out = slot_ptr[s];
    return true;
}

void Ctx::release_all() {
    for (int s = 0; s < SL_COUNT; s++) {
        if (slot_ptr[s]) cudaFree(slot_ptr[s]);
        slot_ptr[s] = nullptr;
        slot_bytes[s] = 0;
    }
}

bool Stager::init(int want_lanes) {
    want_lanes = std::min(std::max(want_lanes, 1), STAGE_LANES_MAX);
    for (; lanes < want_lanes; lanes++)
        for (int b = 0; b < 2; b++) {
            B200_CUDA(cudaHostAlloc(&buf[lanes][b], STAGE_BYTES, cudaHostAllocDefault));
            B200_CUDA(cudaEventCreateWithFlags(&ev[lanes][b], cudaEventDisableTiming));
        }
    return true;
}

void Stager::release() {
    for (int l = 0; l < STAGE_LANES_MAX; l++)
        for (int b = 0; b < 2; b++) {
            if (buf[l][b]) cudaFreeHost(buf[l][b]);
            if (ev[l][b]) cudaEventDestroy(ev[l][b]);
            buf[l][b] = nullptr;
            ev[l][b] = nullptr;
        }
    lanes = 0;
}

namespace {

// ---- engine state -------------------------------------------------------------------------------

// One piece of a resident array: the index range [lo, lo + n) on the device of context `ci`.
struct Shard {
    int ci = 0;
    size_t lo = 0, n = 0;
    void *data = nullptr;    // affine bases, scalars, precomputed multiples, or a window table
    uint8_t *inf = nullptr;  // bases and precomputed multiples
    MsmShape shape = {0, 0, 1, 0};  // precomputed multiples only
};

enum ResidentKind { RK_BASES, RK_SCALARS, RK_MULTIPLES, RK_WINDOW_TABLE };

// Bases, scalars and precomputed multiples are cut by base index exactly like a call's range;
// a fixed-base window table is replicated (batch_exp shards by output range).
struct Resident {
    uint32_t group = 0;
    ResidentKind kind = RK_BASES;
    size_t n = 0;
    std::vector<Shard> shards;
    // window tables (get_window_table, multiexp.tcc:809-846)
    size_t scalar_bits = 0, window = 0, rows = 0, row_len = 0;
};

// A streamed MSM in progress (b200msm_stream_*); lives on the first context
struct Stream {
    uint32_t group;
    size_t n, consumed = 0;
    uint32_t encoding, c;
    int W;                    // elements per base in the stream (1 without precomputation)
    void *d_scalars = nullptr;
    void *d_partials = nullptr;
    size_t partials = 0;
};
constexpr size_t STREAM_MAX_PARTIALS = 256;

std::mutex g_mutex;
std::vector<std::unique_ptr<Ctx>> g_ctxs;  // bound devices, in the order given to init
int g_use = 0;                             // contexts a call spreads over; 0 = all
std::map<uint64_t, Stream> g_streams;
std::map<uint64_t, Resident> g_resident;
uint64_t g_next_handle = 1;

const GroupOps *ops_for(uint32_t group) {
    switch (group) {
        case B200MSM_ALT_BN128_G1: return ops_alt_bn128_g1();
        case B200MSM_ALT_BN128_G2: return ops_alt_bn128_g2();
        case B200MSM_BLS12_377_G1: return ops_bls12_377_g1();
        case B200MSM_BW6_761_G1: return ops_bw6_761_g1();
        case B200MSM_BLS12_377_G2: return ops_bls12_377_g2();
        case B200MSM_BW6_761_G2: return ops_bw6_761_g2();
        case B200MSM_BLS12_381_G1: return ops_bls12_381_g1();
        case B200MSM_BLS12_381_G2: return ops_bls12_381_g2();
    }
    return nullptr;
}

bool bind(Ctx &ctx) {
    B200_CUDA(cudaSetDevice(ctx.device));
    return true;
}

// the first context: single-device entry points (streams, wire format, debug ops) run there
bool ready() {
    if (g_ctxs.empty()) {
        set_error("b200msm_init() has not been called (no CUDA device bound; there is no CPU path)");
        return false;
    }
    return bind(*g_ctxs[0]);
}

size_t active_devices() {
    const size_t all = g_ctxs.size();
    return g_use > 0 ? std::min<size_t>((size_t)g_use, all) : all;
}

const GroupOps *checked_ops(uint32_t group) {
    const GroupOps *ops = ops_for(group);
    if (!ops) set_error("unknown group id %u", group);
    return ops;
}

// The reference's range split (multiexp.tcc:655-678): fewer elements than ranges, or one range,
// is a single call; otherwise `one = n / D` per range and the last takes the remainder.
struct Split {
    size_t D = 1, one = 0, n = 0;
    Split(size_t n_, size_t devices) : n(n_) {
        D = (devices <= 1 || n_ < devices) ? 1 : devices;
        one = n_ / D;
    }
    size_t lo(size_t k) const { return k * one; }
    size_t hi(size_t k) const { return k + 1 == D ? n : (k + 1) * one; }
    size_t count(size_t k) const { return hi(k) - lo(k); }
};

// Runs fn(k) for k < D, range 0 on the calling thread and the others on their own host threads
// (each binds its device itself).  The first failure's message becomes the caller's error.
template <class Fn>
bool fan_out(size_t D, Fn fn) {
    if (D <= 1) return fn((size_t)0);
    std::vector<std::string> errs(D);
    std::vector<char> ok(D, 0);
    std::vector<std::thread> th;
    th.reserve(D - 1);
    for (size_t k = 1; k < D; k++)
        th.emplace_back([&, k] {
            ok[k] = fn(k) ? 1 : 0;
            if (!ok[k]) errs[k] = g_error;
        });
    ok[0] = fn((size_t)0) ? 1 : 0;
    if (!ok[0]) errs[0] = g_error;
    for (auto &t : th) t.join();
    for (size_t k = 0; k < D; k++)
        if (!ok[k]) {
            g_error = errs[k].empty() ? std::string("device call failed") : errs[k];
            return false;
        }
    return true;
}

bool collect_timings(Ctx &ctx) {
    B200_CUDA(cudaEventRecord(ctx.ev[PH_TOTAL + 1], ctx.stream));
    B200_CUDA(cudaEventSynchronize(ctx.ev[PH_TOTAL + 1]));
    for (int p = 0; p < PH_TOTAL; p++)
        B200_CUDA(cudaEventElapsedTime(&ctx.last_ms[p], ctx.ev[p], ctx.ev[p + 1]));
    B200_CUDA(cudaEventElapsedTime(&ctx.last_ms[PH_TOTAL], ctx.ev[0], ctx.ev[PH_TOTAL + 1]));
    ctx.have_timings = true;
    return true;
}

// all events of a run must exist on the stream even when a phase is skipped
bool mark_all(Ctx &ctx, int from) {
    for (int p = from; p <= PH_TOTAL; p++) B200_CUDA(cudaEventRecord(ctx.ev[p], ctx.stream));
    return true;
}

// D2H of the result + phase timings; the only host synchronisation point of a call
bool finish_msm(Ctx &ctx, const GroupOps *ops, const void *d_out, void *out) {
    std::vector<unsigned char> tmp(ops->group_bytes);
    B200_CUDA(cudaMemcpyAsync(tmp.data(), d_out, ops->group_bytes, cudaMemcpyDeviceToHost, ctx.stream));
    if (!collect_timings(ctx)) return false;
    B200_CUDA(cudaStreamSynchronize(ctx.stream));
    memcpy(out, tmp.data(), ops->group_bytes);
    return true;
}

bool run_msm(Ctx &ctx, const GroupOps *ops, const void *d_aff, const uint8_t *d_inf,
             const void *d_scalars, size_t n, void *out, const MsmShape *shape = nullptr) {
    void *d_out;
    if (!ctx.reserve(SL_OUT, ops->group_bytes, &d_out)) return false;
    if (!mark_all(ctx, PH_INGEST + 1)) return false;
    if (!ops->msm(ctx, d_aff, d_inf, d_scalars, n, d_out, shape)) return false;
    return finish_msm(ctx, ops, d_out, out);
}

// How many point ranges a host-buffer multi_exp is cut into so that the upload of range k + 1
// overlaps the computation of range k.
// Ranges of a pipelined call.  What a call pays on top of max(upload, computation) is the first
// range's upload, the last range's computation and a cost per range (its sort launches, sparser
// buckets in the merging accumulation), so the best count grows like the square root of the
// call's length.  Measured per field class (profiles/e2e_ranges_r02d.log): the 8-limb prime field
// gains from 2^19 points on; the merging accumulation of the wider fields is a slower kernel than
// their plain one, so they take fewer ranges; bw6_761 takes none: at its digit sizes a range
// leaves one or two items per bucket, and a slot-kernel variant that merges into live buckets
// measured no better than the register one (2^19: 1 range 35.5 ms, 2 ranges 41.1).
size_t pick_parts(const Ctx &ctx, const GroupOps *ops, size_t n, bool scalars_only, bool pageable) {
    if (scalars_only && ctx.opt_h2d_parts == 0) {
        // a quarter of the bytes: few ranges, so that the buckets stay dense in every range
        return n < ((size_t)1 << 21) ? 1 : 4;
    }
    if (ctx.opt_h2d_parts > 0) return std::min<size_t>((size_t)ctx.opt_h2d_parts, std::max<size_t>(n, 1));
    const bool prime8 = ops->affine_bytes == 64;                           // alt_bn128 G1
    const bool prime24 = ops->affine_bytes == 192 && ops->scalar_bytes == 48;  // bw6_761
    if (!ctx.opt_h2d_shared_buckets) return n < ((size_t)1 << 21) ? 1 : n < ((size_t)1 << 23) ? 3 : 5;
    if (prime24) return 1;
    if (prime8 && !pageable) {  // pageable sources are bound by the host-side staging, not by the pipeline's ends
        if (n < ((size_t)1 << 19)) return 1;
        if (n < ((size_t)1 << 20)) return 2;
        if (n < ((size_t)1 << 22)) return 4;
        if (n < ((size_t)1 << 24)) return 8;
        if (n < ((size_t)1 << 25)) return 12;
        return 16;
    }
    if (n < ((size_t)1 << (prime8 ? 21 : 20))) return 1;
    if (n < ((size_t)1 << 22)) return 2;
    if (n < ((size_t)1 << 23)) return 3;
    return prime8 ? 12 : 6;
}

// Range boundaries of a pipelined call: sizes grow geometrically (x h2d_growth %, default 1.3),
// so the first upload -- the only one nothing overlaps -- is short, while the later, longer MSMs
// still finish after their successor's upload.  bounds has parts + 1 entries.
void pick_bounds(const Ctx &ctx, size_t n, size_t parts, size_t *bounds, bool scalars_only) {
    // With the computation as fast as the upload the LAST range's computation is as exposed as
    // the first range's upload: long calls use (nearly) equal ranges (2^24: 12 x 1.2 51.0 ms,
    // 12 x 1.1 49.3; 2^26: 16 x 1.2 186 ms, 16 equal 169 ms)
    const double g = ctx.opt_h2d_growth > 0 ? ctx.opt_h2d_growth / 100.0
                     : scalars_only ? 2.0
                     : !ctx.opt_h2d_shared_buckets ? 1.3
                     : parts >= 16 ? 1.0 : parts >= 12 ? 1.1 : parts > 5 ? 1.2 : 1.3;
    // weights: geometric growth from the first range; with h2d_taper the trailing ranges shrink
    // again (the last range's computation overlaps no upload either)
    auto weight = [&](size_t k) {
        size_t e = k;
        if (ctx.opt_h2d_taper) e = std::min(k, (parts - 1 - k) * (size_t)ctx.opt_h2d_taper);
        double w = 1;
        for (size_t q = 0; q < e; q++) w *= g;
        return w;
    };
    double total = 0;
    for (size_t k = 0; k < parts; k++) total += weight(k);
    double acc = 0;
    bounds[0] = 0;
    for (size_t k = 0; k + 1 < parts; k++) {
        acc += weight(k);
        size_t b = (size_t)((double)n * (acc / total));
        if (n >= ((size_t)1 << 16)) b &= ~(size_t)255;  // 16-byte aligned ranges in all arrays
        bounds[k + 1] = std::min(std::max(b, bounds[k]), n);
    }
    bounds[parts] = n;
}

// ---- host -> device uploads ------------------------------------------------------------------
bool is_pinned(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        (void)cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

// Queues the upload of `count` records on stream `cs`; returns once every DMA has been issued
// (pageable sources: once the last slice has been copied into a staging buffer).  `keep` <
// `stride` packs the records on the way: a special-form base is X | Y | Z with Z = 1, or the zero
// element (Z = 0), so only the first `keep` bytes (X | Y) of every `stride`-byte element travel --
// a third fewer bytes over PCIe -- and the zero element is sent as the all-zero record, which is
// the device's own affine encoding of infinity ((0, 0) is on none of the curves, b != 0).
// Pinned sources of full records go straight to cudaMemcpyAsync; everything else passes through
// the context's pinned ring, STAGE_LANES host threads copying (and packing) slices.
bool upload(Ctx &ctx, void *dst, const void *src, size_t count, size_t stride, size_t keep,
            cudaStream_t cs) {
    if (count == 0) return true;
    const bool pack = keep < stride;
    const size_t bytes = count * stride;
    if (!pack && (is_pinned(src) || bytes < STAGE_BYTES))
        return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, cs) == cudaSuccess;
    const size_t per = std::max<size_t>(1, STAGE_BYTES / keep);  // records per slice
    const size_t slices = (count + per - 1) / per;
    const int LANES = (int)std::min<size_t>((size_t)std::min(std::max(ctx.opt_stage_lanes, 1), STAGE_LANES_MAX), slices);
    if (!ctx.stager.init(LANES)) return false;
    std::atomic<bool> ok{true};
    Stager &sg = ctx.stager;
    const int device = ctx.device;
    auto lane_fn = [&](int l) {
        if (cudaSetDevice(device) != cudaSuccess) ok = false;
        int b = 0;
        for (size_t s = l; s < slices && ok; s += LANES, b ^= 1) {
            const size_t r0 = s * per, rn = std::min(per, count - r0);
            // the previous DMA out of this buffer must have finished
            if (cudaEventSynchronize(sg.ev[l][b]) != cudaSuccess) ok = false;
            const char *from = (const char *)src + r0 * stride;
            char *to = (char *)sg.buf[l][b];
            if (pack) {
                const size_t zw = (stride - keep) / 8;
                for (size_t r = 0; r < rn; r++) {
                    const char *e = from + r * stride;
                    uint64_t z = 0, w;
                    for (size_t q = 0; q < zw; q++) {
                        memcpy(&w, e + keep + q * 8, 8);
                        z |= w;
                    }
                    if (z != 0)
                        memcpy(to + r * keep, e, keep);
                    else
                        memset(to + r * keep, 0, keep);
                }
            } else {
                memcpy(to, from, rn * keep);
            }
            if (cudaMemcpyAsync((char *)dst + r0 * keep, to, rn * keep, cudaMemcpyHostToDevice, cs) !=
                    cudaSuccess ||
                cudaEventRecord(sg.ev[l][b], cs) != cudaSuccess)
                ok = false;
        }
    };
    if (LANES == 1) {
        lane_fn(0);
        return ok;
    }
    std::vector<std::thread> lanes;
    for (int l = 0; l < LANES; l++) lanes.emplace_back(lane_fn, l);
    for (auto &t : lanes) t.join();
    return ok;
}

// Device -> host copy of a large result array (batch_exp, batch_to_special).  Pageable
// destinations -- a std::vector<G> -- go through the pinned ring as the uploads do: the DMA of
// slice s + LANES runs while slice s is copied out of its staging buffer.  Synchronous: the data
// is in `dst` on return.
bool download(Ctx &ctx, void *dst, const void *src, size_t bytes, cudaStream_t cs) {
    if (bytes == 0) return true;
    if (is_pinned(dst) || bytes < 2 * STAGE_BYTES) {
        B200_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, cs));
        B200_CUDA(cudaStreamSynchronize(cs));
        return true;
    }
    const size_t slices = (bytes + STAGE_BYTES - 1) / STAGE_BYTES;
    const int LANES = (int)std::min<size_t>((size_t)std::min(std::max(ctx.opt_stage_lanes, 1), STAGE_LANES_MAX), slices);
    if (!ctx.stager.init(LANES)) return false;
    B200_CUDA(cudaStreamSynchronize(cs));  // the producer kernels are done: lanes copy on their own streams
    std::atomic<bool> ok{true};
    Stager &sg = ctx.stager;
    const int device = ctx.device;
    auto lane_fn = [&](int l) {
        cudaStream_t ls = nullptr;
        if (cudaSetDevice(device) != cudaSuccess ||
            cudaStreamCreateWithFlags(&ls, cudaStreamNonBlocking) != cudaSuccess) {
            ok = false;
            return;
        }
        auto issue = [&](size_t s, int b) {
            const size_t off = s * STAGE_BYTES, len = std::min(STAGE_BYTES, bytes - off);
            if (cudaMemcpyAsync(sg.buf[l][b], (const char *)src + off, len, cudaMemcpyDeviceToHost, ls) !=
                    cudaSuccess ||
                cudaEventRecord(sg.ev[l][b], ls) != cudaSuccess)
                ok = false;
        };
        int b = 0;
        if ((size_t)l < slices) issue((size_t)l, 0);
        for (size_t s = l; s < slices && ok; s += LANES, b ^= 1) {
            if (s + LANES < slices) issue(s + LANES, b ^ 1);
            if (cudaEventSynchronize(sg.ev[l][b]) != cudaSuccess) ok = false;
            const size_t off = s * STAGE_BYTES, len = std::min(STAGE_BYTES, bytes - off);
            memcpy((char *)dst + off, sg.buf[l][b], len);
        }
        cudaStreamSynchronize(ls);
        cudaStreamDestroy(ls);
    };
    std::vector<std::thread> lanes;
    for (int l = 0; l < LANES; l++) lanes.emplace_back(lane_fn, l);
    for (auto &t : lanes) t.join();
    if (!ok) set_error("device to host copy failed");
    return ok;
}

bool upload_flat(Ctx &ctx, void *dst, const void *src, size_t bytes, cudaStream_t cs) {
    return upload(ctx, dst, src, bytes, 1, 1, cs);
}

// Bases of a host call -> device.  `packed`: X | Y records straight into the device-affine array
// (ingest form 2 then only derives the infinity flags); else the full elements into d_dst.
bool upload_bases(Ctx &ctx, const GroupOps *ops, void *d_dst, const void *bases, size_t count,
                  bool packed, cudaStream_t cs) {
    const size_t gb = ops->group_bytes;
    return upload(ctx, d_dst, bases, count, gb, packed ? ops->affine_bytes : gb, cs);
}

// special-form bases are packed on the way up unless the source is pinned memory (a plain DMA of
// the full records is then faster than any host-side packing; measured, profiles/h2d_probe_*)
bool want_pack(const Ctx &ctx, const void *bases, uint32_t base_form, size_t n) {
    if (base_form != 1 || n == 0) return false;
    if (ctx.opt_h2d_pack == 0) return false;
    if (ctx.opt_h2d_pack == 2) return true;
    return !is_pinned(bases);
}

// Uploads the parts of a pipelined call in order on the copy stream from a helper thread (the
// staging copies block their caller, which must not be the thread that launches the kernels).
struct Uploader {
    std::mutex m;
    std::condition_variable cv;
    int issued = 0;
    bool failed = false;
    std::thread th;
    void wait_for(int k) {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [&] { return issued > k || failed; });
    }
    ~Uploader() {
        if (th.joinable()) th.join();
    }
};

// ---- one MSM from host memory on ONE context ----------------------------------------------------
// bases != nullptr: n reference-layout elements on the host (uploaded and ingested here).
// bases == nullptr: the bases are resident (d_aff_res / d_inf_res) and only the scalars travel --
// the shape of a prover's call: fixed proving key in HBM, fresh witness scalars per proof.
bool msm_host_on(Ctx &ctx, const GroupOps *ops, const void *bases, const void *d_aff_res,
                 const uint8_t *d_inf_res, const void *scalars, size_t n, uint32_t base_form,
                 void *out) {
    if (!bind(ctx)) return false;
    const bool resident = bases == nullptr;
    const size_t gb = ops->group_bytes, sb = ops->scalar_bytes, ab = ops->affine_bytes;
    const bool packed = !resident && want_pack(ctx, bases, base_form, n);
    const size_t rb = packed ? ab : gb;  // bytes per uploaded base record
    const uint32_t form = packed ? 2u : base_form;
    void *d_raw = nullptr, *d_aff = nullptr, *d_inf = nullptr, *d_sc, *d_out, *d_parts;
    if (!resident) {
        if (!ctx.reserve(SL_AFF, n * ab, &d_aff)) return false;
        if (!ctx.reserve(SL_INF, n, &d_inf)) return false;
        if (packed)
            d_raw = d_aff;  // the packed records ARE the device-affine points
        else if (!ctx.reserve(SL_RAW, n * rb, &d_raw))
            return false;
    } else {
        d_aff = const_cast<void *>(d_aff_res);
        d_inf = const_cast<uint8_t *>(d_inf_res);
    }
    if (!ctx.reserve(SL_SCALARS, n * sb, &d_sc)) return false;
    if (!ctx.reserve(SL_OUT, gb, &d_out)) return false;
    size_t parts = pick_parts(ctx, ops, n, resident, !resident && !is_pinned(bases));
    size_t bounds[MAX_H2D_PARTS + 1] = {0};
    if (parts > 1) {  // ranges that came out empty (tiny n) are dropped
        pick_bounds(ctx, n, parts, bounds, resident);
        size_t m = 0;
        for (size_t k = 0; k < parts; k++)
            if (bounds[k + 1] > bounds[k]) bounds[++m] = bounds[k + 1];
        parts = m;
    }
    if (!ctx.reserve(SL_PARTIALS, std::max<size_t>(parts, 1) * gb, &d_parts)) return false;
    B200_CUDA(cudaEventRecord(ctx.ev[0], ctx.stream));
    if (parts <= 1) {
        if (n > 0) {
            if ((!resident && !upload_bases(ctx, ops, d_raw, bases, n, packed, ctx.stream)) ||
                !upload_flat(ctx, d_sc, scalars, n * sb, ctx.stream))
                return set_error("multi_exp: host to device copy failed"), false;
            if (!resident && !ops->ingest(ctx, d_raw, n, form, d_aff, (uint8_t *)d_inf)) return false;
        }
        return run_msm(ctx, ops, d_aff, (const uint8_t *)d_inf, d_sc, n, out);
    }
    // ---- pipelined: upload of range k + 1 (copy stream) under the MSM of range k
    auto lo_of = [&](size_t k) { return bounds[k]; };
    auto hi_of = [&](size_t k) { return bounds[k + 1]; };
    // the compute stream must not overwrite the staging buffers of a previous call that is
    // still in flight: the copy stream starts after everything queued so far
    B200_CUDA(cudaEventRecord(ctx.ev_sync, ctx.stream));
    B200_CUDA(cudaStreamWaitEvent(ctx.copy_stream, ctx.ev_sync, 0));
    Uploader up;
    cudaStream_t cs = ctx.copy_stream;
    cudaEvent_t *evs = ctx.ev_copy;
    Ctx *cp = &ctx;
    up.th = std::thread([&, cp, cs, evs] {
        bool ok = cudaSetDevice(cp->device) == cudaSuccess;
        for (size_t k = 0; k < parts && ok; k++) {
            const size_t lo = lo_of(k), cnt = hi_of(k) - lo;
            ok = upload_flat(*cp, (char *)d_sc + lo * sb, (const char *)scalars + lo * sb, cnt * sb, cs) &&
                 (resident || upload_bases(*cp, ops, (char *)d_raw + lo * rb,
                                           (const char *)bases + lo * gb, cnt, packed, cs)) &&
                 cudaEventRecord(evs[k], cs) == cudaSuccess;
            std::lock_guard<std::mutex> lk(up.m);
            if (ok)
                up.issued = (int)k + 1;
            else
                up.failed = true;
            up.cv.notify_all();
        }
    });
    if (!mark_all(ctx, PH_INGEST + 1)) return false;
    // Shared buckets (default): every range is sorted and accumulated with the digit size of the
    // whole call into ONE bucket set; the bucket reduction and the Horner chain run once, after
    // the last range.  Otherwise (option h2d_shared_buckets = 0) every range is an independent
    // MSM whose latency-bound tail runs under the kernels of range k + 1.
    const bool shared = ctx.opt_h2d_shared_buckets != 0;
    struct TailScope {
        Ctx &c;
        TailScope(Ctx &c_, bool shared_, int range_c) : c(c_) {
            c.defer_tail = !shared_;
            c.range_c = shared_ ? range_c : 0;
        }
        ~TailScope() {
            if (c.tail_pending) cudaStreamSynchronize(c.tail_stream);  // error paths
            c.defer_tail = c.tail_pending = false;
            c.range_c = 0;
        }
    } tail_scope(ctx, shared, (int)ops->window_bits(ctx, n));
    // scratch for the largest range up front: no regrowth (free + malloc + stream syncs) inside
    // the pipeline
    {
        size_t big = 0;
        for (size_t k = 0; k < parts; k++) big = std::max(big, hi_of(k) - lo_of(k));
        if (!ops->msm_prepare(ctx, big, nullptr)) return false;
    }
    for (size_t k = 0; k < parts; k++) {
        const size_t lo = lo_of(k), cnt = hi_of(k) - lo;
        ctx.range_first = k == 0;
        ctx.range_last = k + 1 == parts;
        up.wait_for((int)k);
        if (up.failed) return set_error("multi_exp: host to device copy failed"), false;
        // The range's ingest and counting sort run on the sort stream, ordered only behind the
        // range's own upload: they overlap the accumulation of range k - 1 on the main stream.
        B200_CUDA(cudaStreamWaitEvent(ctx.sort_stream, ctx.ev_copy[k], 0));
        {
            struct StreamSwap {  // the ingest kernels are issued on Ctx::stream
                Ctx &c;
                cudaStream_t keep;
                explicit StreamSwap(Ctx &c_) : c(c_), keep(c_.stream) { c.stream = c.sort_stream; }
                ~StreamSwap() { c.stream = keep; }
            } swap(ctx);
            if (!resident &&
                !ops->ingest(ctx, (char *)d_raw + lo * rb, cnt, form, (char *)d_aff + lo * ab,
                             (uint8_t *)d_inf + lo))
                return false;
        }
        B200_CUDA(cudaEventRecord(ctx.ev[PH_INGEST + 1], ctx.sort_stream));
        ctx.sort_inputs_ready = true;
        const bool msm_ok = ops->msm(ctx, (char *)d_aff + lo * ab, (const uint8_t *)d_inf + lo,
                                     (char *)d_sc + lo * sb, cnt, (char *)d_parts + k * gb, nullptr);
        ctx.sort_inputs_ready = false;
        if (!msm_ok) return false;
        if (ctx.opt_trace) {  // development aid: per-range phase times (serialises the pipeline)
            B200_CUDA(cudaDeviceSynchronize());
            fprintf(stderr, "[b200msm] device %d range %zu: %zu points", ctx.device, k, cnt);
            for (int p = PH_INGEST; p < PH_TOTAL; p++) {
                float ms = 0;
                B200_CUDA(cudaEventElapsedTime(&ms, ctx.ev[p], ctx.ev[p + 1]));
                fprintf(stderr, " %s=%.3f", b200msm_timing_name((size_t)p), ms);
            }
            fprintf(stderr, "\n");
        }
    }
    if (ctx.tail_pending) {
        B200_CUDA(cudaStreamWaitEvent(ctx.stream, ctx.ev_tail_done, 0));
        ctx.tail_pending = false;
    }
    if (shared) {  // the last range reduced the shared buckets: its result is the result
        B200_CUDA(cudaMemcpyAsync(d_out, (char *)d_parts + (parts - 1) * gb, gb,
                                  cudaMemcpyDeviceToDevice, ctx.stream));
    } else if (!ops->sum(ctx, d_parts, parts, d_out)) {
        return false;
    }
    return finish_msm(ctx, ops, d_out, out);
}

// sum of `count` reference-layout points held on the host, on one context
bool sum_on(Ctx &ctx, const GroupOps *ops, const void *points, size_t count, void *out) {
    if (!bind(ctx)) return false;
    void *d_raw, *d_out;
    const size_t gb = ops->group_bytes;
    if (!ctx.reserve(SL_SUM_IN, std::max<size_t>(count, 1) * gb, &d_raw)) return false;
    if (!ctx.reserve(SL_OUT, gb, &d_out)) return false;
    if (count) B200_CUDA(cudaMemcpyAsync(d_raw, points, count * gb, cudaMemcpyHostToDevice, ctx.stream));
    if (!ops->sum(ctx, d_raw, count, d_out)) return false;
    std::vector<unsigned char> tmp(gb);
    B200_CUDA(cudaMemcpyAsync(tmp.data(), d_out, gb, cudaMemcpyDeviceToHost, ctx.stream));
    B200_CUDA(cudaStreamSynchronize(ctx.stream));
    memcpy(out, tmp.data(), gb);
    return true;
}

// Runs part(k, partial_k) for every range of the split on its own device and adds the partial
// points on the first device (multiexp.tcc:681-685).  One range: the partial is the result.
template <class Fn>
bool sharded_msm(const GroupOps *ops, const Split &sp, void *out, Fn part) {
    if (sp.D == 1) return part((size_t)0, out);
    const size_t gb = ops->group_bytes;
    std::vector<unsigned char> partials(sp.D * gb);
    if (!fan_out(sp.D, [&](size_t k) { return part(k, partials.data() + k * gb); })) return false;
    return sum_on(*g_ctxs[0], ops, partials.data(), sp.D, out);
}

void free_resident(Resident &r) {
    for (auto &s : r.shards) {
        if ((size_t)s.ci < g_ctxs.size()) cudaSetDevice(g_ctxs[s.ci]->device);
        if (s.data) cudaFree(s.data);
        if (s.inf) cudaFree(s.inf);
    }
    r.shards.clear();
}

// two resident arrays cut at the same places (same length, same number of devices)?
bool same_cut(const Resident &a, const Resident &b) {
    if (a.shards.size() != b.shards.size()) return false;
    for (size_t k = 0; k < a.shards.size(); k++)
        if (a.shards[k].ci != b.shards[k].ci || a.shards[k].lo != b.shards[k].lo ||
            (a.shards.size() > 1 && a.shards[k].n != b.shards[k].n))
            return false;
    return true;
}

uint32_t pick_precompute_c(const GroupOps *ops, const Ctx &ctx, size_t n) {
    uint32_t c = ops->window_bits(ctx, n);
    while (c < 24 && (uint64_t)n * ((ops->scalar_bits + c - 1) / c) >= (1ull << 31)) c++;
    return c;
}

Ctx *make_ctx(int device, int index) {
    if (cudaSetDevice(device) != cudaSuccess) {
        set_error("cudaSetDevice(%d) failed: %s", device, cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    std::unique_ptr<Ctx> ctx(new Ctx());
    ctx->device = device;
    ctx->index = index;
    auto ok = [&](cudaError_t e) {
        if (e == cudaSuccess) return true;
        set_error("creating the engine context on device %d failed: %s", device, cudaGetErrorString(e));
        return false;
    };
    if (!ok(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking))) return nullptr;
    if (!ok(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking))) return nullptr;
    for (auto &e : ctx->ev)
        if (!ok(cudaEventCreate(&e))) return nullptr;
    for (auto &e : ctx->ev_copy)
        if (!ok(cudaEventCreateWithFlags(&e, cudaEventDisableTiming))) return nullptr;
    if (!ok(cudaEventCreateWithFlags(&ctx->ev_sync, cudaEventDisableTiming))) return nullptr;
    int prio_lo = 0, prio_hi = 0;
    if (!ok(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi))) return nullptr;
    if (!ok(cudaStreamCreateWithPriority(&ctx->tail_stream, cudaStreamNonBlocking, prio_hi))) return nullptr;
    if (!ok(cudaStreamCreateWithFlags(&ctx->acc_stream, cudaStreamNonBlocking))) return nullptr;
    // high priority: the accumulation kernels fill every SM, and the sort's short blocks must get
    // the slots their blocks free instead of waiting for the last wave
    if (!ok(cudaStreamCreateWithPriority(&ctx->sort_stream, cudaStreamNonBlocking, prio_hi))) return nullptr;
    if (!ok(cudaEventCreateWithFlags(&ctx->ev_group_go, cudaEventDisableTiming))) return nullptr;
    for (auto &e : ctx->ev_group)
        if (!ok(cudaEventCreateWithFlags(&e, cudaEventDisableTiming))) return nullptr;
    for (auto &e : ctx->ev_sort)
        if (!ok(cudaEventCreateWithFlags(&e, cudaEventDisableTiming))) return nullptr;
    for (auto &e : ctx->ev_timer)
        if (!ok(cudaEventCreate(&e))) return nullptr;
    if (!ok(cudaEventCreateWithFlags(&ctx->ev_tail_go, cudaEventDisableTiming))) return nullptr;
    if (!ok(cudaEventCreateWithFlags(&ctx->ev_tail_done, cudaEventDisableTiming))) return nullptr;
    if (!ok(cudaHostAlloc((void **)&ctx->host_meta, 64, cudaHostAllocDefault))) return nullptr;
    return ctx.release();
}

void destroy_ctx(Ctx *c) {
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    c->release_all();
    c->stager.release();
    if (c->host_meta) cudaFreeHost(c->host_meta);
    c->host_meta = nullptr;
    for (auto &e : c->ev) cudaEventDestroy(e);
    for (auto &e : c->ev_copy) cudaEventDestroy(e);
    cudaEventDestroy(c->ev_sync);
    for (auto &e : c->ev_timer) cudaEventDestroy(e);
    cudaEventDestroy(c->ev_tail_go);
    cudaEventDestroy(c->ev_tail_done);
    cudaStreamDestroy(c->tail_stream);
    cudaStreamDestroy(c->acc_stream);
    cudaStreamDestroy(c->sort_stream);
    cudaEventDestroy(c->ev_group_go);
    for (auto &e : c->ev_group) cudaEventDestroy(e);
    for (auto &e : c->ev_sort) cudaEventDestroy(e);
    cudaStreamDestroy(c->copy_stream);
    cudaStreamDestroy(c->stream);
}

}  // namespace
}  // namespace b200msm

using namespace b200msm;

extern "C" {

bool b200msm_init_devices(const int *devices, size_t count) {
    std::lock_guard<std::mutex> lk(g_mutex);
    if (!devices || count == 0 || count > 64) return set_error("init_devices: bad device list"), false;
    if (!g_ctxs.empty()) {
        bool same = g_ctxs.size() == count;
        for (size_t k = 0; same && k < count; k++) same = g_ctxs[k]->device == devices[k];
        if (same) return true;
        set_error("already initialised on %zu device(s) (first: %d); call b200msm_shutdown first",
                  g_ctxs.size(), g_ctxs[0]->device);
        return false;
    }
    int have = 0;
    B200_CUDA(cudaGetDeviceCount(&have));
    for (size_t k = 0; k < count; k++)
        if (devices[k] < 0 || devices[k] >= have) {
            set_error("device %d not available (%d CUDA devices); there is no CPU path", devices[k], have);
            return false;
        }
    std::vector<std::unique_ptr<Ctx>> made;
    for (size_t k = 0; k < count; k++) {
        Ctx *c = make_ctx(devices[k], (int)k);
        if (!c) {
            for (auto &m : made) destroy_ctx(m.get());
            return false;
        }
        made.emplace_back(c);
    }
    g_ctxs = std::move(made);
    g_use = 0;
    cudaSetDevice(g_ctxs[0]->device);
    return true;
}

bool b200msm_init(int device) { return b200msm_init_devices(&device, 1); }

int b200msm_init_all(void) {
    int have = 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess || have <= 0) {
        (void)cudaGetLastError();
        set_error("no CUDA device visible; there is no CPU path");
        return 0;
    }
    std::vector<int> ids((size_t)have);
    for (int k = 0; k < have; k++) ids[(size_t)k] = k;
    return b200msm_init_devices(ids.data(), ids.size()) ? have : 0;
}

size_t b200msm_device_count(void) {
    std::lock_guard<std::mutex> lk(g_mutex);
    return g_ctxs.size();
}

void b200msm_shutdown(void) {
    std::lock_guard<std::mutex> lk(g_mutex);
    if (g_ctxs.empty()) return;
    for (auto &c : g_ctxs) {
        cudaSetDevice(c->device);
        cudaDeviceSynchronize();
    }
    for (auto &kv : g_resident) free_resident(kv.second);
    g_resident.clear();
    cudaSetDevice(g_ctxs[0]->device);
    for (auto &kv : g_streams) {
        cudaFree(kv.second.d_scalars);
        cudaFree(kv.second.d_partials);
    }
    g_streams.clear();
    for (auto &c : g_ctxs) destroy_ctx(c.get());
    g_ctxs.clear();
    g_use = 0;
}

const char *b200msm_last_error(void) { return g_error.c_str(); }

size_t b200msm_group_bytes(uint32_t group) {
    const GroupOps *o = ops_for(group);
    return o ? o->group_bytes : 0;
}
size_t b200msm_scalar_bytes(uint32_t group) {
    const GroupOps *o = ops_for(group);
    return o ? o->scalar_bytes : 0;
}

bool b200msm_multi_exp(uint32_t group, const void *bases, size_t bases_bytes, const void *scalars,
                       size_t scalars_bytes, size_t n, uint32_t base_form, void *out,
                       size_t out_bytes) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    if (base_form > 1) return set_error("invalid base_form %u", base_form), false;
    if (bases_bytes != n * ops->group_bytes || scalars_bytes != n * ops->scalar_bytes ||
        out_bytes != ops->group_bytes)
        return set_error("multi_exp: buffer sizes do not match n = %zu", n), false;
    if (!out || (n > 0 && (!bases || !scalars))) return set_error("multi_exp: null buffer"), false;
    const Split sp(n, active_devices());
    const size_t gb = ops->group_bytes, sb = ops->scalar_bytes;
    std::vector<unsigned char> res(gb);
    static const char dummy = 0;  // n == 0: msm_host_on tells host bases from resident ones by the pointer
    if (!sharded_msm(ops, sp, res.data(), [&](size_t k, void *partial) {
            const void *b = n ? (const void *)((const char *)bases + sp.lo(k) * gb) : (const void *)&dummy;
            return msm_host_on(*g_ctxs[k], ops, b, nullptr, nullptr,
                               (const char *)scalars + sp.lo(k) * sb, sp.count(k), base_form, partial);
        }))
        return false;
    memcpy(out, res.data(), gb);
    return true;
}

size_t b200msm_wire_point_bytes(uint32_t group) {
    const GroupOps *o = ops_for(group);
    return o ? o->affine_bytes : 0;
}

bool b200msm_multi_exp_wire(uint32_t group, const void *points, size_t points_bytes,
                            const void *scalars, size_t scalars_bytes, size_t n, uint32_t flags,
                            void *out, size_t out_bytes) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    const size_t pb = ops->affine_bytes, sb = ops->scalar_bytes;
    if (flags & ~(uint32_t)(B200MSM_WIRE_SKIP_CURVE_CHECK | B200MSM_WIRE_SKIP_SUBGROUP_CHECK))
        return set_error("multi_exp_wire: unknown flags %u", flags), false;
    // group_element_read (ffi_serialization.tcc:165) always checks both; skipping is the opt-out
    const uint32_t checks = ((flags & B200MSM_WIRE_SKIP_CURVE_CHECK) ? 0u : 1u) |
                            ((flags & B200MSM_WIRE_SKIP_SUBGROUP_CHECK) ? 0u : 2u);
    if (points_bytes != n * pb || scalars_bytes != n * sb || out_bytes != pb)
        return set_error("multi_exp_wire: buffer sizes do not match n = %zu", n), false;
    if (!out || (n > 0 && (!points || !scalars))) return set_error("multi_exp_wire: null buffer"), false;
    Ctx &ctx = *g_ctxs[0];
    void *d_wp, *d_ws, *d_aff, *d_inf, *d_sc, *d_out, *d_misc;
    if (!ctx.reserve(SL_WIRE_IN, points_bytes, &d_wp)) return false;
    if (!ctx.reserve(SL_WIRE_SC, scalars_bytes, &d_ws)) return false;
    if (!ctx.reserve(SL_AFF, n * ops->affine_bytes, &d_aff)) return false;
    if (!ctx.reserve(SL_INF, n, &d_inf)) return false;
    if (!ctx.reserve(SL_SCALARS, scalars_bytes, &d_sc)) return false;
    if (!ctx.reserve(SL_OUT, ops->group_bytes, &d_out)) return false;
    if (!ctx.reserve(SL_WIRE_MISC, 16 + pb, &d_misc)) return false;
    uint32_t *d_err = (uint32_t *)d_misc;
    void *d_wire_out = (char *)d_misc + 16;
    B200_CUDA(cudaEventRecord(ctx.ev[0], ctx.stream));
    if (!upload_flat(ctx, d_wp, points, points_bytes, ctx.stream) ||
        !upload_flat(ctx, d_ws, scalars, scalars_bytes, ctx.stream))
        return set_error("multi_exp_wire: host to device copy failed"), false;
    if (!ops->wire_decode(ctx, d_wp, d_ws, n, checks, d_aff, (uint8_t *)d_inf, d_sc, d_err)) return false;
    uint32_t err = 0;
    B200_CUDA(cudaMemcpyAsync(&err, d_err, 4, cudaMemcpyDeviceToHost, ctx.stream));
    B200_CUDA(cudaStreamSynchronize(ctx.stream));
    if (err)  // the reference's readers return false (ffi_serialization.tcc:64-75,165)
        return set_error("multi_exp_wire: invalid input (%s%s%s)", err & 1 ? "value >= modulus " : "",
                         err & 2 ? "point not on curve " : "", err & 4 ? "point not in subgroup" : ""),
               false;
    if (!mark_all(ctx, PH_INGEST + 1)) return false;
    if (!ops->msm(ctx, d_aff, (const uint8_t *)d_inf, d_sc, n, d_out, nullptr)) return false;
    if (!ops->wire_encode(ctx, d_out, d_wire_out)) return false;
    std::vector<unsigned char> tmp(pb);
    B200_CUDA(cudaMemcpyAsync(tmp.data(), d_wire_out, pb, cudaMemcpyDeviceToHost, ctx.stream));
    if (!collect_timings(ctx)) return false;
    B200_CUDA(cudaStreamSynchronize(ctx.stream));
    memcpy(out, tmp.data(), pb);
    return true;
}

bool b200msm_bases_upload(uint32_t group, const void *bases, size_t bases_bytes, size_t n,
                          uint32_t base_form, uint64_t *handle) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    if (base_form > 1) return set_error("invalid base_form %u", base_form), false;
    if (!handle || !bases || n == 0 || bases_bytes != n * ops->group_bytes)
        return set_error("bases_upload: bad arguments"), false;
    const Split sp(n, active_devices());
    Resident r;
    r.group = group, r.kind = RK_BASES, r.n = n;
    r.shards.resize(sp.D);
    const size_t gb = ops->group_bytes;
    const bool ok = fan_out(sp.D, [&](size_t k) {
        Ctx &ctx = *g_ctxs[k];
        if (!bind(ctx)) return false;
        Shard &s = r.shards[k];
        s.ci = (int)k, s.lo = sp.lo(k), s.n = sp.count(k);
        const void *src = (const char *)bases + s.lo * gb;
        const bool packed = want_pack(ctx, src, base_form, s.n);
        void *d_raw = nullptr;
        B200_CUDA(cudaMalloc(&s.data, s.n * ops->affine_bytes));
        B200_CUDA(cudaMalloc((void **)&s.inf, s.n));
        if (packed)
            d_raw = s.data;
        else if (!ctx.reserve(SL_RAW, s.n * gb, &d_raw))
            return false;
        if (!upload_bases(ctx, ops, d_raw, src, s.n, packed, ctx.stream))
            return set_error("bases_upload: host to device copy failed"), false;
        if (!ops->ingest(ctx, d_raw, s.n, packed ? 2u : base_form, s.data, s.inf)) return false;
        B200_CUDA(cudaStreamSynchronize(ctx.stream));
        return true;
    });
    if (!ok) {
        free_resident(r);
        return false;
    }
    *handle = g_next_handle++;
    g_resident[*handle] = std::move(r);
    return true;
}

bool b200msm_scalars_upload(uint32_t group, const void *scalars, size_t scalars_bytes, size_t n,
                            uint64_t *handle) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    if (!handle || !scalars || n == 0 || scalars_bytes != n * ops->scalar_bytes)
        return set_error("scalars_upload: bad arguments"), false;
    const Split sp(n, active_devices());
    Resident r;
    r.group = group, r.kind = RK_SCALARS, r.n = n;
    r.shards.resize(sp.D);
    const size_t sb = ops->scalar_bytes;
    const bool ok = fan_out(sp.D, [&](size_t k) {
        Ctx &ctx = *g_ctxs[k];
        if (!bind(ctx)) return false;
        Shard &s = r.shards[k];
        s.ci = (int)k, s.lo = sp.lo(k), s.n = sp.count(k);
        B200_CUDA(cudaMalloc(&s.data, s.n * sb));
        B200_CUDA(cudaMemcpy(s.data, (const char *)scalars + s.lo * sb, s.n * sb, cudaMemcpyHostToDevice));
        return true;
    });
    if (!ok) {
        free_resident(r);
        return false;
    }
    *handle = g_next_handle++;
    g_resident[*handle] = std::move(r);
    return true;
}

bool b200msm_multi_exp_resident(uint32_t group, uint64_t bases_handle, uint64_t scalars_handle,
                                size_t first, size_t n, void *out, size_t out_bytes) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    auto b = g_resident.find(bases_handle), s = g_resident.find(scalars_handle);
    if (b == g_resident.end() || s == g_resident.end() || b->second.kind != RK_BASES ||
        s->second.kind != RK_SCALARS || b->second.group != group || s->second.group != group)
        return set_error("multi_exp_resident: bad handle"), false;
    const Resident &rb = b->second, &rs = s->second;
    if (first + n > rb.n || first + n > rs.n || out_bytes != ops->group_bytes || !out)
        return set_error("multi_exp_resident: range or size mismatch"), false;
    if (!same_cut(rb, rs))
        return set_error("multi_exp_resident: bases and scalars are cut differently across the "
                         "devices (upload both with the same length and device count)"), false;
    const size_t gb = ops->group_bytes;
    // every shard takes its part of [first, first + n)
    struct Piece {
        size_t k, off, cnt;
    };
    std::vector<Piece> pieces;
    for (size_t k = 0; k < rb.shards.size(); k++) {
        const Shard &sh = rb.shards[k];
        const size_t lo = std::max(first, sh.lo), hi = std::min(first + n, sh.lo + sh.n);
        if (hi > lo) pieces.push_back(Piece{k, lo - sh.lo, hi - lo});
    }
    if (pieces.empty()) pieces.push_back(Piece{0, 0, 0});
    std::vector<unsigned char> partials(pieces.size() * gb);
    const bool ok = fan_out(pieces.size(), [&](size_t j) {
        const Piece &p = pieces[j];
        const Shard &sb_ = rb.shards[p.k], &ss = rs.shards[p.k];
        Ctx &ctx = *g_ctxs[sb_.ci];
        if (!bind(ctx)) return false;
        B200_CUDA(cudaEventRecord(ctx.ev[0], ctx.stream));
        return run_msm(ctx, ops, (const char *)sb_.data + p.off * ops->affine_bytes, sb_.inf + p.off,
                       (const char *)ss.data + p.off * ops->scalar_bytes, p.cnt,
                       partials.data() + j * gb);
    });
    if (!ok) return false;
    if (pieces.size() == 1) {
        memcpy(out, partials.data(), gb);
        return true;
    }
    std::vector<unsigned char> res(gb);
    if (!sum_on(*g_ctxs[0], ops, partials.data(), pieces.size(), res.data())) return false;
    memcpy(out, res.data(), gb);
    return true;
}

bool b200msm_multi_exp_resident_bases(uint32_t group, uint64_t bases_handle, const void *scalars,
                                      size_t scalars_bytes, size_t n, void *out, size_t out_bytes) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    auto b = g_resident.find(bases_handle);
    if (b == g_resident.end() || b->second.kind != RK_BASES || b->second.group != group)
        return set_error("multi_exp_resident_bases: bad handle"), false;
    const Resident &rb = b->second;
    if (n != rb.n || scalars_bytes != n * ops->scalar_bytes || out_bytes != ops->group_bytes || !out ||
        !scalars)
        return set_error("multi_exp_resident_bases: need %zu scalars", rb.n), false;
    const size_t gb = ops->group_bytes, sb = ops->scalar_bytes, D = rb.shards.size();
    std::vector<unsigned char> partials(D * gb);
    const bool ok = fan_out(D, [&](size_t k) {
        const Shard &sh = rb.shards[k];
        return msm_host_on(*g_ctxs[sh.ci], ops, nullptr, sh.data, sh.inf,
                           (const char *)scalars + sh.lo * sb, sh.n, 1u, partials.data() + k * gb);
    });
    if (!ok) return false;
    if (D == 1) {
        memcpy(out, partials.data(), gb);
        return true;
    }
    std::vector<unsigned char> res(gb);
    if (!sum_on(*g_ctxs[0], ops, partials.data(), D, res.data())) return false;
    memcpy(out, res.data(), gb);
    return true;
}

// ---- precomputed multiples (multi_exp_stream_with_precompute, multiexp_stream.hpp:27-41) -----

uint32_t b200msm_precompute_window_bits(uint32_t group, size_t n) {
    const GroupOps *ops = ops_for(group);
    if (!ops) return 0;
    // Measured (alt_bn128 G1, 2^24 and 2^20 points): with one bucket set shared by all digits the
    // counting-sort scatter no longer works inside one window's L2-resident slice, and what
    // larger digits save in the accumulation they lose there; the plain MSM's digit size is
    // (near) the optimum here too.
    std::lock_guard<std::mutex> lk(g_mutex);
    if (g_ctxs.empty()) return 0;
    const Split sp(n, active_devices());  // the digit size of one device's share
    return pick_precompute_c(ops, *g_ctxs[0], sp.count(sp.D - 1));
}

size_t b200msm_precompute_digits(uint32_t group, uint32_t c) {
    const GroupOps *ops = ops_for(group);
    if (!ops || c < 2 || c > 24) return 0;
    return (size_t)((ops->scalar_bits + (int)c - 1) / (int)c);  // multiexp_stream.tcc:205
}

bool b200msm_precompute(uint32_t group, uint64_t bases_handle, uint32_t c, uint64_t *table_handle) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    auto b = g_resident.find(bases_handle);
    if (b == g_resident.end() || b->second.kind != RK_BASES || b->second.group != group || !table_handle)
        return set_error("precompute: bad handle"), false;
    const Resident &rb = b->second;
    // one digit size for the whole table: the one a single device would pick for its share
    size_t share = 0;
    for (auto &sh : rb.shards) share = std::max(share, sh.n);
    if (c == 0) c = pick_precompute_c(ops, *g_ctxs[0], share);
    if (c < 2 || c > 24) return set_error("precompute: digit size %u outside [2, 24]", c), false;
    const int W = (ops->scalar_bits + (int)c - 1) / (int)c;
    if ((uint64_t)share * W >= (1ull << 31))
        return set_error("precompute: n * digits = %llu exceeds 2^31 - 1 table entries per device",
                         (unsigned long long)share * W), false;
    Resident r;
    r.group = group, r.kind = RK_MULTIPLES, r.n = rb.n;
    r.shards.resize(rb.shards.size());
    const bool ok = fan_out(rb.shards.size(), [&](size_t k) {
        const Shard &src = rb.shards[k];
        Ctx &ctx = *g_ctxs[src.ci];
        if (!bind(ctx)) return false;
        Shard &s = r.shards[k];
        s.ci = src.ci, s.lo = src.lo, s.n = src.n;
        s.shape = MsmShape{(int)c, W, 1u, (uint32_t)s.n};  // built window-major: table[w * n + i]
        B200_CUDA(cudaMalloc(&s.data, (size_t)W * s.n * ops->affine_bytes));
        B200_CUDA(cudaMalloc((void **)&s.inf, (size_t)W * s.n));  // one flag per multiple
        if (!ops->precompute(ctx, src.data, src.inf, s.n, (int)c, W, s.data, s.inf)) return false;
        B200_CUDA(cudaStreamSynchronize(ctx.stream));
        return true;
    });
    if (!ok) {
        free_resident(r);
        return false;
    }
    *table_handle = g_next_handle++;
    g_resident[*table_handle] = std::move(r);
    return true;
}

bool b200msm_precomputed_upload(uint32_t group, const void *elements, size_t bytes, size_t n,
                                uint32_t c, uint32_t base_form, uint64_t *table_handle) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    if (c < 2 || c > 24) return set_error("precomputed_upload: digit size %u outside [2, 24]", c), false;
    if (base_form > 1) return set_error("invalid base_form %u", base_form), false;
    const int W = (ops->scalar_bits + (int)c - 1) / (int)c;
    if (!table_handle || !elements || n == 0 || bytes != n * (size_t)W * ops->group_bytes)
        return set_error("precomputed_upload: bad arguments (need n * %d elements)", W), false;
    const Split sp(n, active_devices());
    if (sp.count(sp.D - 1) * (size_t)W >= (1ull << 31))
        return set_error("precomputed_upload: too many entries per device"), false;
    Resident r;
    r.group = group, r.kind = RK_MULTIPLES, r.n = n;
    r.shards.resize(sp.D);
    const size_t gb = ops->group_bytes;
    const bool ok = fan_out(sp.D, [&](size_t k) {
        Ctx &ctx = *g_ctxs[k];
        if (!bind(ctx)) return false;
        Shard &s = r.shards[k];
        s.ci = (int)k, s.lo = sp.lo(k), s.n = sp.count(k);
        s.shape = MsmShape{(int)c, W, (uint32_t)W, 1u};  // the reference's stream order: base-major
        const size_t entries = s.n * (size_t)W;
        void *d_raw;
        if (!ctx.reserve(SL_RAW, entries * gb, &d_raw)) return false;
        B200_CUDA(cudaMalloc(&s.data, entries * ops->affine_bytes));
        B200_CUDA(cudaMalloc((void **)&s.inf, entries));  // one flag per multiple
        if (!upload_flat(ctx, d_raw, (const char *)elements + s.lo * (size_t)W * gb, entries * gb, ctx.stream))
            return set_error("precomputed_upload: host to device copy failed"), false;
        if (!ops->ingest(ctx, d_raw, entries, base_form, s.data, s.inf)) return false;
        B200_CUDA(cudaStreamSynchronize(ctx.stream));
        return true;
    });
    if (!ok) {
        free_resident(r);
        return false;
    }
    *table_handle = g_next_handle++;
    g_resident[*table_handle] = std::move(r);
    return true;
}

bool b200msm_multi_exp_precomputed(uint32_t group, uint64_t table_handle, uint64_t scalars_handle,
                                   const void *scalars, size_t scalars_bytes, void *out,
                                   size_t out_bytes) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    auto t = g_resident.find(table_handle);
    if (t == g_resident.end() || t->second.kind != RK_MULTIPLES || t->second.group != group)
        return set_error("multi_exp_precomputed: bad table handle"), false;
    const Resident &rt = t->second;
    const size_t n = rt.n;
    if (out_bytes != ops->group_bytes || !out) return set_error("multi_exp_precomputed: bad out"), false;
    const Resident *rs = nullptr;
    if (scalars_handle != 0) {
        auto sc = g_resident.find(scalars_handle);
        if (sc == g_resident.end() || sc->second.kind != RK_SCALARS || sc->second.group != group ||
            sc->second.n != n || !same_cut(rt, sc->second))
            return set_error("multi_exp_precomputed: bad scalars handle"), false;
        rs = &sc->second;
    } else if (!scalars || scalars_bytes != n * ops->scalar_bytes) {
        return set_error("multi_exp_precomputed: need %zu scalars", n), false;
    }
    const size_t gb = ops->group_bytes, sb = ops->scalar_bytes, D = rt.shards.size();
    std::vector<unsigned char> partials(D * gb);
    const bool ok = fan_out(D, [&](size_t k) {
        const Shard &sh = rt.shards[k];
        Ctx &ctx = *g_ctxs[sh.ci];
        if (!bind(ctx)) return false;
        const void *d_sc = nullptr;
        B200_CUDA(cudaEventRecord(ctx.ev[0], ctx.stream));
        if (rs) {
            d_sc = rs->shards[k].data;
        } else {
            void *d;
            if (!ctx.reserve(SL_SCALARS, sh.n * sb, &d)) return false;
            if (!upload_flat(ctx, d, (const char *)scalars + sh.lo * sb, sh.n * sb, ctx.stream))
                return set_error("multi_exp_precomputed: host to device copy failed"), false;
            d_sc = d;
        }
        return run_msm(ctx, ops, sh.data, sh.inf, d_sc, sh.n, partials.data() + k * gb, &sh.shape);
    });
    if (!ok) return false;
    if (D == 1) {
        memcpy(out, partials.data(), gb);
        return true;
    }
    std::vector<unsigned char> res(gb);
    if (!sum_on(*g_ctxs[0], ops, partials.data(), D, res.data())) return false;
    memcpy(out, res.data(), gb);
    return true;
}

// ---- streamed bases (multi_exp_stream / multi_exp_stream_with_precompute, multiexp_stream.tcc:
// 164-223): the scalars are known up front, the bases arrive in chunks.  One device: the chunks
// arrive one after the other from a single reader. -----------------------------------------------

static size_t stream_element_bytes(const GroupOps *ops, uint32_t encoding) {
    return encoding <= B200MSM_STREAM_MEMORY_NORMAL ? ops->group_bytes : ops->affine_bytes;
}

bool b200msm_stream_begin(uint32_t group, const void *scalars, size_t scalars_bytes, size_t n,
                          uint32_t encoding, uint32_t precompute_c, uint64_t *stream) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    if (encoding > B200MSM_STREAM_BINARY_MONTGOMERY)
        return set_error("stream_begin: unknown encoding %u", encoding), false;
    const uint32_t c = precompute_c & ~(uint32_t)B200MSM_STREAM_REFERENCE_CARRY;
    if (c != 0 && (c < 2 || c > 24))
        return set_error("stream_begin: digit size %u outside [2, 24]", c), false;
    if (!stream || scalars_bytes != n * ops->scalar_bytes || (n > 0 && !scalars))
        return set_error("stream_begin: bad arguments"), false;
    Ctx &ctx = *g_ctxs[0];
    Stream st{group, n};
    st.encoding = encoding;
    st.c = precompute_c;
    st.W = c ? (ops->scalar_bits + (int)c - 1) / (int)c : 1;
    B200_CUDA(cudaMalloc(&st.d_scalars, std::max<size_t>(scalars_bytes, 16)));
    B200_CUDA(cudaMalloc(&st.d_partials, STREAM_MAX_PARTIALS * ops->group_bytes));
    if (n > 0 && (!upload_flat(ctx, st.d_scalars, scalars, scalars_bytes, ctx.stream) ||
                  cudaStreamSynchronize(ctx.stream) != cudaSuccess)) {
        cudaFree(st.d_scalars);
        cudaFree(st.d_partials);
        return set_error("stream_begin: host to device copy failed"), false;
    }
    *stream = g_next_handle++;
    g_streams[*stream] = st;
    return true;
}

bool b200msm_stream_push(uint64_t stream, const void *elements, size_t bytes, size_t count) {
    std::lock_guard<std::mutex> lk(g_mutex);
    auto it = g_streams.find(stream);
    if (it == g_streams.end()) return set_error("stream_push: unknown stream"), false;
    if (!ready()) return false;
    Stream &st = it->second;
    const GroupOps *ops = ops_for(st.group);
    const size_t eb = stream_element_bytes(ops, st.encoding);
    const size_t entries = count * (size_t)st.W;
    if (st.consumed + count > st.n)
        return set_error("stream_push: %zu bases pushed, the stream holds %zu scalars",
                         st.consumed + count, st.n), false;
    if (bytes != entries * eb || (count > 0 && !elements))
        return set_error("stream_push: %zu bytes for %zu x %d elements of %zu bytes", bytes, count,
                         st.W, eb), false;
    if (count == 0) return true;
    if (entries >= (1ull << 31)) return set_error("stream_push: chunk too large"), false;
    Ctx &ctx = *g_ctxs[0];
    void *d_in, *d_aff, *d_inf;
    if (!ctx.reserve(st.encoding <= B200MSM_STREAM_MEMORY_NORMAL ? SL_RAW : SL_STREAM_IN, bytes, &d_in))
        return false;
    if (!ctx.reserve(SL_AFF, entries * ops->affine_bytes, &d_aff)) return false;
    if (!ctx.reserve(SL_INF, entries, &d_inf)) return false;
    if (st.partials == STREAM_MAX_PARTIALS) {  // fold the partial results into one
        if (ctx.tail_pending) {
            B200_CUDA(cudaStreamWaitEvent(ctx.stream, ctx.ev_tail_done, 0));
            ctx.tail_pending = false;
        }
        void *d_tmp;
        if (!ctx.reserve(SL_OUT, ops->group_bytes, &d_tmp)) return false;
        if (!ops->sum(ctx, st.d_partials, st.partials, d_tmp)) return false;
        B200_CUDA(cudaMemcpyAsync(st.d_partials, d_tmp, ops->group_bytes, cudaMemcpyDeviceToDevice,
                                  ctx.stream));
        st.partials = 1;
    }
    B200_CUDA(cudaEventRecord(ctx.ev[0], ctx.stream));
    if (!upload_flat(ctx, d_in, elements, bytes, ctx.stream))
        return set_error("stream_push: host to device copy failed"), false;
    if (st.encoding <= B200MSM_STREAM_MEMORY_NORMAL) {
        if (!ops->ingest(ctx, d_in, entries, st.encoding == B200MSM_STREAM_MEMORY_SPECIAL ? 1u : 0u,
                         d_aff, (uint8_t *)d_inf))
            return false;
    } else {
        if (!ops->stream_decode(ctx, d_in, entries,
                                st.encoding == B200MSM_STREAM_BINARY_MONTGOMERY ? 1u : 0u, d_aff,
                                (uint8_t *)d_inf))
            return false;
    }
    // with precomputed multiples the flags are per stream element (= table entry)
    const uint8_t *d_base_inf = (const uint8_t *)d_inf;
    if (!mark_all(ctx, PH_INGEST + 1)) return false;
    const uint32_t c = st.c & ~(uint32_t)B200MSM_STREAM_REFERENCE_CARRY;
    MsmShape shape{(int)c, st.W, (uint32_t)st.W, 1u};
    shape.drop_top_carry = (st.c & B200MSM_STREAM_REFERENCE_CARRY) != 0;
    // the latency-bound tail of this chunk runs under the kernels of the next one
    ctx.defer_tail = true;
    const bool ok = ops->msm(ctx, d_aff, d_base_inf,
                             (const char *)st.d_scalars + st.consumed * ops->scalar_bytes, count,
                             (char *)st.d_partials + st.partials * ops->group_bytes,
                             c ? &shape : nullptr);
    ctx.defer_tail = false;
    if (!ok) return false;
    st.partials++;
    st.consumed += count;
    return true;
}

bool b200msm_stream_finish(uint64_t stream, void *out, size_t out_bytes) {
    std::lock_guard<std::mutex> lk(g_mutex);
    auto it = g_streams.find(stream);
    if (it == g_streams.end()) return set_error("stream_finish: unknown stream"), false;
    Stream st = it->second;
    const GroupOps *ops = ops_for(st.group);
    bool ok = ready();
    Ctx *ctx = g_ctxs.empty() ? nullptr : g_ctxs[0].get();
    if (ok && ctx->tail_pending) {
        ok = cudaStreamWaitEvent(ctx->stream, ctx->ev_tail_done, 0) == cudaSuccess;
        ctx->tail_pending = false;
    }
    if (ok && (out_bytes != ops->group_bytes || !out)) ok = false, set_error("stream_finish: bad out");
    if (ok && st.consumed != st.n)
        ok = false, set_error("stream_finish: %zu of %zu bases were pushed", st.consumed, st.n);
    if (ok) {
        void *d_out;
        ok = ctx->reserve(SL_OUT, ops->group_bytes, &d_out) &&
             ops->sum(*ctx, st.d_partials, st.partials, d_out) &&
             cudaMemcpyAsync(out, d_out, ops->group_bytes, cudaMemcpyDeviceToHost, ctx->stream) == cudaSuccess &&
             cudaStreamSynchronize(ctx->stream) == cudaSuccess;
        if (!ok && g_error.empty()) set_error("stream_finish: device error");
    }
    if (ctx) cudaStreamSynchronize(ctx->stream);
    cudaFree(st.d_scalars);
    cudaFree(st.d_partials);
    g_streams.erase(it);
    return ok;
}

bool b200msm_release(uint64_t handle) {
    std::lock_guard<std::mutex> lk(g_mutex);
    auto it = g_resident.find(handle);
    if (it == g_resident.end()) return set_error("release: unknown handle"), false;
    free_resident(it->second);
    g_resident.erase(it);
    if (!g_ctxs.empty()) cudaSetDevice(g_ctxs[0]->device);
    return true;
}

bool b200msm_sum(uint32_t group, const void *points, size_t points_bytes, size_t count, void *out,
                 size_t out_bytes) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    if (points_bytes != count * ops->group_bytes || out_bytes != ops->group_bytes || !out ||
        (count > 0 && !points))
        return set_error("sum: buffer sizes do not match count = %zu", count), false;
    return sum_on(*g_ctxs[0], ops, points, count, out);
}

// batch_to_special and batch_exp shard by OUTPUT range with no exchange (multiexp.tcc:898-908)
bool b200msm_batch_to_special(uint32_t group, void *inout, size_t bytes, size_t n) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    if (bytes != n * ops->group_bytes || (n > 0 && !inout))
        return set_error("batch_to_special: buffer size does not match n = %zu", n), false;
    if (n == 0) return true;
    const Split sp(n, active_devices());
    const size_t gb = ops->group_bytes;
    std::vector<void *> d_done(sp.D, nullptr);
    // outputs are written only on success: every device finishes before the first download
    if (!fan_out(sp.D, [&](size_t k) {
            Ctx &ctx = *g_ctxs[k];
            if (!bind(ctx)) return false;
            void *d_raw;
            const size_t cnt = sp.count(k);
            if (!ctx.reserve(SL_RAW, cnt * gb, &d_raw)) return false;
            B200_CUDA(cudaEventRecord(ctx.ev[0], ctx.stream));
            if (!upload_flat(ctx, d_raw, (const char *)inout + sp.lo(k) * gb, cnt * gb, ctx.stream))
                return set_error("batch_to_special: host to device copy failed"), false;
            if (!mark_all(ctx, PH_INGEST + 1)) return false;
            if (!ops->to_special(ctx, d_raw, cnt)) return false;
            if (!mark_all(ctx, PH_ACCUM + 1)) return false;  // the kernels count as "accumulate"
            if (!collect_timings(ctx)) return false;
            d_done[k] = d_raw;
            return true;
        }))
        return false;
    return fan_out(sp.D, [&](size_t k) {
        Ctx &ctx = *g_ctxs[k];
        if (!bind(ctx)) return false;
        return download(ctx, (char *)inout + sp.lo(k) * gb, d_done[k], sp.count(k) * gb, ctx.stream);
    });
}

// ---- fixed-base exponentiation (multiexp.tcc:770-947) -------------------------------------------

namespace {

bool table_shape_ok(size_t scalar_bits, size_t window, size_t num_rows, size_t row_len) {
    return window != 0 && window <= 24 && scalar_bits != 0 && row_len == ((size_t)1 << window) &&
           num_rows == (scalar_bits + window - 1) / window;
}

// out[i] = windowed_exp(table, [coeff *] scalars[i]) for the output range of every device
bool batch_exp_from(const GroupOps *ops, const Resident &rt, const void *scalars, size_t n,
                    const void *coeff, void *out) {
    const Split sp(n, rt.shards.size());
    const size_t gb = ops->group_bytes, sb = ops->scalar_bytes;
    std::vector<void *> d_done(sp.D, nullptr);
    if (!fan_out(sp.D, [&](size_t k) {
            const Shard &sh = rt.shards[k];
            Ctx &ctx = *g_ctxs[sh.ci];
            if (!bind(ctx)) return false;
            const size_t cnt = sp.count(k);
            void *d_sc, *d_out, *d_coeff = nullptr;
            if (!ctx.reserve(SL_SCALARS, cnt * sb, &d_sc)) return false;
            if (!ctx.reserve(SL_BEXP_OUT, cnt * gb, &d_out)) return false;
            B200_CUDA(cudaEventRecord(ctx.ev[0], ctx.stream));
            if (coeff) {
                if (!ctx.reserve(SL_WIRE_MISC, sb, &d_coeff)) return false;
                B200_CUDA(cudaMemcpyAsync(d_coeff, coeff, sb, cudaMemcpyHostToDevice, ctx.stream));
            }
            if (!upload_flat(ctx, d_sc, (const char *)scalars + sp.lo(k) * sb, cnt * sb, ctx.stream))
                return set_error("batch_exp: host to device copy failed"), false;
            if (!mark_all(ctx, PH_INGEST + 1)) return false;
            if (!ops->batch_exp(ctx, sh.data, rt.rows, rt.row_len, (int)rt.window, d_sc, d_coeff, cnt,
                                d_out))
                return false;
            if (!mark_all(ctx, PH_ACCUM + 1)) return false;  // the kernel counts as "accumulate"
            if (!collect_timings(ctx)) return false;
            d_done[k] = d_out;
            return true;
        }))
        return false;
    return fan_out(sp.D, [&](size_t k) {
        Ctx &ctx = *g_ctxs[rt.shards[k].ci];
        if (!bind(ctx)) return false;
        return download(ctx, (char *)out + sp.lo(k) * gb, d_done[k], sp.count(k) * gb, ctx.stream);
    });
}

// uploads (rows != nullptr) or builds (g != nullptr) a window table on every active device
bool make_table(const GroupOps *ops, uint32_t group, size_t scalar_bits, size_t window,
                const void *const *rows, const void *g, size_t num_rows, size_t row_len,
                Resident &r) {
    const size_t D = active_devices();
    const size_t gb = ops->group_bytes, entries = num_rows * row_len, row_bytes = row_len * gb;
    r.group = group, r.kind = RK_WINDOW_TABLE, r.n = entries;
    r.scalar_bits = scalar_bits, r.window = window, r.rows = num_rows, r.row_len = row_len;
    r.shards.resize(D);
    return fan_out(D, [&](size_t k) {
        Ctx &ctx = *g_ctxs[k];
        if (!bind(ctx)) return false;
        Shard &s = r.shards[k];
        s.ci = (int)k, s.lo = 0, s.n = entries;
        void *d_raw;
        if (!ctx.reserve(SL_TABLE_RAW, entries * gb, &d_raw)) return false;
        B200_CUDA(cudaMalloc(&s.data, entries * ops->affine_bytes));
        if (rows) {
            for (size_t q = 0; q < num_rows; q++)
                if (!upload_flat(ctx, (char *)d_raw + q * row_bytes, rows[q], row_bytes, ctx.stream))
                    return set_error("table upload: host to device copy failed"), false;
        } else {
            void *d_g;
            if (!ctx.reserve(SL_WIRE_MISC, gb, &d_g)) return false;
            B200_CUDA(cudaMemcpyAsync(d_g, g, gb, cudaMemcpyHostToDevice, ctx.stream));
            if (!ops->table_build(ctx, d_g, (int)scalar_bits, (int)window, num_rows, row_len, d_raw))
                return false;
        }
        // entries are arbitrary representatives: one batch normalisation, once per table
        if (!ops->table_normalise(ctx, d_raw, entries, s.data)) return false;
        B200_CUDA(cudaStreamSynchronize(ctx.stream));
        return true;
    });
}

}  // namespace

bool b200msm_table_upload(uint32_t group, size_t scalar_bits, size_t window,
                          const void *const *table_rows, size_t num_rows, size_t row_len,
                          uint64_t *handle) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    if (!handle || !table_rows || !table_shape_ok(scalar_bits, window, num_rows, row_len))
        return set_error("table_upload: table shape does not match scalar_bits/window"), false;
    for (size_t q = 0; q < num_rows; q++)
        if (!table_rows[q]) return set_error("table_upload: null table row %zu", q), false;
    Resident r;
    if (!make_table(ops, group, scalar_bits, window, table_rows, nullptr, num_rows, row_len, r)) {
        free_resident(r);
        return false;
    }
    *handle = g_next_handle++;
    g_resident[*handle] = std::move(r);
    return true;
}

bool b200msm_table_build(uint32_t group, size_t scalar_bits, size_t window, const void *g,
                         size_t g_bytes, uint64_t *handle) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    if (!handle || !g || g_bytes != ops->group_bytes || window == 0 || window > 24 || scalar_bits == 0)
        return set_error("table_build: bad arguments"), false;
    const size_t num_rows = (scalar_bits + window - 1) / window, row_len = (size_t)1 << window;
    Resident r;
    if (!make_table(ops, group, scalar_bits, window, nullptr, g, num_rows, row_len, r)) {
        free_resident(r);
        return false;
    }
    *handle = g_next_handle++;
    g_resident[*handle] = std::move(r);
    return true;
}

bool b200msm_table_shape(uint64_t handle, size_t *scalar_bits, size_t *window, size_t *num_rows,
                         size_t *row_len) {
    std::lock_guard<std::mutex> lk(g_mutex);
    auto t = g_resident.find(handle);
    if (t == g_resident.end() || t->second.kind != RK_WINDOW_TABLE)
        return set_error("table_shape: bad handle"), false;
    if (scalar_bits) *scalar_bits = t->second.scalar_bits;
    if (window) *window = t->second.window;
    if (num_rows) *num_rows = t->second.rows;
    if (row_len) *row_len = t->second.row_len;
    return true;
}

bool b200msm_table_download(uint64_t handle, void *const *table_rows, size_t num_rows,
                            size_t row_len) {
    std::lock_guard<std::mutex> lk(g_mutex);
    if (!ready()) return false;
    auto t = g_resident.find(handle);
    if (t == g_resident.end() || t->second.kind != RK_WINDOW_TABLE)
        return set_error("table_download: bad handle"), false;
    const Resident &rt = t->second;
    const GroupOps *ops = ops_for(rt.group);
    if (!table_rows || num_rows != rt.rows || row_len != rt.row_len)
        return set_error("table_download: the table has %zu rows of %zu", rt.rows, rt.row_len), false;
    const Shard &sh = rt.shards[0];
    Ctx &ctx = *g_ctxs[sh.ci];
    if (!bind(ctx)) return false;
    const size_t gb = ops->group_bytes, row_bytes = row_len * gb;
    void *d_raw;
    if (!ctx.reserve(SL_TABLE_RAW, rt.n * gb, &d_raw)) return false;
    if (!ops->table_export(ctx, sh.data, rt.n, d_raw)) return false;
    B200_CUDA(cudaStreamSynchronize(ctx.stream));
    for (size_t q = 0; q < num_rows; q++) {
        if (!table_rows[q]) return set_error("table_download: null table row %zu", q), false;
        B200_CUDA(cudaMemcpy(table_rows[q], (const char *)d_raw + q * row_bytes, row_bytes,
                             cudaMemcpyDeviceToHost));
    }
    return true;
}

bool b200msm_batch_exp_resident(uint32_t group, uint64_t table_handle, const void *scalars,
                                size_t scalars_bytes, size_t n, const void *coeff,
                                size_t coeff_bytes, void *out, size_t out_bytes) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    auto t = g_resident.find(table_handle);
    if (t == g_resident.end() || t->second.kind != RK_WINDOW_TABLE || t->second.group != group)
        return set_error("batch_exp_resident: bad table handle"), false;
    if (scalars_bytes != n * ops->scalar_bytes || out_bytes != n * ops->group_bytes ||
        (n > 0 && (!scalars || !out)) || (coeff && coeff_bytes != ops->scalar_bytes))
        return set_error("batch_exp_resident: buffer sizes do not match n = %zu", n), false;
    if (n == 0) return true;
    return batch_exp_from(ops, t->second, scalars, n, coeff, out);
}

bool b200msm_batch_exp_with_coeff(uint32_t group, size_t scalar_bits, size_t window,
                                  const void *const *table_rows, size_t num_rows, size_t row_len,
                                  const void *coeff, size_t coeff_bytes, const void *scalars,
                                  size_t scalars_bytes, size_t n, void *out, size_t out_bytes) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    if (!table_rows || !table_shape_ok(scalar_bits, window, num_rows, row_len))
        return set_error("batch_exp: table shape does not match scalar_bits/window"), false;
    if (scalars_bytes != n * ops->scalar_bytes || out_bytes != n * ops->group_bytes ||
        (n > 0 && (!scalars || !out)) || (coeff && coeff_bytes != ops->scalar_bytes))
        return set_error("batch_exp: buffer sizes do not match n = %zu", n), false;
    for (size_t q = 0; q < num_rows; q++)
        if (!table_rows[q]) return set_error("batch_exp: null table row %zu", q), false;
    if (n == 0) return true;
    Resident r;
    bool ok = make_table(ops, group, scalar_bits, window, table_rows, nullptr, num_rows, row_len, r) &&
              batch_exp_from(ops, r, scalars, n, coeff, out);
    free_resident(r);
    cudaSetDevice(g_ctxs[0]->device);
    return ok;
}

bool b200msm_batch_exp(uint32_t group, size_t scalar_bits, size_t window,
                       const void *const *table_rows, size_t num_rows, size_t row_len,
                       const void *scalars, size_t scalars_bytes, size_t n, void *out,
                       size_t out_bytes) {
    return b200msm_batch_exp_with_coeff(group, scalar_bits, window, table_rows, num_rows, row_len,
                                        nullptr, 0, scalars, scalars_bytes, n, out, out_bytes);
}

bool b200msm_debug_op(uint32_t group, uint32_t op, const void *a, const void *b, void *out,
                      size_t bytes, size_t n) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    const size_t eb = ops->debug_elem_bytes((int)op);
    if (eb == 0) return set_error("debug_op: unknown op %u", op), false;
    if (bytes != n * eb || (n > 0 && (!a || !out))) return set_error("debug_op: size mismatch"), false;
    if (n == 0) return true;
    Ctx &ctx = *g_ctxs[0];
    void *d_a, *d_b, *d_out;
    if (!ctx.reserve(SL_RAW, bytes, &d_a)) return false;
    if (!ctx.reserve(SL_AFF, bytes, &d_b)) return false;
    if (!ctx.reserve(SL_BEXP_OUT, bytes, &d_out)) return false;
    B200_CUDA(cudaMemcpyAsync(d_a, a, bytes, cudaMemcpyHostToDevice, ctx.stream));
    if (b) B200_CUDA(cudaMemcpyAsync(d_b, b, bytes, cudaMemcpyHostToDevice, ctx.stream));
    if (!ops->debug_op(ctx, (int)op, d_a, d_b, d_out, n)) return false;
    B200_CUDA(cudaStreamSynchronize(ctx.stream));
    B200_CUDA(cudaMemcpy(out, d_out, bytes, cudaMemcpyDeviceToHost));
    return true;
}

bool b200msm_set_option(const char *name, int64_t value) {
    std::lock_guard<std::mutex> lk(g_mutex);
    if (g_ctxs.empty()) return set_error("set_option before init"), false;
    if (!name) return set_error("unknown option (null)"), false;
    if (!strcmp(name, "num_gpus")) {
        if (value < 0 || (size_t)value > g_ctxs.size())
            return set_error("num_gpus: %lld outside [0, %zu bound devices]", (long long)value,
                             g_ctxs.size()), false;
        g_use = (int)value;
        return true;
    }
    struct Opt {
        const char *name;
        int Ctx::*field;
        int64_t lo, hi;
    };
    static const Opt opts[] = {
        {"window_bits", &Ctx::opt_window_bits, 0, 24},
        {"segment", &Ctx::opt_segment, 0, 1 << 20},
        {"window_groups", &Ctx::opt_window_groups, 0, 64},
        {"carry_segment", &Ctx::opt_carry_segment, 0, 1024},
        {"reduce_chunk", &Ctx::opt_reduce_chunk, 0, 12},
        {"reduce_cluster", &Ctx::opt_reduce_cluster, 0, 8},
        {"h2d_growth", &Ctx::opt_h2d_growth, 0, 400},
        {"trace", &Ctx::opt_trace, 0, 1},
        {"h2d_shared_buckets", &Ctx::opt_h2d_shared_buckets, 0, 1},
        {"h2d_parts", &Ctx::opt_h2d_parts, 0, MAX_H2D_PARTS},
        {"h2d_pack", &Ctx::opt_h2d_pack, 0, 2},
        {"stage_lanes", &Ctx::opt_stage_lanes, 1, STAGE_LANES_MAX},
        {"h2d_taper", &Ctx::opt_h2d_taper, 0, 8},
        {"sort_overlap", &Ctx::opt_sort_overlap, 0, 64},
        {"coop_accum", &Ctx::opt_coop_accum, 0, 2},
    };
    for (const Opt &o : opts)
        if (!strcmp(name, o.name)) {
            if (value < o.lo || value > o.hi)
                return set_error("%s: %lld outside [%lld, %lld]", name, (long long)value,
                                 (long long)o.lo, (long long)o.hi), false;
            for (auto &c : g_ctxs) (*c).*(o.field) = (int)value;
            return true;
        }
    return set_error("unknown option %s", name), false;
}

// element-wise maximum over the devices that took part in the most recent call
size_t b200msm_last_timings(float *ms, size_t cap) {
    std::lock_guard<std::mutex> lk(g_mutex);
    bool any = false;
    for (auto &c : g_ctxs) any = any || c->have_timings;
    if (!any) return 0;
    for (size_t p = 0; p < (size_t)PH_COUNT && p < cap; p++) {
        float v = 0;
        for (auto &c : g_ctxs)
            if (c->have_timings) v = std::max(v, c->last_ms[p]);
        ms[p] = v;
    }
    return PH_COUNT;
}

size_t b200msm_device_timings(size_t device_index, float *ms, size_t cap) {
    std::lock_guard<std::mutex> lk(g_mutex);
    if (device_index >= g_ctxs.size() || !g_ctxs[device_index]->have_timings) return 0;
    for (size_t p = 0; p < (size_t)PH_COUNT && p < cap; p++) ms[p] = g_ctxs[device_index]->last_ms[p];
    return PH_COUNT;
}

const char *b200msm_timing_name(size_t phase) {
    static const char *names[PH_COUNT] = {"ingest", "digits", "scan", "scatter", "accumulate",
                                          "carry_rounds", "bucket_reduce", "final", "total"};
    return phase < (size_t)PH_COUNT ? names[phase] : "";
}

bool b200msm_timer_start(void) {
    std::lock_guard<std::mutex> lk(g_mutex);
    if (!ready()) return false;
    B200_CUDA(cudaEventRecord(g_ctxs[0]->ev_timer[0], g_ctxs[0]->stream));
    return true;
}

float b200msm_timer_stop(void) {
    std::lock_guard<std::mutex> lk(g_mutex);
    if (g_ctxs.empty() || cudaSetDevice(g_ctxs[0]->device) != cudaSuccess) return -1.0f;
    Ctx &c = *g_ctxs[0];
    float ms = -1.0f;
    if (cudaEventRecord(c.ev_timer[1], c.stream) != cudaSuccess ||
        cudaEventSynchronize(c.ev_timer[1]) != cudaSuccess ||
        cudaEventElapsedTime(&ms, c.ev_timer[0], c.ev_timer[1]) != cudaSuccess) {
        set_error("timer_stop: %s", cudaGetErrorString(cudaGetLastError()));
        return -1.0f;
    }
    return ms;
}

uint64_t b200msm_launch_count(void) {
    std::lock_guard<std::mutex> lk(g_mutex);
    uint64_t total = 0;
    for (auto &c : g_ctxs) total += c->launches;
    return total;
}

uint32_t b200msm_window_bits(uint32_t group, size_t n) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = ops_for(group);
    if (!ops || g_ctxs.empty()) return 0;
    return ops->window_bits(*g_ctxs[0], n);
}

}  // extern "C"
