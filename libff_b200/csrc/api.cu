// extern "C" boundary of the engine (declared in include/b200msm.h).  Conventions follow the
// reference's ffi layer (ffi/ffi.cpp:16-54,124-260): bool status, exact-size checks, caller-owned
// buffers, outputs written only on success, no exceptions across the boundary.
#include "../../include/b200msm.h"

#include <atomic>
#include <condition_variable>
#include <cstring>
#include <map>
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
            B200_CUDA(cudaStreamSynchronize(stream));
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
            set_error("cudaMalloc of %zu bytes (scratch slot %d) failed: %s", want, (int)s,
                      cudaGetErrorString(e));
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

namespace {

struct Resident {
    uint32_t group;
    bool is_bases;
    size_t n;
    void *data = nullptr;   // affine bases, scalars, or a table of precomputed multiples
    uint8_t *inf = nullptr; // bases and tables
    bool is_table = false;
    MsmShape shape = {0, 0, 1, 0};  // tables only
};

// A streamed MSM in progress (b200msm_stream_*)
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
std::map<uint64_t, Stream> g_streams;

std::mutex g_mutex;
Ctx *g_ctx = nullptr;
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

bool ready() {
    if (!g_ctx) {
        set_error("b200msm_init() has not been called (no CUDA device bound; there is no CPU path)");
        return false;
    }
    B200_CUDA(cudaSetDevice(g_ctx->device));
    return true;
}

const GroupOps *checked_ops(uint32_t group) {
    const GroupOps *ops = ops_for(group);
    if (!ops) set_error("unknown group id %u", group);
    return ops;
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
// overlaps the computation of range k (each range is an independent MSM; the partial results are
// added at the end exactly like the reference's per-chunk partials, multiexp.tcc:681-685).
size_t pick_parts(const Ctx &ctx, size_t n) {
    if (ctx.opt_h2d_parts > 0) return std::min<size_t>((size_t)ctx.opt_h2d_parts, std::max<size_t>(n, 1));
    if (n < ((size_t)1 << 21)) return 1;
    if (n < ((size_t)1 << 22)) return 2;
    if (n < ((size_t)1 << 23)) return 3;
    // shared buckets: no per-range reduction, so more and shorter ranges pay (measured at 2^24:
    // 12 ranges growing by 1.2x, profiles/e2e_sweep_r01e.log)
    return ctx.opt_h2d_shared_buckets ? 12 : 5;
}

// Range boundaries of a pipelined call: sizes grow geometrically (x h2d_growth %, default 1.3),
// so the first upload -- the only one nothing overlaps -- is short, while the later, longer MSMs
// still finish after their successor's upload.  bounds has parts + 1 entries.
void pick_bounds(const Ctx &ctx, size_t n, size_t parts, size_t *bounds) {
    const double g = ctx.opt_h2d_growth > 0 ? ctx.opt_h2d_growth / 100.0
                     : (ctx.opt_h2d_shared_buckets && parts > 5) ? 1.2 : 1.3;
    double total = 0, w = 1;
    for (size_t k = 0; k < parts; k++, w *= g) total += w;
    double acc = 0;
    w = 1;
    bounds[0] = 0;
    for (size_t k = 0; k + 1 < parts; k++, w *= g) {
        acc += w;
        size_t b = (size_t)((double)n * (acc / total));
        if (n >= ((size_t)1 << 16)) b &= ~(size_t)255;  // 16-byte aligned ranges in all arrays
        bounds[k + 1] = std::min(std::max(b, bounds[k]), n);
    }
    bounds[parts] = n;
}

// ---- host -> device uploads ------------------------------------------------------------------
// Pinned (cudaHostAlloc / cudaHostRegister) sources go straight to cudaMemcpyAsync.  Pageable
// sources -- what a std::vector<G> of a libff caller is -- would be staged by the driver at
// ~10 GB/s; instead STAGE_LANES host threads copy slices into the engine's own pinned ring
// buffers and issue the DMA from there, which keeps the PCIe link busy.
constexpr int STAGE_LANES = 4;
constexpr size_t STAGE_BYTES = (size_t)8 << 20;

struct Stager {
    void *buf[STAGE_LANES][2] = {};
    cudaEvent_t ev[STAGE_LANES][2] = {};
    bool ready = false;
    bool init() {
        if (ready) return true;
        for (int l = 0; l < STAGE_LANES; l++)
            for (int b = 0; b < 2; b++) {
                B200_CUDA(cudaHostAlloc(&buf[l][b], STAGE_BYTES, cudaHostAllocDefault));
                B200_CUDA(cudaEventCreateWithFlags(&ev[l][b], cudaEventDisableTiming));
            }
        ready = true;
        return true;
    }
    void release() {
        for (int l = 0; l < STAGE_LANES; l++)
            for (int b = 0; b < 2; b++) {
                if (buf[l][b]) cudaFreeHost(buf[l][b]);
                if (ev[l][b]) cudaEventDestroy(ev[l][b]);
                buf[l][b] = nullptr;
                ev[l][b] = nullptr;
            }
        ready = false;
    }
};
Stager g_stager;

bool is_pinned(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        (void)cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

// Queues the upload of `bytes` bytes on stream `cs`; returns once every DMA has been issued
// (pageable sources: once the last slice has been copied into a staging buffer).
bool upload(int device, void *dst, const void *src, size_t bytes, cudaStream_t cs) {
    if (bytes == 0) return true;
    if (is_pinned(src) || bytes < STAGE_BYTES)
        return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, cs) == cudaSuccess;
    const size_t slices = (bytes + STAGE_BYTES - 1) / STAGE_BYTES;
    std::atomic<bool> ok{true};
    std::thread lanes[STAGE_LANES];
    for (int l = 0; l < STAGE_LANES; l++) {
        lanes[l] = std::thread([&, l] {
            if (cudaSetDevice(device) != cudaSuccess) ok = false;
            int b = 0;
            for (size_t s = l; s < slices && ok; s += STAGE_LANES, b ^= 1) {
                const size_t off = s * STAGE_BYTES, len = std::min(STAGE_BYTES, bytes - off);
                // the previous DMA out of this buffer must have finished
                if (cudaEventSynchronize(g_stager.ev[l][b]) != cudaSuccess) ok = false;
                memcpy(g_stager.buf[l][b], (const char *)src + off, len);
                if (cudaMemcpyAsync((char *)dst + off, g_stager.buf[l][b], len, cudaMemcpyHostToDevice,
                                    cs) != cudaSuccess ||
                    cudaEventRecord(g_stager.ev[l][b], cs) != cudaSuccess)
                    ok = false;
            }
        });
    }
    for (auto &t : lanes) t.join();
    return ok;
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

}  // namespace
}  // namespace b200msm

using namespace b200msm;

extern "C" {

bool b200msm_init(int device) {
    std::lock_guard<std::mutex> lk(g_mutex);
    if (g_ctx) {
        if (g_ctx->device == device) return true;
        set_error("already initialised on device %d", g_ctx->device);
        return false;
    }
    int count = 0;
    B200_CUDA(cudaGetDeviceCount(&count));
    if (device < 0 || device >= count) {
        set_error("device %d not available (%d CUDA devices); there is no CPU path", device, count);
        return false;
    }
    B200_CUDA(cudaSetDevice(device));
    Ctx *ctx = new Ctx();
    ctx->device = device;
    B200_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    B200_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    for (auto &e : ctx->ev) B200_CUDA(cudaEventCreate(&e));
    for (auto &e : ctx->ev_copy) B200_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    B200_CUDA(cudaEventCreateWithFlags(&ctx->ev_sync, cudaEventDisableTiming));
    int prio_lo = 0, prio_hi = 0;
    B200_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    B200_CUDA(cudaStreamCreateWithPriority(&ctx->tail_stream, cudaStreamNonBlocking, prio_hi));
    B200_CUDA(cudaStreamCreateWithFlags(&ctx->acc_stream, cudaStreamNonBlocking));
    B200_CUDA(cudaEventCreateWithFlags(&ctx->ev_group_go, cudaEventDisableTiming));
    for (auto &e : ctx->ev_group) B200_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto &e : ctx->ev_timer) B200_CUDA(cudaEventCreate(&e));
    B200_CUDA(cudaEventCreateWithFlags(&ctx->ev_tail_go, cudaEventDisableTiming));
    B200_CUDA(cudaEventCreateWithFlags(&ctx->ev_tail_done, cudaEventDisableTiming));
    g_ctx = ctx;
    return true;
}

void b200msm_shutdown(void) {
    std::lock_guard<std::mutex> lk(g_mutex);
    if (!g_ctx) return;
    cudaSetDevice(g_ctx->device);
    cudaDeviceSynchronize();
    for (auto &kv : g_resident) {
        cudaFree(kv.second.data);
        if (kv.second.inf) cudaFree(kv.second.inf);
    }
    g_resident.clear();
    for (auto &kv : g_streams) {
        cudaFree(kv.second.d_scalars);
        cudaFree(kv.second.d_partials);
    }
    g_streams.clear();
    cudaStreamSynchronize(g_ctx->tail_stream);
    g_ctx->release_all();
    g_stager.release();
    for (auto &e : g_ctx->ev) cudaEventDestroy(e);
    for (auto &e : g_ctx->ev_copy) cudaEventDestroy(e);
    cudaEventDestroy(g_ctx->ev_sync);
    for (auto &e : g_ctx->ev_timer) cudaEventDestroy(e);
    cudaEventDestroy(g_ctx->ev_tail_go);
    cudaEventDestroy(g_ctx->ev_tail_done);
    cudaStreamDestroy(g_ctx->tail_stream);
    cudaStreamDestroy(g_ctx->acc_stream);
    cudaEventDestroy(g_ctx->ev_group_go);
    for (auto &e : g_ctx->ev_group) cudaEventDestroy(e);
    cudaStreamDestroy(g_ctx->copy_stream);
    cudaStreamDestroy(g_ctx->stream);
    delete g_ctx;
    g_ctx = nullptr;
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
    Ctx &ctx = *g_ctx;
    void *d_raw, *d_aff, *d_inf, *d_sc, *d_out, *d_parts;
    if (!ctx.reserve(SL_RAW, bases_bytes, &d_raw)) return false;
    if (!ctx.reserve(SL_AFF, n * ops->affine_bytes, &d_aff)) return false;
    if (!ctx.reserve(SL_INF, n, &d_inf)) return false;
    if (!ctx.reserve(SL_SCALARS, scalars_bytes, &d_sc)) return false;
    if (!ctx.reserve(SL_OUT, ops->group_bytes, &d_out)) return false;
    size_t parts = pick_parts(ctx, n);
    size_t bounds[MAX_H2D_PARTS + 1] = {0};
    if (parts > 1) {  // ranges that came out empty (tiny n) are dropped
        pick_bounds(ctx, n, parts, bounds);
        size_t m = 0;
        for (size_t k = 0; k < parts; k++)
            if (bounds[k + 1] > bounds[k]) bounds[++m] = bounds[k + 1];
        parts = m;
    }
    if (!ctx.reserve(SL_PARTIALS, std::max<size_t>(parts, 1) * ops->group_bytes, &d_parts)) return false;
    B200_CUDA(cudaEventRecord(ctx.ev[0], ctx.stream));
    if (parts <= 1) {
        if (n > 0) {
            if (!g_stager.init()) return false;
            if (!upload(ctx.device, d_raw, bases, bases_bytes, ctx.stream) ||
                !upload(ctx.device, d_sc, scalars, scalars_bytes, ctx.stream))
                return set_error("multi_exp: host to device copy failed"), false;
            if (!ops->ingest(ctx, d_raw, n, base_form, d_aff, (uint8_t *)d_inf)) return false;
        }
        return run_msm(ctx, ops, d_aff, (const uint8_t *)d_inf, d_sc, n, out);
    }
    // ---- pipelined: upload of range k + 1 (copy stream) under the MSM of range k
    const size_t gb = ops->group_bytes, sb = ops->scalar_bytes, ab = ops->affine_bytes;
    auto lo_of = [&](size_t k) { return bounds[k]; };
    auto hi_of = [&](size_t k) { return bounds[k + 1]; };
    // the compute stream must not overwrite the staging buffers of a previous call that is
    // still in flight: the copy stream starts after everything queued so far
    B200_CUDA(cudaEventRecord(ctx.ev_sync, ctx.stream));
    B200_CUDA(cudaStreamWaitEvent(ctx.copy_stream, ctx.ev_sync, 0));
    if (!g_stager.init()) return false;
    Uploader up;
    const int device = ctx.device;
    cudaStream_t cs = ctx.copy_stream;
    cudaEvent_t *evs = ctx.ev_copy;
    up.th = std::thread([&, device, cs, evs] {
        bool ok = cudaSetDevice(device) == cudaSuccess;
        for (size_t k = 0; k < parts && ok; k++) {
            const size_t lo = lo_of(k), cnt = hi_of(k) - lo;
            ok = upload(device, (char *)d_sc + lo * sb, (const char *)scalars + lo * sb, cnt * sb, cs) &&
                 upload(device, (char *)d_raw + lo * gb, (const char *)bases + lo * gb, cnt * gb, cs) &&
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
    for (size_t k = 0; k < parts; k++) {
        const size_t lo = lo_of(k), cnt = hi_of(k) - lo;
        ctx.range_first = k == 0;
        ctx.range_last = k + 1 == parts;
        up.wait_for((int)k);
        if (up.failed) return set_error("multi_exp: host to device copy failed"), false;
        B200_CUDA(cudaStreamWaitEvent(ctx.stream, ctx.ev_copy[k], 0));
        if (!ops->ingest(ctx, (char *)d_raw + lo * gb, cnt, base_form, (char *)d_aff + lo * ab,
                         (uint8_t *)d_inf + lo))
            return false;
        B200_CUDA(cudaEventRecord(ctx.ev[PH_INGEST + 1], ctx.stream));
        if (!ops->msm(ctx, (char *)d_aff + lo * ab, (const uint8_t *)d_inf + lo, (char *)d_sc + lo * sb,
                      cnt, (char *)d_parts + k * gb, nullptr))
            return false;
        if (ctx.opt_trace) {  // development aid: per-range phase times (serialises the pipeline)
            B200_CUDA(cudaDeviceSynchronize());
            fprintf(stderr, "[b200msm] range %zu: %zu points", k, cnt);
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
    if (flags > 3) return set_error("multi_exp_wire: unknown flags %u", flags), false;
    if (points_bytes != n * pb || scalars_bytes != n * sb || out_bytes != pb)
        return set_error("multi_exp_wire: buffer sizes do not match n = %zu", n), false;
    if (!out || (n > 0 && (!points || !scalars))) return set_error("multi_exp_wire: null buffer"), false;
    Ctx &ctx = *g_ctx;
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
    if (!g_stager.init()) return false;
    if (!upload(ctx.device, d_wp, points, points_bytes, ctx.stream) ||
        !upload(ctx.device, d_ws, scalars, scalars_bytes, ctx.stream))
        return set_error("multi_exp_wire: host to device copy failed"), false;
    if (!ops->wire_decode(ctx, d_wp, d_ws, n, flags, d_aff, (uint8_t *)d_inf, d_sc, d_err)) return false;
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
    Ctx &ctx = *g_ctx;
    Resident r{group, true, n};
    void *d_raw;
    if (!ctx.reserve(SL_RAW, bases_bytes, &d_raw)) return false;
    B200_CUDA(cudaMalloc(&r.data, n * ops->affine_bytes));
    B200_CUDA(cudaMalloc((void **)&r.inf, n));
    if (!g_stager.init()) return false;
    bool ok = upload(ctx.device, d_raw, bases, bases_bytes, ctx.stream) &&
              ops->ingest(ctx, d_raw, n, base_form, r.data, r.inf);
    if (ok && cudaStreamSynchronize(ctx.stream) != cudaSuccess) ok = false, set_error("ingest failed");
    if (!ok) {
        cudaFree(r.data);
        cudaFree(r.inf);
        return false;
    }
    *handle = g_next_handle++;
    g_resident[*handle] = r;
    return true;
}

bool b200msm_scalars_upload(uint32_t group, const void *scalars, size_t scalars_bytes, size_t n,
                            uint64_t *handle) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    if (!handle || !scalars || n == 0 || scalars_bytes != n * ops->scalar_bytes)
        return set_error("scalars_upload: bad arguments"), false;
    Resident r{group, false, n};
    B200_CUDA(cudaMalloc(&r.data, scalars_bytes));
    B200_CUDA(cudaMemcpy(r.data, scalars, scalars_bytes, cudaMemcpyHostToDevice));
    *handle = g_next_handle++;
    g_resident[*handle] = r;
    return true;
}

bool b200msm_multi_exp_resident(uint32_t group, uint64_t bases_handle, uint64_t scalars_handle,
                                size_t first, size_t n, void *out, size_t out_bytes) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    auto b = g_resident.find(bases_handle), s = g_resident.find(scalars_handle);
    if (b == g_resident.end() || s == g_resident.end() || !b->second.is_bases || s->second.is_bases ||
        s->second.is_table ||
        b->second.group != group || s->second.group != group)
        return set_error("multi_exp_resident: bad handle"), false;
    if (first + n > b->second.n || first + n > s->second.n || out_bytes != ops->group_bytes || !out)
        return set_error("multi_exp_resident: range or size mismatch"), false;
    Ctx &ctx = *g_ctx;
    B200_CUDA(cudaEventRecord(ctx.ev[0], ctx.stream));
    return run_msm(ctx, ops, (const char *)b->second.data + first * ops->affine_bytes,
                   b->second.inf + first, (const char *)s->second.data + first * ops->scalar_bytes, n,
                   out);
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
    if (!g_ctx) return 0;
    uint32_t c = ops->window_bits(*g_ctx, n);
    while (c < 24 && (uint64_t)n * ((ops->scalar_bits + c - 1) / c) >= (1ull << 31)) c++;
    return c;
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
    if (b == g_resident.end() || !b->second.is_bases || b->second.group != group || !table_handle)
        return set_error("precompute: bad handle"), false;
    const size_t n = b->second.n;
    if (c == 0) {
        c = ops->window_bits(*g_ctx, n);
        while (c < 24 && (uint64_t)n * ((ops->scalar_bits + c - 1) / c) >= (1ull << 31)) c++;
    }
    if (c < 2 || c > 24) return set_error("precompute: digit size %u outside [2, 24]", c), false;
    const int W = (ops->scalar_bits + (int)c - 1) / (int)c;
    if ((uint64_t)n * W >= (1ull << 31))
        return set_error("precompute: n * digits = %llu exceeds 2^31 - 1 table entries",
                         (unsigned long long)n * W), false;
    Ctx &ctx = *g_ctx;
    Resident r{group, false, n};
    r.is_table = true;
    r.shape = MsmShape{(int)c, W, 1u, (uint32_t)n};  // built window-major: table[w * n + i]
    B200_CUDA(cudaMalloc(&r.data, (size_t)W * n * ops->affine_bytes));
    B200_CUDA(cudaMalloc((void **)&r.inf, n));
    bool ok = cudaMemcpyAsync(r.inf, b->second.inf, n, cudaMemcpyDeviceToDevice, ctx.stream) == cudaSuccess &&
              ops->precompute(ctx, b->second.data, b->second.inf, n, (int)c, W, r.data);
    if (ok && cudaStreamSynchronize(ctx.stream) != cudaSuccess) ok = false, set_error("precompute failed");
    if (!ok) {
        cudaFree(r.data);
        cudaFree(r.inf);
        return false;
    }
    *table_handle = g_next_handle++;
    g_resident[*table_handle] = r;
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
    const size_t entries = n * (size_t)W;
    if (!table_handle || !elements || n == 0 || bytes != entries * ops->group_bytes)
        return set_error("precomputed_upload: bad arguments (need n * %d elements)", W), false;
    if (entries >= (1ull << 31)) return set_error("precomputed_upload: too many entries"), false;
    Ctx &ctx = *g_ctx;
    Resident r{group, false, n};
    r.is_table = true;
    r.shape = MsmShape{(int)c, W, (uint32_t)W, 1u};  // the reference's stream order: base-major
    void *d_raw;
    uint8_t *d_inf_all;
    if (!ctx.reserve(SL_RAW, bytes, &d_raw)) return false;
    if (!ctx.reserve(SL_TABLE_INF, entries, (void **)&d_inf_all)) return false;
    B200_CUDA(cudaMalloc(&r.data, entries * ops->affine_bytes));
    B200_CUDA(cudaMalloc((void **)&r.inf, n));
    if (!g_stager.init()) return false;
    bool ok = upload(ctx.device, d_raw, elements, bytes, ctx.stream) &&
              ops->ingest(ctx, d_raw, entries, base_form, r.data, d_inf_all) &&
              // a base is skipped when its first element is zero (all its multiples are)
              cudaMemcpy2DAsync(r.inf, 1, d_inf_all, (size_t)W, 1, n, cudaMemcpyDeviceToDevice,
                                ctx.stream) == cudaSuccess;
    if (ok && cudaStreamSynchronize(ctx.stream) != cudaSuccess) ok = false, set_error("ingest failed");
    if (!ok) {
        cudaFree(r.data);
        cudaFree(r.inf);
        return false;
    }
    *table_handle = g_next_handle++;
    g_resident[*table_handle] = r;
    return true;
}

bool b200msm_multi_exp_precomputed(uint32_t group, uint64_t table_handle, uint64_t scalars_handle,
                                   const void *scalars, size_t scalars_bytes, void *out,
                                   size_t out_bytes) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    auto t = g_resident.find(table_handle);
    if (t == g_resident.end() || !t->second.is_table || t->second.group != group)
        return set_error("multi_exp_precomputed: bad table handle"), false;
    const size_t n = t->second.n;
    if (out_bytes != ops->group_bytes || !out) return set_error("multi_exp_precomputed: bad out"), false;
    Ctx &ctx = *g_ctx;
    const void *d_sc = nullptr;
    B200_CUDA(cudaEventRecord(ctx.ev[0], ctx.stream));
    if (scalars_handle != 0) {
        auto sc = g_resident.find(scalars_handle);
        if (sc == g_resident.end() || sc->second.is_bases || sc->second.is_table ||
            sc->second.group != group || sc->second.n != n)
            return set_error("multi_exp_precomputed: bad scalars handle"), false;
        d_sc = sc->second.data;
    } else {
        if (!scalars || scalars_bytes != n * ops->scalar_bytes)
            return set_error("multi_exp_precomputed: need %zu scalars", n), false;
        void *d;
        if (!ctx.reserve(SL_SCALARS, scalars_bytes, &d)) return false;
        if (!g_stager.init()) return false;
        if (!upload(ctx.device, d, scalars, scalars_bytes, ctx.stream))
            return set_error("multi_exp_precomputed: host to device copy failed"), false;
        d_sc = d;
    }
    return run_msm(ctx, ops, t->second.data, t->second.inf, d_sc, n, out, &t->second.shape);
}

// ---- streamed bases (multi_exp_stream / multi_exp_stream_with_precompute, multiexp_stream.tcc:
// 164-223): the scalars are known up front, the bases arrive in chunks ------------------------

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
    if (precompute_c != 0 && (precompute_c < 2 || precompute_c > 24))
        return set_error("stream_begin: digit size %u outside [2, 24]", precompute_c), false;
    if (!stream || scalars_bytes != n * ops->scalar_bytes || (n > 0 && !scalars))
        return set_error("stream_begin: bad arguments"), false;
    Ctx &ctx = *g_ctx;
    Stream st{group, n};
    st.encoding = encoding;
    st.c = precompute_c;
    st.W = precompute_c ? (ops->scalar_bits + (int)precompute_c - 1) / (int)precompute_c : 1;
    B200_CUDA(cudaMalloc(&st.d_scalars, std::max<size_t>(scalars_bytes, 16)));
    B200_CUDA(cudaMalloc(&st.d_partials, STREAM_MAX_PARTIALS * ops->group_bytes));
    if (!g_stager.init()) return false;
    if (n > 0 && (!upload(ctx.device, st.d_scalars, scalars, scalars_bytes, ctx.stream) ||
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
    Ctx &ctx = *g_ctx;
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
    if (!upload(ctx.device, d_in, elements, bytes, ctx.stream))
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
    const uint8_t *d_base_inf = (const uint8_t *)d_inf;
    if (st.W > 1) {  // per-base flag = flag of the base's first element
        void *d_bi;
        if (!ctx.reserve(SL_TABLE_INF, count, &d_bi)) return false;
        B200_CUDA(cudaMemcpy2DAsync(d_bi, 1, d_inf, (size_t)st.W, 1, count, cudaMemcpyDeviceToDevice,
                                    ctx.stream));
        d_base_inf = (const uint8_t *)d_bi;
    }
    if (!mark_all(ctx, PH_INGEST + 1)) return false;
    const MsmShape shape{(int)st.c, st.W, (uint32_t)st.W, 1u};
    // the latency-bound tail of this chunk runs under the kernels of the next one
    ctx.defer_tail = true;
    const bool ok = ops->msm(ctx, d_aff, d_base_inf,
                             (const char *)st.d_scalars + st.consumed * ops->scalar_bytes, count,
                             (char *)st.d_partials + st.partials * ops->group_bytes,
                             st.c ? &shape : nullptr);
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
    if (ok && g_ctx->tail_pending) {
        ok = cudaStreamWaitEvent(g_ctx->stream, g_ctx->ev_tail_done, 0) == cudaSuccess;
        g_ctx->tail_pending = false;
    }
    if (ok && (out_bytes != ops->group_bytes || !out)) ok = false, set_error("stream_finish: bad out");
    if (ok && st.consumed != st.n)
        ok = false, set_error("stream_finish: %zu of %zu bases were pushed", st.consumed, st.n);
    if (ok) {
        Ctx &ctx = *g_ctx;
        void *d_out;
        ok = ctx.reserve(SL_OUT, ops->group_bytes, &d_out) &&
             ops->sum(ctx, st.d_partials, st.partials, d_out) &&
             cudaMemcpyAsync(out, d_out, ops->group_bytes, cudaMemcpyDeviceToHost, ctx.stream) == cudaSuccess &&
             cudaStreamSynchronize(ctx.stream) == cudaSuccess;
        if (!ok && g_error.empty()) set_error("stream_finish: device error");
    }
    if (g_ctx) cudaStreamSynchronize(g_ctx->stream);
    cudaFree(st.d_scalars);
    cudaFree(st.d_partials);
    g_streams.erase(it);
    return ok;
}

bool b200msm_release(uint64_t handle) {
    std::lock_guard<std::mutex> lk(g_mutex);
    auto it = g_resident.find(handle);
    if (it == g_resident.end()) return set_error("release: unknown handle"), false;
    if (g_ctx) cudaSetDevice(g_ctx->device);
    cudaFree(it->second.data);
    if (it->second.inf) cudaFree(it->second.inf);
    g_resident.erase(it);
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
    Ctx &ctx = *g_ctx;
    void *d_raw, *d_out;
    if (!ctx.reserve(SL_RAW, points_bytes, &d_raw)) return false;
    if (!ctx.reserve(SL_OUT, ops->group_bytes, &d_out)) return false;
    if (count) B200_CUDA(cudaMemcpyAsync(d_raw, points, points_bytes, cudaMemcpyHostToDevice, ctx.stream));
    if (!ops->sum(ctx, d_raw, count, d_out)) return false;
    std::vector<unsigned char> tmp(ops->group_bytes);
    B200_CUDA(cudaMemcpyAsync(tmp.data(), d_out, ops->group_bytes, cudaMemcpyDeviceToHost, ctx.stream));
    B200_CUDA(cudaStreamSynchronize(ctx.stream));
    memcpy(out, tmp.data(), ops->group_bytes);
    return true;
}

bool b200msm_batch_to_special(uint32_t group, void *inout, size_t bytes, size_t n) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    if (bytes != n * ops->group_bytes || (n > 0 && !inout))
        return set_error("batch_to_special: buffer size does not match n = %zu", n), false;
    if (n == 0) return true;
    Ctx &ctx = *g_ctx;
    void *d_raw;
    if (!ctx.reserve(SL_RAW, bytes, &d_raw)) return false;
    B200_CUDA(cudaMemcpyAsync(d_raw, inout, bytes, cudaMemcpyHostToDevice, ctx.stream));
    if (!ops->to_special(ctx, d_raw, n)) return false;
    // outputs are written only on success: the download is the last step
    B200_CUDA(cudaStreamSynchronize(ctx.stream));
    B200_CUDA(cudaMemcpy(inout, d_raw, bytes, cudaMemcpyDeviceToHost));
    return true;
}

bool b200msm_batch_exp(uint32_t group, size_t scalar_bits, size_t window,
                       const void *const *table_rows, size_t num_rows, size_t row_len,
                       const void *scalars, size_t scalars_bytes, size_t n, void *out,
                       size_t out_bytes) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = checked_ops(group);
    if (!ops || !ready()) return false;
    if (window == 0 || window > 24 || row_len != ((size_t)1 << window) ||
        num_rows != (scalar_bits + window - 1) / window || num_rows == 0 || !table_rows)
        return set_error("batch_exp: table shape does not match scalar_bits/window"), false;
    if (scalars_bytes != n * ops->scalar_bytes || out_bytes != n * ops->group_bytes ||
        (n > 0 && (!scalars || !out)))
        return set_error("batch_exp: buffer sizes do not match n = %zu", n), false;
    if (n == 0) return true;
    Ctx &ctx = *g_ctx;
    const size_t row_bytes = row_len * ops->group_bytes;
    void *d_table, *d_sc, *d_out;
    if (!ctx.reserve(SL_TABLE_RAW, num_rows * row_bytes, &d_table)) return false;
    if (!ctx.reserve(SL_SCALARS, scalars_bytes, &d_sc)) return false;
    if (!ctx.reserve(SL_BEXP_OUT, out_bytes, &d_out)) return false;
    for (size_t r = 0; r < num_rows; r++) {
        if (!table_rows[r]) return set_error("batch_exp: null table row %zu", r), false;
        B200_CUDA(cudaMemcpyAsync((char *)d_table + r * row_bytes, table_rows[r], row_bytes,
                                  cudaMemcpyHostToDevice, ctx.stream));
    }
    B200_CUDA(cudaMemcpyAsync(d_sc, scalars, scalars_bytes, cudaMemcpyHostToDevice, ctx.stream));
    if (!ops->batch_exp(ctx, d_table, num_rows, row_len, (int)window, d_sc, n, d_out)) return false;
    B200_CUDA(cudaStreamSynchronize(ctx.stream));
    B200_CUDA(cudaMemcpy(out, d_out, out_bytes, cudaMemcpyDeviceToHost));
    return true;
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
    Ctx &ctx = *g_ctx;
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
    if (!g_ctx) return set_error("set_option before init"), false;
    if (name && !strcmp(name, "window_bits")) return g_ctx->opt_window_bits = (int)value, true;
    if (name && !strcmp(name, "segment")) return g_ctx->opt_segment = (int)value, true;
    if (name && !strcmp(name, "window_groups")) {
        if (value < 0 || value > 64) return set_error("window_groups out of range"), false;
        return g_ctx->opt_window_groups = (int)value, true;
    }
    if (name && !strcmp(name, "carry_segment")) {
        if (value < 0 || value > 1024) return set_error("carry_segment out of range"), false;
        return g_ctx->opt_carry_segment = (int)value, true;
    }
    if (name && !strcmp(name, "reduce_chunk")) return g_ctx->opt_reduce_chunk = (int)value, true;
    if (name && !strcmp(name, "h2d_growth")) {
        if (value < 0 || value > 400) return set_error("h2d_growth out of range (percent)"), false;
        return g_ctx->opt_h2d_growth = (int)value, true;
    }
    if (name && !strcmp(name, "trace")) return g_ctx->opt_trace = value != 0, true;
    if (name && !strcmp(name, "h2d_shared_buckets")) return g_ctx->opt_h2d_shared_buckets = value != 0, true;
    if (name && !strcmp(name, "h2d_parts")) {
        if (value < 0 || value > MAX_H2D_PARTS) return set_error("h2d_parts out of range"), false;
        return g_ctx->opt_h2d_parts = (int)value, true;
    }
    return set_error("unknown option %s", name ? name : "(null)"), false;
}

size_t b200msm_last_timings(float *ms, size_t cap) {
    std::lock_guard<std::mutex> lk(g_mutex);
    if (!g_ctx || !g_ctx->have_timings) return 0;
    for (size_t p = 0; p < (size_t)PH_COUNT && p < cap; p++) ms[p] = g_ctx->last_ms[p];
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
    B200_CUDA(cudaEventRecord(g_ctx->ev_timer[0], g_ctx->stream));
    return true;
}

float b200msm_timer_stop(void) {
    std::lock_guard<std::mutex> lk(g_mutex);
    if (!g_ctx || cudaSetDevice(g_ctx->device) != cudaSuccess) return -1.0f;
    float ms = -1.0f;
    if (cudaEventRecord(g_ctx->ev_timer[1], g_ctx->stream) != cudaSuccess ||
        cudaEventSynchronize(g_ctx->ev_timer[1]) != cudaSuccess ||
        cudaEventElapsedTime(&ms, g_ctx->ev_timer[0], g_ctx->ev_timer[1]) != cudaSuccess) {
        set_error("timer_stop: %s", cudaGetErrorString(cudaGetLastError()));
        return -1.0f;
    }
    return ms;
}

uint64_t b200msm_launch_count(void) {
    std::lock_guard<std::mutex> lk(g_mutex);
    return g_ctx ? g_ctx->launches : 0;
}

uint32_t b200msm_window_bits(uint32_t group, size_t n) {
    std::lock_guard<std::mutex> lk(g_mutex);
    const GroupOps *ops = ops_for(group);
    if (!ops || !g_ctx) return 0;
    return ops->window_bits(*g_ctx, n);
}

}  // extern "C"
