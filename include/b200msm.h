/* b200msm -- C ABI of the B200-native multi-scalar-multiplication engine.
 *
 * This is the drop-in boundary for the MSM hot path of clearmatics/libff.  libff has no runtime
 * plugin interface; the path's extension point is the compile-time dispatch
 *   libff::multi_exp<G, Fr, Method, Form>()                 multiexp.tcc:643-688 (decl multiexp.hpp:63-73)
 *     -> internal::multi_exp_implementation<...>::multi_exp_inner   multiexp.tcc:198-212, 276-381, 507-633
 *   libff::batch_exp<G, Fr>()                               multiexp.tcc:875-915 (windowed_exp :848-873)
 *   libff::batch_to_special<G>()                            multiexp.tcc:949-974
 * (paths relative to libff/algebra/scalar_multiplication/).  The header
 * libff_b200/include/libff_b200.hpp specialises exactly those templates and forwards to the
 * functions below, so existing call sites stay source-identical.
 *
 * Conventions follow libff's own C ABI (ffi/ffi.h:13-95, ffi/ffi.cpp:16-54,124-260): every
 * function returns `bool` (false on any size mismatch, invalid argument or CUDA error), buffers
 * are caller-owned (pointer, byte-size) pairs whose sizes must match exactly, outputs are written
 * only on success, nothing throws across the boundary.  In addition the last failure's text is
 * kept per thread (b200msm_last_error).
 *
 * Unlike ffi.h (big-endian plain values), these entry points take libff's IN-MEMORY layout so
 * that `&*vec.begin()` can be passed with no host-side conversion:
 *   field element  = Fp_model::mont_repr: little-endian 64-bit limbs, Montgomery form
 *                    (fields/bigint.hpp:28-65, fields/fp.hpp:43)
 *   group element  = X | Y | Z coordinates, contiguous, stride sizeof(GroupT)
 *                    (alt_bn128_g1.hpp:44-46, alt_bn128_g2.hpp:45, bls12_377_g1.hpp:44,
 *                     bw6_761_g1.hpp:37-39): Jacobian, bw6_761 homogeneous projective
 *   scalar         = Fr element in Montgomery form (what std::vector<Fr> holds)
 * Returned points are valid representatives in the reference's coordinate system; they compare
 * equal (operator==) to libff's result and have bit-identical affine coordinates.
 *
 * There is no CPU fallback: every entry point fails with an error if no CUDA device is usable.
 */
#ifndef B200MSM_H
#define B200MSM_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* group ids */
#define B200MSM_ALT_BN128_G1 0u /* 96 B elements, 32 B scalars  */
#define B200MSM_ALT_BN128_G2 1u /* 192 B elements, 32 B scalars */
#define B200MSM_BLS12_377_G1 2u /* 144 B elements, 32 B scalars */
#define B200MSM_BW6_761_G1 3u   /* 288 B elements, 48 B scalars */
/* the remaining groups of the same curve families */
#define B200MSM_BLS12_377_G2 4u /* 288 B elements (Fq2 coordinates), 32 B scalars */
#define B200MSM_BW6_761_G2 5u   /* 288 B elements, 48 B scalars */
#define B200MSM_BLS12_381_G1 6u /* 144 B elements, 32 B scalars */
#define B200MSM_BLS12_381_G2 7u /* 288 B elements (Fq2 coordinates), 32 B scalars */
#define B200MSM_NUM_GROUPS 8u

/* multi_exp_base_form (multiexp.hpp:49-58) */
#define B200MSM_FORM_NORMAL 0u  /* arbitrary Z */
#define B200MSM_FORM_SPECIAL 1u /* Z = 1, or the zero element */

/* Binds the calling process to CUDA device `device` (>= 0) and creates its stream and scratch
 * pool; idempotent for the same device.  Mirrors <curve>_init() (ffi/ffi.cpp:124-128,194-198):
 * call once before anything else, not thread-safe. */
bool b200msm_init(int device);
/* The same for several GPUs of one box (b200msm_init(num_gpus)): `count` device
 * ordinals, one engine context each.  From then on multi_exp / multi_exp_resident[_bases] /
 * multi_exp_precomputed cut their index range into one contiguous range per device exactly as
 * the reference cuts it into `chunks` for its OpenMP threads -- one = n / count, the last range
 * takes the remainder (multiexp.tcc:655-678; n < count or count == 1: a single range) -- run the
 * ranges concurrently, one host thread per device, and add the partial points on the first
 * device (multiexp.tcc:681-685).  batch_exp and batch_to_special shard by output range with no
 * exchange.  The wire-format, streamed and debug entry points use the first device only.
 * A device may be listed more than once (two independent contexts on one GPU; a test aid).
 * init_all binds every visible device and returns their number (0 on failure).
 * Option "num_gpus" (b200msm_set_option) limits later calls to the first k bound devices. */
bool b200msm_init_devices(const int *devices, size_t count);
int b200msm_init_all(void);
size_t b200msm_device_count(void);
void b200msm_shutdown(void);
const char *b200msm_last_error(void);

/* sizeof(GroupT) / sizeof(Fr) for a group id; 0 for an unknown id */
size_t b200msm_group_bytes(uint32_t group);
size_t b200msm_scalar_bytes(uint32_t group);

/* libff::multi_exp<G, Fr, multi_exp_method_BDLO12[_signed], form>(bases, scalars, chunks)
 * (multiexp.tcc:643-688).  bases_bytes == n * sizeof(GroupT), scalars_bytes == n * sizeof(Fr),
 * out_bytes == sizeof(GroupT).  n == 0 yields the zero element (multiexp.tcc:291-309).
 * Host pointers; the copies to and from the device happen inside the call. */
bool b200msm_multi_exp(uint32_t group, const void *bases, size_t bases_bytes, const void *scalars,
                       size_t scalars_bytes, size_t n, uint32_t base_form, void *out,
                       size_t out_bytes);

/* The same computation on inputs kept resident in HBM (a prover's fixed bases; scalars of a
 * batch): upload once, run any sub-range [first, first + n) many times, release. */
bool b200msm_bases_upload(uint32_t group, const void *bases, size_t bases_bytes, size_t n,
                          uint32_t base_form, uint64_t *handle);
bool b200msm_scalars_upload(uint32_t group, const void *scalars, size_t scalars_bytes, size_t n,
                            uint64_t *handle);
bool b200msm_multi_exp_resident(uint32_t group, uint64_t bases_handle, uint64_t scalars_handle,
                                size_t first, size_t n, void *out, size_t out_bytes);
/* The prover's shape of the call: the bases (a proving key) stay resident, the scalars (a
 * witness) come from the host with every call and are uploaded in ranges under the computation.
 * n must equal the number of resident bases. */
bool b200msm_multi_exp_resident_bases(uint32_t group, uint64_t bases_handle, const void *scalars,
                                      size_t scalars_bytes, size_t n, void *out, size_t out_bytes);
bool b200msm_release(uint64_t handle);

/* The same MSM on the byte format of the reference's C ABI (ffi/ffi_serialization.hpp:12-16,
 * ffi/ffi_serialization.tcc:98-187; sizes ffi/ffi.h:13-17,55-59) for callers that are not C++:
 * every base-field value big-endian, left-padded to the in-memory width, plain (not Montgomery);
 * Fq2 as c1 | c0; a point is affine X | Y (b200msm_wire_point_bytes(group) bytes: 64 / 128 / 96 /
 * 192) with (0, 1) for infinity; a scalar is sizeof(Fr) bytes.  Decoding, validation and encoding
 * run on the device.  Like the reference's readers the call fails when a value is >= its modulus
 * or a point fails is_well_formed() (y^2 = x^3 + b) or is_in_safe_subgroup() (r P = 0) -- both
 * checks of group_element_read (ffi_serialization.tcc:165) run by default (flags = 0).  A caller
 * that has validated its points before may opt out of either: */
#define B200MSM_WIRE_SKIP_CURVE_CHECK 1u    /* trust y^2 = x^3 + b */
#define B200MSM_WIRE_SKIP_SUBGROUP_CHECK 2u /* trust r P = 0 (saves n scalar multiplications) */
size_t b200msm_wire_point_bytes(uint32_t group);
bool b200msm_multi_exp_wire(uint32_t group, const void *points, size_t points_bytes,
                            const void *scalars, size_t scalars_bytes, size_t n, uint32_t flags,
                            void *out, size_t out_bytes);

/* Sum of `count` group elements in reference layout: the reduction of per-chunk partials that
 * ends libff::multi_exp (multiexp.tcc:681-685); used to combine the partial results of GPU
 * shards.  points_bytes == count * sizeof(GroupT). */
bool b200msm_sum(uint32_t group, const void *points, size_t points_bytes, size_t count, void *out,
                 size_t out_bytes);

/* libff::batch_to_special<G>(vec) (multiexp.tcc:949-974), in place; bytes == n * sizeof(GroupT).
 * Output is canonical (Z = 1; zero -> (0, 1, 0)) and equals the reference's limb for limb. */
bool b200msm_batch_to_special(uint32_t group, void *inout, size_t bytes, size_t n);

/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  batch_to_special on DEVICE memory that allocates nothing: `d_in` and `d_out` are
 * device pointers to n group elements in reference layout on the CURRENT device (they may be the
 * same buffer), bytes == n * sizeof(GroupT); the kernels are enqueued on `cuda_stream` (a
 * cudaStream_t, NULL = the default stream) and the call returns without synchronising.  The result
 * equals b200msm_batch_to_special's limb for limb.  The inversion tree's scratch is a file-scope
 * `static __device__` array of 6.4 MiB that every GPU loading the library holds until exit, so
 * n <= B200MSM_NOALLOC_MAX_POINTS per call (larger n is refused with an error; split the batch).
 * The call performs no cudaMalloc / cudaFree and needs no b200msm_init.  All calls on a device
 * share the array (it keeps the previous call's contents; nothing in it needs clearing): calls
 * that overlap on the GPU -- different streams or host threads -- race and must be serialised. */
#define B200MSM_NOALLOC_MAX_POINTS 32768u
bool b200msm_batch_to_special_noalloc(uint32_t group, const void *d_in, void *d_out, size_t bytes,
                                      size_t n, void *cuda_stream);

/* libff::batch_exp<G, Fr>(scalar_size, window, table, v) (multiexp.tcc:875-915).  window_table<G>
 * is vector<vector<G>> (multiexp.hpp:101): pass one pointer per row, each row `row_len`
 * (= 1 << window) elements (multiexp.tcc:809-846).  num_rows must equal
 * ceil(scalar_bits / window).  out_bytes == n * sizeof(GroupT). */
bool b200msm_batch_exp(uint32_t group, size_t scalar_bits, size_t window,
                       const void *const *table_rows, size_t num_rows, size_t row_len,
                       const void *scalars, size_t scalars_bytes, size_t n, void *out,
                       size_t out_bytes);
/* libff::batch_exp_with_coeff<G, Fr>(scalar_size, window, table, coeff, v) (multiexp.tcc:917-947):
 * out[i] = (coeff * v[i]) g; `coeff` is one Fr element (Montgomery form), or NULL for batch_exp. */
bool b200msm_batch_exp_with_coeff(uint32_t group, size_t scalar_bits, size_t window,
                                  const void *const *table_rows, size_t num_rows, size_t row_len,
                                  const void *coeff, size_t coeff_bytes, const void *scalars,
                                  size_t scalars_bytes, size_t n, void *out, size_t out_bytes);
/* Fixed-base window tables kept resident (normalised once, replicated on every bound device):
 * table_upload takes a window_table<G> the host already has (get_window_table, multiexp.tcc:
 * 809-846); table_build is get_window_table itself on the device from the base g (one GroupT
 * element): entry [r][j] = j 2^(r window) g, the last row shorter (:815) with a zero tail;
 * table_download returns the table (special form, Z = 1) into caller-owned rows so that the
 * reference's own consumers of a window_table keep working; batch_exp_resident is batch_exp /
 * batch_exp_with_coeff (coeff != NULL) from the resident table.  Release with b200msm_release. */
bool b200msm_table_upload(uint32_t group, size_t scalar_bits, size_t window,
                          const void *const *table_rows, size_t num_rows, size_t row_len,
                          uint64_t *handle);
bool b200msm_table_build(uint32_t group, size_t scalar_bits, size_t window, const void *g,
                         size_t g_bytes, uint64_t *handle);
bool b200msm_table_shape(uint64_t handle, size_t *scalar_bits, size_t *window, size_t *num_rows,
                         size_t *row_len);
bool b200msm_table_download(uint64_t handle, void *const *table_rows, size_t num_rows,
                            size_t row_len);
bool b200msm_batch_exp_resident(uint32_t group, uint64_t table_handle, const void *scalars,
                                size_t scalars_bytes, size_t n, const void *coeff,
                                size_t coeff_bytes, void *out, size_t out_bytes);

/* Element-wise leaves of the path, for parity tests (n elements each, same layout as above):
 * op 0..5 coordinate field add, sub, mul, sqr, neg, inverse (Fp_model, fields/fp.tcc:50-727;
 * Fp2_model for alt_bn128 G2, fields/fp2.tcc:79-168); 6 scalar field Montgomery -> plain
 * (as_bigint, fp.tcc:270-281); 7 scalar field mul; 10 group add, 11 dbl, 12 mixed add (b special),
 * 13 negation (e.g. alt_bn128_g1.cpp:139-326).  `b` may be NULL for unary ops.  bytes is the
 * size of each of a, b, out. */
bool b200msm_debug_op(uint32_t group, uint32_t op, const void *a, const void *b, void *out,
                      size_t bytes, size_t n);

/* ---- Precomputed multiples: multi_exp_stream_with_precompute (multiexp_stream.hpp:27-41,
 * multiexp_stream.tcc:124-162, 195-223).  For every base P_i the table holds the W multiples
 * P_i, 2^c P_i, ..., 2^((W-1) c) P_i with W = (scalar bits + c - 1) / c (b200msm_precompute_digits),
 * so all digits of all scalars feed ONE set of 2^(c-1) buckets.  The table stays resident in HBM
 * (a fixed proving key) and is released with b200msm_release.
 * precompute:          builds the table on the device from resident bases; c = 0 lets the
 *                      engine choose (b200msm_precompute_window_bits).
 * precomputed_upload:  takes the table in the order the reference's stream has it -- for each
 *                      base its W multiples -- as n * W in-memory GroupT elements.
 * multi_exp_precomputed: scalars either resident (scalars_handle != 0) or on the host
 *                      (scalars_handle == 0: n * sizeof(Fr) bytes, Montgomery form).
 * Where W c equals the scalar width the reference drops the carry out of the top signed digit
 * (field_utils.tcc:205-239); here such scalars are folded to r - s, so the result is always
 * the group element sum s_i P_i.  A streamed MSM begun with B200MSM_STREAM_REFERENCE_CARRY or'ed
 * into precompute_c reproduces the reference instead: its signed recoding (a digit of exactly
 * 2^(c-1) carries) with the top carry dropped, bit for bit what multi_exp_stream_with_precompute
 * returns. */
#define B200MSM_STREAM_REFERENCE_CARRY 0x100u
uint32_t b200msm_precompute_window_bits(uint32_t group, size_t n);
size_t b200msm_precompute_digits(uint32_t group, uint32_t c);
bool b200msm_precompute(uint32_t group, uint64_t bases_handle, uint32_t c, uint64_t *table_handle);
bool b200msm_precomputed_upload(uint32_t group, const void *elements, size_t bytes, size_t n,
                                uint32_t c, uint32_t base_form, uint64_t *table_handle);
bool b200msm_multi_exp_precomputed(uint32_t group, uint64_t table_handle, uint64_t scalars_handle,
                                   const void *scalars, size_t scalars_bytes, void *out,
                                   size_t out_bytes);

/* ---- Streamed bases: multi_exp_stream / multi_exp_stream_with_precompute
 * (multiexp_stream.tcc:164-223).  The scalars are given up front, the bases arrive in chunks
 * (what the reference reads from its std::istream through a FIFO); every chunk is uploaded,
 * decoded and reduced on the device while the caller reads the next one.  `encoding`:
 * in-memory GroupT elements (special form Z = 1, or any Z), or the bytes of the reference's
 * uncompressed binary stream encoding (curve_serialization.tcc:79-101: big-endian X then Y,
 * plain or Montgomery values, (0, 1) = zero), decoded on the device.  precompute_c != 0: every
 * base comes as its W multiples (stream layout of multiexp_stream.hpp:30-36).
 * push returns once the chunk has left the caller's buffer.  finish releases the stream. */
#define B200MSM_STREAM_MEMORY_SPECIAL 0u
#define B200MSM_STREAM_MEMORY_NORMAL 1u
#define B200MSM_STREAM_BINARY_PLAIN 2u
#define B200MSM_STREAM_BINARY_MONTGOMERY 3u
bool b200msm_stream_begin(uint32_t group, const void *scalars, size_t scalars_bytes, size_t n,
                          uint32_t encoding, uint32_t precompute_c, uint64_t *stream);
bool b200msm_stream_push(uint64_t stream, const void *elements, size_t bytes, size_t count);
bool b200msm_stream_finish(uint64_t stream, void *out, size_t out_bytes);

/* Tuning and measurement (not part of the reference surface).
 * option: "window_bits" (0 = automatic), "segment" (entries per thread of the accumulation),
 * "h2d_parts" (point ranges a host-buffer multi_exp is cut into so that uploads overlap the
 * computation; 0 = automatic, 1 = no overlap), "h2d_growth" (size ratio of consecutive ranges in
 * percent; 0 = automatic), "h2d_shared_buckets" (1, default: the ranges are sorted with the digit
 * size of the whole call and accumulate into one bucket set that is reduced once; 0: one
 * independent MSM per range), "trace" (1: per-range phase times of a pipelined call on stderr; it
 * serialises the pipeline), "window_groups" (groups the windows are walked in, top window first, so that
 * the reduction of one group runs under the accumulation of the next; 0 = automatic, 1 = one
 * group), "carry_segment" (carries per thread in the later accumulation rounds; 0 = 4),
 * "reduce_chunk" (log2 of the buckets per thread in the bucket reduction; 0 = automatic),
 * "reduce_cluster" (CTAs of the thread-block cluster one window's totals are spread over: 0 =
 * automatic (up to 8), 1 = one CTA per window),
 * "num_gpus" (bound devices a call spreads over; 0 = all), "h2d_pack" (special-form bases leave
 * their constant Z behind on upload, packed to X|Y by host threads: 0 never (default: the plain
 * copy measured faster on a 16-core host), 1 pageable sources, 2 always), "stage_lanes" (host
 * threads / pinned lanes that stage pageable sources, 1..16, default 8), "h2d_taper" (the last
 * upload ranges shrink again by this many growth steps per range; 0 = off, default),
 * "sort_overlap" (groups of windows whose counting sort runs under the previous group's
 * accumulation; 0 = automatic = one sort for all windows, k > 1 = k groups), "coop_accum"
 * (accumulation of the 24-limb field with the accumulator in shared-memory slots: 0 automatic =
 * on, 1 off, 2 on). */
bool b200msm_set_option(const char *name, int64_t value);
/* CUDA-event timings (ms) of the phases of the most recent multi_exp on this thread's engine,
 * measured on the engine's own stream.  Writes up to `cap` values, returns how many phases exist.
 * Order: 0 upload/ingest, 1 digits+histogram, 2 scan, 3 scatter, 4 bucket accumulation
 * (first round), 5 carry rounds, 6 bucket reduction, 7 window reduction + final, 8 total. */
size_t b200msm_last_timings(float *ms, size_t cap);
/* the same for one bound device (last_timings is the element-wise maximum over the devices) */
size_t b200msm_device_timings(size_t device_index, float *ms, size_t cap);
const char *b200msm_timing_name(size_t phase);
/* Device-side stopwatch for benchmarks: CUDA events recorded on the engine's own stream (the
 * stream every kernel and copy of this library is issued on).  timer_stop waits for the stop
 * event and returns the elapsed milliseconds between the two records, or a negative value. */
bool b200msm_timer_start(void);
float b200msm_timer_stop(void);
/* number of kernel launches issued by this library since init */
uint64_t b200msm_launch_count(void);
/* window bits the engine picks for n points of a group */
uint32_t b200msm_window_bits(uint32_t group, size_t n);

#ifdef __cplusplus
}
#endif
#endif /* B200MSM_H */
