/*
 * gjd.h — C ABI of libgjd.so, the B200-native baseline-JPEG (4:4:4) decode path.
 *
 * This is the drop-in boundary for the reference's hot path.  The reference
 * (debesheedas/GPU-JPEG-Decoder) has no FFI of its own; its de-facto interfaces
 * are C++ calls that the CLI and the benchmark drivers make directly.  Each entry
 * point below names the reference interface it replaces (paths relative to the
 * reference repository root).
 *
 * Conventions: plain pointers and sizes only; every function returns 0 (GJD_OK)
 * or a negative gjd_status; no exceptions cross the ABI; no process-global
 * state.  A gjd_ctx belongs to one host thread and one CUDA device.  There is
 * no CPU decode path behind this ABI: without a CUDA device gjd_create fails
 * with GJD_E_NODEVICE and nothing else can be called.
 */
#ifndef GJD_H
#define GJD_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum gjd_status {
    GJD_OK            = 0,
    GJD_E_ARG         = -1,  /* NULL / out-of-range argument                                */
    GJD_E_CUDA        = -2,  /* a CUDA runtime call failed (gjd_last_cuda_error has detail) */
    GJD_E_NOMEM       = -3,  /* host or device allocation failed                            */
    GJD_E_FORMAT      = -4,  /* not a JPEG, truncated, or a marker segment is malformed     */
    GJD_E_UNSUPPORTED = -5,  /* valid JPEG outside the contract: progressive, arithmetic,
                                12-bit, subsampled, 2 or 4 components, multi-scan             */
    GJD_E_CORRUPT     = -6,  /* entropy-coded data is inconsistent (bad code / wrong count /
                                frame header announcing more blocks than the scan can hold) */
    GJD_E_NODEVICE    = -7,  /* no usable CUDA device: the product has no CPU fallback      */
    GJD_E_IO          = -8   /* file could not be read / written (writers only)             */
} gjd_status;

typedef struct gjd_ctx   gjd_ctx;    /* streams + pinned staging + device arenas            */
typedef struct gjd_batch gjd_batch;  /* a parsed batch resident in device memory            */

/* Pixel layouts of gjd_item.rgb.
 * PLANAR      three planes R, G, B of height rows, row_pitch bytes apart, plane_stride between planes —
 *             the order of the reference's .array file (cpp-decoder/src/parser.cpp:203-208).
 * INTERLEAVED R,G,B,R,G,B,... rows of 3*width bytes, row_pitch bytes apart (0 = 3*width); plane_stride is
 *             ignored.  This is the pixel order of the reference's device buffer outputChannels
 *             (cuda-decoder/src/parser.cu:571-573, read back at :715-717) narrowed from int16 to u8 and
 *             cropped, and the body of a binary PPM (testing/jpeglib_output_ppm).
 * INTERLEAVED_I16
 *             the same order with int16 elements (6 bytes per pixel; row_pitch in bytes, 0 = 6*width): element type AND
 *             order of outputChannels (cuda-decoder/src/parser.cu:571-573), for callers that read that buffer as the
 *             reference's write() does (:702-724).  A caller that wants the reference's padded geometry passes
 *             row_pitch = 6 * 8*ceil(width/8); only the width x height pixels are written. */
#define GJD_LAYOUT_PLANAR          0
#define GJD_LAYOUT_INTERLEAVED     1
#define GJD_LAYOUT_INTERLEAVED_I16 2

/* One image of a batch.
 * Replaces the per-image HostData/DeviceData descriptor pair the reference fills with
 * ~26 CUDA calls per image (cuda-decoder/src/parser.h:21-39, benchmark_thoughput/benchmark.cu:49-60). */
typedef struct gjd_item {
    const uint8_t* jpg;          /* in : whole .jpg file in host memory (caller-owned, read-only)      */
    size_t         len;          /* in : its length in bytes                                            */
    uint8_t*       rgb;          /* in : destination of the planar R,G,B u8 pixels, cropped to W x H    */
                                 /*      (caller-owned; host memory unless rgb_on_device != 0).         */
                                 /*      NULL in gjd_batch_prepare = library-owned device buffer.       */
    size_t         plane_stride; /* in : bytes between the R, G and B planes; 0 = row_pitch * height    */
    size_t         row_pitch;    /* in : bytes between rows; 0 = tight.  Bytes between rows/planes are   */
                                 /*      unspecified after the call (regions are copied whole).         */
    int32_t        rgb_on_device;/* in : non-zero: rgb is a device pointer on the context's device      */
    int32_t        width;        /* out: image width  (SOF0)                                            */
    int32_t        height;       /* out: image height (SOF0)                                            */
    int32_t        status;       /* out: per-image gjd_status                                           */
    int32_t        layout;       /* in : GJD_LAYOUT_PLANAR (0), GJD_LAYOUT_INTERLEAVED or _INTERLEAVED_I16 */
    int32_t        reserved;     /* in : 0                                                              */
} gjd_item;

/* Context.  Replaces allocate()/clean() (cuda-decoder/src/parser.cu:324-358, 684-700): all device
 * and pinned memory is arena-allocated once per context and reused across batches.
 * scratch_bytes_hint: device bytes the context may keep for one in-flight sub-batch
 * (0 = default, 6 GiB); batches larger than that are processed as pipelined sub-batches, up to
 * three in flight inside gjd_decode_batch (one being unstuffed by the host threads, one uploading,
 * one in the kernels / reading back), so the context may hold three times the hint. */
int  gjd_create(gjd_ctx** out, int cuda_device, size_t scratch_bytes_hint);
void gjd_destroy(gjd_ctx* ctx);

/* Host-only SOF0 peek so callers can size outputs.
 * Replaces the width/height out-params of extract() (cuda-decoder/src/parser.cu:360-471, SOF0 at :407-416;
 * cpp-decoder/src/parser.cpp:52-60). */
int  gjd_peek(const uint8_t* jpg, size_t len, int32_t* width, int32_t* height);

/* Decode n images end to end: marker parse + byte-unstuffing on the host, H2D, entropy decode,
 * dezigzag+dequant+IDCT+colour on the GPU, D2H into items[i].rgb (skipped for rgb_on_device).
 * Single-component (grayscale) files are decoded too (R = G = B from the reference's colour formula with
 * Cb = Cr = 0); the reference's parser wants exactly three components.
 * Restart markers (DRI/RSTn) are decoded (every restart interval is an independent segment for the
 * entropy kernels; DC predictors restart) — a superset of the reference's parser, which keeps FF Dn as
 * scan data (cpp-decoder/src/parser.cpp:84-96).
 * Synchronous on return.  Returns GJD_OK if the call itself worked; per-image failures are reported
 * in items[i].status and do not stop the batch.
 * Replaces extract()+allocate()+batchDecodeKernel<<<n,256>>>+D2H+clean()
 * (cuda-decoder/benchmark_thoughput/benchmark.cu:49-93, cuda-decoder/src/parser.cu:663-682) and, for n = 1,
 * decodeKernel<<<1,T>>> (cuda-decoder/main.cu:32-39, src/parser.cu:577-611);
 * on the CPU side JPEGParser::extract()+decode() (cpp-decoder/src/parser.cpp:24-103, 144-195). */
int  gjd_decode_batch(gjd_ctx* ctx, gjd_item* items, int32_t n);

/* Resident-batch interface: the split the reference's throughput benchmark makes between the untimed
 * per-image extract()/allocate() loop and the timed batchDecodeKernel launch
 * (cuda-decoder/benchmark_thoughput/benchmark.cu:49-70).
 *   prepare : parse + unstuff + H2D of every scan; allocates coefficient scratch and, for items with
 *             rgb == NULL, a device output buffer.  items[i].width/height/status are filled.
 *   run     : enqueue all GPU stages for the whole batch on the context stream (asynchronous).
 *   sync    : wait; returns GJD_E_CORRUPT if any image failed on the device (per-image detail via
 *             gjd_batch_status).
 *   fetch   : copy image i's pixels to host memory, tight, in the item's layout (planar: W*H bytes per
 *             plane; interleaved: 3*W bytes per row; interleaved int16: 6*W bytes per row). */
int  gjd_batch_prepare(gjd_ctx* ctx, gjd_item* items, int32_t n, gjd_batch** out);
int  gjd_batch_run(gjd_ctx* ctx, gjd_batch* b);
int  gjd_batch_sync(gjd_ctx* ctx, gjd_batch* b);
int  gjd_batch_status(gjd_ctx* ctx, gjd_batch* b, int32_t i);
int  gjd_batch_fetch(gjd_ctx* ctx, gjd_batch* b, int32_t i, uint8_t* rgb_host);
void gjd_batch_free(gjd_ctx* ctx, gjd_batch* b);
/* Sizes of the last prepared batch: [0] pixels (sum W*H), [1] padded pixels (sum pW*pH),
 * [2] unstuffed scan bytes, [3] subsequences, [4] kernel launches per gjd_batch_run. */
int  gjd_batch_info(gjd_ctx* ctx, gjd_batch* b, uint64_t info[5]);

/* Quantised coefficients of one image, int16, zigzag order, absolute DC, layout
 * coeffs[mcu*192 + comp*64 + k] over ceil(W/8)*ceil(H/8) MCUs — the layout of the reference's device
 * buffer after parallelHuffManDecode (cuda-decoder/src/parser.cu:132-208, index at :538).
 * Test hook for the "bit-exact Huffman coefficients" gate.  Grayscale files: coeffs[block*64 + k]. */
int  gjd_decode_coeffs(gjd_ctx* ctx, const uint8_t* jpg, size_t len, int16_t* coeffs_host);

/* Device times (ms, CUDA events on the context stream) of the last gjd_batch_run / the last sub-batch
 * of gjd_decode_batch: [0] entropy sync (cold decode + self-synchronisation + boundary fix-up),
 * [1] count scan + coefficient write pass + DC bases, [2] transform (dequant+IDCT+colour+store),
 * [3] whole GPU pipeline incl. scratch clears.
 * Replaces the cudaEvent pair around the kernel in cuda-decoder/benchmark_thoughput/benchmark.cu:61-70. */
int  gjd_stage_times(gjd_ctx* ctx, float ms[4]);

/* Stand-alone colour conversion of n (Y,Cb,Cr) int16 triples on the GPU with the same device function the
 * transform kernel uses (host pointers in, host pointers out).  Test hook for the exhaustive
 * colour-equivalence check against cpp-decoder/utils/color.cpp:3-18. */
int  gjd_test_color(gjd_ctx* ctx, const int16_t* y, const int16_t* cb, const int16_t* cr, size_t n,
                    uint8_t* r, uint8_t* g, uint8_t* b);
/* Stand-alone dezigzag+dequant+IDCT of nblocks 8x8 blocks (quantised zigzag int16 in, 64-entry u8 zigzag
 * quant table, int16 natural-order samples in [-256,255] out).  Test hook for
 * cpp-decoder/src/idct.cpp:24-133. */
int  gjd_test_idct(gjd_ctx* ctx, const int16_t* coef_zz, size_t nblocks, const uint8_t quant_zz[64],
                   int16_t* samples);

/* gjd_test_idct for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid allocation).  It has NO counterpart in the reference.
 * Same arithmetic and same results as gjd_test_idct, but the device-side scratch (staged
 * coefficients, expanded quantisation table, staged samples, a finished-blocks word) lives in one file-scope
 * `static __device__` array of 512 + 256 * GJD_STATIC_IDCT_MAX_BLOCKS bytes (2 MiB + 512 B), which every GPU that
 * loads the library holds until the process exits.  The call uses no gjd_ctx, runs on the legacy default stream of
 * `cuda_device`, and calls no cudaMalloc / cudaFree / cudaMallocAsync, creates no stream or event.
 *   coef_zz, samples : caller-owned, host or device memory (copied with cudaMemcpyDefault), 64 int16 per block.
 *   nblocks          : at most GJD_STATIC_IDCT_MAX_BLOCKS; more is rejected with GJD_E_ARG and a message in errbuf.
 *   errbuf           : optional (may be NULL): receives a one-line description of any failure.
 * The array is shared by every call and keeps the last call's contents (the word that must start at zero is cleared
 * by each call).  Calls that overlapped on the GPU would race on it: the entry point therefore takes a process-wide
 * lock and is synchronous on return; this is the one piece of process-global state behind this ABI. */
#define GJD_STATIC_IDCT_MAX_BLOCKS 8192
int  gjd_idct_blocks_static(int cuda_device, const int16_t* coef_zz, size_t nblocks, const uint8_t quant_zz[64],
                            int16_t* samples, char* errbuf, size_t errbuf_len);

/* Multi-GPU sharding (host-only, no CUDA): assign n independent images to `parts` devices by greedy
 * longest-processing-time on the padded pixel count 8*ceil(W/8) * 8*ceil(H/8) (the unit both the
 * coefficient traffic and, roughly, the scan bits follow).  Deterministic: ties go to the lower image
 * index, then to the lower part.  part_of[i] receives the device of image i (or -1 when weight[i] is 0,
 * i.e. an image gjd_peek rejected); load[p] (may be NULL) the weight assigned to part p.  The weights are the
 * caller's: gjd_partition_weights (padded pixels) or gjd_partition_costs (padded pixels + a byte term) below.
 * The reference has no multi-GPU path; this replaces "one batchDecodeKernel over everything"
 * (cuda-decoder/benchmark_thoughput/benchmark.cu:56-70) by one such launch per device.  There is no
 * exchange step between shards, hence no collective. */
int  gjd_partition(const uint64_t* weight, int32_t n, int32_t parts, int32_t* part_of, uint64_t* load);
/* Convenience over gjd_peek: weight[i] = padded pixel count of items[i] (0 if the header is unusable). */
int  gjd_partition_weights(const gjd_item* items, int32_t n, uint64_t* weight);
/* The same with a term for the compressed size: cost[i] = padded pixels + byte_weight * items[i].len (0 if the
 * header is unusable).  The entropy stages follow the scan bytes and the IDCT/colour stage the padded pixels, so a
 * set that mixes qualities balances pixels alone and still leaves one device with more bytes (24000 mixed files
 * over 4 parts: bytes max/mean 1.165 with byte_weight 0, 1.000 with any byte_weight from 3 to 8).  The measured
 * ratio of the two rates on B200 is GJD_PARTITION_BYTE_WEIGHT padded pixels per scan byte.
 * No counterpart in the reference (no multi-GPU path). */
#define GJD_PARTITION_BYTE_WEIGHT 5u
int  gjd_partition_costs(const gjd_item* items, int32_t n, uint32_t byte_weight, uint64_t* cost);

/* Pinned host memory helpers (callers that want full-speed H2D/D2H). */
void* gjd_host_alloc(size_t bytes);
void  gjd_host_free(void* p);

/* Output writers on the caller's side of the ABI (host code, no GPU).
 * gjd_write_array replaces JPEGParser::write (cpp-decoder/src/parser.cpp:197-210) and write()
 * (cuda-decoder/src/parser.cu:702-744): "H W\n", then the R, G and B planes as decimal integers each
 * followed by one space, planes separated by '\n', no trailing newline.
 * gjd_write_ppm writes binary P6 (the format of testing/jpeglib_output_ppm). */
int  gjd_write_array(const char* path, int32_t width, int32_t height, const uint8_t* rgb_planar);
int  gjd_write_ppm(const char* path, int32_t width, int32_t height, const uint8_t* rgb_planar);

const char* gjd_strerror(int status);
const char* gjd_last_cuda_error(gjd_ctx* ctx);
/* "gjd <version> sm_100a subseq=<bits> lut=<bits>" */
const char* gjd_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GJD_H */
