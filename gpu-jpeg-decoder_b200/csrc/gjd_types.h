/*
 * gjd_types.h — internal structures shared by the host code and the CUDA kernels.
 *
 * HBM layout of one resident batch (all offsets 256-byte aligned, one arena per kind):
 *
 *   scan   u8    unstuffed entropy-coded bytes of every image, each image starting on a
 *                128-byte boundary and followed by >= 16 zero bytes; restart intervals of an image
 *                follow each other, each starting on a 4-byte boundary (RSTn markers removed).
 *   sinfo  u32x2 one state per subsequence (2^sub_log2 bits of one image's scan), padded per
 *                image to a multiple of GJD_SEQ_SUBS ("sequence" = the subsequences one thread
 *                block synchronises): {p, (n << 8) | pos}.
 *   coef   i16   quantised coefficients, zigzag order, one 128-byte line per 8x8 block:
 *                [(tile*96 + lane*3 + comp) * 64 + k]; tile = 32 consecutive MCUs of one image (one
 *                warp of the transform kernel), so an image's blocks sit in the reference order
 *                mcu*192 + comp*64 + k (cuda-decoder/src/parser.cu:538).  Only the 32-byte sectors up
 *                to the block's last non-zero coefficient are written (and later read); slot 0 is
 *                unused (the DC difference travels in meta).
 *   meta   u32   per block, same index as coef lines: [15:0] DC difference (i16), [21:16] kmax = the
 *                last non-zero zigzag index (0 = DC only: no coef sector written).  Cleared per run.
 *   tilebase i32 [tile][4]: DC predictor of each component at the start of the tile.
 *   rgb    u8    planar R,G,B output per image in the caller's pitch/stride geometry.
 */
#ifndef GJD_TYPES_H
#define GJD_TYPES_H

#include <stdint.h>

#define GJD_SUB_LOG2_MIN   10u      /* subsequence (one thread) = 2^k bits, k chosen per batch:        */
#ifndef GJD_SUB_LOG2_IMAGE_MIN
#define GJD_SUB_LOG2_IMAGE_MIN 9u   /* the shortest a small image takes inside a batch of longer ones (GjdSegment.sub_log2) */
#endif
#define GJD_SUB_LOG2_MAX   13u      /* long enough to amortise the look-back, short enough to fill the GPU */
#define GJD_LOOKBACK_BITS 512u     /* cold-start distance before a subsequence (k_huff_sync)          */
#ifndef GJD_LOOKBACK_SUB_BITS
#define GJD_LOOKBACK_SUB_BITS 512u /* the same for chroma-subsampled scans (MCUs of 4 or 6 blocks)    */
#endif
#define GJD_SEQ_SUBS      256u     /* subsequences per sequence (one thread block) */
#ifndef GJD_LUT_BITS
#define GJD_LUT_BITS      10       /* primary Huffman LUT index width (9 or 10: the secondary LUT covers 6 more bits) */
#endif
#define GJD_LUT_SIZE      (1 << GJD_LUT_BITS)
#define GJD_LUT2_PREFIXES  8       /* secondary LUT: 64-entry sub-tables for the first 8 unresolved prefixes */
#define GJD_LUT2_SIZE     (GJD_LUT2_PREFIXES * 64)
#define GJD_TILE_MCUS     32u      /* MCUs per coefficient tile (one warp)         */
#define GJD_TB_TILES      32u      /* most tiles per transform thread block (one image; GjdDevBatch.tb_tiles) */

/* LUT entry (u16):  [4:0]  total bits = code length + value size (1..27); 0 = code longer than
 *                          GJD_LUT_BITS (or unassigned): then [15:5] = 1 + rank of this prefix among the
 *                          unresolved ones when its 64-entry sub-table exists in lut2 (indexed by the next
 *                          6 bits; same entry format), else 0 = resolve with the canonical tables
 *                   [8:5]  value size (0..15)
 *                   [15:9] slot advance = run + 1 (1..16; DC symbols: 1); 64 = EOB (to the block end) */
#define GJD_ADV_EOB 64u
#define GJD_RING_TABLES 6
#define GJD_RING_STRIDE (GJD_LUT_SIZE * 2u)                  /* bytes per table of the ring        */
#define GJD_RING_BYTES  (GJD_RING_TABLES * GJD_RING_STRIDE)
#define GJD_LUT2_STRIDE  (GJD_LUT2_SIZE * 2u)                 /* bytes per secondary table (= RING_STRIDE / 2, or RING_STRIDE with a 9-bit index) */
/* Offset (from the start of the ring) of the secondary-LUT entry behind link e = (rank + 1) << 5 of the table at
 * ring offset tb, for the window w: sub-table rank of secondary table tb / RING_STRIDE, entry = the 6 bits after
 * the primary index. */
#define GJD_LUT2_OFF(tb, e, w) ((GJD_RING_BYTES - 128u) + (GJD_LUT2_STRIDE == GJD_RING_STRIDE ? (tb) : (tb) >> 1) + \
                                (e) * 4u + (((w) >> (25 - GJD_LUT_BITS)) & 126u))

/* One Huffman table set as the decoders use it, device-resident, built by k_build_luts from a
 * GjdDhtSet: the primary LUTs laid out as a RING in scan order — DC(comp 0), AC(comp 0), DC(comp 1),
 * AC(comp 1), DC(comp 2), AC(comp 2) — so that "next table" is "+1 table, wrap after 6" (a DC symbol
 * and a block end each advance the ring by one).  The canonical tables of the 4 DHT slots
 * (0,1 = DC Th 0,1; 2,3 = AC Th 0,1) serve the rare long codes; slot_of maps ring position -> slot. */
typedef struct GjdLutSet {
    uint16_t lut[GJD_RING_TABLES][GJD_LUT_SIZE];
    uint16_t lut2[GJD_RING_TABLES][GJD_LUT2_SIZE];   /* directly behind lut: one offset space for both */
    uint32_t lim[4][17];             /* lim[L] = first unused code of length L, left-aligned to 16 bits */
    int32_t  off[4][17];             /* valptr[L] - mincode[L]                                   */
    uint8_t  vals[4][256];           /* HUFFVAL                                                  */
    uint8_t  isdc[4];                /* 1 for DC tables (value size = symbol, advance = 1)       */
    uint8_t  slot_of[GJD_RING_TABLES];
    uint8_t  pad[6];
} GjdLutSet;

/* The primary LUT ring as the COUNT pass (k_huff_sync) uses it: same index and the same [4:0] total bits and
 * [15:9] slot advance, but no value size — instead bit 5 (GJD_PAIR_FLAG) marks an entry that covers TWO
 * consecutive symbols of an AC table whose bits all lie inside the index (total and advance are the sums;
 * an EOB as the second symbol contributes GJD_ADV_EOB; [8:6] = advance of the first symbol - 1, at most 7).
 * Pairs cut the symbol loop's trip count by a third
 * on photographic data.  Entries with [4:0] = 0 are the same links into lut2 as in GjdLutSet. */
#define GJD_PAIR_FLAG 0x20u
typedef struct GjdLutCount {
    uint16_t lut[GJD_RING_TABLES][GJD_LUT_SIZE];
} GjdLutCount;

/* Raw DHT content + the scan's table selectors: what the LUT builder consumes (host -> device). */
typedef struct GjdDhtSet {
    uint8_t bits[4][16];
    uint8_t vals[4][256];
    uint8_t isdc[4];
    uint8_t tabsel[8];               /* [c] = slot of component c's DC table, [4+c] = its AC table */
    uint8_t pad[4];
} GjdDhtSet;

/* The quantisation table of each component, zigzag order, widened to i32 for broadcast loads. */
typedef struct GjdQuantSet {
    int32_t q[3][64];
    uint32_t packed[3][32];          /* q[c][2j] | q[c][2j+1] << 24: the words the transform's dp2a dequantisation takes */
} GjdQuantSet;

/* The entropy-coded data of one image's scan.  The entropy kernels (k_huff_sync, k_huff_fix, k_seq_scan, k_huff_write) work
 * on segments; sequences belong to a segment.  Restart intervals (T.81 E.1.4: the decoder state — block in the MCU, zigzag
 * index, DC predictors — is reset at every RSTn) lie back to back inside the segment, markers removed; the table of their
 * start positions makes every restart a hard synchronisation point for the decoders (GjdRstTable, gjd_devfn.cuh).  (Round 1
 * made every interval a segment of its own: one 256-thread block and one LUT copy per interval, most of it idle.) */
typedef struct GjdSegment {
    uint64_t scan_off;      /* byte offset of the unstuffed bytes in the scan arena (4-byte aligned) */
    uint32_t scan_bits;     /* unstuffed length in bits (a trailing partial byte of 1-padding included) */
    uint32_t sub_count;     /* subsequences                                                    */
    uint32_t seq_first;     /* global index of the segment's first sequence                    */
    uint32_t seq_count;
    uint32_t img;           /* owning image                                                    */
    uint32_t blk_first;     /* first 8x8 block of the segment, counted from the image's first block (blocks per MCU * first MCU) */
    uint32_t blk_count;     /* blocks per MCU * MCUs in the segment                            */
    uint16_t lut_id;        /* index into the GjdLutSet array                                  */
    uint8_t  ny;            /* luma blocks per MCU: 1 (4:4:4, single-component), 2 (4:2:2, 4:4:0) or 4 (4:2:0); an MCU is
                               ny luma blocks + Cb + Cr, i.e. ny + 2 blocks (three for the single-component trick)      */
    uint8_t  sub_log2;      /* log2 of this image's subsequence length in bits: the batch's (GjdDevBatch.sub_log2), or shorter for an
                               image too small to fill half a thread block at that length */
    uint32_t rst_first;     /* restart intervals inside the scan: index of the segment's first entry in the rst arena ...     */
    uint32_t rst_count;     /* ... and the number of entries = restart markers = intervals - 1 (0: no restart intervals);
                               entry k - 1 = bit position in the unstuffed scan where interval k starts            */
    uint32_t rst_blocks;    /* blocks per restart interval (restart_interval * blocks per MCU)                       */
    uint32_t pad1;
} GjdSegment;

/* A scan that is unstuffed ON THE DEVICE (k_unstuff_*): the raw entropy-coded bytes as they stand in the file, from the
 * first byte after the SOS header to the end of the file (stuffed FF00 pairs, the EOI marker, whatever follows it).
 * Work is cut into chunks of GJD_UCHUNK bytes counted from the 16-byte boundary at or below raw_off. */
#define GJD_UCHUNK 4096u
typedef struct GjdRaw {
    uint64_t raw_off;       /* byte offset of the first scan byte in the raw arena                    */
    uint32_t raw_len;       /* bytes from there to the end of the file                                 */
    uint32_t chunk_first;   /* global index of its first chunk                                         */
    uint32_t chunk_count;
    uint32_t seg;           /* its segment: scan_bits and sub_count are written by k_unstuff_scan      */
    uint32_t img;
    uint32_t dri;           /* the file has a DRI segment: RSTn markers are expected (and removed) instead of being an error */
} GjdRaw;
/* raw_err word: (position in the scan) << 2 | code; the smallest one before the EOI decides the image's status */
#define GJD_UERR_FORMAT 1u       /* a marker other than RSTn / EOI inside the scan, or no EOI at all   */
#define GJD_UERR_UNSUPPORTED 2u  /* RSTn in a file without DRI                                          */

/* Per-image device descriptor (transform side). */
typedef struct GjdImage {
    uint64_t out_off;       /* byte offset in the rgb arena, or a raw device address           */
    uint64_t plane_stride;  /* bytes between planes                                            */
    uint32_t row_pitch;     /* bytes between rows                                              */
    uint32_t tile_first;    /* global index of the image's first coefficient tile              */
    uint32_t tile_count;
    uint32_t tb_first;      /* global index of the image's first transform block               */
    uint32_t tb_count;      /* ceil(tile_count / tiles per transform block of the batch)       */
    uint32_t mcu_count;
    uint32_t mcus_x;
    uint32_t width, height;
    uint32_t restart_interval; /* MCUs per restart interval (DC predictors reset), 0 = none     */
    uint32_t seg_first;     /* global index of the image's first segment                       */
    uint32_t seg_count;
    uint16_t quant_id;      /* index into the GjdQuantSet array                                */
    uint16_t layout;        /* GJD_LAYOUT_PLANAR / GJD_LAYOUT_INTERLEAVED (include/gjd.h)        */
    uint32_t out_is_ptr;    /* out_off is an absolute device address (caller-owned buffer)     */
    uint32_t gray;          /* single-component image: mcu_count blocks in one DC chain, tiles of 96 blocks,
                               decoded by k_transform_gray (the entropy kernels see three ring positions that
                               all use the one table pair)                                    */
    uint8_t  hs, vs;        /* luma sampling factors (1 or 2 each); chroma is 1x1.  hs * vs > 1: a chroma-subsampled image —
                               MCUs of 8*hs x 8*vs pixels and hs*vs + 2 blocks, decoded by k_dc_sub + k_transform_sub        */
    uint16_t pad1;
} GjdImage;

/* Device-written status bits per image. */
#define GJD_DEV_BADCODE   1u   /* write pass met an unassigned Huffman code                     */
#define GJD_DEV_BADSTATE  2u   /* synchronised state disagrees with the coefficient index       */
#define GJD_DEV_BADCOUNT  4u   /* decoded coefficient count != 192 * MCUs                       */
#define GJD_DEV_NOSYNC    8u   /* a sequence boundary did not converge in one fix-up round      */
#define GJD_DEV_UFORMAT   16u  /* device unstuffing: stray marker inside the scan / no EOI (GJD_E_FORMAT)       */
#define GJD_DEV_UUNSUP    32u  /* device unstuffing: RSTn without DRI (GJD_E_UNSUPPORTED)                        */

#endif
