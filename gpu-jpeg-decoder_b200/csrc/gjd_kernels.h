/*
 * gjd_kernels.h — launchers of the sm_100a kernels (gjd_kernels.cu).  Internal.
 */
#ifndef GJD_KERNELS_H
#define GJD_KERNELS_H

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "gjd_types.h"

/* Device view of one resident batch (see the HBM layout in gjd_types.h). */
struct GjdDevBatch {
    const GjdImage* imgs;
    uint32_t n_img;
    GjdSegment* segs;      /* scan_bits / sub_count of device-unstuffed scans are written by k_unstuff_scan */
    uint32_t n_seg;
    uint32_t* rst;         /* restart-interval start positions of all segments (GjdSegment.rst_first / rst_count); written by the host
                              for host-unstuffed scans and by k_unstuff_write for the others */
    uint8_t* scan;
    const GjdLutSet* luts;
    const GjdLutCount* lutc;   /* pair LUT rings of the count pass, same index as luts */
    const GjdQuantSet* quants;
    uint2* sinfo;          /* [n_seq * GJD_SEQ_SUBS] */
    uint32_t* seq_seg;     /* [n_seq]  sequence -> segment */
    uint32_t* seq_base;    /* [n_seq]  coefficient index (within the segment) at the sequence start */
    uint2* seq_in;         /* [n_seq]  entry state {p, pos} of the sequence's first subsequence (k_huff_sync -> k_huff_fix) */
    uint32_t n_seq;
    uint32_t sub_log2;     /* log2 of the subsequence length in bits chosen for the batch (an image may use a shorter one: GjdSegment.sub_log2) */
    int16_t* coef;         /* [n_tile * 96 * 64] quantised coefficients, one 128-byte line per block */
    uint32_t* meta;        /* [n_tile * 96] per block: DC difference | kmax << 16 */
    int4* tilebase;        /* [n_tile] DC predictors at the tile start (x,y,z = comp 0,1,2) */
    uint32_t* tile_img;    /* [n_tile] tile -> image */
    uint32_t n_tile;
    uint32_t* tb_img;      /* [n_tb] transform block -> image */
    uint32_t n_tb;
    uint32_t tb_tiles;     /* tiles per transform block (1..GJD_TB_TILES, per batch: small batches get more blocks) */
    uint8_t* rgb;          /* output arena (images with out_is_ptr == 0) */
    uint32_t any_gray;        /* some image is single-component: k_transform_gray is launched too */
    uint32_t any_sub;         /* some image is chroma-subsampled: k_dc_sub_tiles + k_transform_sub are launched too */
    uint32_t any_rst;         /* some image has restart intervals */
    uint32_t entropy_mask;    /* which instantiations of k_huff_sync / k_huff_write have work: bit (2 * SUB + RST) */
    int* dcabs;               /* [n_tile * 96] absolute DC per coefficient line (subsampled images only) */
    uint32_t any_interleaved; /* some image wants GJD_LAYOUT_INTERLEAVED: selects the transform kernel variant */
    /* device-side unstuffing (images whose raw scan bytes were uploaded instead of host-unstuffed ones) */
    const GjdRaw* raws;    /* [n_raw] */
    uint32_t n_raw;
    uint32_t n_uchunk;     /* chunks of GJD_UCHUNK raw bytes over all raws */
    const uint8_t* raw;    /* raw arena (>= 16 bytes of slack on both sides of every scan) */
    uint32_t* uchunk_raw;  /* [n_uchunk] chunk -> raw */
    uint32_t* uchunk_cnt;  /* [n_uchunk] bytes the chunk keeps; then (k_unstuff_scan) its exclusive prefix inside the scan */
    uint32_t* uchunk_rst;  /* [n_uchunk] the same for the chunk's restart markers */
    uint32_t* raw_eoi;     /* [n_raw] position of the first EOI in the scan (0xffffffff: none) */
    uint32_t* raw_err;     /* [n_raw] smallest (position << 2 | GJD_UERR_*) (0xffffffff: none) */
    uint32_t* ustatus;     /* [n_img + 1] GJD_DEV_U* bits from the unstuffing (kept across runs of a resident batch), then 0 */
    uint32_t* status;      /* [n_img] GJD_DEV_* bits */
    uint32_t* nosync;      /* [1] sequences whose boundary fix-up ran off the end of the sequence */
};

/* Stage timing events recorded by gjd_launch_pipeline: e[0] start, e[1] after the scratch clears,
 * e[2] after the sync stages, e[3] after the write stages (count scan, coefficient write pass, DC
 * bases), e[4] after the transform. */
struct GjdStageEvents {
    cudaEvent_t e[5];
};

cudaError_t gjd_launch_build_luts(const GjdDhtSet* dht, GjdLutSet* luts, GjdLutCount* lutc, uint32_t n, cudaStream_t s);
cudaError_t gjd_launch_fill_maps(const GjdDevBatch& b, cudaStream_t s);
/* Device-side unstuffing of the batch's raw scans into the scan arena (no-op when n_raw == 0).  Runs once per staged
 * batch, before the first gjd_launch_pipeline; clears the status words it may set. */
cudaError_t gjd_launch_unstuff(const GjdDevBatch& b, cudaStream_t s, int* launches);
/* Clears scratch and runs every decode stage. transform_variant selects the transform kernel
 * implementation (0 = default).  Returns the number of kernel launches in *launches. */
cudaError_t gjd_launch_pipeline(const GjdDevBatch& b, cudaStream_t s, const GjdStageEvents* ev,
                                int transform_variant, int* launches, bool skip_sync = false, bool* attrs_set = nullptr);
/* Re-run of the boundary fix-up only (slow path when *nosync != 0). */
cudaError_t gjd_launch_fix(const GjdDevBatch& b, cudaStream_t s);
cudaError_t gjd_launch_export_coefs(const GjdDevBatch& b, int16_t* out, cudaStream_t s);
cudaError_t gjd_launch_test_color(const int16_t* y, const int16_t* cb, const int16_t* cr, size_t n,
                                  uint8_t* r, uint8_t* g, uint8_t* bl, cudaStream_t s);
cudaError_t gjd_launch_test_idct(const int16_t* coef, size_t nblocks, const int32_t* quant, int16_t* out,
                                 cudaStream_t s);

#endif
