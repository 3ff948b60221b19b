/*
 * gjd_oracle.h — CPU restatement of the reference cpp-decoder hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is linked into, imported by
 * or executed from the product (libgjd.so, the `decoder` CLI, the gjd Python
 * package).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may use it, and only as the checker / the CPU baseline.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this restatement
 *   (a) byte-for-byte against the reference's committed testing/ground_truth
 *       arrays (tests/golden/ground_truth.npz, made by tools/make_golden.py), and
 *   (b) coefficient-for-coefficient and pixel-for-pixel against the unmodified
 *       reference sources compiled into oracle/_ref/ (oracle/Makefile).
 *
 * All citations are relative to /root/reference/.
 */
#ifndef GJD_ORACLE_H
#define GJD_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Parsed container, cpp-decoder/src/parser.cpp:24-103 (JPEGParser::extract). */
typedef struct {
    int width, height;
    uint8_t quant[2][64];        /* zigzag order, as stored in DQT          */
    uint8_t bits[4][16];         /* index: 0=DC0(0x00) 1=DC1(0x01) 2=AC0(0x10) 3=AC1(0x11) */
    uint8_t vals[4][256];
    int nvals[4];
    uint8_t* scan;               /* unstuffed entropy data (malloc'd)       */
    size_t scan_len;
} gjdo_image;

/* 0 on success, <0 on a container the reference parser would not handle.   */
int  gjdo_extract(const uint8_t* jpg, size_t len, gjdo_image* out);
void gjdo_release(gjdo_image* img);

/* Full decode.  Outputs (any may be NULL):
 *   coef : quantised coefficients, int16, layout mcu*192 + comp*64 + k (k = zigzag index),
 *          padded geometry (ceil(W/8)*ceil(H/8) MCUs).     parser.cpp:105-134
 *   ycc  : IDCT output before colour conversion, int16, same layout but natural order
 *   rgb  : planar R,G,B u8, each H*W, cropped.            parser.cpp:172-193, color.cpp:7-18
 * Returns 0 on success.                                                    */
int gjdo_decode(const uint8_t* jpg, size_t len, int* width, int* height,
                int16_t* coef, int16_t* ycc, uint8_t* rgb);

/* Stage functions, exposed for per-stage tests. */
void gjdo_idct_block(int blk[64]);                    /* in: natural order dequantised; idct.cpp:34-133 */
void gjdo_dezigzag(const int zz[64], int nat[64]);    /* idct.cpp:7-32  */
void gjdo_color(int y, int cb, int cr, int* r, int* g, int* b);   /* color.cpp:3-18 */
void gjdo_color_plane(int y, uint8_t* out);            /* all (cb,cr) at one y; see .c */
void gjdo_color_pixels(const int16_t* y, const int16_t* cb, const int16_t* cr, size_t n, uint8_t* out);   /* bulk gjdo_color */
extern const uint8_t gjdo_zigzag_of_natural[64];      /* table[x][y] at idct.cpp:7-18 */

/* `.array` text writer, parser.cpp:197-210.  Returns bytes written or <0. */
long gjdo_write_array(const char* path, int width, int height, const uint8_t* rgb);

#ifdef __cplusplus
}
#endif
#endif
