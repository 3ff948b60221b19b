/*
 * jpeg_coefs.c — dumps the quantised DCT coefficients of a JPEG file as the reference's vendored IJG libjpeg 9e reads
 * them (jpeg_read_coefficients), for every component, over the MCU-padded block grid.
 *
 * TEST INFRASTRUCTURE ONLY (tools/make_golden_sub.py): the pin for the coefficients of chroma-subsampled files, which the
 * reference's own decoder cannot read (its parser ignores the sampling factors, cpp-decoder/src/parser.cpp:52-60).
 * Linked against $(REF)/jpeglib-implementation/libjpeg/lib/libjpeg.a by oracle/Makefile; contains no decoder logic.
 *
 * usage: jpeg_coefs in.jpg out.bin
 * out.bin (little-endian int32 unless noted): width, height, ncomp, then per component:
 *   h_samp, v_samp, blocks per row (padded), block rows (padded), 64 x uint16 quantisation values (natural order),
 *   rows x cols x 64 int16 coefficients (natural order inside a block, as libjpeg stores them).
 */
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
#include <jpeglib.h>

int main(int argc, char** argv) {
    if (argc != 3) { fprintf(stderr, "usage: %s in.jpg out.bin\n", argv[0]); return 2; }
    FILE* in = fopen(argv[1], "rb");
    FILE* out = fopen(argv[2], "wb");
    if (!in || !out) { perror("open"); return 1; }
    struct jpeg_decompress_struct c;
    struct jpeg_error_mgr err;
    c.err = jpeg_std_error(&err);
    jpeg_create_decompress(&c);
    jpeg_stdio_src(&c, in);
    jpeg_read_header(&c, TRUE);
    jvirt_barray_ptr* arrays = jpeg_read_coefficients(&c);
    int32_t hdr[3] = {(int32_t)c.image_width, (int32_t)c.image_height, c.num_components};
    fwrite(hdr, 4, 3, out);
    for (int ci = 0; ci < c.num_components; ci++) {
        jpeg_component_info* comp = &c.comp_info[ci];
        /* the virtual arrays are padded to whole MCUs (jdcoefct.c) */
        int32_t cols = (int32_t)((comp->width_in_blocks + comp->h_samp_factor - 1) / comp->h_samp_factor * comp->h_samp_factor);
        int32_t rows = (int32_t)((comp->height_in_blocks + comp->v_samp_factor - 1) / comp->v_samp_factor * comp->v_samp_factor);
        int32_t ch[4] = {comp->h_samp_factor, comp->v_samp_factor, cols, rows};
        fwrite(ch, 4, 4, out);
        uint16_t q[64];
        for (int i = 0; i < 64; i++) q[i] = comp->quant_table->quantval[i];
        fwrite(q, 2, 64, out);
        for (int32_t r = 0; r < rows; r++) {
            JBLOCKARRAY ba = (*c.mem->access_virt_barray)((j_common_ptr)&c, arrays[ci], (JDIMENSION)r, 1, FALSE);
            for (int32_t x = 0; x < cols; x++) {
                int16_t blk[64];
                for (int i = 0; i < 64; i++) blk[i] = (int16_t)ba[0][x][i];
                fwrite(blk, 2, 64, out);
            }
        }
    }
    jpeg_finish_decompress(&c);
    jpeg_destroy_decompress(&c);
    fclose(in);
    fclose(out);
    return 0;
}
