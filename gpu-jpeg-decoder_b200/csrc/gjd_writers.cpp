/*
 * gjd_writers.cpp — output formats on the caller's side of the ABI (host only, untimed).
 *
 * `.array` text format of JPEGParser::write (cpp-decoder/src/parser.cpp:197-210): "H W\n", then the
 * R, G and B planes, every value followed by one space, planes separated by '\n', nothing after the
 * last plane.  The reference spends as long in its ofstream loop as in the whole decode;
 * this writer formats into one buffer with a 256-entry table and issues a single fwrite.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/gjd.h"

extern "C" int gjd_write_array(const char* path, int32_t width, int32_t height, const uint8_t* rgb) {
    if (!path || !rgb || width <= 0 || height <= 0) return GJD_E_ARG;
    char tab[256][4];
    unsigned char tl[256];
    for (int v = 0; v < 256; v++) tl[v] = (unsigned char)snprintf(tab[v], 4, "%d", v);
    size_t px = (size_t)width * (size_t)height;
    char* buf = (char*)malloc(px * 3 * 4 + 64);
    if (!buf) return GJD_E_NOMEM;
    char* o = buf;
    o += sprintf(o, "%d %d\n", height, width);
    for (int c = 0; c < 3; c++) {
        const uint8_t* p = rgb + (size_t)c * px;
        for (size_t i = 0; i < px; i++) {
            unsigned v = p[i];
            memcpy(o, tab[v], 4);
            o += tl[v];
            *o++ = ' ';
        }
        if (c < 2) *o++ = '\n';
    }
    FILE* f = fopen(path, "wb");
    if (!f) { free(buf); return GJD_E_IO; }
    size_t n = (size_t)(o - buf);
    int rc = fwrite(buf, 1, n, f) == n ? GJD_OK : GJD_E_IO;
    if (fclose(f) != 0) rc = GJD_E_IO;
    free(buf);
    return rc;
}

extern "C" int gjd_write_ppm(const char* path, int32_t width, int32_t height, const uint8_t* rgb) {
    if (!path || !rgb || width <= 0 || height <= 0) return GJD_E_ARG;
    size_t px = (size_t)width * (size_t)height;
    uint8_t* buf = (uint8_t*)malloc(px * 3);
    if (!buf) return GJD_E_NOMEM;
    for (size_t i = 0; i < px; i++) {
        buf[3 * i] = rgb[i];
        buf[3 * i + 1] = rgb[px + i];
        buf[3 * i + 2] = rgb[2 * px + i];
    }
    FILE* f = fopen(path, "wb");
    if (!f) { free(buf); return GJD_E_IO; }
    int rc = GJD_OK;
    if (fprintf(f, "P6\n%d %d\n255\n", width, height) < 0) rc = GJD_E_IO;
    if (rc == GJD_OK && fwrite(buf, 1, px * 3, f) != px * 3) rc = GJD_E_IO;
    if (fclose(f) != 0) rc = GJD_E_IO;
    free(buf);
    return rc;
}
