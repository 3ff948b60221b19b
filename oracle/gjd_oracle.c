/*
 * gjd_oracle.c — plain-C restatement of the reference's CPU decoder (cpp-decoder).
 *
 * TEST INFRASTRUCTURE ONLY (see gjd_oracle.h).  Parity status: PINNED against the
 * reference's testing/ground_truth arrays and against the unmodified reference
 * sources compiled into oracle/_ref/ (tests/test_oracle.py).
 *
 * Build with -ffp-contract=off (the reference is built for baseline x86-64: separate
 * multiply and add, no FMA — color.cpp:10-16 depends on it).
 *
 * Every function cites the reference lines it follows (paths relative to /root/reference/).
 */
#include "gjd_oracle.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---- container parse: cpp-decoder/src/parser.cpp:24-103 -------------------------------- */

static int rd16(const uint8_t* d, size_t n, size_t pos, unsigned* v) {
    /* Stream::getMarker, cpp-decoder/utils/stream.cpp:30-36 */
    if (pos + 2 > n) return -1;
    *v = ((unsigned)d[pos] << 8) | d[pos + 1];
    return 0;
}

static int dht_slot(unsigned header) {
    /* the reference keys trees by the Tc<<4|Th byte and later uses [0],[1],[16],[17]
     * (parser.cpp:66-67,106,115,163-165) */
    switch (header) {
        case 0x00: return 0;
        case 0x01: return 1;
        case 0x10: return 2;
        case 0x11: return 3;
        default:   return -1;
    }
}

static int read_dht(const uint8_t* d, size_t n, size_t* pos, gjdo_image* o) {
    /* parser.cpp:63-67 / :70-74: size word, header byte, size-3 payload bytes;
     * HuffmanTree::createNodes, huffmanTree.cpp:20-32: 16 counts then the symbols. */
    unsigned size;
    if (rd16(d, n, *pos, &size)) return -1;
    *pos += 2;
    if (*pos + 1 > n) return -1;
    unsigned header = d[(*pos)++];
    size_t plen = (size_t)size - 3;
    if (size < 3 + 16 || *pos + plen > n) return -1;
    int slot = dht_slot(header);
    if (slot >= 0) {
        int total = 0;
        for (int i = 0; i < 16; i++) {
            o->bits[slot][i] = d[*pos + i];
            total += d[*pos + i];
        }
        if (total > 256 || 16 + (size_t)total > plen) return -1;
        memcpy(o->vals[slot], d + *pos + 16, (size_t)total);
        o->nvals[slot] = total;
    }
    *pos += plen;
    return 0;
}

int gjdo_extract(const uint8_t* d, size_t n, gjdo_image* o) {
    memset(o, 0, sizeof(*o));
    size_t pos = 0;
    for (;;) {
        unsigned marker, size;
        if (rd16(d, n, pos, &marker)) return -1;
        pos += 2;
        if (marker == 0xffd8) {                                   /* :33-34 */
            continue;
        } else if (marker == 0xffe0) {                            /* :35-38 */
            if (rd16(d, n, pos, &size)) return -1;
            pos += 2;
            pos += (size_t)size - 2;
        } else if (marker == 0xffdb) {                            /* :39-51 two back-to-back DQT */
            if (pos + 3 + 64 > n) return -1;
            pos += 2;                                             /* size, ignored   */
            pos += 1;                                             /* Pq/Tq, ignored  */
            memcpy(o->quant[0], d + pos, 64);
            pos += 64;
            if (rd16(d, n, pos, &marker)) return -1;
            pos += 2;
            if (marker != 0xffdb) return -2;                      /* ":50 Something went wrong" */
            if (pos + 3 + 64 > n) return -1;
            pos += 3;
            memcpy(o->quant[1], d + pos, 64);
            pos += 64;
        } else if (marker == 0xffc0) {                            /* :52-60 */
            if (rd16(d, n, pos, &size)) return -1;
            pos += 2;
            if (size < 7 || pos + size - 2 > n) return -1;
            o->height = (d[pos + 1] << 8) | d[pos + 2];
            o->width  = (d[pos + 3] << 8) | d[pos + 4];
            pos += (size_t)size - 2;
        } else if (marker == 0xffc4) {                            /* :61-77 four DHT segments */
            if (read_dht(d, n, &pos, o)) return -1;
            int count = 1;
            while (count < 4) {
                if (rd16(d, n, pos, &marker)) return -1;
                pos += 2;
                if (marker == 0xffc4) {
                    if (read_dht(d, n, &pos, o)) return -1;
                    count++;
                }
            }
        } else if (marker == 0xffda) {                            /* :78-99 */
            if (rd16(d, n, pos, &size)) return -1;
            pos += 2;
            pos += (size_t)size - 2;
            if (pos > n) return -1;
            o->scan = (uint8_t*)malloc(n - pos + 1);
            if (!o->scan) return -3;
            size_t m = 0;
            uint8_t prev = 0x00;
            for (;;) {
                if (pos >= n) { free(o->scan); o->scan = NULL; return -1; }  /* reference runs off the end */
                uint8_t cur = d[pos++];
                if (prev == 0xff && cur == 0xd9) break;
                if (cur == 0x00) {
                    if (prev != 0xff) o->scan[m++] = cur;
                } else {
                    o->scan[m++] = cur;
                }
                prev = cur;
            }
            if (m > 0) m--;                                       /* :98 pop the 0xff of EOI */
            o->scan_len = m;
            break;
        }
        /* any other word: stepped over two bytes at a time (:31-35 fall-through) */
    }
    return 0;
}

void gjdo_release(gjdo_image* img) {
    free(img->scan);
    img->scan = NULL;
}

/* ---- bit reader: cpp-decoder/utils/stream.cpp:8-22,44-51 -------------------------------- */

typedef struct {
    const uint8_t* data;
    size_t nbits;
    size_t pos;        /* bit position */
    int overrun;
} bitreader;

static int get_bit(bitreader* br) {                               /* stream.cpp:8-13 */
    if (br->pos >= br->nbits) { br->overrun = 1; br->pos++; return 0; }
    int v = (br->data[br->pos >> 3] >> (7 - (br->pos & 7))) & 1;
    br->pos++;
    return v;
}

static int get_nbits(bitreader* br, int n) {                      /* stream.cpp:15-22 */
    int v = 0;
    for (int i = 0; i < n; i++) v = v * 2 + get_bit(br);
    return v & 0xffff;                                            /* returns uint16_t */
}

static int decode_number(int code, int bits) {                    /* stream.cpp:44-51 */
    if (code == 0) return 0;      /* `1 << -1` in the reference; evaluates to "return 0" */
    int l = 1 << (code - 1);
    if (bits >= l) return bits;
    return bits - ((l << 1) - 1);
}

/* ---- Huffman: cpp-decoder/src/huffmanTree.cpp:20-68,85-110 ------------------------------ */

typedef struct {
    int mincode[17], maxcode[17], valptr[17];                     /* per code length 1..16 */
    const uint8_t* vals;
} hufftab;

static void build_hufftab(const uint8_t bits[16], const uint8_t* vals, hufftab* t) {
    /* leftmost-free-leaf insertion in DHT order (huffmanTree.cpp:34-68) is the canonical
     * assignment: consecutive codes within a length, (last+1)<<1 when the length grows. */
    int code = 0, k = 0;
    for (int l = 1; l <= 16; l++) {
        t->valptr[l] = k;
        t->mincode[l] = code;
        code += bits[l - 1];
        k += bits[l - 1];
        t->maxcode[l] = bits[l - 1] ? code - 1 : -1;
        code <<= 1;
    }
    t->vals = vals;
}

static int huff_symbol(const hufftab* t, bitreader* br, int* err) {
    /* traverseTree, huffmanTree.cpp:85-98: one bit per level until a leaf */
    int code = 0;
    for (int l = 1; l <= 16; l++) {
        code = (code << 1) | get_bit(br);
        if (t->maxcode[l] >= 0 && code >= t->mincode[l] && code <= t->maxcode[l])
            return t->vals[t->valptr[l] + code - t->mincode[l]];
    }
    *err = 1;   /* the reference would return an internal node's value here (:91-97) */
    return 0;
}

/* ---- IDCT: cpp-decoder/src/idct.cpp, constants idct.h:4-9 ------------------------------- */

enum { C1 = 2841, C2 = 2676, C3 = 2408, C5 = 1609, C6 = 1108, C7 = 565 };

const uint8_t gjdo_zigzag_of_natural[64] = {                      /* idct.cpp:7-18 */
     0,  1,  5,  6, 14, 15, 27, 28,
     2,  4,  7, 13, 16, 26, 29, 42,
     3,  8, 12, 17, 25, 30, 41, 43,
     9, 11, 18, 24, 31, 40, 44, 53,
    10, 19, 23, 32, 39, 45, 52, 54,
    20, 22, 33, 38, 46, 51, 55, 60,
    21, 34, 37, 47, 50, 56, 59, 61,
    35, 36, 48, 49, 57, 58, 62, 63};

void gjdo_dezigzag(const int zz[64], int nat[64]) {               /* idct.cpp:24-32 */
    for (int i = 0; i < 64; i++) nat[i] = zz[gjdo_zigzag_of_natural[i]];
}

static int clip9(int v) { return v < -256 ? -256 : (v > 255 ? 255 : v); }   /* idct.cpp:20-22 */

static void idct_row(int* b) {                                    /* idct.cpp:34-77 */
    int x0, x1, x2, x3, x4, x5, x6, x7, x8;
    x1 = (int)((unsigned)b[4] << 11); x2 = b[6]; x3 = b[2]; x4 = b[1]; x5 = b[7]; x6 = b[5]; x7 = b[3];
    if (!(x1 | x2 | x3 | x4 | x5 | x6 | x7)) {
        b[0] = b[1] = b[2] = b[3] = b[4] = b[5] = b[6] = b[7] = (int)((unsigned)b[0] << 3);
        return;
    }
    x0 = (int)((unsigned)b[0] << 11) + 128;
    x8 = C7 * (x4 + x5);
    x4 = x8 + (C1 - C7) * x4;
    x5 = x8 - (C1 + C7) * x5;
    x8 = C3 * (x6 + x7);
    x6 = x8 - (C3 - C5) * x6;
    x7 = x8 - (C3 + C5) * x7;
    x8 = x0 + x1;
    x0 -= x1;
    x1 = C6 * (x3 + x2);
    x2 = x1 - (C2 + C6) * x2;
    x3 = x1 + (C2 - C6) * x3;
    x1 = x4 + x6;
    x4 -= x6;
    x6 = x5 + x7;
    x5 -= x7;
    x7 = x8 + x3;
    x8 -= x3;
    x3 = x0 + x2;
    x0 -= x2;
    x2 = (181 * (x4 + x5) + 128) >> 8;
    x4 = (181 * (x4 - x5) + 128) >> 8;
    b[0] = (x7 + x1) >> 8;
    b[1] = (x3 + x2) >> 8;
    b[2] = (x0 + x4) >> 8;
    b[3] = (x8 + x6) >> 8;
    b[4] = (x8 - x6) >> 8;
    b[5] = (x0 - x4) >> 8;
    b[6] = (x3 - x2) >> 8;
    b[7] = (x7 - x1) >> 8;
}

static void idct_col(int* b) {                                    /* idct.cpp:79-122 */
    int x0, x1, x2, x3, x4, x5, x6, x7, x8;
    x1 = (int)((unsigned)b[8 * 4] << 8); x2 = b[8 * 6]; x3 = b[8 * 2]; x4 = b[8 * 1];
    x5 = b[8 * 7]; x6 = b[8 * 5]; x7 = b[8 * 3];
    if (!(x1 | x2 | x3 | x4 | x5 | x6 | x7)) {
        int v = clip9((b[0] + 32) >> 6);
        for (int i = 0; i < 8; i++) b[8 * i] = v;
        return;
    }
    x0 = (int)((unsigned)b[0] << 8) + 8192;
    x8 = C7 * (x4 + x5) + 4;
    x4 = (x8 + (C1 - C7) * x4) >> 3;
    x5 = (x8 - (C1 + C7) * x5) >> 3;
    x8 = C3 * (x6 + x7) + 4;
    x6 = (x8 - (C3 - C5) * x6) >> 3;
    x7 = (x8 - (C3 + C5) * x7) >> 3;
    x8 = x0 + x1;
    x0 -= x1;
    x1 = C6 * (x3 + x2) + 4;
    x2 = (x1 - (C2 + C6) * x2) >> 3;
    x3 = (x1 + (C2 - C6) * x3) >> 3;
    x1 = x4 + x6;
    x4 -= x6;
    x6 = x5 + x7;
    x5 -= x7;
    x7 = x8 + x3;
    x8 -= x3;
    x3 = x0 + x2;
    x0 -= x2;
    x2 = (181 * (x4 + x5) + 128) >> 8;
    x4 = (181 * (x4 - x5) + 128) >> 8;
    b[8 * 0] = clip9((x7 + x1) >> 14);
    b[8 * 1] = clip9((x3 + x2) >> 14);
    b[8 * 2] = clip9((x0 + x4) >> 14);
    b[8 * 3] = clip9((x8 + x6) >> 14);
    b[8 * 4] = clip9((x8 - x6) >> 14);
    b[8 * 5] = clip9((x0 - x4) >> 14);
    b[8 * 6] = clip9((x3 - x2) >> 14);
    b[8 * 7] = clip9((x7 - x1) >> 14);
}

void gjdo_idct_block(int blk[64]) {                               /* idct.cpp:124-133 */
    for (int i = 0; i < 8; i++) idct_row(blk + 8 * i);
    for (int i = 0; i < 8; i++) idct_col(blk + i);
}

/* ---- colour: cpp-decoder/utils/color.cpp:3-18 ------------------------------------------- */

static int clamp255(int c) { return c > 255 ? 255 : (c < 0 ? 0 : c); }      /* color.cpp:3-5 */

void gjdo_color(int y, int cb, int cr, int* R, int* G, int* B) {
    /* called as colorConversion(Y, Cr, Cb, ...) (parser.cpp:193); the int operands are
     * promoted to double, the results rounded to float, `+ 128` is a float add and the
     * cast truncates toward zero (color.cpp:10-16). */
    float r = cr * (2 - 2 * 0.299) + y;
    float b = cb * (2 - 2 * 0.114) + y;
    float g = (y - 0.114 * b - 0.299 * r) / 0.587;
    *R = clamp255((int)(r + 128));
    *G = clamp255((int)(g + 128));
    *B = clamp255((int)(b + 128));
}

/* Bulk form for the exhaustive colour test: all (cb, cr) in [-256,255]^2 at a fixed y,
 * out[((cb+256)*512 + (cr+256))*3 + {0,1,2}] = R,G,B. */
void gjdo_color_plane(int y, uint8_t* out) {
    for (int cb = -256; cb < 256; cb++)
        for (int cr = -256; cr < 256; cr++) {
            int r, g, b;
            gjdo_color(y, cb, cr, &r, &g, &b);
            uint8_t* o = out + ((size_t)(cb + 256) * 512 + (size_t)(cr + 256)) * 3;
            o[0] = (uint8_t)r; o[1] = (uint8_t)g; o[2] = (uint8_t)b;
        }
}

/* Bulk form for composed pins (tools/make_golden_gray.py, make_golden_sub.py): n (y, cb, cr) triples through
 * gjdo_color, out = planar R, G, B (n bytes each). */
void gjdo_color_pixels(const int16_t* y, const int16_t* cb, const int16_t* cr, size_t n, uint8_t* out) {
    for (size_t i = 0; i < n; i++) {
        int r, g, b;
        gjdo_color(y[i], cb[i], cr[i], &r, &g, &b);
        out[i] = (uint8_t)r; out[n + i] = (uint8_t)g; out[2 * n + i] = (uint8_t)b;
    }
}

/* ---- decode: cpp-decoder/src/parser.cpp:105-195 ----------------------------------------- */

static int build_block(bitreader* br, const hufftab* dc, const hufftab* ac, const uint8_t* q,
                       int* pred, int16_t* coef_out, int* samples) {
    /* JPEGParser::buildMCU, parser.cpp:105-142 */
    int err = 0;
    int zz[64] = {0};
    int qz[64] = {0};   /* quantised, for the coefficient dump */
    int code = huff_symbol(dc, br, &err);
    int bits = get_nbits(br, code);
    int dcv = decode_number(code, bits) + *pred;
    zz[0] = dcv * (int)q[0];
    qz[0] = dcv;
    int length = 1;
    while (length < 64) {
        code = huff_symbol(ac, br, &err);
        if (code == 0) break;
        if (code > 15) {
            length += code >> 4;
            code &= 0x0f;
        }
        bits = get_nbits(br, code);
        if (length < 64) {
            int v = decode_number(code, bits);
            zz[length] = v * (int)q[length];
            qz[length] = v;
            length++;
        }
    }
    *pred = dcv;
    if (coef_out) for (int k = 0; k < 64; k++) coef_out[k] = (int16_t)qz[k];
    gjdo_dezigzag(zz, samples);
    gjdo_idct_block(samples);
    return (err || br->overrun) ? -1 : 0;
}

int gjdo_decode(const uint8_t* jpg, size_t len, int* width, int* height,
                int16_t* coef, int16_t* ycc, uint8_t* rgb) {
    gjdo_image img;
    int rc = gjdo_extract(jpg, len, &img);
    if (rc) return rc;
    if (width) *width = img.width;
    if (height) *height = img.height;
    if (!coef && !ycc && !rgb) { gjdo_release(&img); return 0; }

    hufftab tab[4];
    for (int i = 0; i < 4; i++) build_hufftab(img.bits[i], img.vals[i], &tab[i]);

    const int W = img.width, H = img.height;
    const int xb = (W + 7) / 8, yb = (H + 7) / 8;                 /* parser.cpp:150-154 */
    bitreader br = {img.scan, img.scan_len * 8, 0, 0};
    int pred[3] = {0, 0, 0};                                      /* parser.cpp:145-147 */
    int bad = 0;
    int* planes = (int*)malloc(sizeof(int) * 3 * (size_t)xb * yb * 64);
    if (!planes) { gjdo_release(&img); return -3; }

    for (int y = 0; y < yb; y++) {                                /* parser.cpp:161-167 */
        for (int x = 0; x < xb; x++) {
            size_t mcu = (size_t)y * xb + x;
            for (int c = 0; c < 3; c++) {
                /* Y: trees 0x00/0x10 + quant 0; Cb,Cr: trees 0x01/0x11 + quant 1 (:163-165) */
                int hf = c ? 1 : 0;
                int* samples = planes + (mcu * 3 + c) * 64;
                if (build_block(&br, &tab[hf], &tab[2 + hf], img.quant[hf], &pred[c],
                                coef ? coef + (mcu * 3 + c) * 64 : NULL, samples))
                    bad = 1;
                if (ycc) for (int i = 0; i < 64; i++) ycc[(mcu * 3 + c) * 64 + i] = (int16_t)samples[i];
            }
        }
    }

    if (rgb) {                                                    /* parser.cpp:172-193 */
        for (int py = 0; py < H; py++) {
            for (int px = 0; px < W; px++) {
                size_t mcu = (size_t)(py / 8) * xb + (px / 8);
                int idx = (py & 7) * 8 + (px & 7);
                int Y  = planes[(mcu * 3 + 0) * 64 + idx];
                int Cb = planes[(mcu * 3 + 1) * 64 + idx];        /* 2nd scan component */
                int Cr = planes[(mcu * 3 + 2) * 64 + idx];        /* 3rd scan component */
                int r, g, b;
                gjdo_color(Y, Cb, Cr, &r, &g, &b);
                size_t p = (size_t)py * W + px;
                rgb[p] = (uint8_t)r;
                rgb[(size_t)H * W + p] = (uint8_t)g;
                rgb[2 * (size_t)H * W + p] = (uint8_t)b;
            }
        }
    }
    free(planes);
    gjdo_release(&img);
    return bad ? -4 : 0;
}

/* ---- `.array` writer: cpp-decoder/src/parser.cpp:197-210 -------------------------------- */

long gjdo_write_array(const char* path, int width, int height, const uint8_t* rgb) {
    FILE* f = fopen(path, "w");
    if (!f) return -1;
    long n = fprintf(f, "%d %d\n", height, width);
    size_t px = (size_t)width * height;
    for (int c = 0; c < 3; c++) {
        for (size_t i = 0; i < px; i++) n += fprintf(f, "%d ", rgb[c * px + i]);
        if (c < 2) { fputc('\n', f); n++; }
    }
    fclose(f);
    return n;
}
