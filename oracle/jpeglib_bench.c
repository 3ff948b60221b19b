/*
 * jpeglib_bench.c — multi-threaded in-process timing driver over the IJG libjpeg 9e the reference vendors
 * (jpeglib-implementation/libjpeg: include/jpeglib.h + lib/libjpeg.a, linked where they lie).
 *
 * TEST/BENCH INFRASTRUCTURE ONLY (bench.py's second CPU baseline).  No decoder logic.
 * The reference times jpeglib as system("djpeg f > /dev/null") including the process spawn
 * (jpeglib-implementation/benchmark/benchmark.cc:31-39); this driver times the library call sequence djpeg
 * makes — jpeg_mem_src, jpeg_read_header, jpeg_start_decompress, jpeg_read_scanlines to RGB (default
 * JDCT_ISLOW), jpeg_finish_decompress — on files already in memory, which is the fairer baseline.
 *
 * usage: jpeglib_bench <threads> <seconds_budget> <file.jpg>...
 * prints {"threads":T,"images":N,"pixels":P,"busy_s_max":S,"mpix_per_s":P/S/1e6}
 */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "jpeglib.h"

typedef struct { unsigned char* data; unsigned long len; } file_t;
typedef struct { int id, threads, nfiles; file_t* files; double budget, busy; long images, pixels; } worker_t;

static double now(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + ts.tv_nsec * 1e-9;
}

static void* work(void* arg) {
    worker_t* w = (worker_t*)arg;
    unsigned char* row = NULL;
    size_t rowcap = 0;
    long k = w->id;
    do {
        file_t* f = &w->files[k % w->nfiles];
        k += w->threads;
        double t0 = now();
        struct jpeg_decompress_struct cinfo;
        struct jpeg_error_mgr jerr;
        cinfo.err = jpeg_std_error(&jerr);
        jpeg_create_decompress(&cinfo);
        jpeg_mem_src(&cinfo, f->data, f->len);
        jpeg_read_header(&cinfo, TRUE);
        cinfo.out_color_space = JCS_RGB;
        jpeg_start_decompress(&cinfo);
        size_t need = (size_t)cinfo.output_width * cinfo.output_components;
        if (need > rowcap) { row = (unsigned char*)realloc(row, need); rowcap = need; }
        while (cinfo.output_scanline < cinfo.output_height) {
            JSAMPROW r = row;
            jpeg_read_scanlines(&cinfo, &r, 1);
        }
        long px = (long)cinfo.output_width * cinfo.output_height;
        jpeg_finish_decompress(&cinfo);
        jpeg_destroy_decompress(&cinfo);
        w->busy += now() - t0;
        w->images++;
        w->pixels += px;
    } while (w->busy < w->budget);
    free(row);
    return NULL;
}

int main(int argc, char** argv) {
    if (argc < 4) { fprintf(stderr, "usage: %s <threads> <seconds_budget> <file.jpg>...\n", argv[0]); return 2; }
    int threads = atoi(argv[1]);
    double budget = atof(argv[2]);
    int nfiles = argc - 3;
    if (threads < 1) threads = 1;
    file_t* files = (file_t*)calloc((size_t)nfiles, sizeof(file_t));
    for (int i = 0; i < nfiles; i++) {
        FILE* f = fopen(argv[3 + i], "rb");
        if (!f) { fprintf(stderr, "cannot open %s\n", argv[3 + i]); return 2; }
        fseek(f, 0, SEEK_END);
        long n = ftell(f);
        fseek(f, 0, SEEK_SET);
        files[i].data = (unsigned char*)malloc((size_t)n);
        files[i].len = (unsigned long)n;
        if (fread(files[i].data, 1, (size_t)n, f) != (size_t)n) return 2;
        fclose(f);
    }
    worker_t* ws = (worker_t*)calloc((size_t)threads, sizeof(worker_t));
    pthread_t* th = (pthread_t*)calloc((size_t)threads, sizeof(pthread_t));
    for (int t = 0; t < threads; t++) {
        ws[t].id = t; ws[t].threads = threads; ws[t].nfiles = nfiles; ws[t].files = files; ws[t].budget = budget;
        pthread_create(&th[t], NULL, work, &ws[t]);
    }
    double busy = 0;
    long images = 0, pixels = 0;
    for (int t = 0; t < threads; t++) {
        pthread_join(th[t], NULL);
        if (ws[t].busy > busy) busy = ws[t].busy;
        images += ws[t].images;
        pixels += ws[t].pixels;
    }
    printf("{\"threads\":%d,\"images\":%ld,\"pixels\":%ld,\"busy_s_max\":%.4f,\"mpix_per_s\":%.3f}\n", threads, images, pixels,
           busy, pixels / busy / 1e6);
    return 0;
}
