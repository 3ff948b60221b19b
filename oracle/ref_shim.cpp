/*
 * ref_shim.cpp — C entry points over the UNMODIFIED reference cpp-decoder.
 *
 * TEST INFRASTRUCTURE ONLY.  This file contains no decoder logic: it drives the
 * reference's own public object API (JPEGParser ctor / extract / decode,
 * cpp-decoder/src/parser.h:42-69, the same three calls cpp-decoder/main.cpp:11-14 and
 * cpp-decoder/benchmark/benchmark.cc:31-35 make) and taps two values the reference
 * never exposes:
 *
 *  - the dequantised zigzag-order coefficients of every 8x8 block, captured at link
 *    time by `-Wl,--wrap` on IDCT::IDCT(std::vector<int>&) — buildMCU constructs an
 *    IDCT from `arr` right after the entropy decode and the `* quantTables` multiply
 *    (cpp-decoder/src/parser.cpp:136), so the argument of that constructor is exactly
 *    the block in zigzag order;
 *  - the private R/G/B channel vectors and width/height, read by including the reference
 *    header in THIS translation unit under `#define private public` (the reference objects themselves are
 *    compiled untouched from /root/reference by oracle/Makefile).
 *
 * Outputs go to oracle/_ref/ only (git-ignored, travels to the GPU box).
 */
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

/* Pre-include every standard header the reference headers pull in, then open the
 * reference class for reading; the reference objects themselves are compiled untouched. */
#include <algorithm>
#include <cstdint>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <iterator>
#include <unordered_map>
#define private public
#include "src/parser.h"   /* -I$(REF)/cpp-decoder */
#undef private

static thread_local std::vector<int>* g_tap = nullptr;

extern "C" void __real__ZN4IDCTC1ERSt6vectorIiSaIiEE(void* self, std::vector<int>& base);
extern "C" void __wrap__ZN4IDCTC1ERSt6vectorIiSaIiEE(void* self, std::vector<int>& base) {
    if (g_tap) g_tap->insert(g_tap->end(), base.begin(), base.end());
    __real__ZN4IDCTC1ERSt6vectorIiSaIiEE(self, base);
}

extern "C" {

/* Decode `path` with the reference.  On return *deq holds nblocks*64 dequantised zigzag
 * coefficients in decode order (mcu-major, Y/Cb/Cr), *rgb holds 3*H*W ints (R plane,
 * G plane, B plane).  Both are malloc'd; free with gjdref_free.  Returns 0. */
int gjdref_decode(const char* path, int* width, int* height, int32_t** deq, long* ndeq,
                  int32_t** rgb) {
    std::string p(path);
    std::vector<int> tap;
    JPEGParser parser(p);
    parser.extract();
    g_tap = deq ? &tap : nullptr;
    parser.decode();
    g_tap = nullptr;
    *width = parser.width;
    *height = parser.height;
    if (deq) {
        *deq = (int32_t*)malloc(tap.size() * sizeof(int32_t));
        memcpy(*deq, tap.data(), tap.size() * sizeof(int32_t));
        *ndeq = (long)tap.size();
    }
    if (rgb) {
        size_t n = (size_t)parser.width * parser.height;
        *rgb = (int32_t*)malloc(3 * n * sizeof(int32_t));
        memcpy(*rgb, parser.channels->getR().data(), n * sizeof(int32_t));
        memcpy(*rgb + n, parser.channels->getG().data(), n * sizeof(int32_t));
        memcpy(*rgb + 2 * n, parser.channels->getB().data(), n * sizeof(int32_t));
    }
    return 0;
}

/* The two 64-byte quantisation tables the reference parsed (zigzag order). */
int gjdref_quant(const char* path, uint8_t q[128]) {
    std::string p(path);
    JPEGParser parser(p);
    parser.channels = nullptr;   /* the reference ctor leaves it uninitialised; its dtor tests it */
    parser.extract();
    if (parser.quantTables[0].size() < 64 || parser.quantTables[1].size() < 64) return -1;
    memcpy(q, parser.quantTables[0].data(), 64);
    memcpy(q + 64, parser.quantTables[1].data(), 64);
    return 0;
}

void gjdref_free(void* p) { free(p); }

/* extract()+decode() wall time in seconds, timed exactly as
 * cpp-decoder/benchmark/benchmark.cc:31-35 (constructor/file read and write() untimed). */
double gjdref_time_decode(const char* path, long* pixels) {
    std::string p(path);
    JPEGParser parser(p);
    auto t0 = std::chrono::steady_clock::now();
    parser.extract();
    parser.decode();
    auto t1 = std::chrono::steady_clock::now();
    if (pixels) *pixels = (long)parser.width * parser.height;
    return std::chrono::duration<double>(t1 - t0).count();
}

}  // extern "C"
