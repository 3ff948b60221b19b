/*
 * ref_bench.cpp — multi-threaded timing driver over the UNMODIFIED reference cpp-decoder.
 *
 * TEST/BENCH INFRASTRUCTURE ONLY (bench.py --impl reference and the cpu_baseline leg).
 * Contains no decoder logic.  Each worker thread repeats, for the files assigned to it,
 * the timed region of cpp-decoder/benchmark/benchmark.cc:31-35:
 *
 *     JPEGParser parser(path);            // untimed: file read
 *     t0; parser.extract(); parser.decode(); t1;   // timed
 *                                         // write() is not called (untimed there too)
 *
 * usage: ref_bench <threads> <seconds_budget> <file.jpg> [more files...]
 * Files are dealt round-robin to threads; every thread loops over its files until the
 * budget is spent (at least one pass).  Prints one JSON line:
 *   {"threads":T,"images":N,"pixels":P,"busy_s_max":S,"wall_s":W,"mpix_per_s":P/S/1e6}
 * busy_s_max = the largest per-thread sum of timed regions (throughput = pixels / that).
 */
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
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

struct Tally { double busy = 0; long images = 0; long pixels = 0; };

int main(int argc, char** argv) {
    if (argc < 4) {
        fprintf(stderr, "usage: %s <threads> <seconds_budget> <file.jpg>...\n", argv[0]);
        return 2;
    }
    int threads = atoi(argv[1]);
    double budget = atof(argv[2]);
    std::vector<std::string> files(argv + 3, argv + argc);
    if (threads < 1) threads = 1;
    std::vector<Tally> tally(threads);
    std::vector<std::thread> pool;
    auto w0 = std::chrono::steady_clock::now();
    for (int t = 0; t < threads; t++) {
        pool.emplace_back([&, t]() {
            Tally& me = tally[t];
            size_t k = (size_t)t;
            do {
                std::string path = files[k % files.size()];
                k += (size_t)threads;
                JPEGParser* parser = new JPEGParser(path);
                auto t0 = std::chrono::steady_clock::now();
                parser->extract();
                parser->decode();
                auto t1 = std::chrono::steady_clock::now();
                me.busy += std::chrono::duration<double>(t1 - t0).count();
                me.images++;
                me.pixels += (long)parser->width * parser->height;
                delete parser;
            } while (me.busy < budget);
        });
    }
    for (auto& th : pool) th.join();
    auto w1 = std::chrono::steady_clock::now();
    double busy_max = 0; long images = 0, pixels = 0;
    for (auto& t : tally) { if (t.busy > busy_max) busy_max = t.busy; images += t.images; pixels += t.pixels; }
    printf("{\"threads\":%d,\"images\":%ld,\"pixels\":%ld,\"busy_s_max\":%.6f,\"wall_s\":%.6f,\"mpix_per_s\":%.4f}\n",
           threads, images, pixels, busy_max, std::chrono::duration<double>(w1 - w0).count(),
           pixels / busy_max / 1e6);
    return 0;
}
