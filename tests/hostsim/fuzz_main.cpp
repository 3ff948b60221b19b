/*
 * Mutation fuzzer over the product's parser + entropy/transform device functions stepped on the CPU
 * (hostsim.cpp), meant to be built with -fsanitize=address: a damaged file must end in a status code,
 * never in an out-of-bounds access, and never in an endless loop.  Test scaffolding, not product.
 *
 *   fuzz_main <cases> <seed> <file.jpg>...      prints one line: "cases N ok A corrupt B header C unsupported D"
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

extern "C" int hostsim_decode(const uint8_t* jpg, size_t len, int sub_log2_arg, int16_t* coef_out, uint8_t* rgb_out,
                              int32_t* width, int32_t* height, uint32_t stats[9]);
extern "C" int hostsim_parse(const uint8_t* jpg, size_t len, int32_t* w, int32_t* h);

static uint64_t rng_state;
static uint32_t rnd() {
    rng_state ^= rng_state << 13;
    rng_state ^= rng_state >> 7;
    rng_state ^= rng_state << 17;
    return (uint32_t)(rng_state >> 11);
}

/* offset of the first byte after the SOS header, or 0 */
static size_t scan_start(const std::vector<uint8_t>& f) {
    for (size_t i = 2; i + 3 < f.size();) {
        if (f[i] != 0xff) return 0;
        const size_t l = ((size_t)f[i + 2] << 8) | f[i + 3];
        if (f[i + 1] == 0xda) return i + 2 + l;
        i += 2 + l;
    }
    return 0;
}

static void mutate(std::vector<uint8_t>& f, bool in_scan) {
    const size_t ss = scan_start(f);
    size_t lo = 2, hi = f.size();
    if (ss && ss < f.size()) { if (in_scan) lo = ss; else hi = ss; }
    if (hi <= lo) return;
    const size_t at = lo + rnd() % (hi - lo);
    switch (rnd() % 7) {
        case 0: f[at] ^= (uint8_t)(1u << (rnd() & 7)); break;                      /* bit flip */
        case 1: f[at] = (uint8_t)rnd(); break;                                     /* random byte */
        case 2: f[at] = 0xff; break;                                               /* stray marker prefix */
        case 3: {                                                                  /* cut a chunk */
            const size_t n = 1 + rnd() % 400;
            f.erase(f.begin() + at, f.begin() + (at + n < f.size() ? at + n : f.size()));
            break;
        }
        case 4: {                                                                  /* insert random bytes */
            const size_t n = 1 + rnd() % 64;
            std::vector<uint8_t> ins(n);
            for (auto& b : ins) b = (uint8_t)rnd();
            f.insert(f.begin() + at, ins.begin(), ins.end());
            break;
        }
        case 5: f.resize(at); break;                                               /* truncate */
        default: {                                                                 /* run of one value */
            const uint8_t v = (rnd() & 1) ? 0x00 : 0xff;
            for (size_t i = at; i < f.size() && i < at + 1 + rnd() % 32; i++) f[i] = v;
        }
    }
}

int main(int argc, char** argv) {
    if (argc < 4) { fprintf(stderr, "usage: fuzz_main <cases> <seed> <file.jpg>...\n"); return 2; }
    const int cases = atoi(argv[1]);
    rng_state = 0x9e3779b97f4a7c15ull ^ (uint64_t)atoll(argv[2]);
    std::vector<std::vector<uint8_t>> seeds;
    for (int i = 3; i < argc; i++) {
        FILE* fp = fopen(argv[i], "rb");
        if (!fp) { fprintf(stderr, "cannot read %s\n", argv[i]); return 2; }
        std::vector<uint8_t> d;
        uint8_t buf[65536];
        size_t n;
        while ((n = fread(buf, 1, sizeof buf, fp)) > 0) d.insert(d.end(), buf, buf + n);
        fclose(fp);
        seeds.push_back(d);
    }
    long hist[4] = {0, 0, 0, 0}, skipped = 0;
    for (int it = 0; it < cases; it++) {
        std::vector<uint8_t> f = seeds[rnd() % seeds.size()];
        const int nmut = 1 + (int)(rnd() % 3);
        const bool in_scan = (rnd() % 4) != 0;          /* three quarters of the cases damage entropy data */
        for (int m = 0; m < nmut; m++) mutate(f, in_scan);
        /* exact-size heap copy: reads past the end of the file are ASan errors */
        std::vector<uint8_t> exact(f.begin(), f.end());
        int32_t w = 0, h = 0;
        int rc = hostsim_parse(exact.data(), exact.size(), &w, &h);
        if (rc == 0) {
            if ((int64_t)w * h > (int64_t)4 << 20) { skipped++; continue; }   /* damaged SOF: size explosion */
            const size_t pw = ((size_t)w + 7) / 8 * 8, ph = ((size_t)h + 7) / 8 * 8;
            std::vector<int16_t> coef(pw * ph * 3);
            std::vector<uint8_t> rgb((size_t)w * h * 3);
            uint32_t stats[9];
            rc = hostsim_decode(exact.data(), exact.size(), (it & 1) ? 8 : 10, coef.data(), rgb.data(), &w, &h, stats);
        }
        switch (rc) {
            case 0: hist[0]++; break;
            case -6: hist[1]++; break;
            case -4: hist[2]++; break;
            case -5: hist[3]++; break;
            default: fprintf(stderr, "case %d: unexpected status %d\n", it, rc); return 1;
        }
    }
    printf("cases %d ok %ld corrupt %ld header %ld unsupported %ld skipped %ld\n", cases, hist[0], hist[1], hist[2], hist[3],
           skipped);
    return 0;
}
