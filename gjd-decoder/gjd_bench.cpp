/*
 * gjd_bench — the reference's throughput benchmark driver, on top of libgjd's C ABI.
 *
 *   ./gjd_bench <image_dir> [--batch 3000] [--iters 10] [--gpus N] [--e2e] [--tile]
 *
 * Mirrors cuda-decoder/benchmark_thoughput/benchmark.cu (getAllImages :14-22, batches of 3000 :30,
 * 10 iterations :129, only the device work timed :61-70, images/s and MB/s of compressed file bytes
 * :95-111) without Google Benchmark, and adds what the reference's driver lacks:
 *   --gpus N  the directory is sharded over N devices with gjd_partition (greedy LPT on padded pixels + 5 x file bytes),
 *             one host thread + one gjd_ctx per device, no exchange between shards;
 *   --e2e     a second figure through gjd_decode_batch with host buffers (parse, unstuff, H2D, D2H timed);
 *   --tile    repeat the directory cyclically until a batch is full (the reference stops at the last file).
 * Prints one JSON line.
 */
#include <dirent.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <string>
#include <thread>
#include <vector>

#include "../include/gjd.h"

namespace {

bool has_jpeg_ext(const std::string& n) {
    size_t dot = n.find_last_of('.');
    if (dot == std::string::npos) return false;
    std::string e = n.substr(dot + 1);
    for (auto& c : e) c = (char)tolower(c);
    return e == "jpg" || e == "jpeg";
}

/* getAllImages (benchmark.cu:14-22), sorted so that runs are reproducible */
std::vector<std::string> list_images(const std::string& dir) {
    std::vector<std::string> out;
    DIR* d = opendir(dir.c_str());
    if (!d) return out;
    while (dirent* e = readdir(d))
        if (has_jpeg_ext(e->d_name)) out.push_back(dir + "/" + e->d_name);
    closedir(d);
    std::sort(out.begin(), out.end());
    return out;
}

bool read_file(const std::string& path, std::vector<uint8_t>& out) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? (size_t)n : 0);
    bool ok = n >= 0 && fread(out.data(), 1, out.size(), f) == out.size();
    fclose(f);
    return ok;
}

struct ShardResult {
    int rc = 0;
    double dev_ms = 0, e2e_s = 0;      /* summed over iterations */
    uint64_t pixels = 0, bytes = 0, images = 0, failed = 0, e2e_failed = 0;
    std::string err;
};

void run_shard(int device, const std::vector<const std::vector<uint8_t>*>& files, int batch, int iters, bool e2e,
               ShardResult* r) {
    gjd_ctx* ctx = nullptr;
    r->rc = gjd_create(&ctx, device, 0);
    if (r->rc) { r->err = gjd_strerror(r->rc); return; }
    for (size_t first = 0; first < files.size() && r->rc == 0; first += (size_t)batch) {
        const int32_t n = (int32_t)std::min<size_t>((size_t)batch, files.size() - first);
        std::vector<gjd_item> items((size_t)n);
        uint64_t bytes = 0;
        for (int32_t i = 0; i < n; i++) {
            memset(&items[i], 0, sizeof(gjd_item));
            items[i].jpg = files[first + i]->data();
            items[i].len = files[first + i]->size();
            bytes += items[i].len;
        }
        /* == the untimed extract()/allocate() loop + DeviceData upload (benchmark.cu:49-60) */
        gjd_batch* b = nullptr;
        r->rc = gjd_batch_prepare(ctx, items.data(), n, &b);
        if (r->rc) { r->err = std::string(gjd_strerror(r->rc)) + " " + gjd_last_cuda_error(ctx); break; }
        uint64_t info[5];
        gjd_batch_info(ctx, b, info);
        for (int it = -2; it < iters; it++) {           /* two warm-up launches */
            r->rc = gjd_batch_run(ctx, b);              /* == batchDecodeKernel<<<batch,256>>> (:66) */
            if (r->rc == 0) {
                int s = gjd_batch_sync(ctx, b);
                if (s != GJD_OK && s != GJD_E_CORRUPT) r->rc = s;
            }
            if (r->rc) { r->err = std::string(gjd_strerror(r->rc)) + " " + gjd_last_cuda_error(ctx); break; }
            if (it < 0) continue;
            float ms[4];
            gjd_stage_times(ctx, ms);                   /* == the cudaEvent pair (:61-70) */
            r->dev_ms += ms[3];
        }
        for (int32_t i = 0; i < n; i++) r->failed += gjd_batch_status(ctx, b, i) != GJD_OK;
        gjd_batch_free(ctx, b);                         /* == clean() (:88-93) */
        if (r->rc) break;
        r->pixels += info[0];
        r->bytes += bytes;
        r->images += (uint64_t)n;
        if (e2e) {
            /* output offsets and the buffer size come from the SAME loop over the items whose header parses: an image
             * that parses but later fails to unstuff still owns its region */
            std::vector<size_t> off((size_t)n, (size_t)-1);
            size_t total = 0;
            for (int32_t i = 0; i < n; i++) {
                int32_t w = 0, h = 0;
                if (gjd_peek(items[i].jpg, items[i].len, &w, &h) == GJD_OK) {
                    off[(size_t)i] = total;
                    total += (size_t)w * h * 3;
                }
            }
            uint8_t* host = (uint8_t*)gjd_host_alloc(total + 16);
            if (!host) { r->rc = GJD_E_NOMEM; r->err = "gjd_host_alloc"; break; }
            for (int32_t i = 0; i < n; i++) items[i].rgb = off[(size_t)i] == (size_t)-1 ? nullptr : host + off[(size_t)i];
            int rc = gjd_decode_batch(ctx, items.data(), n);     /* warm-up: sizes the staging buffers */
            auto t0 = std::chrono::steady_clock::now();
            for (int it = 0; it < iters && rc == GJD_OK; it++) rc = gjd_decode_batch(ctx, items.data(), n);
            r->e2e_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            gjd_host_free(host);
            if (rc != GJD_OK) { r->rc = rc; r->err = std::string("e2e: ") + gjd_strerror(rc) + " " + gjd_last_cuda_error(ctx); break; }
            uint64_t bad = 0;
            for (int32_t i = 0; i < n; i++) bad += items[i].status != GJD_OK;
            r->e2e_failed += bad;
        }
    }
    gjd_destroy(ctx);
}

}  // namespace

int main(int argc, char** argv) {
    std::string dir;
    int batch = 3000, iters = 10, gpus = 1;
    bool e2e = false, tile = false;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        if (a == "--batch" && i + 1 < argc) batch = atoi(argv[++i]);
        else if (a == "--iters" && i + 1 < argc) iters = atoi(argv[++i]);
        else if (a == "--gpus" && i + 1 < argc) gpus = atoi(argv[++i]);
        else if (a == "--e2e") e2e = true;
        else if (a == "--tile") tile = true;
        else if (dir.empty()) dir = a;
    }
    if (dir.empty() || batch < 1 || iters < 1 || gpus < 1) {
        fprintf(stderr, "usage: gjd_bench <image_dir> [--batch 3000] [--iters 10] [--gpus N] [--e2e] [--tile]\n");
        return 1;
    }
    std::vector<std::string> paths = list_images(dir);
    if (paths.empty()) { fprintf(stderr, "gjd_bench: no .jpg files in %s\n", dir.c_str()); return 2; }
    std::vector<std::vector<uint8_t>> data(paths.size());
    for (size_t i = 0; i < paths.size(); i++)
        if (!read_file(paths[i], data[i])) { fprintf(stderr, "gjd_bench: cannot read %s\n", paths[i].c_str()); return 2; }

    /* the work list: every file once, or the directory tiled up to gpus * batch images */
    std::vector<const std::vector<uint8_t>*> work;
    const size_t total = tile ? std::max(paths.size(), (size_t)batch * (size_t)gpus) : paths.size();
    for (size_t i = 0; i < total; i++) work.push_back(&data[i % data.size()]);

    /* shard by padded pixels plus a term for the scan bytes: no exchange step between devices */
    std::vector<gjd_item> items(work.size());
    for (size_t i = 0; i < work.size(); i++) {
        memset(&items[i], 0, sizeof(gjd_item));
        items[i].jpg = work[i]->data();
        items[i].len = work[i]->size();
    }
    std::vector<uint64_t> weight(work.size()), load((size_t)gpus);
    std::vector<int32_t> part(work.size());
    gjd_partition_costs(items.data(), (int32_t)items.size(), GJD_PARTITION_BYTE_WEIGHT, weight.data());
    gjd_partition(weight.data(), (int32_t)items.size(), gpus, part.data(), load.data());
    std::vector<std::vector<const std::vector<uint8_t>*>> shard((size_t)gpus);
    size_t rejected = 0;
    for (size_t i = 0; i < work.size(); i++) {
        if (part[i] < 0) { rejected++; continue; }
        shard[(size_t)part[i]].push_back(work[i]);
    }

    std::vector<ShardResult> res((size_t)gpus);
    std::vector<std::thread> th;
    auto w0 = std::chrono::steady_clock::now();
    for (int g = 0; g < gpus; g++) th.emplace_back(run_shard, g, std::cref(shard[(size_t)g]), batch, iters, e2e, &res[(size_t)g]);
    for (auto& t : th) t.join();
    double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - w0).count();

    double dev_ms_max = 0, e2e_max = 0;
    uint64_t pixels = 0, bytes = 0, images = 0, failed = 0, e2e_failed = 0;
    for (int g = 0; g < gpus; g++) {
        if (res[(size_t)g].rc) { fprintf(stderr, "gjd_bench: device %d: %s\n", g, res[(size_t)g].err.c_str()); return 3; }
        dev_ms_max = std::max(dev_ms_max, res[(size_t)g].dev_ms);
        e2e_max = std::max(e2e_max, res[(size_t)g].e2e_s);
        pixels += res[(size_t)g].pixels; bytes += res[(size_t)g].bytes; images += res[(size_t)g].images; failed += res[(size_t)g].failed;
        e2e_failed += res[(size_t)g].e2e_failed;
    }
    const double s = dev_ms_max * 1e-3;   /* all devices run concurrently: the slowest shard bounds the job */
    printf("{\"images\": %llu, \"failed\": %llu, \"rejected\": %zu, \"gpus\": %d, \"batch\": %d, \"iters\": %d, "
           "\"device_ms_per_iter\": %.3f, \"images_per_s\": %.1f, \"compressed_MB_per_s\": %.1f, \"MPixel_per_s\": %.1f",
           (unsigned long long)images, (unsigned long long)failed, rejected, gpus, batch, iters, dev_ms_max / iters,
           images * (double)iters / s, bytes * (double)iters / s / 1e6, pixels * (double)iters / s / 1e6);
    if (e2e) printf(", \"e2e_MPixel_per_s\": %.1f, \"e2e_failed\": %llu", pixels * (double)iters / e2e_max / 1e6, (unsigned long long)e2e_failed);
    printf(", \"wall_s\": %.2f}\n", wall);
    return 0;
}
