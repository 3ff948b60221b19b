/*
 * gjd_api.cu — host runtime behind the C ABI of include/gjd.h: context, worker pool, pinned staging,
 * device arenas, sub-batch pipeline.  Host C++ only (no kernels here); compiled by nvcc for the
 * CUDA runtime headers.
 *
 * Replaces the per-image extract()/allocate()/clean() storm of the reference
 * (cuda-decoder/src/parser.cu:324-471, 684-700; ~26 CUDA API calls per image) with O(1) API calls
 * per sub-batch: one staging copy, one descriptor upload, one pipeline of launches, merged D2H copies.
 */
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>   /* header-only; ranges cost nothing unless a profiler is attached */

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <new>
#include <string>
#include <chrono>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/gjd.h"
#include "gjd_kernels.h"
#include "gjd_parse.h"
#include "gjd_types.h"

#define GJD_VERSION_STR "0.1"
#define GJD_STR2(x) #x
#define GJD_STR(x) GJD_STR2(x)

namespace {

/* NVTX range per stage of a call (parse / stage+unstuff / H2D / GPU stages / D2H), after the reference's
 * nvtxRangePush around its decode (cuda-decoder/benchmark/benchmark.cu:9,41,70; README.md:147-166). */
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

/* Every entry point runs on the context's device and leaves the caller's current device as it found it
 * (a multi-GPU host such as PyTorch keeps its own notion of the current device). */
struct DeviceGuard {
    int prev = -1;
    bool ok = false;
    explicit DeviceGuard(int device) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        ok = (prev == device) || cudaSetDevice(device) == cudaSuccess;
        if (prev == device) prev = -1;   /* nothing to restore */
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

/* No C++ exception crosses the C ABI: allocation failures of the host containers become GJD_E_NOMEM. */
#define GJD_API_TRY try {
#define GJD_API_CATCH } catch (const std::bad_alloc&) { return GJD_E_NOMEM; } catch (...) { return GJD_E_CUDA; }

/* ---------------------------------------------------------------- worker pool --------- */

class Pool {
public:
    explicit Pool(int n) {
        for (int i = 0; i < n; i++) th_.emplace_back([this] { loop(); });
    }
    ~Pool() {
        {
            std::lock_guard<std::mutex> l(m_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto& t : th_) t.join();
    }
    /* fn(i) for i in [0,n), dynamically chunked; the caller participates. */
    void parallel_for(size_t n, size_t chunk, const std::function<void(size_t)>& fn) {
        if (n == 0) return;
        if (th_.empty() || n <= chunk) {
            for (size_t i = 0; i < n; i++) fn(i);
            return;
        }
        {
            std::lock_guard<std::mutex> l(m_);
            fn_ = &fn; n_ = n; chunk_ = chunk; next_.store(0); pending_ = (int)th_.size(); gen_++;
        }
        cv_.notify_all();
        work();
        {
            std::unique_lock<std::mutex> l(m_);
            done_.wait(l, [this] { return pending_ == 0; });
            fn_ = nullptr;
        }
        if (failed_.exchange(false)) throw std::bad_alloc();   /* a worker ran out of memory: reported by the caller's thread */
    }
    int size() const { return (int)th_.size() + 1; }

private:
    void work() {
        for (;;) {
            size_t b = next_.fetch_add(chunk_);
            if (b >= n_) break;
            size_t e = std::min(n_, b + chunk_);
            for (size_t i = b; i < e; i++) {
                try { (*fn_)(i); } catch (...) { failed_.store(true); }
            }
        }
    }
    void loop() {
        uint64_t seen = 0;
        for (;;) {
            {
                std::unique_lock<std::mutex> l(m_);
                cv_.wait(l, [&] { return stop_ || gen_ != seen; });
                if (stop_) return;
                seen = gen_;
            }
            work();
            {
                std::lock_guard<std::mutex> l(m_);
                if (--pending_ == 0) done_.notify_one();
            }
        }
    }
    std::vector<std::thread> th_;
    std::mutex m_;
    std::condition_variable cv_, done_;
    const std::function<void(size_t)>* fn_ = nullptr;
    size_t n_ = 0, chunk_ = 1;
    std::atomic<size_t> next_{0};
    std::atomic<bool> failed_{false};
    int pending_ = 0;
    uint64_t gen_ = 0;
    bool stop_ = false;
};

/* ---------------------------------------------------------------- buffers ------------- */

/* Device arena that only grows.  Stream-ordered (cudaMallocAsync / cudaFreeAsync on the owning slot's stream): growing
 * one slot's arena does not synchronise the device while the other slots' kernels and copies are in flight, which a
 * cudaFree would.  The context raises the pool's release threshold so freed blocks stay with the process. */
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes, cudaStream_t s) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFreeAsync(p, s);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 4096;
        cudaError_t e = cudaMallocAsync(&p, want, s);
        if (e != cudaSuccess) { cudaGetLastError(); e = cudaMallocAsync(&p, bytes, s); want = bytes; }
        if (e == cudaSuccess) cap = want; else p = nullptr;
        return e;
    }
    void release(cudaStream_t s) { if (p) cudaFreeAsync(p, s); p = nullptr; cap = 0; }
};

struct PinBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 4096;
        cudaError_t e = cudaHostAlloc(&p, want, cudaHostAllocDefault);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct ItemGeom {
    int32_t status = GJD_OK;   /* host-side status (parse / unstuff) */
    int32_t width = 0, height = 0;
    size_t out_bytes = 0;      /* bytes of the image's output region */
    size_t pitch = 0, pstride = 0;
    int32_t layout = GJD_LAYOUT_PLANAR;
    bool jpg_pinned = false;   /* the caller's JPEG bytes lie in page-locked memory: DMA straight from them */
};

}  // namespace

/* sub-batches in flight inside gjd_decode_batch: one being staged by the host, one uploading, one in the kernels */
#define GJD_SLOTS 3

/* ---------------------------------------------------------------- batch (one slot) ---- */

struct gjd_batch {
    int32_t n = 0;
    std::vector<GjdParsed> parsed;
    std::vector<ItemGeom> geom;
    std::vector<GjdImage> imgs;      /* device descriptors, only for items with status OK */
    std::vector<GjdSegment> segs;    /* restart intervals of those images (one per image without restarts) */
    std::vector<int32_t> img_of;     /* item -> index into imgs, or -1 */
    std::vector<GjdDhtSet> dhts;
    std::vector<GjdQuantSet> quants;
    std::vector<GjdRaw> raws;        /* scans unstuffed on the device */
    std::vector<uint32_t> rst;       /* restart-interval start positions (bits) of all segments */
    PinBuf h_scan, h_raw, h_meta, h_status;
    DevBuf d_raw, d_uchunk, d_rawmeta, d_ustatus, d_dcabs;
    DevBuf d_scan, d_meta, d_luts, d_lutc, d_sinfo, d_seqmaps, d_coef, d_blkmeta, d_tilebase, d_tileimg, d_tbimg, d_rgb, d_status;
    GjdDevBatch view{};
    GjdStageEvents ev{};
    bool ev_ok = false;
    cudaStream_t stream = nullptr;
    size_t scan_alloc = 0, rgb_alloc = 0;
    uint64_t pixels = 0, padded_pixels = 0, scan_bytes = 0, subseqs = 0;
    int launches = 0, unstuff_launches = 0;
    bool ran = false;

    void release() {
        h_scan.release(); h_raw.release(); h_meta.release(); h_status.release();
        for (DevBuf* d : {&d_raw, &d_uchunk, &d_rawmeta, &d_ustatus, &d_dcabs, &d_scan, &d_meta, &d_luts, &d_lutc, &d_sinfo, &d_seqmaps, &d_coef, &d_blkmeta, &d_tilebase, &d_tileimg,
                          &d_tbimg, &d_rgb, &d_status})
            d->release(stream);
        if (ev_ok) for (auto& e : ev.e) cudaEventDestroy(e);
        ev_ok = false;
        if (stream) { cudaStreamSynchronize(stream); cudaStreamDestroy(stream); }
        stream = nullptr;
    }
};

struct gjd_ctx {
    int device = 0;
    size_t scratch = 0;
    Pool* pool = nullptr;
    gjd_batch slot[GJD_SLOTS];
    gjd_batch* last = nullptr;     /* batch whose stage events gjd_stage_times reports */
    std::string cuda_err;
    int transform_variant = 0;
    uint64_t min_subseqs = 1200000;   /* ~4 waves of 148 SMs x 2048 threads */
    uint32_t max_tb_tiles = GJD_TB_TILES;   /* GJD_TB_TILES_MAX: measurement knob (power of two, 4..128) */
    int force_sub_log2 = 0;           /* GJD_SUB_LOG2: fixed subsequence length (tests, experiments) */
    bool device_unstuff = true;       /* GJD_DEVICE_UNSTUFF=0: unstuff every scan on the host (A/B, tests) */
    bool kernel_attrs_set = false;    /* the > 48 KB dynamic shared memory opt-in of k_huff_write (per context) */
};

namespace {

int cuda_fail(gjd_ctx* ctx, cudaError_t e, const char* what) {
    if (ctx) ctx->cuda_err = std::string(what) + ": " + cudaGetErrorString(e);
    return e == cudaErrorMemoryAllocation ? GJD_E_NOMEM : GJD_E_CUDA;
}
#define GJD_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return cuda_fail(ctx, e_, #call); } while (0)

int batch_init(gjd_ctx* ctx, gjd_batch* b) {
    if (!b->stream) GJD_TRY(cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking));
    if (!b->ev_ok) {
        for (auto& e : b->ev.e) GJD_TRY(cudaEventCreate(&e));
        b->ev_ok = true;
    }
    return GJD_OK;
}

uint64_t fnv1a(const void* p, size_t n) {
    const uint8_t* s = (const uint8_t*)p;
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < n; i++) { h ^= s[i]; h *= 1099511628211ull; }
    return h;
}

/* Estimated device bytes one image needs (scan + sinfo + coefficient lines + block meta + rgb). */
size_t estimate_bytes(const gjd_item& it, const ItemGeom& g) {
    size_t pw = align_up((size_t)g.width, 16), ph = align_up((size_t)g.height, 16);
    return pw * ph * 6 + pw * ph / 8 + pw * ph / 4 + g.out_bytes + it.len * 4 + 65536;
}

/*
 * Parse + unstuff items [first, first+n) into the batch's pinned staging, build descriptors, and
 * enqueue the H2D copies + table/map kernels on the batch stream.
 * pre: parsed/geom for the items are already filled by parse_all (indexed from `first`).
 */
int batch_stage(gjd_ctx* ctx, gjd_batch* b, gjd_item* items, const GjdParsed* parsed, const ItemGeom* geom,
                int32_t n) {
    int rc = batch_init(ctx, b);
    if (rc) return rc;
    b->n = n;
    b->ran = false;
    b->geom.assign(geom, geom + n);
    b->img_of.assign(n, -1);
    b->imgs.clear();
    b->segs.clear();
    b->dhts.clear();
    b->quants.clear();

    /* Where each scan is unstuffed.  On the device (k_unstuff_*), the default: the raw bytes from the first scan byte to
     * the end of the file are uploaded, straight from the caller's buffer when that is page-locked (adjacent files merge
     * into one copy), through the pinned h_raw staging otherwise.  On the host (gjd_unstuff_segments) with
     * GJD_DEVICE_UNSTUFF=0. */
    b->raws.clear();
    b->rst.clear();
    std::vector<uint8_t> mode(n, 0);                 /* 1 = device-unstuffed */
    std::vector<size_t> soff(n + 1, 0);              /* scan arena (device) */
    std::vector<size_t> hoff(n + 1, 0);              /* h_scan staging: host-unstuffed scans only */
    std::vector<size_t> roff(n, 0);                  /* raw arena: first scan byte of device-unstuffed scans */
    struct Copy { const uint8_t* src; size_t dst, len; };
    std::vector<Copy> dma;                           /* user buffer -> raw arena */
    size_t raw_staged = 16, raw_total = 0;           /* 16 bytes of slack in front of the arena */
    for (int32_t i = 0; i < n; i++) {
        size_t room = 0, hroom = 0;
        if (b->geom[i].status == GJD_OK) {
            const size_t rawlen = (size_t)(parsed[i].file_end - parsed[i].scan_begin);
            mode[i] = ctx->device_unstuff;
            if (mode[i]) {
                room = align_up(rawlen + 64, 128);
                if (!b->geom[i].jpg_pinned) { roff[i] = raw_staged; raw_staged += align_up(rawlen, 16) + 16; }
            } else {
                room = hroom = align_up(items[i].len + 64, 128);
            }
        }
        soff[i + 1] = soff[i] + room;
        hoff[i + 1] = hoff[i] + hroom;
    }
    raw_total = raw_staged;
    for (int32_t i = 0; i < n; i++) {                /* page-locked inputs: runs of files that follow each other in memory */
        if (!mode[i] || !b->geom[i].jpg_pinned) continue;
        const bool joins = !dma.empty() && i > 0 && mode[i - 1] && b->geom[i - 1].jpg_pinned &&
                           items[i].jpg == items[i - 1].jpg + items[i - 1].len;
        if (!joins) {
            raw_total = align_up(raw_total, 16) + 16;
            dma.push_back({parsed[i].scan_begin, raw_total, 0});
        }
        Copy& c = dma.back();
        roff[i] = c.dst + (size_t)(parsed[i].scan_begin - c.src);
        c.len = (size_t)(parsed[i].file_end - c.src);
        raw_total = c.dst + c.len;
    }
    raw_total += 64;
    b->scan_alloc = soff[n] + 256;
    GJD_TRY(b->h_scan.ensure(hoff[n] + 256));
    GJD_TRY(b->h_raw.ensure(raw_staged + 64));
    uint8_t* hs = (uint8_t*)b->h_scan.p;
    uint8_t* hr = (uint8_t*)b->h_raw.p;

    std::vector<long> ulen(n, 0);                    /* unstuffed length (device-unstuffed scans: the raw length, an upper bound) */
    std::vector<std::vector<uint32_t>> rst_off(n), rst_len(n);   /* restart intervals (images with DRI only) */
    nvtxRangePushA("gjd: stage scans (host unstuffing / raw copies into pinned staging)");
    ctx->pool->parallel_for((size_t)n, 4, [&](size_t i) {
        if (b->geom[i].status != GJD_OK) return;
        const GjdParsed& p = parsed[i];
        if (mode[i]) {
            ulen[i] = (long)(p.file_end - p.scan_begin);
            if (!b->geom[i].jpg_pinned) memcpy(hr + roff[i], p.scan_begin, (size_t)ulen[i]);
            return;
        }
        uint8_t* dst = hs + hoff[i];
        long m = p.restart_interval ? gjd_unstuff_segments(p.scan_begin, p.file_end, dst, &rst_off[i], &rst_len[i])
                                    : gjd_unstuff(p.scan_begin, p.file_end, dst);
        if (m < 0) { b->geom[i].status = (int32_t)m; return; }
        memset(dst + m, 0, 16);
        ulen[i] = m;
    });
    nvtxRangePop();
    NvtxRange r_desc("gjd: descriptors + H2D + table kernels");

    /* subsequence length for this batch: the longest that still leaves >= GJD_MIN_SUBSEQS threads.  Where most of the bits
     * are chroma-subsampled scans an eighth of that is enough: half of the cold decodes of a 4:2:0 subsequence start in the
     * wrong block of the MCU and are repeated, which costs less the fewer subsequence boundaries there are (1000 x 1024^2
     * 4:2:0: sync pass 1.93 ms at 2^10 bits, 0.92 ms at 2^13; the 4:4:4 set of the same size is fastest at 2^10-2^11;
     * profiles/r02I_sub_log2.txt) */
    uint32_t sub_log2 = GJD_SUB_LOG2_MIN;
    {
        uint64_t total_bits = 0, sub_bits_total = 0;
        for (int32_t i = 0; i < n; i++) {
            const uint64_t bits = (uint64_t)std::max<long>(ulen[i], 0) * 8;
            total_bits += bits;
            if (b->geom[i].status == GJD_OK && (uint32_t)parsed[i].hs * parsed[i].vs > 1u) sub_bits_total += bits;
        }
        const bool mostly_sub = sub_bits_total * 2 > total_bits;
        const uint64_t want = mostly_sub ? std::max<uint64_t>(ctx->min_subseqs / 8, 1) : ctx->min_subseqs;
        /* 2^13 only for the subsampled case: 4:4:4 batches of any size are fastest at 2^12 (C3 14.76 ms against 15.03, 4000
         * images of the C5 set 23.30 against 24.30; profiles/r02I_sub_log2.txt, r02Z_c5_sub_log2.txt) */
        const uint32_t longest = mostly_sub ? GJD_SUB_LOG2_MAX : GJD_SUB_LOG2_MAX - 1u;
        while (sub_log2 < longest && (total_bits >> (sub_log2 + 1)) >= want) sub_log2++;
        if (ctx->force_sub_log2) sub_log2 = (uint32_t)ctx->force_sub_log2;
    }

    /* tiles per transform block: GJD_TB_TILES for big batches (one cold, unprefetched tile in 32); halved while
     * the batch would give the 148 SMs x 8 resident blocks fewer than four waves of blocks */
    uint32_t tb_tiles = ctx->max_tb_tiles;
    {
        uint64_t tiles = 0;
        for (int32_t i = 0; i < n; i++) {
            if (b->geom[i].status != GJD_OK) continue;
            const uint32_t mw = 8u * parsed[i].hs, mh = 8u * parsed[i].vs;
            const uint64_t mc = (uint64_t)((b->geom[i].width + mw - 1) / mw) * (uint64_t)((b->geom[i].height + mh - 1) / mh);
            tiles += (mc * (parsed[i].components == 1 ? 1u : parsed[i].hs * parsed[i].vs + 2u) + 95u) / 96u;
        }
        while (tb_tiles > 4u && tiles / tb_tiles < 4u * 148u * 8u) tb_tiles >>= 1;
    }

    /* descriptors */
    std::unordered_map<uint64_t, std::vector<uint32_t>> lut_index, quant_index;
    uint32_t seq = 0, tile = 0, tb = 0, uchunk = 0;
    size_t rgb_off = 0;
    int last_lut = -1, last_quant = -1;            /* most images of a batch share their tables with the one before */
    b->pixels = b->padded_pixels = b->scan_bytes = b->subseqs = 0;
    for (int32_t i = 0; i < n; i++) {
        ItemGeom& g = b->geom[i];
        if (g.status != GJD_OK) continue;
        const GjdParsed& p = parsed[i];
        if ((uint64_t)ulen[i] * 8 >= 0xffff0000ull) { g.status = GJD_E_UNSUPPORTED; continue; }
        {   /* coefficient slot indices (sinfo counts, seq_base, the write pass's idx0) are 32-bit */
            const uint32_t mw = 8u * p.hs, mh = 8u * p.vs;
            const uint64_t mc = (uint64_t)((g.width + mw - 1) / mw) * (uint64_t)((g.height + mh - 1) / mh);
            if (mc * (p.components == 1 ? 1u : p.hs * p.vs + 2u) * 64u >= 0xffffffffull) { g.status = GJD_E_UNSUPPORTED; continue; }
        }
        GjdImage d;
        memset(&d, 0, sizeof(d));
        d.width = (uint32_t)g.width;
        d.height = (uint32_t)g.height;
        d.gray = p.components == 1 ? 1u : 0u;
        d.hs = (uint8_t)(d.gray ? 1u : p.hs);
        d.vs = (uint8_t)(d.gray ? 1u : p.vs);
        d.mcus_x = (d.width + 8u * d.hs - 1u) / (8u * d.hs);
        d.mcu_count = d.mcus_x * ((d.height + 8u * d.vs - 1u) / (8u * d.vs));
        d.tile_first = tile;
        const uint32_t bpm = d.gray ? 1u : (uint32_t)d.hs * d.vs + 2u;   /* 8x8 blocks per MCU */
        d.tile_count = (d.mcu_count * bpm + 95u) / 96u;              /* a tile is 96 coefficient lines */
        d.tb_first = tb;
        d.tb_count = (d.tile_count + tb_tiles - 1) / tb_tiles;
        d.row_pitch = (uint32_t)g.pitch;
        d.plane_stride = g.pstride;
        d.layout = (uint16_t)g.layout;
        if (items[i].rgb_on_device && items[i].rgb) {
            d.out_is_ptr = 1;
            d.out_off = (uint64_t)(uintptr_t)items[i].rgb;
        } else {
            d.out_is_ptr = 0;
            d.out_off = rgb_off;
            rgb_off += align_up(g.out_bytes, 256);
        }
        d.restart_interval = p.restart_interval;
        const uint32_t nseg = p.restart_interval ? (d.mcu_count + p.restart_interval - 1) / p.restart_interval : 1u;
        if (p.restart_interval && !mode[i] && rst_off[i].size() != nseg) { g.status = GJD_E_CORRUPT; continue; }
        uint16_t lut_id = 0;
        /* Huffman table set: dedup by content */
        {
            GjdDhtSet key = p.dht;              /* the LUT ring depends on the scan's table selectors too */
            memcpy(key.tabsel, p.tabsel, 8);
            memset(key.pad, 0, sizeof(key.pad));
            int found = last_lut >= 0 && memcmp(&b->dhts[last_lut], &key, sizeof(GjdDhtSet)) == 0 ? last_lut : -1;
            uint64_t h = found < 0 ? fnv1a(&key, sizeof(GjdDhtSet)) : 0;
            auto& cand = lut_index[h];
            if (found < 0)
                for (uint32_t id : cand)
                    if (memcmp(&b->dhts[id], &key, sizeof(GjdDhtSet)) == 0) { found = (int)id; break; }
            if (found < 0) {
                if (b->dhts.size() >= 65535) { g.status = GJD_E_NOMEM; continue; }
                found = (int)b->dhts.size();
                b->dhts.push_back(key);
                cand.push_back((uint32_t)found);
            }
            lut_id = (uint16_t)found;
            last_lut = found;
        }
        {
            GjdQuantSet q;
            for (int c = 0; c < 3; c++)
                for (int k = 0; k < 64; k++) q.q[c][k] = p.quant[p.qsel[c]][k];
            for (int c = 0; c < 3; c++)
                for (int j = 0; j < 32; j++)
                    q.packed[c][j] = (uint32_t)q.q[c][2 * j] | ((uint32_t)q.q[c][2 * j + 1] << 24);
            int found = last_quant >= 0 && memcmp(&b->quants[last_quant], &q, sizeof(q)) == 0 ? last_quant : -1;
            uint64_t h = found < 0 ? fnv1a(&q, sizeof(q)) : 0;
            auto& cand = quant_index[h];
            if (found < 0)
                for (uint32_t id : cand)
                    if (memcmp(&b->quants[id], &q, sizeof(q)) == 0) { found = (int)id; break; }
            if (found < 0) {
                if (b->quants.size() >= 65535) { g.status = GJD_E_NOMEM; continue; }
                found = (int)b->quants.size();
                b->quants.push_back(q);
                cand.push_back((uint32_t)found);
            }
            d.quant_id = (uint16_t)found;
            last_quant = found;
        }
        d.seg_first = (uint32_t)b->segs.size();
        d.seg_count = 1;
        {   /* one segment per image; its restart intervals lie back to back inside it (table of their start positions) */
            GjdSegment sg;
            memset(&sg, 0, sizeof(sg));
            sg.scan_off = soff[i];
            sg.scan_bits = (uint32_t)ulen[i] * 8u;
            /* an image that would fill less than half a thread block (GJD_SEQ_SUBS subsequences) at the batch's length takes a
             * shorter one: a 256x256 picture in a big batch is 32 subsequences of 2^12 bits, i.e. one block of 256 threads of
             * which 224 wait; at 2^10 bits it is 128 that each do a quarter of the work */
            uint32_t sl = sub_log2;
            if (!ctx->force_sub_log2)
                while (sl > GJD_SUB_LOG2_IMAGE_MIN && (((uint64_t)sg.scan_bits + (1ull << sl) - 1) >> sl) < GJD_SEQ_SUBS / 2u) sl--;
            sg.sub_log2 = (uint8_t)sl;
            sg.sub_count = std::max<uint32_t>(1u, (uint32_t)(((uint64_t)sg.scan_bits + (1ull << sl) - 1) >> sl));
            sg.seq_first = seq;
            sg.seq_count = (sg.sub_count + GJD_SEQ_SUBS - 1) / GJD_SEQ_SUBS;
            sg.img = (uint32_t)b->imgs.size();
            sg.blk_first = 0;
            sg.blk_count = bpm * d.mcu_count;
            if (p.restart_interval && nseg > 1) {
                sg.rst_first = (uint32_t)b->rst.size();
                sg.rst_count = nseg - 1;
                sg.rst_blocks = bpm * p.restart_interval;
                /* (device-unstuffed scans: k_unstuff_write fills the table in) */
                for (uint32_t k = 1; k < nseg; k++) b->rst.push_back(mode[i] ? 0u : rst_off[i][k] * 8u);
            }
            sg.lut_id = lut_id;
            sg.ny = (uint8_t)(d.gray ? 1u : (uint32_t)d.hs * d.vs);
            if (mode[i]) {
                GjdRaw r;
                memset(&r, 0, sizeof(r));
                r.raw_off = roff[i];
                r.raw_len = (uint32_t)ulen[i];
                r.chunk_first = uchunk;
                r.chunk_count = (uint32_t)(((roff[i] & 15u) + (size_t)ulen[i] + GJD_UCHUNK - 1) / GJD_UCHUNK);
                if (r.chunk_count == 0) r.chunk_count = 1;
                r.seg = (uint32_t)b->segs.size();
                r.img = (uint32_t)b->imgs.size();
                r.dri = p.restart_interval != 0;
                uchunk += r.chunk_count;
                b->raws.push_back(r);
            }
            seq += sg.seq_count;
            b->subseqs += sg.sub_count;
            b->segs.push_back(sg);
        }
        tile += d.tile_count;
        tb += d.tb_count;
        b->img_of[i] = (int32_t)b->imgs.size();
        b->imgs.push_back(d);
        b->pixels += (uint64_t)d.width * d.height;
        b->padded_pixels += (uint64_t)d.mcu_count * 64 * d.hs * d.vs;   /* per full-resolution plane */
        b->scan_bytes += (uint64_t)ulen[i];
    }
    b->rgb_alloc = rgb_off;

    const uint32_t n_img = (uint32_t)b->imgs.size();
    GjdDevBatch& v = b->view;
    memset(&v, 0, sizeof(v));
    v.n_img = n_img;
    v.n_seg = (uint32_t)b->segs.size();
    for (const GjdImage& im : b->imgs) {
        v.any_interleaved |= im.layout != GJD_LAYOUT_PLANAR;
        v.any_gray |= im.gray;
        v.any_sub |= (uint32_t)im.hs * im.vs > 1u;
        v.any_rst |= im.restart_interval != 0u;
        /* an image with restart intervals may still hold a segment without a restart table (one interval): both kernels */
        const uint32_t sub = !im.gray && (uint32_t)im.hs * im.vs > 1u ? 2u : 0u;
        v.entropy_mask |= (1u << sub) | (im.restart_interval != 0u ? 1u << (sub + 1u) : 0u);
    }
    v.n_seq = seq;
    v.sub_log2 = sub_log2;
    v.n_tile = tile;
    v.n_tb = tb;
    v.tb_tiles = tb_tiles;
    if (n_img == 0) return GJD_OK;

    /* one pinned block for descriptors + DHT sets + quant sets, one H2D copy */
    size_t o_img = 0;
    size_t o_seg = align_up(o_img + sizeof(GjdImage) * n_img, 256);
    size_t o_dht = align_up(o_seg + sizeof(GjdSegment) * b->segs.size(), 256);
    size_t o_q = align_up(o_dht + sizeof(GjdDhtSet) * b->dhts.size(), 256);
    size_t o_raw = align_up(o_q + sizeof(GjdQuantSet) * b->quants.size(), 256);
    size_t o_rst = align_up(o_raw + sizeof(GjdRaw) * b->raws.size(), 256);
    size_t meta = align_up(o_rst + sizeof(uint32_t) * b->rst.size(), 256);
    GJD_TRY(b->h_meta.ensure(meta));
    if (!b->raws.empty()) memcpy((uint8_t*)b->h_meta.p + o_raw, b->raws.data(), sizeof(GjdRaw) * b->raws.size());
    if (!b->rst.empty()) memcpy((uint8_t*)b->h_meta.p + o_rst, b->rst.data(), sizeof(uint32_t) * b->rst.size());
    memcpy((uint8_t*)b->h_meta.p + o_img, b->imgs.data(), sizeof(GjdImage) * n_img);
    memcpy((uint8_t*)b->h_meta.p + o_seg, b->segs.data(), sizeof(GjdSegment) * b->segs.size());
    memcpy((uint8_t*)b->h_meta.p + o_dht, b->dhts.data(), sizeof(GjdDhtSet) * b->dhts.size());
    memcpy((uint8_t*)b->h_meta.p + o_q, b->quants.data(), sizeof(GjdQuantSet) * b->quants.size());

    GJD_TRY(b->d_meta.ensure(meta, b->stream));
    GJD_TRY(b->d_scan.ensure(b->scan_alloc, b->stream));
    GJD_TRY(b->d_luts.ensure(sizeof(GjdLutSet) * b->dhts.size(), b->stream));
    GJD_TRY(b->d_lutc.ensure(sizeof(GjdLutCount) * b->dhts.size(), b->stream));
    GJD_TRY(b->d_sinfo.ensure((size_t)seq * GJD_SEQ_SUBS * sizeof(uint2), b->stream));
    GJD_TRY(b->d_seqmaps.ensure((size_t)seq * 4 * sizeof(uint32_t), b->stream));
    GJD_TRY(b->d_coef.ensure((size_t)tile * 96 * 64 * sizeof(int16_t), b->stream));
    GJD_TRY(b->d_tbimg.ensure((size_t)tb * sizeof(uint32_t), b->stream));
    GJD_TRY(b->d_blkmeta.ensure((size_t)tile * 96 * sizeof(uint32_t), b->stream));
    GJD_TRY(b->d_tilebase.ensure((size_t)tile * sizeof(int4), b->stream));
    GJD_TRY(b->d_tileimg.ensure((size_t)tile * sizeof(uint32_t), b->stream));
    if (v.any_sub) GJD_TRY(b->d_dcabs.ensure((size_t)tile * 96 * sizeof(int), b->stream));
    GJD_TRY(b->d_rgb.ensure(std::max<size_t>(b->rgb_alloc, 256), b->stream));
    GJD_TRY(b->d_status.ensure((size_t)(n_img + 1) * sizeof(uint32_t), b->stream));
    GJD_TRY(b->d_ustatus.ensure((size_t)(n_img + 1) * sizeof(uint32_t), b->stream));
    GJD_TRY(b->d_raw.ensure(raw_total + 64, b->stream));
    GJD_TRY(b->d_uchunk.ensure((size_t)(uchunk + 1) * 3 * sizeof(uint32_t), b->stream));
    GJD_TRY(b->d_rawmeta.ensure((b->raws.size() + 1) * 2 * sizeof(uint32_t), b->stream));
    GJD_TRY(b->h_status.ensure((size_t)(n_img + 1) * sizeof(uint32_t)));

    v.imgs = (const GjdImage*)((uint8_t*)b->d_meta.p + o_img);
    v.segs = (GjdSegment*)((uint8_t*)b->d_meta.p + o_seg);
    v.scan = (uint8_t*)b->d_scan.p;
    v.rst = (uint32_t*)((uint8_t*)b->d_meta.p + o_rst);
    v.raws = (const GjdRaw*)((uint8_t*)b->d_meta.p + o_raw);
    v.n_raw = (uint32_t)b->raws.size();
    v.n_uchunk = uchunk;
    v.raw = (const uint8_t*)b->d_raw.p;
    v.uchunk_raw = (uint32_t*)b->d_uchunk.p;
    v.uchunk_cnt = (uint32_t*)b->d_uchunk.p + uchunk;
    v.uchunk_rst = (uint32_t*)b->d_uchunk.p + 2 * (size_t)uchunk;
    v.raw_eoi = (uint32_t*)b->d_rawmeta.p;
    v.raw_err = (uint32_t*)b->d_rawmeta.p + b->raws.size();
    v.ustatus = (uint32_t*)b->d_ustatus.p;
    v.luts = (const GjdLutSet*)b->d_luts.p;
    v.lutc = (const GjdLutCount*)b->d_lutc.p;
    v.quants = (const GjdQuantSet*)((uint8_t*)b->d_meta.p + o_q);
    v.sinfo = (uint2*)b->d_sinfo.p;
    v.seq_seg = (uint32_t*)b->d_seqmaps.p;
    v.seq_base = (uint32_t*)b->d_seqmaps.p + seq;
    v.seq_in = (uint2*)((uint32_t*)b->d_seqmaps.p + 2 * (size_t)seq);
    v.coef = (int16_t*)b->d_coef.p;
    v.tb_img = (uint32_t*)b->d_tbimg.p;
    v.meta = (uint32_t*)b->d_blkmeta.p;
    v.tilebase = (int4*)b->d_tilebase.p;
    v.tile_img = (uint32_t*)b->d_tileimg.p;
    v.dcabs = (int*)b->d_dcabs.p;
    v.rgb = (uint8_t*)b->d_rgb.p;
    v.status = (uint32_t*)b->d_status.p;
    v.nosync = (uint32_t*)b->d_status.p + n_img;

    {   /* host-unstuffed scans: staging -> scan arena, runs of neighbours in one copy */
        int32_t i = 0;
        while (i < n) {
            if (hoff[i + 1] == hoff[i]) { i++; continue; }
            int32_t j = i + 1;
            while (j < n && hoff[j + 1] != hoff[j] && soff[j] - soff[i] == hoff[j] - hoff[i]) j++;
            GJD_TRY(cudaMemcpyAsync((uint8_t*)b->d_scan.p + soff[i], hs + hoff[i], hoff[j] - hoff[i], cudaMemcpyHostToDevice, b->stream));
            i = j;
        }
    }
    if (raw_staged > 16)
        GJD_TRY(cudaMemcpyAsync(b->d_raw.p, b->h_raw.p, raw_staged, cudaMemcpyHostToDevice, b->stream));
    for (const auto& c : dma)   /* page-locked JPEG bytes: no staging copy at all */
        GJD_TRY(cudaMemcpyAsync((uint8_t*)b->d_raw.p + c.dst, c.src, c.len, cudaMemcpyHostToDevice, b->stream));
    GJD_TRY(cudaMemcpyAsync(b->d_meta.p, b->h_meta.p, meta, cudaMemcpyHostToDevice, b->stream));
    GJD_TRY(gjd_launch_build_luts((const GjdDhtSet*)((uint8_t*)b->d_meta.p + o_dht), (GjdLutSet*)b->d_luts.p, (GjdLutCount*)b->d_lutc.p,
                                  (uint32_t)b->dhts.size(), b->stream));
    GJD_TRY(gjd_launch_fill_maps(v, b->stream));
    GJD_TRY(gjd_launch_unstuff(v, b->stream, &b->unstuff_launches));
    return GJD_OK;
}

int batch_enqueue_run(gjd_ctx* ctx, gjd_batch* b) {
    if (b->view.n_img == 0) { b->launches = 0; b->ran = true; return GJD_OK; }
    GJD_TRY(gjd_launch_pipeline(b->view, b->stream, &b->ev, ctx->transform_variant, &b->launches, false, &ctx->kernel_attrs_set));
    GJD_TRY(cudaMemcpyAsync(b->h_status.p, b->d_status.p, (size_t)(b->view.n_img + 1) * sizeof(uint32_t),
                            cudaMemcpyDeviceToHost, b->stream));
    b->ran = true;
    ctx->last = b;
    return GJD_OK;
}

/* Wait for the batch.  If a sequence boundary did not converge in the single fix-up round of the
 * pipeline (a chain crossing a whole 256-subsequence sequence without merging: never seen on real
 * streams, kept for correctness), iterate the fix-up to its fixed point and redo the later stages. */
int batch_wait(gjd_ctx* ctx, gjd_batch* b) {
    if (!b->stream) return GJD_OK;
    NvtxRange r("gjd: wait for the sub-batch");
    GJD_TRY(cudaStreamSynchronize(b->stream));
    if (!b->ran || b->view.n_img == 0) return GJD_OK;
    uint32_t* hst = (uint32_t*)b->h_status.p;
    if (hst[b->view.n_img] == 0) return GJD_OK;
    for (int guard = 0; guard < 65536 && hst[b->view.n_img] != 0; guard++) {
        GJD_TRY(cudaMemsetAsync(b->view.nosync, 0, sizeof(uint32_t), b->stream));
        GJD_TRY(gjd_launch_fix(b->view, b->stream));
        GJD_TRY(cudaMemcpyAsync(hst + b->view.n_img, b->view.nosync, sizeof(uint32_t), cudaMemcpyDeviceToHost,
                                b->stream));
        GJD_TRY(cudaStreamSynchronize(b->stream));
    }
    int launches = 0;
    GJD_TRY(gjd_launch_pipeline(b->view, b->stream, nullptr, ctx->transform_variant, &launches, true, &ctx->kernel_attrs_set));
    GJD_TRY(cudaMemcpyAsync(b->h_status.p, b->d_status.p, (size_t)(b->view.n_img + 1) * sizeof(uint32_t),
                            cudaMemcpyDeviceToHost, b->stream));
    GJD_TRY(cudaStreamSynchronize(b->stream));
    return 1;   /* outputs were regenerated: the caller repeats its D2H copies */
}

void set_item_status(gjd_batch* b, gjd_item* items) {
    const uint32_t* hst = (const uint32_t*)b->h_status.p;
    for (int32_t i = 0; i < b->n; i++) {
        int32_t st = b->geom[i].status;
        if (st == GJD_OK && b->img_of[i] >= 0 && b->ran && hst[b->img_of[i]] != 0) {
            /* what the device unstuffing found comes first, with the codes the host loop gives (gjd_unstuff_segments) */
            const uint32_t bits = hst[b->img_of[i]];
            st = (bits & GJD_DEV_UFORMAT) ? GJD_E_FORMAT : (bits & GJD_DEV_UUNSUP) ? GJD_E_UNSUPPORTED : GJD_E_CORRUPT;
        }
        b->geom[i].status = st;
        if (items) {
            items[i].status = st;
            items[i].width = b->geom[i].width;
            items[i].height = b->geom[i].height;
        }
    }
}

/* Header parse of every item (parallel). */
void parse_all(gjd_ctx* ctx, const gjd_item* items, int32_t n, std::vector<GjdParsed>& parsed,
               std::vector<ItemGeom>& geom) {
    NvtxRange r("gjd: parse headers");
    parsed.resize(n);
    geom.assign(n, ItemGeom());
    ctx->pool->parallel_for((size_t)n, 16, [&](size_t i) {
        ItemGeom& g = geom[i];
        if (!items[i].jpg) { g.status = GJD_E_ARG; return; }
        int rc = gjd_parse_headers(items[i].jpg, items[i].len, &parsed[i]);
        if (rc == GJD_OK && ctx->device_unstuff) {
            cudaPointerAttributes at;
            if (cudaPointerGetAttributes(&at, items[i].jpg) == cudaSuccess) g.jpg_pinned = at.type == cudaMemoryTypeHost;
            else cudaGetLastError();
        }
        g.width = parsed[i].width;
        g.height = parsed[i].height;
        if (rc) { g.status = rc; return; }
        g.layout = items[i].layout;
        if (g.layout == GJD_LAYOUT_INTERLEAVED || g.layout == GJD_LAYOUT_INTERLEAVED_I16) {
            const size_t bpp = g.layout == GJD_LAYOUT_INTERLEAVED ? 3 : 6;        /* bytes per pixel */
            g.pitch = items[i].row_pitch ? items[i].row_pitch : bpp * (size_t)g.width;
            g.pstride = 0;
            if (g.pitch < bpp * (size_t)g.width || g.pitch > 0xffffffffull) { g.status = GJD_E_ARG; return; }
            if (bpp == 6 && ((g.pitch & 1u) || ((uintptr_t)items[i].rgb & 1u))) { g.status = GJD_E_ARG; return; }   /* int16 elements */
            g.out_bytes = g.pitch * (size_t)g.height;
            return;
        }
        if (g.layout != GJD_LAYOUT_PLANAR) { g.status = GJD_E_ARG; return; }
        g.pitch = items[i].row_pitch ? items[i].row_pitch : (size_t)g.width;
        g.pstride = items[i].plane_stride ? items[i].plane_stride : g.pitch * (size_t)g.height;
        if (g.pitch < (size_t)g.width || g.pstride < g.pitch * (size_t)g.height || g.pitch > 0xffffffffull) {
            g.status = GJD_E_ARG;
            return;
        }
        g.out_bytes = 2 * g.pstride + g.pitch * (size_t)g.height;
    });
}

}  // namespace

/* ---------------------------------------------------------------- C ABI --------------- */

extern "C" {

const char* gjd_version(void) {
    return "gjd " GJD_VERSION_STR " sm_100a subseq=2^" GJD_STR(GJD_SUB_LOG2_MIN) "..2^" GJD_STR(GJD_SUB_LOG2_MAX) " lut=" GJD_STR(GJD_LUT_BITS);
}

const char* gjd_strerror(int st) {
    switch (st) {
        case GJD_OK: return "ok";
        case GJD_E_ARG: return "invalid argument";
        case GJD_E_CUDA: return "CUDA runtime error";
        case GJD_E_NOMEM: return "out of memory";
        case GJD_E_FORMAT: return "not a JPEG or malformed/truncated marker segment";
        case GJD_E_UNSUPPORTED: return "JPEG outside the 8-bit baseline single-scan contract (4:4:4, 4:2:2, 4:4:0, 4:2:0 or grayscale)";
        case GJD_E_CORRUPT: return "corrupt entropy-coded data";
        case GJD_E_NODEVICE: return "no usable CUDA device (libgjd has no CPU decode path)";
        case GJD_E_IO: return "file I/O error";
        default: return "unknown status";
    }
}

const char* gjd_last_cuda_error(gjd_ctx* ctx) { return ctx ? ctx->cuda_err.c_str() : ""; }

int gjd_create(gjd_ctx** out, int device, size_t scratch_bytes_hint) {
    if (!out) return GJD_E_ARG;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) return GJD_E_NODEVICE;
    if (device < 0 || device >= count) return GJD_E_ARG;
    DeviceGuard guard(device);
    if (!guard.ok) return GJD_E_NODEVICE;
    GJD_API_TRY
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return GJD_E_NODEVICE;
    if (prop.major < 10) return GJD_E_NODEVICE;   /* kernels are built for sm_100a only */
    gjd_ctx* ctx = new (std::nothrow) gjd_ctx();
    if (!ctx) return GJD_E_NOMEM;
    ctx->device = device;
    ctx->scratch = scratch_bytes_hint ? scratch_bytes_hint : ((size_t)6 << 30);
    int threads = (int)std::thread::hardware_concurrency();
    if (const char* e = getenv("GJD_HOST_THREADS")) threads = atoi(e);
    threads = std::max(1, std::min(threads, 64));
    ctx->pool = new Pool(threads - 1);
    if (const char* e = getenv("GJD_TRANSFORM_VARIANT")) ctx->transform_variant = atoi(e);
    if (const char* e = getenv("GJD_SUB_LOG2")) {   /* 8..16: beyond the 10..13 the library picks from, for the tests */
        int k = atoi(e);
        if (k >= 8 && k <= 16) ctx->force_sub_log2 = k;
    }
    if (const char* e = getenv("GJD_TB_TILES_MAX")) {
        const int v = atoi(e);
        if (v >= 4 && v <= 128 && (v & (v - 1)) == 0) ctx->max_tb_tiles = (uint32_t)v;
    }
    if (const char* e = getenv("GJD_DEVICE_UNSTUFF")) ctx->device_unstuff = atoi(e) != 0;
    if (const char* e = getenv("GJD_MIN_SUBSEQS")) ctx->min_subseqs = (uint64_t)std::max(1ll, atoll(e));
    {   /* the arenas are stream-ordered allocations: keep what they free inside the process */
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
            uint64_t keep = ~0ull;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
    }
    *out = ctx;
    return GJD_OK;
    GJD_API_CATCH
}

void gjd_destroy(gjd_ctx* ctx) {
    if (!ctx) return;
    DeviceGuard guard(ctx->device);
    for (auto& s : ctx->slot) {
        if (s.stream) cudaStreamSynchronize(s.stream);
        s.release();
    }
    delete ctx->pool;
    delete ctx;
}

int gjd_peek(const uint8_t* jpg, size_t len, int32_t* width, int32_t* height) {
    if (!jpg) return GJD_E_ARG;
    GJD_API_TRY
    GjdParsed p;
    int rc = gjd_parse_headers(jpg, len, &p);
    if (width) *width = p.width;
    if (height) *height = p.height;
    return rc;
    GJD_API_CATCH
}

void* gjd_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) return nullptr;
    return p;
}
void gjd_host_free(void* p) { if (p) cudaFreeHost(p); }

/* D2H of every image of a staged+run batch into the callers' buffers; adjacent regions merge. */
static int enqueue_d2h(gjd_ctx* ctx, gjd_batch* b, gjd_item* items) {
    NvtxRange r("gjd: D2H pixels");
    const uint8_t* dbase = (const uint8_t*)b->d_rgb.p;
    uint8_t* run_dst = nullptr;
    const uint8_t* run_src = nullptr;
    size_t run_len = 0;
    for (int32_t i = 0; i < b->n; i++) {
        int32_t k = b->img_of[i];
        if (k < 0) continue;
        const GjdImage& d = b->imgs[k];
        if (d.out_is_ptr || !items[i].rgb) continue;
        const uint8_t* src = dbase + d.out_off;
        size_t len = b->geom[i].out_bytes;
        if (run_len && run_dst + run_len == items[i].rgb && run_src + run_len == src) {
            run_len += len;
            continue;
        }
        if (run_len) GJD_TRY(cudaMemcpyAsync(run_dst, run_src, run_len, cudaMemcpyDeviceToHost, b->stream));
        run_dst = items[i].rgb; run_src = src; run_len = len;
    }
    if (run_len) GJD_TRY(cudaMemcpyAsync(run_dst, run_src, run_len, cudaMemcpyDeviceToHost, b->stream));
    return GJD_OK;
}

int gjd_decode_batch(gjd_ctx* ctx, gjd_item* items, int32_t n) {
    if (!ctx || (n > 0 && !items) || n < 0) return GJD_E_ARG;
    DeviceGuard guard(ctx->device);
    if (!guard.ok) return GJD_E_NODEVICE;
    if (n == 0) return GJD_OK;
    GJD_API_TRY
    NvtxRange whole("gjd_decode_batch");
    std::vector<GjdParsed> parsed;
    std::vector<ItemGeom> geom;
    parse_all(ctx, items, n, parsed, geom);

    /* sub-batches under the scratch budget, rotating through the GJD_SLOTS slots.  The call is bound by the
     * device-to-host link, which idles until the first sub-batch has been unstuffed, uploaded and decoded:
     * the first sub-batch gets 1/16 of the budget and the following ones double up to the full budget. */
    struct Span { int32_t first, count; };
    std::vector<Span> spans;
    {
        size_t acc = 0, cap = std::max<size_t>(ctx->scratch >> 4, (size_t)64 << 20);
        int32_t first = 0;
        for (int32_t i = 0; i < n; i++) {
            size_t need = geom[i].status == GJD_OK ? estimate_bytes(items[i], geom[i]) : 0;
            if (i > first && acc + need > std::min(cap, ctx->scratch)) {
                spans.push_back({first, i - first});
                first = i;
                acc = 0;
                cap *= 2;
            }
            acc += need;
        }
        spans.push_back({first, n - first});
    }
    int rc = GJD_OK;
    int32_t pending[GJD_SLOTS];      /* span index in flight on each slot */
    for (int s = 0; s < GJD_SLOTS; s++) pending[s] = -1;
    auto finish = [&](int s) -> int {
        if (pending[s] < 0) return GJD_OK;
        gjd_batch* b = &ctx->slot[s];
        const Span& sp = spans[pending[s]];
        int w = batch_wait(ctx, b);
        if (w < 0) return w;
        if (w == 1) {   /* outputs regenerated on the slow path */
            int r2 = enqueue_d2h(ctx, b, items + sp.first);
            if (r2) return r2;
            if (cudaStreamSynchronize(b->stream) != cudaSuccess) return GJD_E_CUDA;
        }
        set_item_status(b, items + sp.first);
        pending[s] = -1;
        return GJD_OK;
    };
    for (size_t k = 0; k < spans.size() && rc == GJD_OK; k++) {
        int s = (int)(k % GJD_SLOTS);
        rc = finish(s);
        if (rc) break;
        gjd_batch* b = &ctx->slot[s];
        const Span& sp = spans[k];
        rc = batch_stage(ctx, b, items + sp.first, parsed.data() + sp.first, geom.data() + sp.first, sp.count);
        if (rc) break;
        rc = batch_enqueue_run(ctx, b);
        if (rc) break;
        rc = enqueue_d2h(ctx, b, items + sp.first);
        pending[s] = (int32_t)k;
    }
    for (int s = 0; s < GJD_SLOTS; s++) {
        int r2 = finish(s);
        if (rc == GJD_OK) rc = r2;
    }
    if (rc != GJD_OK)
        for (int32_t i = 0; i < n; i++)
            if (items[i].status == GJD_OK && geom[i].status != GJD_OK) items[i].status = geom[i].status;
    return rc;
    GJD_API_CATCH
}

int gjd_batch_prepare(gjd_ctx* ctx, gjd_item* items, int32_t n, gjd_batch** out) {
    if (!ctx || !out || (n > 0 && !items) || n < 0) return GJD_E_ARG;
    *out = nullptr;
    DeviceGuard guard(ctx->device);
    if (!guard.ok) return GJD_E_NODEVICE;
    gjd_batch* b = new (std::nothrow) gjd_batch();
    if (!b) return GJD_E_NOMEM;
    try {
    NvtxRange whole("gjd_batch_prepare");
    std::vector<ItemGeom> geom;
    parse_all(ctx, items, n, b->parsed, geom);
    /* items[i].rgb == NULL (or a host pointer): the pixels land in the batch's own device arena, read with gjd_batch_fetch */
    int rc = batch_stage(ctx, b, items, b->parsed.data(), geom.data(), n);
    if (rc == GJD_OK && cudaStreamSynchronize(b->stream) != cudaSuccess) rc = cuda_fail(ctx, cudaGetLastError(), "prepare");
    if (rc == GJD_OK && b->view.n_raw) {   /* what the device unstuffing found is known already: report it from prepare */
        cudaError_t e = cudaMemcpyAsync(b->h_status.p, b->view.ustatus, (size_t)b->view.n_img * sizeof(uint32_t), cudaMemcpyDeviceToHost, b->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(b->stream);
        if (e != cudaSuccess) rc = cuda_fail(ctx, e, "prepare");
        const uint32_t* hst = (const uint32_t*)b->h_status.p;
        for (int32_t i = 0; rc == GJD_OK && i < n; i++)
            if (b->geom[i].status == GJD_OK && b->img_of[i] >= 0 && hst[b->img_of[i]])
                b->geom[i].status = (hst[b->img_of[i]] & GJD_DEV_UFORMAT) ? GJD_E_FORMAT
                                    : (hst[b->img_of[i]] & GJD_DEV_UUNSUP) ? GJD_E_UNSUPPORTED : GJD_E_CORRUPT;
    }
    if (rc != GJD_OK) {
        b->release();
        delete b;
        return rc;
    }
    set_item_status(b, items);
    b->parsed.clear();
    b->parsed.shrink_to_fit();
    *out = b;
    return GJD_OK;
    } catch (...) {
        b->release();
        delete b;
        return GJD_E_NOMEM;
    }
}

int gjd_batch_run(gjd_ctx* ctx, gjd_batch* b) {
    if (!ctx || !b) return GJD_E_ARG;
    DeviceGuard guard(ctx->device);
    if (!guard.ok) return GJD_E_NODEVICE;
    return batch_enqueue_run(ctx, b);
}

int gjd_batch_sync(gjd_ctx* ctx, gjd_batch* b) {
    if (!ctx || !b) return GJD_E_ARG;
    DeviceGuard guard(ctx->device);
    if (!guard.ok) return GJD_E_NODEVICE;
    int w = batch_wait(ctx, b);
    if (w < 0) return w;
    set_item_status(b, nullptr);
    for (int32_t i = 0; i < b->n; i++)
        if (b->geom[i].status == GJD_E_CORRUPT) return GJD_E_CORRUPT;
    return GJD_OK;
}

int gjd_batch_status(gjd_ctx* ctx, gjd_batch* b, int32_t i) {
    if (!ctx || !b || i < 0 || i >= b->n) return GJD_E_ARG;
    return b->geom[i].status;
}

int gjd_batch_fetch(gjd_ctx* ctx, gjd_batch* b, int32_t i, uint8_t* rgb_host) {
    if (!ctx || !b || !rgb_host || i < 0 || i >= b->n) return GJD_E_ARG;
    DeviceGuard guard(ctx->device);
    if (!guard.ok) return GJD_E_NODEVICE;
    int32_t k = b->img_of[i];
    if (k < 0) return b->geom[i].status ? b->geom[i].status : GJD_E_ARG;
    const GjdImage& d = b->imgs[k];
    const uint8_t* src = d.out_is_ptr ? (const uint8_t*)(uintptr_t)d.out_off : (const uint8_t*)b->d_rgb.p + d.out_off;
    if (d.layout != GJD_LAYOUT_PLANAR) {
        const size_t row = (d.layout == GJD_LAYOUT_INTERLEAVED ? 3 : 6) * (size_t)d.width;
        GJD_TRY(cudaMemcpy2DAsync(rgb_host, row, src, d.row_pitch, row, d.height, cudaMemcpyDeviceToHost, b->stream));
    } else {
        for (int c = 0; c < 3; c++)
            GJD_TRY(cudaMemcpy2DAsync(rgb_host + (size_t)c * d.width * d.height, d.width, src + (size_t)c * d.plane_stride,
                                      d.row_pitch, d.width, d.height, cudaMemcpyDeviceToHost, b->stream));
    }
    GJD_TRY(cudaStreamSynchronize(b->stream));
    return GJD_OK;
}

void gjd_batch_free(gjd_ctx* ctx, gjd_batch* b) {
    if (!b) return;
    DeviceGuard guard(ctx ? ctx->device : 0);
    if (ctx && ctx->last == b) ctx->last = nullptr;
    if (b->stream) cudaStreamSynchronize(b->stream);
    b->release();
    delete b;
}

int gjd_batch_info(gjd_ctx* ctx, gjd_batch* b, uint64_t info[5]) {
    if (!ctx || !b || !info) return GJD_E_ARG;
    info[0] = b->pixels; info[1] = b->padded_pixels; info[2] = b->scan_bytes; info[3] = b->subseqs;
    info[4] = (uint64_t)b->launches;
    return GJD_OK;
}

int gjd_stage_times(gjd_ctx* ctx, float ms[4]) {
    if (!ctx || !ms) return GJD_E_ARG;
    gjd_batch* b = ctx->last;
    if (!b || !b->ev_ok || !b->ran || b->view.n_img == 0) return GJD_E_ARG;
    DeviceGuard guard(ctx->device);
    if (!guard.ok) return GJD_E_NODEVICE;
    GJD_TRY(cudaEventSynchronize(b->ev.e[4]));
    GJD_TRY(cudaEventElapsedTime(&ms[0], b->ev.e[1], b->ev.e[2]));
    GJD_TRY(cudaEventElapsedTime(&ms[1], b->ev.e[2], b->ev.e[3]));
    GJD_TRY(cudaEventElapsedTime(&ms[2], b->ev.e[3], b->ev.e[4]));
    GJD_TRY(cudaEventElapsedTime(&ms[3], b->ev.e[0], b->ev.e[4]));
    return GJD_OK;
}

int gjd_decode_coeffs(gjd_ctx* ctx, const uint8_t* jpg, size_t len, int16_t* coeffs_host) {
    if (!ctx || !jpg || !coeffs_host) return GJD_E_ARG;
    DeviceGuard guard(ctx->device);
    if (!guard.ok) return GJD_E_NODEVICE;
    gjd_item it;
    memset(&it, 0, sizeof(it));
    it.jpg = jpg;
    it.len = len;
    gjd_batch* b = nullptr;
    int rc = gjd_batch_prepare(ctx, &it, 1, &b);
    if (rc) return rc;
    if (it.status != GJD_OK) { rc = it.status; gjd_batch_free(ctx, b); return rc; }
    rc = gjd_batch_run(ctx, b);
    if (rc == GJD_OK) rc = gjd_batch_sync(ctx, b);
    if (rc == GJD_OK || rc == GJD_E_CORRUPT) {
        int rc0 = rc;
        size_t ncoef = (size_t)b->imgs[0].mcu_count * (b->imgs[0].gray ? 64 : ((size_t)b->imgs[0].hs * b->imgs[0].vs + 2) * 64);
        DevBuf tmp;
        cudaError_t e = tmp.ensure(ncoef * sizeof(int16_t), b->stream);
        if (e == cudaSuccess) e = gjd_launch_export_coefs(b->view, (int16_t*)tmp.p, b->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(coeffs_host, tmp.p, ncoef * sizeof(int16_t), cudaMemcpyDeviceToHost, b->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(b->stream);
        tmp.release(b->stream);
        rc = e == cudaSuccess ? rc0 : cuda_fail(ctx, e, "export_coefs");
    }
    gjd_batch_free(ctx, b);
    return rc;
}

int gjd_test_color(gjd_ctx* ctx, const int16_t* y, const int16_t* cb, const int16_t* cr, size_t n,
                   uint8_t* r, uint8_t* g, uint8_t* bl) {
    if (!ctx || !y || !cb || !cr || !r || !g || !bl) return GJD_E_ARG;
    DeviceGuard guard(ctx->device);
    if (!guard.ok) return GJD_E_NODEVICE;
    if (n == 0) return GJD_OK;
    DevBuf in, out;
    GJD_TRY(in.ensure(n * 6, 0));
    GJD_TRY(out.ensure(n * 3, 0));
    int16_t* di = (int16_t*)in.p;
    uint8_t* dout = (uint8_t*)out.p;
    int rc = GJD_OK;
    cudaError_t e = cudaMemcpy(di, y, n * 2, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(di + n, cb, n * 2, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(di + 2 * n, cr, n * 2, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = gjd_launch_test_color(di, di + n, di + 2 * n, n, dout, dout + n, dout + 2 * n, 0);
    if (e == cudaSuccess) e = cudaMemcpy(r, dout, n, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(g, dout + n, n, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(bl, dout + 2 * n, n, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) rc = cuda_fail(ctx, e, "gjd_test_color");
    in.release(0);
    out.release(0);
    return rc;
}

int gjd_test_idct(gjd_ctx* ctx, const int16_t* coef_zz, size_t nblocks, const uint8_t quant_zz[64], int16_t* samples) {
    if (!ctx || !coef_zz || !quant_zz || !samples) return GJD_E_ARG;
    DeviceGuard guard(ctx->device);
    if (!guard.ok) return GJD_E_NODEVICE;
    if (nblocks == 0) return GJD_OK;
    DevBuf in, out, q;
    GJD_TRY(in.ensure(nblocks * 128, 0));
    GJD_TRY(out.ensure(nblocks * 128, 0));
    GJD_TRY(q.ensure(256, 0));
    int32_t hq[64];
    for (int i = 0; i < 64; i++) hq[i] = quant_zz[i];
    int rc = GJD_OK;
    cudaError_t e = cudaMemcpy(in.p, coef_zz, nblocks * 128, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(q.p, hq, 256, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = gjd_launch_test_idct((const int16_t*)in.p, nblocks, (const int32_t*)q.p, (int16_t*)out.p, 0);
    if (e == cudaSuccess) e = cudaMemcpy(samples, out.p, nblocks * 128, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) rc = cuda_fail(ctx, e, "gjd_test_idct");
    in.release(0); out.release(0); q.release(0);
    return rc;
}

}  /* extern "C" */
