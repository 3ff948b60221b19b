/*
 * gjd_kernels.cu — the sm_100a kernels of the decode path.
 *
 *   k_build_luts    DHT (BITS/HUFFVAL) -> flattened lookup tables, one block per table set
 *   k_fill_maps     sequence -> image and tile -> image maps
 *   k_huff_sync     look-back decode of every subsequence + entry/exit state validation inside a sequence
 *   k_huff_fix      boundary fix-up between sequences
 *   k_seq_scan      exclusive scan of per-sequence coefficient counts (per image)
 *   k_huff_write    re-decode from the synchronised states; every thread stages the blocks it owns in
 *                   shared memory and writes them as whole 32-byte sectors (+ DC difference, kmax)
 *   k_dc_tilesum    DC differences summed per 32-MCU tile
 *   k_dc_tilebase   DC predictor at the start of every tile (scan of the tile sums per image)
 *   k_transform     per 8x8 block: load the coefficient sectors, DC prefix, dequantise, dezigzag, integer
 *                   IDCT in registers; samples meet in shared memory for colour conversion + RGB store
 *
 * Stage structure after the reference's decodeImage / parallelHuffManDecode
 * (cuda-decoder/src/parser.cu:613-661, 132-208); none of its code is reused.
 */
#include "gjd_kernels.h"

#include <nvtx3/nvToolsExt.h>

#include "gjd_devfn.cuh"
#include "../../include/gjd.h"

#define GJD_SYNC_THREADS 256
static_assert(GJD_SYNC_THREADS == GJD_SEQ_SUBS, "one thread per subsequence of a sequence");

/* ------------------------------------------------------------------ small helpers ----- */

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) >= o) v += t;
    }
    return v;
}

/* Exclusive scan over a 256-thread block; *total gets the block sum. ws: 8 words of shared memory. */
__device__ __forceinline__ uint32_t block_excl_scan_256(uint32_t v, uint32_t* ws, uint32_t* total) {
    uint32_t incl = warp_incl_scan(v);
    uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    __syncthreads();   /* ws may still be read from a previous call */
    if (lane == 31) ws[wid] = incl;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        uint32_t t = ws[w];
        if (w < (int)wid) base += t;
        tot += t;
    }
    *total = tot;
    return base + incl - v;
}

__device__ __forceinline__ int warp_incl_scan_i(int v, uint32_t lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, o);
        if ((int)lane >= o) v += t;
    }
    return v;
}

/* The count pass's view: the pair LUT ring in place of the primary one, everything else as in the set. */
__device__ __forceinline__ void copy_count_lut_to_smem(GjdLutSet* dst, const GjdLutSet* src, const GjdLutCount* cnt) {
    static_assert(sizeof(GjdLutCount) == GJD_RING_BYTES && sizeof(GjdLutCount) % 16 == 0, "ring size");
    const uint4* c = reinterpret_cast<const uint4*>(cnt);
    const uint4* s = reinterpret_cast<const uint4*>(src);
    uint4* d = reinterpret_cast<uint4*>(dst);
    for (uint32_t i = threadIdx.x; i < sizeof(GjdLutSet) / 16; i += blockDim.x)
        d[i] = __ldg(i < sizeof(GjdLutCount) / 16 ? c + i : s + i);
}

/* the first `bytes` of the set: all of it, or only the two LUT levels (the canonical tables then stay in global memory) */
__device__ __forceinline__ void copy_lut_to_smem(GjdLutSet* dst, const GjdLutSet* src, uint32_t bytes = sizeof(GjdLutSet)) {
    static_assert(sizeof(GjdLutSet) % 16 == 0, "GjdLutSet must be a multiple of 16 bytes");
    const uint4* s = reinterpret_cast<const uint4*>(src);
    uint4* d = reinterpret_cast<uint4*>(dst);
    for (uint32_t i = threadIdx.x; i < bytes / 16; i += blockDim.x) d[i] = __ldg(s + i);
}

#ifdef GJD_SCAN_TMA
/* A/B build: the entropy kernels read the scan through per-thread TMA rings (GjdScanRing, gjd_devfn.cuh) */
#define GJD_RING_BYTES_PER_BLOCK (2u * GJD_SEQ_SUBS * GJD_RING_CHUNK)
__device__ __forceinline__ GjdScanRing gjd_ring_setup(uint8_t* ring, uint64_t* bars, const uint8_t* arena, uint64_t seg_off) {
    GjdScanRing r;
    r.arena = arena;
    r.base = seg_off;
    r.slots = (uint32_t)__cvta_generic_to_shared(ring + threadIdx.x * GJD_RING_CHUNK);
    r.stride = GJD_SEQ_SUBS * GJD_RING_CHUNK;
    r.bars = (uint32_t)__cvta_generic_to_shared(bars + 2u * threadIdx.x);
    r.cur = 0xffffffffu;
    r.phase = 0u;
    r.inflight = 0u;
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(r.bars));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(r.bars + 8u));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    return r;
}
#endif

/* ------------------------------------------------------------------ tables and maps --- */

__global__ void __launch_bounds__(256) k_build_luts(const GjdDhtSet* __restrict__ dht, GjdLutSet* __restrict__ luts,
                                                    GjdLutCount* __restrict__ lutc) {
    const GjdDhtSet& in = dht[blockIdx.x];
    GjdLutSet& out = luts[blockIdx.x];
    __shared__ uint32_t lim[4][17];
    __shared__ int32_t off[4][17];
    if (threadIdx.x < 4) {
        gjd_canonical(in.bits[threadIdx.x], lim[threadIdx.x], off[threadIdx.x]);
        for (int L = 0; L <= 16; L++) {
            out.lim[threadIdx.x][L] = lim[threadIdx.x][L];
            out.off[threadIdx.x][L] = off[threadIdx.x][L];
        }
        out.isdc[threadIdx.x] = in.isdc[threadIdx.x];
    }
    if (threadIdx.x < GJD_RING_TABLES) out.slot_of[threadIdx.x] = (uint8_t)gjd_ring_slot(in.tabsel, threadIdx.x);
    for (uint32_t i = threadIdx.x; i < 4 * 256; i += blockDim.x) out.vals[i >> 8][i & 255] = in.vals[i >> 8][i & 255];
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < GJD_RING_TABLES * GJD_LUT_SIZE; i += blockDim.x) {
        uint32_t r = i / GJD_LUT_SIZE, j = i % GJD_LUT_SIZE;
        uint32_t t = gjd_ring_slot(in.tabsel, r);
        out.lut[r][j] = gjd_lut_entry(lim[t], off[t], in.vals[t], in.isdc[t], j);
        lutc[blockIdx.x].lut[r][j] = gjd_lutc_entry(lim[t], off[t], in.vals[t], in.isdc[t], j);
    }
    for (uint32_t i = threadIdx.x; i < GJD_RING_TABLES * GJD_LUT2_SIZE; i += blockDim.x) {
        uint32_t r = i / GJD_LUT2_SIZE, j = i % GJD_LUT2_SIZE;
        uint32_t t = gjd_ring_slot(in.tabsel, r);
        out.lut2[r][j] = gjd_lut2_entry(lim[t], off[t], in.vals[t], in.isdc[t], j);
    }
}

__global__ void __launch_bounds__(128) k_fill_maps(GjdDevBatch b) {
    /* blocks [0, n_img): tile and transform-block maps; blocks [n_img, n_img + n_seg): sequence -> segment */
    if (blockIdx.x < b.n_img) {
        uint32_t img = blockIdx.x;
        const GjdImage& d = b.imgs[img];
        for (uint32_t i = threadIdx.x; i < d.tile_count; i += blockDim.x) b.tile_img[d.tile_first + i] = img;
        for (uint32_t i = threadIdx.x; i < d.tb_count; i += blockDim.x) b.tb_img[d.tb_first + i] = img;
    } else if (blockIdx.x < b.n_img + b.n_seg) {
        uint32_t sg = blockIdx.x - b.n_img;
        const GjdSegment& d = b.segs[sg];
        for (uint32_t i = threadIdx.x; i < d.seq_count; i += blockDim.x) b.seq_seg[d.seq_first + i] = sg;
    } else {
        uint32_t r = blockIdx.x - b.n_img - b.n_seg;
        const GjdRaw& d = b.raws[r];
        for (uint32_t i = threadIdx.x; i < d.chunk_count; i += blockDim.x) b.uchunk_raw[d.chunk_first + i] = r;
    }
}

/* ------------------------------------------------------------------ device unstuffing -- */
/*
 * FF00 unstuffing with marker detection on the device, replacing the host's byte loop
 * (cpp-decoder/src/parser.cpp:78-99, cuda-decoder/src/parser.cu:446-464; host twin gjd_unstuff_segments) for scans
 * without restart intervals.  The host uploads the raw bytes from the first scan byte to the end of the file.
 *   k_unstuff_count  one block per 4 KB chunk, 16 bytes per thread: bytes kept per chunk; position of every EOI and of
 *                    every marker that does not belong in the scan (atomicMin per scan)
 *   k_unstuff_scan   one block per scan: the first EOI ends the scan (everything after it is ignored, including
 *                    markers); an offending marker before it, or no EOI, fails the image with the status the host loop
 *                    gives; exclusive scan of the chunk counts; unstuffed length -> scan_bits / sub_count of the segment
 *                    (the host sized the sequences from the raw length, an upper bound); 16 zero bytes behind the scan
 *   k_unstuff_write  one block per chunk: the kept bytes, compacted in shared memory, go to the scan arena as whole
 *                    32-bit words (byte stores only at the chunk's unaligned ends)
 */
/* limit: positions >= limit are outside the scan (raw_len, or the EOI position in the write pass) */
__device__ __forceinline__ GjdUnstuffLane gjd_unstuff_lane(const GjdDevBatch& b, const GjdRaw& r, uint32_t chunk, uint32_t limit) {
    const uint64_t a = (r.raw_off & ~(uint64_t)15) + (uint64_t)chunk * GJD_UCHUNK + threadIdx.x * 16u;   /* arena offset of byte 0 */
    return gjd_unstuff_lane_at(b.raw, r.raw_off, r.raw_len, a, limit, r.dri != 0u);
}

__global__ void __launch_bounds__(256) k_unstuff_count(GjdDevBatch b) {
    __shared__ uint32_t ws[8];
    const uint32_t ri = b.uchunk_raw[blockIdx.x];
    const GjdRaw& r = b.raws[ri];
    const uint32_t chunk = blockIdx.x - r.chunk_first;
    const GjdUnstuffLane L = gjd_unstuff_lane(b, r, chunk, r.raw_len);
    uint32_t mk = L.m.mark;
    while (mk) {                                   /* rare: a marker inside the scan */
        const uint32_t i = __ffs(mk) - 1u;
        mk &= mk - 1u;
        const uint32_t nb = i < 15u ? gjd_lane_byte(L, i + 1u) : L.next;
        const uint32_t pos = L.pos0 + i;
        if (nb == 0xd9u) atomicMin(b.raw_eoi + ri, pos);
        else atomicMin(b.raw_err + ri, gjd_unstuff_err_word(pos, nb));
    }
    /* bytes kept (<= 4096 per chunk) and restart markers (<= 2048) in one sum: kept | markers << 16 */
    uint32_t kept = __popc(L.valid & ~L.m.drop) | ((uint32_t)__popc(L.m.rst) << 16);
    kept = __reduce_add_sync(0xffffffffu, kept);
    if ((threadIdx.x & 31u) == 0u) ws[threadIdx.x >> 5] = kept;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) t += ws[w];
        b.uchunk_cnt[blockIdx.x] = t & 0xffffu;
        b.uchunk_rst[blockIdx.x] = t >> 16;
    }
}

__global__ void __launch_bounds__(256) k_unstuff_scan(GjdDevBatch b, uint32_t* __restrict__ ustatus) {
    __shared__ uint32_t ws[8];
    const uint32_t ri = blockIdx.x, tid = threadIdx.x;
    const GjdRaw& r = b.raws[ri];
    GjdSegment& sg = b.segs[r.seg];
    uint32_t eoi = b.raw_eoi[ri];
    const uint32_t err = b.raw_err[ri];
    __syncthreads();                                /* every thread has read raw_eoi before it is rewritten below */
    const bool failed = eoi == 0xffffffffu || (err != 0xffffffffu && (err >> 2) < eoi);
    if (failed) {
        if (tid == 0) {
            const bool unsup = err != 0xffffffffu && (eoi == 0xffffffffu || (err >> 2) < eoi) && (err & 3u) == GJD_UERR_UNSUPPORTED;
            ustatus[r.img] = unsup ? GJD_DEV_UUNSUP : GJD_DEV_UFORMAT;
            sg.scan_bits = 0;
            sg.sub_count = 1;
            b.raw_eoi[ri] = 0;                      /* k_unstuff_write: nothing to write */
        }
        if (tid < 16) b.scan[sg.scan_off + tid] = 0;
        return;
    }
    /* exclusive scan of the chunk counts up to the chunk that holds the EOI */
    const uint32_t ce = (uint32_t)(((r.raw_off & 15u) + eoi) / GJD_UCHUNK);
    uint32_t* cnt = b.uchunk_cnt + r.chunk_first;
    uint32_t* rcnt = b.uchunk_rst + r.chunk_first;
    uint32_t carry = 0, rcarry = 0;
    for (uint32_t base = 0; base < ce; base += 256u) {
        const uint32_t c = base + tid;
        const uint32_t v = c < ce ? cnt[c] : 0u, rv = c < ce ? rcnt[c] : 0u;
        uint32_t blocksum, rsum;
        const uint32_t excl = block_excl_scan_256(v, ws, &blocksum);
        const uint32_t rexcl = block_excl_scan_256(rv, ws, &rsum);
        if (c < ce) { cnt[c] = carry + excl; rcnt[c] = rcarry + rexcl; }
        carry += blocksum;
        rcarry += rsum;
    }
    /* the chunk with the EOI keeps only what lies before it */
    const GjdUnstuffLane L = gjd_unstuff_lane(b, r, ce, eoi);
    uint32_t blocksum;
    block_excl_scan_256(__popc(L.valid & ~L.m.drop) | ((uint32_t)__popc(L.m.rst) << 16), ws, &blocksum);
    const uint32_t total = carry + (blocksum & 0xffffu), markers = rcarry + (blocksum >> 16);
    if (tid == 0) {
        cnt[ce] = carry;
        rcnt[ce] = rcarry;
        sg.scan_bits = total * 8u;
        const uint32_t S = 1u << sg.sub_log2;
        sg.sub_count = max(1u, (total * 8u + S - 1u) >> sg.sub_log2);
        /* as many RSTn as the frame size and DRI call for (the host loop: GJD_E_CORRUPT otherwise) */
        if (markers != sg.rst_count) ustatus[r.img] = GJD_DEV_BADCOUNT;
    }
    if (tid < 16) b.scan[sg.scan_off + total + tid] = 0;     /* the bit window reads up to 12 bytes past the end */
}

__global__ void __launch_bounds__(256) k_unstuff_write(GjdDevBatch b) {
    __shared__ uint32_t ws[8];
    __shared__ __align__(16) uint8_t stage[GJD_UCHUNK + 16];
    const uint32_t ri = b.uchunk_raw[blockIdx.x], tid = threadIdx.x;
    const GjdRaw& r = b.raws[ri];
    const uint32_t chunk = blockIdx.x - r.chunk_first;
    const uint32_t eoi = b.raw_eoi[ri];
    const int64_t cstart = (int64_t)chunk * GJD_UCHUNK - (int64_t)(r.raw_off & 15u);    /* position of the chunk's first byte */
    if (cstart >= (int64_t)eoi) return;             /* block-uniform: behind the end of the scan (or a failed image) */
    const GjdUnstuffLane L = gjd_unstuff_lane(b, r, chunk, eoi);
    uint32_t keep = L.valid & ~L.m.drop;
    uint32_t total;
    uint32_t o = block_excl_scan_256(__popc(keep) | ((uint32_t)__popc(L.m.rst) << 16), ws, &total);
    total &= 0xffffu;
    if (L.m.rst) {                                  /* restart markers: number them, check the 0..7 cycle, note where the
                                                       next interval starts in the unstuffed scan */
        const GjdSegment& sg = b.segs[r.seg];
        uint32_t ord = b.uchunk_rst[blockIdx.x] + (o >> 16);
        const uint32_t at0 = b.uchunk_cnt[blockIdx.x] + (o & 0xffffu);
        for (uint32_t mk = L.m.rst; mk; mk &= mk - 1u, ord++) {
            const uint32_t i = __ffs(mk) - 1u;
            const uint32_t nb = i < 15u ? gjd_lane_byte(L, i + 1u) : L.next;
            if (((nb - 0xd0u) & 7u) != (ord & 7u)) atomicOr(b.ustatus + r.img, GJD_DEV_UFORMAT);
            if (ord < sg.rst_count) b.rst[sg.rst_first + ord] = (at0 + __popc(keep & ((1u << i) - 1u))) * 8u;
        }
    }
    o &= 0xffffu;
    if (keep == 0xffffu && (o & 3u) == 0u) {
#pragma unroll
        for (int j = 0; j < 4; j++) *reinterpret_cast<uint32_t*>(stage + o + 4 * j) = L.w[j];
    } else {
        while (keep) {
            const uint32_t i = __ffs(keep) - 1u;
            keep &= keep - 1u;
            stage[o++] = (uint8_t)gjd_lane_byte(L, i);
        }
    }
    __syncthreads();
    uint8_t* dst = b.scan + b.segs[r.seg].scan_off + b.uchunk_cnt[blockIdx.x];
    const uint32_t head = min(total, (uint32_t)(-(intptr_t)dst) & 3u);      /* bytes up to the first aligned word of dst */
    const uint32_t nword = (total - head) >> 2;
    if (tid < head) dst[tid] = stage[tid];
    const uint32_t tail0 = head + 4u * nword;
    if (tid >= 32u && tid - 32u < total - tail0) dst[tail0 + tid - 32u] = stage[tail0 + tid - 32u];
    uint32_t* dw = reinterpret_cast<uint32_t*>(dst + head);
    const uint32_t* sw = reinterpret_cast<const uint32_t*>(stage);
    const uint32_t sel = 0x3210u + 0x1111u * head;                          /* bytes head..head+3 of the pair (lo, hi) */
    for (uint32_t k = tid; k < nword; k += 256u) dw[k] = __byte_perm(sw[k], sw[k + 1], sel);
}

/* ------------------------------------------------------------------ entropy: sync ----- */
/*
 * One block per sequence (256 subsequences of S = 2^sub_log2 bits of one image), one thread per
 * subsequence i.
 * Phase 1 (look-back): the thread starts cold (component 0, zigzag 0) GJD_LOOKBACK_BITS before its
 * subsequence; by the time it reaches the boundary i*S it has almost always fallen into step with the
 * true symbol sequence (self-synchronisation; 99.8 % within 512 bits on the benchmark images).  Its
 * state at the boundary is its entry state Ein_i; from there it decodes its own subsequence, counting
 * coefficient slots, to the exit state Eout_i.  Subsequence 0 of an image starts from the true state.
 * Phase 2 (validation): Ein_i must equal Eout_{i-1}.  A thread whose entry state differs re-decodes its
 * subsequence from Eout_{i-1}; rounds repeat until nothing changes (the first wrong thread of a chain
 * is always repaired, so this terminates).  The first thread of a sequence cannot see its predecessor
 * (another thread block): it leaves Ein in seq_in for k_huff_fix.
 * After the reference's parallelHuffManDecode (cuda-decoder/src/parser.cu:132-173), which continues
 * every thread through whole slices instead.
 */
#ifndef GJD_SYNC_MINB
#define GJD_SYNC_MINB 6            /* resident blocks per SM of the count pass: 40 registers; 8 (32 registers) spills since the decoders carry ring and restart state: 1.95 vs 2.00 ms (profiles/r02l_win_ahead_ab.txt) */
#endif
/* RST: segments with restart intervals inside the scan; SUB: chroma-subsampled scans (MCUs of 4 or 6 blocks, checkpoints).
 * Four kernels instead of four paths in one (launched only for batches that hold such images), so that the registers and
 * spills of one instantiation are not the others': the 4:4:4 kernel without restarts is the one the benchmark runs. */
template <bool RST, bool SUB>
__global__ void __launch_bounds__(GJD_SYNC_THREADS, GJD_SYNC_MINB) k_huff_sync(GjdDevBatch b) {
    __shared__ GjdLutSet lut;
    __shared__ uint32_t sp[GJD_SEQ_SUBS];
    __shared__ uint32_t sn[GJD_SEQ_SUBS];
#if GJD_SYNC_CHK
    __shared__ GjdChkPt cps[SUB ? GJD_SYNC_CHK : 1][SUB ? GJD_SEQ_SUBS : 1];
#endif

    const uint32_t g = blockIdx.x, tid = threadIdx.x;
    const GjdSegment& d = b.segs[b.seq_seg[g]];
    if ((d.rst_count != 0u) != RST || (d.ny > 1u) != SUB) return;    /* block-uniform: another instantiation's */
    const uint32_t ls = g - d.seq_first;
    const uint32_t sub = ls * GJD_SEQ_SUBS + tid;
    const uint32_t sub_count = d.sub_count, scan_bits = d.scan_bits;
    const uint32_t S = 1u << d.sub_log2;
#ifdef GJD_SCAN_TMA
    __shared__ __align__(128) uint8_t ring[GJD_RING_BYTES_PER_BLOCK];
    __shared__ uint64_t ring_bars[2 * GJD_SEQ_SUBS];
    GjdScanRing scan = gjd_ring_setup(ring, ring_bars, b.scan, d.scan_off);
#else
    const uint32_t* scan = reinterpret_cast<const uint32_t*>(b.scan + d.scan_off);
#endif
    copy_count_lut_to_smem(&lut, b.luts + d.lut_id, b.lutc + d.lut_id);
    const GjdLutSmem slut(&lut);
    __syncthreads();

    const bool active = sub < sub_count;
    const uint32_t start = sub << d.sub_log2;
    const uint32_t end = active ? min(start + S, scan_bits) : 0u;
    /* the same code for both MCU shapes (block-uniform choice): 4:4:4 / single-component, or ny luma blocks + Cb + Cr */
    auto body = [&](auto ring, auto rst) {
        GjdState st;
        st.p = 0; st.n = 0; st.pos = 0;
        uint32_t in_p = 0, in_pos = 0;
        if (active) {
            if (sub != 0u) {
                const uint32_t lookback = decltype(ring)::sub ? GJD_LOOKBACK_SUB_BITS : GJD_LOOKBACK_BITS;
                st.p = start > lookback ? start - lookback : 0u;
                gjd_count_span_pairs(scan, slut, start, scan_bits, st, ring, rst);
            }
            in_p = st.p; in_pos = st.pos;
            st.n = 0;
#if GJD_SYNC_CHK
            /* checkpoints for chroma-subsampled scans only: with four luma blocks per MCU a cold decode lands in the wrong
             * block of the MCU half of the time (10 % for 4:2:2, never on the 4:4:4 benchmark images), and all but 2 % of the
             * repeats are back on the first decode's path at the first checkpoint (tests/hostsim stats) */
            if constexpr (decltype(ring)::sub) {
                GjdChkRecord rec = {&cps[0][tid], start + S / (GJD_SYNC_CHK + 1u), S / (GJD_SYNC_CHK + 1u), 0u};
                gjd_count_span_pairs(scan, slut, end, scan_bits, st, ring, rst, &rec);
            } else
#endif
            gjd_count_span_pairs(scan, slut, end, scan_bits, st, ring, rst);
        }
        sp[tid] = st.p;
        sn[tid] = (st.n << GJD_POS_BITS) | st.pos;
        if (tid == 0 && ls != 0u) b.seq_in[g] = make_uint2(in_p, in_pos);

        for (uint32_t r = 0; r < GJD_SEQ_SUBS; r++) {
            __syncthreads();
            bool bad = false;
            if (active && tid != 0u) {
                const uint32_t pp = sp[tid - 1], ppos = sn[tid - 1] & GJD_POS_MASK;
                bad = pp != in_p || ppos != in_pos;
                if (bad) {
                    st.p = in_p = pp; st.pos = in_pos = ppos; st.n = 0;
#if GJD_SYNC_CHK
                    if constexpr (decltype(ring)::sub) {
                        GjdChkCompare cmp = {&cps[0][tid], start + S / (GJD_SYNC_CHK + 1u), S / (GJD_SYNC_CHK + 1u), 0u, 0u, false};
                        gjd_count_span_pairs(scan, slut, end, scan_bits, st, ring, rst, &cmp);
                        if (cmp.hit) gjd_chk_finish(cmp, st, sp[tid], sn[tid]);
                    } else
#endif
                    gjd_count_span_pairs(scan, slut, end, scan_bits, st, ring, rst);
                }
            }
            if (!__syncthreads_or(bad)) break;
            if (bad) {
                sp[tid] = st.p;
                sn[tid] = (st.n << GJD_POS_BITS) | st.pos;
            }
        }
    };
    /* block-uniform: with or without restart intervals inside the scan */
    auto with_rst = [&](auto ring) {
        if constexpr (RST) body(ring, GjdRstTable(b.rst + d.rst_first, d.rst_count)); else body(ring, GjdNoRst());
    };
    if constexpr (SUB) with_rst(GjdRingSub(d.ny)); else with_rst(GjdRing444());
    if (active) b.sinfo[(size_t)g * GJD_SEQ_SUBS + tid] = make_uint2(sp[tid], sn[tid]);
#ifdef GJD_SCAN_TMA
    scan.drain();
#endif
}

/*
 * One thread per sequence that is not the first of its image: if the entry state of the sequence's first
 * subsequence differs from the exit state of the previous sequence, continue from that exit state until
 * the chain merges with what k_huff_sync left.  A chain that crosses the whole sequence without merging
 * is counted in *nosync; the host then repeats this kernel.
 */
__global__ void __launch_bounds__(128) k_huff_fix(GjdDevBatch b) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= b.n_seq) return;
    const GjdSegment& d = b.segs[b.seq_seg[g]];
    uint32_t ls = g - d.seq_first;
    if (ls == 0) return;
    uint2* si = b.sinfo + (size_t)g * GJD_SEQ_SUBS;
    const uint2 prev = si[-1];
    const uint2 ein = b.seq_in[g];
    if (prev.x == ein.x && (prev.y & GJD_POS_MASK) == ein.y) return;
    b.seq_in[g] = make_uint2(prev.x, prev.y & GJD_POS_MASK);
    const uint32_t* scan = reinterpret_cast<const uint32_t*>(b.scan + d.scan_off);
    GjdLutMem glut;
    glut.set = b.luts + d.lut_id;
    glut.badcode = nullptr;
    const uint32_t scan_bits = d.scan_bits;
    GjdState st;
    st.p = prev.x;
    st.pos = prev.y & GJD_POS_MASK;
    bool merged = false;
    for (uint32_t t = 0; t < GJD_SEQ_SUBS; t++) {
        uint32_t sub = ls * GJD_SEQ_SUBS + t;
        if (sub >= d.sub_count) { merged = true; break; }
        st.n = 0;
        uint32_t end = min((sub + 1) << d.sub_log2, scan_bits);
        if (d.rst_count) {
            const GjdRstTable rst(b.rst + d.rst_first, d.rst_count);
            if (d.ny > 1u) gjd_count_span(scan, glut, end, scan_bits, st, GjdRingSub(d.ny), rst);
            else gjd_count_span(scan, glut, end, scan_bits, st, GjdRing444(), rst);
        } else if (d.ny > 1u) gjd_count_span(scan, glut, end, scan_bits, st, GjdRingSub(d.ny));
        else gjd_count_span(scan, glut, end, scan_bits, st);
        uint2 old = si[t];
        si[t] = make_uint2(st.p, (st.n << GJD_POS_BITS) | st.pos);
        if (old.x == st.p && (old.y & GJD_POS_MASK) == st.pos) { merged = true; break; }
    }
    if (!merged && ls + 1 < d.seq_count) {
        atomicAdd(b.nosync, 1u);
    }
}

/* ------------------------------------------------------------------ entropy: scan ----- */
/*
 * One block per segment: coefficient count of every sequence, exclusive scan over the segment's
 * sequences -> seq_base.  Also checks the segment's total against 64 * its block count.
 */
__global__ void __launch_bounds__(256) k_seq_scan(GjdDevBatch b) {
    __shared__ uint32_t ws[8];
    __shared__ uint32_t tot_s[256];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
    const GjdSegment& d = b.segs[blockIdx.x];
    const uint32_t img = d.img;
    uint32_t carry = 0;
    for (uint32_t base = 0; base < d.seq_count; base += 256) {
        /* totals of up to 256 sequences: a warp per sequence, coalesced over its 256 subsequence records */
        for (uint32_t i = wid; i < 256u; i += 8u) {
            const uint32_t s = base + i;
            if (s >= d.seq_count) break;                      /* warp-uniform */
            const uint2* si = b.sinfo + (size_t)(d.seq_first + s) * GJD_SEQ_SUBS;
            /* (a device-unstuffed scan may be shorter than the raw length its sequences were sized from: the last
             * sequences can be empty) */
            const uint32_t cnt = s * GJD_SEQ_SUBS < d.sub_count ? min(GJD_SEQ_SUBS, d.sub_count - s * GJD_SEQ_SUBS) : 0u;
            uint32_t t = 0;
            for (uint32_t k = lane; k < cnt; k += 32u) t += si[k].y >> GJD_POS_BITS;
            t = __reduce_add_sync(0xffffffffu, t);
            if (lane == 0) tot_s[i] = t;
        }
        __syncthreads();
        const uint32_t s = base + tid;
        const uint32_t total = s < d.seq_count ? tot_s[tid] : 0u;
        uint32_t blocksum;
        uint32_t excl = block_excl_scan_256(total, ws, &blocksum);
        if (s < d.seq_count) b.seq_base[d.seq_first + s] = carry + excl;
        carry += blocksum;
        __syncthreads();                                      /* tot_s is rewritten by the next chunk */
    }
    if (tid == 0 && carry != 64u * d.blk_count) atomicOr(b.status + img, GJD_DEV_BADCOUNT);
}

/* ------------------------------------------------------------------ entropy: write ---- */

#define GJD_WRITE_THREADS 256
static_assert(GJD_WRITE_THREADS == GJD_SEQ_SUBS, "one thread per subsequence of a sequence");

/*
 * Staging of the block a thread is decoding: 64 int16 per thread in shared memory, interleaved across
 * the block's threads in 16-byte chunks (chunk j of thread t at buf[j][t]) so that the flush is eight
 * conflict-free 128-bit loads.  flush() moves the 32-byte sectors up to kmax to the coefficient line
 * of the block (whole sectors only: no read-modify-write in L2/HBM), re-zeroes them, and writes meta.
 */
struct GjdWriteSink {
    uint4* mine;       /* &buf[0][tid] */
    uint4* dst;        /* coefficient line of the next block this thread flushes (blocks are flushed in order) */
    uint32_t* mp;      /* its meta word */
    __device__ __forceinline__ void store(uint32_t k, int v) const {
        reinterpret_cast<int16_t*>(mine + (k >> 3) * GJD_WRITE_THREADS)[k & 7u] = (int16_t)v;
    }
    __device__ __forceinline__ void sector(uint32_t sct) const {
        uint4* s0 = mine + (2u * sct) * GJD_WRITE_THREADS;
        uint4* s1 = mine + (2u * sct + 1u) * GJD_WRITE_THREADS;
        const uint4 a = *s0, c = *s1;
        dst[2u * sct] = a;
        dst[2u * sct + 1u] = c;
        *s0 = make_uint4(0u, 0u, 0u, 0u);
        *s1 = make_uint4(0u, 0u, 0u, 0u);
    }
    /* gjd_write_span flushes consecutive blocks starting at (idx0 + 63) >> 6: running pointers replace the
     * per-block address arithmetic (blk itself is only the caller's bound check) */
    __device__ __forceinline__ void flush(uint32_t, int dcdiff, uint32_t kmax) {
        *mp = ((uint32_t)dcdiff & 0xffffu) | (kmax << 16);
        if (kmax != 0u) {
            sector(0u);
            if (kmax >= 16u) {
                sector(1u);
                if (kmax >= 32u) {
                    sector(2u);
                    if (kmax >= 48u) sector(3u);
                }
            }
        }
        mp += 1;
        dst += 8;
    }
};

/*
 * One block per sequence.  Thread i starts from the synchronised end state of subsequence i-1, at
 * coefficient index seq_base + (block-exclusive scan of the counts), and writes the blocks it owns
 * (gjd_write_span).
 */
#if GJD_LUT_BITS < 10
/* experiment: a 9-bit primary LUT, only the two LUT levels in shared memory (12 KB; the canonical tables of
 * the rare slow path are read from global memory) -> 44 KB per block, five blocks per SM */
#define GJD_WRITE_LUT_BYTES (GJD_RING_BYTES + GJD_RING_TABLES * GJD_LUT2_STRIDE)
#define GJD_WRITE_MINB 5
#else
#define GJD_WRITE_LUT_BYTES sizeof(GjdLutSet)
#define GJD_WRITE_MINB 4
#endif
#ifndef GJD_WRITE_SMEM_EXTRA
#define GJD_WRITE_SMEM_EXTRA 0u    /* experiment knob: unused shared memory per block, lowers the resident blocks per SM (4 -> 3 at 20 KB, -> 2 at 56 KB) */
#endif
#ifdef GJD_SCAN_TMA
#define GJD_WRITE_SMEM (GJD_WRITE_LUT_BYTES + 8u * GJD_WRITE_THREADS * sizeof(uint4) + 2u * GJD_SEQ_SUBS * GJD_RING_CHUNK + 16u * GJD_SEQ_SUBS)
#else
#define GJD_WRITE_SMEM (GJD_WRITE_LUT_BYTES + 8u * GJD_WRITE_THREADS * sizeof(uint4) + GJD_WRITE_SMEM_EXTRA)   /* dynamic: > 48 KB with the whole set */
#endif
/* (RST, SUB: as for k_huff_sync — one kernel per combination, launched where the batch holds such images) */
template <bool RST, bool SUB>
__global__ void __launch_bounds__(GJD_WRITE_THREADS, GJD_WRITE_MINB) k_huff_write(GjdDevBatch b) {
    extern __shared__ uint4 dyn_smem[];
    static_assert(GJD_WRITE_LUT_BYTES % 16 == 0 && GJD_WRITE_LUT_BYTES <= sizeof(GjdLutSet), "LUT part of the set");
    GjdLutSet& lut = *reinterpret_cast<GjdLutSet*>(dyn_smem);
    uint4 (*buf)[GJD_WRITE_THREADS] = reinterpret_cast<uint4 (*)[GJD_WRITE_THREADS]>(dyn_smem + GJD_WRITE_LUT_BYTES / 16);
    __shared__ uint32_t ws[8];
    __shared__ uint32_t badcode;
    const uint32_t g = blockIdx.x, tid = threadIdx.x;
    const GjdSegment& d = b.segs[b.seq_seg[g]];
    if ((d.rst_count != 0u) != RST || (d.ny > 1u) != SUB) return;    /* block-uniform: another instantiation's */
    const uint32_t img = d.img;
    const uint32_t ls = g - d.seq_first;
    const uint32_t sub = ls * GJD_SEQ_SUBS + tid;
    const uint32_t scan_bits = d.scan_bits;
#ifdef GJD_SCAN_TMA
    uint8_t* ring = reinterpret_cast<uint8_t*>(dyn_smem) + GJD_WRITE_LUT_BYTES + 8u * GJD_WRITE_THREADS * sizeof(uint4);
    GjdScanRing scan = gjd_ring_setup(ring, reinterpret_cast<uint64_t*>(ring + GJD_RING_BYTES_PER_BLOCK), b.scan, d.scan_off);
#else
    const uint32_t* scan = reinterpret_cast<const uint32_t*>(b.scan + d.scan_off);
#endif
    const bool active = sub < d.sub_count;
    const uint2* si = b.sinfo + (size_t)g * GJD_SEQ_SUBS;
    copy_lut_to_smem(&lut, b.luts + d.lut_id, GJD_WRITE_LUT_BYTES);
    const GjdLutSmem slut(&lut, &badcode, GJD_WRITE_LUT_BYTES < sizeof(GjdLutSet) ? b.luts + d.lut_id : &lut);
    if (tid == 0) badcode = 0u;
#pragma unroll
    for (int j = 0; j < 8; j++) buf[j][tid] = make_uint4(0u, 0u, 0u, 0u);

    uint32_t n = active ? (si[tid].y >> GJD_POS_BITS) : 0u;
    uint32_t blocksum;
    uint32_t excl = block_excl_scan_256(n, ws, &blocksum);   /* has the __syncthreads the LUT copy needs */

    GjdState st;
    st.p = 0; st.n = 0; st.pos = 0;
    if (active && sub != 0u) {
        uint2 prev = si[(int)tid - 1];
        st.p = prev.x;
        st.pos = prev.y & GJD_POS_MASK;
    }
    uint32_t bad = 0;
    GjdWriteSink sink;
    sink.mine = &buf[0][tid];
    const size_t blk0 = (size_t)b.imgs[img].tile_first * 96u + d.blk_first;   /* the segment's first block */
    const uint32_t idx0 = b.seq_base[g] + excl;
    const size_t first = blk0 + ((idx0 + 63u) >> 6);          /* the first block this thread owns */
    sink.dst = reinterpret_cast<uint4*>(b.coef) + first * 8u;
    sink.mp = b.meta + first;
    const uint32_t mcu_slots = ((uint32_t)d.ny + 2u) * 64u;       /* coefficient slots per MCU: 192 for 4:4:4 and single-component scans */
    if (active && idx0 % mcu_slots != st.pos && st.p < scan_bits) bad |= GJD_DEV_BADSTATE;
    uint32_t end = min((sub + 1) << d.sub_log2, scan_bits);
    /* block-uniform choices: the MCU shape, and restart intervals inside the scan */
    if constexpr (RST) {
        const GjdRstTable rst(b.rst + d.rst_first, d.rst_count);
        if constexpr (SUB) gjd_write_span(scan, slut, end, scan_bits, st, idx0, d.blk_count, active, sink, bad, GjdRingSub(d.ny), rst, d.rst_blocks);
        else gjd_write_span(scan, slut, end, scan_bits, st, idx0, d.blk_count, active, sink, bad, GjdRing444(), rst, d.rst_blocks);
    } else if constexpr (SUB) gjd_write_span(scan, slut, end, scan_bits, st, idx0, d.blk_count, active, sink, bad, GjdRingSub(d.ny));
    else gjd_write_span(scan, slut, end, scan_bits, st, idx0, d.blk_count, active, sink, bad);
#ifdef GJD_SCAN_TMA
    scan.drain();
#endif
    if (bad) atomicOr(b.status + img, bad);
    __syncthreads();
    if (tid == 0 && badcode) atomicOr(b.status + img, GJD_DEV_BADCODE);
}

/* ------------------------------------------------------------------ DC predictors ----- */
/*
 * tilebase[tile] = the three DC predictors at the start of the tile (cpp-decoder/src/parser.cpp:108-111,141:
 * three running predictors, never reset; restart intervals are applied in the transform).  Two steps so that
 * one huge image (BASELINE config 4: 32768 tiles) is not summed by a single block:
 *   k_dc_tilesum   one thread per tile of the batch: sums the DC differences of the tile's 96 meta words
 *                  (32 MCUs x 3 components, 96 blocks of a single-component image, or 16 / 24 MCUs of a chroma-subsampled one)
 *   k_dc_tilebase  one block per image: exclusive scan of its tile sums, in place
 */
__global__ void __launch_bounds__(256) k_dc_tilesum(GjdDevBatch b) {
    const uint32_t t = blockIdx.x * 256u + threadIdx.x;
    if (t >= b.n_tile) return;
    const uint4* mt = reinterpret_cast<const uint4*>(b.meta) + (size_t)t * 24u;
    const GjdImage& im = b.imgs[b.tile_img[t]];
    const uint32_t ny = im.gray ? 1u : (uint32_t)im.hs * im.vs;
    int sum[3] = {0, 0, 0};
    if (ny == 4u) {            /* 4:2:0: lines Y Y Y Y Cb Cr, 16 MCUs per tile */
#pragma unroll
        for (int i = 0; i < 24; i++) {
            const uint4 v = __ldg(mt + i);
            const int d4[4] = {(int)(int16_t)(v.x & 0xffffu), (int)(int16_t)(v.y & 0xffffu), (int)(int16_t)(v.z & 0xffffu), (int)(int16_t)(v.w & 0xffffu)};
#pragma unroll
            for (int k = 0; k < 4; k++) sum[(i * 4 + k) % 6 < 4 ? 0 : (i * 4 + k) % 6 - 3] += d4[k];
        }
    } else if (ny == 2u) {     /* 4:2:2 / 4:4:0: lines Y Y Cb Cr, 24 MCUs per tile */
#pragma unroll
        for (int i = 0; i < 24; i++) {
            const uint4 v = __ldg(mt + i);
            sum[0] += (int)(int16_t)(v.x & 0xffffu) + (int)(int16_t)(v.y & 0xffffu);
            sum[1] += (int)(int16_t)(v.z & 0xffffu);
            sum[2] += (int)(int16_t)(v.w & 0xffffu);
        }
    } else {
#pragma unroll
        for (int i = 0; i < 24; i++) {
            const uint4 v = __ldg(mt + i);
            sum[(i * 4) % 3] += (int)(int16_t)(v.x & 0xffffu);
            sum[(i * 4 + 1) % 3] += (int)(int16_t)(v.y & 0xffffu);
            sum[(i * 4 + 2) % 3] += (int)(int16_t)(v.z & 0xffffu);
            sum[(i * 4 + 3) % 3] += (int)(int16_t)(v.w & 0xffffu);
        }
        if (im.gray) { sum[0] += sum[1] + sum[2]; sum[1] = sum[2] = 0; }   /* one DC chain over the 96 blocks */
    }
    b.tilebase[t] = make_int4(sum[0], sum[1], sum[2], 0);
}

__global__ void __launch_bounds__(256) k_dc_tilebase(GjdDevBatch b) {
    __shared__ int ws[3][8];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
    const GjdImage& d = b.imgs[blockIdx.x];
    int4* tbase = b.tilebase + d.tile_first;
    int carry[3] = {0, 0, 0};
    /* 2048 tiles per round: a thread takes 8 consecutive tiles, the block scans the three per-thread sums at once */
    for (uint32_t base = 0; base < d.tile_count; base += 2048u) {
        const uint32_t t0 = base + tid * 8u;
        int4 v[8];
        int sum[3] = {0, 0, 0};
#pragma unroll
        for (int j = 0; j < 8; j++) {
            v[j] = t0 + j < d.tile_count ? tbase[t0 + j] : make_int4(0, 0, 0, 0);
            sum[0] += v[j].x; sum[1] += v[j].y; sum[2] += v[j].z;
        }
        int incl[3];
#pragma unroll
        for (int c = 0; c < 3; c++) incl[c] = warp_incl_scan_i(sum[c], lane);
        __syncthreads();   /* ws may still be read from the previous round */
        if (lane == 31u) { ws[0][wid] = incl[0]; ws[1][wid] = incl[1]; ws[2][wid] = incl[2]; }
        __syncthreads();
        int run[3];
#pragma unroll
        for (int c = 0; c < 3; c++) {
            int before = 0, tot = 0;
#pragma unroll
            for (int w = 0; w < 8; w++) {
                const int x = ws[c][w];
                if (w < (int)wid) before += x;
                tot += x;
            }
            run[c] = carry[c] + before + incl[c] - sum[c];
            carry[c] += tot;
        }
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (t0 + j < d.tile_count) tbase[t0 + j] = make_int4(run[0], run[1], run[2], 0);
            run[0] += v[j].x; run[1] += v[j].y; run[2] += v[j].z;
        }
    }
}

/* ------------------------------------------------------------------ transform --------- */

#define GJD_TRANSFORM_THREADS 96

/* DC of (mcu, component c) in an image with restart intervals (T.81 E.1.4: the predictors restart from 0
 * every restart_interval MCUs).  incl = inclusive warp scan of the tile's DC differences, tb = the plain
 * running sums at the tile start (k_dc_tilebase ignores restarts).  A lane whose interval starts inside the
 * tile subtracts the scan value just before the interval start; the others need the plain sum at their
 * interval start ms0 (the same for all of them), read from the tile that holds ms0.  Warp-converged. */
__device__ __noinline__ int gjd_restart_dc(const int4* __restrict__ tilebase_img, const uint32_t* __restrict__ meta_img,
                                           uint32_t ri, uint32_t ltile, uint32_t mcu, uint32_t c, uint32_t lane,
                                           int incl, int4 tb) {
    const uint32_t m0 = ltile * GJD_TILE_MCUS;
    const uint32_t ms = mcu - mcu % ri;                 /* first MCU of this lane's interval */
    const int before = __shfl_sync(0xffffffffu, incl, ms > m0 ? ms - m0 - 1u : 0u);
    const int p_tile = c == 0 ? tb.x : c == 1 ? tb.y : tb.z;
    const uint32_t ms0 = m0 - m0 % ri;                  /* interval running at the tile start */
    int p_ms0 = p_tile;
    if (ms0 != m0) {                                    /* warp-uniform */
        const uint32_t j = ms0 / GJD_TILE_MCUS, r = ms0 % GJD_TILE_MCUS;
        const int4 tj = tilebase_img[j];
        int v = 0;
        if (lane < r) v = (int)(int16_t)(__ldg(meta_img + (size_t)j * 96u + lane * 3u + c) & 0xffffu);
        p_ms0 = (c == 0 ? tj.x : c == 1 ? tj.y : tj.z) + __reduce_add_sync(0xffffffffu, v);
    }
    if (ms > m0) return incl - before;
    if (ms == m0) return incl;
    return incl + p_tile - p_ms0;
}

/* natural index -> zigzag index as a compile-time function (GJD_ZZ_TABLE) */
__host__ __device__ constexpr int gjd_zz_of_nat(int i) {
    constexpr int zz[64] = GJD_ZZ_TABLE;
    return zz[i];
}
template <int I> struct GjdZz { static constexpr int value = gjd_zz_of_nat(I); };

/* v[natural i] = coefficient at zigzag(i) of the block held as 32 packed words (two int16 each), times
 * q[zigzag(i)]; the dezigzag is a compile-time renaming of registers.  Coefficients with zigzag index
 * >= KLIM are known to be zero (kmax < KLIM for the whole warp): they become literal zeros, and the
 * compiler prunes the butterflies of gjd_idct8x8 that only see zeros (e.g. KLIM = 16: rows 5-7 vanish,
 * row 4 is DC-only).  Exactness is untouched: it is the same arithmetic on the same values. */
template <int I, int KLIM> struct GjdUnpack {
    static __device__ __forceinline__ void run(const uint32_t* w, const uint32_t* qq, int* v) {
        GjdUnpack<I - 1, KLIM>::run(w, qq, v);
        constexpr int k = GjdZz<I>::value;
        if (k < KLIM) {
            /* one IDP.2A (FMA pipe) selects the 16-bit half AND multiplies: qq[j] = q[2j] | q[2j+1] << 24, so
             * dp2a.lo sees the bytes (q[2j], 0) against (lo, hi) and dp2a.hi sees (0, q[2j+1]) */
            int r;
            if (k & 1) asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(w[k >> 1]), "r"(qq[k >> 1]), "r"(0));
            else asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(w[k >> 1]), "r"(qq[k >> 1]), "r"(0));
            v[I] = r;
        } else {
            v[I] = 0;
        }
    }
};
template <int KLIM> struct GjdUnpack<0, KLIM> {
    static __device__ __forceinline__ void run(const uint32_t*, const uint32_t*, int*) {}
};

/* the quantisation table of one component as GjdUnpack wants it: 32 words q[2j] | q[2j+1] << 24 */
__device__ __forceinline__ uint32_t gjd_pack_quant(const int* q, uint32_t j) {
    return (uint32_t)q[2u * j] | ((uint32_t)q[2u * j + 1u] << 24);
}

/* dequantise + dezigzag + IDCT of one block (w: packed zigzag coefficients, dc: absolute DC), then park
 * the 64 samples (biased, u16) in shared memory rows dst[r * RSTRIDE] */
template <int KLIM, int RSTRIDE = 32>
__device__ __forceinline__ void gjd_block_to_samples(const uint32_t* w, const uint32_t* qq, int dc, uint4* dst) {
    int v[64];
    GjdUnpack<63, KLIM>::run(w, qq, v);
    v[0] = dc * (int)(qq[0] & 0xffu);
    gjd_idct8x8(v);
#pragma unroll
    for (int r = 0; r < 8; r++)
        dst[r * RSTRIDE] = make_uint4((uint32_t)v[8 * r + 1] * 65536u + (uint32_t)v[8 * r],
                                 (uint32_t)v[8 * r + 3] * 65536u + (uint32_t)v[8 * r + 2],
                                 (uint32_t)v[8 * r + 5] * 65536u + (uint32_t)v[8 * r + 4],
                                 (uint32_t)v[8 * r + 7] * 65536u + (uint32_t)v[8 * r + 6]);
}

/* 8 pixels of one row as int16 R,G,B triples (GJD_LAYOUT_INTERLEAVED_I16: element type and order of the reference's
 * outputChannels, cuda-decoder/src/parser.cu:571-573): 48 bytes per lane.  r8/g8/b8 hold the 8 clamped bytes of each
 * channel; n = pixels of this row inside the image (1..8); vec = the 48 bytes are 16-byte aligned and n == 8. */
__device__ __forceinline__ void gjd_store_row_i16(uint8_t* px, uint64_t r8, uint64_t g8, uint64_t b8, uint32_t n, bool vec) {
    if (vec) {
        uint32_t w[12];
#pragma unroll
        for (int e = 0; e < 24; e += 2) {   /* element e = pixel e / 3, channel e % 3 */
            const uint64_t lo = (e % 3 == 0 ? r8 : e % 3 == 1 ? g8 : b8) >> (8 * (e / 3));
            const uint64_t hi = ((e + 1) % 3 == 0 ? r8 : (e + 1) % 3 == 1 ? g8 : b8) >> (8 * ((e + 1) / 3));
            w[e / 2] = ((uint32_t)lo & 0xffu) | (((uint32_t)hi & 0xffu) << 16);
        }
#pragma unroll
        for (int j = 0; j < 3; j++) __stcs(reinterpret_cast<uint4*>(px) + j, make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]));
    } else {
        int16_t* o = reinterpret_cast<int16_t*>(px);
#pragma unroll 1
        for (uint32_t x = 0; x < n; x++) {
            o[3u * x] = (int16_t)((r8 >> (8u * x)) & 0xffu);
            o[3u * x + 1u] = (int16_t)((g8 >> (8u * x)) & 0xffu);
            o[3u * x + 2u] = (int16_t)((b8 >> (8u * x)) & 0xffu);
        }
    }
}

/*
 * One thread block (3 warps) per b.tb_tiles (<= GJD_TB_TILES) tiles of one image; per tile, warp c takes component c
 * and lane l the 8x8 block of MCU l:
 *   transform phase  load the block's coefficient sectors (only those the write pass produced: kmax from
 *                    meta), warp-scan the DC differences onto the tile's DC base, dequantise + dezigzag
 *                    into 64 registers, integer IDCT in registers, park the 64 samples (biased, 9 bits,
 *                    as u16) in shared memory [comp][row][lane] — conflict-free 128-bit stores.
 *   colour phase     the warps split the 8 rows (GJD_COLOR_ROWS_*): lane l reads row r of Y, Cb, Cr of MCU l (three
 *                    128-bit loads), converts 8 pixels and stores 8 bytes per plane; a warp covers 256
 *                    contiguous bytes of an image row.
 * The sample buffer is double-buffered, so one __syncthreads per tile separates the phases.
 * Splitting the components over warps keeps the straight-line code (unpack + IDCT, ~1100 instructions)
 * inside the 32 KB instruction cache; the fused one-thread-per-MCU variant (4500 instructions) was
 * instruction-fetch bound (profiles/r01b).
 * EXPORT = true is the coefficient test hook: blocks are written out in the reference layout
 * mcu*192 + comp*64 + k with the absolute DC in slot 0 instead of being transformed.
 */
/* Rows of the colour phase per warp: [0, B1) for the luma warp, [B1, B2) and [B2, 8) for the chroma warps.  The three warps
 * meet at one barrier per tile, so a tile takes as long as its slowest warp, and on photographic tiles that is the luma
 * warp (full 64-coefficient transform, ~6 colour rows' worth of instructions, against ~3 for the pruned chroma one): 16 % of
 * the kernel's warp time was barrier wait with the 2/3/3 split (profiles/r02n_k_transform.hotspots.txt).  Measured on C3
 * (profiles/r02p_transform_split_ab.txt): 3/3/2 6.85 ms, 2/3/3 6.82, 1/3/4 6.77, 0/4/4 6.72.  A per-tile choice from the
 * classes the warps actually took (a byte per warp in shared memory) cost 0.4 ms: one more live value across the butterfly
 * spills 28 bytes inside it at the 72-register cap. */
#ifndef GJD_COLOR_ROWS_B1
#define GJD_COLOR_ROWS_B1 0u
#define GJD_COLOR_ROWS_B2 4u
#endif
#ifndef GJD_TRANSFORM_HOIST
#define GJD_TRANSFORM_HOIST 1      /* image geometry in registers + reciprocal multiply for the MCU row (0: re-read per tile, divide) */
#endif
#ifndef GJD_TRANSFORM_MINB
#define GJD_TRANSFORM_MINB 8       /* resident blocks per SM of the planar transform: 80 registers, 24.4 KB of shared memory.  Nine blocks (72
                                      registers, quantisation words from global memory) were 0.4 % faster in round 1; with the image geometry
                                      kept in registers the ninth block costs spills in the butterfly: 6.71 against 6.39 ms (profiles/r02w) */
#endif
#ifndef GJD_TRANSFORM_MINB_ALT
#define GJD_TRANSFORM_MINB_ALT 8   /* the GJD_TRANSFORM_VARIANT=1 instantiation (A/B): 80 registers, quantisation words in shared memory */
#endif
template <bool EXPORT, int MINB, bool ILV>
__global__ void __launch_bounds__(GJD_TRANSFORM_THREADS, MINB) k_transform(GjdDevBatch b, int16_t* __restrict__ coef_out) {
    /* MINB >= 9 (A/B builds): the packed quantisation words are read from global memory (GjdQuantSet.packed) — the
     * 384 bytes of sq are what keeps a ninth block off the SM */
    constexpr bool QGLOBAL = MINB >= 9;
    __shared__ uint32_t sq[QGLOBAL ? 1 : 3][32];
    __shared__ uint4 smp[2][3][8][32];

    const uint32_t tid = threadIdx.x, c = tid >> 5, lane = tid & 31u;
    const GjdImage& d = b.imgs[b.tb_img[blockIdx.x]];
    if (d.gray || (uint32_t)d.hs * d.vs > 1u) return;   /* k_transform_gray's / k_transform_sub's */
    const uint32_t ltb = blockIdx.x - d.tb_first;
    {
        const int* qg = reinterpret_cast<const int*>(b.quants + d.quant_id);
        if (!QGLOBAL) sq[c][lane] = gjd_pack_quant(qg + c * 64u, lane);
    }
    __syncthreads();

    const uint32_t W = d.width, H = d.height;
    uint8_t* out = d.out_is_ptr ? reinterpret_cast<uint8_t*>(d.out_off) : b.rgb + d.out_off;
    const size_t pitch = d.row_pitch, pstride = d.plane_stride;
    /* ILV = false: the batch holds planar images only (the common case keeps the shorter kernel) */
    const bool interleaved = ILV && d.layout != GJD_LAYOUT_PLANAR;
    const bool wide = ILV && d.layout == GJD_LAYOUT_INTERLEAVED_I16;      /* int16 elements: 6 bytes per pixel */
    const bool aligned = (((size_t)out | pitch | (interleaved ? 0 : pstride)) & (wide ? 15u : 7u)) == 0;
    const uint32_t* q = QGLOBAL ? b.quants[d.quant_id].packed[c] : sq[c];
#if GJD_TRANSFORM_HOIST
    /* the image's geometry in registers: behind the per-tile __syncthreads the compiler has to re-read d.* from memory, and
     * the division by mcus_x (I2F/MUFU/fix-up, 20 instructions per tile and lane) becomes a multiply-high and one fix-up */
    const uint32_t mcus_x = d.mcus_x, tile_count = d.tile_count, mcu_count = d.mcu_count, tile_first = d.tile_first;
    const uint32_t mcus_x_inv = 0xffffffffu / mcus_x;
#else
    const uint32_t& mcus_x = d.mcus_x; const uint32_t& tile_count = d.tile_count; const uint32_t& mcu_count = d.mcu_count;
    const uint32_t& tile_first = d.tile_first;
#endif


#pragma unroll 1
    for (uint32_t t = 0; t < b.tb_tiles; t++) {
        const uint32_t ltile = ltb * b.tb_tiles + t;
        if (ltile >= tile_count) break;
        const uint32_t mcu = ltile * GJD_TILE_MCUS + lane;
        const bool valid = mcu < mcu_count;
        const size_t blk = (size_t)(tile_first + ltile) * 96u + lane * 3u + c;

        /* ---- transform phase: component c of MCU `lane` ---- */
        if (t + 1u < b.tb_tiles && ltile + 1u < tile_count) {
            /* the next tile's coefficient line and meta word on their way to L2 while this tile computes */
            asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const uint4*>(b.coef) + (blk + 96u) * 8u));
            if (lane < 12u && c == 0u) asm volatile("prefetch.global.L2 [%0];" ::"l"(b.meta + blk + 96u - lane * 3u + lane * 8u));
        }
        const uint32_t m = valid ? __ldg(b.meta + blk) : 0u;
        const uint32_t kmax = (m >> 16) & 63u;
        const uint32_t nchunk = kmax ? (kmax >> 3) + 1u : 0u;   /* 16-byte chunks up to the last coefficient */
        const uint4* src = reinterpret_cast<const uint4*>(b.coef) + blk * 8u;
        uint32_t w[32];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            uint4 r = make_uint4(0u, 0u, 0u, 0u);
            if ((uint32_t)j < nchunk) r = __ldcs(src + j);
            w[4 * j] = r.x; w[4 * j + 1] = r.y; w[4 * j + 2] = r.z; w[4 * j + 3] = r.w;
        }
        const int4 tb = b.tilebase[tile_first + ltile];
        int dc = warp_incl_scan_i((int)(int16_t)(m & 0xffffu), lane);
        if (d.restart_interval == 0u) {
            dc += c == 0 ? tb.x : c == 1 ? tb.y : tb.z;
        } else {
            dc = gjd_restart_dc(b.tilebase + tile_first, b.meta + (size_t)tile_first * 96u, d.restart_interval,
                                ltile, mcu, c, lane, dc, tb);
        }
        if (EXPORT) {
            if (valid) {
                uint4* o = reinterpret_cast<uint4*>(coef_out + (size_t)mcu * 192u + c * 64u);
                w[0] = (w[0] & 0xffff0000u) | ((uint32_t)dc & 0xffffu);
#pragma unroll
                for (int j = 0; j < 8; j++) o[j] = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
            }
            continue;
        }
        /* warp-uniform choice: most chroma tiles of photographic images have no coefficient past k = 15 */
        if (__reduce_max_sync(0xffffffffu, kmax) < 16u) gjd_block_to_samples<16>(w, q, dc, &smp[t & 1u][c][0][lane]);
        else gjd_block_to_samples<64>(w, q, dc, &smp[t & 1u][c][0][lane]);
        __syncthreads();

        /* ---- colour phase: some of the 8 rows of MCU `lane` ---- */
#if GJD_TRANSFORM_HOIST
        uint32_t by = __umulhi(mcu, mcus_x_inv), bx = mcu - by * mcus_x;     /* floor or floor - 1 (mcu < 2^27) */
        if (bx >= mcus_x) { by++; bx -= mcus_x; }
#else
        const uint32_t by = mcu / mcus_x, bx = mcu - by * mcus_x;
#endif
        const uint32_t x0 = bx * 8u, y0 = by * 8u;
        const bool fast = aligned && (x0 + 8u <= W);
        const uint32_t r0 = c == 0u ? 0u : c == 1u ? GJD_COLOR_ROWS_B1 : GJD_COLOR_ROWS_B2;
        const uint32_t rend = c == 0u ? GJD_COLOR_ROWS_B1 : c == 1u ? GJD_COLOR_ROWS_B2 : 8u;
        /* rows of this warp that exist in the image (none for lanes past the last MCU) */
        const uint32_t rows_left = (valid && y0 < H) ? H - y0 : 0u;
        const uint32_t r1 = min(rend, rows_left);
        uint8_t* rowp = out + (size_t)(y0 + r0) * pitch + (interleaved ? (size_t)x0 * (wide ? 6u : 3u) : (size_t)x0);
#pragma unroll 1
        for (uint32_t r = r0; r < r1; r++, rowp += pitch) {
            const uint4 y4 = smp[t & 1u][0][r][lane], b4 = smp[t & 1u][1][r][lane], r4 = smp[t & 1u][2][r][lane];
            const uint32_t yw[4] = {y4.x, y4.y, y4.z, y4.w}, bw[4] = {b4.x, b4.y, b4.z, b4.w}, rw[4] = {r4.x, r4.y, r4.z, r4.w};
            /* 8 pixels on 16-bit lanes (gjd_color_row8): R/B in 2^-16, G in 2^-32 fixed point with the one-sided danger zone —
             * round 1's symmetric zone sent every neutral pixel (Cb = Cr = 0) through the IEEE sequence */
            uint32_t pr[2], pg[2], pb[2];
            gjd_color_row8(yw, bw, rw, pr, pg, pb);
            const uint64_t g8 = ((uint64_t)pg[1] << 32) | pg[0];
            if (wide) {          /* block-uniform: int16 R,G,B triples, 48 bytes per lane */
                gjd_store_row_i16(rowp, ((uint64_t)pr[1] << 32) | pr[0], g8, ((uint64_t)pb[1] << 32) | pb[0], min(8u, W - x0), fast);
                continue;
            }
            if (interleaved) {   /* block-uniform: R,G,B,R,G,B,... 24 bytes per lane, 768 contiguous bytes per warp */
                uint8_t* px = rowp;
                const uint32_t g0 = (uint32_t)g8, g1 = (uint32_t)(g8 >> 32);
                if (fast) {
                    /* word j of the 24 bytes: __byte_perm picks from {a: bytes 0-3, b: bytes 4-7} */
                    const uint32_t w0 = __byte_perm(__byte_perm(pr[0], g0, 0x1040), pb[0], 0x3410);   /* R0 G0 B0 R1 */
                    const uint32_t w1 = __byte_perm(__byte_perm(g0, pb[0], 0x2051), pr[0], 0x3610);   /* G1 B1 R2 G2 */
                    const uint32_t w2 = __byte_perm(__byte_perm(pb[0], pr[0], 0x3072), g0, 0x3710);   /* B2 R3 G3 B3 */
                    const uint32_t w3 = __byte_perm(__byte_perm(pr[1], g1, 0x1040), pb[1], 0x3410);
                    const uint32_t w4 = __byte_perm(__byte_perm(g1, pb[1], 0x2051), pr[1], 0x3610);
                    const uint32_t w5 = __byte_perm(__byte_perm(pb[1], pr[1], 0x3072), g1, 0x3710);
                    __stcs(reinterpret_cast<uint2*>(px), make_uint2(w0, w1));
                    __stcs(reinterpret_cast<uint2*>(px + 8), make_uint2(w2, w3));
                    __stcs(reinterpret_cast<uint2*>(px + 16), make_uint2(w4, w5));
                } else {
                    const uint64_t r8 = ((uint64_t)pr[1] << 32) | pr[0], b8 = ((uint64_t)pb[1] << 32) | pb[0];
#pragma unroll 1
                    for (uint32_t x = 0; x < 8u && x0 + x < W; x++) {
                        px[3u * x] = (uint8_t)(r8 >> (8u * x));
                        px[3u * x + 1u] = (uint8_t)(g8 >> (8u * x));
                        px[3u * x + 2u] = (uint8_t)(b8 >> (8u * x));
                    }
                }
                continue;
            }
            uint8_t* row = rowp;
            if (fast) {
                __stcs(reinterpret_cast<uint2*>(row), make_uint2(pr[0], pr[1]));
                __stcs(reinterpret_cast<uint2*>(row + pstride), make_uint2((uint32_t)g8, (uint32_t)(g8 >> 32)));
                __stcs(reinterpret_cast<uint2*>(row + 2 * pstride), make_uint2(pb[0], pb[1]));
            } else {
                const uint64_t r8 = ((uint64_t)pr[1] << 32) | pr[0], b8 = ((uint64_t)pb[1] << 32) | pb[0];
#pragma unroll 1
                for (uint32_t x = 0; x < 8u && x0 + x < W; x++) {
                    row[x] = (uint8_t)(r8 >> (8u * x));
                    row[pstride + x] = (uint8_t)(g8 >> (8u * x));
                    row[2 * pstride + x] = (uint8_t)(b8 >> (8u * x));
                }
            }
        }
    }
}

/* ------------------------------------------------------------------ transform, round 2 -- */
/*
 * k_transform2 — GJD_TRANSFORM_VARIANT=2, a measured alternative to k_transform above (which ships).  Same result, 8 % fewer
 * instructions, and slower: 7.32 ms against 7.07 ms on C3 (profiles/r02g_transform_ab.txt).  Both kernels are bound by the
 * ALU pipe (70 % of its peak at 68-76 % issue-active, profiles/r02e_k_transform2.summary.txt), and what the sort saves are
 * mostly IMADs of pruned butterflies (FMA pipe), while its queue bookkeeping, votes and selects are ALU instructions.
 * Kept because it is the variant that wins once the shifts and clamps stop dominating (e.g. low-quality images whose
 * blocks end before zigzag index 10).  What it does differently:
 *
 *  1. Blocks are SORTED BY COST before they are transformed.  A lane used to own the block of its MCU, so a warp paid
 *     for the most expensive of its 32 blocks: every luma warp ran the full 64-coefficient butterfly, every chroma
 *     warp the 16-coefficient one although 93 % of the chroma blocks of the benchmark images end before zigzag index
 *     10.  Now a thread block takes a WINDOW of two tiles (192 blocks): phase A reads the meta words in natural order
 *     (warp c = component c, lane = MCU: the DC prefix scan needs that order), classes every block by kmax
 *     (< 10, < 28, the rest) and appends it to the class's queue in shared memory (warp-aggregated atomics); phase B
 *     hands the concatenated queues out 32 at a time, so a warp gets blocks of one class and runs the butterfly pruned
 *     for it (gjd_block_transform<10|28|64>: ~450 / ~690 / ~920 instructions).  The samples go to the block's own
 *     place in shared memory, wherever it was transformed; the colour phase reads them in MCU order as before.
 *  2. Two finished samples per packed operation at the end of the column pass, rows 0 and 4 handed over without shifts
 *     (gjd_idct8x8_raw, gjd_pack_samples).
 *  3. Colour conversion on 16-bit lanes, 8 pixels per call (gjd_color_row8): no per-pixel shift/add/clamp, and neutral
 *     pixels (Cb = Cr = 0) no longer take the IEEE path — round 1's symmetric danger zone sent 8.7 % of the rows of
 *     the benchmark images through it.
 *  4. MCU coordinates advance incrementally (one division per thread block instead of one per tile).
 * Shared memory: 24 KB of samples (two tiles, single-buffered: the barriers of the sort separate the phases anyway)
 * + 1.5 KB of queues -> 8 blocks per SM.
 */
#define GJD_TW 2u                      /* tiles per sort window */
#define GJD_WIN_BLOCKS (GJD_TW * 96u)
#define GJD_CLS0_KLIM 10
#define GJD_CLS1_KLIM 28

/* First half of a block's transform: load the coefficient chunks that can hold a coefficient below zigzag index KLIM,
 * dequantise + dezigzag (GjdUnpack), row pass.  One instantiation per class: the row pass prunes itself on the literal
 * zeros.  Lanes without a block run along with kmax = 0. */
template <int KLIM>
__device__ __forceinline__ void gjd_block_rows(const uint4* __restrict__ src, uint32_t kmax, int dc,
                                               const uint32_t* __restrict__ qq, int v[64]) {
    constexpr int NCH = (KLIM + 7) / 8;            /* 16-byte chunks that can hold a coefficient below KLIM */
    const uint32_t nchunk = kmax ? (kmax >> 3) + 1u : 0u;
    uint32_t w[4 * NCH];
#pragma unroll
    for (int j = 0; j < NCH; j++) {
        uint4 r = make_uint4(0u, 0u, 0u, 0u);
        if ((uint32_t)j < nchunk) r = __ldcs(src + j);
        w[4 * j] = r.x; w[4 * j + 1] = r.y; w[4 * j + 2] = r.z; w[4 * j + 3] = r.w;
    }
    GjdUnpack<63, KLIM>::run(w, qq, v);
    v[0] = dc * (int)(qq[0] & 0xffu);
    gjd_idct_rows_raw(v);
}

/* samples of a finished block, parked at dst[r * 32] (r = row) */
__device__ __forceinline__ void gjd_block_park(const int v[64], uint4* dst, bool active) {
    if (active) {
#pragma unroll
        for (int r = 0; r < 8; r++)
            dst[r * 32] = make_uint4(gjd_pack_samples(v[8 * r + 1], v[8 * r]), gjd_pack_samples(v[8 * r + 3], v[8 * r + 2]),
                                     gjd_pack_samples(v[8 * r + 5], v[8 * r + 4]), gjd_pack_samples(v[8 * r + 7], v[8 * r + 6]));
    }
}

/* (bx, by) of an MCU `n` further on in raster order; block-uniform branch */
__device__ __forceinline__ void gjd_mcu_advance(uint32_t& bx, uint32_t& by, uint32_t n, uint32_t mcus_x) {
    bx += n;
    if (mcus_x >= n) {
        if (bx >= mcus_x) { bx -= mcus_x; by++; }
    } else {
        const uint32_t q = bx / mcus_x;
        by += q;
        bx -= q * mcus_x;
    }
}

template <int MINB, bool ILV>
__global__ void __launch_bounds__(GJD_TRANSFORM_THREADS, MINB) k_transform2(GjdDevBatch b) {
    __shared__ uint4 smp[GJD_TW][3][8][32];
    __shared__ int s_dc[GJD_WIN_BLOCKS];           /* absolute DC of every block of the window, natural order */
    __shared__ uint8_t s_km[GJD_WIN_BLOCKS];       /* its kmax */
    __shared__ uint8_t s_q[3][GJD_WIN_BLOCKS];     /* per class: the blocks (natural index) in arrival order */
    __shared__ uint32_t s_cnt[2][4];               /* queue lengths, double-buffered over the windows */

    const uint32_t tid = threadIdx.x, c = tid >> 5, lane = tid & 31u;
    const GjdImage& d = b.imgs[b.tb_img[blockIdx.x]];
    if (d.gray || (uint32_t)d.hs * d.vs > 1u) return;   /* k_transform_gray's / k_transform_sub's */
    const uint32_t ltb = blockIdx.x - d.tb_first;
    if (tid < 8u) (&s_cnt[0][0])[tid] = 0u;
    __syncthreads();

    const uint32_t W = d.width, H = d.height, mcus_x = d.mcus_x;
    uint8_t* out = d.out_is_ptr ? reinterpret_cast<uint8_t*>(d.out_off) : b.rgb + d.out_off;
    const size_t pitch = d.row_pitch, pstride = d.plane_stride;
    const bool interleaved = ILV && d.layout != GJD_LAYOUT_PLANAR;
    const bool wide = ILV && d.layout == GJD_LAYOUT_INTERLEAVED_I16;      /* int16 elements: 6 bytes per pixel */
    const bool aligned = (((size_t)out | pitch | (interleaved ? 0 : pstride)) & (wide ? 15u : 7u)) == 0;
    const uint32_t bpp = interleaved ? (wide ? 6u : 3u) : 1u;
    const GjdQuantSet& qs = b.quants[d.quant_id];
    const uint4* coef4 = reinterpret_cast<const uint4*>(b.coef) + (size_t)d.tile_first * 96u * 8u;   /* the image's first line */
    const uint32_t* meta_img = b.meta + (size_t)d.tile_first * 96u;
    const int4* tilebase_img = b.tilebase + d.tile_first;

    const uint32_t tile0 = ltb * b.tb_tiles;
    const uint32_t tile_end = min(tile0 + b.tb_tiles, d.tile_count);
    /* raster position of this lane's MCU in the first tile of the window */
    uint32_t by0 = (tile0 * GJD_TILE_MCUS + lane) / mcus_x, bx0 = tile0 * GJD_TILE_MCUS + lane - by0 * mcus_x;
    const uint32_t lt_mask = (1u << lane) - 1u;

#pragma unroll 1
    for (uint32_t wt = tile0, win = 0; wt < tile_end; wt += GJD_TW, win++) {
        const uint32_t nt = min(GJD_TW, tile_end - wt);
        uint32_t* cnt = s_cnt[win & 1u];

        /* ---- phase A: natural order.  DC prefix, class, queue ---- */
#pragma unroll 1
        for (uint32_t t = 0; t < nt; t++) {
            const uint32_t ltile = wt + t;
            const uint32_t mcu = ltile * GJD_TILE_MCUS + lane;
            const bool valid = mcu < d.mcu_count;
            const uint32_t lblk = ltile * 96u + lane * 3u + c;            /* block number inside the image */
            if (ltile + GJD_TW < tile_end) {
                /* the lines and meta words of the next window on their way to L2 while this one computes */
                asm volatile("prefetch.global.L2 [%0];" ::"l"(coef4 + (size_t)(lblk + GJD_WIN_BLOCKS) * 8u));
                if (lane < 12u && c == 0u) asm volatile("prefetch.global.L2 [%0];" ::"l"(meta_img + (ltile + GJD_TW) * 96u + lane * 8u));
            }
            const uint32_t m = valid ? __ldg(meta_img + lblk) : 0u;
            const uint32_t kmax = (m >> 16) & 63u;
            int dc = warp_incl_scan_i((int)(int16_t)(m & 0xffffu), lane);
            const int4 tb = tilebase_img[ltile];
            if (d.restart_interval == 0u) dc += c == 0u ? tb.x : c == 1u ? tb.y : tb.z;
            else dc = gjd_restart_dc(tilebase_img, meta_img, d.restart_interval, ltile, mcu, c, lane, dc, tb);
            s_dc[t * 96u + tid] = dc;
            s_km[t * 96u + tid] = (uint8_t)kmax;
            const uint32_t cls = kmax < (uint32_t)GJD_CLS0_KLIM ? 0u : kmax < (uint32_t)GJD_CLS1_KLIM ? 1u : 2u;
            const uint32_t m0 = __ballot_sync(0xffffffffu, valid && cls == 0u), m1 = __ballot_sync(0xffffffffu, valid && cls == 1u),
                           m2 = __ballot_sync(0xffffffffu, valid && cls == 2u);
            /* one shared-memory atomic per warp and tile: lane k reserves the warp's blocks of class k */
            uint32_t base = 0;
            if (lane < 3u) base = atomicAdd(&cnt[lane], __popc(lane == 0u ? m0 : lane == 1u ? m1 : m2));
            const uint32_t mine = cls == 0u ? m0 : cls == 1u ? m1 : m2;
            const uint32_t at = __shfl_sync(0xffffffffu, base, cls) + __popc(mine & lt_mask);
            if (valid) s_q[cls][at] = (uint8_t)(t * 96u + tid);
        }
        if (tid < 4u) s_cnt[(win + 1u) & 1u][tid] = 0u;    /* the next window's counters: idle since the barrier before last */
        __syncthreads();

        /* ---- phase B: the queues, cheapest class first, 32 blocks per warp and round ---- */
        {
            const uint32_t n0 = cnt[0], n1 = cnt[1], total = n0 + n1 + cnt[2];
#pragma unroll 1
            for (uint32_t r = 0; r < GJD_TW; r++) {
                const uint32_t slot = r * 96u + tid;
                const bool act = slot < total;
                const uint32_t cls = slot < n0 ? 0u : slot < n0 + n1 ? 1u : 2u;
                const uint32_t vmax = __reduce_max_sync(0xffffffffu, act ? cls + 1u : 0u);
                if (vmax == 0u) continue;                                   /* warp-uniform: this warp has no block in this round */
                const uint32_t item = act ? s_q[cls][slot - (cls == 0u ? 0u : cls == 1u ? n0 : n0 + n1)] : 0u;
                const uint32_t t = item >= 96u ? 1u : 0u, ci = (item - t * 96u) >> 5, li = item & 31u;
                const uint32_t kmax = act ? s_km[item] : 0u;
                const int dc = s_dc[item];
                const uint4* src = coef4 + (size_t)((wt + t) * 96u + li * 3u + ci) * 8u;
                const uint32_t* q = qs.packed[ci];
                uint4* dst = &smp[t][ci][0][li];
                /* the row pass is per class; the two expensive classes share ONE column pass (code size: the instruction
                 * cache holds 32 KB, r02d measured 2.2 stall cycles per issue on instruction fetch with three full copies) */
                int v[64];
                if (vmax == 1u) {
                    gjd_block_rows<GJD_CLS0_KLIM>(src, kmax, dc, q, v);
                    gjd_idct_cols_raw<true>(v);
                } else {
                    if (vmax == 2u) gjd_block_rows<GJD_CLS1_KLIM>(src, kmax, dc, q, v);
                    else gjd_block_rows<64>(src, kmax, dc, q, v);
                    gjd_idct_cols_raw<false>(v);
                }
                gjd_block_park(v, dst, act);
            }
        }
        __syncthreads();

        /* ---- phase C: colour, in MCU order.  The window's 16 tile rows are dealt out as runs of consecutive rows of one
         * tile: warp 0 rows 0-5 of tile 0, warp 1 rows 6-7 of tile 0 and 0-3 of tile 1, warp 2 rows 4-7 of tile 1 (in
         * phase B warp 2 ends up with the most expensive class, so it gets the fewest rows) ---- */
        {
            uint32_t by1 = by0, bx1 = bx0;
            gjd_mcu_advance(bx1, by1, GJD_TILE_MCUS, mcus_x);
#pragma unroll 1
            for (uint32_t seg = 0; seg < 2u; seg++) {
                /* (tile, first row, one past the last row) of this warp's run number `seg`; empty runs have rb >= re */
                const uint32_t t = c == 0u ? 0u : c == 1u ? seg : 1u;
                const uint32_t rb = c == 0u ? (seg ? 8u : 0u) : c == 1u ? (seg ? 0u : 6u) : (seg ? 8u : 4u);
                const uint32_t re = c == 0u ? (seg ? 8u : 6u) : c == 1u ? (seg ? 4u : 8u) : 8u;
                if (rb >= re || t >= nt) continue;                          /* warp-uniform */
                const uint32_t bx = t ? bx1 : bx0, by = t ? by1 : by0;
                const uint32_t x0 = bx * 8u, y0 = by * 8u;
                const bool mvalid = (wt + t) * GJD_TILE_MCUS + lane < d.mcu_count && y0 < H;
                const uint32_t rows_left = mvalid ? H - y0 : 0u;            /* rows of this MCU that exist in the image */
                const uint32_t r1 = min(re, rows_left);
                const bool fast = aligned && (x0 + 8u <= W);
                uint8_t* rowp = out + (size_t)(y0 + rb) * pitch + (size_t)x0 * bpp;
                const uint4* sp = &smp[t][0][rb][lane];
#pragma unroll 1
                for (uint32_t r = rb; r < r1; r++, rowp += pitch, sp += 32) {
                    const uint4 y4 = sp[0], b4 = sp[8 * 32], r4 = sp[16 * 32];
                    const uint32_t yw[4] = {y4.x, y4.y, y4.z, y4.w}, bw[4] = {b4.x, b4.y, b4.z, b4.w}, rw[4] = {r4.x, r4.y, r4.z, r4.w};
                    uint32_t pr[2], pg[2], pb[2];
                    gjd_color_row8(yw, bw, rw, pr, pg, pb);
                    if (ILV && wide) {       /* block-uniform: int16 R,G,B triples, 48 bytes per lane */
                        gjd_store_row_i16(rowp, ((uint64_t)pr[1] << 32) | pr[0], ((uint64_t)pg[1] << 32) | pg[0], ((uint64_t)pb[1] << 32) | pb[0],
                                          min(8u, W - x0), fast);
                    } else if (ILV && interleaved) {   /* block-uniform: R,G,B,R,G,B,... 24 bytes per lane, 768 contiguous bytes per warp */
                        if (fast) {
                            /* word j of the 24 bytes: __byte_perm picks from {a: bytes 0-3, b: bytes 4-7} */
                            const uint32_t w0 = __byte_perm(__byte_perm(pr[0], pg[0], 0x1040), pb[0], 0x3410);   /* R0 G0 B0 R1 */
                            const uint32_t w1 = __byte_perm(__byte_perm(pg[0], pb[0], 0x2051), pr[0], 0x3610);   /* G1 B1 R2 G2 */
                            const uint32_t w2 = __byte_perm(__byte_perm(pb[0], pr[0], 0x3072), pg[0], 0x3710);   /* B2 R3 G3 B3 */
                            const uint32_t w3 = __byte_perm(__byte_perm(pr[1], pg[1], 0x1040), pb[1], 0x3410);
                            const uint32_t w4 = __byte_perm(__byte_perm(pg[1], pb[1], 0x2051), pr[1], 0x3610);
                            const uint32_t w5 = __byte_perm(__byte_perm(pb[1], pr[1], 0x3072), pg[1], 0x3710);
                            __stcs(reinterpret_cast<uint2*>(rowp), make_uint2(w0, w1));
                            __stcs(reinterpret_cast<uint2*>(rowp + 8), make_uint2(w2, w3));
                            __stcs(reinterpret_cast<uint2*>(rowp + 16), make_uint2(w4, w5));
                        } else {
                            const uint64_t r8 = ((uint64_t)pr[1] << 32) | pr[0], g8 = ((uint64_t)pg[1] << 32) | pg[0], b8 = ((uint64_t)pb[1] << 32) | pb[0];
#pragma unroll 1
                            for (uint32_t x = 0; x < 8u && x0 + x < W; x++) {
                                rowp[3u * x] = (uint8_t)(r8 >> (8u * x));
                                rowp[3u * x + 1u] = (uint8_t)(g8 >> (8u * x));
                                rowp[3u * x + 2u] = (uint8_t)(b8 >> (8u * x));
                            }
                        }
                    } else if (fast) {
                        __stcs(reinterpret_cast<uint2*>(rowp), make_uint2(pr[0], pr[1]));
                        __stcs(reinterpret_cast<uint2*>(rowp + pstride), make_uint2(pg[0], pg[1]));
                        __stcs(reinterpret_cast<uint2*>(rowp + 2 * pstride), make_uint2(pb[0], pb[1]));
                    } else {
                        const uint64_t r8 = ((uint64_t)pr[1] << 32) | pr[0], g8 = ((uint64_t)pg[1] << 32) | pg[0], b8 = ((uint64_t)pb[1] << 32) | pb[0];
#pragma unroll 1
                        for (uint32_t x = 0; x < 8u && x0 + x < W; x++) {
                            rowp[x] = (uint8_t)(r8 >> (8u * x));
                            rowp[pstride + x] = (uint8_t)(g8 >> (8u * x));
                            rowp[2 * pstride + x] = (uint8_t)(b8 >> (8u * x));
                        }
                    }
                }
            }
            gjd_mcu_advance(bx0, by0, GJD_TW * GJD_TILE_MCUS, mcus_x);
        }
    }
}

/* ------------------------------------------------------------------ transform: grayscale -- */
/*
 * Single-component images (beyond the reference's parser, which wants three components).
 * One thread per 8x8 block, 96 consecutive blocks (one "tile" of the coefficient layout) per pass:
 * block-wide inclusive scan of the DC differences (one predictor chain, restarted every restart_interval
 * blocks), dequantise + IDCT in registers, then the reference's colour formula with Cb = Cr = 0
 * (cpp-decoder/utils/color.cpp:10-16 gives R = G = B = clamp(Y + 128) up to its float rounding, which
 * gjd_color reproduces) and 8-byte stores to the three planes (or interleaved rows).
 */
template <bool EXPORT>
__global__ void __launch_bounds__(GJD_TRANSFORM_THREADS, 4) k_transform_gray(GjdDevBatch b, int16_t* __restrict__ coef_out) {
    __shared__ uint32_t sq[32];
    __shared__ int wsum[3];
    __shared__ int sincl[GJD_TRANSFORM_THREADS];
    __shared__ int s_red;
    const uint32_t tid = threadIdx.x, wid = tid >> 5, lane = tid & 31u;
    const GjdImage& d = b.imgs[b.tb_img[blockIdx.x]];
    if (!d.gray) return;
    const uint32_t ltb = blockIdx.x - d.tb_first;
    if (tid < 32u) sq[tid] = gjd_pack_quant(reinterpret_cast<const int*>(b.quants + d.quant_id), tid);
    const uint32_t W = d.width, H = d.height, ri = d.restart_interval;
    uint8_t* out = d.out_is_ptr ? reinterpret_cast<uint8_t*>(d.out_off) : b.rgb + d.out_off;
    const size_t pitch = d.row_pitch, pstride = d.plane_stride;
    const bool interleaved = d.layout != GJD_LAYOUT_PLANAR;
    const bool wide = d.layout == GJD_LAYOUT_INTERLEAVED_I16;
    const bool aligned = (((size_t)out | pitch | (interleaved ? 0 : pstride)) & (wide ? 15u : 7u)) == 0;
    const uint32_t* meta_img = b.meta + (size_t)d.tile_first * 96u;

#pragma unroll 1
    for (uint32_t t = 0; t < b.tb_tiles; t++) {
        const uint32_t ltile = ltb * b.tb_tiles + t;
        if (ltile >= d.tile_count) break;
        const uint32_t m0 = ltile * 96u, blkno = m0 + tid;       /* block number inside the image */
        const bool valid = blkno < d.mcu_count;
        const size_t blk = (size_t)d.tile_first * 96u + blkno;
        const uint32_t m = valid ? __ldg(b.meta + blk) : 0u;
        const uint32_t kmax = (m >> 16) & 63u;
        const uint32_t nchunk = kmax ? (kmax >> 3) + 1u : 0u;   /* 16-byte chunks up to the last coefficient */
        const uint4* src = reinterpret_cast<const uint4*>(b.coef) + blk * 8u;
        uint32_t w[32];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            uint4 r = make_uint4(0u, 0u, 0u, 0u);
            if ((uint32_t)j < nchunk) r = __ldcs(src + j);
            w[4 * j] = r.x; w[4 * j + 1] = r.y; w[4 * j + 2] = r.z; w[4 * j + 3] = r.w;
        }
        /* inclusive scan of the DC differences over the 96 blocks of the tile */
        int incl = warp_incl_scan_i((int)(int16_t)(m & 0xffffu), lane);
        __syncthreads();                       /* previous pass done with wsum / sincl */
        if (lane == 31u) wsum[wid] = incl;
        __syncthreads();
        if (wid >= 1u) incl += wsum[0];
        if (wid >= 2u) incl += wsum[1];
        const int p_tile = b.tilebase[d.tile_first + ltile].x;    /* plain running sum at the tile start */
        int dc = incl + p_tile;
        if (ri != 0u) {                        /* T.81 E.1.4: the predictor restarts every ri blocks */
            sincl[tid] = incl;
            const uint32_t ms = blkno - blkno % ri, ms0 = m0 - m0 % ri;
            int p_ms0 = p_tile;
            if (ms0 != m0) {                   /* block-uniform: the running interval began in an earlier tile */
                const uint32_t j = ms0 / 96u, r = ms0 % 96u;
                int v = tid < r ? (int)(int16_t)(__ldg(meta_img + (size_t)j * 96u + tid) & 0xffffu) : 0;
                v = __reduce_add_sync(0xffffffffu, v);
                if (tid == 0u) s_red = 0;
                __syncthreads();
                if (lane == 0u) atomicAdd(&s_red, v);
                __syncthreads();
                p_ms0 = b.tilebase[d.tile_first + j].x + s_red;
            } else {
                __syncthreads();
            }
            if (ms > m0) dc = incl - sincl[ms - m0 - 1u];
            else if (ms == m0) dc = incl;
            else dc = incl + p_tile - p_ms0;
        }
        if (EXPORT) {
            if (valid) {   /* blocks in the reference's slot layout, as if they were the Y blocks of an image */
                uint4* o = reinterpret_cast<uint4*>(coef_out + (size_t)blkno * 64u);
                w[0] = (w[0] & 0xffff0000u) | ((uint32_t)dc & 0xffffu);
#pragma unroll
                for (int j = 0; j < 8; j++) o[j] = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
            }
            continue;
        }
        int v[64];
        GjdUnpack<63, 64>::run(w, sq, v);
        v[0] = dc * (int)(sq[0] & 0xffu);
        gjd_idct8x8(v);
        if (!valid) continue;
        const uint32_t by = blkno / d.mcus_x, bx = blkno - by * d.mcus_x;
        const uint32_t x0 = bx * 8u, y0 = by * 8u;
        const bool fast = aligned && (x0 + 8u <= W);
        uint8_t* rowp = out + (size_t)y0 * pitch + (interleaved ? (size_t)x0 * (wide ? 6u : 3u) : (size_t)x0);
#pragma unroll
        for (int r = 0; r < 8; r++) {
            if (y0 + (uint32_t)r < H) {
                uint64_t r8 = 0, g8 = 0, b8 = 0;
#pragma unroll
                for (int x = 0; x < 8; x++) {
                    int R, G, B;
                    gjd_color(v[8 * r + x], 256, 256, R, G, B);
                    r8 |= (uint64_t)R << (8 * x); g8 |= (uint64_t)G << (8 * x); b8 |= (uint64_t)B << (8 * x);
                }
                uint8_t* row = rowp + (size_t)r * pitch;
                if (wide) {
                    gjd_store_row_i16(row, r8, g8, b8, min(8u, W - x0), fast);
                } else if (fast && !interleaved) {
                    __stcs(reinterpret_cast<uint2*>(row), make_uint2((uint32_t)r8, (uint32_t)(r8 >> 32)));
                    __stcs(reinterpret_cast<uint2*>(row + pstride), make_uint2((uint32_t)g8, (uint32_t)(g8 >> 32)));
                    __stcs(reinterpret_cast<uint2*>(row + 2 * pstride), make_uint2((uint32_t)b8, (uint32_t)(b8 >> 32)));
                } else {
#pragma unroll 1
                    for (uint32_t x = 0; x < 8u && x0 + x < W; x++) {
                        const uint8_t R = (uint8_t)(r8 >> (8u * x)), G = (uint8_t)(g8 >> (8u * x)), B = (uint8_t)(b8 >> (8u * x));
                        if (interleaved) { row[3u * x] = R; row[3u * x + 1u] = G; row[3u * x + 2u] = B; }
                        else { row[x] = R; row[pstride + x] = G; row[2 * pstride + x] = B; }
                    }
                }
            }
        }
    }
}

/* ------------------------------------------------------------------ chroma-subsampled images -- */
/*
 * 4:2:2, 4:4:0 and 4:2:0 (what data_preprocessing/image_converter.py:14-19 exists to avoid; the reference's
 * parser ignores the sampling factors and garbles such files).  An MCU is hs*vs luma blocks (row-major inside the MCU),
 * then Cb, then Cr (T.81 A.2.3); the entropy kernels run unchanged apart from the ring stepper (GjdRingSub) and write one
 * 128-byte coefficient line per block in scan order, so line = mcu * (hs*vs + 2) + block-in-MCU.  Two more kernels:
 *   k_dc_sub_tiles   the three DC chains — luma blocks in scan order, Cb, Cr — per tile and in parallel on top of the tile bases
 *                    of k_dc_tilebase (segmented scans, heads at the restart intervals): absolute DC of every line -> dcabs
 *   k_transform_sub  one thread per line: dequantise, dezigzag, IDCT (the same gjd_idct8x8 as everything else), samples
 *                    parked in shared memory; then 8-pixel row pieces: luma sample + the chroma sample that covers it
 *                    (sample replication, i.e. libjpeg's h2v1 / h2v2 "non-fancy" upsampling), the reference's colour
 *                    formula (gjd_color_row8, the packed form of the 4:4:4 kernel), 8-byte stores.
 * Pin (no reference output exists): coefficients against jpeg_read_coefficients of the reference's vendored libjpeg 9e,
 * pixels against the oracle's IDCT + replication + the oracle's colour function (tools/make_golden_sub.py).
 */
/* Inclusive SEGMENTED scan over a block of WARPS warps: v summed from the nearest thread at or below with head = true (from
 * `carry`, the running value in front of the block, when there is none).  Returns the thread's value; *carry_out = the
 * running value behind the block.  wv / wf: WARPS words of shared memory each. */
template <uint32_t WARPS>
__device__ __forceinline__ int block_seg_scan(int v, bool head, int carry, int* wv, uint32_t* wf, int* carry_out) {
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint32_t heads = __ballot_sync(0xffffffffu, head);
    const uint32_t below = heads & (0xffffffffu >> (31u - lane));           /* heads at lanes <= lane */
    const int first = below ? 31 - __clz(below) : 0;                        /* this lane's head (lane 0 stands in when none) */
    const int incl = warp_incl_scan_i(v, lane);
    const int before = __shfl_sync(0xffffffffu, incl, first > 0 ? first - 1 : 0);
    const int seg = incl - (first > 0 ? before : 0);                        /* sum since the head / since the warp's first lane */
    __syncthreads();                                                         /* wv / wf may still be read from the previous call */
    if (lane == 31u) { wv[wid] = seg; wf[wid] = heads != 0u; }
    __syncthreads();
    int run = carry, mine = carry;
#pragma unroll
    for (uint32_t w = 0; w < WARPS; w++) {
        if (w == wid) mine = run;
        run = wf[w] ? wv[w] : run + wv[w];
    }
    *carry_out = run;
    return below ? seg : mine + seg;
}

/* Absolute DC of every line of a chroma-subsampled image, per tile and in parallel (one huge 4:2:0 image is 16384 tiles, which
 * one block per image would walk in 6144 rounds).  The plain running sum of each chain at the start of a tile comes from
 * k_dc_tilesum / k_dc_tilebase (as for 4:4:4); the 96 lines of a tile are one segmented scan over three warps — threads in the
 * order luma lines, Cb lines, Cr lines (the order of k_transform_sub), which is the order of each chain.
 * Restart intervals (T.81 E.1.4: the predictors restart with the interval, i.e. at the first line of each chain in it) add
 * heads inside the tile; the lines in front of a chain's first head belong to the interval that is open at the tile start, and
 * take "plain sum at the tile start minus plain sum at that interval's start" — the latter from the tile that holds the
 * interval's first line: its base plus the chain's lines in front of that line (at most 95, summed by the block). */
#define GJD_DCSUB_TILES 8u
__global__ void __launch_bounds__(GJD_TRANSFORM_THREADS) k_dc_sub_tiles(GjdDevBatch b, int* __restrict__ dcabs) {
    __shared__ int wv[3];
    __shared__ uint32_t wf[3];
    __shared__ int part[3];
    const uint32_t tid = threadIdx.x;
    for (uint32_t k = 0; k < GJD_DCSUB_TILES; k++) {
        const uint32_t t = blockIdx.x * GJD_DCSUB_TILES + k;
        if (t >= b.n_tile) break;                               /* block-uniform */
        const GjdImage& d = b.imgs[b.tile_img[t]];
        const uint32_t ny = (uint32_t)d.hs * d.vs;
        if (d.gray || ny == 1u) continue;                       /* block-uniform: k_transform's / k_transform_gray's */
        const uint32_t nblk = ny + 2u, mcus_tile = 96u / nblk, nluma = mcus_tile * ny;
        uint32_t ml, bp;
        if (tid < nluma) { ml = tid / ny; bp = tid - ml * ny; }
        else { const uint32_t q = tid - nluma, cc = q >= mcus_tile ? 1u : 0u; ml = q - cc * mcus_tile; bp = ny + cc; }
        const uint32_t c = bp < ny ? 0u : bp - ny + 1u;
        const uint32_t l0 = (t - d.tile_first) * 96u;           /* first line of the tile, inside the image */
        const uint32_t j = l0 + ml * nblk + bp;                 /* this thread's line */
        const size_t line0 = (size_t)d.tile_first * 96u;
        const bool valid = j < d.mcu_count * nblk;
        const int diff = valid ? (int)(int16_t)(__ldg(b.meta + line0 + j) & 0xffffu) : 0;
        const int4 tb = b.tilebase[t];
        int carry = c == 0u ? tb.x : c == 1u ? tb.y : tb.z;     /* what the lines in front of the chain's first head take */
        bool head = tid == 0u || tid == nluma || tid == nluma + mcus_tile;
        bool after_head = false;
        const uint32_t rb = d.restart_interval * nblk;          /* lines per restart interval (0: none) */
        if (rb != 0u) {                                         /* block-uniform */
            const uint32_t off = c == 0u ? 0u : ny + c - 1u;    /* the chain's first line inside an interval */
            const uint32_t r0 = l0 % rb;
            const uint32_t hc = l0 + (r0 <= off ? off - r0 : rb - r0 + off);   /* the chain's first head at or behind the tile start */
            head = head || (j % rb) == off;
            after_head = j >= hc;
            const uint32_t s0 = l0 - r0;                        /* first line of the interval that is open at the tile start */
            if (s0 < l0) {
                const uint32_t ts = s0 / 96u, js = ts * 96u + tid;     /* the tile that holds it; its lines in front of s0 */
                __syncthreads();
                if (tid < 3u) part[tid] = 0;
                __syncthreads();
                if (js < s0) {
                    const uint32_t bps = js % nblk;
                    atomicAdd(&part[bps < ny ? 0u : bps - ny + 1u], (int)(int16_t)(__ldg(b.meta + line0 + js) & 0xffffu));
                }
                __syncthreads();
                const int4 tbs = b.tilebase[d.tile_first + ts];
                carry -= (c == 0u ? tbs.x : c == 1u ? tbs.y : tbs.z) + part[c];
            } else {
                carry = 0;                                      /* (an interval starts with the tile: every line is behind a head) */
            }
        }
        int unused;
        const int pre = block_seg_scan<3u>(diff, head, 0, wv, wf, &unused);
        if (valid) dcabs[line0 + j] = pre + (after_head ? 0 : carry);
    }
}

/* 8 pixels of one image row, R/G/B as 8 clamped bytes each (two words), in any of the three layouts */
__device__ __forceinline__ void gjd_store_pixels8(uint8_t* rowp, size_t pstride, const uint32_t* pr, const uint32_t* pg, const uint32_t* pb,
                                                  uint32_t n, bool fast, bool interleaved, bool wide) {
    const uint64_t r8 = ((uint64_t)pr[1] << 32) | pr[0], g8 = ((uint64_t)pg[1] << 32) | pg[0], b8 = ((uint64_t)pb[1] << 32) | pb[0];
    if (wide) {
        gjd_store_row_i16(rowp, r8, g8, b8, n, fast);
    } else if (fast && !interleaved) {
        __stcs(reinterpret_cast<uint2*>(rowp), make_uint2(pr[0], pr[1]));
        __stcs(reinterpret_cast<uint2*>(rowp + pstride), make_uint2(pg[0], pg[1]));
        __stcs(reinterpret_cast<uint2*>(rowp + 2 * pstride), make_uint2(pb[0], pb[1]));
    } else if (fast) {
        /* word j of the 24 bytes: __byte_perm picks from {a: bytes 0-3, b: bytes 4-7} */
        const uint32_t w0 = __byte_perm(__byte_perm(pr[0], pg[0], 0x1040), pb[0], 0x3410);   /* R0 G0 B0 R1 */
        const uint32_t w1 = __byte_perm(__byte_perm(pg[0], pb[0], 0x2051), pr[0], 0x3610);   /* G1 B1 R2 G2 */
        const uint32_t w2 = __byte_perm(__byte_perm(pb[0], pr[0], 0x3072), pg[0], 0x3710);   /* B2 R3 G3 B3 */
        const uint32_t w3 = __byte_perm(__byte_perm(pr[1], pg[1], 0x1040), pb[1], 0x3410);
        const uint32_t w4 = __byte_perm(__byte_perm(pg[1], pb[1], 0x2051), pr[1], 0x3610);
        const uint32_t w5 = __byte_perm(__byte_perm(pb[1], pr[1], 0x3072), pg[1], 0x3710);
        __stcs(reinterpret_cast<uint2*>(rowp), make_uint2(w0, w1));
        __stcs(reinterpret_cast<uint2*>(rowp + 8), make_uint2(w2, w3));
        __stcs(reinterpret_cast<uint2*>(rowp + 16), make_uint2(w4, w5));
    } else {
#pragma unroll 1
        for (uint32_t x = 0; x < n; x++) {
            const uint8_t R = (uint8_t)(r8 >> (8u * x)), G = (uint8_t)(g8 >> (8u * x)), B = (uint8_t)(b8 >> (8u * x));
            if (interleaved) { rowp[3u * x] = R; rowp[3u * x + 1u] = G; rowp[3u * x + 2u] = B; }
            else { rowp[x] = R; rowp[pstride + x] = G; rowp[2 * pstride + x] = B; }
        }
    }
}

/*
 * One thread block (3 warps) per b.tb_tiles tiles of one image; a tile is 96 coefficient lines = 16 MCUs of 6 blocks (4:2:0)
 * or 24 MCUs of 4 (4:2:2, 4:4:0).
 *   transform phase  thread -> line: the tile's luma lines first, then its Cb lines, then its Cr lines, so that whole warps
 *                    hold chroma blocks and take the pruned transform (4:2:0: warps 0-1 luma, warp 2 chroma); samples parked
 *                    as rows of 8 (uint4) at smp[row][line in scan order], double-buffered: one barrier per tile.
 *   colour phase     thread -> 8-pixel column of the tile's pixel rows (MCU, luma block column): consecutive lanes write
 *                    consecutive 8-byte pieces of an image row; per row the luma row of its block and the half (hs = 2) or
 *                    whole chroma row that covers it, replicated on 16-bit lanes (PRMT), through the packed colour function
 *                    of the 4:4:4 kernel (gjd_color_row8) and the same stores.
 */
#ifndef GJD_SUB_MINB
#define GJD_SUB_MINB 8
#endif
template <bool EXPORT>
__global__ void __launch_bounds__(GJD_TRANSFORM_THREADS, GJD_SUB_MINB) k_transform_sub(GjdDevBatch b, const int* __restrict__ dcabs,
                                                                                        int16_t* __restrict__ coef_out) {
    __shared__ uint4 smp[2][8][96];                         /* [tile parity][row][line of the tile]: 8 samples (biased, 9 bits, u16) */
    const uint32_t tid = threadIdx.x;
    const GjdImage& d = b.imgs[b.tb_img[blockIdx.x]];
    const uint32_t hs = d.hs, vs = d.vs, ny = hs * vs;
    if (d.gray || ny == 1u) return;
    const uint32_t nblk = ny + 2u;
    const uint32_t ltb = blockIdx.x - d.tb_first;
    const uint32_t W = d.width, H = d.height;
    uint8_t* out = d.out_is_ptr ? reinterpret_cast<uint8_t*>(d.out_off) : b.rgb + d.out_off;
    const size_t pitch = d.row_pitch, pstride = d.plane_stride;
    const bool interleaved = d.layout != GJD_LAYOUT_PLANAR;
    const bool wide = d.layout == GJD_LAYOUT_INTERLEAVED_I16;
    const bool aligned = (((size_t)out | pitch | (interleaved ? 0 : pstride)) & (wide ? 15u : 7u)) == 0;
    const uint32_t mcu_count = d.mcu_count, tile_count = d.tile_count, mcus_x = d.mcus_x;
    const uint32_t mcus_x_inv = 0xffffffffu / mcus_x;
    const uint32_t nlines = mcu_count * nblk;
    const size_t line0 = (size_t)d.tile_first * 96u;
    const uint32_t mcus_tile = 96u / nblk;                  /* 96 = 24 x 4 = 16 x 6: a tile holds whole MCUs */
    /* transform phase: this thread's line of the tile */
    const uint32_t nluma = mcus_tile * ny;
    uint32_t tml, tbp;                                      /* MCU in the tile, block in the MCU */
    if (tid < nluma) { tml = tid / ny; tbp = tid - tml * ny; }
    else { const uint32_t k = tid - nluma, cc = k >= mcus_tile ? 1u : 0u; tml = k - cc * mcus_tile; tbp = ny + cc; }
    const uint32_t tline = tml * nblk + tbp;
    const uint32_t c = tbp < ny ? 0u : tbp - ny + 1u;
    const uint32_t* qq = b.quants[d.quant_id].packed[c];
    /* colour phase: this thread's column of 8 pixels and its first row */
    const uint32_t cols = mcus_tile * hs, rpi = GJD_TRANSFORM_THREADS / cols;   /* 32/3 (4:2:0), 48/2 (4:2:2), 24/4 (4:4:0) */
    const uint32_t row0 = tid / cols, col = tid - row0 * cols;
    const uint32_t cml = hs == 2u ? col >> 1 : col, cbx = hs == 2u ? col & 1u : 0u;
    const uint32_t vshift = vs - 1u;                        /* chroma row = pixel row of the MCU >> vshift */

#pragma unroll 1
    for (uint32_t t = 0; t < b.tb_tiles; t++) {
        const uint32_t ltile = ltb * b.tb_tiles + t;
        if (ltile >= tile_count) break;
        const uint32_t j = ltile * 96u + tline;             /* line inside the image */
        const bool valid = j < nlines;
        const uint32_t m = valid ? __ldg(b.meta + line0 + j) : 0u;
        const uint32_t kmax = (m >> 16) & 63u;
        const uint32_t nchunk = kmax ? (kmax >> 3) + 1u : 0u;
        const uint4* src = reinterpret_cast<const uint4*>(b.coef) + (line0 + j) * 8u;
        uint32_t w[32];
#pragma unroll
        for (int q = 0; q < 8; q++) {
            uint4 r = make_uint4(0u, 0u, 0u, 0u);
            if ((uint32_t)q < nchunk) r = __ldcs(src + q);
            w[4 * q] = r.x; w[4 * q + 1] = r.y; w[4 * q + 2] = r.z; w[4 * q + 3] = r.w;
        }
        const int dc = valid ? dcabs[line0 + j] : 0;
        if (EXPORT) {                                       /* gjd_decode_coeffs: lines in scan order, absolute DC in slot 0 */
            if (valid) {
                uint4* o = reinterpret_cast<uint4*>(coef_out + (size_t)j * 64u);
                w[0] = (w[0] & 0xffff0000u) | ((uint32_t)dc & 0xffffu);
#pragma unroll
                for (int q = 0; q < 8; q++) o[q] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
            }
            continue;
        }
        if (__reduce_max_sync(0xffffffffu, kmax) < 16u) gjd_block_to_samples<16, 96>(w, qq, dc, &smp[t & 1u][0][tline]);
        else gjd_block_to_samples<64, 96>(w, qq, dc, &smp[t & 1u][0][tline]);
        __syncthreads();

        const uint32_t mcu = ltile * mcus_tile + cml;
        if (row0 >= rpi || mcu >= mcu_count) continue;      /* (96 = rpi * cols for the three shapes: no thread is left over) */
        uint32_t my = __umulhi(mcu, mcus_x_inv), mx = mcu - my * mcus_x;     /* floor or floor - 1 (mcu < 2^27) */
        if (mx >= mcus_x) { my++; mx -= mcus_x; }
        const uint32_t x0 = (mx * hs + cbx) * 8u, ytop = my * vs * 8u;
        if (x0 >= W) continue;
        const bool fast = aligned && x0 + 8u <= W;
        const uint32_t npx = min(8u, W - x0);
        const uint4* sm = &smp[t & 1u][0][cml * nblk];
        uint8_t* rowp = out + (size_t)(ytop + row0) * pitch + (size_t)x0 * (interleaved ? (wide ? 6u : 3u) : 1u);
#pragma unroll 1
        for (uint32_t row = row0; row < vs * 8u && ytop + row < H; row += rpi, rowp += (size_t)rpi * pitch) {
            const uint32_t r = row & 7u, yb = (row >> 3) * hs + cbx, cy = row >> vshift;
            const uint4 y4 = sm[r * 96u + yb], b4 = sm[cy * 96u + ny], r4 = sm[cy * 96u + ny + 1u];
            const uint32_t yw[4] = {y4.x, y4.y, y4.z, y4.w};
            uint32_t bw[4], rw[4];
            if (hs == 2u) {     /* block-uniform: each chroma sample covers two pixels; this luma block sees one half of the chroma row */
                const uint32_t bl = cbx ? b4.z : b4.x, bh = cbx ? b4.w : b4.y, rl = cbx ? r4.z : r4.x, rh = cbx ? r4.w : r4.y;
                bw[0] = __byte_perm(bl, 0u, 0x1010); bw[1] = __byte_perm(bl, 0u, 0x3232);
                bw[2] = __byte_perm(bh, 0u, 0x1010); bw[3] = __byte_perm(bh, 0u, 0x3232);
                rw[0] = __byte_perm(rl, 0u, 0x1010); rw[1] = __byte_perm(rl, 0u, 0x3232);
                rw[2] = __byte_perm(rh, 0u, 0x1010); rw[3] = __byte_perm(rh, 0u, 0x3232);
            } else {
                bw[0] = b4.x; bw[1] = b4.y; bw[2] = b4.z; bw[3] = b4.w;
                rw[0] = r4.x; rw[1] = r4.y; rw[2] = r4.z; rw[3] = r4.w;
            }
            uint32_t pr[2], pg[2], pb[2];
            gjd_color_row8(yw, bw, rw, pr, pg, pb);
            gjd_store_pixels8(rowp, pstride, pr, pg, pb, npx, fast, interleaved, wide);
        }
    }
}

/* ------------------------------------------------------------------ test hooks -------- */

__global__ void k_test_color(const int16_t* __restrict__ y, const int16_t* __restrict__ cb,
                             const int16_t* __restrict__ cr, size_t n, uint8_t* __restrict__ r,
                             uint8_t* __restrict__ g, uint8_t* __restrict__ bl) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int R, G, B;
    gjd_color((int)y[i] + 256, (int)cb[i] + 256, (int)cr[i] + 256, R, G, B);
    r[i] = (uint8_t)R; g[i] = (uint8_t)G; bl[i] = (uint8_t)B;
}

/* dequantise + dezigzag + IDCT of stand-alone blocks, through the same unpack as k_transform */
__global__ void __launch_bounds__(GJD_TRANSFORM_THREADS) k_test_idct(const int16_t* __restrict__ coef, size_t nblocks,
                                                                      const int32_t* __restrict__ quant,
                                                                      int16_t* __restrict__ out) {
    __shared__ uint32_t sq[32];
    if (threadIdx.x < 32) sq[threadIdx.x] = gjd_pack_quant(quant, threadIdx.x);
    __syncthreads();
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nblocks) return;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(coef + i * 64);
    uint32_t w[32];
#pragma unroll
    for (int j = 0; j < 32; j++) w[j] = src[j];
    int v[64];
    GjdUnpack<63, 64>::run(w, sq, v);
    v[0] = (int)(int16_t)(w[0] & 0xffffu) * (int)(sq[0] & 0xffu);
    gjd_idct8x8(v);
#pragma unroll
    for (int j = 0; j < 64; j++) out[i * 64 + j] = (int16_t)(v[j] - 256);
}

/* ------------------------------------------------------------------ launchers --------- */

#define GJD_CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return e_; } while (0)

cudaError_t gjd_launch_build_luts(const GjdDhtSet* dht, GjdLutSet* luts, GjdLutCount* lutc, uint32_t n, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    k_build_luts<<<n, 256, 0, s>>>(dht, luts, lutc);
    return cudaGetLastError();
}

cudaError_t gjd_launch_fill_maps(const GjdDevBatch& b, cudaStream_t s) {
    if (b.n_img == 0) return cudaSuccess;
    k_fill_maps<<<b.n_img + b.n_seg + b.n_raw, 128, 0, s>>>(b);
    return cudaGetLastError();
}

cudaError_t gjd_launch_unstuff(const GjdDevBatch& b, cudaStream_t s, int* launches) {
    if (launches) *launches = 0;
    GJD_CK(cudaMemsetAsync(b.ustatus, 0, (size_t)(b.n_img + 1) * sizeof(uint32_t), s));
    if (b.n_raw == 0) return cudaSuccess;
    nvtxRangePushA("gjd stage 0: unstuff scans on the device");
    GJD_CK(cudaMemsetAsync(b.raw_eoi, 0xff, (size_t)b.n_raw * 2 * sizeof(uint32_t), s));    /* raw_eoi and raw_err */
    k_unstuff_count<<<b.n_uchunk, 256, 0, s>>>(b);
    k_unstuff_scan<<<b.n_raw, 256, 0, s>>>(b, b.ustatus);
    k_unstuff_write<<<b.n_uchunk, 256, 0, s>>>(b);
    nvtxRangePop();
    if (launches) *launches = 3;
    return cudaGetLastError();
}

cudaError_t gjd_launch_fix(const GjdDevBatch& b, cudaStream_t s) {
    k_huff_fix<<<(b.n_seq + 127) / 128, 128, 0, s>>>(b);
    return cudaGetLastError();
}

cudaError_t gjd_launch_pipeline(const GjdDevBatch& b, cudaStream_t s, const GjdStageEvents* ev,
                                int transform_variant, int* launches, bool skip_sync, bool* attrs_set) {
    (void)transform_variant;
    int n = 0;
    if (b.n_img == 0) { if (launches) *launches = 0; return cudaSuccess; }
    if (ev) GJD_CK(cudaEventRecord(ev->e[0], s));
    GJD_CK(cudaMemsetAsync(b.meta, 0, (size_t)b.n_tile * 96 * sizeof(uint32_t), s));
    /* status starts from what the device unstuffing found (all zero for host-unstuffed batches); nosync = 0 */
    GJD_CK(cudaMemcpyAsync(b.status, b.ustatus, (size_t)(b.n_img + 1) * sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
    if (ev) GJD_CK(cudaEventRecord(ev->e[1], s));
    /* NVTX ranges bracket the launches of each stage (host side, like the reference's range around its kernel) */
    if (!skip_sync) {
        nvtxRangePushA("gjd stage 1: entropy sync");
        if (b.entropy_mask & 1u) { k_huff_sync<false, false><<<b.n_seq, GJD_SYNC_THREADS, 0, s>>>(b); n++; }
        if (b.entropy_mask & 2u) { k_huff_sync<true, false><<<b.n_seq, GJD_SYNC_THREADS, 0, s>>>(b); n++; }
        if (b.entropy_mask & 4u) { k_huff_sync<false, true><<<b.n_seq, GJD_SYNC_THREADS, 0, s>>>(b); n++; }
        if (b.entropy_mask & 8u) { k_huff_sync<true, true><<<b.n_seq, GJD_SYNC_THREADS, 0, s>>>(b); n++; }
        k_huff_fix<<<(b.n_seq + 127) / 128, 128, 0, s>>>(b); n++;
        nvtxRangePop();
    }
    if (ev) GJD_CK(cudaEventRecord(ev->e[2], s));
    nvtxRangePushA("gjd stage 2: count scan + coefficient write + DC bases");
    k_seq_scan<<<b.n_seg, 256, 0, s>>>(b); n++;
    if (!attrs_set || !*attrs_set) {   /* opt in to > 48 KB of dynamic shared memory; the caller's gjd_ctx remembers it */
        GJD_CK(cudaFuncSetAttribute(k_huff_write<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GJD_WRITE_SMEM));
        GJD_CK(cudaFuncSetAttribute(k_huff_write<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GJD_WRITE_SMEM));
        GJD_CK(cudaFuncSetAttribute(k_huff_write<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GJD_WRITE_SMEM));
        GJD_CK(cudaFuncSetAttribute(k_huff_write<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GJD_WRITE_SMEM));
        if (attrs_set) *attrs_set = true;
    }
    if (b.entropy_mask & 1u) { k_huff_write<false, false><<<b.n_seq, GJD_WRITE_THREADS, GJD_WRITE_SMEM, s>>>(b); n++; }
    if (b.entropy_mask & 2u) { k_huff_write<true, false><<<b.n_seq, GJD_WRITE_THREADS, GJD_WRITE_SMEM, s>>>(b); n++; }
    if (b.entropy_mask & 4u) { k_huff_write<false, true><<<b.n_seq, GJD_WRITE_THREADS, GJD_WRITE_SMEM, s>>>(b); n++; }
    if (b.entropy_mask & 8u) { k_huff_write<true, true><<<b.n_seq, GJD_WRITE_THREADS, GJD_WRITE_SMEM, s>>>(b); n++; }
    k_dc_tilesum<<<(b.n_tile + 255u) / 256u, 256, 0, s>>>(b); n++;
    k_dc_tilebase<<<b.n_img, 256, 0, s>>>(b); n++;
    nvtxRangePop();
    if (ev) GJD_CK(cudaEventRecord(ev->e[3], s));
    nvtxRangePushA("gjd stage 3: dequant + IDCT + colour");
    if (transform_variant == 2) {   /* the cost-sorted kernel (measured: 7.32 ms against 7.07 ms on C3, profiles/r02g_transform_ab.txt) */
        if (b.any_interleaved) k_transform2<8, true><<<b.n_tb, GJD_TRANSFORM_THREADS, 0, s>>>(b);
        else k_transform2<8, false><<<b.n_tb, GJD_TRANSFORM_THREADS, 0, s>>>(b);
    } else if (b.any_interleaved) k_transform<false, 8, true><<<b.n_tb, GJD_TRANSFORM_THREADS, 0, s>>>(b, nullptr);
    else k_transform<false, GJD_TRANSFORM_MINB, false><<<b.n_tb, GJD_TRANSFORM_THREADS, 0, s>>>(b, nullptr);
    n++;
    if (b.any_gray) { k_transform_gray<false><<<b.n_tb, GJD_TRANSFORM_THREADS, 0, s>>>(b, nullptr); n++; }
    if (b.any_sub) {
        k_dc_sub_tiles<<<(b.n_tile + GJD_DCSUB_TILES - 1u) / GJD_DCSUB_TILES, GJD_TRANSFORM_THREADS, 0, s>>>(b, b.dcabs); n++;
        k_transform_sub<false><<<b.n_tb, GJD_TRANSFORM_THREADS, 0, s>>>(b, b.dcabs, nullptr); n++;
    }
    nvtxRangePop();
    if (ev) GJD_CK(cudaEventRecord(ev->e[4], s));
    if (launches) *launches = n;
    return cudaGetLastError();
}

cudaError_t gjd_launch_export_coefs(const GjdDevBatch& b, int16_t* out, cudaStream_t s) {
    /* single-image batch: out is indexed by the image's own MCU numbers */
    if (b.any_gray) k_transform_gray<true><<<b.n_tb, GJD_TRANSFORM_THREADS, 0, s>>>(b, out);
    else if (b.any_sub) k_transform_sub<true><<<b.n_tb, GJD_TRANSFORM_THREADS, 0, s>>>(b, b.dcabs, out);
    else k_transform<true, 6, false><<<b.n_tb, GJD_TRANSFORM_THREADS, 0, s>>>(b, out);
    return cudaGetLastError();
}

cudaError_t gjd_launch_test_color(const int16_t* y, const int16_t* cb, const int16_t* cr, size_t n,
                                  uint8_t* r, uint8_t* g, uint8_t* bl, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    k_test_color<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(y, cb, cr, n, r, g, bl);
    return cudaGetLastError();
}

cudaError_t gjd_launch_test_idct(const int16_t* coef, size_t nblocks, const int32_t* quant, int16_t* out,
                                 cudaStream_t s) {
    if (nblocks == 0) return cudaSuccess;
    k_test_idct<<<(unsigned)((nblocks + GJD_TRANSFORM_THREADS - 1) / GJD_TRANSFORM_THREADS), GJD_TRANSFORM_THREADS, 0, s>>>(coef, nblocks, quant, out);
    return cudaGetLastError();
}

/* ------------------------------------------------------------------ no-allocation entry point ---- */
/*
 * gjd_idct_blocks_static (include/gjd.h): what gjd_test_idct does, for callers that cannot call cudaMalloc.
 * gjd_test_idct keeps its device-side scratch (the staged coefficients, the expanded quantisation table, the staged
 * samples) in three stream-ordered allocations; here that scratch is one file-scope device array, so the call allocates
 * nothing.  Every GPU that loads libgjd.so holds the array until the process exits.
 *
 *   [0, 256)                     header; word 0 = blocks the kernel finished (cleared by every call, checked after it)
 *   [256, 512)                   the quantisation table, 64 int32 in zigzag order
 *   [512, 512 + MAX*128)         coefficients, int16 zigzag, 128 bytes per block
 *   [512 + MAX*128, +MAX*128)    samples, int16 natural order
 *
 * The array keeps the last call's contents; the one thing that must start at zero (the header word) is cleared per call.
 * All calls share it: two calls that overlap on the GPU would race, so the entry point holds a process-wide lock and is
 * synchronous on return.  In host code g_idct_static is only the host-side stand-in of the symbol — the device address
 * comes from cudaGetSymbolAddress, once per device, with that device current.
 */
#define GJD_STATIC_HDR   256u
#define GJD_STATIC_QOFF  256u
#define GJD_STATIC_COEF  512u
#define GJD_STATIC_BYTES (GJD_STATIC_COEF + 2u * (size_t)GJD_STATIC_IDCT_MAX_BLOCKS * 128u)
#define GJD_STATIC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_idct_static[GJD_STATIC_BYTES];

/* k_test_idct with the quantisation table and the finished-blocks word in the static array (named directly: device code
 * may); coef and out are addresses inside the same array, handed in by the host from cudaGetSymbolAddress */
__global__ void __launch_bounds__(GJD_TRANSFORM_THREADS) k_idct_blocks_static(const int16_t* __restrict__ coef, uint32_t nblocks,
                                                                               int16_t* __restrict__ out) {
    __shared__ uint32_t sq[32];
    __shared__ uint32_t finished;
    if (threadIdx.x < 32) sq[threadIdx.x] = gjd_pack_quant(reinterpret_cast<const int*>(g_idct_static + GJD_STATIC_QOFF), threadIdx.x);
    if (threadIdx.x == 0) finished = 0u;
    __syncthreads();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nblocks) {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(coef + (size_t)i * 64);
        uint32_t w[32];
#pragma unroll
        for (int j = 0; j < 32; j++) w[j] = src[j];
        int v[64];
        GjdUnpack<63, 64>::run(w, sq, v);
        v[0] = (int)(int16_t)(w[0] & 0xffffu) * (int)(sq[0] & 0xffu);
        gjd_idct8x8(v);
#pragma unroll
        for (int j = 0; j < 64; j++) out[(size_t)i * 64 + j] = (int16_t)(v[j] - 256);
        atomicAdd(&finished, 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0 && finished) atomicAdd(reinterpret_cast<uint32_t*>(g_idct_static), finished);
}

#include <cstdio>
#include <mutex>

namespace {
std::mutex g_static_mu;                                        /* one call at a time: the array is shared */
unsigned char* g_static_dev[GJD_STATIC_MAX_DEVICES] = {};      /* device address of g_idct_static, per device */

int static_fail(char* errbuf, size_t errbuf_len, int rc, const char* what, cudaError_t e) {
    if (errbuf && errbuf_len)
        snprintf(errbuf, errbuf_len, "gjd_idct_blocks_static: %s%s%s", what, e != cudaSuccess ? ": " : "",
                 e != cudaSuccess ? cudaGetErrorString(e) : "");
    return rc;
}
}  /* namespace */

extern "C" int gjd_idct_blocks_static(int cuda_device, const int16_t* coef_zz, size_t nblocks, const uint8_t quant_zz[64],
                                      int16_t* samples, char* errbuf, size_t errbuf_len) {
    if (errbuf && errbuf_len) errbuf[0] = 0;
    if (!coef_zz || !quant_zz || !samples) return static_fail(errbuf, errbuf_len, GJD_E_ARG, "NULL argument", cudaSuccess);
    if (cuda_device < 0 || cuda_device >= GJD_STATIC_MAX_DEVICES)
        return static_fail(errbuf, errbuf_len, GJD_E_ARG, "device index out of range", cudaSuccess);
    if (nblocks > (size_t)GJD_STATIC_IDCT_MAX_BLOCKS) {
        if (errbuf && errbuf_len)
            snprintf(errbuf, errbuf_len, "gjd_idct_blocks_static: %zu blocks exceed the static scratch, GJD_STATIC_IDCT_MAX_BLOCKS = %d "
                     "(split the call, or use gjd_test_idct, which allocates)", nblocks, (int)GJD_STATIC_IDCT_MAX_BLOCKS);
        return GJD_E_ARG;
    }
    if (nblocks == 0) return GJD_OK;
    std::lock_guard<std::mutex> lock(g_static_mu);
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e != cudaSuccess) return static_fail(errbuf, errbuf_len, GJD_E_NODEVICE, "no usable CUDA device", e);
    if (prev != cuda_device && (e = cudaSetDevice(cuda_device)) != cudaSuccess)
        return static_fail(errbuf, errbuf_len, GJD_E_NODEVICE, "cudaSetDevice", e);
    int rc = GJD_OK;
    const char* what = "";
    do {
        unsigned char*& base = g_static_dev[cuda_device];
        if (!base) {   /* set-up, once per device, with that device current */
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_idct_static);
            if (e != cudaSuccess || !p) { what = "cudaGetSymbolAddress"; break; }
            base = static_cast<unsigned char*>(p);
        }
        int16_t* d_coef = reinterpret_cast<int16_t*>(base + GJD_STATIC_COEF);
        int16_t* d_out = d_coef + (size_t)GJD_STATIC_IDCT_MAX_BLOCKS * 64;
        int32_t hq[64];
        for (int i = 0; i < 64; i++) hq[i] = quant_zz[i];
        /* everything on the legacy default stream, synchronous copies: no stream, event or buffer is created */
        what = "clear header";
        if ((e = cudaMemset(base, 0, GJD_STATIC_HDR)) != cudaSuccess) break;
        what = "copy quantisation table";
        if ((e = cudaMemcpy(base + GJD_STATIC_QOFF, hq, sizeof hq, cudaMemcpyHostToDevice)) != cudaSuccess) break;
        what = "copy coefficients";
        if ((e = cudaMemcpy(d_coef, coef_zz, nblocks * 128, cudaMemcpyDefault)) != cudaSuccess) break;
        what = "launch";
        k_idct_blocks_static<<<(unsigned)((nblocks + GJD_TRANSFORM_THREADS - 1) / GJD_TRANSFORM_THREADS), GJD_TRANSFORM_THREADS>>>(
            d_coef, (uint32_t)nblocks, d_out);
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        what = "copy samples";
        if ((e = cudaMemcpy(samples, d_out, nblocks * 128, cudaMemcpyDefault)) != cudaSuccess) break;
        what = "read header";
        uint32_t finished = 0;
        if ((e = cudaMemcpy(&finished, base, sizeof finished, cudaMemcpyDeviceToHost)) != cudaSuccess) break;
        if (finished != (uint32_t)nblocks) { rc = static_fail(errbuf, errbuf_len, GJD_E_CUDA, "the kernel did not finish every block", cudaSuccess); }
    } while (0);
    if (e != cudaSuccess) rc = static_fail(errbuf, errbuf_len, GJD_E_CUDA, what, e);
    if (prev != cuda_device && prev >= 0) cudaSetDevice(prev);
    return rc;
}
