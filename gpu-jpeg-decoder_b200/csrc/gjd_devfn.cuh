/*
 * gjd_devfn.cuh — the arithmetic of the hot path as __host__ __device__ inline functions.
 *
 * The CUDA kernels (gjd_kernels.cu) are built from these.  They also compile as plain C++
 * (GJD_HD expands to nothing without __CUDACC__) so that tests/hostsim can single-step the very
 * same functions on the build machine, which has no GPU.  The host build is test scaffolding
 * only: libgjd.so has no CPU decode path.
 *
 * Semantics follow the reference's CPU decoder (paths relative to the reference root):
 *   entropy decode  cpp-decoder/src/parser.cpp:105-134, utils/stream.cpp:8-22,44-51
 *   dezigzag        cpp-decoder/src/idct.cpp:7-32
 *   IDCT            cpp-decoder/src/idct.cpp:34-133, constants idct.h:4-9
 *   colour          cpp-decoder/utils/color.cpp:3-18
 */
#ifndef GJD_DEVFN_CUH
#define GJD_DEVFN_CUH

#include <stdint.h>
#include "gjd_types.h"

#ifndef GJD_WRITE_SYMS
#define GJD_WRITE_SYMS 2          /* symbols per trip of the write pass's warp-voted loop */
#endif

#ifndef GJD_DBG_SLOW
#define GJD_DBG_SLOW() ((void)0)  /* ... and the symbols that leave the LUTs for the canonical tables */
#endif
#ifndef GJD_DBG_SYM
#define GJD_DBG_SYM() ((void)0)   /* test harnesses may count decoded symbols */
#endif

#ifdef __CUDACC__
#define GJD_HD __host__ __device__ __forceinline__
#define GJD_HD_COLD __host__ __device__ __noinline__   /* rare paths: kept out of line (code size) */
#else
#define GJD_HD inline
#define GJD_HD_COLD inline
#endif

/* ------------------------------------------------------------------ Huffman tables ---- */

/* Pack one decoded symbol into the LUT entry format of gjd_types.h. */
GJD_HD uint32_t gjd_make_entry(uint32_t code_len, uint32_t sym, uint32_t isdc) {
    uint32_t size = sym & 15u;
    uint32_t adv = isdc ? 1u : ((sym == 0u) ? GJD_ADV_EOB : (sym >> 4) + 1u);
    return (code_len + size) | (size << 5) | (adv << 9);
}
#define GJD_ENTRY_BAD 0x10000u   /* bit 16 (slow path only): no code assigned to these bits */
/* what the decoders do with an unassigned code: consume 16 bits, no value, advance one slot — so that
 * "zigzag index 0" and "the ring is at a DC table" stay equivalent even on garbage */
#define GJD_ENTRY_UNASSIGNED (16u | (1u << 9))

/* Canonical code tables from BITS[16]: lim[L] = first unused code of length L left-aligned to
 * 16 bits (non-decreasing in L), off[L] = valptr[L] - mincode[L].
 * (The reference's leftmost-free-leaf insertion, huffmanTree.cpp:34-68, yields this assignment.) */
GJD_HD void gjd_canonical(const uint8_t bits[16], uint32_t lim[17], int32_t off[17]) {
    uint32_t code = 0, k = 0;
    lim[0] = 0; off[0] = 0;
    for (int L = 1; L <= 16; L++) {
        off[L] = (int32_t)k - (int32_t)code;
        code += bits[L - 1];
        k += bits[L - 1];
        lim[L] = code << (16 - L);
        code <<= 1;
    }
}

/* Resolve a 16-bit window against the canonical tables, lengths from..to (16 unless fewer bits are known). */
GJD_HD uint32_t gjd_resolve(const uint32_t* lim, const int32_t* off, const uint8_t* vals,
                            uint32_t isdc, uint32_t win16, int from, int to = 16) {
    for (int L = from; L <= to; L++) {
        if (win16 < lim[L]) {
            uint32_t idx = (uint32_t)(off[L] + (int32_t)(win16 >> (16 - L))) & 255u;
            return gjd_make_entry((uint32_t)L, vals[idx], isdc);
        }
    }
    /* unassigned code: consume 16 bits, advance one slot; flagged for the write pass */
    return 16u | (1u << 9) | GJD_ENTRY_BAD;
}

/* Primary LUT entry for index i (GJD_LUT_BITS leading bits).  Codes longer than GJD_LUT_BITS occupy the
 * top of the code space (lim[] is non-decreasing), i.e. the prefixes >= lim[GJD_LUT_BITS] >> 6: the first
 * GJD_LUT2_PREFIXES of them link to a 64-entry sub-table of the secondary LUT ((rank + 1) << 5, low five
 * bits 0); 0 = not resolvable from tables (gjd_lookup_slow). */
GJD_HD uint16_t gjd_lut_entry(const uint32_t* lim, const int32_t* off, const uint8_t* vals,
                              uint32_t isdc, uint32_t i) {
    uint32_t win16 = i << (16 - GJD_LUT_BITS);
    for (int L = 1; L <= GJD_LUT_BITS; L++) {
        if (win16 < lim[L]) {
            uint32_t idx = (uint32_t)(off[L] + (int32_t)(win16 >> (16 - L))) & 255u;
            return (uint16_t)gjd_make_entry((uint32_t)L, vals[idx], isdc);
        }
    }
    uint32_t rank = i - (lim[GJD_LUT_BITS] >> (16 - GJD_LUT_BITS));
    return rank < (uint32_t)GJD_LUT2_PREFIXES ? (uint16_t)((rank + 1u) << 5) : (uint16_t)0;
}

/* Secondary LUT entry i2 = rank * 64 + (the 6 bits after the primary index): GJD_LUT_BITS + 6 bits of the
 * window are known (all 16 with a 10-bit index, so every assigned code resolves);
 * 0 = longer or unassigned code (left to gjd_lookup_slow, which flags the latter). */
GJD_HD uint16_t gjd_lut2_entry(const uint32_t* lim, const int32_t* off, const uint8_t* vals,
                               uint32_t isdc, uint32_t i2) {
    uint32_t j = (lim[GJD_LUT_BITS] >> (16 - GJD_LUT_BITS)) + (i2 >> 6);
    if (j >= (uint32_t)GJD_LUT_SIZE) return 0;
    uint32_t e = gjd_resolve(lim, off, vals, isdc, ((j << 6) | (i2 & 63u)) << (10 - GJD_LUT_BITS), GJD_LUT_BITS + 1,
                             GJD_LUT_BITS + 6);
    return (e & GJD_ENTRY_BAD) ? (uint16_t)0 : (uint16_t)e;
}

/* Count-pass LUT entry for index i (GjdLutCount): the single-symbol entry without its size field, or — for an
 * AC table, when the first symbol is not EOB and a second symbol of the same table fits completely into the
 * remaining index bits — the pair entry. */
GJD_HD uint16_t gjd_lutc_entry(const uint32_t* lim, const int32_t* off, const uint8_t* vals,
                               uint32_t isdc, uint32_t i) {
    const uint32_t e1 = gjd_lut_entry(lim, off, vals, isdc, i);
    if ((e1 & 31u) == 0u) return (uint16_t)e1;
    const uint32_t tot1 = e1 & 31u, adv1 = e1 >> 9;
    const uint32_t single = tot1 | (adv1 << 9);
    if (isdc || adv1 > 8u || tot1 >= (uint32_t)GJD_LUT_BITS) return (uint16_t)single;   /* EOB (64) and runs > 7: no pair */
    const uint32_t e2 = gjd_lut_entry(lim, off, vals, 0u, (i << tot1) & (uint32_t)(GJD_LUT_SIZE - 1));
    const uint32_t tot2 = e2 & 31u;
    if (tot2 == 0u || tot1 + tot2 > (uint32_t)GJD_LUT_BITS) return (uint16_t)single;
    return (uint16_t)((tot1 + tot2) | GJD_PAIR_FLAG | ((adv1 - 1u) << 6) | ((adv1 + (e2 >> 9)) << 9));
}

/* DHT slot behind ring position r (0..5): even = DC table, odd = AC table of component r/2. */
GJD_HD uint32_t gjd_ring_slot(const uint8_t tabsel[8], uint32_t r) {
    return (uint32_t)tabsel[(r & 1u) * 4u + (r >> 1)] & 3u;
}

/* ------------------------------------------------------------------ byte unstuffing --- */
/*
 * On-device FF00 unstuffing with marker detection.  The byte loop it replaces
 * (cpp-decoder/src/parser.cpp:78-99, cuda-decoder/src/parser.cu:446-464, host twin gjd_unstuff_segments) walks the scan
 * once, serially; what it does to a byte depends only on the byte and its two neighbours:
 *     00 after FF                 -> dropped (stuffing)
 *     FF before FF                -> dropped (fill byte, T.81 B.1.1.2)
 *     FF before D9                -> the scan ends here (EOI)
 *     FF before D0..D7            -> restart marker
 *     FF before anything else     -> a marker that has no business inside a scan
 * so 16 bytes per thread can be classified independently.  Masks have bit i = byte i of the 16 (memory order).
 */
struct GjdUnstuffMasks {
    uint32_t drop;   /* stuffed zeros, fill FFs and (scans with restart intervals) both bytes of every RSTn   */
    uint32_t mark;   /* FFs followed by a byte that is neither 00 nor FF: EOI or a stray marker (RSTn: only in
                        scans without restart intervals, where it is an error)                                */
    uint32_t rst;    /* scans with restart intervals: FFs that start an RSTn marker                            */
};

/* bit 7 of every byte of w that equals 0xFF / 0x00, gathered into a 4-bit mask (bit i = byte i) */
GJD_HD uint32_t gjd_bytes_ff4(uint32_t w) {
    const uint32_t m = (((w & 0x7f7f7f7fu) + 0x01010101u) & w & 0x80808080u) >> 7;
    return ((m * 0x01020408u) >> 24) & 15u;
}
GJD_HD uint32_t gjd_bytes_004(uint32_t w) {
    const uint32_t m = (~(((w & 0x7f7f7f7fu) + 0x7f7f7f7fu) | w) & 0x80808080u) >> 7;
    return ((m * 0x01020408u) >> 24) & 15u;
}

/* w[4] = 16 bytes in memory order (little-endian words); valid = mask of the bytes that belong to the scan;
 * prev_ff = the byte before them is a scan byte equal to FF; next = the scan byte after them, next_valid = it exists
 * (an FF that is the last byte of the data is no marker: gjd_unstuff_segments stops there with "no EOI"). */
GJD_HD GjdUnstuffMasks gjd_unstuff_classify(const uint32_t w[4], uint32_t valid, uint32_t prev_ff, uint32_t next,
                                            uint32_t next_valid, bool allow_rst = false) {
    uint32_t ff = 0, z = 0, dn = 0;
    for (int j = 0; j < 4; j++) {
        ff |= gjd_bytes_ff4(w[j]) << (4 * j);
        z |= gjd_bytes_004(w[j]) << (4 * j);
        dn |= gjd_bytes_004((w[j] ^ 0xd0d0d0d0u) & 0xf8f8f8f8u) << (4 * j);      /* bytes D0..D7 */
    }
    ff &= valid;
    z &= valid;
    dn &= valid;
    const uint32_t nv = next_valid ? 1u : 0u;
    const uint32_t ff_next = (ff >> 1) | ((nv && next == 0xffu ? 1u : 0u) << 15);   /* byte i+1 is FF */
    const uint32_t z_next = (z >> 1) | ((nv && next == 0x00u ? 1u : 0u) << 15);     /* byte i+1 is 00 */
    const uint32_t has_next = (valid >> 1) | (nv << 15);                            /* byte i+1 exists */
    const uint32_t ff_prev = ((ff << 1) | (prev_ff ? 1u : 0u)) & 0xffffu;           /* byte i-1 is FF */
    GjdUnstuffMasks m;
    m.drop = (z & ff_prev) | (ff & ff_next);
    m.mark = ff & ~ff_next & ~z_next & has_next;
    m.rst = 0;
    if (allow_rst) {
        const uint32_t dn_next = (dn >> 1) | ((nv && (next & 0xf8u) == 0xd0u ? 1u : 0u) << 15);   /* byte i+1 is D0..D7 */
        m.rst = m.mark & dn_next;
        m.mark &= ~dn_next;
        m.drop |= m.rst | (dn & ff_prev);
    }
    return m;
}

/* One thread's share of a raw scan: 16 bytes at arena offset a (a multiple of 16).  raw_off / raw_len: where the scan's
 * first byte lies in the arena and how many bytes follow it to the end of the file.  limit: positions >= limit are
 * outside the scan for this pass (raw_len in the counting pass; the EOI position once it is known). */
struct GjdUnstuffLane {
    uint32_t w[4];      /* the 16 raw bytes */
    uint32_t valid;     /* those of them that are scan bytes at a position < limit */
    uint32_t pos0;      /* position in the scan of byte 0 (wraps for the bytes in front of the scan) */
    uint32_t next;      /* the byte after the 16 (0 when the file ends there) */
    GjdUnstuffMasks m;
};

#ifdef __CUDA_ARCH__
#define GJD_LDG8(p) ((uint32_t)__ldg(p))
#else
#define GJD_LDG8(p) ((uint32_t)*(p))
#endif

GJD_HD GjdUnstuffLane gjd_unstuff_lane_at(const uint8_t* raw, uint64_t raw_off, uint32_t raw_len, uint64_t a, uint32_t limit,
                                          bool allow_rst = false) {
    GjdUnstuffLane L;
    const int64_t p0 = (int64_t)a - (int64_t)raw_off;                                 /* position in the scan of byte 0 */
    L.pos0 = (uint32_t)p0;
    L.valid = 0;
    L.next = 0;
    L.w[0] = L.w[1] = L.w[2] = L.w[3] = 0;
    L.m.drop = L.m.mark = L.m.rst = 0;
    if (p0 >= (int64_t)limit || p0 + 16 <= 0) return L;
#ifdef __CUDA_ARCH__
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(raw + a));
    L.w[0] = v.x; L.w[1] = v.y; L.w[2] = v.z; L.w[3] = v.w;
#else
    for (int j = 0; j < 4; j++)
        L.w[j] = (uint32_t)raw[a + 4 * j] | ((uint32_t)raw[a + 4 * j + 1] << 8) | ((uint32_t)raw[a + 4 * j + 2] << 16) |
                 ((uint32_t)raw[a + 4 * j + 3] << 24);
#endif
    const uint32_t lo = p0 < 0 ? (uint32_t)(-p0) : 0u;                                /* first byte of the scan among the 16 */
    const uint32_t hi = (int64_t)limit - p0 < 16 ? (uint32_t)((int64_t)limit - p0) : 16u;
    L.valid = (0xffffu >> (16u - hi)) & (0xffffu << lo) & 0xffffu;
    const uint32_t prev_ff = (p0 > 0 && GJD_LDG8(raw + a - 1) == 0xffu) ? 1u : 0u;
    /* neighbours are bytes of the FILE, also at or past `limit` (the D9 of the EOI itself lies there) */
    const uint32_t next_valid = p0 + 16 < (int64_t)raw_len ? 1u : 0u;
    L.next = next_valid ? GJD_LDG8(raw + a + 16) : 0u;
    const uint32_t hi_file = (int64_t)raw_len - p0 < 16 ? (uint32_t)((int64_t)raw_len - p0) : 16u;
    const uint32_t valid_file = (0xffffu >> (16u - hi_file)) & (0xffffu << lo) & 0xffffu;
    L.m = gjd_unstuff_classify(L.w, valid_file, prev_ff, L.next, next_valid, allow_rst);
    L.m.drop &= L.valid;
    L.m.mark &= L.valid;
    L.m.rst &= L.valid;
    return L;
}

GJD_HD uint32_t gjd_lane_byte(const GjdUnstuffLane& L, uint32_t i) {
    return (L.w[i >> 2] >> (8u * (i & 3u))) & 0xffu;
}

/* the raw_err word of a marker with second byte nb at scan position pos (not an EOI) */
GJD_HD uint32_t gjd_unstuff_err_word(uint32_t pos, uint32_t nb) {
    return (pos << 2) | ((nb >= 0xd0u && nb <= 0xd7u) ? GJD_UERR_UNSUPPORTED : GJD_UERR_FORMAT);
}

/* ------------------------------------------------------------------ bit reader -------- */

GJD_HD uint32_t gjd_bswap32(uint32_t x) {
#ifdef __CUDA_ARCH__
    return __byte_perm(x, 0, 0x0123);
#else
    return (x >> 24) | ((x >> 8) & 0xff00u) | ((x << 8) & 0xff0000u) | (x << 24);
#endif
}

/* Two cached big-endian words of the scan (wi = index of w0) plus the word after them, still in memory
 * byte order: it is loaded one refill ahead of its use so that the load latency hides behind the symbols
 * decoded in between (the byte swap waits for the data, so it is deferred to the next refill). */
#ifndef GJD_WIN_AHEAD
#define GJD_WIN_AHEAD 1            /* refills between the load of a word and its first use; 2 was measured: write 6.30 -> 6.92 ms (profiles/r02l_win_ahead_ab.txt) */
#endif
struct GjdWin {
    uint32_t wi, w0, w1, w2raw;
#if GJD_WIN_AHEAD > 1
    uint32_t w3raw;                /* the word after w2raw: loaded two refills (64 bits of symbols) ahead of its use */
#endif
};

/* Where the words of the scan come from.  The decoders are templates over the source: a plain pointer (global memory on
 * the device, the test harness's buffer on the host), or GjdScanRing, the TMA-staged variant of the A/B build
 * -DGJD_SCAN_TMA=1 (profiles/r02_tma_scan_ab.txt). */
GJD_HD uint32_t gjd_scan_word(const uint32_t*& scan, uint32_t i) { return scan[i]; }

#if defined(__CUDACC__) && defined(GJD_SCAN_TMA)
/* Per-thread double-buffered ring in shared memory, filled by 1-D TMA bulk copies (cp.async.bulk ... mbarrier::
 * complete_tx::bytes; SASS UBLKCP) of GJD_RING_CHUNK bytes: BASELINE.json's "compressed scans staged into shared memory by
 * TMA bulk copies".  Every thread walks its own subsequence, so every thread owns two chunks and two mbarriers; the copy
 * of chunk c+1 is issued when the reader enters chunk c (the words of chunk c-1 are all in registers by then: GjdWin
 * holds three), so it has a whole chunk's worth of symbols to land. */
#ifndef GJD_RING_CHUNK
#define GJD_RING_CHUNK 32u
#endif
struct GjdScanRing {
    const uint8_t* arena;     /* scan arena (256-byte aligned) */
    uint64_t base;            /* byte offset of the segment's first scan byte in the arena (4-byte aligned) */
    uint32_t slots;           /* shared address of this thread's two chunks (slot s at slots + s * GJD_RING_CHUNK * blockDim.x) */
    uint32_t bars;            /* shared address of this thread's two mbarriers (8 bytes each, adjacent) */
    uint32_t stride;          /* bytes between the two slots */
    uint32_t cur;             /* chunk the reader is in (arena byte offset / GJD_RING_CHUNK), 0xffffffff = none */
    uint32_t phase;           /* bit s = parity the next wait on slot s uses */
    uint32_t inflight;        /* bit s = a copy into slot s has been issued and not waited for */
    __device__ __forceinline__ void issue(uint32_t chunk) {
        const uint32_t s = chunk & 1u;
        const uint32_t bar = bars + 8u * s, dst = slots + s * stride;
        const uint8_t* src = arena + (uint64_t)chunk * GJD_RING_CHUNK;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(GJD_RING_CHUNK) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(src), "r"(GJD_RING_CHUNK), "r"(bar) : "memory");
        inflight |= 1u << s;
    }
    __device__ __forceinline__ void wait(uint32_t s) {
        const uint32_t bar = bars + 8u * s, par = (phase >> s) & 1u;
        uint32_t done;
        do {
            asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p;}"
                         : "=r"(done) : "r"(bar), "r"(par) : "memory");
        } while (!done);
        phase ^= 1u << s;
        inflight &= ~(1u << s);
    }
    __device__ __forceinline__ uint32_t word(uint32_t i) {
        const uint64_t a = base + 4ull * i;
        const uint32_t chunk = (uint32_t)(a / GJD_RING_CHUNK);
        if (chunk != cur) {
            if (chunk != cur + 1u || !((inflight >> (chunk & 1u)) & 1u)) {
                /* a jump (start of a span): drain what is in flight, fetch this chunk */
                if (inflight & 1u) wait(0u);
                if (inflight & 2u) wait(1u);
                issue(chunk);
            }
            wait(chunk & 1u);
            cur = chunk;
            issue(chunk + 1u);
        }
        uint32_t v;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(slots + (chunk & 1u) * stride + (uint32_t)(a % GJD_RING_CHUNK)));
        return v;
    }
    /* leave nothing in flight (before the thread block's shared memory goes away) */
    __device__ __forceinline__ void drain() {
        if (inflight & 1u) wait(0u);
        if (inflight & 2u) wait(1u);
    }
};
__device__ __forceinline__ uint32_t gjd_scan_word(GjdScanRing& r, uint32_t i) { return r.word(i); }
#endif

template <class Src>
GJD_HD void gjd_win_init(GjdWin& c, Src& scan, uint32_t p) {
    c.wi = p >> 5;
    c.w0 = gjd_bswap32(gjd_scan_word(scan, c.wi));
    c.w1 = gjd_bswap32(gjd_scan_word(scan, c.wi + 1));
    c.w2raw = gjd_scan_word(scan, c.wi + 2);
#if GJD_WIN_AHEAD > 1
    c.w3raw = gjd_scan_word(scan, c.wi + 3);
#endif
}

/* 32 bits of the scan starting at bit p (MSB first).  p may advance by at most 32 bits per call.
 * Reads up to 4 * (GJD_WIN_AHEAD + 2) bytes past the start of the word that holds p (the scan arena is padded). */
template <class Src>
GJD_HD uint32_t gjd_peek32(GjdWin& c, Src& scan, uint32_t p) {
    uint32_t wi = p >> 5;
    if (wi != c.wi) {
        c.w0 = c.w1;
        c.w1 = gjd_bswap32(c.w2raw);
#if GJD_WIN_AHEAD > 1
        c.w2raw = c.w3raw;
        c.w3raw = gjd_scan_word(scan, wi + 3);
#else
        c.w2raw = gjd_scan_word(scan, wi + 2);
#endif
        c.wi = wi;
    }
#ifdef __CUDA_ARCH__
    return __funnelshift_l(c.w1, c.w0, p);   /* shift amount taken mod 32 */
#else
    uint32_t sh = p & 31u;
    return sh ? (c.w0 << sh) | (c.w1 >> (32u - sh)) : c.w0;
#endif
}

/* ------------------------------------------------------------------ entropy decode ---- */

/* Decoder state between symbols as stored in sinfo ({p, n << GJD_POS_BITS | pos}): p = bit position in the image's unstuffed
 * scan, n = coefficient slots advanced so far, pos = block-in-MCU * 64 + zigzag index of the next slot (0..191 for
 * 4:4:4, where the block in the MCU is the component; up to 383 for 4:2:0).
 * This is the (p, n, c, z) state of the reference's decodeSequence (cuda-decoder/src/parser.cu:50-121). */
struct GjdState {
    uint32_t p, n, pos;
};
#define GJD_POS_BITS 9u
#define GJD_POS_MASK 511u

/* Inside the loops the state is (p, q, tb): q = slot counter with q & 63 = zigzag index (starts at
 * pos, never wraps), tb = byte offset of the current table in the LUT ring. */
GJD_HD uint32_t gjd_ring_of_pos(uint32_t pos) {
    return ((pos >> 6) * 2u + ((pos & 63u) ? 1u : 0u)) * GJD_RING_STRIDE;
}

/* JPEG EXTEND of the `size` value bits that follow a code of `len` bits in the 32-bit window w
 * (Stream::decodeNumber, cpp-decoder/utils/stream.cpp:44-51). */
GJD_HD int gjd_extend(uint32_t w, uint32_t len, uint32_t size) {
#ifdef __CUDA_ARCH__
    /* branch-free; shifts by 32 give 0 on the GPU, so size == 0 yields 0 */
    const int t = (int)(w << len);                       /* value bits left-aligned */
    const uint32_t vb = (uint32_t)t >> (32u - size);
    const int neg = ~(t >> 31);                          /* all ones when the leading value bit is 0 */
    return (int)vb + (neg & (int)(1u - (1u << size)));
#else
    if (size == 0u) return 0;
    uint32_t vb = (w << len) >> (32u - size);
    return (vb >> (size - 1u)) ? (int)vb : (int)vb - (int)((1u << size) - 1u);
#endif
}

/* Where the decoders read the LUT ring from: a GjdLutSet in generic (global or host) memory ... */
struct GjdLutMem {
    const GjdLutSet* set;
    uint32_t* badcode;      /* set to 1 when an unassigned code is met (write pass only; may be null) */
    GJD_HD uint32_t entry(uint32_t off) const {
        return *reinterpret_cast<const uint16_t*>(reinterpret_cast<const char*>(set->lut) + off);
    }
};
#ifdef __CUDACC__
/* ... or a copy in shared memory, addressed by its 32-bit shared-window offset (one LDS, no generic
 * address arithmetic in the loop). */
struct GjdLutSmem {
    const GjdLutSet* set;
    uint32_t* badcode;
    uint32_t base;
    /* cold: where the slow path finds the canonical tables when s holds only the LUT levels */
    __device__ __forceinline__ explicit GjdLutSmem(const GjdLutSet* s, uint32_t* bc = nullptr, const GjdLutSet* cold = nullptr)
        : set(cold ? cold : s), badcode(bc) {
        /* opaque move: keeps the window offset in a register instead of re-deriving it per symbol */
        asm volatile("mov.u32 %0, %1;" : "=r"(base) : "r"((uint32_t)__cvta_generic_to_shared(s->lut)));
    }
    __device__ __forceinline__ uint32_t entry(uint32_t off) const {
        uint16_t v;
        asm("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(base + off));
        return v;
    }
};
#endif

/* Rare path: the code is longer than GJD_LUT_BITS, or unassigned.  An unassigned code comes back as
 * GJD_ENTRY_UNASSIGNED: the write pass flags it, and both passes treat it alike. */
GJD_HD_COLD uint32_t gjd_lookup_slow(const GjdLutSet* lut, uint32_t tb, uint32_t w, uint32_t room, uint32_t* badcode) {
    GJD_DBG_SLOW();
    uint32_t slot = lut->slot_of[tb / GJD_RING_STRIDE];
    uint32_t e = gjd_resolve(lut->lim[slot], lut->off[slot], lut->vals[slot], lut->isdc[slot], w >> 16, GJD_LUT_BITS + 1);
    if (e & GJD_ENTRY_BAD) {
        /* only a symbol that lies inside the scan counts (the padding after the last MCU does not);
         * benign race: every writer stores the same value */
        if (badcode && room >= 16u) *badcode = 1u;
        e = GJD_ENTRY_UNASSIGNED;
    }
    return e;
}

/* One symbol at window w with the ring at tb: the 16-bit LUT entry.  room = bits left in the scan.
 * Primary LUT on the first GJD_LUT_BITS bits; a long code follows its link into the secondary LUT. */
template <class Lut>
GJD_HD uint32_t gjd_lookup(const Lut& lut, uint32_t tb, uint32_t w, uint32_t room) {
    uint32_t e = lut.entry(tb + ((w >> (31 - GJD_LUT_BITS)) & (GJD_RING_STRIDE - 2u)));
    if ((e & 31u) == 0u) {
        /* sub-table ((e >> 5) - 1) of secondary table tb / RING_STRIDE, entry = next 6 bits */
        if (e) e = lut.entry(GJD_LUT2_OFF(tb, e, w));
        if ((e & 31u) == 0u) e = gjd_lookup_slow(lut.set, tb, w, room, lut.badcode);
    }
    return e;
}

/* Ring step for a symbol decoded at zigzag index z whose successor index is zn (>= 64: block done). */
GJD_HD uint32_t gjd_ring_next(uint32_t tb, uint32_t z, uint32_t zn) {
    if (z == 0u || zn >= 64u) {
        tb += GJD_RING_STRIDE;
        if (tb == GJD_RING_BYTES) tb = 0u;
    }
    return tb;
}

/* How the table ring moves from block to block.  The decoders are templates over it.
 *   GjdRing444  one block of every component per MCU (4:4:4; also single-component scans, whose one table pair fills all
 *               three ring positions): DC -> AC -> next component's DC, the stepping above.  Stateless.
 *   GjdRingSub  chroma-subsampled scans (T.81 A.2.3): an MCU is ny = H*V luma blocks, then Cb, then Cr.  The ring
 *               still holds six tables; after a luma block the ring goes BACK to the luma DC table while luma blocks of
 *               the MCU remain (yl counts them).  pos = block-in-MCU * 64 + zigzag index, up to (ny + 2) * 64 - 1.
 * Only whole blocks change anything, so the per-symbol path of the 4:4:4 decoders is untouched by the second one. */
struct GjdRing444 {
    static constexpr bool sub = false;
    GJD_HD uint32_t of_pos(uint32_t pos) { return gjd_ring_of_pos(pos); }
    GJD_HD uint32_t next(uint32_t tb, uint32_t z, uint32_t zn) { return gjd_ring_next(tb, z, zn); }
    GJD_HD uint32_t slots() const { return 192u; }
};
struct GjdRingSub {
    static constexpr bool sub = true;
    uint32_t ny, yl;
    GJD_HD explicit GjdRingSub(uint32_t ny_) : ny(ny_), yl(0) {}
    GJD_HD uint32_t of_pos(uint32_t pos) {
        const uint32_t bp = pos >> 6;
        const uint32_t comp = bp < ny ? 0u : bp - ny + 1u;
        yl = bp < ny ? ny - 1u - bp : 0u;
        return (comp * 2u + ((pos & 63u) ? 1u : 0u)) * GJD_RING_STRIDE;
    }
    GJD_HD uint32_t next(uint32_t tb, uint32_t z, uint32_t zn) {
        /* straight-line (selects): in a warp some lane completes a block on nearly every trip of the symbol loop, so a branch
         * here would run both of its sides every time */
        const bool done = zn >= 64u;                        /* block complete (never on its DC symbol: that advances one slot) */
        const bool more = done && tb <= GJD_RING_STRIDE && yl != 0u;      /* another luma block of this MCU follows */
        uint32_t nxt = (tb | GJD_RING_STRIDE) + GJD_RING_STRIDE;          /* the next component's DC table */
        const bool wrap = nxt == GJD_RING_BYTES;
        nxt = wrap ? 0u : nxt;
        yl = more ? yl - 1u : (done && wrap) ? ny - 1u : yl;
        return more ? 0u : done ? nxt : z == 0u ? tb + GJD_RING_STRIDE : tb;
    }
    GJD_HD uint32_t slots() const { return (ny + 2u) * 64u; }
};

/* Restart intervals inside one scan (T.81 B.2.4.4, E.1.4; outside the reference's parser, which keeps FF Dn as data).
 * The unstuffed scan of an image holds its intervals back to back, RSTn markers removed: interval k >= 1 starts at the
 * bit position start[k - 1] (a byte boundary); the 0-7 one-bits in front of it pad the previous interval.  A decoder
 * — synchronised or not — that reaches such a position, or whose next symbol would run across it, drops what is left
 * of the interval and restarts there at block 0 of an MCU, zigzag index 0: restart positions are hard synchronisation
 * points.  The decoders are templates over this policy: GjdNoRst compiles to exactly the code without restart intervals
 * (limit == scan_bits), GjdRstTable carries the table.  A block cut short by a restart counts as ended (corrupt data;
 * the write pass flags it). */
struct GjdNoRst {
    static constexpr bool enabled = false;
    GJD_HD uint32_t seek(uint32_t, uint32_t scan_bits) { return scan_bits; }
    GJD_HD uint32_t advance(uint32_t scan_bits) { return scan_bits; }
    GJD_HD uint32_t interval() const { return 0u; }
};
struct GjdRstTable {
    static constexpr bool enabled = true;
    const uint32_t* start;      /* start[k - 1] = bit position of interval k */
    uint32_t count;             /* restart markers = intervals - 1 */
    uint32_t next;              /* start[next] = the boundary in front of the decoder (next == count: the end of the scan) */
    GJD_HD GjdRstTable(const uint32_t* s, uint32_t n) : start(s), count(n), next(0) {}
    /* the first boundary > p: the decoder at p is inside the interval that ends there */
    GJD_HD uint32_t seek(uint32_t p, uint32_t scan_bits) {
        uint32_t lo = 0, hi = count;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (start[mid] <= p) lo = mid + 1u; else hi = mid;
        }
        next = lo;
        return lo < count ? start[lo] : scan_bits;
    }
    GJD_HD uint32_t advance(uint32_t scan_bits) {
        next++;
        return next < count ? start[next] : scan_bits;
    }
    GJD_HD uint32_t interval() const { return next; }   /* index of the interval the decoder is in */
};

/* Decode symbols that START before end_bits, beginning at state s.  A symbol that would run past
 * scan_bits is padding and is not counted.
 * Follows JPEGParser::buildMCU (cpp-decoder/src/parser.cpp:105-134): DC size/diff; AC run/size with
 * EOB = 0x00, "length += run", value bits always consumed, coefficient kept only if index < 64. */
template <class Src, class Lut, class Ring = GjdRing444, class Rst = GjdNoRst>
GJD_HD void gjd_count_span(Src& scan, const Lut& lut, uint32_t end_bits, uint32_t scan_bits,
                           GjdState& s, Ring ring = Ring(), Rst rst = Rst()) {
    if (s.p >= end_bits) return;
    GjdWin win;
    gjd_win_init(win, scan, s.p);
    uint32_t p = s.p, z = s.pos & 63u, tb = ring.of_pos(s.pos), nb = 0;
    uint32_t limit = rst.seek(p, scan_bits);
    while (p < end_bits) {
        GJD_DBG_SYM();
        const uint32_t w = gjd_peek32(win, scan, p);
        const uint32_t e = gjd_lookup(lut, tb, w, limit - p);
        const uint32_t pn = p + (e & 31u);
        if (pn > limit) {
            if (!Rst::enabled || limit >= scan_bits) break;
            p = limit;                                  /* restart: the rest of the interval is padding */
            if (z != 0u) nb++;
            z = 0u;
            tb = ring.of_pos(0u);
            limit = rst.advance(scan_bits);
            continue;
        }
        p = pn;
        const uint32_t zn = z + (e >> 9);
        tb = ring.next(tb, z, zn);
        z = zn;
        if (zn >= 64u) { z = 0u; nb++; }
    }
    const uint32_t adv = nb * 64u + z - (s.pos & 63u);      /* slots advanced (mod 2^32 arithmetic is exact) */
    s.p = p;
    s.n += adv;
    s.pos = (s.pos + adv) % ring.slots();
}

/* Any single symbol from the canonical tables alone (code lengths 1..16): what gjd_lookup returns, without
 * touching a LUT.  Rare paths of the pair-counting loop only. */
GJD_HD_COLD uint32_t gjd_lookup_canonical(const GjdLutSet* lut, uint32_t tb, uint32_t w) {
    GJD_DBG_SLOW();
    const uint32_t slot = lut->slot_of[tb / GJD_RING_STRIDE];
    const uint32_t e = gjd_resolve(lut->lim[slot], lut->off[slot], lut->vals[slot], lut->isdc[slot], w >> 16, 1);
    return (e & GJD_ENTRY_BAD) ? GJD_ENTRY_UNASSIGNED : e;
}

/* gjd_count_span with the pair LUT (lut.entry() reads a GjdLutCount ring followed by lut2; lut.set gives the
 * canonical tables).  Same states as gjd_count_span for every input:
 *   - pairs are taken only while p + GJD_LUT_BITS <= end_bits, so the second symbol of a pair always STARTS
 *     before end_bits; the last few symbols of the span are decoded one at a time from the canonical tables;
 *   - when a pair would complete the block, the first symbol is checked alone: if it already completes the
 *     block (a coefficient at index 63, or garbage), the second symbol belongs to the next block's DC table
 *     and the pair is dropped. */
/* Checkpoints inside a span (k_huff_sync).  A thread that has to decode its subsequence again, from the state its
 * predecessor really ended in, almost always falls back into the path of its first decode within a few hundred bits; what
 * follows would be the same symbols again.  So the first decode leaves its state at a few fixed places of the span, and a
 * repeat stops at the first of them it reaches in the same state (same bit position, same slot in the MCU).  The places
 * cost the symbol loop nothing: a checkpoint is one more fence of the loop that is already there for the end of the span
 * and for restart boundaries, and the loop leaves at "the first symbol boundary less than GJD_LUT_BITS in front of the
 * fence" — the same boundary for every decode that is on the same path.
 *   GjdNoChk   no checkpoints: exactly the code without them.
 *   visit(k, p, n, pos) -> true: stop here (the caller finishes the state from what the first decode left). */
struct GjdNoChk {
    static constexpr bool enabled = false;
    GJD_HD uint32_t fence(uint32_t f) const { return f; }
    GJD_HD bool visit(uint32_t, uint32_t, uint32_t) { return false; }
};

/* The checkpoints of k_huff_sync: GJD_SYNC_CHK evenly spaced places inside a subsequence (3: its quarter points), {p, n << GJD_POS_BITS | pos} like sinfo, checkpoint k
 * of a thread at cp[k * GJD_SEQ_SUBS]. */
#ifndef GJD_SYNC_CHK
#define GJD_SYNC_CHK 3
#endif
struct GjdChkPt { uint32_t x, y; };
struct GjdChkRecord {            /* first decode: leave the state at every checkpoint */
    static constexpr bool enabled = true;
    GjdChkPt* cp;
    uint32_t next, step, k;
    GJD_HD uint32_t fence(uint32_t f) const { return (k < (uint32_t)GJD_SYNC_CHK && next < f) ? next : f; }
    GJD_HD bool visit(uint32_t p, uint32_t n, uint32_t pos) {
        cp[k * GJD_SEQ_SUBS].x = p; cp[k * GJD_SEQ_SUBS].y = (n << GJD_POS_BITS) | pos;
        k++; next += step;
        return false;
    }
};
struct GjdChkCompare {           /* repeat decode: stop at the first checkpoint reached in the state the earlier decode had there */
    static constexpr bool enabled = true;
    GjdChkPt* cp;
    uint32_t next, step, k;
    uint32_t n_old;              /* slots the earlier decode had counted at the checkpoint that matched */
    bool hit;
    GJD_HD uint32_t fence(uint32_t f) const { return (k < (uint32_t)GJD_SYNC_CHK && next < f) ? next : f; }
    GJD_HD bool visit(uint32_t p, uint32_t n, uint32_t pos) {
        const uint32_t ox = cp[k * GJD_SEQ_SUBS].x, oy = cp[k * GJD_SEQ_SUBS].y;
        cp[k * GJD_SEQ_SUBS].x = p; cp[k * GJD_SEQ_SUBS].y = (n << GJD_POS_BITS) | pos;     /* (the same place, the new count) */
        if (ox == p && (oy & GJD_POS_MASK) == pos) { n_old = oy >> GJD_POS_BITS; hit = true; return true; }
        k++; next += step;
        return false;
    }
};
/* After a repeat decode that stopped at checkpoint c.k with st = its state there: the rest of the earlier decode stands —
 * its exit {exit_p, exit_n_pos} and its later checkpoints — with the slot counts shifted by what the new start changed. */
GJD_HD void gjd_chk_finish(const GjdChkCompare& c, GjdState& st, uint32_t exit_p, uint32_t exit_n_pos) {
    const uint32_t delta = st.n - c.n_old;
    for (uint32_t k = c.k + 1u; k < (uint32_t)GJD_SYNC_CHK; k++) c.cp[k * GJD_SEQ_SUBS].y += delta << GJD_POS_BITS;
    st.p = exit_p;
    st.pos = exit_n_pos & GJD_POS_MASK;
    st.n = (exit_n_pos >> GJD_POS_BITS) + delta;
}

template <class Src, class Lut, class Ring = GjdRing444, class Rst = GjdNoRst, class Chk = GjdNoChk>
GJD_HD void gjd_count_span_pairs(Src& scan, const Lut& lut, uint32_t end_bits, uint32_t scan_bits,
                                 GjdState& s, Ring ring = Ring(), Rst rst = Rst(), Chk* chk = nullptr) {
    if (s.p >= end_bits) return;
    GjdWin win;
    gjd_win_init(win, scan, s.p);
    uint32_t p = s.p, z = s.pos & 63u, tb = ring.of_pos(s.pos), nb = 0;
    uint32_t limit = rst.seek(p, scan_bits);
    if (Rst::enabled) {
        /* Scans with restart intervals: ONE flat loop, a pair step or a single step per trip.  The nest of loops below sends
         * a lane that reaches a restart boundary through the single steps and back into the fast loop on its own, and the
         * lanes of a warp do not meet again until the span ends: 17 of 32 lanes were active per instruction on images with
         * one MCU row per interval (profiles/r02Kdri_k_huff_sync.hotspots.txt).  Here they meet at the end of every trip.
         * Same states as the nest for every input (tests/test_hostsim.py: pair counting against the single-symbol walk). */
        bool run = true;
        while (run) {
            GJD_DBG_SYM();
            const uint32_t fence = limit < end_bits ? limit : end_bits;
            const uint32_t w = gjd_peek32(win, scan, p);
            uint32_t pn = p, zn = z;
            bool single = true, step = false;
            if (p + (uint32_t)GJD_LUT_BITS <= fence) {
                uint32_t e = lut.entry(tb + ((w >> (31 - GJD_LUT_BITS)) & (GJD_RING_STRIDE - 2u)));
                if ((e & 31u) == 0u) {
                    if (e) e = lut.entry(GJD_LUT2_OFF(tb, e, w));
                    if ((e & 31u) == 0u) e = gjd_lookup_slow(lut.set, tb, w, limit - p, lut.badcode);
                    e &= ~0x1E0u;
                }
                pn = p + (e & 31u);
                if (pn <= limit) {       /* (else: a long code across the boundary, left to the single step) */
                    zn = z + (e >> 9);
                    single = zn >= 64u && (e & GJD_PAIR_FLAG) && z + ((e >> 6) & 7u) >= 63u;   /* the first symbol alone completes the block */
                    step = !single;
                }
            }
            if (single) {
                if (p >= end_bits) run = false;
                else {
                    const uint32_t e = gjd_lookup_canonical(lut.set, tb, w);
                    pn = p + (e & 31u);
                    if (pn <= limit) { zn = z + (e >> 9); step = true; }
                    else if (limit >= scan_bits) run = false;     /* padding at the end of the scan */
                    else {                                       /* restart: the rest of the interval is padding */
                        p = limit;
                        if (z != 0u) nb++;
                        z = 0u;
                        tb = ring.of_pos(0u);
                        limit = rst.advance(scan_bits);
                    }
                }
            }
            if (step) {
                p = pn;
                tb = ring.next(tb, z, zn);
                z = zn;
                if (zn >= 64u) { z = 0u; nb++; }
            }
        }
    } else
    for (;;) {
        /* fast steps: only where a whole LUT index lies in front of both the end of the span and the next restart */
        const uint32_t fence0 = (Rst::enabled && limit < end_bits) ? limit : end_bits;
        const uint32_t fence = Chk::enabled ? chk->fence(fence0) : fence0;
        if (Chk::enabled && fence < fence0 && p + (uint32_t)GJD_LUT_BITS > fence) {
            /* at a checkpoint (the loop below left because of it, or the span starts this close to it) */
            const uint32_t adv = nb * 64u + z - (s.pos & 63u);
            if (chk->visit(p, s.n + adv, (s.pos + adv) % ring.slots())) {
                s.p = p; s.n += adv; s.pos = (s.pos + adv) % ring.slots();
                return;
            }
            continue;
        }
        while (p + (uint32_t)GJD_LUT_BITS <= fence) {
            GJD_DBG_SYM();
            const uint32_t w = gjd_peek32(win, scan, p);
            uint32_t e = lut.entry(tb + ((w >> (31 - GJD_LUT_BITS)) & (GJD_RING_STRIDE - 2u)));
            if ((e & 31u) == 0u) {   /* long code: single-symbol entries (with a size field) from lut2 / the tables */
                if (e) e = lut.entry(GJD_LUT2_OFF(tb, e, w));
                if ((e & 31u) == 0u) e = gjd_lookup_slow(lut.set, tb, w, limit - p, lut.badcode);
                e &= ~0x1E0u;
            }
            uint32_t pn = p + (e & 31u);
            if (pn > limit) break;   /* (a long code only) left to the single steps below */
            uint32_t zn = z + (e >> 9);
            if (zn >= 64u && (e & GJD_PAIR_FLAG) && z + ((e >> 6) & 7u) >= 63u) {
                /* the first symbol alone completes the block (rare): take it alone, from the canonical tables */
                const uint32_t e1 = gjd_lookup_canonical(lut.set, tb, w);
                pn = p + (e1 & 31u);
                zn = z + (e1 >> 9);
            }
            p = pn;
            tb = ring.next(tb, z, zn);
            z = zn;
            if (zn >= 64u) { z = 0u; nb++; }
        }
        if (Chk::enabled && fence < fence0 && p + (uint32_t)GJD_LUT_BITS > fence) continue;   /* a checkpoint, not the end */
        /* single steps from the canonical tables: the last symbols of the span, or of a restart interval */
        bool again = false;
        while (p < end_bits) {
            GJD_DBG_SYM();
            const uint32_t w = gjd_peek32(win, scan, p);
            const uint32_t e = gjd_lookup_canonical(lut.set, tb, w);
            const uint32_t pn = p + (e & 31u);
            if (pn > limit) {
                if (!Rst::enabled || limit >= scan_bits) break;
                p = limit;                              /* restart: the rest of the interval is padding */
                if (z != 0u) nb++;
                z = 0u;
                tb = ring.of_pos(0u);
                limit = rst.advance(scan_bits);
                again = true;
                break;
            }
            p = pn;
            const uint32_t zn = z + (e >> 9);
            tb = ring.next(tb, z, zn);
            z = zn;
            if (zn >= 64u) { z = 0u; nb++; }
            if (Rst::enabled && p + (uint32_t)GJD_LUT_BITS <= fence) { again = true; break; }   /* a long code sent us here */
        }
        if (!Rst::enabled || !again) break;
    }
    const uint32_t adv = nb * 64u + z - (s.pos & 63u);
    s.p = p;
    s.n += adv;
    s.pos = (s.pos + adv) % ring.slots();
}

/* Warp-wide "any lane still running": keeps the 32 lanes of a warp in one converged loop on the GPU
 * (every lane of the warp must call it); the identity on the host. */
#ifdef __CUDA_ARCH__
#define GJD_WARP_ANY(x) __any_sync(0xffffffffu, (x))
#else
#define GJD_WARP_ANY(x) (x)
#endif

/*
 * Write pass of one subsequence.  The thread starts at the synchronised state s (bit position p,
 * pos = comp*64 + zigzag) whose next coefficient slot is idx0, and OWNS every 8x8 block whose DC symbol
 * starts in [s.p, end_bits): it first skips the tail of the block it starts inside (that block belongs
 * to the thread before it), then decodes its blocks completely — running past end_bits to finish the
 * last one.  Per block it calls sink.store(k, value) for every non-zero AC coefficient (zigzag index
 * k, increasing) and then sink.flush(block_index, dc_difference, kmax) with kmax = the last non-zero
 * zigzag index (0 = DC only).  Both phases are flat loops over symbols driven by a warp vote, so the
 * lanes of a warp stay converged: ALL 32 lanes call this (enable = false for lanes without a
 * subsequence).  Same buildMCU semantics as gjd_count_span (cpp-decoder/src/parser.cpp:105-134).
 */
template <class Src, class Lut, class Sink, class Ring = GjdRing444, class Rst = GjdNoRst>
GJD_HD void gjd_write_span(Src& scan, const Lut& lut, uint32_t end_bits, uint32_t scan_bits,
                           GjdState s, uint32_t idx0, uint32_t nblocks, bool enable, Sink& sink, uint32_t& bad, Ring ring = Ring(),
                           Rst rst = Rst(), uint32_t rst_blocks = 0) {
    uint32_t run = (enable && s.p < end_bits) ? 1u : 0u;
    GjdWin win;
    win = GjdWin();
    if (run) gjd_win_init(win, scan, s.p);
    uint32_t p = s.p, z = s.pos & 63u, tb = ring.of_pos(s.pos);
    uint32_t blk = (idx0 + 63u) >> 6;
    uint32_t limit = run ? rst.seek(p, scan_bits) : scan_bits;

    /* phase A: the tail of a block that belongs to the previous thread */
    uint32_t skip = (run && z != 0u) ? 1u : 0u;
    while (GJD_WARP_ANY(skip != 0u)) {
        if (skip) {
            const uint32_t w = gjd_peek32(win, scan, p);
            const uint32_t e = gjd_lookup(lut, tb, w, limit - p);
            const uint32_t pn = p + (e & 31u);
            if (Rst::enabled && p < end_bits && pn > limit && limit < scan_bits) {
                /* the block was cut short by a restart (corrupt data; its owner flags it): this thread starts with the interval */
                p = limit; z = 0u; tb = ring.of_pos(0u);
                limit = rst.advance(scan_bits);
                skip = 0u;
            } else if (p >= end_bits || pn > limit) {
                skip = 0u; run = 0u;                 /* owns nothing (a truncated block is its owner's to report) */
            } else {
                p = pn;
                const uint32_t zn = z + (e >> 9);
                tb = ring.next(tb, z, zn);
                z = zn >= 64u ? 0u : zn;
                if (z == 0u) skip = 0u;
            }
        }
    }

    /* phase B: the blocks this thread owns.  Two symbols per trip of the warp-voted loop, ONE flush point per
     * trip: with 32 lanes nearly every trip sees some lane complete a block, so a flush inside the symbol step
     * runs its ~45 instructions for 2-3 active lanes on 95 % of the trips (ncu source page, r01m: 37 % of the
     * pass's warp instructions).  A lane completes at most one block per trip: the symbol after a completed
     * block is the next block's DC, which neither completes a block nor touches the staging buffer, so the
     * completed block can wait there (DC difference parked in fdc, kmax left alone) until the end of the trip.
     * (A restart counts as a step that completes the block it cuts short, if any.) */
    int dcdiff = 0, fdc = 0;
    uint32_t kmax = 0, fl = 0;
    while (GJD_WARP_ANY(run != 0u)) {
#ifdef __CUDACC__
#pragma unroll
#endif
        for (int half = 0; half < GJD_WRITE_SYMS; half++) {
            /* (more than two symbols per trip: past a completed block only the next DC symbol may be taken) */
            if (run && !(GJD_WRITE_SYMS > 2 && fl && z != 0u)) {
                const uint32_t w = gjd_peek32(win, scan, p);     /* (reads at most one word into the padding) */
                const uint32_t e = gjd_lookup(lut, tb, w, limit - p);
                const uint32_t total = e & 31u;
                const uint32_t pn = p + total;
                const bool stop = z == 0u && p >= end_bits;
                if (Rst::enabled && !stop && pn > limit && limit < scan_bits) {
                    /* restart: what is left of the interval is padding; the next interval starts a new MCU, and it must
                     * start at the block the restart interval says (T.81 E.1.4) */
                    if (z != 0u) {                               /* a block cut short: corrupt, still flushed (staging stays clean) */
                        bad |= GJD_DEV_BADCOUNT;
                        if (fl || blk >= nblocks) run = 0u;      /* (one flush per trip; cannot happen on valid data) */
                        else { fl = 1u; fdc = dcdiff; blk++; }
                    }
                    p = limit; z = 0u; tb = ring.of_pos(0u);
                    limit = rst.advance(scan_bits);
                    if (blk != rst.interval() * rst_blocks) { bad |= GJD_DEV_BADCOUNT; run = 0u; }
                } else if (stop || pn > limit) {                 /* padding at a block start is the normal end */
                    if (z != 0u) bad |= GJD_DEV_BADCOUNT;
                    run = 0u;
                } else {
                    p = pn;
                    const uint32_t size = (e >> 5) & 15u;
                    const uint32_t zn = z + (e >> 9);            /* next zigzag index; >= 64: block complete */
                    const int v = gjd_extend(w, total - size, size);
                    if (z == 0u) dcdiff = v;
                    else if (size != 0u && zn <= 64u) { sink.store(zn - 1u, v); kmax = zn - 1u; }
                    tb = ring.next(tb, z, zn);
                    z = zn;
                    if (zn >= 64u) {
                        z = 0u;
                        if (blk >= nblocks) { bad |= GJD_DEV_BADCOUNT; run = 0u; }
                        else { fl = 1u; fdc = dcdiff; }
                        blk++;
                    }
                }
            }
        }
        if (fl) {
            sink.flush(blk - 1u, fdc, kmax);
            kmax = 0u;
            fl = 0u;
        }
    }
}

/* ------------------------------------------------------------------ IDCT -------------- */

/* natural index (row*8+col) -> zigzag index; cpp-decoder/src/idct.cpp:7-18 */
#define GJD_ZZ_TABLE                                                                        \
    { 0,  1,  5,  6, 14, 15, 27, 28,  2,  4,  7, 13, 16, 26, 29, 42,  3,  8, 12, 17, 25, 30, 41, 43, \
      9, 11, 18, 24, 31, 40, 44, 53, 10, 19, 23, 32, 39, 45, 52, 54, 20, 22, 33, 38, 46, 51, 55, 60, \
     21, 34, 37, 47, 50, 56, 59, 61, 35, 36, 48, 49, 57, 58, 62, 63 }

#define GJD_C1 2841
#define GJD_C2 2676
#define GJD_C3 2408
#define GJD_C5 1609
#define GJD_C6 1108
#define GJD_C7 565

/* Row pass on b[0..7] (stride 1); general path only — the reference's all-zero shortcut
 * (idct.cpp:38-41) is value-identical: ((b0<<11)+128)>>8 == b0<<3.
 * PRE: what the column pass is handed for this row.  0: the reference's b >> 8.  1 (row 0) and 2 (row 4): the column
 * pass only ever uses these rows as b*256 (x0 = b0*256 + 8192, x1 = b4*256, idct.cpp:83-88), and (s >> 8) * 256 is
 * s & ~255 for every 32-bit s; row 0 also carries the 8192, which is a multiple of 256 and enters all eight sums of
 * the row once: ((s + 8192) >> 8) * 256 = (s >> 8) * 256 + 8192.  Same values, one instruction less per sample. */
template <int PRE>
GJD_HD void gjd_idct_row_t(int& b0, int& b1, int& b2, int& b3, int& b4, int& b5, int& b6, int& b7) {
    int x0, x1, x2, x3, x4, x5, x6, x7, x8;
    x1 = b4 * 2048; x2 = b6; x3 = b2; x4 = b1; x5 = b7; x6 = b5; x7 = b3;
    x0 = b0 * 2048 + (PRE == 1 ? 128 + 8192 : 128);
    x8 = GJD_C7 * (x4 + x5);
    x4 = x8 + (GJD_C1 - GJD_C7) * x4;
    x5 = x8 - (GJD_C1 + GJD_C7) * x5;
    x8 = GJD_C3 * (x6 + x7);
    x6 = x8 - (GJD_C3 - GJD_C5) * x6;
    x7 = x8 - (GJD_C3 + GJD_C5) * x7;
    x8 = x0 + x1;
    x0 -= x1;
    x1 = GJD_C6 * (x3 + x2);
    x2 = x1 - (GJD_C2 + GJD_C6) * x2;
    x3 = x1 + (GJD_C2 - GJD_C6) * x3;
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
    if (PRE == 0) {
        b0 = (x7 + x1) >> 8; b1 = (x3 + x2) >> 8; b2 = (x0 + x4) >> 8; b3 = (x8 + x6) >> 8;
        b4 = (x8 - x6) >> 8; b5 = (x0 - x4) >> 8; b6 = (x3 - x2) >> 8; b7 = (x7 - x1) >> 8;
    } else {
        b0 = (x7 + x1) & ~255; b1 = (x3 + x2) & ~255; b2 = (x0 + x4) & ~255; b3 = (x8 + x6) & ~255;
        b4 = (x8 - x6) & ~255; b5 = (x0 - x4) & ~255; b6 = (x3 - x2) & ~255; b7 = (x7 - x1) & ~255;
    }
}
GJD_HD void gjd_idct_row(int& b0, int& b1, int& b2, int& b3, int& b4, int& b5, int& b6, int& b7) {
    gjd_idct_row_t<0>(b0, b1, b2, b3, b4, b5, b6, b7);
}

/* clip(v >> 14) to [-256,255] (idct.cpp:20-22), returned with a +256 bias: [0,511]. */
GJD_HD int gjd_descale_clip_biased(int v) {
#ifdef __CUDA_ARCH__
    return __viaddmin_s32_relu(v >> 14, 256, 511);
#else
    int t = (v >> 14) + 256;
    t = t > 511 ? 511 : t;
    return t < 0 ? 0 : t;
#endif
}

/* Column pass (idct.cpp:79-122; shortcut :83-86 is value-identical to the general path).
 * RAW = false: outputs are clip(.)+256, i.e. in [0,511], from rows as the reference hands them over.
 * RAW = true : b0 and b4 come pre-multiplied from gjd_idct_row_t<1> / <2> (b0*256 + 8192, b4*256), and the outputs are
 *              the eight sums BEFORE the final >> 14 and clip (gjd_pack_samples finishes two of them at a time). */
template <bool RAW>
GJD_HD void gjd_idct_col_t(int& b0, int& b1, int& b2, int& b3, int& b4, int& b5, int& b6, int& b7) {
    int x0, x1, x2, x3, x4, x5, x6, x7, x8;
    x1 = RAW ? b4 : b4 * 256; x2 = b6; x3 = b2; x4 = b1; x5 = b7; x6 = b5; x7 = b3;
    x0 = RAW ? b0 : b0 * 256 + 8192;
    x8 = GJD_C7 * (x4 + x5) + 4;
    x4 = (x8 + (GJD_C1 - GJD_C7) * x4) >> 3;
    x5 = (x8 - (GJD_C1 + GJD_C7) * x5) >> 3;
    x8 = GJD_C3 * (x6 + x7) + 4;
    x6 = (x8 - (GJD_C3 - GJD_C5) * x6) >> 3;
    x7 = (x8 - (GJD_C3 + GJD_C5) * x7) >> 3;
    x8 = x0 + x1;
    x0 -= x1;
    x1 = GJD_C6 * (x3 + x2) + 4;
    x2 = (x1 - (GJD_C2 + GJD_C6) * x2) >> 3;
    x3 = (x1 + (GJD_C2 - GJD_C6) * x3) >> 3;
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
    if (RAW) {
        b0 = x7 + x1; b1 = x3 + x2; b2 = x0 + x4; b3 = x8 + x6;
        b4 = x8 - x6; b5 = x0 - x4; b6 = x3 - x2; b7 = x7 - x1;
    } else {
        b0 = gjd_descale_clip_biased(x7 + x1);
        b1 = gjd_descale_clip_biased(x3 + x2);
        b2 = gjd_descale_clip_biased(x0 + x4);
        b3 = gjd_descale_clip_biased(x8 + x6);
        b4 = gjd_descale_clip_biased(x8 - x6);
        b5 = gjd_descale_clip_biased(x0 - x4);
        b6 = gjd_descale_clip_biased(x3 - x2);
        b7 = gjd_descale_clip_biased(x7 - x1);
    }
}
GJD_HD void gjd_idct_col(int& b0, int& b1, int& b2, int& b3, int& b4, int& b5, int& b6, int& b7) {
    gjd_idct_col_t<false>(b0, b1, b2, b3, b4, b5, b6, b7);
}

/* ---- packed 16-bit helpers (sm_100a: PRMT, I2IP.S16.S32.SAT, VIADDMNMX.S16x2.RELU; plain C on the host) ---- */
GJD_HD uint32_t gjd_prmt(uint32_t a, uint32_t b, uint32_t sel) {
#ifdef __CUDA_ARCH__
    return __byte_perm(a, b, sel);
#else
    const uint64_t ab = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int i = 0; i < 4; i++) r |= (uint32_t)((ab >> (8 * ((sel >> (4 * i)) & 7u))) & 0xffu) << (8 * i);
    return r;
#endif
}
/* per 16-bit lane: max(min(a + b, c), 0), signed lanes that wrap like the hardware's */
GJD_HD uint32_t gjd_viaddmin_s16x2_relu(uint32_t a, uint32_t b, uint32_t c) {
#ifdef __CUDA_ARCH__
    return __viaddmin_s16x2_relu(a, b, c);
#else
    uint32_t r = 0;
    for (int h = 0; h < 2; h++) {
        int16_t t = (int16_t)((a >> (16 * h)) + (b >> (16 * h)));
        const int16_t m = (int16_t)(c >> (16 * h));
        t = t > m ? m : t;
        t = t < 0 ? (int16_t)0 : t;
        r |= (uint32_t)(uint16_t)t << (16 * h);
    }
    return r;
#endif
}
/* (sat_s16(hi) << 16) | (sat_s16(lo) & 0xffff) */
GJD_HD uint32_t gjd_pack_sat_s16(int hi, int lo) {
#ifdef __CUDA_ARCH__
    uint32_t d;
    asm("cvt.pack.sat.s16.s32 %0, %1, %2;" : "=r"(d) : "r"(hi), "r"(lo));
    return d;
#else
    const int h = hi > 32767 ? 32767 : (hi < -32768 ? -32768 : hi), l = lo > 32767 ? 32767 : (lo < -32768 ? -32768 : lo);
    return ((uint32_t)(uint16_t)(int16_t)h << 16) | (uint32_t)(uint16_t)(int16_t)l;
#endif
}

/* per 16-bit lane: max(min(a, c), 0) */
GJD_HD uint32_t gjd_vimin_s16x2_relu(uint32_t a, uint32_t c) {
#ifdef __CUDA_ARCH__
    return __vimin_s16x2_relu(a, c);
#else
    uint32_t r = 0;
    for (int h = 0; h < 2; h++) {
        int16_t t = (int16_t)(a >> (16 * h));
        const int16_t m = (int16_t)(c >> (16 * h));
        t = t > m ? m : t;
        t = t < 0 ? (int16_t)0 : t;
        r |= (uint32_t)(uint16_t)t << (16 * h);
    }
    return r;
#endif
}

/* Two column-pass sums -> two finished samples in one word: clip(s >> 14) + 256 for both (idct.cpp:20-22, 113-120),
 * s1 in the high half.  (s >> 14) + 256 is one LEA.HI each; the saturation to 16 bits cannot change a value that is
 * clamped to [0,511] next, and no 16-bit lane ever adds (a lane add after the saturation could wrap). */
GJD_HD uint32_t gjd_pack_samples(int s1, int s0) {
    return gjd_vimin_s16x2_relu(gjd_pack_sat_s16((s1 >> 14) + 256, (s0 >> 14) + 256), 0x01ff01ffu);
}

/* Full 8x8: v[] natural order, dequantised in, biased samples (clip+256) out. */
GJD_HD void gjd_idct8x8(int v[64]) {
#pragma unroll
    for (int r = 0; r < 8; r++)
        gjd_idct_row(v[8 * r], v[8 * r + 1], v[8 * r + 2], v[8 * r + 3], v[8 * r + 4], v[8 * r + 5],
                     v[8 * r + 6], v[8 * r + 7]);
#pragma unroll
    for (int c = 0; c < 8; c++)
        gjd_idct_col(v[c], v[8 + c], v[16 + c], v[24 + c], v[32 + c], v[40 + c], v[48 + c], v[56 + c]);
}

/* The same with the shortcuts of gjd_idct_row_t<1,2> / gjd_idct_col_t<true>, in two steps so that k_transform2 can share one
 * column pass between its classes: after both, v[] holds the 64 column-pass sums before the final >> 14 and clip (natural
 * order); gjd_pack_samples turns them into samples. */
GJD_HD void gjd_idct_rows_raw(int v[64]) {
    gjd_idct_row_t<1>(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
#pragma unroll
    for (int r = 1; r < 8; r++) {
        if (r == 4) gjd_idct_row_t<2>(v[32], v[33], v[34], v[35], v[36], v[37], v[38], v[39]);
        else gjd_idct_row_t<0>(v[8 * r], v[8 * r + 1], v[8 * r + 2], v[8 * r + 3], v[8 * r + 4], v[8 * r + 5], v[8 * r + 6], v[8 * r + 7]);
    }
}
/* TOP4: rows 4-7 are known to be zero (no coefficient at or past zigzag index 10): their inputs are literal zeros here, so the
 * butterflies that only see zeros fall away at compile time — same values as the general pass on the same data */
template <bool TOP4>
GJD_HD void gjd_idct_cols_raw(int v[64]) {
#pragma unroll
    for (int c = 0; c < 8; c++) {
        if (TOP4) {
            int z4 = 0, z5 = 0, z6 = 0, z7 = 0;
            gjd_idct_col_t<true>(v[c], v[8 + c], v[16 + c], v[24 + c], z4, z5, z6, z7);
            v[32 + c] = z4; v[40 + c] = z5; v[48 + c] = z6; v[56 + c] = z7;
        } else {
            gjd_idct_col_t<true>(v[c], v[8 + c], v[16 + c], v[24 + c], v[32 + c], v[40 + c], v[48 + c], v[56 + c]);
        }
    }
}
GJD_HD void gjd_idct8x8_raw(int v[64]) {
    gjd_idct_rows_raw(v);
    gjd_idct_cols_raw<false>(v);
}

/* ------------------------------------------------------------------ colour ------------ */
/*
 * The reference computes (color.cpp:10-16, called as colorConversion(Y, Cr, Cb)):
 *     float r = Cr*(2-2*0.299) + Y;  float b = Cb*(2-2*0.114) + Y;      (double math, rounded to float)
 *     float g = (Y - 0.114*b - 0.299*r)/0.587;                           (double math on the floats)
 *     R = clamp((int)(r+128)) ...                                        (float add, truncation)
 * With Y,Cb,Cr integers in [-256,255]:
 *   - r+128 and b+128 are never within rounding distance of a non-integer boundary (fractions are
 *     multiples of 1/500 and 1/250) and truncation vs floor only differs below 0, where the clamp
 *     wins — so R and B equal clamp(Y + 128 + floor(K*C / 2^20)) with K = ceil(1.402*2^20),
 *     ceil(1.772*2^20).  Proven over all 2^18 (Y,C) pairs (tests/test_color.py).
 *   - g is a rational with denominator 587000: float rounding can flip the truncation when the exact
 *     value is within 2.8e-5 of an integer (843 of the 2^27 triples).  The fast path evaluates G in
 *     2^-20 fixed point and, when the fraction is within 2^-14 of an integer, falls back to the
 *     reference's own IEEE sequence (gjd_color_g_exact).  Proven over all 2^27 triples.
 * Inputs to the fast path carry the +256 bias of gjd_descale_clip_biased.
 */
#define GJD_KR 1470104            /* ceil(1.402 * 2^20) */
#define GJD_KB 1858077            /* ceil(1.772 * 2^20) */
#define GJD_KGB 360853            /* round(0.114*1.772/0.587 * 2^20) */
#define GJD_KGR 748826            /* round(0.299*1.402/0.587 * 2^20) */
#define GJD_GDELTA 64             /* danger half-width, 2^-14 in 2^-20 units */
#define GJD_CONST_R (-256 * GJD_KR - (128 << 20))
#define GJD_CONST_B (-256 * GJD_KB - (128 << 20))
#define GJD_CONST_G (256 * GJD_KGB + 256 * GJD_KGR - (128 << 20) + GJD_GDELTA)

GJD_HD int gjd_addmin_relu(int a, int b, int c) {   /* max(min(a+b, c), 0) */
#ifdef __CUDA_ARCH__
    return __viaddmin_s32_relu(a, b, c);
#else
    int t = a + b;
    t = t > c ? c : t;
    return t < 0 ? 0 : t;
#endif
}

/* The reference's IEEE sequence for G, on unbiased inputs. No FMA contraction anywhere. */
GJD_HD_COLD int gjd_color_g_exact(int y, int cb, int cr) {
#ifdef __CUDA_ARCH__
    float r = __double2float_rn(__dadd_rn(__dmul_rn((double)cr, (2 - 2 * 0.299)), (double)y));
    float b = __double2float_rn(__dadd_rn(__dmul_rn((double)cb, (2 - 2 * 0.114)), (double)y));
    double t = __dsub_rn(__dsub_rn((double)y, __dmul_rn(0.114, (double)b)), __dmul_rn(0.299, (double)r));
    float g = __double2float_rn(__ddiv_rn(t, 0.587));
    int G = __float2int_rz(__fadd_rn(g, 128.0f));
#else
    volatile double rd = (double)cr * (2 - 2 * 0.299);
    volatile double rs = rd + (double)y;
    float r = (float)rs;
    volatile double bd = (double)cb * (2 - 2 * 0.114);
    volatile double bs = bd + (double)y;
    float b = (float)bs;
    volatile double m1 = 0.114 * (double)b;
    volatile double m2 = 0.299 * (double)r;
    volatile double t1 = (double)y - m1;
    volatile double t2 = t1 - m2;
    float g = (float)(t2 / 0.587);
    volatile float gs = g + 128.0f;
    int G = (int)gs;
#endif
    return G > 255 ? 255 : (G < 0 ? 0 : G);
}

/* Fast path on biased inputs (yb, cbb, crb in [0,511]).  Returns the fraction field of G left-aligned
 * (tg << 12): G sits in the danger zone and must be recomputed with gjd_color_g_exact (843 of the 2^27
 * triples) when it is below GJD_FRAC_DANGER. */
#define GJD_FRAC_DANGER ((2u * GJD_GDELTA) << 12)
GJD_HD uint32_t gjd_color_fast_frac(int yb, int cbb, int crb, int& R, int& G, int& B) {
    int tr = crb * GJD_KR + GJD_CONST_R;
    int tb = cbb * GJD_KB + GJD_CONST_B;
    int tg = crb * (-GJD_KGR) + GJD_CONST_G;
    tg = cbb * (-GJD_KGB) + tg;
    R = gjd_addmin_relu(tr >> 20, yb, 255);
    B = gjd_addmin_relu(tb >> 20, yb, 255);
    G = gjd_addmin_relu(tg >> 20, yb, 255);
    return (uint32_t)tg * 4096u;
}
/* d = sat_u8(b) | sat_u8(a) << 8 | (c & 0xffff) << 16  (PTX cvt.pack.sat.u8.s32.b32 = SASS I2IP.U8.S32.SAT):
 * the 0..255 clamp of two pixels and their packing in one instruction. */
GJD_HD uint32_t gjd_pack_sat_u8(int a, int b, uint32_t c) {
#ifdef __CUDA_ARCH__
    uint32_t d;
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#else
    const uint32_t sa = a < 0 ? 0u : (a > 255 ? 255u : (uint32_t)a), sb = b < 0 ? 0u : (b > 255 ? 255u : (uint32_t)b);
    return sb | (sa << 8) | ((c & 0xffffu) << 16);
#endif
}

/* gjd_color_fast_frac without the clamps: R, G, B before saturation (for gjd_pack_sat_u8). */
GJD_HD uint32_t gjd_color_fast_raw(int yb, int cbb, int crb, int& R, int& G, int& B) {
    int tr = crb * GJD_KR + GJD_CONST_R;
    int tb = cbb * GJD_KB + GJD_CONST_B;
    int tg = crb * (-GJD_KGR) + GJD_CONST_G;
    tg = cbb * (-GJD_KGB) + tg;
    R = (tr >> 20) + yb;
    B = (tb >> 20) + yb;
    G = (tg >> 20) + yb;
    return (uint32_t)tg * 4096u;
}

GJD_HD bool gjd_color_fast(int yb, int cbb, int crb, int& R, int& G, int& B) {
    return gjd_color_fast_frac(yb, cbb, crb, R, G, B) < GJD_FRAC_DANGER;
}

GJD_HD void gjd_color(int yb, int cbb, int crb, int& R, int& G, int& B) {
    if (gjd_color_fast(yb, cbb, crb, R, G, B)) G = gjd_color_g_exact(yb - 256, cbb - 256, crb - 256);
}

/* ------------------------------------------------------------------ colour, 8 pixels of a row at a time ----------
 * What the transform kernels run (the scalar gjd_color above stays as the per-pixel statement of the same function).
 * Samples arrive as the IDCT parks them: two biased 9-bit samples per word (x odd in the high half).
 *   R, B   2^-16 fixed point: hi16(Cr' * KR16 + CR16) = floor(1.402 Cr) - 128 for every Cr' = Cr + 256 in [0,511]
 *          (all 512 values equal the 2^-20 form above), so the two high halves of two pixels are ONE byte permute away
 *          from a packed pair that a single VIADDMNMX.S16x2.RELU adds to the packed Y pair and clamps to [0,255]: no
 *          per-pixel shift, add or clamp, and Y is never unpacked.
 *   G      the 2^-20 fixed point of gjd_color_fast_raw (two IMADs per pixel), its integer parts packed the same way; the
 *          fractions' unsigned minimum over the row flags the danger zone.  A flagged row re-runs the reference's IEEE
 *          sequence (gjd_color_g_exact) for its flagged pixels — EXCEPT neutral ones (Cb = Cr = 0): their exact G is the
 *          integer Y + 128 the fast path already gives, and they sit in the middle of the zone by construction, so round 1
 *          sent every gray pixel through the double-precision divide.
 *   A 2^-32 form of G in 64 bits (no shifts, a one-sided zone that excludes the integers) was measured and is slower: the
 *   kernels are bound by the ALU pipe, and the 64-bit carries are ALU instructions (profiles/r02f_transform_ab.txt).
 * Proven against cpp-decoder/utils/color.cpp:3-18 for all 2^27 triples on the CPU (tests/test_hostsim.py::
 * test_colour_exhaustive_packed_rows) and on the GPU (tests/test_gpu_parity.py::test_colour_exhaustive_on_device). */
#define GJD_KR16 91881
#define GJD_CR16 (-31910037)
#define GJD_KB16 116130
#define GJD_CB16 (-38117862)

/* yw, bw, rw: four words each = the 8 samples of one row of Y, Cb, Cr.  r, g, b: two words each = 8 clamped bytes. */
GJD_HD void gjd_color_row8(const uint32_t yw[4], const uint32_t bw[4], const uint32_t rw[4], uint32_t r[2], uint32_t g[2], uint32_t b[2]) {
    uint32_t r2[4], g2[4], b2[4], fmin = 0xffffffffu;
#ifdef __CUDACC__
#pragma unroll
#endif
    for (int j = 0; j < 4; j++) {
        const uint32_t cb0 = bw[j] & 0xffffu, cb1 = bw[j] >> 16, cr0 = rw[j] & 0xffffu, cr1 = rw[j] >> 16;
        const uint32_t tr0 = cr0 * (uint32_t)GJD_KR16 + (uint32_t)GJD_CR16, tr1 = cr1 * (uint32_t)GJD_KR16 + (uint32_t)GJD_CR16;
        const uint32_t tb0 = cb0 * (uint32_t)GJD_KB16 + (uint32_t)GJD_CB16, tb1 = cb1 * (uint32_t)GJD_KB16 + (uint32_t)GJD_CB16;
        r2[j] = gjd_viaddmin_s16x2_relu(gjd_prmt(tr0, tr1, 0x7632u), yw[j], 0x00ff00ffu);
        b2[j] = gjd_viaddmin_s16x2_relu(gjd_prmt(tb0, tb1, 0x7632u), yw[j], 0x00ff00ffu);
        int tg0 = (int)cr0 * (-GJD_KGR) + GJD_CONST_G, tg1 = (int)cr1 * (-GJD_KGR) + GJD_CONST_G;
        tg0 = (int)cb0 * (-GJD_KGB) + tg0;
        tg1 = (int)cb1 * (-GJD_KGB) + tg1;
        g2[j] = gjd_viaddmin_s16x2_relu(gjd_prmt((uint32_t)(tg0 >> 20), (uint32_t)(tg1 >> 20), 0x5410u), yw[j], 0x00ff00ffu);
        const uint32_t f0 = (uint32_t)tg0 * 4096u, f1 = (uint32_t)tg1 * 4096u;
        fmin = fmin < f0 ? fmin : f0;
        fmin = fmin < f1 ? fmin : f1;
    }
    if (fmin < GJD_FRAC_DANGER) {        /* some pixel's G is within 2^-14 of an integer */
#ifdef __CUDACC__
#pragma unroll
#endif
        for (int x = 0; x < 8; x++) {    /* (unrolled: run-time indices would push the arrays into local memory) */
            const int sh = (x & 1) * 16;
            const uint32_t cbb = (bw[x >> 1] >> sh) & 0xffffu, crb = (rw[x >> 1] >> sh) & 0xffffu;
            const uint32_t f = (uint32_t)((int)crb * (-GJD_KGR) + (int)cbb * (-GJD_KGB) + GJD_CONST_G) * 4096u;
            if (f < GJD_FRAC_DANGER && (cbb != 256u || crb != 256u)) {
                const uint32_t yb = (yw[x >> 1] >> sh) & 0xffffu;
                const uint32_t G = (uint32_t)gjd_color_g_exact((int)yb - 256, (int)cbb - 256, (int)crb - 256);
                g2[x >> 1] = (g2[x >> 1] & ~(0xffffu << sh)) | (G << sh);
            }
        }
    }
    for (int h = 0; h < 2; h++) {        /* bytes 0 and 2 of two words -> four pixels in one word */
        r[h] = gjd_prmt(r2[2 * h], r2[2 * h + 1], 0x6420u);
        g[h] = gjd_prmt(g2[2 * h], g2[2 * h + 1], 0x6420u);
        b[h] = gjd_prmt(b2[2 * h], b2[2 * h + 1], 0x6420u);
    }
}

#endif /* GJD_DEVFN_CUH */
