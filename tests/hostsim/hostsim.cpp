/*
 * hostsim.cpp — TEST SCAFFOLDING: steps the product's __host__ __device__ functions
 * (gpu-jpeg-decoder_b200/csrc/gjd_devfn.cuh) and host parser on the CPU, thread by thread, in the
 * same order the CUDA kernels of gjd_kernels.cu do (cold decode, lock-step synchronisation rounds,
 * boundary fix-up, count scan, coefficient write pass with block ownership and sector flushes, DC
 * prefix and the per-MCU transform).  It exists because the build machine has no GPU: it lets `pytest -m "not gpu"` check
 * the algorithm and every arithmetic function against the oracle before a kernel ever runs.
 * It is not part of libgjd.so and nothing in the product can reach it.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

static unsigned long long g_syms_dbg = 0, g_slow_dbg = 0;
#define GJD_DBG_SYM() (g_syms_dbg++)
#define GJD_DBG_SLOW() (g_slow_dbg++)
#include "../../gpu-jpeg-decoder_b200/csrc/gjd_devfn.cuh"
#include "../../gpu-jpeg-decoder_b200/csrc/gjd_parse.h"
#include "../../include/gjd.h"

namespace {

/* host twin of GjdWriteSink (gjd_kernels.cu): a 64-entry staging block per thread, flushed as whole
 * 32-byte sectors up to kmax */
struct WriteSink {
    int16_t stage[64];
    int16_t* coef;      /* image coefficient lines, 64 int16 per block */
    uint32_t* meta;
    uint8_t* written;   /* per block: times flushed (must end up exactly 1) */
    void store(uint32_t k, int v) { stage[k] = (int16_t)v; }
    void flush(uint32_t blk, int dcdiff, uint32_t kmax) {
        meta[blk] = ((uint32_t)dcdiff & 0xffffu) | (kmax << 16);
        written[blk]++;
        if (kmax == 0u) return;
        uint32_t n = ((kmax >> 4) + 1u) * 16u;
        memcpy(coef + (size_t)blk * 64, stage, n * sizeof(int16_t));
        memset(stage, 0, n * sizeof(int16_t));
    }
};

const int kZZ[64] = GJD_ZZ_TABLE;

}  // namespace

extern "C" {

/* stats: [0] max synchronisation rounds inside a sequence, [1] fix-up subsequences decoded,
 *        [2] device status bits, [3] nosync count, [4] subsequences, [5] sequences,
 *        [6] subsequences re-decoded in validation rounds (all rounds), [7] of those, in round 1 */
int hostsim_decode(const uint8_t* jpg, size_t len, int sub_log2_arg, int16_t* coef_out, uint8_t* rgb_out,
                   int32_t* width, int32_t* height, uint32_t stats[9]) {
    const uint32_t sub_log2 = sub_log2_arg ? (uint32_t)sub_log2_arg : GJD_SUB_LOG2_MIN;
    const uint32_t S = 1u << sub_log2;
    GjdParsed p;
    int rc = gjd_parse_headers(jpg, len, &p);
    if (rc) return rc;
    if (width) *width = p.width;
    if (height) *height = p.height;
    const size_t raw = (size_t)(p.file_end - p.scan_begin);
    std::vector<uint8_t> scanbuf(raw + raw / 2 + 128, 0);
    std::vector<uint32_t> seg_off, seg_len;
    /* like batch_stage: restart markers are looked for only where the header announced a restart interval (an RSTn in a
     * scan without DRI is GJD_E_UNSUPPORTED whatever its number) */
    long ulen = p.restart_interval ? gjd_unstuff_segments(p.scan_begin, p.file_end, scanbuf.data(), &seg_off, &seg_len)
                                   : gjd_unstuff(p.scan_begin, p.file_end, scanbuf.data());
    if (ulen < 0) return (int)ulen;
    if (!p.restart_interval) { seg_off.assign(1, 0u); seg_len.assign(1, (uint32_t)ulen); }
    memset(scanbuf.data() + ulen, 0, 16);
    /* word-aligned copy */
    std::vector<uint32_t> scanw((scanbuf.size() + 3) / 4 + 4, 0);
    memcpy(scanw.data(), scanbuf.data(), scanbuf.size());
    const uint32_t W = (uint32_t)p.width, H = (uint32_t)p.height;
    const bool gray = p.components == 1;
    const uint32_t hs = gray ? 1u : p.hs, vs = gray ? 1u : p.vs, ny = hs * vs;
    const bool subsamp = ny > 1u;                                  /* chroma-subsampled: ny luma blocks + Cb + Cr per MCU */
    const uint32_t mcus_x = (W + 8 * hs - 1) / (8 * hs), mcu_count = mcus_x * ((H + 8 * vs - 1) / (8 * vs));
    const uint32_t bpm = gray ? 1u : ny + 2u;                  /* 8x8 blocks per MCU */
    const uint32_t mcu_slots = (gray ? 3u : bpm) * 64u;
    const uint32_t tile_count = (mcu_count * bpm + 95) / 96;
    const uint32_t ri = p.restart_interval;
    if (!ri && seg_off.size() != 1) return GJD_E_UNSUPPORTED;          /* RSTn without DRI */
    if (ri && seg_off.size() != (mcu_count + ri - 1) / ri) return GJD_E_CORRUPT;

    /* k_build_luts */
    GjdLutSet* lut = (GjdLutSet*)calloc(1, sizeof(GjdLutSet));
    for (int t = 0; t < 4; t++) {
        gjd_canonical(p.dht.bits[t], lut->lim[t], lut->off[t]);
        memcpy(lut->vals[t], p.dht.vals[t], 256);
        lut->isdc[t] = p.dht.isdc[t];
    }
    for (uint32_t r = 0; r < GJD_RING_TABLES; r++) {
        uint32_t t = gjd_ring_slot(p.tabsel, r);
        lut->slot_of[r] = (uint8_t)t;
        for (uint32_t i = 0; i < GJD_LUT_SIZE; i++)
            lut->lut[r][i] = gjd_lut_entry(lut->lim[t], lut->off[t], lut->vals[t], lut->isdc[t], i);
        for (uint32_t i = 0; i < GJD_LUT2_SIZE; i++)
            lut->lut2[r][i] = gjd_lut2_entry(lut->lim[t], lut->off[t], lut->vals[t], lut->isdc[t], i);
    }
    uint32_t badcode = 0;
    GjdLutMem hlut;
    hlut.set = lut;
    hlut.badcode = &badcode;
    /* k_huff_sync's view (copy_count_lut_to_smem): the pair LUT ring in place of the primary one */
    GjdLutSet* lutS = (GjdLutSet*)malloc(sizeof(GjdLutSet));
    memcpy(lutS, lut, sizeof(GjdLutSet));
    for (uint32_t r = 0; r < GJD_RING_TABLES; r++) {
        uint32_t t = lut->slot_of[r];
        for (uint32_t i = 0; i < GJD_LUT_SIZE; i++)
            lutS->lut[r][i] = gjd_lutc_entry(lut->lim[t], lut->off[t], lut->vals[t], lut->isdc[t], i);
    }
    GjdLutMem clut;
    clut.set = lutS;
    clut.badcode = nullptr;
    uint32_t max_rounds = 0, fix_decodes = 0, nosync = 0, status = 0, sub_total = 0, seq_total = 0, redec = 0, redec1 = 0, chk_hits = 0;
    std::vector<int16_t> coef((size_t)tile_count * 96 * 64, (int16_t)0x7e7e);   /* stale garbage where nothing is written */
    std::vector<uint32_t> meta((size_t)tile_count * 96, 0);
    std::vector<uint8_t> written((size_t)tile_count * 96, 0);
    /* the entropy kernels run per segment (restart interval); one segment without restarts */
    /* one segment per image; restart intervals lie back to back inside it (GjdRstTable: their start positions) */
    std::vector<uint32_t> rst_bits;
    for (size_t k = 1; k < seg_off.size(); k++) rst_bits.push_back(seg_off[k] * 8u);
    const GjdRstTable rstt(rst_bits.data(), (uint32_t)rst_bits.size());
    const uint32_t rst_blocks = bpm * ri;
    const bool has_rst = !rst_bits.empty();
    /* the four instantiations the kernels choose from (MCU shape x restart intervals) */
    auto count_pairs = [&](const uint32_t*& sc, const GjdLutMem& l, uint32_t end, uint32_t bits, GjdState& st) {
        if (has_rst) { if (subsamp) gjd_count_span_pairs(sc, l, end, bits, st, GjdRingSub(ny), rstt); else gjd_count_span_pairs(sc, l, end, bits, st, GjdRing444(), rstt); }
        else if (subsamp) gjd_count_span_pairs(sc, l, end, bits, st, GjdRingSub(ny)); else gjd_count_span_pairs(sc, l, end, bits, st);
    };
#if GJD_SYNC_CHK
    /* the checkpointed forms k_huff_sync uses for a thread's own subsequence (first decode / repeat) */
    auto count_pairs_rec = [&](const uint32_t*& sc, const GjdLutMem& l, uint32_t end, uint32_t bits, GjdState& st, GjdChkRecord* c) {
        if (has_rst) { if (subsamp) gjd_count_span_pairs(sc, l, end, bits, st, GjdRingSub(ny), rstt, c); else gjd_count_span_pairs(sc, l, end, bits, st, GjdRing444(), rstt, c); }
        else if (subsamp) gjd_count_span_pairs(sc, l, end, bits, st, GjdRingSub(ny), GjdNoRst(), c); else gjd_count_span_pairs(sc, l, end, bits, st, GjdRing444(), GjdNoRst(), c);
    };
    auto count_pairs_cmp = [&](const uint32_t*& sc, const GjdLutMem& l, uint32_t end, uint32_t bits, GjdState& st, GjdChkCompare* c) {
        if (has_rst) { if (subsamp) gjd_count_span_pairs(sc, l, end, bits, st, GjdRingSub(ny), rstt, c); else gjd_count_span_pairs(sc, l, end, bits, st, GjdRing444(), rstt, c); }
        else if (subsamp) gjd_count_span_pairs(sc, l, end, bits, st, GjdRingSub(ny), GjdNoRst(), c); else gjd_count_span_pairs(sc, l, end, bits, st, GjdRing444(), GjdNoRst(), c);
    };
#endif
    auto count_one = [&](const uint32_t*& sc, const GjdLutMem& l, uint32_t end, uint32_t bits, GjdState& st) {
        if (has_rst) { if (subsamp) gjd_count_span(sc, l, end, bits, st, GjdRingSub(ny), rstt); else gjd_count_span(sc, l, end, bits, st, GjdRing444(), rstt); }
        else if (subsamp) gjd_count_span(sc, l, end, bits, st, GjdRingSub(ny)); else gjd_count_span(sc, l, end, bits, st);
    };
    {
    const uint32_t* scan = scanw.data();
    const uint32_t scan_bits = (uint32_t)ulen * 8u;
    const uint32_t blk0 = 0;
    const uint32_t blk_count = bpm * mcu_count;
    const uint32_t sub_count = std::max<uint32_t>(1u, (uint32_t)(((uint64_t)scan_bits + S - 1) / S));
    const uint32_t seq_count = (sub_count + GJD_SEQ_SUBS - 1) / GJD_SEQ_SUBS;
    sub_total += sub_count; seq_total += seq_count;
    std::vector<uint32_t> sp((size_t)seq_count * GJD_SEQ_SUBS, 0), sn((size_t)seq_count * GJD_SEQ_SUBS, 0);

    /* k_huff_sync: one "block" per sequence */
    g_syms_dbg = 0; g_slow_dbg = 0;
    std::vector<uint32_t> seq_in_p(seq_count, 0), seq_in_pos(seq_count, 0);
    for (uint32_t g = 0; g < seq_count; g++) {
        uint32_t* bp = &sp[(size_t)g * GJD_SEQ_SUBS];
        uint32_t* bn = &sn[(size_t)g * GJD_SEQ_SUBS];
        std::vector<uint32_t> in_p(GJD_SEQ_SUBS, 0), in_pos(GJD_SEQ_SUBS, 0), endb(GJD_SEQ_SUBS, 0);
        std::vector<char> active(GJD_SEQ_SUBS, 0);
#if GJD_SYNC_CHK
        std::vector<GjdChkPt> cps((size_t)GJD_SYNC_CHK * GJD_SEQ_SUBS);     /* [checkpoint][thread], like the kernel's shared array */
#endif
        for (uint32_t tid = 0; tid < GJD_SEQ_SUBS; tid++) {
            uint32_t sub = g * GJD_SEQ_SUBS + tid;
            active[tid] = sub < sub_count;
            GjdState st; st.p = 0; st.n = 0; st.pos = 0;
            if (active[tid]) {
                uint32_t start = sub << sub_log2;
                endb[tid] = std::min(start + S, scan_bits);
                if (sub != 0) {
                    const uint32_t lookback = subsamp ? GJD_LOOKBACK_SUB_BITS : GJD_LOOKBACK_BITS;
                    st.p = start > lookback ? start - lookback : 0u;
                    count_pairs(scan, clut, start, scan_bits, st);
                }
                in_p[tid] = st.p; in_pos[tid] = st.pos;
                st.n = 0;
#if GJD_SYNC_CHK
                if (subsamp) {      /* like the kernel: checkpoints for chroma-subsampled scans only */
                    GjdChkRecord rec = {&cps[tid], start + S / (GJD_SYNC_CHK + 1u), S / (GJD_SYNC_CHK + 1u), 0u};
                    count_pairs_rec(scan, clut, endb[tid], scan_bits, st, &rec);
                } else
#endif
                count_pairs(scan, clut, endb[tid], scan_bits, st);
            }
            bp[tid] = st.p;
            bn[tid] = (st.n << GJD_POS_BITS) | st.pos;
        }
        seq_in_p[g] = in_p[0]; seq_in_pos[g] = in_pos[0];
        for (uint32_t r = 0; r < GJD_SEQ_SUBS; r++) {
            /* all threads read their predecessor first, then the repaired ones write (two barriers) */
            std::vector<uint32_t> np(GJD_SEQ_SUBS), nn(GJD_SEQ_SUBS);
            std::vector<char> badv(GJD_SEQ_SUBS, 0);
            bool any = false;
            for (uint32_t tid = 1; tid < GJD_SEQ_SUBS; tid++) {
                if (!active[tid]) continue;
                uint32_t pp = bp[tid - 1], ppos = bn[tid - 1] & GJD_POS_MASK;
                if (pp != in_p[tid] || ppos != in_pos[tid]) {
                    GjdState st; st.p = in_p[tid] = pp; st.pos = in_pos[tid] = ppos; st.n = 0;
#if GJD_SYNC_CHK
                    if (!subsamp) count_pairs(scan, clut, endb[tid], scan_bits, st);
                    else {
                    const uint32_t start = (g * GJD_SEQ_SUBS + tid) << sub_log2;
                    GjdChkCompare cmp = {&cps[tid], start + S / (GJD_SYNC_CHK + 1u), S / (GJD_SYNC_CHK + 1u), 0u, 0u, false};
                    count_pairs_cmp(scan, clut, endb[tid], scan_bits, st, &cmp);
                    if (cmp.hit) {
                        gjd_chk_finish(cmp, st, bp[tid], bn[tid]);
                        /* the short cut must give what the whole repeat gives */
                        GjdState full; full.p = pp; full.pos = ppos; full.n = 0;
                        count_pairs(scan, clut, endb[tid], scan_bits, full);
                        if (full.p != st.p || full.pos != st.pos || full.n != st.n) { fprintf(stderr, "hostsim: checkpoint short cut differs\n"); abort(); }
                        chk_hits++;
                    }
                    }
#else
                    count_pairs(scan, clut, endb[tid], scan_bits, st);
#endif
                    np[tid] = st.p; nn[tid] = (st.n << GJD_POS_BITS) | st.pos;
                    badv[tid] = 1; any = true; redec++; if (r == 0) redec1++;
                }
            }
            if (!any) break;
            max_rounds = std::max(max_rounds, r + 1);
            for (uint32_t tid = 1; tid < GJD_SEQ_SUBS; tid++)
                if (badv[tid]) { bp[tid] = np[tid]; bn[tid] = nn[tid]; }
        }
    }
    if (getenv("HOSTSIM_DEBUG"))
        fprintf(stderr, "hostsim sync pass: %u subsequences, %llu loop trips, %llu of them through the canonical tables, %u repeats (%u cut short)\n",
                sub_count, g_syms_dbg, g_slow_dbg, redec, chk_hits);
    /* k_huff_fix, iterated to its fixed point like batch_wait does */
    for (int iter = 0; iter < 65536; iter++) {
        uint32_t ns = 0;
        for (uint32_t g = 1; g < seq_count; g++) {
            size_t base = (size_t)g * GJD_SEQ_SUBS;
            uint32_t pp = sp[base - 1], ppos = sn[base - 1] & GJD_POS_MASK;
            if (pp == seq_in_p[g] && ppos == seq_in_pos[g]) continue;
            seq_in_p[g] = pp; seq_in_pos[g] = ppos;
            GjdState st; st.p = pp; st.pos = ppos;
            bool merged = false;
            for (uint32_t t = 0; t < GJD_SEQ_SUBS; t++) {
                uint32_t sub = g * GJD_SEQ_SUBS + t;
                if (sub >= sub_count) { merged = true; break; }
                st.n = 0;
                uint32_t end = std::min((sub + 1) << sub_log2, scan_bits);
                count_one(scan, hlut, end, scan_bits, st);
                fix_decodes++;
                bool match = sp[base + t] == st.p && (sn[base + t] & GJD_POS_MASK) == st.pos;
                sp[base + t] = st.p;
                sn[base + t] = (st.n << GJD_POS_BITS) | st.pos;
                if (match) { merged = true; break; }
            }
            if (!merged && g + 1 < seq_count) ns++;
        }
        nosync += ns;
        if (ns == 0) break;
    }
    /* k_seq_scan */
    std::vector<uint32_t> seq_base(seq_count, 0);
    uint32_t carry = 0;
    for (uint32_t g = 0; g < seq_count; g++) {
        seq_base[g] = carry;
        uint32_t cnt = std::min<uint32_t>(GJD_SEQ_SUBS, sub_count - g * GJD_SEQ_SUBS);
        for (uint32_t t = 0; t < cnt; t++) carry += sn[(size_t)g * GJD_SEQ_SUBS + t] >> GJD_POS_BITS;
    }
    badcode = 0;   /* only the write pass reports unassigned codes */
    if (carry != 64u * blk_count) status |= GJD_DEV_BADCOUNT;
    /* k_huff_write */
    for (uint32_t g = 0; g < seq_count; g++) {
        uint32_t excl = 0;
        for (uint32_t tid = 0; tid < GJD_SEQ_SUBS; tid++) {
            uint32_t sub = g * GJD_SEQ_SUBS + tid;
            if (sub >= sub_count) break;
            size_t a = (size_t)g * GJD_SEQ_SUBS + tid;
            GjdState st; st.p = 0; st.n = 0; st.pos = 0;
            if (sub != 0) { st.p = sp[a - 1]; st.pos = sn[a - 1] & GJD_POS_MASK; }
            uint32_t tbad = 0;
            WriteSink sink;
            memset(sink.stage, 0, sizeof(sink.stage));
            sink.coef = coef.data() + (size_t)blk0 * 64; sink.meta = meta.data() + blk0; sink.written = written.data() + blk0;
            uint32_t idx0 = seq_base[g] + excl;
            if (idx0 % mcu_slots != st.pos && st.p < scan_bits) tbad |= GJD_DEV_BADSTATE;
            uint32_t end = std::min((sub + 1) << sub_log2, scan_bits);
            if (has_rst) {
                if (subsamp) gjd_write_span(scan, hlut, end, scan_bits, st, idx0, blk_count, true, sink, tbad, GjdRingSub(ny), rstt, rst_blocks);
                else gjd_write_span(scan, hlut, end, scan_bits, st, idx0, blk_count, true, sink, tbad, GjdRing444(), rstt, rst_blocks);
            } else if (subsamp) gjd_write_span(scan, hlut, end, scan_bits, st, idx0, blk_count, true, sink, tbad, GjdRingSub(ny));
            else gjd_write_span(scan, hlut, end, scan_bits, st, idx0, blk_count, true, sink, tbad);
            for (int k = 0; k < 64; k++) if (sink.stage[k]) tbad |= GJD_DEV_BADSTATE;   /* staging must be left clean */
            status |= tbad;
            excl += sn[a] >> GJD_POS_BITS;
        }
    }
    }   /* the segment */
    for (uint32_t bk = 0; bk < bpm * mcu_count; bk++) if (written[bk] != 1) status |= GJD_DEV_BADCOUNT;
    if (badcode) status |= GJD_DEV_BADCODE;
    /* k_dc_tilebase + k_transform */
    int q[3][64];
    for (int c = 0; c < 3; c++)
        for (int k = 0; k < 64; k++) q[c][k] = p.quant[p.qsel[c]][k];
    int pred[3] = {0, 0, 0};
    /* k_transform_gray: one block per MCU, one DC chain, colour formula with Cb = Cr = 0 */
    for (uint32_t bkn = 0; gray && bkn < mcu_count; bkn++) {
        if (ri && bkn % ri == 0) pred[0] = 0;
        uint32_t m = meta[bkn];
        uint32_t kmax = (m >> 16) & 63u;
        uint32_t nsec = kmax ? (kmax >> 4) + 1u : 0u;
        int zq[64] = {0};
        for (uint32_t k = 0; k < nsec * 16u; k++) zq[k] = coef[(size_t)bkn * 64 + k];
        pred[0] += (int)(int16_t)(m & 0xffffu);
        int v[64];
        for (int i = 0; i < 64; i++) v[i] = zq[kZZ[i]] * q[0][kZZ[i]];
        v[0] = pred[0] * q[0][0];
        if (coef_out)
            for (int k = 0; k < 64; k++) coef_out[(size_t)bkn * 64 + k] = (int16_t)(k == 0 ? pred[0] : zq[k]);
        gjd_idct8x8(v);
        if (rgb_out) {
            uint32_t by = bkn / mcus_x, bx = bkn % mcus_x;
            for (int r = 0; r < 8; r++)
                for (int x = 0; x < 8; x++) {
                    uint32_t py = by * 8 + r, px = bx * 8 + x;
                    if (py >= H || px >= W) continue;
                    int R, G, B;
                    gjd_color(v[8 * r + x], 256, 256, R, G, B);
                    size_t o = (size_t)py * W + px;
                    rgb_out[o] = (uint8_t)R;
                    rgb_out[(size_t)W * H + o] = (uint8_t)G;
                    rgb_out[2 * (size_t)W * H + o] = (uint8_t)B;
                }
        }
    }
    /* k_dc_sub + k_transform_sub: ny luma blocks + Cb + Cr per MCU, chroma samples replicated over hs x vs pixels */
    for (uint32_t mcu = 0; subsamp && mcu < mcu_count; mcu++) {
        int smp[6][64];
        if (ri && mcu % ri == 0) pred[0] = pred[1] = pred[2] = 0;
        for (uint32_t bpos = 0; bpos < bpm; bpos++) {
            const uint32_t c = bpos < ny ? 0u : bpos - ny + 1u;
            size_t bk = (size_t)mcu * bpm + bpos;
            uint32_t m = meta[bk];
            uint32_t kmax = (m >> 16) & 63u;
            uint32_t nsec = kmax ? (kmax >> 4) + 1u : 0u;
            int zq[64] = {0};
            for (uint32_t k = 0; k < nsec * 16u; k++) zq[k] = coef[bk * 64 + k];
            pred[c] += (int)(int16_t)(m & 0xffffu);
            int v[64];
            for (int i = 0; i < 64; i++) v[i] = zq[kZZ[i]] * q[c][kZZ[i]];
            v[0] = pred[c] * q[c][0];
            if (coef_out)
                for (int k = 0; k < 64; k++) coef_out[bk * 64 + k] = (int16_t)(k == 0 ? pred[c] : zq[k]);
            gjd_idct8x8(v);
            memcpy(smp[bpos], v, sizeof(v));
        }
        if (rgb_out) {
            uint32_t my = mcu / mcus_x, mx = mcu % mcus_x;
            for (uint32_t yy = 0; yy < 8 * vs; yy++)
                for (uint32_t xx = 0; xx < 8 * hs; xx++) {
                    uint32_t py = my * 8 * vs + yy, px = mx * 8 * hs + xx;
                    if (py >= H || px >= W) continue;
                    const int* yb = smp[(yy / 8) * hs + xx / 8];
                    int R, G, B;
                    gjd_color(yb[(yy % 8) * 8 + xx % 8], smp[ny][(yy / vs) * 8 + xx / hs], smp[ny + 1][(yy / vs) * 8 + xx / hs], R, G, B);
                    size_t o = (size_t)py * W + px;
                    rgb_out[o] = (uint8_t)R;
                    rgb_out[(size_t)W * H + o] = (uint8_t)G;
                    rgb_out[2 * (size_t)W * H + o] = (uint8_t)B;
                }
        }
    }
    for (uint32_t mcu = 0; !gray && !subsamp && mcu < mcu_count; mcu++) {
        uint32_t samples[3][32];   /* as k_transform parks them: two biased samples per word, row r = words 4r..4r+3 */
        if (ri && mcu % ri == 0) pred[0] = pred[1] = pred[2] = 0;   /* T.81 E.1.4 */
        for (int c = 0; c < 3; c++) {
            size_t bk = (size_t)mcu * 3 + c;
            uint32_t m = meta[bk];
            uint32_t kmax = (m >> 16) & 63u;
            uint32_t nsec = kmax ? (kmax >> 4) + 1u : 0u;
            int zq[64] = {0};
            for (uint32_t k = 0; k < nsec * 16u; k++) zq[k] = coef[bk * 64 + k];
            pred[c] += (int)(int16_t)(m & 0xffffu);
            int v[64];
            for (int i = 0; i < 64; i++) v[i] = zq[kZZ[i]] * q[c][kZZ[i]];
            v[0] = pred[c] * q[c][0];
            if (coef_out)
                for (int k = 0; k < 64; k++) coef_out[(size_t)mcu * 192 + c * 64 + k] = (int16_t)(k == 0 ? pred[c] : zq[k]);
            gjd_idct8x8_raw(v);
            for (int j = 0; j < 32; j++) samples[c][j] = gjd_pack_samples(v[2 * j + 1], v[2 * j]);
        }
        if (rgb_out) {
            uint32_t by = mcu / mcus_x, bx = mcu % mcus_x;
            for (int r = 0; r < 8; r++) {
                uint32_t pr[2], pg[2], pb[2];
                gjd_color_row8(&samples[0][4 * r], &samples[1][4 * r], &samples[2][4 * r], pr, pg, pb);
                for (int x = 0; x < 8; x++) {
                    uint32_t py = by * 8 + r, px = bx * 8 + x;
                    if (py >= H || px >= W) continue;
                    size_t o = (size_t)py * W + px;
                    rgb_out[o] = (uint8_t)(pr[x >> 2] >> (8 * (x & 3)));
                    rgb_out[(size_t)W * H + o] = (uint8_t)(pg[x >> 2] >> (8 * (x & 3)));
                    rgb_out[2 * (size_t)W * H + o] = (uint8_t)(pb[x >> 2] >> (8 * (x & 3)));
                }
            }
        }
    }
    if (stats) {
        stats[0] = max_rounds; stats[1] = fix_decodes; stats[2] = status; stats[3] = nosync;
        stats[4] = sub_total; stats[5] = seq_total; stats[6] = redec; stats[7] = redec1; stats[8] = chk_hits;
    }
    free(lut);
    free(lutS);
    return status ? GJD_E_CORRUPT : GJD_OK;
}

/* gjd_color on unbiased inputs, all (y, cb, cr) in [-256,255]^3 with y fixed: out[cb+256][cr+256][3] */
void hostsim_color_plane(int y, uint8_t* out) {
    for (int cb = -256; cb < 256; cb++)
        for (int cr = -256; cr < 256; cr++) {
            int R, G, B;
            gjd_color(y + 256, cb + 256, cr + 256, R, G, B);
            uint8_t* o = out + ((size_t)(cb + 256) * 512 + (cr + 256)) * 3;
            o[0] = (uint8_t)R; o[1] = (uint8_t)G; o[2] = (uint8_t)B;
        }
}

/* the same plane through gjd_color_row8, the packed 8-pixels-per-call form k_transform runs */
void hostsim_color_plane_row8(int y, uint8_t* out) {
    for (int cb = -256; cb < 256; cb++)
        for (int cr0 = -256; cr0 < 256; cr0 += 8) {
            uint32_t yw[4], bw[4], rw[4], r[2], g[2], b[2];
            for (int j = 0; j < 4; j++) {
                yw[j] = (uint32_t)(y + 256) * 0x10001u;
                bw[j] = (uint32_t)(cb + 256) * 0x10001u;
                rw[j] = (uint32_t)(cr0 + 2 * j + 256) | ((uint32_t)(cr0 + 2 * j + 1 + 256) << 16);
            }
            gjd_color_row8(yw, bw, rw, r, g, b);
            for (int x = 0; x < 8; x++) {
                uint8_t* o = out + ((size_t)(cb + 256) * 512 + (cr0 + x + 256)) * 3;
                o[0] = (uint8_t)(r[x >> 2] >> (8 * (x & 3))); o[1] = (uint8_t)(g[x >> 2] >> (8 * (x & 3))); o[2] = (uint8_t)(b[x >> 2] >> (8 * (x & 3)));
            }
        }
}

/* dequantised natural-order block in, clip(.) samples (unbiased) out */
void hostsim_idct(const int32_t* in, int32_t* out, size_t nblocks) {
    for (size_t b = 0; b < nblocks; b++) {
        int v[64];
        for (int i = 0; i < 64; i++) v[i] = in[b * 64 + i];
        gjd_idct8x8(v);
        for (int i = 0; i < 64; i++) out[b * 64 + i] = v[i] - 256;
    }
}

/* the same through gjd_idct8x8_raw + gjd_pack_samples (the form k_transform runs) */
void hostsim_idct_packed(const int32_t* in, int32_t* out, size_t nblocks) {
    for (size_t b = 0; b < nblocks; b++) {
        int v[64];
        for (int i = 0; i < 64; i++) v[i] = in[b * 64 + i];
        gjd_idct8x8_raw(v);
        for (int j = 0; j < 32; j++) {
            const uint32_t w = gjd_pack_samples(v[2 * j + 1], v[2 * j]);
            out[b * 64 + 2 * j] = (int)(w & 0xffffu) - 256;
            out[b * 64 + 2 * j + 1] = (int)(w >> 16) - 256;
        }
    }
}

int hostsim_parse(const uint8_t* jpg, size_t len, int32_t* w, int32_t* h) {
    GjdParsed p;
    int rc = gjd_parse_headers(jpg, len, &p);
    if (w) *w = p.width;
    if (h) *h = p.height;
    return rc;
}

/* Every 16-bit window against every ring table of the image's DHTs: the two-level LUT (gjd_lookup) must
 * give the entry the canonical tables give (gjd_resolve from length 1).  out[0] = windows that differ,
 * out[1] = windows that are assigned codes yet miss both LUT levels (served by gjd_lookup_slow),
 * out[2] = windows with a code longer than GJD_LUT_BITS. */
int hostsim_lut_check(const uint8_t* jpg, size_t len, uint32_t out[3]) {
    GjdParsed p;
    int rc = gjd_parse_headers(jpg, len, &p);
    if (rc) return rc;
    GjdLutSet* lut = (GjdLutSet*)calloc(1, sizeof(GjdLutSet));
    for (int t = 0; t < 4; t++) {
        gjd_canonical(p.dht.bits[t], lut->lim[t], lut->off[t]);
        memcpy(lut->vals[t], p.dht.vals[t], 256);
        lut->isdc[t] = p.dht.isdc[t];
    }
    out[0] = out[1] = out[2] = 0;
    for (uint32_t r = 0; r < GJD_RING_TABLES; r++) {
        uint32_t t = gjd_ring_slot(p.tabsel, r);
        lut->slot_of[r] = (uint8_t)t;
        for (uint32_t i = 0; i < GJD_LUT_SIZE; i++)
            lut->lut[r][i] = gjd_lut_entry(lut->lim[t], lut->off[t], lut->vals[t], lut->isdc[t], i);
        for (uint32_t i = 0; i < GJD_LUT2_SIZE; i++)
            lut->lut2[r][i] = gjd_lut2_entry(lut->lim[t], lut->off[t], lut->vals[t], lut->isdc[t], i);
        GjdLutMem m;
        uint32_t badcode = 0;
        m.set = lut; m.badcode = &badcode;
        for (uint32_t win = 0; win < 65536u; win++) {
            const uint32_t w = win << 16, tb = r * GJD_RING_STRIDE;
            uint32_t want = gjd_resolve(lut->lim[t], lut->off[t], lut->vals[t], lut->isdc[t], win, 1);
            if (want & GJD_ENTRY_BAD) want = GJD_ENTRY_UNASSIGNED;
            if (gjd_lookup(m, tb, w, 64u) != want) out[0]++;
            uint32_t e = m.entry(tb + ((w >> (31 - GJD_LUT_BITS)) & (GJD_RING_STRIDE - 2u)));
            if ((e & 31u) == 0u) {
                if (want != GJD_ENTRY_UNASSIGNED) out[2]++;
                uint32_t e2 = e ? m.entry(GJD_LUT2_OFF(tb, e, w)) : 0u;
                if ((e2 & 31u) == 0u && want != GJD_ENTRY_UNASSIGNED) out[1]++;
            }
        }
    }
    free(lut);
    return 0;
}

/* gjd_count_span_pairs against gjd_count_span: from every subsequence start of every segment, both from the
 * true state (chained single-symbol decode) and from a cold state `lookback` bits earlier, for spans of
 * 2^sub_log2 bits.  out[0] = spans compared, out[1] = spans whose exit state (p, n, pos) differs,
 * out[2] = symbols decoded by the single-symbol walk, out[3] = loop trips of the pair walk. */
static unsigned long long g_trips = 0;
int hostsim_pair_check(const uint8_t* jpg, size_t len, int sub_log2, uint32_t lookback, unsigned long long out[4]) {
    GjdParsed p;
    int rc = gjd_parse_headers(jpg, len, &p);
    if (rc) return rc;
    const size_t raw = (size_t)(p.file_end - p.scan_begin);
    std::vector<uint8_t> scanbuf(raw + raw / 2 + 128, 0);
    std::vector<uint32_t> seg_off, seg_len;
    long ulen = gjd_unstuff_segments(p.scan_begin, p.file_end, scanbuf.data(), &seg_off, &seg_len);
    if (ulen < 0) return (int)ulen;
    std::vector<uint32_t> scanw((scanbuf.size() + 3) / 4 + 4, 0);
    memcpy(scanw.data(), scanbuf.data(), scanbuf.size());
    GjdLutSet* lut = (GjdLutSet*)calloc(1, sizeof(GjdLutSet));
    GjdLutSet* lutS = (GjdLutSet*)calloc(1, sizeof(GjdLutSet));
    for (int t = 0; t < 4; t++) {
        gjd_canonical(p.dht.bits[t], lut->lim[t], lut->off[t]);
        memcpy(lut->vals[t], p.dht.vals[t], 256);
        lut->isdc[t] = p.dht.isdc[t];
    }
    for (uint32_t r = 0; r < GJD_RING_TABLES; r++) {
        uint32_t t = gjd_ring_slot(p.tabsel, r);
        lut->slot_of[r] = (uint8_t)t;
        for (uint32_t i = 0; i < GJD_LUT_SIZE; i++)
            lut->lut[r][i] = gjd_lut_entry(lut->lim[t], lut->off[t], lut->vals[t], lut->isdc[t], i);
        for (uint32_t i = 0; i < GJD_LUT2_SIZE; i++)
            lut->lut2[r][i] = gjd_lut2_entry(lut->lim[t], lut->off[t], lut->vals[t], lut->isdc[t], i);
    }
    memcpy(lutS, lut, sizeof(GjdLutSet));
    for (uint32_t r = 0; r < GJD_RING_TABLES; r++) {
        uint32_t t = lut->slot_of[r];
        for (uint32_t i = 0; i < GJD_LUT_SIZE; i++)
            lutS->lut[r][i] = gjd_lutc_entry(lut->lim[t], lut->off[t], lut->vals[t], lut->isdc[t], i);
    }
    GjdLutMem one, two;
    one.set = lut; one.badcode = nullptr;
    two.set = lutS; two.badcode = nullptr;
    const uint32_t S = 1u << sub_log2;
    out[0] = out[1] = out[2] = out[3] = 0;
    {   /* the whole scan, its restart intervals inside it: both walks must make the same restart transitions */
        const size_t sg = 0;
        const uint32_t* scan = scanw.data();
        const uint32_t bits = (uint32_t)ulen * 8u;
        std::vector<uint32_t> rst_bits;
        for (size_t k = 1; k < seg_off.size(); k++) rst_bits.push_back(seg_off[k] * 8u);
        const GjdRstTable rstt(rst_bits.data(), (uint32_t)rst_bits.size());
        const uint32_t ny = p.components == 1 ? 1u : p.hs * p.vs;
        GjdState truth; truth.p = 0; truth.n = 0; truth.pos = 0;
        for (uint32_t start = 0; start < bits; start += S) {
            const uint32_t end = std::min(start + S, bits);
            for (int cold = 0; cold < 2; cold++) {
                GjdState a, b;
                if (cold) { a.p = start > lookback ? start - lookback : 0u; a.n = 0; a.pos = 0; }
                else a = truth;
                b = a;
                g_syms_dbg = 0;
                if (ny > 1u) gjd_count_span(scan, one, end, bits, a, GjdRingSub(ny), rstt); else gjd_count_span(scan, one, end, bits, a, GjdRing444(), rstt);
                out[2] += g_syms_dbg;
                g_syms_dbg = 0;
                if (ny > 1u) gjd_count_span_pairs(scan, two, end, bits, b, GjdRingSub(ny), rstt); else gjd_count_span_pairs(scan, two, end, bits, b, GjdRing444(), rstt);
                out[3] += g_syms_dbg;
                out[0]++;
                if (a.p != b.p || a.n != b.n || a.pos != b.pos) { out[1]++; if (getenv("GJD_PAIR_DEBUG")) fprintf(stderr, "seg %zu start %u end %u bits %u cold %d: single p=%u n=%u pos=%u  pair p=%u n=%u pos=%u\n", sg, start, end, bits, cold, a.p, a.n, a.pos, b.p, b.n, b.pos); }
                if (!cold) truth = a;
            }
        }
    }
    (void)g_trips;
    free(lut); free(lutS);
    return 0;
}

/* The device unstuffing (k_unstuff_count / k_unstuff_scan / k_unstuff_write, gjd_kernels.cu) stepped on the CPU with the
 * product's own lane functions: the scan bytes of `jpg` are placed at arena offset 32 + misalign, classified 16 bytes per
 * "thread", 4 KB per "block", and compacted.  Returns the unstuffed length (bytes in out, which needs the raw length + 64)
 * or the negative gjd_status the kernels would report.  Must equal gjd_unstuff() for every input without DRI. */
long hostsim_unstuff_device(const uint8_t* jpg, size_t len, uint32_t misalign, uint8_t* out, uint32_t* rst_out, uint32_t* nrst_out) {
    GjdParsed p;
    int rc = gjd_parse_headers(jpg, len, &p);
    if (rc) return rc;
    const bool dri = p.restart_interval != 0;
    const uint32_t raw_len = (uint32_t)(p.file_end - p.scan_begin);
    const uint64_t raw_off = 32 + (misalign & 15u);
    std::vector<uint8_t> arena(raw_off + raw_len + 64, 0xff);       /* slack filled with FF: must never be looked at */
    memcpy(arena.data() + raw_off, p.scan_begin, raw_len);
    const uint64_t base16 = raw_off & ~(uint64_t)15;
    const uint32_t chunks = std::max<uint32_t>(1u, (uint32_t)(((raw_off & 15u) + raw_len + GJD_UCHUNK - 1) / GJD_UCHUNK));
    uint32_t eoi = 0xffffffffu, err = 0xffffffffu;
    std::vector<uint32_t> cnt(chunks, 0), rcnt(chunks, 0);
    for (uint32_t c = 0; c < chunks; c++)                            /* k_unstuff_count */
        for (uint32_t tid = 0; tid < 256; tid++) {
            const GjdUnstuffLane L = gjd_unstuff_lane_at(arena.data(), raw_off, raw_len, base16 + (uint64_t)c * GJD_UCHUNK + tid * 16u, raw_len, dri);
            for (uint32_t mk = L.m.mark; mk; mk &= mk - 1u) {
                const uint32_t i = (uint32_t)__builtin_ctz(mk);
                const uint32_t nb = i < 15u ? gjd_lane_byte(L, i + 1u) : L.next;
                const uint32_t pos = L.pos0 + i;
                if (nb == 0xd9u) eoi = std::min(eoi, pos);
                else err = std::min(err, gjd_unstuff_err_word(pos, nb));
            }
            cnt[c] += (uint32_t)__builtin_popcount(L.valid & ~L.m.drop);
            rcnt[c] += (uint32_t)__builtin_popcount(L.m.rst);
        }
    /* k_unstuff_scan */
    if (eoi == 0xffffffffu || (err != 0xffffffffu && (err >> 2) < eoi)) {
        const bool unsup = err != 0xffffffffu && (eoi == 0xffffffffu || (err >> 2) < eoi) && (err & 3u) == GJD_UERR_UNSUPPORTED;
        return unsup ? GJD_E_UNSUPPORTED : GJD_E_FORMAT;
    }
    const uint32_t ce = (uint32_t)(((raw_off & 15u) + eoi) / GJD_UCHUNK);
    std::vector<uint32_t> base(chunks, 0), rbase(chunks, 0);
    uint32_t carry = 0, rcarry = 0;
    for (uint32_t c = 0; c < ce; c++) { base[c] = carry; carry += cnt[c]; rbase[c] = rcarry; rcarry += rcnt[c]; }
    base[ce] = carry;
    rbase[ce] = rcarry;
    uint32_t total = carry, markers = rcarry;
    for (uint32_t tid = 0; tid < 256; tid++) {
        const GjdUnstuffLane L = gjd_unstuff_lane_at(arena.data(), raw_off, raw_len, base16 + (uint64_t)ce * GJD_UCHUNK + tid * 16u, eoi, dri);
        total += (uint32_t)__builtin_popcount(L.valid & ~L.m.drop);
        markers += (uint32_t)__builtin_popcount(L.m.rst);
    }
    /* k_unstuff_write */
    bool bad_cycle = false;
    uint32_t nrst = 0;
    for (uint32_t c = 0; c < chunks; c++) {
        const int64_t cstart = (int64_t)c * GJD_UCHUNK - (int64_t)(raw_off & 15u);
        if (cstart >= (int64_t)eoi) continue;
        uint32_t o = base[c], ord = rbase[c];
        for (uint32_t tid = 0; tid < 256; tid++) {
            const GjdUnstuffLane L = gjd_unstuff_lane_at(arena.data(), raw_off, raw_len, base16 + (uint64_t)c * GJD_UCHUNK + tid * 16u, eoi, dri);
            const uint32_t keep0 = L.valid & ~L.m.drop;
            for (uint32_t mk = L.m.rst; mk; mk &= mk - 1u, ord++) {
                const uint32_t i = (uint32_t)__builtin_ctz(mk);
                const uint32_t nb = i < 15u ? gjd_lane_byte(L, i + 1u) : L.next;
                if (((nb - 0xd0u) & 7u) != (ord & 7u)) bad_cycle = true;
                if (rst_out) rst_out[ord] = (o + (uint32_t)__builtin_popcount(keep0 & ((1u << i) - 1u))) * 8u;
                nrst = std::max(nrst, ord + 1u);
            }
            for (uint32_t keep = keep0; keep; keep &= keep - 1u) out[o++] = (uint8_t)gjd_lane_byte(L, (uint32_t)__builtin_ctz(keep));
        }
        if (c < ce && o != base[c + 1]) return -100;                 /* the count pass and the write pass disagree */
    }
    if (bad_cycle) return GJD_E_FORMAT;
    if (nrst != markers) return -101;
    if (nrst_out) *nrst_out = nrst;
    memset(out + total, 0, 16);
    return (long)total;
}

/* the host byte loop on the same input (what the device path has to reproduce); restart intervals: their start
 * positions in bits, like the device's table */
long hostsim_unstuff_host(const uint8_t* jpg, size_t len, uint8_t* out, uint32_t* rst_out, uint32_t* nrst_out) {
    GjdParsed p;
    int rc = gjd_parse_headers(jpg, len, &p);
    if (rc) return rc;
    if (!p.restart_interval) { if (nrst_out) *nrst_out = 0; return gjd_unstuff(p.scan_begin, p.file_end, out); }
    std::vector<uint32_t> so, sl;
    long m = gjd_unstuff_segments(p.scan_begin, p.file_end, out, &so, &sl);
    if (m >= 0) {
        for (size_t k = 1; k < so.size(); k++) if (rst_out) rst_out[k - 1] = so[k] * 8u;
        if (nrst_out) *nrst_out = (uint32_t)so.size() - 1u;
    }
    return m;
}

}  /* extern "C" */
