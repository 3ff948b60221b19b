/*
 * gjd_parse.cpp — host JPEG marker parse + scan unstuffing.  See gjd_parse.h.
 */
#include "gjd_parse.h"

#include <string.h>

#include <vector>

#include "../../include/gjd.h"

namespace {

inline uint32_t be16(const uint8_t* p) { return ((uint32_t)p[0] << 8) | p[1]; }

/* DQT payload: one or more (Pq|Tq, 64 bytes) tables. */
int parse_dqt(const uint8_t* p, size_t n, GjdParsed* o) {
    while (n > 0) {
        uint32_t pq = p[0] >> 4, tq = p[0] & 15;
        if (tq > 3) return GJD_E_FORMAT;
        if (pq != 0) return GJD_E_UNSUPPORTED;   /* 16-bit tables: not baseline 8-bit */
        if (n < 65) return GJD_E_FORMAT;
        memcpy(o->quant[tq], p + 1, 64);
        o->quant_seen[tq] = 1;
        p += 65;
        n -= 65;
    }
    return GJD_OK;
}

/* DHT payload: one or more (Tc|Th, BITS[16], HUFFVAL) tables. */
int parse_dht(const uint8_t* p, size_t n, GjdParsed* o) {
    while (n > 0) {
        if (n < 17) return GJD_E_FORMAT;
        uint32_t tc = p[0] >> 4, th = p[0] & 15;
        if (tc > 1 || th > 3) return GJD_E_FORMAT;
        uint32_t total = 0, code = 0;
        for (int i = 0; i < 16; i++) {
            total += p[1 + i];
            code = (code + p[1 + i]) << 1;
            if (code > (2u << (i + 1))) return GJD_E_FORMAT;   /* oversubscribed code space */
        }
        if (total > 256 || n < 17 + total) return GJD_E_FORMAT;
        if (th <= 1) {   /* baseline allows Th 0/1 only; other tables are ignored unless SOS selects them */
            int slot = (int)(tc * 2 + th);
            memcpy(o->dht.bits[slot], p + 1, 16);
            memset(o->dht.vals[slot], 0, 256);
            memcpy(o->dht.vals[slot], p + 17, total);
            o->dht.isdc[slot] = tc == 0;
            o->dht_seen[slot] = 1;
        }
        p += 17 + total;
        n -= 17 + total;
    }
    return GJD_OK;
}

}  // namespace

int gjd_parse_headers(const uint8_t* d, size_t len, GjdParsed* o) {
    if (!d || !o) return GJD_E_ARG;
    if (len < 4 || d[0] != 0xFF || d[1] != 0xD8) return GJD_E_FORMAT;
    memset(&o->dht, 0, sizeof(o->dht));
    o->file_end = d + len;
    uint8_t comp_id[3] = {0, 0, 0};
    bool have_sof = false;
    size_t pos = 2;
    for (;;) {
        /* next marker: FF, optional fill FFs, code */
        if (pos + 2 > len) return GJD_E_FORMAT;
        if (d[pos] != 0xFF) return GJD_E_FORMAT;
        while (pos + 1 < len && d[pos + 1] == 0xFF) pos++;
        if (pos + 2 > len) return GJD_E_FORMAT;
        uint32_t m = d[pos + 1];
        pos += 2;
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;   /* standalone */
        if (m == 0xD9) return GJD_E_FORMAT;                                /* EOI before SOS */
        if (pos + 2 > len) return GJD_E_FORMAT;
        uint32_t seg = be16(d + pos);
        if (seg < 2 || pos + seg > len) return GJD_E_FORMAT;
        const uint8_t* body = d + pos + 2;
        size_t blen = seg - 2;
        switch (m) {
            case 0xDB: { int rc = parse_dqt(body, blen, o); if (rc) return rc; break; }
            case 0xC4: { int rc = parse_dht(body, blen, o); if (rc) return rc; break; }
            case 0xC0: {
                if (have_sof) return GJD_E_FORMAT;
                if (blen < 6) return GJD_E_FORMAT;
                uint32_t prec = body[0], nf = body[5];
                o->height = (int32_t)be16(body + 1);
                o->width = (int32_t)be16(body + 3);
                if (prec != 8 || (nf != 3 && nf != 1)) return GJD_E_UNSUPPORTED;
                if (blen < 6 + 3 * nf) return GJD_E_FORMAT;
                if (o->width <= 0 || o->height <= 0) return GJD_E_UNSUPPORTED;   /* DNL-defined height */
                o->components = nf;
                for (uint32_t c = 0; c < nf; c++) {
                    comp_id[c] = body[6 + 3 * c];
                    /* luma 1x1, 2x1, 1x2 or 2x2 over 1x1 chroma; a single component has one block per MCU whatever its
                     * factors say (T.81 A.2.2) */
                    if (nf == 3) {
                        const uint32_t h = body[7 + 3 * c] >> 4, v = body[7 + 3 * c] & 15u;
                        if (c == 0) {
                            if (h < 1 || h > 2 || v < 1 || v > 2) return GJD_E_UNSUPPORTED;
                            o->hs = h;
                            o->vs = v;
                        } else if (h != 1 || v != 1) {
                            return GJD_E_UNSUPPORTED;
                        }
                    }
                    uint32_t tq = body[8 + 3 * c];
                    if (tq > 3) return GJD_E_FORMAT;
                    o->qsel[c] = (uint8_t)tq;
                }
                if (nf == 1) o->qsel[1] = o->qsel[2] = o->qsel[0];
                have_sof = true;
                break;
            }
            case 0xC1: case 0xC2: case 0xC3: case 0xC5: case 0xC6: case 0xC7:
            case 0xC9: case 0xCA: case 0xCB: case 0xCD: case 0xCE: case 0xCF:
                return GJD_E_UNSUPPORTED;   /* extended / progressive / lossless / arithmetic */
            case 0xDD:
                if (blen < 2) return GJD_E_FORMAT;
                o->restart_interval = be16(body);
                break;
            case 0xDA: {
                if (!have_sof) return GJD_E_FORMAT;
                if (blen < 1) return GJD_E_FORMAT;
                uint32_t ns = body[0];
                if (ns != o->components) return GJD_E_UNSUPPORTED;   /* non-interleaved / multi-scan */
                if (blen < 1 + 2 * ns + 3) return GJD_E_FORMAT;
                for (uint32_t c = 0; c < ns; c++) {
                    if (body[1 + 2 * c] != comp_id[c]) return GJD_E_UNSUPPORTED;
                    uint32_t td = body[2 + 2 * c] >> 4, ta = body[2 + 2 * c] & 15;
                    if (td > 1 || ta > 1) return GJD_E_UNSUPPORTED;
                    if (!o->dht_seen[td] || !o->dht_seen[2 + ta]) return GJD_E_FORMAT;
                    o->tabsel[c] = (uint8_t)td;
                    o->tabsel[4 + c] = (uint8_t)(2 + ta);
                }
                if (ns == 1) {   /* grayscale: the one table pair serves all three ring positions */
                    o->tabsel[1] = o->tabsel[2] = o->tabsel[0];
                    o->tabsel[5] = o->tabsel[6] = o->tabsel[4];
                }
                const uint8_t* sp = body + 1 + 2 * ns;
                if (sp[0] != 0 || sp[1] != 63 || sp[2] != 0) return GJD_E_UNSUPPORTED;
                for (uint32_t c = 0; c < 3; c++)
                    if (!o->quant_seen[o->qsel[c]]) return GJD_E_FORMAT;
                o->scan_begin = d + pos + seg;
                {
                    /* a block costs at least 2 bits (a 1-bit DC size code and a 1-bit EOB): a frame header that
                     * announces more blocks than the remaining bytes can hold is damaged — refused here so that it
                     * cannot ask for gigabytes of coefficient memory */
                    const uint32_t mw = 8u * o->hs, mh = 8u * o->vs;
                    const uint64_t mcus = (uint64_t)((o->width + mw - 1) / mw) * (uint64_t)((o->height + mh - 1) / mh);
                    const uint64_t blocks = mcus * (o->components == 1 ? 1u : o->hs * o->vs + 2u);
                    if (blocks * 2u > (uint64_t)(o->file_end - o->scan_begin) * 8u) return GJD_E_CORRUPT;
                }
                return GJD_OK;
            }
            default:
                break;   /* APPn, COM, anything else with a length: skipped */
        }
        pos += seg;
    }
}

#if defined(__SSE2__)
#include <emmintrin.h>
#endif

/* Copy [p, end) to o up to (not including) the first 0xFF; returns where it stopped.  One pass instead of
 * memchr + memcpy per ~256-byte run: 32 bytes per step, stored before they are tested, so up to 32 bytes past
 * the copied prefix are scribbled on (the destination's slack, see gjd_parse.h). */
static inline const uint8_t* copy_until_ff(const uint8_t* p, const uint8_t* end, uint8_t*& o) {
#if defined(__SSE2__)
    const __m128i ff = _mm_set1_epi8((char)0xFF);
    while (p + 32 <= end) {
        const __m128i a = _mm_loadu_si128((const __m128i*)p), b = _mm_loadu_si128((const __m128i*)(p + 16));
        _mm_storeu_si128((__m128i*)o, a);
        _mm_storeu_si128((__m128i*)(o + 16), b);
        const unsigned m = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(a, ff)) |
                           ((unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(b, ff)) << 16);
        if (m) {
            const unsigned k = (unsigned)__builtin_ctz(m);
            o += k;
            return p + k;
        }
        p += 32;
        o += 32;
    }
#endif
    while (p < end && *p != 0xFF) *o++ = *p++;
    return p;
}

long gjd_unstuff_segments(const uint8_t* p, const uint8_t* end, uint8_t* dst, std::vector<uint32_t>* seg_off,
                          std::vector<uint32_t>* seg_len) {
    uint8_t* o = dst;
    uint8_t* seg = dst;          /* start of the current restart interval */
    uint32_t expect = 0;         /* RSTn numbers cycle 0..7 (T.81 E.1.4) */
    if (seg_off) seg_off->clear();
    if (seg_len) seg_len->clear();
    for (;;) {
        if (p >= end) return GJD_E_FORMAT;
        const uint8_t* q = copy_until_ff(p, end, o);
        if (q + 1 >= end) return GJD_E_FORMAT;         /* no EOI: the reference runs off the buffer */
        uint8_t m = q[1];
        if (m == 0x00) {
            *o++ = 0xFF;
            p = q + 2;
        } else if (m == 0xD9) {
            if (seg_off) seg_off->push_back((uint32_t)(seg - dst));
            if (seg_len) seg_len->push_back((uint32_t)(o - seg));
            return (long)(o - dst);
        } else if (m == 0xFF) {
            p = q + 1;   /* fill byte */
        } else if (m >= 0xD0 && m <= 0xD7) {
            if (!seg_off) return GJD_E_UNSUPPORTED;          /* caller did not announce a restart interval */
            if ((uint32_t)(m - 0xD0) != expect) return GJD_E_FORMAT;
            expect = (expect + 1u) & 7u;
            seg_off->push_back((uint32_t)(seg - dst));
            if (seg_len) seg_len->push_back((uint32_t)(o - seg));
            seg = o;
            p = q + 2;
        } else {
            return GJD_E_FORMAT;
        }
    }
}

long gjd_unstuff(const uint8_t* p, const uint8_t* end, uint8_t* dst) {
    return gjd_unstuff_segments(p, end, dst, nullptr, nullptr);
}

/* ---- multi-GPU sharding: greedy LPT over padded pixel counts ------------------------ */
#include <algorithm>
#include <numeric>
#include <queue>
#include <vector>
#include "../../include/gjd.h"

extern "C" int gjd_partition(const uint64_t* weight, int32_t n, int32_t parts, int32_t* part_of, uint64_t* load) {
    if (n < 0 || parts < 1 || (n > 0 && (!weight || !part_of))) return GJD_E_ARG;
    std::vector<int32_t> order(n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return weight[a] > weight[b]; });
    /* min-heap on (load, part): the lightest part takes the next-heaviest image */
    typedef std::pair<uint64_t, int32_t> Slot;
    std::priority_queue<Slot, std::vector<Slot>, std::greater<Slot>> heap;
    for (int32_t p = 0; p < parts; p++) heap.push(Slot(0, p));
    std::vector<uint64_t> acc(parts, 0);
    for (int32_t i : order) {
        if (weight[i] == 0) { part_of[i] = -1; continue; }
        Slot s = heap.top();
        heap.pop();
        part_of[i] = s.second;
        acc[s.second] = s.first + weight[i];
        heap.push(Slot(acc[s.second], s.second));
    }
    if (load) std::copy(acc.begin(), acc.end(), load);
    return GJD_OK;
}

extern "C" int gjd_partition_weights(const gjd_item* items, int32_t n, uint64_t* weight) {
    if (n < 0 || (n > 0 && (!items || !weight))) return GJD_E_ARG;
    for (int32_t i = 0; i < n; i++) {
        GjdParsed p;
        weight[i] = 0;
        if (items[i].jpg && gjd_parse_headers(items[i].jpg, items[i].len, &p) == GJD_OK) {
            const uint32_t mw = 8u * p.hs, mh = 8u * p.vs;       /* MCU size: the padded frame is whole MCUs */
            weight[i] = (uint64_t)(mw * ((p.width + mw - 1) / mw)) * (uint64_t)(mh * ((p.height + mh - 1) / mh));
        }
    }
    return GJD_OK;
}

extern "C" int gjd_partition_costs(const gjd_item* items, int32_t n, uint32_t byte_weight, uint64_t* cost) {
    const int rc = gjd_partition_weights(items, n, cost);
    if (rc != GJD_OK) return rc;
    for (int32_t i = 0; i < n; i++)
        if (cost[i]) cost[i] += (uint64_t)byte_weight * (uint64_t)items[i].len;
    return GJD_OK;
}
