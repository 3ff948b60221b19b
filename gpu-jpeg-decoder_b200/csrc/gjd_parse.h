/*
 * gjd_parse.h — host-side JPEG container parse (the role of JPEGParser::extract,
 * cpp-decoder/src/parser.cpp:24-103, and extract(), cuda-decoder/src/parser.cu:360-471).
 */
#ifndef GJD_PARSE_H
#define GJD_PARSE_H

#include <stddef.h>
#include <stdint.h>
#include <vector>

#include "gjd_types.h"

struct GjdParsed {
    int32_t width = 0, height = 0;
    const uint8_t* scan_begin = nullptr;   /* first entropy-coded byte (after the SOS header)      */
    const uint8_t* file_end = nullptr;     /* one past the last byte of the file                    */
    GjdDhtSet dht;                         /* slots 0=DC0 1=DC1 2=AC0 3=AC1                        */
    uint8_t dht_seen[4] = {0, 0, 0, 0};
    uint8_t quant[4][64];                  /* by Tq, zigzag order as stored                         */
    uint8_t quant_seen[4] = {0, 0, 0, 0};
    uint8_t tabsel[8] = {0};               /* [c] DC slot, [4+c] AC slot of component c             */
    uint8_t qsel[4] = {0};                 /* [c] Tq of component c                                 */
    uint32_t restart_interval = 0;
    uint32_t components = 3;               /* 3 (Y, Cb, Cr) or 1 (grayscale)                         */
    uint32_t hs = 1, vs = 1;               /* luma sampling factors (1 or 2 each; chroma is always 1x1): 1x1 = 4:4:4,
                                              2x1 = 4:2:2, 1x2 = 4:4:0, 2x2 = 4:2:0                   */
};

/* Parse every marker segment up to and including the SOS header.  Returns a gjd_status.
 * A superset of what the reference parser accepts: APPn/COM/unknown segments are
 * skipped by length, DQT/DHT may hold several tables in any order, selectors in SOF0/SOS are
 * honoured; single-component (grayscale) frames and luma sampling factors of 1 or 2 (4:2:2, 4:4:0, 4:2:0) are accepted;
 * everything else outside the 8-bit baseline single-scan contract is rejected with
 * GJD_E_UNSUPPORTED instead of being mis-decoded.  DRI is accepted: restart_interval = MCUs per interval. */
int gjd_parse_headers(const uint8_t* jpg, size_t len, GjdParsed* out);

/* Copy the entropy-coded segment [src, end) to dst, removing the stuffed 0x00 after every 0xFF and
 * stopping at EOI (FF D9), exactly like the byte loop at cpp-decoder/src/parser.cpp:84-99 (and its
 * twin cuda-decoder/src/parser.cu:450-462).  dst needs (end - src) + 32 bytes (the copy runs 32 bytes at a time
 * and stores before it tests).  Returns the unstuffed
 * length, or a negative gjd_status (truncated scan, RSTn, any other marker inside the scan). */
long gjd_unstuff(const uint8_t* src, const uint8_t* end, uint8_t* dst);

/* The same with restart markers (DRI/RSTn, T.81 B.2.4.4 + E.1.4; outside the reference's parser, which
 * keeps FF Dn as data, cpp-decoder/src/parser.cpp:84-96): every RSTn ends a restart interval, is checked
 * against the 0..7 cycle and removed; the intervals follow each other without gaps in dst.
 * seg_off/seg_len receive the byte offset and length of every interval (one entry for a scan without
 * restarts).  dst needs (end - src) + 32 bytes.  Returns the bytes used in dst. */
long gjd_unstuff_segments(const uint8_t* src, const uint8_t* end, uint8_t* dst, std::vector<uint32_t>* seg_off,
                          std::vector<uint32_t>* seg_len);

#endif
