// Host marker walk + Huffman lookup-table construction (see jpeg_stream.h).
#include "jpeg_stream.h"

#include <string>

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace b200jpeg {

namespace {

const uint8_t kZigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20,
    13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46,
    53, 60, 61, 54, 47, 55, 62, 63};

inline unsigned be16(const uint8_t* p) { return (unsigned(p[0]) << 8) | p[1]; }

bool read_dqt(const uint8_t* s, size_t n, StreamHeader* hdr)
{
    size_t q = 0;
    while (q < n) {
        const int pq = s[q] >> 4, tq = s[q] & 15;
        ++q;
        if (tq > 3 || pq > 1)
            return false;
        const size_t need = pq ? 128 : 64;
        if (q + need > n)
            return false;
        for (int i = 0; i < 64; ++i) {
            hdr->qt[tq][kZigzag[i]] = pq ? uint16_t(be16(s + q + 2 * i)) : uint16_t(s[q + i]);
        }
        q += need;
        hdr->qt_present[tq] = true;
    }
    return true;
}

bool read_dht(const uint8_t* s, size_t n, StreamHeader* hdr)
{
    size_t q = 0;
    while (q < n) {
        if (q + 17 > n)
            return false;
        const int tc = s[q] >> 4, th = s[q] & 15;
        if (tc > 1 || th > 3)
            return false;
        HuffSpec& hs = hdr->huff[tc][th];
        int total = 0;
        hs.bits[0] = 0;
        for (int l = 1; l <= 16; ++l) {
            hs.bits[l] = s[q + l];
            total += s[q + l];
        }
        q += 17;
        if (total > 256 || q + size_t(total) > n)
            return false;
        std::memset(hs.vals, 0, sizeof(hs.vals));
        std::memcpy(hs.vals, s + q, size_t(total));
        q += size_t(total);
        // Kraft check: a canonical code must not overflow any length
        int code = 0;
        for (int l = 1; l <= 16; ++l) {
            code += hs.bits[l];
            if (code > (1 << l))
                return false;
            code <<= 1;
        }
        hs.present = true;
    }
    return true;
}

} // namespace

StreamStatus parse_stream(const uint8_t* d, size_t n, StreamHeader* hdr)
{
    for (int c = 0; c < 2; ++c)
        for (int t = 0; t < 4; ++t)
            hdr->huff[c][t].present = false;
    if (n < 4 || d[0] != 0xFF || d[1] != 0xD8)
        return kStreamNotJpeg;
    size_t p = 2;
    bool have_sof = false;
    StreamStatus verdict = kStreamOk;
    for (;;) {
        if (p + 2 > n || d[p] != 0xFF)
            return kStreamCorrupt;
        while (p < n && d[p] == 0xFF)
            ++p; // optional fill bytes before a marker (test/parsers/jpeg_test.cpp:361 exercises them)
        if (p >= n)
            return kStreamCorrupt;
        const int m = d[p++];
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7))
            continue;
        if (m == 0xD9)
            return kStreamCorrupt;
        if (p + 2 > n)
            return kStreamCorrupt;
        const size_t len = be16(d + p);
        if (len < 2 || p + len > n)
            return kStreamCorrupt;
        const uint8_t* s = d + p + 2;
        const size_t sl = len - 2;
        if (m == 0xDB) {
            if (!read_dqt(s, sl, hdr))
                return kStreamCorrupt;
        } else if (m == 0xC4) {
            if (!read_dht(s, sl, hdr))
                return kStreamCorrupt;
        } else if (m == 0xDD) {
            if (sl < 2)
                return kStreamCorrupt;
            hdr->restart_interval = int(be16(s));
        } else if (m == 0xE0) {
            if (sl >= 5 && std::memcmp(s, "JFIF", 5) == 0)
                hdr->has_jfif = true;
        } else if (m == 0xEE) {
            if (sl >= 12 && std::memcmp(s, "Adobe", 5) == 0)
                hdr->adobe_transform = s[11];
        } else if (m >= 0xC0 && m <= 0xCF && m != 0xC8 && m != 0xCC) {
            if (have_sof || sl < 6)
                return kStreamCorrupt;
            have_sof = true;
            hdr->sof = m;
            hdr->precision = s[0];
            hdr->height = int(be16(s + 1));
            hdr->width = int(be16(s + 3));
            hdr->ncomp = s[5];
            if (hdr->ncomp < 1 || hdr->ncomp > kMaxComp || sl < 6 + 3 * size_t(hdr->ncomp))
                return hdr->ncomp > kMaxComp ? kStreamUnsupported : kStreamCorrupt;
            for (int c = 0; c < hdr->ncomp; ++c) {
                hdr->comp_id[c] = s[6 + 3 * c];
                hdr->h[c] = s[7 + 3 * c] >> 4;
                hdr->v[c] = s[7 + 3 * c] & 15;
                hdr->tq[c] = s[8 + 3 * c];
                if (hdr->h[c] < 1 || hdr->h[c] > 4 || hdr->v[c] < 1 || hdr->v[c] > 4 || hdr->tq[c] > 3)
                    return kStreamCorrupt;
            }
            // SOF0 baseline and SOF1 extended-sequential Huffman decode identically at 8 bits (SOF1 merely
            // allows 4 tables per class and 12-bit samples; those are declined below / here)
            if (m == 0xC2 && hdr->precision == 8)
                verdict = kStreamProgressive; // parse_progressive lists its scans
            else if ((m != 0xC0 && m != 0xC1) || hdr->precision != 8)
                verdict = kStreamUnsupported; // lossless / arithmetic / 12-bit
            if (hdr->width == 0 || hdr->height == 0)
                verdict = kStreamUnsupported; // DNL-defined height
            if (verdict == kStreamProgressive && hdr->ncomp != 1 && hdr->ncomp != 3)
                verdict = kStreamUnsupported;
        } else if (m == 0xDA) {
            if (!have_sof || sl < 1)
                return kStreamCorrupt;
            const int ns = s[0];
            if (ns < 1 || ns > 4 || sl < 1 + 2 * size_t(ns) + 3)
                return kStreamCorrupt;
            hdr->nscans = 1;
            hdr->scan_offset = p + len;
            hdr->scan_length = n - hdr->scan_offset;
            if (verdict != kStreamOk)
                return verdict;
            if (ns != hdr->ncomp)
                return (ns == 1 && hdr->ncomp == 3) ? kStreamMultiScan : kStreamUnsupported; // one scan per component: parse_multiscan
            for (int c = 0; c < ns; ++c) {
                if (s[1 + 2 * c] != hdr->comp_id[c])
                    return kStreamUnsupported; // scan order differs from frame order
                hdr->td[c] = s[2 + 2 * c] >> 4;
                hdr->ta[c] = s[2 + 2 * c] & 15;
                if (hdr->td[c] > 1 || hdr->ta[c] > 1)
                    return kStreamUnsupported; // table ids 2 and 3 (legal in SOF1 only) are not carried to the device
            }
            break;
        }
        p += len;
    }
    if (hdr->ncomp != 1 && hdr->ncomp != 3)
        return kStreamUnsupported; // CMYK / YCCK / 2-component
    if (hdr->ncomp == 3) {
        // libjpeg's colour-space guess (jdapimin.c default_decompress_parms): components tagged R,G,B
        // or an Adobe transform of 0 mean "already RGB" -- not produced by the encoders in scope.
        if (hdr->adobe_transform == 0)
            return kStreamUnsupported;
        if (!hdr->has_jfif && hdr->adobe_transform < 0 && hdr->comp_id[0] == 'R' && hdr->comp_id[1] == 'G' && hdr->comp_id[2] == 'B')
            return kStreamUnsupported;
    }
    if (hdr->ncomp == 1) {
        hdr->h[0] = hdr->v[0] = 1; // a one-component scan is never interleaved: MCU == one block
    }
    hdr->hmax = hdr->vmax = 1;
    hdr->blocks_per_mcu = 0;
    for (int c = 0; c < hdr->ncomp; ++c) {
        if (hdr->h[c] > hdr->hmax)
            hdr->hmax = hdr->h[c];
        if (hdr->v[c] > hdr->vmax)
            hdr->vmax = hdr->v[c];
        hdr->blocks_per_mcu += hdr->h[c] * hdr->v[c];
    }
    if (hdr->blocks_per_mcu > kMaxBlocksPerMcu)
        return kStreamCorrupt;
    for (int c = 0; c < hdr->ncomp; ++c) {
        if (!hdr->qt_present[hdr->tq[c]] || !hdr->huff[0][hdr->td[c]].present || !hdr->huff[1][hdr->ta[c]].present)
            return kStreamCorrupt;
        if (hdr->h[c] == 3 || hdr->v[c] == 3 || hdr->hmax % hdr->h[c] || hdr->vmax % hdr->v[c])
            return kStreamUnsupported; // fractional upsampling ratios
    }
    hdr->mcus_x = (hdr->width + 8 * hdr->hmax - 1) / (8 * hdr->hmax);
    hdr->mcus_y = (hdr->height + 8 * hdr->vmax - 1) / (8 * hdr->vmax);
    if (hdr->scan_length == 0)
        return kStreamCorrupt;
    return kStreamOk;
}

StreamStatus parse_multiscan(const uint8_t* d, size_t n, StreamHeader* frame, StreamHeader* scans)
{
    StreamHeader cur; // tables and restart interval as they stand at the current position
    for (int c = 0; c < 2; ++c)
        for (int t = 0; t < 4; ++t)
            cur.huff[c][t].present = false;
    if (n < 4 || d[0] != 0xFF || d[1] != 0xD8)
        return kStreamNotJpeg;
    size_t p = 2;
    bool have_sof = false, seen[3] = {false, false, false};
    int n_seen = 0;
    while (n_seen < 3) {
        if (p + 2 > n || d[p] != 0xFF)
            return kStreamCorrupt;
        while (p < n && d[p] == 0xFF)
            ++p;
        if (p >= n)
            return kStreamCorrupt;
        const int m = d[p++];
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7))
            continue;
        if (m == 0xD9 || p + 2 > n)
            return kStreamCorrupt;
        const size_t len = be16(d + p);
        if (len < 2 || p + len > n)
            return kStreamCorrupt;
        const uint8_t* s = d + p + 2;
        const size_t sl = len - 2;
        if (m == 0xDB) {
            if (!read_dqt(s, sl, &cur))
                return kStreamCorrupt;
        } else if (m == 0xC4) {
            if (!read_dht(s, sl, &cur))
                return kStreamCorrupt;
        } else if (m == 0xDD) {
            if (sl < 2)
                return kStreamCorrupt;
            cur.restart_interval = int(be16(s));
        } else if (m == 0xE0) {
            if (sl >= 5 && std::memcmp(s, "JFIF", 5) == 0)
                cur.has_jfif = true;
        } else if (m == 0xEE) {
            if (sl >= 12 && std::memcmp(s, "Adobe", 5) == 0)
                cur.adobe_transform = s[11];
        } else if (m == 0xC0 || m == 0xC1) {
            if (have_sof || sl < 6 + 9 || s[0] != 8 || s[5] != 3)
                return have_sof ? kStreamCorrupt : kStreamUnsupported;
            have_sof = true;
            cur.sof = m;
            cur.precision = 8;
            cur.height = int(be16(s + 1));
            cur.width = int(be16(s + 3));
            cur.ncomp = 3;
            if (cur.width == 0 || cur.height == 0)
                return kStreamUnsupported;
            for (int c = 0; c < 3; ++c) {
                cur.comp_id[c] = s[6 + 3 * c];
                cur.h[c] = s[7 + 3 * c] >> 4;
                cur.v[c] = s[7 + 3 * c] & 15;
                cur.tq[c] = s[8 + 3 * c];
                if (cur.h[c] < 1 || cur.h[c] > 4 || cur.v[c] < 1 || cur.v[c] > 4 || cur.tq[c] > 3)
                    return kStreamCorrupt;
            }
            cur.hmax = std::max(cur.h[0], std::max(cur.h[1], cur.h[2]));
            cur.vmax = std::max(cur.v[0], std::max(cur.v[1], cur.v[2]));
        } else if (m >= 0xC2 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
            return kStreamUnsupported;
        } else if (m == 0xDA) {
            if (!have_sof || sl < 1 + 2 + 3)
                return kStreamCorrupt;
            if (s[0] != 1)
                return kStreamUnsupported; // a mix of interleaved and single-component scans
            int c = -1;
            for (int k = 0; k < 3; ++k)
                if (s[1] == cur.comp_id[k])
                    c = k;
            if (c < 0)
                return kStreamCorrupt;
            if (seen[c])
                return kStreamUnsupported;
            const int td = s[2] >> 4, ta = s[2] & 15;
            if (td > 3 || ta > 3 || !cur.huff[0][td].present || !cur.huff[1][ta].present || !cur.qt_present[cur.tq[c]])
                return kStreamCorrupt;
            StreamHeader& sc = scans[c];
            sc = StreamHeader();
            for (int cl = 0; cl < 2; ++cl)
                for (int t = 0; t < 4; ++t)
                    sc.huff[cl][t].present = false;
            sc.sof = cur.sof;
            sc.precision = 8;
            sc.ncomp = 1;
            sc.width = (cur.width * cur.h[c] + cur.hmax - 1) / cur.hmax;   // the component's own sample grid (T.81 A.1.1)
            sc.height = (cur.height * cur.v[c] + cur.vmax - 1) / cur.vmax;
            sc.comp_id[0] = cur.comp_id[c];
            sc.h[0] = sc.v[0] = 1;
            sc.hmax = sc.vmax = 1;
            sc.blocks_per_mcu = 1;
            sc.mcus_x = (sc.width + 7) / 8;
            sc.mcus_y = (sc.height + 7) / 8;
            sc.tq[0] = 0;
            std::memcpy(sc.qt[0], cur.qt[cur.tq[c]], sizeof(sc.qt[0]));
            sc.qt_present[0] = true;
            sc.td[0] = sc.ta[0] = 0; // the tables in force now, under ids of our own
            sc.huff[0][0] = cur.huff[0][td];
            sc.huff[1][0] = cur.huff[1][ta];
            sc.restart_interval = cur.restart_interval;
            sc.nscans = 1;
            sc.scan_offset = p + len;
            // the entropy-coded segment ends at the first marker that is neither a stuffed FF00, a fill FF nor RSTn
            size_t q = sc.scan_offset;
            for (;;) {
                const void* f = q < n ? std::memchr(d + q, 0xFF, n - q) : nullptr;
                if (!f) {
                    q = n;
                    break;
                }
                q = size_t(static_cast<const uint8_t*>(f) - d);
                if (q + 1 >= n) {
                    q = n;
                    break;
                }
                const int x = d[q + 1];
                if (x == 0x00 || (x >= 0xD0 && x <= 0xD7)) {
                    q += 2;
                } else if (x == 0xFF) {
                    q += 1;
                } else {
                    break;
                }
            }
            sc.scan_length = std::min(n, q + 2) - sc.scan_offset; // including the marker, which is what stops K0
            if (q <= sc.scan_offset)
                return kStreamCorrupt;
            seen[c] = true;
            ++n_seen;
            p = q;
            continue;
        }
        p += len;
    }
    *frame = cur;
    frame->nscans = 3;
    frame->scan_offset = frame->scan_length = 0;
    if (frame->adobe_transform == 0)
        return kStreamUnsupported;
    if (!frame->has_jfif && frame->adobe_transform < 0 && frame->comp_id[0] == 'R' && frame->comp_id[1] == 'G' && frame->comp_id[2] == 'B')
        return kStreamUnsupported;
    frame->blocks_per_mcu = 0;
    for (int c = 0; c < 3; ++c) {
        frame->blocks_per_mcu += frame->h[c] * frame->v[c];
        if (frame->h[c] == 3 || frame->v[c] == 3 || frame->hmax % frame->h[c] || frame->vmax % frame->v[c])
            return kStreamUnsupported;
    }
    if (frame->blocks_per_mcu > kMaxBlocksPerMcu)
        return kStreamCorrupt;
    frame->mcus_x = (frame->width + 8 * frame->hmax - 1) / (8 * frame->hmax);
    frame->mcus_y = (frame->height + 8 * frame->vmax - 1) / (8 * frame->vmax);
    return kStreamOk;
}

namespace {

// End of an entropy-coded segment that starts at `from`: the first marker that is neither a stuffed FF00, a fill FF nor
// RSTn (or the end of the data).
size_t segment_end(const uint8_t* d, size_t n, size_t from)
{
    size_t q = from;
    for (;;) {
        const void* f = q < n ? std::memchr(d + q, 0xFF, n - q) : nullptr;
        if (!f)
            return n;
        q = size_t(static_cast<const uint8_t*>(f) - d);
        if (q + 1 >= n)
            return n;
        const int x = d[q + 1];
        if (x == 0x00 || (x >= 0xD0 && x <= 0xD7))
            q += 2;
        else if (x == 0xFF)
            q += 1;
        else
            return q;
    }
}

int max_jpeg_scans()
{
    static const int v = [] {
        const char* e = std::getenv("NVIMGCODEC_MAX_JPEG_SCANS");
        const int x = e ? std::atoi(e) : 0;
        return x > 0 ? x : 256;
    }();
    return v;
}

} // namespace

int sof_marker_of(const uint8_t* d, size_t n)
{
    if (n < 4 || d[0] != 0xFF || d[1] != 0xD8)
        return 0;
    size_t p = 2;
    while (p + 4 <= n && d[p] == 0xFF) {
        while (p < n && d[p] == 0xFF)
            ++p;
        if (p >= n)
            return 0;
        const int m = d[p++];
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7))
            continue;
        if (m == 0xDA || m == 0xD9 || p + 2 > n)
            return 0;
        if (m >= 0xC0 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC)
            return m;
        const size_t len = be16(d + p);
        if (len < 2)
            return 0;
        p += len;
    }
    return 0;
}

StreamStatus parse_progressive(const uint8_t* d, size_t n, StreamHeader* frame, ProgStream* out)
{
    out->scans.clear();
    out->tables.clear();
    StreamHeader cur;
    for (int c = 0; c < 2; ++c)
        for (int t = 0; t < 4; ++t)
            cur.huff[c][t].present = false;
    if (n < 4 || d[0] != 0xFF || d[1] != 0xD8)
        return kStreamNotJpeg;
    size_t p = 2;
    bool have_sof = false;
    int coef_bits[kMaxComp][64]; // bits still missing per coefficient, -1 = never sent (jdphuff.c coef_bits)
    int table_of[2][4] = {{-1, -1, -1, -1}, {-1, -1, -1, -1}}; // tables already copied for the specs in force
    for (int c = 0; c < kMaxComp; ++c)
        for (int k = 0; k < 64; ++k) {
            coef_bits[c][k] = -1;
        }
    for (;;) {
        if (p + 2 > n)
            break; // no EOI: what was listed stands if the progression is complete
        if (d[p] != 0xFF)
            return kStreamCorrupt;
        while (p < n && d[p] == 0xFF)
            ++p;
        if (p >= n)
            break;
        const int m = d[p++];
        if (m == 0xD9)
            break;
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7))
            continue;
        if (p + 2 > n)
            return kStreamCorrupt;
        const size_t len = be16(d + p);
        if (len < 2 || p + len > n)
            return kStreamCorrupt;
        const uint8_t* s = d + p + 2;
        const size_t sl = len - 2;
        if (m == 0xDB) {
            if (!out->scans.empty())
                return kStreamUnsupported; // quantisation tables redefined between scans
            if (!read_dqt(s, sl, &cur))
                return kStreamCorrupt;
        } else if (m == 0xC4) {
            if (!read_dht(s, sl, &cur))
                return kStreamCorrupt;
            for (size_t q = 0; q < sl;) { // the tables this segment (re)defined need a fresh copy
                const int tc = s[q] >> 4, th = s[q] & 15;
                int total = 0;
                for (int l = 1; l <= 16; ++l)
                    total += s[q + l];
                table_of[tc][th] = -1;
                q += 17 + size_t(total);
            }
        } else if (m == 0xDD) {
            if (sl < 2)
                return kStreamCorrupt;
            cur.restart_interval = int(be16(s));
        } else if (m == 0xE0) {
            if (sl >= 5 && std::memcmp(s, "JFIF", 5) == 0)
                cur.has_jfif = true;
        } else if (m == 0xEE) {
            if (sl >= 12 && std::memcmp(s, "Adobe", 5) == 0)
                cur.adobe_transform = s[11];
        } else if (m == 0xC2) {
            if (have_sof || sl < 6)
                return kStreamCorrupt;
            have_sof = true;
            cur.sof = m;
            cur.precision = s[0];
            cur.height = int(be16(s + 1));
            cur.width = int(be16(s + 3));
            cur.ncomp = s[5];
            if (cur.precision != 8 || cur.width == 0 || cur.height == 0 || (cur.ncomp != 1 && cur.ncomp != 3))
                return kStreamUnsupported;
            if (sl < 6 + 3 * size_t(cur.ncomp))
                return kStreamCorrupt;
            for (int c = 0; c < cur.ncomp; ++c) {
                cur.comp_id[c] = s[6 + 3 * c];
                cur.h[c] = s[7 + 3 * c] >> 4;
                cur.v[c] = s[7 + 3 * c] & 15;
                cur.tq[c] = s[8 + 3 * c];
                if (cur.h[c] < 1 || cur.h[c] > 4 || cur.v[c] < 1 || cur.v[c] > 4 || cur.tq[c] > 3)
                    return kStreamCorrupt;
            }
            if (cur.ncomp == 1)
                cur.h[0] = cur.v[0] = 1;
            cur.hmax = cur.vmax = 1;
            cur.blocks_per_mcu = 0;
            for (int c = 0; c < cur.ncomp; ++c) {
                cur.hmax = std::max(cur.hmax, cur.h[c]);
                cur.vmax = std::max(cur.vmax, cur.v[c]);
                cur.blocks_per_mcu += cur.h[c] * cur.v[c];
            }
            if (cur.blocks_per_mcu > kMaxBlocksPerMcu)
                return kStreamCorrupt;
            for (int c = 0; c < cur.ncomp; ++c)
                if (cur.h[c] == 3 || cur.v[c] == 3 || cur.hmax % cur.h[c] || cur.vmax % cur.v[c])
                    return kStreamUnsupported;
            cur.mcus_x = (cur.width + 8 * cur.hmax - 1) / (8 * cur.hmax);
            cur.mcus_y = (cur.height + 8 * cur.vmax - 1) / (8 * cur.vmax);
        } else if (m >= 0xC0 && m <= 0xCF && m != 0xC8 && m != 0xCC) {
            return have_sof ? kStreamCorrupt : kStreamUnsupported;
        } else if (m == 0xDA) {
            if (!have_sof || sl < 1)
                return kStreamCorrupt;
            if (int(out->scans.size()) >= max_jpeg_scans())
                return kStreamUnsupported;
            const int ns = s[0];
            if (ns < 1 || ns > 4 || sl < 1 + 2 * size_t(ns) + 3)
                return kStreamCorrupt;
            ProgScan sc;
            sc.ncomp = ns;
            int td[kMaxComp] = {0, 0, 0, 0}, ta[kMaxComp] = {0, 0, 0, 0};
            for (int i = 0; i < ns; ++i) {
                int c = -1;
                for (int k = 0; k < cur.ncomp; ++k)
                    if (s[1 + 2 * i] == cur.comp_id[k])
                        c = k;
                td[i] = s[2 + 2 * i] >> 4;
                ta[i] = s[2 + 2 * i] & 15;
                if (c < 0 || td[i] > 3 || ta[i] > 3 || (i > 0 && c <= sc.comp[i - 1]))
                    return kStreamCorrupt;
                sc.comp[i] = c;
            }
            sc.ss = s[1 + 2 * ns];
            sc.se = s[2 + 2 * ns];
            sc.ah = s[3 + 2 * ns] >> 4;
            sc.al = s[3 + 2 * ns] & 15;
            // jdphuff.c start_pass_phuff_decoder
            bool bad = sc.ss == 0 ? sc.se != 0 : (sc.se < sc.ss || sc.se > 63 || ns != 1);
            bad = bad || (sc.ah != 0 && sc.ah - 1 != sc.al) || sc.al > 13;
            if (bad)
                return kStreamCorrupt;
            if (ns != 1 && ns != cur.ncomp)
                return kStreamUnsupported; // a scan that interleaves only some of the components
            for (int i = 0; i < ns; ++i) {
                const int c = sc.comp[i];
                if (!cur.qt_present[cur.tq[c]])
                    return kStreamCorrupt;
                if (sc.ss && coef_bits[c][0] < 0)
                    return kStreamUnsupported; // AC before DC (libjpeg warns and goes on)
                for (int k = sc.ss; k <= sc.se; ++k) {
                    if (sc.ah != (coef_bits[c][k] < 0 ? 0 : coef_bits[c][k]))
                        return kStreamUnsupported; // out-of-order refinement (libjpeg warns and goes on)
                    coef_bits[c][k] = sc.al;
                }
            }
            if (!(sc.ss == 0 && sc.ah != 0)) { // every scan type but DC refinement decodes Huffman symbols
                for (int i = 0; i < ns; ++i) {
                    const int cls = sc.ss ? 1 : 0, id = sc.ss ? ta[i] : td[i];
                    if (!cur.huff[cls][id].present)
                        return kStreamCorrupt;
                    if (table_of[cls][id] < 0) {
                        table_of[cls][id] = int(out->tables.size());
                        out->tables.push_back(cur.huff[cls][id]);
                    }
                    sc.table[i] = table_of[cls][id];
                }
            }
            sc.restart_interval = cur.restart_interval;
            if (ns == 1) {
                const int c = sc.comp[0];
                const int cw = (cur.width * cur.h[c] + cur.hmax - 1) / cur.hmax, ch = (cur.height * cur.v[c] + cur.vmax - 1) / cur.vmax;
                sc.wb = (cw + 7) / 8;
                sc.hb = (ch + 7) / 8;
                sc.mcus = sc.wb * sc.hb;
            } else {
                sc.mcus = cur.mcus_x * cur.mcus_y;
            }
            sc.offset = p + len;
            const size_t q = segment_end(d, n, sc.offset);
            sc.length = std::min(n, q + 2) - sc.offset;
            out->scans.push_back(sc);
            p = q;
            continue;
        }
        p += len;
    }
    if (!have_sof || out->scans.empty())
        return kStreamCorrupt;
    for (int c = 0; c < cur.ncomp; ++c)
        for (int k = 0; k < 64; ++k)
            if (coef_bits[c][k] != 0)
                return kStreamUnsupported; // incomplete progression
    if (cur.ncomp == 3) {
        if (cur.adobe_transform == 0)
            return kStreamUnsupported;
        if (!cur.has_jfif && cur.adobe_transform < 0 && cur.comp_id[0] == 'R' && cur.comp_id[1] == 'G' && cur.comp_id[2] == 'B')
            return kStreamUnsupported;
    }
    *frame = cur;
    frame->nscans = int(out->scans.size());
    frame->restart_interval = 0;
    frame->scan_offset = frame->scan_length = 0;
    return kStreamOk;
}

void build_prog_table(const HuffSpec& spec, ProgHuffDev* out)
{
    std::memset(out, 0, sizeof(*out));
    int code = 0, k = 0;
    for (int l = 1; l <= 16; ++l) {
        out->valoff[l] = k - code;
        const int cnt = spec.bits[l];
        for (int i = 0; i < cnt; ++i, ++code, ++k) {
            if (l <= 9) {
                const int first = code << (9 - l), span = 1 << (9 - l);
                for (int j = 0; j < span; ++j)
                    out->lut[first + j] = uint16_t((l << 8) | spec.vals[k]);
            }
        }
        out->maxcode[l] = cnt ? code - 1 : -1;
        code <<= 1;
    }
    out->maxcode[0] = -1;
    out->maxcode[17] = 0x7fffffff;
    std::memcpy(out->vals, spec.vals, 256);
}

// What a decoded symbol does, packed for the device (kernels.cu decode_subsequence):
//   bits [0,5)  code length + magnitude bits = how far the bit position advances (<= 31);
//               0 = not in the first-level table
//   bits [5,9)  number of magnitude bits that follow the code
//   bits [9,16) how far the zig-zag index advances: DC -> 1; AC run/size -> run + 1;
//               ZRL (0xF0) -> 16; EOB (0x00) -> 64 (ends the block)
uint16_t pack_symbol(bool is_dc, int len, int sym)
{
    const int size = sym & 15;
    int skip;
    if (is_dc)
        skip = 1;
    else if (size)
        skip = (sym >> 4) + 1;
    else
        skip = ((sym >> 4) == 15) ? 16 : 64;
    return uint16_t((len + size) | (size << 5) | (skip << 9));
}

void build_huff_table(const HuffSpec& spec, bool is_dc, HuffTableDev* out)
{
    std::memset(out, 0, sizeof(*out));
    int code = 0, k = 0;
    uint32_t first16 = 0x10000u;
    for (int l = 1; l <= 16; ++l) {
        out->valoff[l] = k - code;
        const int cnt = spec.present ? spec.bits[l] : 0;
        if (cnt && l > kLutBits && first16 == 0x10000u)
            first16 = uint32_t(code) << (16 - l);
        for (int i = 0; i < cnt; ++i, ++code, ++k) {
            out->sym[k] = pack_symbol(is_dc, l, spec.vals[k]);
            if (l <= kLutBits) {
                const int first = code << (kLutBits - l);
                const int span = 1 << (kLutBits - l);
                for (int j = 0; j < span; ++j)
                    out->lut[first + j] = out->sym[k];
            }
        }
        out->maxcode[l] = cnt ? code - 1 : -1;
        code <<= 1;
    }
    out->maxcode[0] = -1;
    out->maxcode[17] = 0x7fffffff;
    out->valoff[0] = 0;
    out->valoff[17] = 0;
    // second level: every 16-bit window >= first16 belongs to a long code (or is unassigned: entry 0)
    out->first16 = first16;
    out->n2 = 0;
    if (first16 < 0x10000u && 0x10000u - first16 <= uint32_t(kLut2Size)) {
        out->n2 = 0x10000u - first16;
        code = 0;
        k = 0;
        for (int l = 1; l <= 16; ++l) {
            const int cnt = spec.present ? spec.bits[l] : 0;
            for (int i = 0; i < cnt; ++i, ++code, ++k) {
                if (l <= kLutBits)
                    continue;
                const uint32_t left = uint32_t(code) << (16 - l);
                for (uint32_t j = 0; j < (1u << (16 - l)); ++j)
                    if (left + j >= first16 && left + j - first16 < out->n2)
                        out->lut2[left + j - first16] = out->sym[k];
            }
            code <<= 1;
        }
    }
}

// Canonical byte string of the four table specs an image uses (presence, code-length counts, the symbols
// those counts cover): equal strings <=> equal device tables, so it is the cache key itself -- a hash of it
// alone could be made to collide by a crafted DHT and would decode a neighbour with the wrong tables.
void huff_set_key(const StreamHeader& hdr, std::string* key)
{
    key->clear();
    for (int cls = 0; cls < 2; ++cls)
        for (int id = 0; id < 2; ++id) {
            const HuffSpec& s = hdr.huff[cls][id];
            key->push_back(s.present ? char(1) : char(0));
            if (!s.present)
                continue;
            int n = 0;
            for (int l = 1; l <= 16; ++l)
                n += s.bits[l];
            key->append(reinterpret_cast<const char*>(s.bits + 1), 16);
            key->append(reinterpret_cast<const char*>(s.vals), size_t(n > 256 ? 256 : n));
        }
}

void build_huff_set(const StreamHeader& hdr, HuffSetDev* out)
{
    for (int cls = 0; cls < 2; ++cls)
        for (int id = 0; id < 2; ++id)
            build_huff_table(hdr.huff[cls][id], cls == 0, &out->t[cls * 2 + id]);
}

} // namespace b200jpeg
