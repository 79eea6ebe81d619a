// Baseline JPEG decode on the GPU (the decode half of the input pipeline):
//     image = Image.open(path).convert("RGB")        ref:Stage1/train_projection_stage1.py:99, ref:Stage2/dataset.py:86-90
// i.e. what Pillow's libjpeg-turbo does for a baseline (SOF0 / SOF1 8-bit Huffman) file: entropy decode, de-quantise,
// the "islow" integer IDCT (jidctint.c), "fancy" triangle chroma up-sampling (jdsample.c h2v1 / h2v2), the JFIF
// YCbCr -> RGB integer conversion (jdcolor.c); grayscale files come out with R = G = B as `convert("RGB")` gives.
// Every stage is integer arithmetic with libjpeg's constants and rounding: the result is bit-exact with Pillow
// (tests/test_jpeg_gpu.py).
//
// Split of the work:
//   host  (pt_jpeg_plan / pt_jpeg_pack, run by the data-loader workers): marker parsing, Huffman look-up tables,
//         removal of the FF00 byte stuffing and cutting the entropy-coded data at restart markers — a linear scan
//         of the compressed bytes, no pixel work;
//   device (pt_jpeg_decode_batch): [1] Huffman decode, one warp per restart segment of every image of the batch in
//         ONE launch (a segment is inherently sequential: lane 0 walks the bit stream, the warp stages the tables
//         in shared memory; files without restart markers decode as one segment per image, the batch's images side
//         by side) -> int16 coefficients; [2] de-quantise + IDCT, 8 threads per block -> uint8 component planes;
//         [3] up-sample + colour convert -> packed uint8 RGB [H, W, 3], which pt_image_preprocess takes.
// Not handled (rejected with an error, nothing falls back): progressive / arithmetic / lossless / 12-bit files,
// CMYK / YCCK, multi-scan baseline files, sampling factors other than 1x1, 2x1, 2x2 chroma sub-sampling.
#include <algorithm>
#include <vector>

#include "pt_common.cuh"

namespace pt {

constexpr uint32_t kJpegMagic = 0x4a504754u;     // "JPGT"
constexpr int kLookBits = 9;

struct HuffDev {                   // one Huffman table as the device walks it (jdhuff.c d_derived_tbl, restated)
    uint16_t look[1 << kLookBits]; // 9-bit prefix -> (code length << 8) | symbol, 0 when the code is longer
    int32_t maxcode[18];           // largest code of length l (-1: none); [17] = sentinel
    int32_t valoff[17];            // huffval index of the first code of length l minus that code
    uint8_t huffval[256];
};

struct JpegImg {                   // per image, in the blob (host and device read it)
    int32_t width, height, ncomp;
    int32_t hs[3], vs[3], hmax, vmax;
    int32_t mcux, mcuy;            // MCUs per row / column
    int32_t bw[3], bh[3];          // blocks per row / column of each component, whole MCUs
    int32_t dw[3], dh[3];          // real (un-padded) sample columns / rows of each component
    int32_t qsel[3], dcsel[3], acsel[3];
    int32_t restart_interval, nseg, seg0;
    int32_t pad0;
    int64_t tables_off;            // blob offset of: uint16 quant[4][64] (natural order), HuffDev huff[8] (dc 0-3, ac 0-3)
    int64_t coef_off[3];           // workspace byte offsets: int16 [bh][bw][64]
    int64_t plane_off[3];          // workspace byte offsets: uint8 [bh*8][bw*8]
    int64_t rgb_off;               // destination byte offset: uint8 [height][width][3]
};

struct JpegSeg {                   // one restart segment
    int32_t img, first_mcu, n_mcu;
    int32_t mode;                  // 0: decoded by one warp front to back; 1: decoded in parallel sub-sequences
    int64_t byte_off, byte_len;    // blob offset (8-byte aligned) and length of the un-stuffed bit stream
    int32_t sub0, nsub;            // mode 1: first entry in the SubInfo array, sub-sequences of kSubBits bits
    int32_t total_blocks, pad;     // blocks the segment holds (n_mcu * blocks per MCU)
};

constexpr int kSubBits = 1024;     // sub-sequence length of the parallel Huffman decode
constexpr int kSubThreads = 256;   // sub-sequences synchronised inside one CTA
constexpr int64_t kParallelMinBytes = 8192;     // shorter segments take the one-warp path

struct JpegCta { int32_t seg, first_sub; };     // work list of the per-sub-sequence kernels

struct SubInfo {                   // device scratch, one per sub-sequence
    int32_t p_start, p_end;        // bit positions in the segment: first symbol decoded here / first symbol of the next
    int32_t st_start, st_end;      // parser state there: (block slot in the MCU << 8) | zig-zag index
    int32_t nblk, blk0;            // blocks completed inside; blocks completed before (exclusive scan)
};

struct JpegBlobHeader {
    uint32_t magic; int32_t n, total_segs, pad;
    int64_t blob_bytes, workspace_bytes, coef_bytes, rgb_bytes;
    int64_t imgs_off, segs_off;
    int64_t cta_off, pseg_off;     // JpegCta list, list of the mode-1 segment indices (capacities fixed by the plan)
    int32_t n_cta, n_pseg, total_subs, pad1;
    int64_t sub_off;               // workspace offset of the SubInfo array
};

static const uint8_t kZigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
                                    41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22,
                                    15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55,
                                    62, 63};

static inline int64_t up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

// ---- host: marker parsing ---------------------------------------------------------------------------------------
struct Parsed {
    JpegImg im;
    uint16_t quant[4][64];
    bool have_q[4];
    uint8_t bits[8][17];
    uint8_t vals[8][256];
    bool have_h[8];
    int64_t scan_begin, scan_end;     // entropy-coded data in the file: [begin, end)
};

static int parse_file(const uint8_t* f, int64_t n, int idx, Parsed& P) {
    memset(&P, 0, sizeof(P));
    PT_REQUIRE(f != nullptr && n >= 4 && f[0] == 0xFF && f[1] == 0xD8, "jpeg[%d]: not a JPEG file (no SOI marker)", idx);
    int64_t p = 2;
    bool have_sof = false, have_sos = false;
    int adobe_transform = -1;
    int comp_id[3] = {0, 0, 0};
    while (p + 4 <= n && !have_sos) {
        PT_REQUIRE(f[p] == 0xFF, "jpeg[%d]: marker expected at byte %lld", idx, (long long)p);
        while (p < n && f[p] == 0xFF) ++p;                 // fill bytes
        PT_REQUIRE(p < n, "jpeg[%d]: truncated", idx);
        const int m = f[p++];
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
        PT_REQUIRE(m != 0xD9, "jpeg[%d]: EOI before any scan", idx);
        PT_REQUIRE(p + 2 <= n, "jpeg[%d]: truncated", idx);
        const int64_t len = (f[p] << 8) | f[p + 1];
        PT_REQUIRE(len >= 2 && p + len <= n, "jpeg[%d]: bad segment length", idx);
        const uint8_t* s = f + p + 2;
        const int64_t sl = len - 2;
        if (m == 0xC0 || m == 0xC1) {
            PT_REQUIRE(sl >= 6 && s[0] == 8, "jpeg[%d]: only 8-bit samples are handled", idx);
            P.im.height = (s[1] << 8) | s[2];
            P.im.width = (s[3] << 8) | s[4];
            P.im.ncomp = s[5];
            PT_REQUIRE(P.im.width > 0 && P.im.height > 0, "jpeg[%d]: zero image size (DNL files are not handled)", idx);
            PT_REQUIRE(P.im.ncomp == 1 || P.im.ncomp == 3, "jpeg[%d]: %d components (CMYK / YCCK are not handled)", idx, P.im.ncomp);
            PT_REQUIRE(sl >= 6 + 3 * P.im.ncomp, "jpeg[%d]: short SOF", idx);
            for (int c = 0; c < P.im.ncomp; ++c) {
                comp_id[c] = s[6 + 3 * c];
                P.im.hs[c] = s[7 + 3 * c] >> 4; P.im.vs[c] = s[7 + 3 * c] & 15;
                P.im.qsel[c] = s[8 + 3 * c];
                PT_REQUIRE(P.im.qsel[c] < 4, "jpeg[%d]: bad quantisation table index", idx);
            }
            have_sof = true;
        } else if (m >= 0xC2 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
            set_error("jpeg[%d]: SOF%d (progressive / lossless / arithmetic) files are not handled, baseline only", idx, m - 0xC0);
            return PT_ERR_INVALID;
        } else if (m == 0xDB) {
            int64_t q = 0;
            while (q < sl) {
                const int pq = s[q] >> 4, tq = s[q] & 15;
                PT_REQUIRE(tq < 4 && q + 1 + (pq ? 128 : 64) <= sl, "jpeg[%d]: bad DQT", idx);
                for (int k = 0; k < 64; ++k)
                    P.quant[tq][kZigzag[k]] = pq ? static_cast<uint16_t>((s[q + 1 + 2 * k] << 8) | s[q + 2 + 2 * k]) : s[q + 1 + k];
                P.have_q[tq] = true;
                q += 1 + (pq ? 128 : 64);
            }
        } else if (m == 0xC4) {
            int64_t q = 0;
            while (q < sl) {
                PT_REQUIRE(q + 17 <= sl, "jpeg[%d]: bad DHT", idx);
                const int tc = s[q] >> 4, th = s[q] & 15;
                PT_REQUIRE(tc < 2 && th < 4, "jpeg[%d]: bad Huffman table id", idx);
                const int t = tc * 4 + th;
                int cnt = 0;
                P.bits[t][0] = 0;
                for (int i = 1; i <= 16; ++i) { P.bits[t][i] = s[q + i]; cnt += s[q + i]; }
                PT_REQUIRE(cnt <= 256 && q + 17 + cnt <= sl, "jpeg[%d]: bad DHT counts", idx);
                memcpy(P.vals[t], s + q + 17, cnt);
                P.have_h[t] = true;
                q += 17 + cnt;
            }
        } else if (m == 0xDD) {
            PT_REQUIRE(sl >= 2, "jpeg[%d]: bad DRI", idx);
            P.im.restart_interval = (s[0] << 8) | s[1];
        } else if (m == 0xEE) {
            if (sl >= 12 && memcmp(s, "Adobe", 5) == 0) adobe_transform = s[11];
        } else if (m == 0xDA) {
            PT_REQUIRE(have_sof, "jpeg[%d]: SOS before SOF", idx);
            PT_REQUIRE(sl >= 1 && s[0] == P.im.ncomp && sl >= 1 + 2 * P.im.ncomp + 3,
                       "jpeg[%d]: a scan with %d of %d components (multi-scan files are not handled)", idx, sl >= 1 ? s[0] : 0, P.im.ncomp);
            for (int c = 0; c < P.im.ncomp; ++c) {
                PT_REQUIRE(s[1 + 2 * c] == comp_id[c], "jpeg[%d]: scan component order differs from the frame's", idx);
                P.im.dcsel[c] = s[2 + 2 * c] >> 4; P.im.acsel[c] = s[2 + 2 * c] & 15;
                PT_REQUIRE(P.im.dcsel[c] < 4 && P.im.acsel[c] < 4 && P.have_h[P.im.dcsel[c]] && P.have_h[4 + P.im.acsel[c]],
                           "jpeg[%d]: the scan names a Huffman table the file does not define", idx);
                PT_REQUIRE(P.have_q[P.im.qsel[c]], "jpeg[%d]: missing quantisation table", idx);
            }
            {   // the device stages two DC and two AC tables per image (luma-side / chroma-side)
                int nd = 1, na = 1;
                for (int c = 1; c < P.im.ncomp; ++c) {
                    bool sd = false, sa = false;
                    for (int k = 0; k < c; ++k) { sd |= P.im.dcsel[k] == P.im.dcsel[c]; sa |= P.im.acsel[k] == P.im.acsel[c]; }
                    nd += !sd; na += !sa;
                }
                PT_REQUIRE(nd <= 2 && na <= 2, "jpeg[%d]: three distinct Huffman tables of one class in a scan are not handled", idx);
            }
            have_sos = true;
            P.scan_begin = p + len;
        }
        p += len;
    }
    PT_REQUIRE(have_sos, "jpeg[%d]: no scan found", idx);
    PT_REQUIRE(!(P.im.ncomp == 3 && adobe_transform == 0), "jpeg[%d]: Adobe RGB (untransformed) files are not handled", idx);
    JpegImg& im = P.im;
    if (im.ncomp == 1) { im.hs[0] = im.vs[0] = 1; }       // a single-component scan is never interleaved
    im.hmax = im.vmax = 1;
    for (int c = 0; c < im.ncomp; ++c) { im.hmax = std::max(im.hmax, im.hs[c]); im.vmax = std::max(im.vmax, im.vs[c]); }
    for (int c = 0; c < im.ncomp; ++c) {
        const bool luma = c == 0;
        PT_REQUIRE(im.hs[c] >= 1 && im.vs[c] >= 1, "jpeg[%d]: zero sampling factor", idx);
        PT_REQUIRE(luma ? (im.hs[c] == im.hmax && im.vs[c] == im.vmax) : (im.hs[c] == 1 && im.vs[c] == 1),
                   "jpeg[%d]: sampling %dx%d of component %d is not handled (4:4:4, 4:2:2, 4:2:0 only)", idx, im.hs[c], im.vs[c], c);
    }
    PT_REQUIRE((im.hmax == 1 && im.vmax == 1) || (im.hmax == 2 && im.vmax <= 2),
               "jpeg[%d]: %dx%d chroma sub-sampling is not handled (4:4:4, 4:2:2, 4:2:0 only)", idx, im.hmax, im.vmax);
    im.mcux = (im.width + 8 * im.hmax - 1) / (8 * im.hmax);
    im.mcuy = (im.height + 8 * im.vmax - 1) / (8 * im.vmax);
    for (int c = 0; c < im.ncomp; ++c) {
        im.bw[c] = im.mcux * im.hs[c]; im.bh[c] = im.mcuy * im.vs[c];
        im.dw[c] = (im.width * im.hs[c] + im.hmax - 1) / im.hmax;
        im.dh[c] = (im.height * im.vs[c] + im.vmax - 1) / im.vmax;
    }
    const int64_t mcus = static_cast<int64_t>(im.mcux) * im.mcuy;
    im.nseg = im.restart_interval > 0 ? static_cast<int32_t>((mcus + im.restart_interval - 1) / im.restart_interval) : 1;
    // end of the entropy-coded data: the first marker that is neither a stuffed FF00 nor RSTn
    int64_t e = P.scan_begin;
    while (e + 1 < n) {
        if (f[e] == 0xFF && f[e + 1] != 0x00 && !(f[e + 1] >= 0xD0 && f[e + 1] <= 0xD7) && f[e + 1] != 0xFF) break;
        ++e;
    }
    P.scan_end = (e + 1 < n) ? e : n;
    return PT_OK;
}

// jdhuff.c jpeg_make_d_derived_tbl, restated: canonical codes from the (bits, huffval) lists
static int build_huff(const uint8_t bits[17], const uint8_t* vals, HuffDev& H, int idx) {
    memset(&H, 0, sizeof(H));
    int huffsize[257], huffcode[257], nsym = 0;
    for (int l = 1; l <= 16; ++l)
        for (int i = 0; i < bits[l]; ++i) huffsize[nsym++] = l;
    PT_REQUIRE(nsym <= 256, "jpeg[%d]: bad Huffman table", idx);
    huffsize[nsym] = 0;
    int code = 0, si = huffsize[0], p = 0;
    while (huffsize[p]) {
        while (huffsize[p] == si) huffcode[p++] = code++;
        PT_REQUIRE(code <= (1 << si), "jpeg[%d]: bad Huffman code lengths", idx);
        code <<= 1; ++si;
    }
    p = 0;
    for (int l = 1; l <= 16; ++l) {
        if (bits[l]) {
            H.valoff[l] = p - huffcode[p];
            p += bits[l];
            H.maxcode[l] = huffcode[p - 1];
        } else {
            H.maxcode[l] = -1;
        }
    }
    H.maxcode[17] = 0xFFFFF;
    p = 0;
    for (int l = 1; l <= kLookBits; ++l)
        for (int i = 0; i < bits[l]; ++i, ++p) {
            const int first = huffcode[p] << (kLookBits - l);
            for (int k = 0; k < (1 << (kLookBits - l)); ++k)
                H.look[first + k] = static_cast<uint16_t>((l << 8) | vals[p]);
        }
    memcpy(H.huffval, vals, nsym);
    return PT_OK;
}

constexpr int64_t kTablesBytes = 4 * 64 * 2 + 8 * static_cast<int64_t>(sizeof(HuffDev));

struct Layout {
    std::vector<Parsed> P;
    JpegBlobHeader h;
};

// Sizes and offsets of everything: blob = [header][JpegImg x n][JpegSeg x total][per image: tables, segment bytes]
static int layout_batch(const uint8_t* const* files, const int64_t* sizes, int n, Layout& L) {
    PT_REQUIRE(files && sizes && n > 0 && n <= 4096, "pt_jpeg: bad batch (n = %d)", n);
    L.P.resize(n);
    int64_t total_segs = 0;
    for (int i = 0; i < n; ++i) {
        const int rc = parse_file(files[i], sizes[i], i, L.P[i]);
        if (rc != PT_OK) return rc;
        L.P[i].im.seg0 = static_cast<int32_t>(total_segs);
        total_segs += L.P[i].im.nseg;
    }
    PT_REQUIRE(total_segs < (1ll << 30), "pt_jpeg: too many restart segments");
    JpegBlobHeader& h = L.h;
    memset(&h, 0, sizeof(h));
    h.magic = kJpegMagic; h.n = n; h.total_segs = static_cast<int32_t>(total_segs);
    int64_t o = up(sizeof(JpegBlobHeader), 16);
    h.imgs_off = o; o = up(o + static_cast<int64_t>(n) * sizeof(JpegImg), 16);
    h.segs_off = o; o = up(o + total_segs * sizeof(JpegSeg), 16);
    // the un-stuffed lengths are known only after packing: capacities from the stuffed lengths
    int64_t cap_subs = 0, cap_cta = 0;
    for (int i = 0; i < n; ++i) {
        const int64_t len = L.P[i].scan_end - L.P[i].scan_begin;
        const int64_t subs = len * 8 / kSubBits + 2 * L.P[i].im.nseg;
        cap_subs += subs;
        cap_cta += subs / kSubThreads + L.P[i].im.nseg;
    }
    PT_REQUIRE(cap_subs < (1ll << 31), "pt_jpeg: batch too large");
    h.cta_off = o; o = up(o + cap_cta * sizeof(JpegCta), 16);
    h.pseg_off = o; o = up(o + total_segs * 4, 16);
    int64_t ws = 0, rgb = 0;
    for (int i = 0; i < n; ++i) {
        JpegImg& im = L.P[i].im;
        im.tables_off = o; o = up(o + kTablesBytes, 16);
        // un-stuffed data is never longer than the stuffed data; every segment start is aligned to 8 bytes
        // (+ room behind the last byte: the parallel parse reads up to one sub-sequence past a segment's end)
        o = up(o + (L.P[i].scan_end - L.P[i].scan_begin) + 8ll * im.nseg + kSubBits / 8 + 64, 16);
        for (int c = 0; c < im.ncomp; ++c) { im.coef_off[c] = ws; ws = up(ws + 128ll * im.bw[c] * im.bh[c], 256); }
        im.rgb_off = rgb; rgb += 3ll * im.width * im.height;
    }
    h.coef_bytes = ws;                              // cleared before every decode: only non-zero coefficients are stored
    for (int i = 0; i < n; ++i) {
        JpegImg& im = L.P[i].im;
        for (int c = 0; c < im.ncomp; ++c) { im.plane_off[c] = ws; ws = up(ws + 64ll * im.bw[c] * im.bh[c], 256); }
    }
    h.sub_off = ws; ws = up(ws + cap_subs * static_cast<int64_t>(sizeof(SubInfo)), 256);
    h.blob_bytes = o; h.workspace_bytes = ws; h.rgb_bytes = rgb;
    return PT_OK;
}

}  // namespace pt

using namespace pt;

extern "C" int pt_jpeg_plan(const uint8_t* const* files, const int64_t* sizes, int n, int32_t* widths, int32_t* heights,
                            int64_t* rgb_offsets, int64_t* blob_bytes, int64_t* workspace_bytes) {
    PT_REQUIRE(widths && heights && rgb_offsets && blob_bytes && workspace_bytes, "pt_jpeg_plan: null output");
    Layout L;
    const int rc = layout_batch(files, sizes, n, L);
    if (rc != PT_OK) return rc;
    for (int i = 0; i < n; ++i) {
        widths[i] = L.P[i].im.width; heights[i] = L.P[i].im.height; rgb_offsets[i] = L.P[i].im.rgb_off;
    }
    rgb_offsets[n] = L.h.rgb_bytes;
    *blob_bytes = L.h.blob_bytes;
    *workspace_bytes = L.h.workspace_bytes;
    return PT_OK;
}

extern "C" int pt_jpeg_pack(const uint8_t* const* files, const int64_t* sizes, int n, uint8_t* blob, int64_t blob_bytes) {
    PT_REQUIRE(blob != nullptr && (reinterpret_cast<uintptr_t>(blob) & 15) == 0, "pt_jpeg_pack: blob must be 16-byte aligned");
    Layout L;
    int rc = layout_batch(files, sizes, n, L);
    if (rc != PT_OK) return rc;
    PT_REQUIRE(blob_bytes >= L.h.blob_bytes, "pt_jpeg_pack: blob too small (see pt_jpeg_plan)");
    memset(blob, 0, static_cast<size_t>(L.h.blob_bytes));
    memcpy(blob, &L.h, sizeof(L.h));
    JpegImg* imgs = reinterpret_cast<JpegImg*>(blob + L.h.imgs_off);
    JpegSeg* segs = reinterpret_cast<JpegSeg*>(blob + L.h.segs_off);
    JpegBlobHeader* hdr = reinterpret_cast<JpegBlobHeader*>(blob);
    JpegCta* ctas = reinterpret_cast<JpegCta*>(blob + L.h.cta_off);
    int32_t* pseg = reinterpret_cast<int32_t*>(blob + L.h.pseg_off);
    for (int i = 0; i < n; ++i) {
        const Parsed& P = L.P[i];
        const JpegImg& im = P.im;
        imgs[i] = im;
        uint8_t* t = blob + im.tables_off;
        memcpy(t, P.quant, sizeof(P.quant));
        HuffDev* H = reinterpret_cast<HuffDev*>(t + sizeof(P.quant));
        for (int k = 0; k < 8; ++k)
            if (P.have_h[k]) { rc = build_huff(P.bits[k], P.vals[k], H[k], i); if (rc != PT_OK) return rc; }
        // entropy-coded data: drop the zero after a stuffed FF, cut at RSTn, skip fill bytes
        const uint8_t* f = files[i];
        const int64_t mcus = static_cast<int64_t>(im.mcux) * im.mcuy;
        int64_t o = up(im.tables_off + kTablesBytes, 16);
        int seg = 0;
        int64_t start = o, q = P.scan_begin;
        auto close = [&](int64_t end_o) -> int {
            JpegSeg& s = segs[im.seg0 + seg];
            s.img = i;
            s.first_mcu = seg * (im.restart_interval > 0 ? im.restart_interval : 0);
            s.n_mcu = static_cast<int32_t>(im.restart_interval > 0
                                               ? std::min<int64_t>(im.restart_interval, mcus - s.first_mcu) : mcus);
            s.byte_off = start; s.byte_len = end_o - start;
            int bpm = 0;
            for (int c = 0; c < im.ncomp; ++c) bpm += im.hs[c] * im.vs[c];
            s.total_blocks = s.n_mcu * bpm;
            PT_REQUIRE(s.byte_len < (1ll << 27), "jpeg[%d]: entropy-coded segment too long", i);
            s.mode = s.byte_len >= kParallelMinBytes ? 1 : 0;
            if (s.mode) {
                s.nsub = static_cast<int32_t>((s.byte_len * 8 + kSubBits - 1) / kSubBits);
                s.sub0 = hdr->total_subs;
                hdr->total_subs += s.nsub;
                pseg[hdr->n_pseg++] = im.seg0 + seg;
                for (int fs = 0; fs < s.nsub; fs += kSubThreads) { ctas[hdr->n_cta].seg = im.seg0 + seg; ctas[hdr->n_cta++].first_sub = fs; }
            }
            return PT_OK;
        };
        while (q < P.scan_end) {
            const uint8_t b = f[q++];
            if (b != 0xFF) { blob[o++] = b; continue; }
            if (q >= P.scan_end) break;
            const uint8_t m = f[q];
            if (m == 0x00) { blob[o++] = 0xFF; ++q; }
            else if (m >= 0xD0 && m <= 0xD7) {
                ++q;
                PT_REQUIRE(seg + 1 < im.nseg, "jpeg[%d]: more restart markers than the restart interval allows", i);
                rc = close(o);
                if (rc != PT_OK) return rc;
                ++seg;
                o = up(o, 8); start = o;
            }                                           // else: a fill byte (FF FF): drop this FF, look at the next
        }
        rc = close(o);
        if (rc != PT_OK) return rc;
        PT_REQUIRE(seg + 1 == im.nseg, "jpeg[%d]: %d restart segments found, %d expected (truncated file?)", i, seg + 1, im.nseg);
    }
    return PT_OK;
}

// ---- device [1]: Huffman decode --------------------------------------------------------------------------------
namespace pt {

__constant__ uint8_t c_zigzag[64];

struct BitReader {
    const uint32_t* p;     // next aligned word of the un-stuffed stream
    const uint32_t* end;   // first word past the segment (+ slack): a corrupt stream reads zeros from there on
    uint64_t buf;          // left-aligned bit buffer
    int cnt;               // valid bits in buf
    __device__ __forceinline__ void refill() {          // after this, cnt >= 33
        if (cnt <= 32) {
            const uint32_t w = p < end ? __byte_perm(__ldg(p), 0, 0x0123) : 0u;      // big-endian bit order
            ++p;
            buf |= static_cast<uint64_t>(w) << (32 - cnt);
            cnt += 32;
        }
    }
    __device__ __forceinline__ uint32_t peek(int n) const { return static_cast<uint32_t>(buf >> (64 - n)); }
    __device__ __forceinline__ void skip(int n) { buf <<= n; cnt -= n; }
};

__device__ __forceinline__ int huff_symbol(BitReader& br, const HuffDev& H) {
    const uint32_t lk = H.look[br.peek(kLookBits)];
    if (lk) { br.skip(lk >> 8); return lk & 0xFF; }
    int l = kLookBits + 1;
    int code = br.peek(l);
    while (l <= 16 && code > H.maxcode[l]) { ++l; code = br.peek(l); }
    if (l > 16) { br.skip(16); return 0; }              // corrupt stream: libjpeg warns and yields a zero
    br.skip(l);
    return H.huffval[(code + H.valoff[l]) & 0xFF];
}

__device__ __forceinline__ int huff_extend(int v, int s) { return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v; }

// One warp per restart segment.  Lanes stage the four Huffman tables of the image in shared memory; lane 0 decodes
// the segment's MCUs in order and scatters the non-zero coefficients (natural order, int16) into the cleared
// coefficient arrays.
__global__ void __launch_bounds__(32)
jpeg_huffman_kernel(const uint8_t* __restrict__ blob, int64_t imgs_off, int64_t segs_off, uint8_t* __restrict__ ws) {
    __shared__ HuffDev sh[4];          // dc luma-side, dc chroma-side, ac luma-side, ac chroma-side as the scan names them
    __shared__ int sel_dc[3], sel_ac[3];
    const JpegSeg seg = reinterpret_cast<const JpegSeg*>(blob + segs_off)[blockIdx.x];
    if (seg.mode != 0) return;         // long segments: the parallel kernels below
    const JpegImg& im = reinterpret_cast<const JpegImg*>(blob + imgs_off)[seg.img];
    const HuffDev* tab = reinterpret_cast<const HuffDev*>(blob + im.tables_off + 4 * 64 * 2);
    // at most two distinct DC and two distinct AC tables are in use (luma / chroma)
    int dc_ids[2] = {im.dcsel[0], im.dcsel[0]}, ac_ids[2] = {im.acsel[0], im.acsel[0]};
    for (int c = 1; c < im.ncomp; ++c) {
        if (im.dcsel[c] != dc_ids[0]) dc_ids[1] = im.dcsel[c];
        if (im.acsel[c] != ac_ids[0]) ac_ids[1] = im.acsel[c];
    }
    {
        constexpr int W = sizeof(HuffDev) / 4;
        for (int t = 0; t < 4; ++t) {
            const uint32_t* src = reinterpret_cast<const uint32_t*>(tab + (t < 2 ? dc_ids[t] : 4 + ac_ids[t - 2]));
            uint32_t* dst = reinterpret_cast<uint32_t*>(&sh[t]);
            for (int i = threadIdx.x; i < W; i += 32) dst[i] = __ldg(src + i);
        }
        if (threadIdx.x < 3) {
            const int c = threadIdx.x < im.ncomp ? threadIdx.x : 0;
            sel_dc[threadIdx.x] = im.dcsel[c] == dc_ids[0] ? 0 : 1;
            sel_ac[threadIdx.x] = im.acsel[c] == ac_ids[0] ? 2 : 3;
        }
    }
    __syncwarp();
    if (threadIdx.x != 0) return;
    BitReader br;
    br.p = reinterpret_cast<const uint32_t*>(blob + seg.byte_off);
    br.end = br.p + (seg.byte_len + 3) / 4 + 2;
    br.buf = 0; br.cnt = 0;
    int pred[3] = {0, 0, 0};
    const int ncomp = im.ncomp;
    int mx = seg.first_mcu % im.mcux, my = seg.first_mcu / im.mcux;
    for (int m = 0; m < seg.n_mcu; ++m) {
        for (int c = 0; c < ncomp; ++c) {
            const HuffDev& Hd = sh[sel_dc[c]];
            const HuffDev& Ha = sh[sel_ac[c]];
            int16_t* coef = reinterpret_cast<int16_t*>(ws + im.coef_off[c]);
            const int hs = im.hs[c], vs = im.vs[c], bw = im.bw[c];
            for (int v = 0; v < vs; ++v)
                for (int h = 0; h < hs; ++h) {
                    int16_t* blk = coef + (static_cast<int64_t>(my * vs + v) * bw + (mx * hs + h)) * 64;
                    br.refill();
                    int s = min(huff_symbol(br, Hd), 16);          // (a damaged table may name any byte)
                    if (s) {
                        br.refill();
                        const int d = huff_extend(br.peek(s), s);
                        br.skip(s);
                        pred[c] += d;
                    }
                    if (pred[c]) blk[0] = static_cast<int16_t>(pred[c]);
                    for (int k = 1; k < 64;) {
                        br.refill();
                        const int rs = huff_symbol(br, Ha);
                        const int r = rs >> 4;
                        s = rs & 15;
                        if (s) {
                            k += r;
                            br.refill();
                            const int val = huff_extend(br.peek(s), s);
                            br.skip(s);
                            if (k < 64) blk[c_zigzag[k]] = static_cast<int16_t>(val);
                            ++k;
                        } else if (r == 15) {
                            k += 16;
                        } else {
                            break;
                        }
                    }
                }
        }
        if (++mx == im.mcux) { mx = 0; ++my; }
    }
}

// ---- device [1b]: Huffman decode of long segments in parallel sub-sequences ------------------------------------
// A file without restart markers is one bit stream that only parses front to back.  Huffman streams re-synchronise,
// though: a decoder started at a wrong position falls into step with the true parse after a few symbols.  So
// (Klein & Wiseman 2003; Weissenberger & Schmidt 2018 for JPEG): cut the stream into sub-sequences of kSubBits bits,
//   A  every thread parses its sub-sequence from a guessed state; then, repeatedly, re-parses it from the state in
//      which its left neighbour ENDED, until no end state in the CTA changes (state = bit position, block slot in the
//      MCU, zig-zag index: everything the further parse depends on; DC prediction is handled separately);
//   B  the true state is carried across the CTA boundaries: all boundaries in parallel first (a CTA's last end state
//      is almost always already true), then one thread per segment walks them left to right, re-parsing from the
//      true state until the stored end state is met again wherever the chain does not hold yet;
//   C  exclusive scan of the blocks completed per sub-sequence -> first block index of each;
//   D  every thread parses its sub-sequence once more from its now-true start state and stores the coefficients (DC
//      as differences);
//   E  running sums turn the DC differences of each component into DC values.
struct SegCtx {
    const uint32_t* words;         // un-stuffed stream
    const uint32_t* end;           // first word past it (+ slack)
    int bpm;                       // blocks per MCU
    int slot_comp[6], slot_h[6], slot_v[6];
};

__device__ __forceinline__ void seg_context(const JpegImg& im, const uint8_t* blob, const JpegSeg& seg, SegCtx& cx) {
    cx.words = reinterpret_cast<const uint32_t*>(blob + seg.byte_off);
    cx.end = cx.words + (seg.byte_len + 3) / 4 + 2;
    cx.bpm = 0;
    for (int c = 0; c < im.ncomp; ++c)
        for (int v = 0; v < im.vs[c]; ++v)
            for (int h = 0; h < im.hs[c]; ++h) { cx.slot_comp[cx.bpm] = c; cx.slot_h[cx.bpm] = h; cx.slot_v[cx.bpm] = v; ++cx.bpm; }
}

struct PosReader {                 // BitReader that knows its bit position in the segment
    const uint32_t* base; const uint32_t* p; const uint32_t* end; uint64_t buf; int cnt;
    __device__ __forceinline__ void seek(const uint32_t* b, const uint32_t* e, int bit) {
        base = b; end = e; p = b + (bit >> 5); buf = 0; cnt = 0;
        refill(); skip(bit & 31); refill();
    }
    __device__ __forceinline__ void refill() {
        if (cnt <= 32) {
            const uint32_t w = p < end ? __byte_perm(__ldg(p), 0, 0x0123) : 0u;     // past the segment: zeros
            ++p;
            buf |= static_cast<uint64_t>(w) << (32 - cnt);
            cnt += 32;
        }
    }
    __device__ __forceinline__ uint32_t peek(int n) const { return static_cast<uint32_t>(buf >> (64 - n)); }
    __device__ __forceinline__ void skip(int n) { buf <<= n; cnt -= n; }
    __device__ __forceinline__ int pos() const { return static_cast<int>(p - base) * 32 - cnt; }
};

__device__ __forceinline__ int huff_symbol_p(PosReader& br, const HuffDev& H) {
    const uint32_t lk = H.look[br.peek(kLookBits)];
    if (lk) { br.skip(lk >> 8); return lk & 0xFF; }
    int l = kLookBits + 1;
    int code = br.peek(l);
    while (l <= 16 && code > H.maxcode[l]) { ++l; code = br.peek(l); }
    if (l > 16) { br.skip(16); return 0; }
    br.skip(l);
    return H.huffval[(code + H.valoff[l]) & 0xFF];
}

// Parse from (p, state) while symbols START before bit `limit`; returns the end position / state and the number of
// blocks completed.  WRITE: also store the coefficients, the first block being number blk of the segment.
template <bool WRITE>
__device__ __forceinline__ void parse_sub(const SegCtx& cx, const HuffDev* sh, const int* sel_dc, const int* sel_ac,
                                          int p, int st, int limit, int& p_end, int& st_end, int& nblk,
                                          const JpegImg* im, const JpegSeg* seg, uint8_t* ws, int blk) {
    PosReader br;
    br.seek(cx.words, cx.end, p);
    int slot = st >> 8, k = st & 0xFF, done = 0;
    int16_t* out = nullptr;
    [[maybe_unused]] auto locate = [&]() {
        const int mcu = seg->first_mcu + blk / cx.bpm;
        const int c = cx.slot_comp[slot];
        const int my = mcu / im->mcux, mx = mcu - my * im->mcux;
        out = reinterpret_cast<int16_t*>(ws + im->coef_off[c]) +
              (static_cast<int64_t>(my * im->vs[c] + cx.slot_v[slot]) * im->bw[c] + (mx * im->hs[c] + cx.slot_h[slot])) * 64;
    };
    if (WRITE) {
        if (blk >= seg->total_blocks) { p_end = p; st_end = st; nblk = 0; return; }
        locate();
    }
    while (br.pos() < limit) {
        const int c = cx.slot_comp[slot];
        br.refill();
        if (k == 0) {
            const int s = min(huff_symbol_p(br, sh[sel_dc[c]]), 16);   // (a damaged table may name any byte)
            if (s) {
                br.refill();
                const int d = huff_extend(br.peek(s), s);
                br.skip(s);
                if (WRITE && d) out[0] = static_cast<int16_t>(d);
            }
            k = 1;
        } else {
            const int rs = huff_symbol_p(br, sh[sel_ac[c]]);
            const int r = rs >> 4, s = rs & 15;
            if (s) {
                k += r;
                br.refill();
                const int val = huff_extend(br.peek(s), s);
                br.skip(s);
                if (WRITE && k < 64) out[c_zigzag[k]] = static_cast<int16_t>(val);
                ++k;
            } else if (r == 15) {
                k += 16;
            } else {
                k = 64;
            }
        }
        if (k >= 64) {
            k = 0; ++done;
            if (++slot == cx.bpm) slot = 0;
            if (WRITE) {
                if (++blk >= seg->total_blocks) break;
                locate();
            }
        }
    }
    p_end = br.pos(); st_end = (slot << 8) | k; nblk = done;
}

// shared by kernels A, B, D: the image's four Huffman tables into shared memory
__device__ __forceinline__ void stage_tables(const uint8_t* blob, const JpegImg& im, HuffDev* sh, int* sel_dc, int* sel_ac) {
    const HuffDev* tab = reinterpret_cast<const HuffDev*>(blob + im.tables_off + 4 * 64 * 2);
    int dc_ids[2] = {im.dcsel[0], im.dcsel[0]}, ac_ids[2] = {im.acsel[0], im.acsel[0]};
    for (int c = 1; c < im.ncomp; ++c) {
        if (im.dcsel[c] != dc_ids[0]) dc_ids[1] = im.dcsel[c];
        if (im.acsel[c] != ac_ids[0]) ac_ids[1] = im.acsel[c];
    }
    constexpr int W = sizeof(HuffDev) / 4;
    for (int t = 0; t < 4; ++t) {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(tab + (t < 2 ? dc_ids[t] : 4 + ac_ids[t - 2]));
        uint32_t* dst = reinterpret_cast<uint32_t*>(&sh[t]);
        for (int i = threadIdx.x; i < W; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    if (threadIdx.x < 3) {
        const int c = threadIdx.x < im.ncomp ? threadIdx.x : 0;
        sel_dc[threadIdx.x] = im.dcsel[c] == dc_ids[0] ? 0 : 1;
        sel_ac[threadIdx.x] = im.acsel[c] == ac_ids[0] ? 2 : 3;
    }
    __syncthreads();
}

// A: synchronise the sub-sequences of one CTA with each other
__global__ void __launch_bounds__(kSubThreads)
jpeg_sync_kernel(const uint8_t* __restrict__ blob, int64_t imgs_off, int64_t segs_off, int64_t cta_off, SubInfo* __restrict__ subs) {
    __shared__ HuffDev sh[4];
    __shared__ int sel_dc[3], sel_ac[3];
    __shared__ int e_p[kSubThreads], e_st[kSubThreads];
    const JpegCta w = reinterpret_cast<const JpegCta*>(blob + cta_off)[blockIdx.x];
    const JpegSeg& seg = reinterpret_cast<const JpegSeg*>(blob + segs_off)[w.seg];
    const JpegImg& im = reinterpret_cast<const JpegImg*>(blob + imgs_off)[seg.img];
    stage_tables(blob, im, sh, sel_dc, sel_ac);
    SegCtx cx;
    seg_context(im, blob, seg, cx);
    const int t = threadIdx.x, sub = w.first_sub + t;
    const bool valid = sub < seg.nsub;
    const int limit = (sub + 1) * kSubBits;
    int p0 = sub * kSubBits, st0 = 0, pe = p0, se = 0, nb = 0;
    if (valid) parse_sub<false>(cx, sh, sel_dc, sel_ac, p0, st0, limit, pe, se, nb, nullptr, nullptr, nullptr, 0);
    e_p[t] = pe; e_st[t] = se;
    for (int it = 0; it < kSubThreads; ++it) {
        __syncthreads();
        const int pp = t > 0 ? e_p[t - 1] : p0, ps = t > 0 ? e_st[t - 1] : st0;
        const bool redo = valid && t > 0 && (pp != p0 || ps != st0);
        __syncthreads();
        if (redo) {
            p0 = pp; st0 = ps;
            parse_sub<false>(cx, sh, sel_dc, sel_ac, p0, st0, limit, pe, se, nb, nullptr, nullptr, nullptr, 0);
            e_p[t] = pe; e_st[t] = se;
        }
        if (!__syncthreads_or(redo)) break;
    }
    if (valid) {
        SubInfo& o = subs[seg.sub0 + sub];
        o.p_start = p0; o.st_start = st0; o.p_end = pe; o.st_end = se; o.nblk = nb; o.blk0 = 0;
    }
}

// B1: all CTA boundaries at once.  The end state of a CTA's last sub-sequence is almost always already the true one
// (the parse falls into step within a few sub-sequences of the CTA's guessed start), so every boundary can be
// carried over independently: one warp per boundary, lane 0 re-parses from the left neighbour's end state until the
// stored end state is met again (at most to the end of its own CTA).  Kernel B2 below then only has to confirm the
// chain — and repairs the rare boundary whose left neighbour did change.
__global__ void __launch_bounds__(32)
jpeg_stitch_par_kernel(const uint8_t* __restrict__ blob, int64_t imgs_off, int64_t segs_off, int64_t cta_off, SubInfo* __restrict__ subs) {
    __shared__ HuffDev sh[4];
    __shared__ int sel_dc[3], sel_ac[3];
    const JpegCta w = reinterpret_cast<const JpegCta*>(blob + cta_off)[blockIdx.x];
    if (w.first_sub == 0) return;
    const JpegSeg& seg = reinterpret_cast<const JpegSeg*>(blob + segs_off)[w.seg];
    const JpegImg& im = reinterpret_cast<const JpegImg*>(blob + imgs_off)[seg.img];
    stage_tables(blob, im, sh, sel_dc, sel_ac);
    if (threadIdx.x != 0) return;
    SubInfo* S = subs + seg.sub0;
    SegCtx cx;
    seg_context(im, blob, seg, cx);
    int pp = S[w.first_sub - 1].p_end, ps = S[w.first_sub - 1].st_end;
    const int stop = min(seg.nsub, w.first_sub + kSubThreads);
    for (int cur = w.first_sub; cur < stop; ++cur) {
        if (S[cur].p_start == pp && S[cur].st_start == ps) break;
        int pe, se, nb;
        parse_sub<false>(cx, sh, sel_dc, sel_ac, pp, ps, (cur + 1) * kSubBits, pe, se, nb, nullptr, nullptr, nullptr, 0);
        const bool same_end = pe == S[cur].p_end && se == S[cur].st_end;
        S[cur].p_start = pp; S[cur].st_start = ps; S[cur].p_end = pe; S[cur].st_end = se; S[cur].nblk = nb;
        if (same_end) break;
        pp = pe; ps = se;
    }
}

// B2: confirm / carry the true state across the CTA boundaries of a segment left to right, then C: exclusive scan of
// the block counts
__global__ void __launch_bounds__(32)
jpeg_stitch_kernel(const uint8_t* __restrict__ blob, int64_t imgs_off, int64_t segs_off, int64_t pseg_off, SubInfo* __restrict__ subs) {
    __shared__ HuffDev sh[4];
    __shared__ int sel_dc[3], sel_ac[3];
    const int si = reinterpret_cast<const int32_t*>(blob + pseg_off)[blockIdx.x];
    const JpegSeg& seg = reinterpret_cast<const JpegSeg*>(blob + segs_off)[si];
    const JpegImg& im = reinterpret_cast<const JpegImg*>(blob + imgs_off)[seg.img];
    stage_tables(blob, im, sh, sel_dc, sel_ac);
    SubInfo* S = subs + seg.sub0;
    if (threadIdx.x == 0) {
        SegCtx cx;
        seg_context(im, blob, seg, cx);
        for (int first = kSubThreads; first < seg.nsub; first += kSubThreads) {
            int pp = S[first - 1].p_end, ps = S[first - 1].st_end;
            for (int cur = first; cur < seg.nsub; ++cur) {
                if (S[cur].p_start == pp && S[cur].st_start == ps) break;          // in step with the true parse
                int pe, se, nb;
                parse_sub<false>(cx, sh, sel_dc, sel_ac, pp, ps, (cur + 1) * kSubBits, pe, se, nb, nullptr, nullptr, nullptr, 0);
                const bool same_end = pe == S[cur].p_end && se == S[cur].st_end;
                S[cur].p_start = pp; S[cur].st_start = ps; S[cur].p_end = pe; S[cur].st_end = se; S[cur].nblk = nb;
                if (same_end) break;
                pp = pe; ps = se;
            }
        }
    }
    __syncwarp();
    // exclusive scan of nblk over the segment's sub-sequences (warp scan, 32 at a time)
    int carry = 0;
    for (int base = 0; base < seg.nsub; base += 32) {
        const int i = base + threadIdx.x;
        const int v = i < seg.nsub ? S[i].nblk : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, inc, o); if (static_cast<int>(threadIdx.x) >= o) inc += u; }
        if (i < seg.nsub) S[i].blk0 = carry + inc - v;
        carry += __shfl_sync(0xffffffffu, inc, 31);
    }
}

// D: the final parse, storing coefficients
__global__ void __launch_bounds__(kSubThreads)
jpeg_write_kernel(const uint8_t* __restrict__ blob, int64_t imgs_off, int64_t segs_off, int64_t cta_off,
                  const SubInfo* __restrict__ subs, uint8_t* __restrict__ ws) {
    __shared__ HuffDev sh[4];
    __shared__ int sel_dc[3], sel_ac[3];
    const JpegCta w = reinterpret_cast<const JpegCta*>(blob + cta_off)[blockIdx.x];
    const JpegSeg& seg = reinterpret_cast<const JpegSeg*>(blob + segs_off)[w.seg];
    const JpegImg& im = reinterpret_cast<const JpegImg*>(blob + imgs_off)[seg.img];
    stage_tables(blob, im, sh, sel_dc, sel_ac);
    const int sub = w.first_sub + threadIdx.x;
    if (sub >= seg.nsub) return;
    SegCtx cx;
    seg_context(im, blob, seg, cx);
    const SubInfo in = subs[seg.sub0 + sub];
    int pe, se, nb;
    parse_sub<true>(cx, sh, sel_dc, sel_ac, in.p_start, in.st_start, (sub + 1) * kSubBits, pe, se, nb, &im, &seg, ws, in.blk0);
}

// E: DC differences -> DC values, one CTA per (mode-1 segment, component), blocks in decode order
__global__ void __launch_bounds__(1024)
jpeg_dc_kernel(const uint8_t* __restrict__ blob, int64_t imgs_off, int64_t segs_off, int64_t pseg_off, uint8_t* __restrict__ ws) {
    __shared__ int wsum[32];
    __shared__ int carry_s;
    const int si = reinterpret_cast<const int32_t*>(blob + pseg_off)[blockIdx.x];
    const JpegSeg& seg = reinterpret_cast<const JpegSeg*>(blob + segs_off)[si];
    const JpegImg& im = reinterpret_cast<const JpegImg*>(blob + imgs_off)[seg.img];
    const int c = blockIdx.y;
    if (c >= im.ncomp) return;
    const int hs = im.hs[c], vs = im.vs[c], per = hs * vs, bw = im.bw[c];
    const int n = seg.n_mcu * per;
    int16_t* coef = reinterpret_cast<int16_t*>(ws + im.coef_off[c]);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + threadIdx.x;
        int16_t* ptr = nullptr;
        int v = 0;
        if (i < n) {
            const int mcu = seg.first_mcu + i / per, r = i % per;
            const int my = mcu / im.mcux, mx = mcu - my * im.mcux;
            ptr = coef + (static_cast<int64_t>(my * vs + r / hs) * bw + (mx * hs + r % hs)) * 64;
            v = *ptr;
        }
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int w = wsum[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += u; }
            wsum[lane] = w;
        }
        __syncthreads();
        const int total = carry_s + (warp ? wsum[warp - 1] : 0) + inc;
        if (ptr) *ptr = static_cast<int16_t>(total);
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = total;
        __syncthreads();
    }
}

// ---- device [2]: de-quantise + IDCT (jidctint.c jpeg_idct_islow, restated) ---------------------------------------
constexpr int kConstBits = 13, kPass1Bits = 2;
#define PT_FIX_0_298631336 2446
#define PT_FIX_0_390180644 3196
#define PT_FIX_0_541196100 4433
#define PT_FIX_0_765366865 6270
#define PT_FIX_0_899976223 7373
#define PT_FIX_1_175875602 9633
#define PT_FIX_1_501321110 12299
#define PT_FIX_1_847759065 15137
#define PT_FIX_1_961570560 16069
#define PT_FIX_2_053119869 16819
#define PT_FIX_2_562915447 20995
#define PT_FIX_3_072711026 25172

__device__ __forceinline__ int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

// the 1-D transform both passes share: in[0..7] -> out[0..7], un-descaled
__device__ __forceinline__ void idct_1d(const int (&in)[8], int (&out)[8], int dc_shift) {
    int z2 = in[2], z3 = in[6];
    int z1 = (z2 + z3) * PT_FIX_0_541196100;
    int tmp2 = z1 + z3 * (-PT_FIX_1_847759065);
    int tmp3 = z1 + z2 * PT_FIX_0_765366865;
    z2 = in[0]; z3 = in[4];
    int tmp0 = (z2 + z3) << dc_shift;
    int tmp1 = (z2 - z3) << dc_shift;
    const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    tmp0 = in[7]; tmp1 = in[5]; tmp2 = in[3]; tmp3 = in[1];
    z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2;
    int z4 = tmp1 + tmp3;
    const int z5 = (z3 + z4) * PT_FIX_1_175875602;
    tmp0 *= PT_FIX_0_298631336; tmp1 *= PT_FIX_2_053119869; tmp2 *= PT_FIX_3_072711026; tmp3 *= PT_FIX_1_501321110;
    z1 *= -PT_FIX_0_899976223; z2 *= -PT_FIX_2_562915447; z3 *= -PT_FIX_1_961570560; z4 *= -PT_FIX_0_390180644;
    z3 += z5; z4 += z5;
    tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
    out[0] = tmp10 + tmp3; out[7] = tmp10 - tmp3;
    out[1] = tmp11 + tmp2; out[6] = tmp11 - tmp2;
    out[2] = tmp12 + tmp1; out[5] = tmp12 - tmp1;
    out[3] = tmp13 + tmp0; out[4] = tmp13 - tmp0;
}

// 8 threads per 8x8 block (column j in pass 1, row j in pass 2), 8 blocks per CTA of 64 threads
__global__ void __launch_bounds__(64)
jpeg_idct_kernel(const int16_t* __restrict__ coef, const uint16_t* __restrict__ quant, uint8_t* __restrict__ plane,
                 int bw, int nblocks) {
    __shared__ int wsp[8][8][9];
    const int g = threadIdx.x >> 3, j = threadIdx.x & 7;
    const int b = blockIdx.x * 8 + g;
    if (b >= nblocks) return;               // whole 8-thread groups leave together
    const int16_t* cb = coef + static_cast<int64_t>(b) * 64;
    int in[8], out[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) in[k] = static_cast<int>(cb[k * 8 + j]) * static_cast<int>(__ldg(quant + k * 8 + j));
    idct_1d(in, out, kConstBits);
#pragma unroll
    for (int k = 0; k < 8; ++k) wsp[g][k][j] = descale(out[k], kConstBits - kPass1Bits);
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 8; ++k) in[k] = wsp[g][j][k];
    idct_1d(in, out, kConstBits);
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        int v = descale(out[k], kConstBits + kPass1Bits + 3) + 128;
        v = min(255, max(0, v));
        if (k < 4) lo |= static_cast<uint32_t>(v) << (8 * k); else hi |= static_cast<uint32_t>(v) << (8 * (k - 4));
    }
    const int by = b / bw, bx = b - by * bw;
    uint2* dst = reinterpret_cast<uint2*>(plane + (static_cast<int64_t>(by) * 8 + j) * (bw * 8) + bx * 8);
    *dst = make_uint2(lo, hi);
}

// ---- device [3]: chroma up-sampling (jdsample.c, fancy) + YCbCr -> RGB (jdcolor.c) ----------------------------
struct ColorArgs {
    int width, height, ncomp, fx, fy;     // chroma factors: 1 or 2
    int pw0, pw1, dwc, dhc;               // plane row strides (luma, chroma), real chroma columns / rows
    const uint8_t* y; const uint8_t* cb; const uint8_t* cr;
    uint8_t* rgb;
};

__device__ __forceinline__ int chroma_at(const uint8_t* __restrict__ pl, int pw, int n, int nrows, int fx, int fy, int x, int y) {
    if (fx == 1) return pl[static_cast<int64_t>(y) * pw + x];
    const int cx = x >> 1;
    // jdsample.c jinit_upsampler: the triangle filters are used only when the component is more than two samples
    // wide; narrower components are replicated (h2v1_upsample / h2v2_upsample)
    if (n <= 2) return pl[static_cast<int64_t>(fy == 2 ? y >> 1 : y) * pw + cx];
    if (fy == 1) {                        // h2v1_fancy_upsample
        const uint8_t* in = pl + static_cast<int64_t>(y) * pw;
        const int v = in[cx];
        if (cx == 0) return (x & 1) ? (v * 3 + in[1] + 2) >> 2 : v;
        if (cx == n - 1) return (x & 1) ? v : (v * 3 + in[cx - 1] + 1) >> 2;
        return (x & 1) ? (v * 3 + in[cx + 1] + 2) >> 2 : (v * 3 + in[cx - 1] + 1) >> 2;
    }
    // h2v2_fancy_upsample: 3/4 nearer row + 1/4 further row, then the same along the row, in sixteenths
    const int cy = y >> 1;
    int oy = (y & 1) ? cy + 1 : cy - 1;
    oy = min(nrows - 1, max(0, oy));
    const uint8_t* in0 = pl + static_cast<int64_t>(cy) * pw;
    const uint8_t* in1 = pl + static_cast<int64_t>(oy) * pw;
    const int cur = in0[cx] * 3 + in1[cx];
    if (cx == 0) return (x & 1) ? (cur * 3 + (in0[1] * 3 + in1[1]) + 7) >> 4 : (cur * 4 + 8) >> 4;
    if (cx == n - 1) return (x & 1) ? (cur * 4 + 7) >> 4 : (cur * 3 + (in0[cx - 1] * 3 + in1[cx - 1]) + 8) >> 4;
    return (x & 1) ? (cur * 3 + (in0[cx + 1] * 3 + in1[cx + 1]) + 7) >> 4
                   : (cur * 3 + (in0[cx - 1] * 3 + in1[cx - 1]) + 8) >> 4;
}

__global__ void __launch_bounds__(256)
jpeg_color_kernel(const ColorArgs a) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= a.width) return;
    const int Y = a.y[static_cast<int64_t>(y) * a.pw0 + x];
    int r = Y, g = Y, b = Y;
    if (a.ncomp == 3) {
        const int cb = chroma_at(a.cb, a.pw1, a.dwc, a.dhc, a.fx, a.fy, x, y) - 128;
        const int cr = chroma_at(a.cr, a.pw1, a.dwc, a.dhc, a.fx, a.fy, x, y) - 128;
        // jdcolor.c build_ycc_rgb_table: FIX(x) = (int)(x * 65536 + 0.5), ONE_HALF = 32768, arithmetic shifts
        r = Y + ((91881 * cr + 32768) >> 16);
        b = Y + ((116130 * cb + 32768) >> 16);
        g = Y + ((-22554 * cb + 32768 - 46802 * cr) >> 16);
        r = min(255, max(0, r)); g = min(255, max(0, g)); b = min(255, max(0, b));
    }
    uint8_t* o = a.rgb + (static_cast<int64_t>(y) * a.width + x) * 3;
    o[0] = static_cast<uint8_t>(r); o[1] = static_cast<uint8_t>(g); o[2] = static_cast<uint8_t>(b);
}

}  // namespace pt

extern "C" int pt_jpeg_decode_batch(const uint8_t* host_blob, const uint8_t* dev_blob, uint8_t* dst_rgb,
                                    void* workspace, int64_t workspace_bytes, void* stream) {
    PT_REQUIRE(host_blob && dev_blob && dst_rgb && workspace, "pt_jpeg_decode_batch: null pointer");
    const JpegBlobHeader& h = *reinterpret_cast<const JpegBlobHeader*>(host_blob);
    PT_REQUIRE(h.magic == kJpegMagic && h.n > 0, "pt_jpeg_decode_batch: host_blob was not written by pt_jpeg_pack");
    PT_REQUIRE(workspace_bytes >= h.workspace_bytes, "pt_jpeg_decode_batch: workspace too small (see pt_jpeg_plan)");
    PT_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0 && (reinterpret_cast<uintptr_t>(dev_blob) & 15) == 0,
               "pt_jpeg_decode_batch: workspace must be 256-byte and dev_blob 16-byte aligned");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    static bool zz_done[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { set_error("pt_jpeg_decode_batch: bad device"); return PT_ERR_CUDA; }
    if (!zz_done[dev]) {
        if (cudaMemcpyToSymbol(c_zigzag, kZigzag, 64) != cudaSuccess) return check_launch("jpeg zigzag table");
        zz_done[dev] = true;
    }
    uint8_t* ws = static_cast<uint8_t*>(workspace);
    if (cudaMemsetAsync(ws, 0, static_cast<size_t>(h.coef_bytes), st) != cudaSuccess) return check_launch("jpeg coefficient clear");
    if (h.n_pseg < h.total_segs) jpeg_huffman_kernel<<<h.total_segs, 32, 0, st>>>(dev_blob, h.imgs_off, h.segs_off, ws);
    if (h.n_pseg > 0) {
        SubInfo* subs = reinterpret_cast<SubInfo*>(ws + h.sub_off);
        jpeg_sync_kernel<<<h.n_cta, kSubThreads, 0, st>>>(dev_blob, h.imgs_off, h.segs_off, h.cta_off, subs);
        if (h.n_cta > h.n_pseg)
            jpeg_stitch_par_kernel<<<h.n_cta, 32, 0, st>>>(dev_blob, h.imgs_off, h.segs_off, h.cta_off, subs);
        jpeg_stitch_kernel<<<h.n_pseg, 32, 0, st>>>(dev_blob, h.imgs_off, h.segs_off, h.pseg_off, subs);
        jpeg_write_kernel<<<h.n_cta, kSubThreads, 0, st>>>(dev_blob, h.imgs_off, h.segs_off, h.cta_off, subs, ws);
        jpeg_dc_kernel<<<dim3(h.n_pseg, 3), 1024, 0, st>>>(dev_blob, h.imgs_off, h.segs_off, h.pseg_off, ws);
    }
    const JpegImg* imgs = reinterpret_cast<const JpegImg*>(host_blob + h.imgs_off);
    for (int i = 0; i < h.n; ++i) {
        const JpegImg& im = imgs[i];
        for (int c = 0; c < im.ncomp; ++c) {
            const int nb = im.bw[c] * im.bh[c];
            jpeg_idct_kernel<<<(nb + 7) / 8, 64, 0, st>>>(
                reinterpret_cast<const int16_t*>(ws + im.coef_off[c]),
                reinterpret_cast<const uint16_t*>(dev_blob + im.tables_off) + 64 * im.qsel[c], ws + im.plane_off[c],
                im.bw[c], nb);
        }
        ColorArgs a;
        a.width = im.width; a.height = im.height; a.ncomp = im.ncomp;
        a.fx = im.hmax; a.fy = im.vmax;
        a.pw0 = im.bw[0] * 8;
        a.pw1 = im.ncomp == 3 ? im.bw[1] * 8 : 0;
        a.dwc = im.ncomp == 3 ? im.dw[1] : 0; a.dhc = im.ncomp == 3 ? im.dh[1] : 0;
        a.y = ws + im.plane_off[0];
        a.cb = im.ncomp == 3 ? ws + im.plane_off[1] : nullptr;
        a.cr = im.ncomp == 3 ? ws + im.plane_off[2] : nullptr;
        a.rgb = dst_rgb + im.rgb_off;
        dim3 grid((im.width + 255) / 256, im.height);
        jpeg_color_kernel<<<grid, 256, 0, st>>>(a);
    }
    return check_launch("jpeg_decode_batch");
}
