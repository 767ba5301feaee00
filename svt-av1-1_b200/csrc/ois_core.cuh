// ois_core.cuh -- the arithmetic of the open-loop intra search (OpenLoopIntraSearchLcu, reference
// Source/Lib/Codec/EbMotionEstimation.c:7419-8365; predictors EbIntraPrediction.c:3718-3800, 5230-5430 with the leaves of
// ASM_AVX2/EbIntraPrediction_Intrinsic_AVX2.c) as small host/device functions: one neighbour sample, one sample of the
// projected main reference, one predicted pixel, and the per-CU candidate-list rules. ois.cu maps them onto warps; the same
// functions compile for the host (tests/host_emul/ois_emul.cpp), where a serial emulation of the kernel is checked against
// the oracle without a GPU.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define OIS_HD __host__ __device__ __forceinline__
#define OIS_UNROLL _Pragma("unroll")
#else
#define OIS_HD static inline
#define OIS_UNROLL
#endif

namespace svtme {
namespace ois {

enum { kPlanar = 0, kDc = 1, kHor = 10, kVer = 26, kNumModes = 35, kMaxCand = 18, kCus = 85 };

struct Picture { const uint8_t* pix; int stride, w, h; };          // pix = sample (0, 0) of the (padded) source plane

struct Params {                                                     // what OpenLoopIntraSearchLcu reads from PCS / SCS
    int32_t slice_type;            // 0 = I, 1 = P, 2 = B
    int32_t temporal_layer_index, is_used_as_reference;
    int32_t cu8x8_mode;            // 1: no 8x8 CUs on inter pictures
    int32_t input_resolution;      // 3 = INPUT_SIZE_4K_RANGE
    int32_t limit_ois_to_dc;
    int32_t enc_mode;              // one quirk reads it (valid flag of a vertical stage-1 winner, ME.c:7604-7612)
};

// raster-scan CU index 1..84 -> position and size inside the SB
OIS_HD void cu_geometry(int cu, int& x, int& y, int& n) {
    if (cu < 5) { n = 32; x = ((cu - 1) & 1) * 32; y = ((cu - 1) >> 1) * 32; }
    else if (cu < 21) { n = 16; x = ((cu - 5) & 3) * 16; y = ((cu - 5) >> 2) * 16; }
    else { n = 8; x = ((cu - 21) & 7) * 8; y = ((cu - 21) >> 3) * 8; }
}

// Neighbour array r[0 .. 4n] of the CU at (x0, y0) (UpdateNeighborSamplesArrayOpenLoop, "regular" order):
// r[0..2n-1] = left column bottom-up, r[2n] = top-left, r[2n+1..4n] = top row. Samples outside the picture are 128; the
// left column / top row stop at the picture's bottom / right edge.
OIS_HD int neighbour(const Picture& p, int x0, int y0, int n, int idx) {
    if (idx < 2 * n) {
        const int y = 2 * n - 1 - idx;
        if (x0 == 0 || y0 + y >= p.h) return 128;
        return p.pix[(long long)(y0 + y) * p.stride + x0 - 1];
    }
    if (idx == 2 * n) return (x0 != 0 && y0 != 0) ? p.pix[(long long)(y0 - 1) * p.stride + x0 - 1] : 128;
    const int x = idx - 2 * n - 1;
    if (y0 == 0 || x0 + x >= p.w) return 128;
    return p.pix[(long long)(y0 - 1) * p.stride + x0 + x];
}

OIS_HD int angle_of(int mode) {
    const int d = mode < 18 ? mode - 10 : mode - 26;                // distance from pure horizontal / vertical
    const int m = d < 0 ? -d : d;
    const int mag = (int)((0x2069544D245080ull >> (6 * m)) & 63);    // 0 2 5 9 13 17 21 26 32, six bits each (no jump table on the device)
    return (mode < 18 ? -d : d) < 0 ? -mag : mag;                   // horizontal modes: 2 -> +32 ... 17 -> -26; vertical: 18 -> -32 ... 34 -> +32
}
OIS_HD int inv_angle_of(int mode) {                                 // only for the negative angles (modes 11..25)
    const int d = mode < 18 ? mode - 10 : 26 - mode;                // 1..8
    const unsigned long long t = d <= 4 ? 0x13B0E38CCD000ull : 0x8004EC30C1E2ull;   // 4096 1638 910 630 | 482 390 315 256, 13 bits each
    return -(int)((t >> (13 * ((d - 1) & 3))) & 8191);
}

// Main reference of an angular mode: ref[i], i in [-n, 2n]; i = 0 is the top-left sample, i > 0 runs along the top row
// (modes >= 18) or down the left column (modes < 18), i < 0 is the side reference projected by the inverse angle.
OIS_HD int ref_sample(const uint8_t* r, int n, int mode, int i) {
    const bool vertical = mode >= 18;
    const uint8_t* left = r + 2 * n - 1;          // left[-y]: sample left of row y
    const uint8_t* top = r + 2 * n + 1;           // top[x]; top[-1] = top-left
    if (i >= 0) return vertical ? top[i - 1] : (i == 0 ? top[-1] : left[-(i - 1)]);
    const int k = (i * inv_angle_of(mode) + 128) >> 8;
    return vertical ? left[-(k - 1)] : top[k - 1];
}
// the i >= 0 part does not depend on the angle: the word-wise kernel keeps it per CU (one buffer per direction) and only
// projects the i < 0 part per mode
OIS_HD int straight_sample(const uint8_t* r, int n, bool vertical, int i) { return vertical ? r[2 * n + i] : r[2 * n - i]; }
OIS_HD int ref_first(int n, int mode) { const int a = angle_of(mode); return a < 0 ? (n * a) >> 5 : 0; }   // lowest index that is read

OIS_HD int clip8(int v) { return v < 0 ? 0 : v > 255 ? 255 : v; }

// One predicted pixel. r: the neighbour array; ref: the mode's main reference with ref[0] = top-left (angular modes only);
// dc: (sum of the n top + n left neighbours + n) >> (log2 n + 1) (DC mode only).
OIS_HD int pred_pixel(const uint8_t* r, const uint8_t* ref, int n, int lg, int mode, int x, int y, int dc) {
    const uint8_t* left = r + 2 * n - 1;
    const uint8_t* top = r + 2 * n + 1;
    if (mode == kPlanar)
        return ((n - 1 - x) * left[-y] + (x + 1) * top[n] + (n - 1 - y) * top[x] + (y + 1) * left[-n] + n) >> (lg + 1);
    if (mode == kDc) {
        if (n < 32) {
            if (x == 0 && y == 0) return (left[0] + 2 * dc + top[0] + 2) >> 2;
            if (y == 0) return (top[x] + 3 * dc + 2) >> 2;
            if (x == 0) return (left[-y] + 3 * dc + 2) >> 2;
        }
        return dc;
    }
    const bool vertical = mode >= 18;
    const int a = angle_of(mode);
    if (a == 0 && n < 32) {                                          // boundary smoothing of the pure vertical / horizontal modes
        if (vertical && x == 0) return clip8(top[0] + ((left[-y] - top[-1]) >> 1));
        if (!vertical && y == 0) return clip8(left[0] + ((top[x] - top[-1]) >> 1));
    }
    const int j = vertical ? y : x, i = vertical ? x : y;
    const int pos = (j + 1) * a, idx = pos >> 5, f = pos & 31;
    return f ? ((32 - f) * ref[i + idx + 1] + f * ref[i + idx + 2] + 16) >> 5 : ref[i + idx + 1];
}

// ---- word-wise prediction (the fast kernel): a lane predicts K adjacent pixels at once --------------------------------------
// A CU of n x n pixels is cut into words of K pixels (K = 4, 8x8 CUs: 2) that run along the MINOR axis i of the mode at a
// fixed MAJOR coordinate j: vertical modes (>= 18), planar and DC: i = x, j = y (a piece of a row); horizontal modes (2..17):
// i = y, j = x (a piece of a column, compared with the transposed source word). With that choice every angular mode is the
// same code: the K pixels interpolate K + 1 consecutive samples of the main reference with one fraction.
OIS_HD int word_pixels(int lg) { return lg == 3 ? 2 : 4; }
OIS_HD int words_per_lane(int lg) { return ((1 << lg) << lg) / word_pixels(lg) / 32; }           // 8x8: 1, 16x16: 2, 32x32: 8
OIS_HD void word_pos(int lg, int lane, int k, int& j, int& i0) {
    const int wi = lane + 32 * k;
    j = wi & ((1 << lg) - 1);
    i0 = (wi >> lg) * word_pixels(lg);
}
// K source pixels from (x, y), stepping (dx, dy); byte t of the result = pixel t, unused bytes 0
OIS_HD uint32_t source_pixels(const uint8_t* src, int stride, int pixels, int j, int i0, bool transposed) {
    uint32_t w = 0;
    for (int t = 0; t < pixels; t++) {
        const int x = transposed ? j : i0 + t, y = transposed ? i0 + t : j;
        w |= (uint32_t)src[(long long)y * stride + x] << (8 * t);
    }
    return w;
}
OIS_HD uint32_t source_word(const uint8_t* src, int stride, int lg, int j, int i0, bool transposed) {
    return source_pixels(src, stride, word_pixels(lg), j, i0, transposed);
}
OIS_HD uint32_t sad_word(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __vsadu4(a, b);
#else
    uint32_t s = 0;
    for (int t = 0; t < 4; t++) { const int d = (int)((a >> (8 * t)) & 255) - (int)((b >> (8 * t)) & 255); s += (uint32_t)(d < 0 ? -d : d); }
    return s;
#endif
}
// planar / DC word of row j from column i0
OIS_HD uint32_t flat_word(const uint8_t* r, int lg, int mode, int j, int i0, int dc) {
    uint32_t w = 0;
    for (int t = 0; t < word_pixels(lg); t++) w |= (uint32_t)pred_pixel(r, nullptr, 1 << lg, lg, mode, i0 + t, j, dc) << (8 * t);
    return w;
}
// angular word. refbuf: 4-byte aligned, refbuf[32 + i] = ref[i] of ref_sample(); a = angle_of(mode). Two pixels share one
// 32-bit multiply-add (16-bit fields: 255 * 32 + 16 fits); a fraction of 0 reproduces the sample itself, so there is no branch.
OIS_HD uint32_t angular_pixels(const uint8_t* refbuf, const uint8_t* r, int lg, bool vertical, int a, int j, int i0, int pixels) {
    const int n = 1 << lg;
    const int pos = (j + 1) * a, idx = pos >> 5;
    const uint32_t f = (uint32_t)(pos & 31);
    const int b = 32 + i0 + idx + 1;
    const uint32_t* w = reinterpret_cast<const uint32_t*>(refbuf) + (b >> 2);
    const uint32_t w0 = w[0], w1 = w[1];
    const int s = (b & 3) * 8;
#if defined(__CUDA_ARCH__)
    const uint32_t lo = __funnelshift_r(w0, w1, s);
#else
    const uint32_t lo = (uint32_t)((((uint64_t)w1 << 32) | w0) >> s);
#endif
    const uint32_t b4 = (w1 >> s) & 0xFFu;
    const uint32_t e02 = lo & 0x00FF00FFu, e13 = (lo >> 8) & 0x00FF00FFu, e24 = (e02 >> 16) | (b4 << 16);
    const uint32_t p02 = ((e02 * (32u - f) + e13 * f + 0x00100010u) >> 5) & 0x00FF00FFu;
    const uint32_t p13 = ((e13 * (32u - f) + e24 * f + 0x00100010u) >> 5) & 0x00FF00FFu;
    uint32_t p = p02 | (p13 << 8);
    if (a == 0 && lg < 5 && i0 == 0) {                              // boundary smoothing of the pure modes: first pixel of the line
        const int side = vertical ? r[2 * n - 1 - j] : r[2 * n + 1 + j];
        p = (p & ~0xFFu) | (uint32_t)clip8((int)refbuf[33] + ((side - (int)refbuf[32]) >> 1));
    }
    return pixels == 2 ? (p & 0xFFFFu) : p;
}
OIS_HD uint32_t angular_word(const uint8_t* refbuf, const uint8_t* r, int lg, bool vertical, int a, int j, int i0) {
    return angular_pixels(refbuf, r, lg, vertical, a, j, i0, word_pixels(lg));
}

// ---- line-per-lane prediction (the shipped kernel): lane gl of a group of n lanes owns line j = gl of the CU (a row for the
// vertical modes, planar and DC, a column for the horizontal ones) as n / 4 words of 4 pixels; 32 / n modes run side by side
OIS_HD uint32_t dc_line_word(const uint8_t* r, int lg, int j, int i0, int dc) {
    const int n = 1 << lg;
    const uint8_t* left = r + 2 * n - 1;
    const uint8_t* top = r + 2 * n + 1;
    uint32_t w = (uint32_t)dc * 0x01010101u;
    if (lg < 5) {                                                    // the filtered first row and first column of 8x8 / 16x16 CUs
        if (j == 0) {
            w = 0;
            for (int t = 0; t < 4; t++) w |= (uint32_t)((top[i0 + t] + 3 * dc + 2) >> 2) << (8 * t);
        }
        if (i0 == 0) w = (w & ~0xFFu) | (uint32_t)(j == 0 ? (left[0] + 2 * dc + top[0] + 2) >> 2 : (left[-j] + 3 * dc + 2) >> 2);
    }
    return w;
}
OIS_HD uint32_t planar_line_word(const uint8_t* r, int lg, int j, int i0) {
    uint32_t w = 0;
    for (int t = 0; t < 4; t++) w |= (uint32_t)pred_pixel(r, nullptr, 1 << lg, lg, kPlanar, i0 + t, j, 0) << (8 * t);
    return w;
}

// ---- candidate lists (per CU, after the SADs) --------------------------------------------------------------------------
OIS_HD uint32_t pack(uint32_t distortion, int valid, int mode) {     // byte image of OisCandidate_t (EbCodingUnit.h:303-313)
    return (distortion & 0xFFFFFu) | ((uint32_t)(valid & 1) << 20) | ((uint32_t)mode << 24);
}
OIS_HD int islice_mode(int i) { return (int)((0x2248268A040ull >> (6 * i)) & 63); }        // iSliceModesArray[0..6] = 0 1 10 26 2 18 34
OIS_HD int stage1_mode(int i) { return (int)((0x1E58E1A248268Aull >> (6 * i)) & 63); }     // stage1ModesArray = 10 26 2 18 34 6 14 22 30
OIS_HD int points_modes(int point) { return 2 * point + 1; }          // numberOfOisModePoints = intraSearchInMd[point][depth 1..3] = 1, 3, 5, 7, 9

// OIS point from the DC SAD and the ME SAD of the same square (GetInterIntraSadDistance + GetOisPoint, ME.c:7747-7826)
OIS_HD int ois_point(const Params& q, uint32_t me, uint32_t dc_sad) {
    const bool is_4k = q.input_resolution == 3;
    const int set = is_4k ? ((q.temporal_layer_index == 0 || q.is_used_as_reference) ? 2 : 1) : 2;
    const int tl = q.temporal_layer_index;
    const int32_t diff = (int32_t)(me - dc_sad) * 100;
    const int32_t dist = dc_sad ? diff / (int32_t)dc_sad : 0;
    int t0, t1, t2, t3;                                               // OisPointTh[set][layer], ME.c:28-58
    if (set == 2) { t0 = -400; t1 = -300; t2 = -200; t3 = 0; }
    else if (set == 1) { t0 = tl < 2 ? -150 : tl == 2 ? -125 : -50; t1 = tl < 3 ? 0 : 50; t2 = tl < 2 ? 150 : tl == 2 ? 100 : 200; t3 = tl < 2 ? 200 : tl == 2 ? 150 : 300; }
    else { t0 = -20; t1 = 50; t2 = tl < 2 ? 150 : tl == 2 ? 100 : 200; t3 = tl < 2 ? 200 : tl == 2 ? 150 : 300; }
    if (dc_sad == 0 || me == 0 || dist <= t0) return 0;
    return dist <= t1 ? 1 : dist <= t2 ? 2 : dist <= t3 ? 3 : 4;
}

// InjectIntraCandidatesBasedOnBestMode (ME.c:7525-7745): list k of the stage-1 winner `best`; slot = stage-1 index its SAD comes from
OIS_HD int inject_mode(int best, int k) {
    if (k == 0) return best;
    if (k == 1) return kDc;
    if (k == 2) return kPlanar;
    const int j = k - 3;                                              // 0..5
    switch (best) {
    case 2:    return j < 3 ? 3 + j : 4 + j;                          // 3 4 5 7 8 9
    case kHor: { const int t[6] = { 9, 11, 8, 12, 7, 13 }; return t[j]; }
    case 18:   { const int t[6] = { 17, 19, 16, 20, 15, 21 }; return t[j]; }
    case kVer: { const int t[6] = { 25, 27, 24, 28, 23, 29 }; return t[j]; }
    case 34:   { const int t[6] = { 33, 32, 29, 31, 27, 28 }; return t[j]; }
    case 6:    { const int t[6] = { 7, 5, 4, 8, 3, 9 }; return t[j]; }
    case 14:   { const int t[6] = { 13, 15, 12, 16, 11, 17 }; return t[j]; }
    case 22:   { const int t[6] = { 21, 23, 20, 24, 19, 25 }; return t[j]; }
    default:   { const int t[6] = { 29, 31, 28, 32, 27, 33 }; return t[j]; }   // 30
    }
}
OIS_HD int inject_slot(int best) { return best == kHor ? 0 : best == kVer ? 1 : best == 2 ? 2 : best == 18 ? 3 : best == 34 ? 4 : best == 6 ? 5 : best == 14 ? 6 : best == 22 ? 7 : 8; }
OIS_HD int inject_canonical(int best) {                               // the switch's `default` is the list of mode 30
    return (best == 2 || best == kHor || best == 18 || best == kVer || best == 34 || best == 6 || best == 14 || best == 22) ? best : 30;
}

// ---- phase 2 of a CU, shared by the kernel and its host emulation --------------------------------------------------------
// sad: the CU's SAD row as phase 1 left it -- intra pictures: [i] = islice_mode(i) (32x32: [0] = planar only); layer-0 inter
// pictures: [m] = mode m, all 35; other inter pictures: [0..8] = stage1_mode(i) for the first points_modes(point) entries,
// [9] = DC. point: the OIS point (0 when the path has none).
OIS_HD bool all_modes_path(const Params& q) { return q.slice_type != 0 && q.temporal_layer_index == 0 && q.input_resolution != 3; }

// the stage-1 / intra-picture winner of the CU, -1 when the CU ran no such search or nothing beat the initial 32 * 32 * 255
OIS_HD int cu_winner(const Params& q, int n, const uint32_t* sad, int point) {
    int count = 0;
    if (q.slice_type == 0) count = n == 32 ? 0 : 7;
    else if (!all_modes_path(q) && !q.limit_ois_to_dc && point > 0) count = points_modes(point);
    uint32_t best = 32 * 32 * 255;
    int bm = -1;
    for (int i = 0; i < count; i++)
        if (sad[i] < best) { best = sad[i]; bm = q.slice_type == 0 ? islice_mode(i) : stage1_mode(i); }
    return bm;
}

// c[18] (byte images of OisCandidate_t, zero where the reference writes nothing) and total_intra_luma_mode of the CU;
// bm: the winner in force for this CU (its own, else the last earlier one, else planar)
OIS_HD int build_lists(const Params& q, int n, const uint32_t* sad, int point, int bm, uint32_t* c) {
    for (int i = 0; i < kMaxCand; i++) c[i] = 0;
    if (q.slice_type == 0) {
        if (n == 32) { c[0] = pack(sad[0], 1, kPlanar); return 0; }
        c[0] = pack(sad[0], 1, kPlanar); c[1] = pack(0, 0, kDc);                       // InjectIntraCandidatesBasedOnBestModeIslice
        if (bm == kPlanar || bm == kDc) return 2;
        c[2] = pack(0, 0, bm == 2 ? 2 : bm == kHor ? kHor : bm == 18 ? 18 : bm == kVer ? kVer : 34);
        c[3] = pack(0, 0, bm == 2 ? 4 : bm == kHor ? 6 : bm == 18 ? 14 : bm == kVer ? 22 : 32);
        c[4] = pack(0, 0, bm == 2 ? 6 : bm == kHor ? 14 : bm == 18 ? 22 : 30);
        return 5;
    }
    if (all_modes_path(q)) {
        // SortIntraModesOpenLoop + SortOisCandidateOpenLoop (ME.c:7828-7913): modes 0..17 seed the list, a later mode replaces
        // the FIRST worst entry when strictly better, then the reference's swap sort (its order among equal SADs is kept)
        // (every index into d / m below is a compile-time constant once unrolled, so that both arrays live in registers)
        uint32_t d[kMaxCand], m[kMaxCand];
        OIS_UNROLL
        for (int i = 0; i < kMaxCand; i++) { d[i] = sad[i]; m[i] = (uint32_t)i; }
        for (int mode = kMaxCand; mode < kNumModes; mode++) {
            uint32_t worst_d = d[0];
            int worst = 0;
            OIS_UNROLL
            for (int i = 1; i < kMaxCand; i++) if (d[i] > worst_d) { worst_d = d[i]; worst = i; }
            const uint32_t s = sad[mode];
            if (s < worst_d) {
                OIS_UNROLL
                for (int i = 0; i < kMaxCand; i++) if (i == worst) { d[i] = s; m[i] = (uint32_t)mode; }
            }
        }
        OIS_UNROLL
        for (int i = 0; i < kMaxCand; i++) {
            OIS_UNROLL
            for (int j = i + 1; j < kMaxCand; j++)
                if (d[i] > d[j]) { const uint32_t t = d[i]; d[i] = d[j]; d[j] = t; const uint32_t u = m[i]; m[i] = m[j]; m[j] = u; }
        }
        OIS_UNROLL
        for (int i = 0; i < kMaxCand; i++) c[i] = pack(d[i], 0, (int)m[i]);
        return kMaxCand;
    }
    if (q.limit_ois_to_dc) { c[0] = pack(sad[9], 1, kDc); return 1; }              // OpenLoopIntraDC
    if (point == 0) { c[0] = pack(sad[9], 0, kDc); return 1; }                     // OIS_VERY_FAST_MODE
    const int b = inject_canonical(bm);
    // quirk: with ENCODER_MODE_CLEANUP every case of the reference's switch sets valid = (layer > 1) -- except the vertical
    // winner's, which kept its enc-mode test (ME.c:7604-7612): TRUE above enc-mode 1
    const int valid = (b == kVer && q.enc_mode > 1) ? 1 : q.temporal_layer_index > 1;
    c[0] = pack(sad[inject_slot(b)], valid, b);
    for (int k = 1; k < 9; k++) c[k] = pack(0, 0, inject_mode(b, k));
    return points_modes(point);
}

}  // namespace ois
}  // namespace svtme
