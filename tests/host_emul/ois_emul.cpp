// ois_emul.cpp -- serial host emulation of ois_kernel (svt-av1-1_b200/csrc/ois.cu): the same ois_core.cuh functions, called in
// the kernel's order (phase 1: neighbours -> per-mode main reference -> per-pixel prediction and SAD; phase 2a: winners; phase
// 2b: look-back + candidate lists), so that the arithmetic the kernel shares with this file is checked against the oracle on
// machines without a GPU. Test infrastructure only: nothing in the product links it.
#include <string.h>
#include "../../svt-av1-1_b200/csrc/ois_core.cuh"

using namespace svtme::ois;

extern "C" void ois_emul_picture(const uint8_t* pix, int stride, int w, int h, const Params* prm, const uint32_t* me_sad,
                                 uint32_t* out_cand, uint8_t* out_total) {
    const Params q = *prm;
    const Picture pic = { pix, stride, w, h };
    const int sb_cols = (w + 63) / 64, sb_rows = (h + 63) / 64;
    const bool all_modes = all_modes_path(q);
    for (int sb = 0; sb < sb_cols * sb_rows; sb++) {
        static uint32_t s_sad[kCus][kNumModes];
        int8_t s_point[kCus], s_best[kCus];
        memset(s_sad, 0xAB, sizeof(s_sad));                              // shared memory is not initialised either
        for (int i = 0; i < kCus; i++) s_point[i] = s_best[i] = -1;
        const int ox = (sb % sb_cols) * 64, oy = (sb / sb_cols) * 64;
        for (int cu = 1; cu < kCus; cu++) {
            int cx, cy, n;
            cu_geometry(cu, cx, cy, n);
            const int x0 = ox + cx, y0 = oy + cy, lg = n == 8 ? 3 : n == 16 ? 4 : 5;
            if (x0 + n > pic.w || y0 + n > pic.h) continue;
            if (q.slice_type != 0 && q.cu8x8_mode == 1 && n == 8) continue;
            uint8_t r[132], refbuf[104], *ref = refbuf + 32;
            memset(r, 0xCD, sizeof(r)); memset(refbuf, 0xEF, sizeof(refbuf));
            for (int i = 0; i <= 4 * n; i++) r[i] = (uint8_t)neighbour(pic, x0, y0, n, i);
            int dc = 0;
            for (int i = 0; i < 2 * n; i++) dc += i < n ? r[2 * n + 1 + i] : r[2 * n - 1 - (i - n)];
            dc = (dc + n) >> (lg + 1);
            const uint8_t* src = pic.pix + (long long)y0 * pic.stride + x0;
            auto mode_sad = [&](int mode) -> uint32_t {
                if (mode >= 2) for (int i = ref_first(n, mode); i <= 2 * n; i++) ref[i] = (uint8_t)ref_sample(r, n, mode, i);
                uint32_t sum = 0;
                for (int px = 0; px < n * n; px++) {
                    const int x = px & (n - 1), y = px >> lg;
                    const int d = (int)src[(long long)y * pic.stride + x] - pred_pixel(r, ref, n, lg, mode, x, y, dc);
                    sum += (uint32_t)(d < 0 ? -d : d);
                }
                return sum;
            };
            if (q.slice_type == 0) {
                if (n == 32) s_sad[cu][0] = mode_sad(kPlanar);
                else for (int i = 0; i < 7; i++) s_sad[cu][i] = mode_sad(islice_mode(i));
                s_point[cu] = 0;
            } else if (all_modes) {
                for (int m = 0; m < kNumModes; m++) s_sad[cu][m] = mode_sad(m);
                s_point[cu] = 0;
            } else {
                const uint32_t dcs = mode_sad(kDc);
                int point = 0;
                if (!q.limit_ois_to_dc) {
                    point = ois_point(q, me_sad[(size_t)sb * kCus + cu], dcs);
                    if (point) for (int i = 0; i < points_modes(point); i++) s_sad[cu][i] = mode_sad(stage1_mode(i));
                }
                s_sad[cu][9] = dcs; s_point[cu] = (int8_t)point;
            }
        }
        for (int cu = 1; cu < kCus; cu++) {
            int cx, cy, n;
            cu_geometry(cu, cx, cy, n);
            if (s_point[cu] >= 0) s_best[cu] = (int8_t)cu_winner(q, n, s_sad[cu], s_point[cu]);
        }
        for (int cu = 0; cu < kCus; cu++) {
            uint32_t* c = out_cand + ((size_t)sb * kCus + cu) * kMaxCand;
            int total = 0;
            if (cu >= 1 && s_point[cu] >= 0) {
                int cx, cy, n;
                cu_geometry(cu, cx, cy, n);
                int bm = s_best[cu];
                for (int k = cu - 1; bm < 0 && k >= 1; k--) bm = s_best[k];
                if (bm < 0) bm = kPlanar;
                total = build_lists(q, n, s_sad[cu], s_point[cu], bm, c);
            } else {
                for (int i = 0; i < kMaxCand; i++) c[i] = 0;
            }
            out_total[(size_t)sb * kCus + cu] = (uint8_t)total;
        }
    }
}

// The word-wise kernel (ois_words_kernel), lane by lane: source words in "registers" per lane, the main reference in a word
// aligned buffer whose unwritten bytes hold junk, flat_word / angular_word / sad_word per word, the sum over lanes and words.
template <int LG>
static uint32_t words_mode_sad(const Picture& pic, int x0, int y0, const uint8_t* r, uint32_t (*refwords)[26], int mode, int dc) {
    constexpr int n = 1 << LG, W = (n * n) / (LG == 3 ? 2 : 4) / 32;
    static_assert(W == (LG == 3 ? 1 : LG == 4 ? 2 : 8), "words per lane");
    const uint8_t* src = pic.pix + (long long)y0 * pic.stride + x0;
    // [0] / [1]: the vertical / horizontal main reference; i >= 0 was filled once for the CU, i < 0 is projected per mode and
    // whatever an earlier mode left below ref_first stays there
    uint8_t* refbuf = reinterpret_cast<uint8_t*>(refwords[mode >= 18 ? 0 : 1]);
    if (mode >= 2 && angle_of(mode) < 0)
        for (int lane = 0; lane < 32; lane++) { const int i = ref_first(n, mode) + lane; if (i < 0) refbuf[32 + i] = (uint8_t)ref_sample(r, n, mode, i); }
    uint32_t sum = 0;
    for (int lane = 0; lane < 32; lane++)
        for (int k = 0; k < W; k++) {
            int j, i0;
            word_pos(LG, lane, k, j, i0);
            if (k >= words_per_lane(LG)) return ~0u;
            const uint32_t srow = source_word(src, pic.stride, LG, j, i0, false), scol = source_word(src, pic.stride, LG, j, i0, true);
            if (mode < 2) sum += sad_word(flat_word(r, LG, mode, j, i0, dc), srow);
            else sum += sad_word(angular_word(refbuf, r, LG, mode >= 18, angle_of(mode), j, i0), mode >= 18 ? srow : scol);
        }
    return sum;
}

extern "C" void ois_emul_picture_words(const uint8_t* pix, int stride, int w, int h, const Params* prm, const uint32_t* me_sad,
                                       uint32_t* out_cand, uint8_t* out_total) {
    const Params q = *prm;
    const Picture pic = { pix, stride, w, h };
    const int sb_cols = (w + 63) / 64, sb_rows = (h + 63) / 64;
    const bool all_modes = all_modes_path(q);
    for (int sb = 0; sb < sb_cols * sb_rows; sb++) {
        static uint32_t s_sad[kCus][kNumModes];
        int8_t s_point[kCus], s_best[kCus];
        memset(s_sad, 0xAB, sizeof(s_sad));
        for (int i = 0; i < kCus; i++) s_point[i] = s_best[i] = -1;
        const int ox = (sb % sb_cols) * 64, oy = (sb / sb_cols) * 64;
        for (int cu = 1; cu < kCus; cu++) {
            int cx, cy, n;
            cu_geometry(cu, cx, cy, n);
            const int x0 = ox + cx, y0 = oy + cy, lg = n == 8 ? 3 : n == 16 ? 4 : 5;
            if (x0 + n > pic.w || y0 + n > pic.h) continue;
            if (q.slice_type != 0 && q.cu8x8_mode == 1 && n == 8) continue;
            uint8_t r[132];
            uint32_t refwords[2][26];
            memset(r, 0xCD, sizeof(r)); memset(refwords, 0xEF, sizeof(refwords));
            for (int i = 0; i <= 4 * n; i++) r[i] = (uint8_t)neighbour(pic, x0, y0, n, i);
            for (int i = 0; i <= 2 * n; i++) {
                reinterpret_cast<uint8_t*>(refwords[0])[32 + i] = (uint8_t)straight_sample(r, n, true, i);
                reinterpret_cast<uint8_t*>(refwords[1])[32 + i] = (uint8_t)straight_sample(r, n, false, i);
            }
            int dc = 0;
            for (int i = 0; i < 2 * n; i++) dc += i < n ? r[2 * n + 1 + i] : r[2 * n - 1 - (i - n)];
            dc = (dc + n) >> (lg + 1);
            auto mode_sad = [&](int mode) -> uint32_t {
                return lg == 5 ? words_mode_sad<5>(pic, x0, y0, r, refwords, mode, dc)
                     : lg == 4 ? words_mode_sad<4>(pic, x0, y0, r, refwords, mode, dc) : words_mode_sad<3>(pic, x0, y0, r, refwords, mode, dc);
            };
            if (q.slice_type == 0) {
                if (n == 32) s_sad[cu][0] = mode_sad(kPlanar);
                else for (int i = 0; i < 7; i++) s_sad[cu][i] = mode_sad(islice_mode(i));
                s_point[cu] = 0;
            } else if (all_modes) {
                for (int m = 0; m < kNumModes; m++) s_sad[cu][m] = mode_sad(m);
                s_point[cu] = 0;
            } else {
                const uint32_t dcs = mode_sad(kDc);
                int point = 0;
                if (!q.limit_ois_to_dc) {
                    point = ois_point(q, me_sad[(size_t)sb * kCus + cu], dcs);
                    if (point) for (int i = 0; i < points_modes(point); i++) s_sad[cu][i] = mode_sad(stage1_mode(i));
                }
                s_sad[cu][9] = dcs; s_point[cu] = (int8_t)point;
            }
        }
        for (int cu = 1; cu < kCus; cu++) {
            int cx, cy, n;
            cu_geometry(cu, cx, cy, n);
            if (s_point[cu] >= 0) s_best[cu] = (int8_t)cu_winner(q, n, s_sad[cu], s_point[cu]);
        }
        for (int cu = 0; cu < kCus; cu++) {
            uint32_t* c = out_cand + ((size_t)sb * kCus + cu) * kMaxCand;
            int total = 0;
            if (cu >= 1 && s_point[cu] >= 0) {
                int cx, cy, n;
                cu_geometry(cu, cx, cy, n);
                int bm = s_best[cu];
                for (int k = cu - 1; bm < 0 && k >= 1; k--) bm = s_best[k];
                if (bm < 0) bm = kPlanar;
                total = build_lists(q, n, s_sad[cu], s_point[cu], bm, c);
            } else {
                for (int i = 0; i < kMaxCand; i++) c[i] = 0;
            }
            out_total[(size_t)sb * kCus + cu] = (uint8_t)total;
        }
    }
}

// The line-per-lane kernel (ois_lines_kernel), pass by pass and lane by lane: n lanes per mode, 32 / n modes side by side, each
// group of lanes with its own pair of main references; the staged path (DC, then the stage-1 directions the OIS point asks for)
// runs its DC pass on group 0 alone, exactly like the kernel.
template <int LG>
static void lines_search_cu(const Picture& pic, const Params& q, const uint32_t* me_sad, int sb, int cu, int x0, int y0, uint32_t* sad_row,
                            int8_t* s_point, bool all_modes) {
    constexpr int n = 1 << LG, M = 32 / n, W = n / 4;
    uint8_t r[132];
    uint32_t refs[4][2][26];
    memset(r, 0xCD, sizeof(r)); memset(refs, 0xEF, sizeof(refs));
    for (int i = 0; i <= 4 * n; i++) r[i] = (uint8_t)neighbour(pic, x0, y0, n, i);
    int dc = 0;
    for (int i = 0; i < 2 * n; i++) dc += i < n ? r[2 * n + 1 + i] : r[2 * n - 1 - (i - n)];
    dc = (dc + n) >> (LG + 1);
    for (int g = 0; g < M; g++)
        for (int i = 0; i <= 2 * n; i++) {
            reinterpret_cast<uint8_t*>(refs[g][0])[32 + i] = (uint8_t)straight_sample(r, n, true, i);
            reinterpret_cast<uint8_t*>(refs[g][1])[32 + i] = (uint8_t)straight_sample(r, n, false, i);
        }
    const uint8_t* src = pic.pix + (long long)y0 * pic.stride + x0;
    const bool intra = q.slice_type == 0, staged = !intra && !all_modes;
    int count = intra ? (n == 32 ? 1 : 7) : all_modes ? kNumModes : 1, point = 0, base = 0;
    while (base < count) {
        const bool dc_pass = staged && base == 0;
        uint32_t group_sum[4] = { 0, 0, 0, 0 };
        for (int g = 0; g < M; g++) {
            const int p = base + g;
            const bool active = dc_pass ? g == 0 : p < count;
            int mode = kDc, slot = 9;
            if (active && !dc_pass) {
                slot = staged ? p - 1 : p;
                mode = intra ? islice_mode(p) : all_modes ? p : stage1_mode(p - 1);
            }
            const bool angular = mode >= 2, vertical = mode >= 18;
            const int a = angular ? angle_of(mode) : 0;
            uint8_t* refbuf = reinterpret_cast<uint8_t*>(refs[g][vertical ? 0 : 1]);
            if (active && a < 0)
                for (int gl = 0; gl < n; gl++) { const int i = ref_first(n, mode) + gl; if (i < 0) refbuf[32 + i] = (uint8_t)ref_sample(r, n, mode, i); }
            if (!active) continue;
            for (int gl = 0; gl < n; gl++)
                for (int k = 0; k < W; k++) {
                    const uint32_t srow = source_pixels(src, pic.stride, 4, gl, 4 * k, false), scol = source_pixels(src, pic.stride, 4, gl, 4 * k, true);
                    if (angular) group_sum[g] += sad_word(angular_pixels(refbuf, r, LG, vertical, a, gl, 4 * k, 4), vertical ? srow : scol);
                    else if (mode == kDc) group_sum[g] += sad_word(dc_line_word(r, LG, gl, 4 * k, dc), srow);
                    else group_sum[g] += sad_word(planar_line_word(r, LG, gl, 4 * k), srow);
                }
            sad_row[slot] = group_sum[g];
        }
        if (dc_pass) {
            base = 1;
            if (!q.limit_ois_to_dc) {
                point = ois_point(q, me_sad[(size_t)sb * kCus + cu], group_sum[0]);
                if (point) count = 1 + points_modes(point);
            }
        } else {
            base += M;
        }
    }
    s_point[cu] = (int8_t)point;
}

extern "C" void ois_emul_picture_lines(const uint8_t* pix, int stride, int w, int h, const Params* prm, const uint32_t* me_sad,
                                       uint32_t* out_cand, uint8_t* out_total) {
    const Params q = *prm;
    const Picture pic = { pix, stride, w, h };
    const int sb_cols = (w + 63) / 64, sb_rows = (h + 63) / 64;
    const bool all_modes = all_modes_path(q);
    for (int sb = 0; sb < sb_cols * sb_rows; sb++) {
        static uint32_t s_sad[kCus][kNumModes];
        int8_t s_point[kCus], s_best[kCus];
        memset(s_sad, 0xAB, sizeof(s_sad));
        for (int i = 0; i < kCus; i++) s_point[i] = s_best[i] = -1;
        const int ox = (sb % sb_cols) * 64, oy = (sb / sb_cols) * 64;
        for (int cu = 1; cu < kCus; cu++) {
            int cx, cy, n;
            cu_geometry(cu, cx, cy, n);
            const int x0 = ox + cx, y0 = oy + cy;
            if (x0 + n > pic.w || y0 + n > pic.h) continue;
            if (q.slice_type != 0 && q.cu8x8_mode == 1 && n == 8) continue;
            if (n == 32) lines_search_cu<5>(pic, q, me_sad, sb, cu, x0, y0, s_sad[cu], s_point, all_modes);
            else if (n == 16) lines_search_cu<4>(pic, q, me_sad, sb, cu, x0, y0, s_sad[cu], s_point, all_modes);
            else lines_search_cu<3>(pic, q, me_sad, sb, cu, x0, y0, s_sad[cu], s_point, all_modes);
        }
        for (int cu = 1; cu < kCus; cu++) {
            int cx, cy, n;
            cu_geometry(cu, cx, cy, n);
            if (s_point[cu] >= 0) s_best[cu] = (int8_t)cu_winner(q, n, s_sad[cu], s_point[cu]);
        }
        for (int cu = 0; cu < kCus; cu++) {
            uint32_t* c = out_cand + ((size_t)sb * kCus + cu) * kMaxCand;
            int total = 0;
            if (cu >= 1 && s_point[cu] >= 0) {
                int cx, cy, n;
                cu_geometry(cu, cx, cy, n);
                int bm = s_best[cu];
                for (int k = cu - 1; bm < 0 && k >= 1; k--) bm = s_best[k];
                if (bm < 0) bm = kPlanar;
                total = build_lists(q, n, s_sad[cu], s_point[cu], bm, c);
            } else {
                for (int i = 0; i < kMaxCand; i++) c[i] = 0;
            }
            out_total[(size_t)sb * kCus + cu] = (uint8_t)total;
        }
    }
}
