/* oracle/enc_driver.c -- TEST INFRASTRUCTURE ONLY.
 * Runs the UNMODIFIED reference encoder (every process, through its public API, Source/API/EbApi.h) from
 * oracle/_ref/libsvtref.so on a raw 8-bit yuv420p file and dumps the motion-estimation stage of that encode through the
 * MotionEstimateLcu wrapper in ref_driver.c: BASELINE configs[0] ("run the encoder on CPU, dump MotionEstimateLcu
 * MVs/SADs as golden"). usage: enc_dump in.yuv width height frames enc_mode out.medump [--bitstream out.obu] [--cuda-me lib.so]
 *   out.medump      "-" = no dump
 *   --txq-dump      record forward transform + quantisation blocks of the EncDec stage (ref_txq.c wrappers)
 *   --bitstream     write the packets the encoder returns, in order
 *   --cuda-txq      same for the secondary family: every 8-bit Av1EstimateTransform / EncDec Av1QuantizeInvQuantize call of the
 *                   encode is served by svttxq_fwd_txfm_quant (include/svttxq.h), one block per call
 *   --cuda-me       encoder-in-the-loop: dlopen libsvtme_cuda.so and serve every MotionEstimateLcu call of the encode from
 *                   its SB facade (include/svtme_lcu.h) instead of the reference's C function. This hook is the whole
 *                   reference-side binding; the bitstream must not change.
 */
#define _GNU_SOURCE
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <dlfcn.h>
#include "EbApi.h"
#include "svtme_lcu.h"
#include "svttxq.h"
#include <pthread.h>

int ref_me_dump_open(const char *path);
void ref_me_dump_close(void);
int ref_txq_dump_open(const char *path, int every, int cap);
void ref_txq_dump_close(void);
typedef int (*ref_me_hook_fn)(void *user, const int32_t *hdr, const uint8_t *cur, int cur_stride, const uint8_t *ref0, int ref0_stride,
                              const uint8_t *ref1, int ref1_stride, void *row_out);
void ref_me_hook_set(ref_me_hook_fn fn, void *user);
void ref_me_timing_enable(int on);
void ref_me_timing_get(long long *ns, long long *calls);
#include <time.h>

/* ---- the reference-side binding: MotionEstimateLcu -> svtme_lcu_motion_estimate ------------------------------------ */
static struct {
    void *lib;
    SvtMeLcuSession *session;
    void (*default_params)(SvtMeParams *, int32_t, int32_t, int32_t, int32_t, int32_t, int32_t);
    int32_t (*create)(SvtMeLcuSession **, int32_t, int32_t, int32_t, int32_t);
    void (*destroy)(SvtMeLcuSession *);
    const char *(*last_error)(const SvtMeLcuSession *);
    int32_t (*lcu)(SvtMeLcuSession *, const SvtMeParams *, const SvtMeLcuPicture *, const SvtMeLcuPicture *, const SvtMeLcuPicture *, int32_t, SvtMePuResult *);
    void (*stats)(const SvtMeLcuSession *, int64_t *, int64_t *, int64_t *);
} g_cuda;

static int cuda_me_hook(void *user, const int32_t *h, const uint8_t *cur, int cur_stride, const uint8_t *ref0, int ref0_stride,
                        const uint8_t *ref1, int ref1_stride, void *row_out) {
    (void)user;
    SvtMeParams p;                                        /* what MotionEstimateLcu reads from PCS / SCS / MeContext */
    g_cuda.default_params(&p, h[5], h[6], h[7], h[8], h[9], h[10]);
    p.same_ref_poc = h[7] == 2 && h[11] == h[12];
    p.enable_hme_flag = h[13]; p.enable_hme_level0 = h[14]; p.enable_hme_level1 = h[15]; p.enable_hme_level2 = h[16];
    p.use_subpel = h[17]; p.num_pus = h[18]; p.cu8x8_mode = h[19];
    p.num_hme_regions_w = h[20]; p.num_hme_regions_h = h[21]; p.hme_l0_total_w = h[22]; p.hme_l0_total_h = h[23];
    for (int i = 0; i < 2; i++) {
        p.hme_l0_w[i] = h[24 + i]; p.hme_l0_h[i] = h[26 + i]; p.hme_l1_w[i] = h[28 + i]; p.hme_l1_h[i] = h[30 + i];
        p.hme_l2_w[i] = h[32 + i]; p.hme_l2_h[i] = h[34 + i];
    }
    p.search_area_w = h[36]; p.search_area_h = h[37];
    const SvtMeLcuPicture c = {(uint64_t)h[1], cur, cur_stride, 0}, r0 = {(uint64_t)h[11], ref0, ref0_stride, 0}, r1 = {(uint64_t)h[12], ref1, ref1_stride, 0};
    const int32_t rc = g_cuda.lcu(g_cuda.session, &p, &c, &r0, h[7] == 2 ? &r1 : NULL, h[2], (SvtMePuResult *)row_out);
    if (rc != 0) fprintf(stderr, "svtme_lcu_motion_estimate: 0x%08x %s\n", (unsigned)rc, g_cuda.last_error(g_cuda.session));
    return rc;
}

static int cuda_me_open(const char *path, int w, int h) {
    g_cuda.lib = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    if (!g_cuda.lib) { fprintf(stderr, "dlopen %s: %s\n", path, dlerror()); return -1; }
#define SYM(field, name) do { *(void **)&g_cuda.field = dlsym(g_cuda.lib, name); if (!g_cuda.field) { fprintf(stderr, "missing %s\n", name); return -1; } } while (0)
    SYM(default_params, "svtme_default_params"); SYM(create, "svtme_lcu_session_create"); SYM(destroy, "svtme_lcu_session_destroy");
    SYM(last_error, "svtme_lcu_last_error"); SYM(lcu, "svtme_lcu_motion_estimate"); SYM(stats, "svtme_lcu_session_stats");
#undef SYM
    const int32_t rc = g_cuda.create(&g_cuda.session, 0, w, h, 16);
    if (rc != 0) { fprintf(stderr, "svtme_lcu_session_create failed: 0x%08x (no CUDA device?)\n", (unsigned)rc); return -1; }
    ref_me_hook_set(cuda_me_hook, NULL);
    return 0;
}

/* ---- the binding of the secondary family: Av1EstimateTransform + Av1QuantizeInvQuantize -> svttxq_fwd_txfm_quant ------------
 * One block per call, because that is the granularity of the reference's call sites; this proves results, not speed
 * (the batched, plane-level use is INTEGRATION.md section 4). */
typedef int (*ref_txq_hook_fn)(void *user, const int16_t *residual, int w, int h, int tx_type, const int16_t *q10, int32_t *coeff,
                               int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob, uint64_t *energy);
void ref_txq_hook_set(ref_txq_hook_fn fn, void *user);
void ref_txq_hook_stats(long *transforms, long *quants);
static struct {
    void *lib;
    SvtTxqContext *ctx;
    pthread_mutex_t mu;
    int32_t (*create)(SvtTxqContext **, int32_t);
    void (*destroy)(SvtTxqContext *);
    const char *(*last_error)(const SvtTxqContext *);
    int32_t (*run)(SvtTxqContext *, const int16_t *, int32_t, int32_t, int32_t, int32_t, int32_t, int32_t, const SvtTxqQuant *,
                   int32_t *, int32_t *, int32_t *, uint16_t *, uint64_t *);
} g_txq;

static int cuda_txq_hook(void *user, const int16_t *residual, int w, int h, int tx_type, const int16_t *q10, int32_t *coeff,
                         int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob, uint64_t *energy) {
    (void)user;
    SvtTxqQuant q;
    if (q10) {
        for (int i = 0; i < 2; i++) { q.zbin[i] = q10[i]; q.round[i] = q10[2 + i]; q.quant[i] = q10[4 + i]; q.quant_shift[i] = q10[6 + i]; q.dequant[i] = q10[8 + i]; }
    } else {                                              /* transform only: any legal quantiser, its outputs are dropped */
        for (int i = 0; i < 2; i++) { q.zbin[i] = 16; q.round[i] = 8; q.quant[i] = -1; q.quant_shift[i] = 16384; q.dequant[i] = 16; }
    }
    pthread_mutex_lock(&g_txq.mu);
    const int32_t rc = g_txq.run(g_txq.ctx, residual, w, w, h, w, h, tx_type, &q, coeff, q10 ? qcoeff : NULL, q10 ? dqcoeff : NULL,
                                 q10 ? eob : NULL, energy);
    if (rc != 0) fprintf(stderr, "svttxq_fwd_txfm_quant: 0x%08x %s\n", (unsigned)rc, g_txq.last_error(g_txq.ctx));
    pthread_mutex_unlock(&g_txq.mu);
    return rc;
}

static int cuda_txq_open(const char *path) {
    g_txq.lib = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    if (!g_txq.lib) { fprintf(stderr, "dlopen %s: %s\n", path, dlerror()); return -1; }
#define SYM(field, name) do { *(void **)&g_txq.field = dlsym(g_txq.lib, name); if (!g_txq.field) { fprintf(stderr, "missing %s\n", name); return -1; } } while (0)
    SYM(create, "svttxq_create"); SYM(destroy, "svttxq_destroy"); SYM(last_error, "svttxq_last_error"); SYM(run, "svttxq_fwd_txfm_quant");
#undef SYM
    pthread_mutex_init(&g_txq.mu, NULL);
    const int32_t rc = g_txq.create(&g_txq.ctx, 0);
    if (rc != 0) { fprintf(stderr, "svttxq_create failed: 0x%08x (no CUDA device?)\n", (unsigned)rc); return -1; }
    ref_txq_hook_set(cuda_txq_hook, NULL);
    return 0;
}

int main(int argc, char **argv) {
    if (argc < 7) { fprintf(stderr, "usage: %s in.yuv width height frames enc_mode out.medump|- [--bitstream out.obu] [--txq-dump out.txqdump] [--cuda-me lib.so]\n", argv[0]); return 2; }
    const char *bitstream_path = NULL, *cuda_me = NULL, *txq_dump = NULL, *cuda_txq = NULL;
    for (int i = 7; i + 1 < argc; i += 2) {
        if (!strcmp(argv[i], "--bitstream")) bitstream_path = argv[i + 1];
        else if (!strcmp(argv[i], "--cuda-me")) cuda_me = argv[i + 1];
        else if (!strcmp(argv[i], "--txq-dump")) txq_dump = argv[i + 1];
        else if (!strcmp(argv[i], "--cuda-txq")) cuda_txq = argv[i + 1];
        else { fprintf(stderr, "unknown option %s\n", argv[i]); return 2; }
    }
    const int w = atoi(argv[2]), h = atoi(argv[3]), n = atoi(argv[4]), mode = atoi(argv[5]);
    FILE *f = fopen(argv[1], "rb");
    if (!f) { perror(argv[1]); return 2; }
    EbComponentType *enc = NULL;
    EbSvtAv1EncConfiguration cfg;
    if (eb_init_handle(&enc, NULL, &cfg) != EB_ErrorNone) { fprintf(stderr, "eb_init_handle failed\n"); return 1; }
    cfg.source_width = (uint32_t)w; cfg.source_height = (uint32_t)h;
    cfg.enc_mode = (uint8_t)mode;
    cfg.framesToBeEncoded = (uint64_t)n;
    cfg.encoder_bit_depth = 8;
    cfg.frame_rate = 30 << 16; cfg.frame_rate_numerator = 0; cfg.frame_rate_denominator = 0;
    if (eb_svt_enc_set_parameter(enc, &cfg) != EB_ErrorNone) { fprintf(stderr, "eb_svt_enc_set_parameter failed\n"); return 1; }
    if (strcmp(argv[6], "-") && ref_me_dump_open(argv[6]) != 0) { perror(argv[6]); return 2; }
    if (txq_dump && ref_txq_dump_open(txq_dump, 0, 0) != 0) { perror(txq_dump); return 2; }
    if (cuda_me && cuda_me_open(cuda_me, w, h) != 0) return 3;
    if (cuda_txq && cuda_txq_open(cuda_txq) != 0) return 3;
    FILE *bs = bitstream_path ? fopen(bitstream_path, "wb") : NULL;
    if (bitstream_path && !bs) { perror(bitstream_path); return 2; }
    if (eb_init_encoder(enc) != EB_ErrorNone) { fprintf(stderr, "eb_init_encoder failed\n"); return 1; }
    {   /* the patched library (integration/svt_cuda_me.patch) creates its CUDA session on first use: do it ahead of the timed encode,
           like cuda_me_open above does for the hook */
        int (*prepare)(uint32_t, uint32_t) = (int (*)(uint32_t, uint32_t))dlsym(RTLD_DEFAULT, "svt_cuda_me_open");
        if (prepare) prepare((uint32_t)w, (uint32_t)h);
    }
    ref_me_timing_enable(1);
    struct timespec t_begin, t_end;
    clock_gettime(CLOCK_MONOTONIC, &t_begin);

    const size_t ysz = (size_t)w * h, csz = ysz / 4;
    uint8_t *frame = (uint8_t *)malloc(ysz + 2 * csz);
    EbSvtEncInput io; memset(&io, 0, sizeof(io));
    io.luma = frame; io.cb = frame + ysz; io.cr = frame + ysz + csz;
    io.yStride = (uint32_t)w; io.cbStride = io.crStride = (uint32_t)w / 2;
    EbBufferHeaderType in; memset(&in, 0, sizeof(in));
    in.size = sizeof(in); in.p_buffer = (uint8_t *)&io;
    int sent = 0, done = 0; size_t bytes = 0; int packets = 0;
    while (!done) {
        if (sent < n && fread(frame, 1, ysz + 2 * csz, f) == ysz + 2 * csz) {
            in.n_filled_len = (uint32_t)(ysz + 2 * csz); in.flags = 0; in.pts = sent; in.pic_type = EB_INVALID_PICTURE;
            eb_svt_enc_send_picture(enc, &in);
            sent++;
        } else if (sent >= 0) {
            EbBufferHeaderType eos; memset(&eos, 0, sizeof(eos));
            eos.size = sizeof(eos); eos.flags = EB_BUFFERFLAG_EOS; eos.pic_type = EB_INVALID_PICTURE;
            eb_svt_enc_send_picture(enc, &eos);
            sent = -1;
        }
        for (;;) {                                        /* drain what is ready; block once everything has been sent */
            EbBufferHeaderType *out = NULL;
            const EbErrorType rc = eb_svt_get_packet(enc, &out, sent < 0);
            if (rc != EB_ErrorNone || !out) break;
            bytes += out->n_filled_len; packets++;
            if (bs) fwrite(out->p_buffer, 1, out->n_filled_len, bs);
            const int eos = (out->flags & EB_BUFFERFLAG_EOS) != 0;
            eb_svt_release_out_buffer(&out);
            if (eos) { done = 1; break; }
        }
    }
    clock_gettime(CLOCK_MONOTONIC, &t_end);
    {   /* one machine-readable line for profiles/encoder_timing.py: encode wall clock, time inside MotionEstimateLcu over all threads */
        long long ns = 0, calls = 0;
        ref_me_timing_get(&ns, &calls);
        fprintf(stderr, "enc_dump: timing wall_s %.3f me_thread_s %.3f me_calls %lld frames %d\n",
                (double)(t_end.tv_sec - t_begin.tv_sec) + 1e-9 * (double)(t_end.tv_nsec - t_begin.tv_nsec), 1e-9 * (double)ns, calls, n);
    }
    ref_me_dump_close();
    ref_txq_dump_close();
    if (bs) fclose(bs);
    fprintf(stderr, "enc_dump: %d packets, %zu bytes\n", packets, bytes);
    eb_deinit_encoder(enc);
    eb_deinit_handle(enc);
    if (g_cuda.session) {
        int64_t pics = 0, ups = 0, rows = 0;
        g_cuda.stats(g_cuda.session, &pics, &ups, &rows);
        fprintf(stderr, "enc_dump: cuda-me pictures %lld uploads %lld rows %lld\n", (long long)pics, (long long)ups, (long long)rows);
        ref_me_hook_set(NULL, NULL);
        g_cuda.destroy(g_cuda.session);
    }
    if (g_txq.ctx) {
        long t = 0, q = 0;
        ref_txq_hook_stats(&t, &q);
        fprintf(stderr, "enc_dump: cuda-txq transforms %ld quantisations %ld\n", t, q);
        ref_txq_hook_set(NULL, NULL);
        g_txq.destroy(g_txq.ctx);
    }
    free(frame); fclose(f);
    return 0;
}
