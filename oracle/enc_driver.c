/* oracle/enc_driver.c -- TEST INFRASTRUCTURE ONLY.
 * Runs the UNMODIFIED reference encoder (every process, through its public API, Source/API/EbApi.h) from
 * oracle/_ref/libsvtref.so on a raw 8-bit yuv420p file and dumps the motion-estimation stage of that encode through the
 * MotionEstimateLcu wrapper in ref_driver.c: BASELINE configs[0] ("run the encoder on CPU, dump MotionEstimateLcu
 * MVs/SADs as golden"). usage: enc_dump in.yuv width height frames enc_mode out.medump [hierarchical_levels]
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "EbApi.h"

int ref_me_dump_open(const char *path);
void ref_me_dump_close(void);

int main(int argc, char **argv) {
    if (argc < 7) { fprintf(stderr, "usage: %s in.yuv width height frames enc_mode out.medump [hierarchical_levels]\n", argv[0]); return 2; }
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
    if (argc > 7) cfg.hierarchical_levels = (uint32_t)atoi(argv[7]);
    if (eb_svt_enc_set_parameter(enc, &cfg) != EB_ErrorNone) { fprintf(stderr, "eb_svt_enc_set_parameter failed\n"); return 1; }
    if (ref_me_dump_open(argv[6]) != 0) { perror(argv[6]); return 2; }
    if (eb_init_encoder(enc) != EB_ErrorNone) { fprintf(stderr, "eb_init_encoder failed\n"); return 1; }

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
            const int eos = (out->flags & EB_BUFFERFLAG_EOS) != 0;
            eb_svt_release_out_buffer(&out);
            if (eos) { done = 1; break; }
        }
    }
    ref_me_dump_close();
    fprintf(stderr, "enc_dump: %d packets, %zu bytes\n", packets, bytes);
    eb_deinit_encoder(enc);
    eb_deinit_handle(enc);
    free(frame); fclose(f);
    return 0;
}
