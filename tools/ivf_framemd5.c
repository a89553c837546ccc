/*
 * ivf_framemd5.c -- drives libavcodec's "vp9" decoder (reference libavcodec/vp9.c:1965)
 * through the public avcodec API and prints one MD5 per output frame.  The same file is
 * linked twice: by oracle/Makefile against the UNMODIFIED reference objects (-> the
 * framemd5 oracle and timed CPU baseline, oracle/_ref/ffvp9_ref) and by
 * ffmpeg_hybrid_b200/host/Makefile against the overlay build whose reconstruction runs in
 * libvp9cuda.so (-> ffvp9_cuda).  It contains no decoding logic of its own.
 *
 * It stands in for `ffmpeg -i S.ivf -f framemd5 -` (reference tests/fate-run.sh:172-174):
 * the hash covers the visible w x h crop of each plane, packed without stride,
 * 10-bit samples as little-endian 16-bit -- the byte string libavcodec/rawenc.c:52-60
 * hands to the framemd5 muxer.  IVF parsing follows libavformat/ivfdec.c:29-57.
 *
 * usage: ffvp9_ref [-threads N] [-thread_type frame|slice] [-null] [-raw out.yuv] [-loop K] in.ivf
 *   -loop K feeds the packet list K times to the one decoder instance (the stream starts with a
 *   key frame, so this is a K times longer stream); with K > 1 the first pass is the warm-up
 *   and "fps=" is the rate of the K - 1 passes after it (bench.py: steady-state decode rate).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "libavcodec/avcodec.h"
#include "libavcodec/vp9data.h"     /* -tables: the default probability tables of the VP9 specification */
#include "libavutil/frame.h"
#include "libavutil/md5.h"
#include "libavutil/pixdesc.h"

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

static unsigned rl32(const unsigned char *p) { return p[0] | p[1] << 8 | p[2] << 16 | (unsigned)p[3] << 24; }

static void print_tables(const ProbContext *d, const uint8_t *cf)
{
    printf("{\"skip\": [%d, %d, %d], \"intra\": [%d, %d, %d, %d], \"partition\": [", d->skip[0], d->skip[1], d->skip[2],
           d->intra[0], d->intra[1], d->intra[2], d->intra[3]);
    for (int k = 0; k < 48; k++)
        printf("%s%d", k ? ", " : "", (&d->partition[0][0][0])[k]);
    printf("], \"coef\": [");
    for (int k = 0; k < 4 * 2 * 2 * 6 * 6 * 3; k++)
        printf("%s%d", k ? ", " : "", cf[k]);
    printf("]}\n");
    fflush(stdout);
}

/* -probe: packets come in on stdin (IVF), one at a time; after each one is decoded the SAVED probability context 0
 * of the decoder (VP9Context.prob_ctx[0], i.e. after backward adaptation, vp9prob.c) goes out as one JSON line.
 * A stream generator that writes differential probability updates (vp9.c:399-455) for frame N+1 reads it to know
 * exactly what the decoder holds (ffmpeg_hybrid_b200/streamgen.py). */
static int probe(AVCodecContext *c)
{
    unsigned char hdr[32];
    AVPacket *pkt = av_packet_alloc();
    AVFrame *fr = av_frame_alloc();
    if (fread(hdr, 1, 32, stdin) != 32 || memcmp(hdr, "DKIF", 4)) return 2;
    while (fread(hdr, 1, 12, stdin) == 12) {
        unsigned fsz = rl32(hdr);
        if (av_new_packet(pkt, fsz) < 0 || fread(pkt->data, 1, fsz, stdin) != fsz) return 2;
        if (avcodec_send_packet(c, pkt) < 0) { fprintf(stderr, "probe: decode error\n"); return 1; }
        av_packet_unref(pkt);
        while (avcodec_receive_frame(c, fr) >= 0) av_frame_unref(fr);
        const VP9Context *s = c->priv_data;
        /* coef[4][2][2][6][6][3] of the saved context is laid out like ff_vp9_default_coef_probs */
        print_tables(&s->prob_ctx[0].p, &s->prob_ctx[0].coef[0][0][0][0][0][0]);
    }
    return 0;
}

static int nframes;
static int do_hash = 1, do_probe;
static FILE *raw;

static void emit(AVFrame *f)
{
    const AVPixFmtDescriptor *d = av_pix_fmt_desc_get(f->format);
    int bps = d->comp[0].depth > 8 ? 2 : 1;
    if (do_hash || raw) {
        struct AVMD5 *md5 = av_md5_alloc();
        unsigned char dig[16];
        av_md5_init(md5);
        for (int p = 0; p < 3; p++) {
            int w = p ? AV_CEIL_RSHIFT(f->width, d->log2_chroma_w) : f->width;
            int h = p ? AV_CEIL_RSHIFT(f->height, d->log2_chroma_h) : f->height;
            for (int y = 0; y < h; y++) {
                const unsigned char *row = f->data[p] + (ptrdiff_t)y * f->linesize[p];
                av_md5_update(md5, row, w * bps);
                if (raw)
                    fwrite(row, 1, w * bps, raw);
            }
        }
        av_md5_final(md5, dig);
        av_free(md5);
        printf("%d %d %d %s ", nframes, f->width, f->height, d->name);
        for (int i = 0; i < 16; i++)
            printf("%02x", dig[i]);
        printf("\n");
    }
    nframes++;
}

int main(int argc, char **argv)
{
    int threads = 1, ttype = 0, loops = 1;
    const char *in = NULL;
    for (int i = 1; i < argc; i++) {
        if (!strcmp(argv[i], "-threads") && i + 1 < argc) threads = atoi(argv[++i]);
        else if (!strcmp(argv[i], "-thread_type") && i + 1 < argc)
            ttype = !strcmp(argv[++i], "slice") ? FF_THREAD_SLICE : FF_THREAD_FRAME;
        else if (!strcmp(argv[i], "-tables")) {
            /* the decoder's default probabilities (ff_vp9_default_probs, ff_vp9_default_coef_probs, reference
             * libavcodec/vp9data.c) as JSON: what a stream generator needs to write differential probability
             * updates (vp9.c:399-455) that land on the values it wants */
            print_tables(&ff_vp9_default_probs, &ff_vp9_default_coef_probs[0][0][0][0][0][0]);
            return 0;
        }
        else if (!strcmp(argv[i], "-probe")) do_probe = 1;
        else if (!strcmp(argv[i], "-null")) do_hash = 0;
        else if (!strcmp(argv[i], "-loop") && i + 1 < argc) loops = atoi(argv[++i]);
        else if (!strcmp(argv[i], "-raw") && i + 1 < argc) raw = fopen(argv[++i], "wb");
        else in = argv[i];
    }
    if (do_probe) {
        const AVCodec *codec = avcodec_find_decoder(AV_CODEC_ID_VP9);
        AVCodecContext *c = avcodec_alloc_context3(codec);
        c->thread_count = 1;
        c->flags |= AV_CODEC_FLAG_BITEXACT;
        if (avcodec_open2(c, codec, NULL) < 0) { fprintf(stderr, "open failed\n"); return 2; }
        return probe(c);
    }
    if (!in) {
        fprintf(stderr, "usage: %s [-threads N] [-thread_type frame|slice] [-null] [-raw f] in.ivf\n", argv[0]);
        return 2;
    }
    FILE *fp = fopen(in, "rb");
    if (!fp) { perror(in); return 2; }
    fseek(fp, 0, SEEK_END);
    long sz = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    unsigned char *buf = malloc(sz + AV_INPUT_BUFFER_PADDING_SIZE);
    if (fread(buf, 1, sz, fp) != (size_t)sz) return 2;
    memset(buf + sz, 0, AV_INPUT_BUFFER_PADDING_SIZE);
    fclose(fp);
    if (sz < 32 || memcmp(buf, "DKIF", 4)) { fprintf(stderr, "not IVF\n"); return 2; }

    const AVCodec *codec = avcodec_find_decoder(AV_CODEC_ID_VP9);
    if (!codec) { fprintf(stderr, "no vp9 decoder\n"); return 2; }
    AVCodecContext *c = avcodec_alloc_context3(codec);
    c->thread_count = threads;
    if (ttype) c->thread_type = ttype;
    c->flags |= AV_CODEC_FLAG_BITEXACT;
    c->err_recognition |= AV_EF_EXPLODE;
    if (avcodec_open2(c, codec, NULL) < 0) { fprintf(stderr, "open failed\n"); return 2; }

    AVPacket *pkt = av_packet_alloc();
    AVFrame *fr = av_frame_alloc();
    int rc = 0, npkt = 0;
    double t0 = now_s(), t1 = t0;
    int frames_warm = 0;
    for (int pass = 0; pass < (loops < 1 ? 1 : loops); pass++) {
        long pos = 32;                  /* IVF file header */
        if (pass == 1) {                /* the timed region starts with the second pass */
            t1 = now_s();
            frames_warm = nframes;
        }
        while (pos + 12 <= sz) {
            unsigned fsz = rl32(buf + pos);
            if (pos + 12 + fsz > sz) { fprintf(stderr, "truncated IVF\n"); rc = 1; break; }
            /* the packet payload must be padded: copy into an av_new_packet */
            if (av_new_packet(pkt, fsz) < 0) return 2;
            memcpy(pkt->data, buf + pos + 12, fsz);
            pkt->pts = pkt->dts = rl32(buf + pos + 4);
            pos += 12 + fsz;
            int r = avcodec_send_packet(c, pkt);
            av_packet_unref(pkt);
            /* a damaged packet is reported and skipped, the following packets are still decoded and the
             * decoder is drained at the end (what fftools does): the exit code says that something failed */
            if (r < 0) { fprintf(stderr, "decode error %d at packet %d\n", r, npkt); rc = 1; }
            npkt++;
            /* under frame threading a packet's error surfaces here, some packets later: report it and go on */
            for (;;) {
                r = avcodec_receive_frame(c, fr);
                if (r >= 0) { emit(fr); av_frame_unref(fr); continue; }
                if (r == AVERROR(EAGAIN) || r == AVERROR_EOF) break;
                fprintf(stderr, "decode error %d\n", r); rc = 1;
            }
        }
    }
    {
        avcodec_send_packet(c, NULL);
        for (int guard = 0; guard < 1000; guard++) {
            int r = avcodec_receive_frame(c, fr);
            if (r >= 0) { emit(fr); av_frame_unref(fr); continue; }
            if (r == AVERROR_EOF || r == AVERROR(EAGAIN)) break;
            fprintf(stderr, "decode error %d at drain\n", r); rc = 1;
        }
    }
    double t2 = now_s();
    if (loops > 1)   /* frames still inside the decoder at t1 are counted with the timed passes: a steady-state rate */
        fprintf(stderr, "frames=%d seconds=%.6f fps=%.3f threads=%d first_pass_fps=%.3f loops=%d\n", nframes - frames_warm,
                t2 - t1, (nframes - frames_warm) / (t2 - t1), threads, frames_warm / (t1 - t0), loops);
    else
        fprintf(stderr, "frames=%d seconds=%.6f fps=%.3f threads=%d\n", nframes, t2 - t0, nframes / (t2 - t0), threads);
    if (raw) fclose(raw);
    av_frame_free(&fr);
    av_packet_free(&pkt);
    avcodec_free_context(&c);
    free(buf);
    return rc;
}
