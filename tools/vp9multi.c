/*
 * vp9multi.c -- ONE process decoding many VP9 streams, or the closed GOPs of one stream, on the
 * GPUs of a box (the path shards at stream / closed-GOP granularity,
 * no inter-GPU traffic).
 *
 *   streams : every input file is a job; job i runs on GPU i mod N with a decoder of its own.
 *   -gops   : the single input is cut at its key frames into closed GOPs (a key frame resets every
 *             reference slot and probability context, vp9.c:880-905, so a GOP decodes on its own);
 *             worker r owns GPU r mod N and decodes GOPs r, r + W, r + 2W, ... with one decoder,
 *             flushed between GOPs.  The per-frame MD5 lines are put back into stream order, so the
 *             output is byte-identical to a single decoder's (`ffvp9_ref in.ivf`).
 *
 * Container handling follows the reference: IVF (libavformat/ivfdec.c:29-100); superframe packets
 * are left to the decoder's own BSF (bsf/vp9_superframe_split.c:41-137); "key frame" is read from
 * the first frame header of the packet (frame_marker, profile, show_existing_frame, frame_type:
 * vp9.c:527-552).  The device of a decoder is chosen with $VP9CUDA_DEVICE at avcodec_open2() time
 * (vp9cuda_glue.c: ff_vp9cuda_init), which is why the decoders are opened one after the other by
 * the main thread.  Linked against the reference objects (oracle/_ref/ffvp9_multi_ref) the same
 * scheduler runs on the CPU, which is how the CPU tests check it.
 *
 * usage: ffvp9_multi [-gpus N] [-workers W] [-threads T] [-thread_type frame|slice] [-gops] [-null]
 *                    in.ivf [more.ivf ...]
 *   stdout: streams mode "== <file>" then its framemd5 lines, per input in order; -gops: the lines of the
 *   whole stream.  stderr: "jobs=.. frames=.. seconds=.. fps=.." (wall clock around all jobs).
 */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "libavcodec/avcodec.h"
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

typedef struct Stream {
    const char *name;
    unsigned char *buf;
    long size;
    long *pkt_pos;              /* offset of every packet's 12-byte IVF frame header */
    int n_pkt;
} Stream;

typedef struct Job {            /* a run of packets that starts with a key frame */
    const Stream *st;
    int first, last;            /* packets [first, last) */
    char *out;                  /* framemd5 lines, frame numbers relative to the job */
    size_t out_len, out_cap;
    int n_frames, failed;
} Job;

typedef struct Worker {
    AVCodecContext *c;
    Job **jobs;
    int n_jobs;
    int do_hash;
    pthread_t th;
} Worker;

static int load_stream(Stream *s, const char *path)
{
    FILE *fp = fopen(path, "rb");
    if (!fp) { perror(path); return -1; }
    fseek(fp, 0, SEEK_END);
    s->size = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    s->buf = malloc(s->size + AV_INPUT_BUFFER_PADDING_SIZE);
    if (!s->buf || fread(s->buf, 1, s->size, fp) != (size_t)s->size) return -1;
    memset(s->buf + s->size, 0, AV_INPUT_BUFFER_PADDING_SIZE);
    fclose(fp);
    if (s->size < 32 || memcmp(s->buf, "DKIF", 4)) { fprintf(stderr, "%s: not IVF\n", path); return -1; }
    s->name = path;
    int cap = 64;
    s->pkt_pos = malloc(cap * sizeof(long));
    for (long pos = 32; pos + 12 <= s->size;) {
        unsigned fsz = rl32(s->buf + pos);
        if (pos + 12 + fsz > s->size) { fprintf(stderr, "%s: truncated\n", path); return -1; }
        if (s->n_pkt == cap) s->pkt_pos = realloc(s->pkt_pos, (cap *= 2) * sizeof(long));
        s->pkt_pos[s->n_pkt++] = pos;
        pos += 12 + fsz;
    }
    return 0;
}

/* does the packet start with a key frame?  vp9.c:527-552 */
static int packet_is_key(const unsigned char *p, unsigned n)
{
    if (n < 1 || (p[0] >> 6) != 2)
        return 0;
    int bit = 2, profile = (p[0] >> 5 & 1) | (p[0] >> 4 & 1) << 1;
    bit += 2;
    if (profile == 3) bit++;                        /* reserved bit */
#define BIT(k) (p[(k) >> 3] >> (7 - ((k) & 7)) & 1)
    if (BIT(bit)) return 0;                         /* show_existing_frame */
    return !BIT(bit + 1);                           /* frame_type: 0 = key frame */
#undef BIT
}

static void job_printf(Job *j, const char *line)
{
    size_t n = strlen(line);
    if (j->out_len + n + 1 > j->out_cap) {
        j->out_cap = (j->out_cap + n + 1) * 2;
        j->out = realloc(j->out, j->out_cap);
    }
    memcpy(j->out + j->out_len, line, n + 1);
    j->out_len += n;
}

static void emit(Job *j, AVFrame *f, int do_hash)
{
    char line[160];
    const AVPixFmtDescriptor *d = av_pix_fmt_desc_get(f->format);
    if (do_hash) {
        int bps = d->comp[0].depth > 8 ? 2 : 1;
        struct AVMD5 *md5 = av_md5_alloc();
        unsigned char dig[16];
        av_md5_init(md5);
        for (int p = 0; p < 3; p++) {
            int w = p ? AV_CEIL_RSHIFT(f->width, d->log2_chroma_w) : f->width;
            int h = p ? AV_CEIL_RSHIFT(f->height, d->log2_chroma_h) : f->height;
            for (int y = 0; y < h; y++)
                av_md5_update(md5, f->data[p] + (ptrdiff_t)y * f->linesize[p], w * bps);
        }
        av_md5_final(md5, dig);
        av_free(md5);
        int n = snprintf(line, sizeof(line), "%d %d %s ", f->width, f->height, d->name);
        for (int i = 0; i < 16; i++) n += snprintf(line + n, sizeof(line) - n, "%02x", dig[i]);
        snprintf(line + n, sizeof(line) - n, "\n");
        job_printf(j, line);
    }
    j->n_frames++;
}

static void run_job(Worker *w, Job *j)
{
    AVPacket *pkt = av_packet_alloc();
    AVFrame *fr = av_frame_alloc();
    for (int k = j->first; k < j->last; k++) {
        const unsigned char *h = j->st->buf + j->st->pkt_pos[k];
        unsigned fsz = rl32(h);
        if (av_new_packet(pkt, fsz) < 0) { j->failed = 1; break; }
        memcpy(pkt->data, h + 12, fsz);
        pkt->pts = pkt->dts = k;
        if (avcodec_send_packet(w->c, pkt) < 0) j->failed = 1;
        av_packet_unref(pkt);
        while (avcodec_receive_frame(w->c, fr) >= 0) { emit(j, fr, w->do_hash); av_frame_unref(fr); }
    }
    /* drain, then make the decoder ready for the next job (a key frame follows) */
    avcodec_send_packet(w->c, NULL);
    for (int guard = 0; guard < 1000; guard++) {
        int r = avcodec_receive_frame(w->c, fr);
        if (r >= 0) { emit(j, fr, w->do_hash); av_frame_unref(fr); continue; }
        if (r != AVERROR_EOF && r != AVERROR(EAGAIN)) j->failed = 1;
        break;
    }
    avcodec_flush_buffers(w->c);
    av_frame_free(&fr);
    av_packet_free(&pkt);
}

static void *worker_main(void *arg)
{
    Worker *w = arg;
    for (int i = 0; i < w->n_jobs; i++)
        run_job(w, w->jobs[i]);
    return NULL;
}

int main(int argc, char **argv)
{
    int gpus = 1, workers = 0, threads = 1, ttype = 0, gops = 0, do_hash = 1, n_in = 0;
    const char *in[256];
    for (int i = 1; i < argc; i++) {
        if (!strcmp(argv[i], "-gpus") && i + 1 < argc) gpus = atoi(argv[++i]);
        else if (!strcmp(argv[i], "-workers") && i + 1 < argc) workers = atoi(argv[++i]);
        else if (!strcmp(argv[i], "-threads") && i + 1 < argc) threads = atoi(argv[++i]);
        else if (!strcmp(argv[i], "-thread_type") && i + 1 < argc)
            ttype = !strcmp(argv[++i], "slice") ? FF_THREAD_SLICE : FF_THREAD_FRAME;
        else if (!strcmp(argv[i], "-gops")) gops = 1;
        else if (!strcmp(argv[i], "-null")) do_hash = 0;
        else if (n_in < 256) in[n_in++] = argv[i];
    }
    if (!n_in || gpus < 1 || (gops && n_in != 1)) {
        fprintf(stderr, "usage: %s [-gpus N] [-workers W] [-threads T] [-thread_type frame|slice] [-gops] [-null] in.ivf ...\n", argv[0]);
        return 2;
    }
    Stream *st = calloc(n_in, sizeof(*st));
    for (int i = 0; i < n_in; i++)
        if (load_stream(&st[i], in[i]) < 0) return 2;

    /* jobs */
    int n_jobs = 0, cap = 16;
    Job *jobs = calloc(cap, sizeof(*jobs));
    if (gops) {
        for (int k = 0; k < st[0].n_pkt; k++) {
            const unsigned char *h = st[0].buf + st[0].pkt_pos[k];
            if (k == 0 || packet_is_key(h + 12, rl32(h))) {
                if (k == 0 && !packet_is_key(h + 12, rl32(h))) { fprintf(stderr, "stream does not start with a key frame\n"); return 2; }
                if (n_jobs == cap) { jobs = realloc(jobs, (cap *= 2) * sizeof(*jobs)); memset(jobs + n_jobs, 0, (cap - n_jobs) * sizeof(*jobs)); }
                if (n_jobs) jobs[n_jobs - 1].last = k;
                jobs[n_jobs].st = &st[0]; jobs[n_jobs].first = k; jobs[n_jobs].last = st[0].n_pkt;
                n_jobs++;
            }
        }
    } else {
        for (int i = 0; i < n_in; i++) {
            if (n_jobs == cap) { jobs = realloc(jobs, (cap *= 2) * sizeof(*jobs)); memset(jobs + n_jobs, 0, (cap - n_jobs) * sizeof(*jobs)); }
            jobs[n_jobs].st = &st[i]; jobs[n_jobs].first = 0; jobs[n_jobs].last = st[i].n_pkt;
            n_jobs++;
        }
    }
    if (workers < 1) workers = gops ? gpus : n_jobs;
    if (workers > n_jobs) workers = n_jobs;

    /* one decoder per worker, worker r on GPU r mod N; job k goes to worker k mod W */
    const AVCodec *codec = avcodec_find_decoder(AV_CODEC_ID_VP9);
    Worker *wk = calloc(workers, sizeof(*wk));
    for (int r = 0; r < workers; r++) {
        char dev[16];
        snprintf(dev, sizeof(dev), "%d", r % gpus);
        setenv("VP9CUDA_DEVICE", dev, 1);
        wk[r].c = avcodec_alloc_context3(codec);
        wk[r].c->thread_count = threads;
        if (ttype) wk[r].c->thread_type = ttype;
        wk[r].c->flags |= AV_CODEC_FLAG_BITEXACT;
        wk[r].c->err_recognition |= AV_EF_EXPLODE;
        if (avcodec_open2(wk[r].c, codec, NULL) < 0) { fprintf(stderr, "open failed (worker %d)\n", r); return 2; }
        wk[r].do_hash = do_hash;
        wk[r].jobs = calloc(n_jobs, sizeof(Job *));
        for (int k = r; k < n_jobs; k += workers)
            wk[r].jobs[wk[r].n_jobs++] = &jobs[k];
    }
    double t0 = now_s();
    for (int r = 0; r < workers; r++)
        pthread_create(&wk[r].th, NULL, worker_main, &wk[r]);
    for (int r = 0; r < workers; r++)
        pthread_join(wk[r].th, NULL);
    double t1 = now_s();

    int frames = 0, failed = 0, idx = 0;
    for (int k = 0; k < n_jobs; k++) {
        if (!gops) { printf("== %s\n", jobs[k].st->name); idx = 0; }
        /* prefix the running frame number, as the single-stream tool prints it */
        for (char *p = jobs[k].out; p && *p;) {
            char *e = strchr(p, '\n');
            printf("%d %.*s\n", idx++, (int)(e - p), p);
            p = e + 1;
        }
        frames += jobs[k].n_frames;
        failed |= jobs[k].failed;
    }
    fprintf(stderr, "jobs=%d workers=%d gpus=%d frames=%d seconds=%.6f fps=%.3f\n", n_jobs, workers, gpus, frames, t1 - t0,
            frames / (t1 - t0));
    for (int r = 0; r < workers; r++)
        avcodec_free_context(&wk[r].c);
    return failed;
}
