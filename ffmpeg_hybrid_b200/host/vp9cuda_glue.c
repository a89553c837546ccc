/*
 * vp9cuda_glue.c -- libavcodec-side record emitters for the vp9cuda back end.
 *
 * Compiled inside libavcodec (it needs VP9Context / VP9TileData).  It owns no pixel math:
 * ff_vp9cuda_emit_block() walks one coded block exactly the way intra_recon()/inter_recon()
 * do (reference vp9recon.c:235-364, 655-764) but writes records + moves the dequantised
 * coefficients into the pinned lane ring instead of calling VP9DSPContext slots.
 *
 * Two sinks sit behind the same function table (include/vp9cuda.h):
 *   - libvp9cuda.so, dlopen()ed from $VP9CUDA_LIB (default "libvp9cuda.so"): the product path;
 *   - a record dump to $VP9CUDA_DUMP (.v9r): no reconstruction at all, used on GPU-less
 *     machines to cut replay fixtures for the tests.  Decoded pictures are then undefined.
 * There is no software reconstruction in either.
 */
#include <dlfcn.h>
#include <errno.h>
#include <pthread.h>
#include <stdatomic.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "libavutil/buffer.h"
#include "libavutil/frame.h"
#include "libavutil/imgutils.h"
#include "libavutil/log.h"
#include "libavutil/mem.h"

#include "avcodec.h"
#include "packet.h"
#include "vp9data.h"
#include "vp9dec.h"
#include "vp9cuda_glue.h"
#include "vp9cuda.h"

typedef struct SinkApi {
    int  (*create)(Vp9CudaCtx **, int, int, int, int, int);
    void (*destroy)(Vp9CudaCtx *);
    int  (*flush)(Vp9CudaCtx *);
    int  (*begin_frame)(Vp9CudaCtx *, const Vp9CudaFrameHeader *, Vp9CudaFrameRing *);
    int  (*tile_ring)(Vp9CudaCtx *, uint32_t, int, Vp9CudaLaneRing *);
    int  (*submit_tile)(Vp9CudaCtx *, uint32_t, int, const Vp9CudaLaneCounts *);
    int  (*finish_frame)(Vp9CudaCtx *, uint32_t, uint8_t *const[3], const ptrdiff_t[3], uint64_t *);
    int  (*abort_frame)(Vp9CudaCtx *, uint32_t);
    int  (*wait)(Vp9CudaCtx *, uint64_t);
    void *(*host_alloc)(size_t);             /* pinned host memory (NULL in the dump sink) */
    void (*host_free)(void *);
} SinkApi;

/* One per decoder instance.  Under frame threading (pthread_frame.c) every worker has its own
 * VP9Context and its own VP9CudaGlue, and all of them share this: the back end's context (one GPU
 * stream set, one reference pool) and the pool of pinned output pictures. */
typedef struct VP9CudaShared {
    SinkApi api;
    void *dl;
    Vp9CudaCtx *ctx;
    atomic_int refs;
    pthread_mutex_t pool_mu;
    AVBufferPool *frame_pool;    /* output pictures in pinned memory: the copy-out is a real asynchronous DMA */
    size_t frame_pool_size;
    int pinned_frames;
} VP9CudaShared;

struct VP9CudaGlue {
    VP9CudaShared *sh;
    int frame_mt;                /* libavcodec frame threading: one frame per worker, several in flight */
    int device;                  /* GPU of this decoder: $VP9CUDA_DEVICE as it was when the decoder was opened, so that
                                    one process can open decoders on several GPUs (tools/vp9multi.c) */
    Vp9CudaFrameHeader hdr;
    Vp9CudaFrameRing fring;
    Vp9CudaLaneRing ring[VP9CUDA_MAX_LANES];
    Vp9CudaLaneCounts cnt[VP9CUDA_MAX_LANES];
    int n_lanes;
    int in_frame;
    atomic_int overflow;
    unsigned sb_cols;
    /* intra dependency levels: per plane one (tag << 16 | level) word per 4x4 cell; a word whose
     * tag is not the current frame's means "no intra block of this frame here" */
    uint32_t *lvl[3];
    int lvl_w[3], lvl_h[3];
    size_t lvl_cap[3];
    uint32_t lvl_tag;
    /* without frame threading one frame is kept in flight: the picture handed out by the NEXT
     * ff_vp9cuda_deliver call */
    int pipeline;
    AVFrame *pending;
    int pending_valid;
    uint64_t pending_ticket, last_ticket;
};

/* ------------------------------------------------------------------------------------------
 * dump sink: same calls, written to a file (format read by ffmpeg_hybrid_b200/v9r.py and
 * oracle/replay_model.cc).  One frame at a time: no frame threading.
 * ---------------------------------------------------------------------------------------- */
enum { CAP_BLK = 64, CAP_TX = 384, CAP_ITX = 384, CAP_COEF = 6144 };

typedef struct DumpSet {
    Vp9CudaFrameHeader hdr;
    Vp9CudaLaneRing ring[VP9CUDA_MAX_LANES];
    Vp9CudaLaneCounts cnt[VP9CUDA_MAX_LANES];
    unsigned lane_sbs[VP9CUDA_MAX_LANES];
    Vp9CudaSbIndex *sb_index;
    Vp9CudaLoopFilterSb *lf;
    unsigned n_sb, n_sb_cap;
    uint32_t frame_id;
    uint64_t seq;
    int busy;
} DumpSet;

/* like the back end: several frames may be open (frame threading), they are written to the file
 * in begin_frame order */
typedef struct DumpCtx {
    FILE *fp;                    /* NULL: count only ($VP9CUDA_DUMP=/dev/null, the entropy-only decoder) */
    pthread_mutex_t mu;
    pthread_cond_t cv;
    DumpSet *set;
    int n_sets;
    uint32_t next_id;
    uint64_t next_begin, next_finish;
    uint64_t aborted[64];
    int n_aborted;
} DumpCtx;

static int dump_create(Vp9CudaCtx **out, int dev, int w, int h, int bpp, int nfl)
{
    DumpCtx *d = av_mallocz(sizeof(*d));
    const char *path = getenv("VP9CUDA_DUMP");
    if (!d)
        return -ENOMEM;
    d->n_sets = FFMAX(nfl, 1);
    d->set = av_calloc(d->n_sets, sizeof(*d->set));
    if (!d->set) {
        av_free(d);
        return -ENOMEM;
    }
    if (strcmp(path, "/dev/null")) {
        d->fp = fopen(path, "wb");
        if (!d->fp) {
            av_free(d->set);
            av_free(d);
            return -EIO;
        }
        fwrite("V9R3", 1, 4, d->fp);
        {
            uint32_t sz[8] = { sizeof(Vp9CudaFrameHeader), sizeof(Vp9CudaLaneCounts),
                               sizeof(Vp9CudaInterBlock), sizeof(Vp9CudaTx), sizeof(Vp9CudaIntraTx),
                               sizeof(Vp9CudaSbIndex), sizeof(Vp9CudaLoopFilterSb), VP9CUDA_ABI_VERSION };
            fwrite(sz, 4, 8, d->fp);
        }
    }
    pthread_mutex_init(&d->mu, NULL);
    pthread_cond_init(&d->cv, NULL);
    d->next_id = 1;
    *out = (Vp9CudaCtx *)d;
    return 0;
}

static void dump_destroy(Vp9CudaCtx *c)
{
    DumpCtx *d = (DumpCtx *)c;
    for (int i = 0; i < d->n_sets; i++) {
        DumpSet *q = &d->set[i];
        for (int l = 0; l < VP9CUDA_MAX_LANES; l++) {
            av_freep(&q->ring[l].blk);
            av_freep(&q->ring[l].tx);
            av_freep(&q->ring[l].itx);
            av_freep(&q->ring[l].coef);
        }
        av_freep(&q->sb_index);
        av_freep(&q->lf);
    }
    av_free(d->set);
    if (d->fp)
        fclose(d->fp);
    pthread_mutex_destroy(&d->mu);
    pthread_cond_destroy(&d->cv);
    av_free(d);
}

static DumpSet *dump_set_of(DumpCtx *d, uint32_t id)
{
    for (int i = 0; i < d->n_sets; i++)
        if (d->set[i].busy && d->set[i].frame_id == id)
            return &d->set[i];
    return NULL;
}

static void dump_advance(DumpCtx *d)      /* d->mu held */
{
    d->next_finish++;
    for (int i = 0; i < d->n_aborted;) {
        if (d->aborted[i] == d->next_finish) {
            d->aborted[i] = d->aborted[--d->n_aborted];
            d->next_finish++;
            i = 0;
        } else {
            i++;
        }
    }
    pthread_cond_broadcast(&d->cv);
}

static void dump_release(DumpCtx *d, DumpSet *q)      /* d->mu held; the frame is not written */
{
    q->busy = 0;
    if (q->seq == d->next_finish)
        dump_advance(d);
    else if (d->n_aborted < 64)
        d->aborted[d->n_aborted++] = q->seq;
    pthread_cond_broadcast(&d->cv);
}

static int dump_flush(Vp9CudaCtx *c)
{
    DumpCtx *d = (DumpCtx *)c;
    pthread_mutex_lock(&d->mu);
    for (int i = 0; i < d->n_sets; i++)
        if (d->set[i].busy)
            dump_release(d, &d->set[i]);
    pthread_mutex_unlock(&d->mu);
    return 0;
}

static int dump_begin_frame(Vp9CudaCtx *c, const Vp9CudaFrameHeader *h, Vp9CudaFrameRing *fr)
{
    DumpCtx *d = (DumpCtx *)c;
    DumpSet *q = NULL;
    unsigned n_sb = ((h->w + 63) >> 6) * ((h->h + 63) >> 6);
    uint32_t sb0 = 0;
    /* per-superblock capacities as in the library (csrc/vp9cuda.cu: cap_scale): 4:4:4 carries twice the samples */
    const unsigned k = (h->ss_h && h->ss_v) ? 1 : 2;

    pthread_mutex_lock(&d->mu);
    while (!q) {
        for (int i = 0; i < d->n_sets && !q; i++)
            if (!d->set[i].busy)
                q = &d->set[i];
        if (!q)
            pthread_cond_wait(&d->cv, &d->mu);
    }
    q->busy = 1;
    q->frame_id = d->next_id++;
    q->seq = d->next_begin++;
    pthread_mutex_unlock(&d->mu);

    q->hdr = *h;
    if (n_sb > q->n_sb_cap) {
        av_freep(&q->sb_index);
        av_freep(&q->lf);
        q->sb_index = av_malloc_array(n_sb, sizeof(*q->sb_index));
        q->lf = av_malloc_array(n_sb, sizeof(*q->lf));
        q->n_sb_cap = n_sb;
    }
    memset(q->sb_index, 0, n_sb * sizeof(*q->sb_index));
    memset(q->lf, 0, n_sb * sizeof(*q->lf));
    q->n_sb = n_sb;
    for (int l = 0; l < h->n_lanes; l++) {
        unsigned sbs = h->lane_sbs[l];
        Vp9CudaLaneRing *r = &q->ring[l];
        if (sbs * k > q->lane_sbs[l]) {
            av_freep(&r->blk);
            av_freep(&r->tx);
            av_freep(&r->itx);
            av_freep(&r->coef);
            r->blk = av_malloc_array(sbs * CAP_BLK, sizeof(*r->blk));
            r->tx = av_malloc_array(sbs * k * CAP_TX, sizeof(*r->tx));
            r->itx = av_malloc_array(sbs * k * CAP_ITX, sizeof(*r->itx));
            r->coef = av_malloc_array(sbs * k * CAP_COEF, sizeof(Vp9CudaCoef32));
            q->lane_sbs[l] = sbs * k;
        }
        r->blk_cap = sbs * CAP_BLK;
        r->tx_cap = sbs * k * CAP_TX;
        r->itx_cap = sbs * k * CAP_ITX;
        r->coef_cap = sbs * k * CAP_COEF;
        r->coef_base = sb0 * k * CAP_COEF;
        sb0 += sbs;
        memset(&q->cnt[l], 0, sizeof(q->cnt[l]));
    }
    fr->sb_index = q->sb_index;
    fr->lf = q->lf;
    fr->n_sb = n_sb;
    fr->frame_id = q->frame_id;
    return 0;
}

static int dump_tile_ring(Vp9CudaCtx *c, uint32_t id, int lane, Vp9CudaLaneRing *r)
{
    DumpCtx *d = (DumpCtx *)c;
    DumpSet *q;
    pthread_mutex_lock(&d->mu);
    q = dump_set_of(d, id);
    pthread_mutex_unlock(&d->mu);
    if (!q)
        return -EINVAL;
    *r = q->ring[lane];
    return 0;
}

static int dump_submit_tile(Vp9CudaCtx *c, uint32_t id, int lane, const Vp9CudaLaneCounts *n)
{
    DumpCtx *d = (DumpCtx *)c;
    DumpSet *q;
    pthread_mutex_lock(&d->mu);
    q = dump_set_of(d, id);
    pthread_mutex_unlock(&d->mu);
    if (!q)
        return -EINVAL;
    q->cnt[lane] = *n;
    return 0;
}

static int dump_finish_frame(Vp9CudaCtx *c, uint32_t id, uint8_t *const out[3],
                             const ptrdiff_t st[3], uint64_t *ticket)
{
    DumpCtx *d = (DumpCtx *)c;
    DumpSet *q;
    int ceb;
    pthread_mutex_lock(&d->mu);
    q = dump_set_of(d, id);
    if (!q) {
        pthread_mutex_unlock(&d->mu);
        return -EINVAL;
    }
    while (d->next_finish != q->seq)
        pthread_cond_wait(&d->cv, &d->mu);
    pthread_mutex_unlock(&d->mu);
    ceb = q->hdr.bpp > 8 ? sizeof(Vp9CudaCoef32) : sizeof(Vp9CudaCoef16);
    if (d->fp) {
        fwrite("FRAM", 1, 4, d->fp);
        fwrite(&q->hdr, sizeof(q->hdr), 1, d->fp);
        fwrite(&q->n_sb, 4, 1, d->fp);
        for (int l = 0; l < q->hdr.n_lanes; l++) {
            const Vp9CudaLaneRing *r = &q->ring[l];
            const Vp9CudaLaneCounts *n = &q->cnt[l];
            fwrite(n, sizeof(*n), 1, d->fp);
            fwrite(&r->coef_base, 4, 1, d->fp);
            fwrite(r->blk, sizeof(*r->blk), n->blk, d->fp);
            fwrite(r->tx, sizeof(*r->tx), n->tx, d->fp);
            fwrite(r->itx, sizeof(*r->itx), n->itx, d->fp);
            fwrite(r->coef, ceb, n->coef, d->fp);
        }
        fwrite(q->sb_index, sizeof(*q->sb_index), q->n_sb, d->fp);
        fwrite(q->lf, sizeof(*q->lf), q->n_sb, d->fp);
    }
    *ticket = id;
    pthread_mutex_lock(&d->mu);
    q->busy = 0;
    dump_advance(d);
    pthread_mutex_unlock(&d->mu);
    return 0;
}

static int dump_abort_frame(Vp9CudaCtx *c, uint32_t id)
{
    DumpCtx *d = (DumpCtx *)c;
    DumpSet *q;
    pthread_mutex_lock(&d->mu);
    q = dump_set_of(d, id);
    if (q)
        dump_release(d, q);
    pthread_mutex_unlock(&d->mu);
    return q ? 0 : -EINVAL;
}

static int dump_wait(Vp9CudaCtx *c, uint64_t t) { return 0; }
/* output pictures come from the same size-stable pool as with the back end (plain memory here), so
 * that the entropy-only decoder differs from the real one in the missing reconstruction only */
static void *dump_host_alloc(size_t n) { return av_malloc(n); }
static void dump_host_free(void *p) { av_free(p); }

/* ------------------------------------------------------------------------------------------ */

static VP9CudaGlue *glue_of(const VP9Context *s) { return s->cuda; }

/* ------------------------------------------------------------------------------------------
 * get_buffer2 in pinned memory.  The back end copies every finished picture into the decoder's
 * AVFrame; into pageable memory that copy blocks the calling thread, into pinned memory it is an
 * asynchronous DMA that runs beside the next frame's entropy decode.  Installed only when the
 * application left the default allocator in place.  Buffers may outlive the decoder, so the
 * free callback reaches the library through a process-wide pointer (the library stays loaded).
 * ---------------------------------------------------------------------------------------- */
static void (*pinned_host_free)(void *);

static void pinned_buffer_free(void *opaque, uint8_t *data)
{
    pinned_host_free(data);
}

static AVBufferRef *pinned_buffer_alloc(void *opaque, size_t size)
{
    VP9CudaShared *sh = opaque;
    AVBufferRef *ref;
    uint8_t *p = sh->api.host_alloc(size);
    if (!p)
        return NULL;
    ref = av_buffer_create(p, size, pinned_buffer_free, NULL, 0);
    if (!ref)
        sh->api.host_free(p);
    return ref;
}

static int vp9cuda_get_buffer2(AVCodecContext *avctx, AVFrame *frame, int flags)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g = s->cuda;
    VP9CudaShared *sh = g ? g->sh : NULL;
    int w = frame->width, h = frame->height, ret, i;
    int align[AV_NUM_DATA_POINTERS], linesize[4];
    ptrdiff_t ls[4];
    size_t plane[4], total = 0, off = 0;

    if (!sh || !sh->pinned_frames)
        return avcodec_default_get_buffer2(avctx, frame, flags);
    avcodec_align_dimensions2(avctx, &w, &h, align);
    if ((ret = av_image_fill_linesizes(linesize, frame->format, w)) < 0)
        return ret;
    for (i = 0; i < 4; i++)
        ls[i] = linesize[i] = FFALIGN(linesize[i], 64);
    if ((ret = av_image_fill_plane_sizes(plane, frame->format, h, ls)) < 0)
        return ret;
    for (i = 0; i < 4; i++)
        total += FFALIGN(plane[i], 64);
    total += 64;
    pthread_mutex_lock(&sh->pool_mu);
    if (!sh->frame_pool || sh->frame_pool_size < total) {
        /* one pool, sized for the largest picture seen (streams with resize frames alternate) */
        av_buffer_pool_uninit(&sh->frame_pool);
        sh->frame_pool = av_buffer_pool_init2(total, sh, pinned_buffer_alloc, NULL);
        sh->frame_pool_size = sh->frame_pool ? total : 0;
    }
    frame->buf[0] = sh->frame_pool ? av_buffer_pool_get(sh->frame_pool) : NULL;
    pthread_mutex_unlock(&sh->pool_mu);
    if (!frame->buf[0])
        return AVERROR(ENOMEM);
    for (i = 0; i < 4; i++) {
        frame->linesize[i] = linesize[i];
        frame->data[i] = plane[i] ? frame->buf[0]->data + off : NULL;
        off += FFALIGN(plane[i], 64);
    }
    frame->extended_data = frame->data;
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * the shared half: library handle + back-end context, created when the first frame header is
 * known (the context is sized for that frame; the first frame of a stream is a key frame)
 * ---------------------------------------------------------------------------------------- */
static void shared_unref(VP9CudaShared **psh)
{
    VP9CudaShared *sh = *psh;
    *psh = NULL;
    if (!sh || atomic_fetch_sub(&sh->refs, 1) != 1)
        return;
    av_buffer_pool_uninit(&sh->frame_pool);
    if (sh->ctx)
        sh->api.destroy(sh->ctx);
    if (sh->dl && !pinned_host_free)    /* pinned pictures handed to the application may still be alive */
        dlclose(sh->dl);
    pthread_mutex_destroy(&sh->pool_mu);
    av_free(sh);
}

static int shared_create(AVCodecContext *avctx, VP9CudaGlue *g)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaShared *sh = av_mallocz(sizeof(*sh));
    int ret, dev = 0, nfl;

    if (!sh)
        return AVERROR(ENOMEM);
    atomic_init(&sh->refs, 1);
    pthread_mutex_init(&sh->pool_mu, NULL);
    if (getenv("VP9CUDA_DUMP")) {
        sh->api = (SinkApi){ dump_create, dump_destroy, dump_flush, dump_begin_frame, dump_tile_ring,
                             dump_submit_tile, dump_finish_frame, dump_abort_frame, dump_wait,
                             dump_host_alloc, dump_host_free };
    } else {
        const char *lib = getenv("VP9CUDA_LIB");
        sh->dl = dlopen(lib ? lib : "libvp9cuda.so", RTLD_NOW | RTLD_LOCAL);
        if (!sh->dl) {
            av_log(avctx, AV_LOG_ERROR, "vp9cuda: cannot load back end: %s\n", dlerror());
            shared_unref(&sh);
            return AVERROR(ENODEV);
        }
#define SYM(f) do { *(void **)&sh->api.f = dlsym(sh->dl, "vp9cuda_" #f); \
                    if (!sh->api.f) { av_log(avctx, AV_LOG_ERROR, "vp9cuda: missing symbol %s\n", #f); \
                                      shared_unref(&sh); return AVERROR(ENODEV); } } while (0)
        SYM(create); SYM(destroy); SYM(flush); SYM(begin_frame); SYM(tile_ring); SYM(submit_tile);
        SYM(finish_frame); SYM(abort_frame); SYM(wait); SYM(host_alloc); SYM(host_free);
#undef SYM
    }
    dev = g->device;
    /* ring sets: one per frame a worker may hold between begin_frame and finish_frame, plus one */
    nfl = g->frame_mt ? FFMIN(avctx->thread_count + 1, 32) : 2;
    ret = sh->api.create(&sh->ctx, dev, s->w, s->h, s->s.h.bpp, nfl);
    if (ret < 0) {
        av_log(avctx, AV_LOG_ERROR, "vp9cuda: back end refused to start (%d)\n", ret);
        sh->ctx = NULL;
        shared_unref(&sh);
        return AVERROR(ENODEV);
    }
    sh->pinned_frames = sh->api.host_alloc &&
                        (!getenv("VP9CUDA_PINNED_FRAMES") || atoi(getenv("VP9CUDA_PINNED_FRAMES")));
    if (sh->pinned_frames)
        pinned_host_free = sh->api.host_free;
    g->sh = sh;
    return 0;
}

int ff_vp9cuda_init(AVCodecContext *avctx)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g;

    if (getenv("VP9CUDA_DISABLE"))       /* the overlay then is the plain software decoder (A/B timing) */
        return 0;
    g = av_mallocz(sizeof(*g));
    if (!g)
        return AVERROR(ENOMEM);
    g->frame_mt = !!(avctx->active_thread_type & FF_THREAD_FRAME);
    g->device = getenv("VP9CUDA_DEVICE") ? atoi(getenv("VP9CUDA_DEVICE")) : 0;
    atomic_init(&g->overflow, 0);
    /* one worker per frame already keeps many frames in flight; otherwise hold one picture back */
    g->pipeline = !g->frame_mt && (!getenv("VP9CUDA_PIPELINE") || atoi(getenv("VP9CUDA_PIPELINE")));
    g->pending = av_frame_alloc();
    if (!g->pending) {
        av_free(g);
        return AVERROR(ENOMEM);
    }
    s->cuda = g;
    if (!g->frame_mt) {
        /* fail at open time when the library or the GPU is missing: there is no software fallback.
         * (Frame-threaded workers share one back end; the first frame creates it.) */
        void *dl;
        if (getenv("VP9CUDA_DUMP"))
            return 0;
        dl = dlopen(getenv("VP9CUDA_LIB") ? getenv("VP9CUDA_LIB") : "libvp9cuda.so", RTLD_NOW | RTLD_LOCAL);
        if (!dl) {
            av_log(avctx, AV_LOG_ERROR, "vp9cuda: cannot load back end: %s\n", dlerror());
            return AVERROR(ENODEV);
        }
        dlclose(dl);
    }
    return 0;
}

/* vp9_decode_free, first thing: no copy-out may still be writing into a picture that is about to
 * be unreferenced */
void ff_vp9cuda_quiesce(AVCodecContext *avctx)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g = s->cuda;
    if (!g || !g->sh)
        return;
    if (g->in_frame) {
        g->sh->api.abort_frame(g->sh->ctx, g->fring.frame_id);
        g->in_frame = 0;
    }
    if (g->last_ticket)
        g->sh->api.wait(g->sh->ctx, g->last_ticket);
    if (g->pending_valid)
        av_frame_unref(g->pending);
    g->pending_valid = 0;
}

void ff_vp9cuda_uninit(AVCodecContext *avctx)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g = s->cuda;
    if (!g)
        return;
    ff_vp9cuda_quiesce(avctx);
    av_frame_free(&g->pending);
    shared_unref(&g->sh);
    for (int p = 0; p < 3; p++)
        av_freep(&g->lvl[p]);
    av_free(g);
    s->cuda = NULL;
}

/* vp9_decode_update_thread_context (vp9.c:1923-1962): the next worker decodes the next frame on
 * the same back end */
int ff_vp9cuda_update_thread_context(AVCodecContext *dst, const AVCodecContext *src)
{
    VP9Context *s = dst->priv_data;
    const VP9Context *ssrc = src->priv_data;
    if (!s->cuda || !ssrc->cuda || s->cuda->sh == ssrc->cuda->sh || !ssrc->cuda->sh)
        return 0;
    shared_unref(&s->cuda->sh);
    atomic_fetch_add(&ssrc->cuda->sh->refs, 1);
    s->cuda->sh = ssrc->cuda->sh;
    return 0;
}

void ff_vp9cuda_flush(AVCodecContext *avctx)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g = s->cuda;
    if (!g)
        return;
    g->in_frame = 0;                     /* the back end's flush aborts every open frame */
    if (g->sh)
        g->sh->api.flush(g->sh->ctx);
    if (g->pending_valid)
        av_frame_unref(g->pending);
    g->pending_valid = 0;
    g->last_ticket = 0;
}

/* vp9_decode_frame, after decode_frame_header and before vp9_frame_alloc: make sure the back end
 * exists (its buffers are sized from this first header) and that CUR_FRAME comes from the pinned
 * pool.  Under frame threading update_context_from_user() re-installs the application's
 * get_buffer2 in the worker's context before every packet, hence per frame. */
int ff_vp9cuda_prepare_frame(AVCodecContext *avctx)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g = s->cuda;
    int ret;

    if (g->in_frame) {                   /* a frame this context began and never finished */
        g->sh->api.abort_frame(g->sh->ctx, g->fring.frame_id);
        g->in_frame = 0;
    }
    if (s->ss_h != s->ss_v) {
        av_log(avctx, AV_LOG_ERROR, "vp9cuda: 4:2:2 and 4:4:0 are not implemented (4:2:0 and 4:4:4 are)\n");
        return AVERROR_PATCHWELCOME;
    }
    if (!g->sh && (ret = shared_create(avctx, g)) < 0)
        return ret;
    if (g->sh->pinned_frames && avctx->get_buffer2 == avcodec_default_get_buffer2)
        avctx->get_buffer2 = vp9cuda_get_buffer2;
    return 0;
}

int ff_vp9cuda_begin_frame(AVCodecContext *avctx)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g = s->cuda;
    VP9CudaShared *sh = g->sh;
    Vp9CudaFrameHeader *h = &g->hdr;
    int ret;

    memset(h, 0, sizeof(*h));
    h->w = s->w;
    h->h = s->h;
    h->bpp = s->s.h.bpp;
    h->ss_h = s->ss_h;
    h->ss_v = s->ss_v;
    h->keyframe = s->s.h.keyframe;
    h->intraonly = s->s.h.intraonly;
    h->lossless = s->s.h.lossless;
    h->invisible = s->s.h.invisible;
    h->lf_level = s->s.h.filter.level;
    h->refreshrefmask = s->s.h.refreshrefmask;
    /* key and intra-only frames predict from nothing: refidx / mvscale are leftovers there */
    for (int i = 0; i < 3 && !s->s.h.keyframe && !s->s.h.intraonly; i++) {
        h->ref_slot[i] = s->s.h.refidx[i];
        h->mvscale[i][0] = s->mvscale[i][0];
        h->mvscale[i][1] = s->mvscale[i][1];
        /* vp9.c:857-872 leaves mvstep as it was for unscaled / unusable references */
        if (s->mvscale[i][0] && s->mvscale[i][0] != REF_INVALID_SCALE) {
            h->mvstep[i][0] = s->mvstep[i][0];
            h->mvstep[i][1] = s->mvstep[i][1];
        }
    }
    memcpy(h->lim_lut, s->filter_lut.lim_lut, 64);
    memcpy(h->mblim_lut, s->filter_lut.mblim_lut, 64);
    g->sb_cols = s->sb_cols;
    g->n_lanes = s->active_tile_cols;
    h->n_lanes = g->n_lanes;
    if (g->n_lanes == 1) {
        h->lane_sbs[0] = s->sb_cols * s->sb_rows;
    } else {
        for (int i = 0; i < g->n_lanes; i++) {
            unsigned a = (i * s->sb_cols) >> s->s.h.tiling.log2_tile_cols;
            unsigned b = ((i + 1) * s->sb_cols) >> s->s.h.tiling.log2_tile_cols;
            h->lane_sbs[i] = (FFMIN(b, s->sb_cols) - FFMIN(a, s->sb_cols)) * s->sb_rows;
        }
    }
    ret = sh->api.begin_frame(sh->ctx, h, &g->fring);
    if (ret == -EINVAL) {
        av_log(avctx, AV_LOG_ERROR, "vp9cuda: %dx%d %d-bit frames are not supported by the back end\n",
               s->w, s->h, s->s.h.bpp);
        return AVERROR_PATCHWELCOME;
    }
    if (ret < 0)
        return AVERROR_EXTERNAL;
    g->in_frame = 1;
    atomic_store(&g->overflow, 0);
    for (int i = 0; i < g->n_lanes; i++) {
        ret = sh->api.tile_ring(sh->ctx, g->fring.frame_id, i, &g->ring[i]);
        if (ret < 0)
            return AVERROR_EXTERNAL;
        memset(&g->cnt[i], 0, sizeof(g->cnt[i]));
    }
    for (int p = 0; p < 3; p++) {
        int w = (s->cols * 2) >> (p ? s->ss_h : 0), hh = (s->rows * 2) >> (p ? s->ss_v : 0);
        size_t n = (size_t)w * hh;
        if (n > g->lvl_cap[p]) {
            av_freep(&g->lvl[p]);
            g->lvl[p] = av_calloc(n, sizeof(uint32_t));
            if (!g->lvl[p])
                return AVERROR(ENOMEM);
            g->lvl_cap[p] = n;
        }
        g->lvl_w[p] = w;
        g->lvl_h[p] = hh;
    }
    if (++g->lvl_tag > 0xffff) {               /* tags wrapped: forget everything */
        for (int p = 0; p < 3; p++)
            memset(g->lvl[p], 0, g->lvl_cap[p] * sizeof(uint32_t));
        g->lvl_tag = 1;
    }
    return 0;
}

/* a packet failed between begin_frame and end_frame (vp9.c `fail:` and the early returns of the
 * tile set-up): give the ring set back, the frames behind it must not wait for this one */
void ff_vp9cuda_abort_frame(AVCodecContext *avctx)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g = s->cuda;
    if (g && g->in_frame) {
        g->sh->api.abort_frame(g->sh->ctx, g->fring.frame_id);
        g->in_frame = 0;
    }
}

/* Move one transform block's non-zero dequantised coefficients into the lane's coefficient
 * stream and re-zero the source, which is what every itxfm_add slot does
 * (vp9dsp_template.c:1168,1182).  decode_coeffs wrote coef[scan[i]] for i < eob only
 * (vp9block.c:816-920), so walking that prefix of the scan finds every non-zero value. */
static av_always_inline uint32_t take_coefs(VP9CudaGlue *g, int lane, int16_t *src, const int16_t *scan,
                                            int eob, int bytesperpixel, uint16_t *ncoef)
{
    Vp9CudaLaneRing *r = &g->ring[lane];
    Vp9CudaLaneCounts *c = &g->cnt[lane];
    uint32_t off = c->coef;
    int n = 0, i;
    if ((uint64_t)off + eob > r->coef_cap) {
        atomic_store(&g->overflow, 1);
        *ncoef = 0;
        return 0;
    }
    if (bytesperpixel == 1) {
        Vp9CudaCoef16 *o = (Vp9CudaCoef16 *)r->coef + off;
        for (i = 0; i < eob; i++) {
            int rc = scan[i], v = src[rc];
            o[n].val = v;
            o[n].pos = rc;
            n += v != 0;
            src[rc] = 0;
        }
    } else {
        Vp9CudaCoef32 *o = (Vp9CudaCoef32 *)r->coef + off;
        int32_t *blk = (int32_t *)src;
        for (i = 0; i < eob; i++) {
            int rc = scan[i], v = blk[rc];
            o[n].val = v;
            o[n].pos = rc;
            n += v != 0;
            blk[rc] = 0;
        }
    }
    c->coef = off + n;
    *ncoef = n;
    return r->coef_base + off;
}

static av_always_inline int lvl_get(const VP9CudaGlue *g, int p, int x, int y)
{
    uint32_t v = g->lvl[p][(size_t)y * g->lvl_w[p] + x];
    return (v >> 16) == g->lvl_tag ? (int)(v & 0xffff) : -1;
}

/* Edges the predictor of `mode` reads once check_intra_mode() has replaced the mode for the edges
 * that do not exist (mode_conv, vp9recon.c:49-70): bit 0 left column, bit 1 top row, bit 2 top-left
 * sample, bit 3 the top-right extension.  D135 / D117 / D153 keep their mode and read synthetic
 * values (127 / 129) in the place of a missing edge. */
static av_always_inline int intra_needs(int mode, int have_left, int have_top)
{
    switch (mode) {
    case VERT_PRED:
        return have_top ? 2 : 0;
    case DIAG_DOWN_LEFT_PRED:
    case VERT_LEFT_PRED:
        return have_top ? 2 | 8 : 0;
    case HOR_PRED:
    case HOR_UP_PRED:
        return have_left ? 1 : 0;
    case DC_PRED:
        return (have_left ? 1 : 0) | (have_top ? 2 : 0);
    case TM_VP8_PRED:
        return have_left ? (have_top ? 1 | 2 | 4 : 1) : (have_top ? 2 : 0);
    default:
        return (have_left ? 1 : 0) | (have_top ? 2 : 0) | (have_left && have_top ? 4 : 0);
    }
}

/* Dependency depth of one intra transform block (n x n cells at cell x,y of plane p) and
 * registration of its cells.  The neighbour set is what the block's predictor reads
 * (check_intra_mode, vp9recon.c:37-221, by mode: intra_needs): the left column, the top-left
 * cell, the top row, and one more cell to the right for 4x4 blocks with have_right whose mode
 * reads the top-right extension.  All of them precede the block in coding order and lie in the
 * same tile column, so tile threads never race. */
static av_always_inline int intra_level(VP9CudaGlue *g, int p, int x, int y, int n, int needs, int topright)
{
    const int w = g->lvl_w[p], h = g->lvl_h[p];
    int lv = 0, i, j, xe, ye;
    if (needs & 1) {
        ye = FFMIN(y + n, h);
        for (j = y; j < ye; j++)
            lv = FFMAX(lv, lvl_get(g, p, x - 1, j) + 1);
    }
    if (needs & 4)
        lv = FFMAX(lv, lvl_get(g, p, x - 1, y - 1) + 1);
    if (needs & 2) {
        xe = FFMIN(x + ((needs & 8) && topright ? 2 : n), w);
        for (i = x; i < xe; i++)
            lv = FFMAX(lv, lvl_get(g, p, i, y - 1) + 1);
    }
    xe = FFMIN(x + n, w);
    ye = FFMIN(y + n, h);
    for (j = y; j < ye; j++)
        for (i = x; i < xe; i++)
            g->lvl[p][(size_t)j * w + i] = g->lvl_tag << 16 | lv;
    return lv;
}

void ff_vp9cuda_emit_block(VP9TileData *td)
{
    const VP9Context *s = td->s;
    VP9CudaGlue *g = glue_of(s);
    const VP9Block *b = td->b;
    const int lane = td - s->td;
    Vp9CudaLaneRing *r = &g->ring[lane];
    Vp9CudaLaneCounts *c = &g->cnt[lane];
    Vp9CudaSbIndex *sbi = &g->fring.sb_index[(td->row >> 3) * g->sb_cols + (td->col >> 3)];
    const int bytesperpixel = s->bytesperpixel;
    const int row = td->row, col = td->col;
    const int lossless = s->s.h.lossless ? 0x80 : 0;
    const int16_t *const *yscans = ff_vp9_scans[4 * s->s.h.lossless + b->tx];     /* as decode_coeffs, vp9block.c:978-981 */
    const int16_t *uvscan = ff_vp9_scans[b->uvtx][DCT_DCT];
    int w4 = ff_vp9_bwh_tab[1][b->bs][0] << 1, h4 = ff_vp9_bwh_tab[1][b->bs][1] << 1;
    int end_x = FFMIN(2 * (s->cols - col), w4);
    int end_y = FFMIN(2 * (s->rows - row), h4);
    int step1d = 1 << b->tx, step = 1 << (b->tx * 2);
    int uvstep1d = 1 << b->uvtx, uvstep = 1 << (b->uvtx * 2);
    int x, y, n, p;

    if (b->intra) {
        const int have_left_blk = col > td->tile_col_start;
        unsigned first = c->itx;
        unsigned n_tx = ((end_x + step1d - 1) >> b->tx) * ((end_y + step1d - 1) >> b->tx) +
                        2 * (((end_x >> s->ss_h) + uvstep1d - 1) >> b->uvtx) *
                            (((end_y >> s->ss_v) + uvstep1d - 1) >> b->uvtx);
        Vp9CudaIntraTx *o;
        if (c->itx + n_tx > r->itx_cap) {
            atomic_store(&g->overflow, 1);
            return;
        }
        o = &r->itx[c->itx];
        for (n = 0, y = 0; y < end_y; y += step1d) {
            for (x = 0; x < end_x; x += step1d, n += step, o++) {
                int mode = b->mode[b->bs > BS_8x8 && b->tx == TX_4X4 ? y * 2 + x : 0];
                int eob = b->skip ? 0 : b->tx > TX_8X8 ? AV_RN16A(&td->eob[n]) : td->eob[n];
                int txtp = ff_vp9_intra_txfm_type[mode];
                o->x4 = col * 2 + x;
                o->y4 = row * 2 + y;
                o->eob = eob;
                o->plane = 0;
                o->tx = b->tx;
                o->mode = mode;
                o->txtp = txtp | lossless;
                o->flags = ((have_left_blk || x > 0) ? VP9CUDA_ITX_HAVE_LEFT : 0) |
                           (x < w4 - 1 ? VP9CUDA_ITX_HAVE_RIGHT : 0);
                o->pad = 0;
                o->ncoef = 0;
                o->level = intra_level(g, 0, o->x4, o->y4, step1d,
                                       intra_needs(mode, o->flags & VP9CUDA_ITX_HAVE_LEFT, o->y4 > 0),
                                       b->tx == TX_4X4 && (o->flags & VP9CUDA_ITX_HAVE_RIGHT));
                o->coef_off = eob ? take_coefs(g, lane, td->block + 16 * n * bytesperpixel, yscans[txtp],
                                               eob, bytesperpixel, &o->ncoef) : 0;
            }
        }
        w4 >>= s->ss_h;
        end_x >>= s->ss_h;
        end_y >>= s->ss_v;
        for (p = 0; p < 2; p++) {
            for (n = 0, y = 0; y < end_y; y += uvstep1d) {
                for (x = 0; x < end_x; x += uvstep1d, n += uvstep, o++) {
                    int eob = b->skip ? 0 : b->uvtx > TX_8X8 ? AV_RN16A(&td->uveob[p][n]) : td->uveob[p][n];
                    o->x4 = ((col * 2) >> s->ss_h) + x;
                    o->y4 = ((row * 2) >> s->ss_v) + y;
                    o->eob = eob;
                    o->plane = 1 + p;
                    o->tx = b->uvtx;
                    o->mode = b->uvmode;
                    o->txtp = DCT_DCT | lossless;
                    o->flags = ((have_left_blk || x > 0) ? VP9CUDA_ITX_HAVE_LEFT : 0) |
                               (x < w4 - 1 ? VP9CUDA_ITX_HAVE_RIGHT : 0);
                    o->pad = 0;
                    o->ncoef = 0;
                    o->level = intra_level(g, 1 + p, o->x4, o->y4, uvstep1d,
                                           intra_needs(b->uvmode, o->flags & VP9CUDA_ITX_HAVE_LEFT, o->y4 > 0),
                                           b->uvtx == TX_4X4 && (o->flags & VP9CUDA_ITX_HAVE_RIGHT));
                    o->coef_off = eob ? take_coefs(g, lane, td->uvblock[p] + 16 * n * bytesperpixel, uvscan,
                                                   eob, bytesperpixel, &o->ncoef) : 0;
                }
            }
        }
        c->itx = o - r->itx;
        if (!sbi->itx_count) {
            sbi->itx_start = first;
            sbi->lane = lane;
        }
        sbi->itx_count += c->itx - first;
    } else {
        Vp9CudaInterBlock *o;
        unsigned first = c->tx;
        /* same refusal as inter_recon(), vp9recon.c:661-669 */
        if (s->mvscale[b->ref[0]][0] == REF_INVALID_SCALE ||
            (b->comp && s->mvscale[b->ref[1]][0] == REF_INVALID_SCALE)) {
            if (!s->td->error_info) {
                s->td->error_info = AVERROR_INVALIDDATA;
                av_log(NULL, AV_LOG_ERROR, "Bitstream not supported, "
                                           "reference frame has invalid dimensions\n");
            }
            return;
        }
        if (c->blk >= r->blk_cap) {
            atomic_store(&g->overflow, 1);
            return;
        }
        o = &r->blk[c->blk++];
        o->col = col;
        o->row = row;
        o->bs = b->bs;
        o->filter = b->filter;
        o->comp = b->comp;
        o->ref[0] = b->ref[0];
        o->ref[1] = b->comp ? b->ref[1] : 0;
        memset(o->pad, 0, sizeof(o->pad));
        for (n = 0; n < 4; n++) {
            o->mv[n][0][0] = b->mv[n][0].x;
            o->mv[n][0][1] = b->mv[n][0].y;
            /* b->mv[n][1] is stale for single-reference blocks: keep the records deterministic */
            o->mv[n][1][0] = b->comp ? b->mv[n][1].x : 0;
            o->mv[n][1][1] = b->comp ? b->mv[n][1].y : 0;
        }
        o->reserved = 0;
        if (!b->skip) {
            Vp9CudaTx *t;
            for (n = 0, y = 0; y < end_y; y += step1d) {
                for (x = 0; x < end_x; x += step1d, n += step) {
                    int eob = b->tx > TX_8X8 ? AV_RN16A(&td->eob[n]) : td->eob[n];
                    if (!eob)
                        continue;
                    if (c->tx >= r->tx_cap) {
                        atomic_store(&g->overflow, 1);
                        return;
                    }
                    t = &r->tx[c->tx++];
                    t->x4 = col * 2 + x;
                    t->y4 = row * 2 + y;
                    t->eob = eob;
                    t->plane = 0;
                    t->tx = b->tx;
                    t->txtp = DCT_DCT | lossless;
                    t->pad = 0;
                    t->coef_off = take_coefs(g, lane, td->block + 16 * n * bytesperpixel, yscans[DCT_DCT],
                                             eob, bytesperpixel, &t->ncoef);
                }
            }
            end_x >>= s->ss_h;
            end_y >>= s->ss_v;
            for (p = 0; p < 2; p++) {
                for (n = 0, y = 0; y < end_y; y += uvstep1d) {
                    for (x = 0; x < end_x; x += uvstep1d, n += uvstep) {
                        int eob = b->uvtx > TX_8X8 ? AV_RN16A(&td->uveob[p][n]) : td->uveob[p][n];
                        if (!eob)
                            continue;
                        if (c->tx >= r->tx_cap) {
                            atomic_store(&g->overflow, 1);
                            return;
                        }
                        t = &r->tx[c->tx++];
                        t->x4 = ((col * 2) >> s->ss_h) + x;
                        t->y4 = ((row * 2) >> s->ss_v) + y;
                        t->eob = eob;
                        t->plane = 1 + p;
                        t->tx = b->uvtx;
                        t->txtp = DCT_DCT | lossless;
                        t->pad = 0;
                        t->coef_off = take_coefs(g, lane, td->uvblock[p] + 16 * n * bytesperpixel, uvscan,
                                                 eob, bytesperpixel, &t->ncoef);
                    }
                }
            }
        }
        if (c->tx > first) {
            if (!sbi->tx_count) {
                sbi->tx_start = first;
                sbi->lane = lane;
            }
            sbi->tx_count += c->tx - first;
        }
    }
}

void ff_vp9cuda_emit_lf(VP9Context *s, const VP9Filter *lflvl, int row, int col)
{
    VP9CudaGlue *g = s->cuda;
    Vp9CudaLoopFilterSb *o = &g->fring.lf[(row >> 3) * g->sb_cols + (col >> 3)];
    const int w = FFMIN(8, s->cols - col), h = FFMIN(8, s->rows - row);
    memcpy(o, lflvl, sizeof(*o));
    /* VP9Filter.level keeps stale values outside the frame (never read: the masks end at the frame's
     * edge); clear them so that the record stream does not depend on the decoder's history */
    if (w < 8 || h < 8)
        for (int y = 0; y < 8; y++)
            for (int x = 0; x < 8; x++)
                if (x >= w || y >= h)
                    o->level[y * 8 + x] = 0;
}

/* a lane finished a superblock row (decode_tiles vp9.c:1434, decode_tiles_mt vp9.c:1517): what it
 * has produced so far is final and may start its way to the GPU */
void ff_vp9cuda_row_done(VP9TileData *td)
{
    const VP9Context *s = td->s;
    VP9CudaGlue *g = glue_of(s);
    const int lane = td - s->td;
    if (g->in_frame && !atomic_load(&g->overflow))
        g->sh->api.submit_tile(g->sh->ctx, g->fring.frame_id, lane, &g->cnt[lane]);
}

int ff_vp9cuda_end_frame(AVCodecContext *avctx)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g = s->cuda;
    VP9CudaShared *sh = g->sh;
    AVFrame *f = s->s.frames[CUR_FRAME].tf.f;
    uint8_t *planes[3] = { f->data[0], f->data[1], f->data[2] };
    ptrdiff_t strides[3] = { f->linesize[0], f->linesize[1], f->linesize[2] };
    uint64_t ticket;
    int ret;

    if (!g->in_frame)
        return 0;
    if (atomic_load(&g->overflow)) {
        av_log(avctx, AV_LOG_ERROR, "vp9cuda: lane ring overflow\n");
        return AVERROR_BUG;              /* vp9_decode_frame's error path aborts the frame */
    }
    for (int i = 0; i < g->n_lanes; i++) {
        ret = sh->api.submit_tile(sh->ctx, g->fring.frame_id, i, &g->cnt[i]);
        if (ret < 0)
            return AVERROR_EXTERNAL;
    }
    g->in_frame = 0;                     /* finish_frame consumes the ring set whatever it returns */
    ret = sh->api.finish_frame(sh->ctx, g->fring.frame_id, planes, strides, &ticket);
    if (ret < 0)
        return AVERROR_EXTERNAL;
    g->last_ticket = ticket;
    if (g->pipeline)
        return 0;                    /* ff_vp9cuda_deliver waits when it hands the picture out */
    /* frame threading: this worker waits for its own picture while the others decode on */
    ret = sh->api.wait(sh->ctx, ticket);
    return ret < 0 ? AVERROR_EXTERNAL : 0;
}

int ff_vp9cuda_deliver(AVCodecContext *avctx, AVFrame *frame, int *got_frame, AVFrame *src, const AVPacket *pkt)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g = s->cuda;
    int ret;

    if (!g->pipeline) {
        if (!src)
            return 0;
        /* a frame's own ticket was waited for in end_frame; show_existing_frame shows an older one */
        if ((ret = av_frame_ref(frame, src)) < 0)
            return ret;
        if (pkt) {
            frame->pts     = pkt->pts;
            frame->pkt_dts = pkt->dts;
        }
        *got_frame = 1;
        return 0;
    }
    if (g->pending_valid) {
        if (g->pending_ticket && g->sh->api.wait(g->sh->ctx, g->pending_ticket) < 0)
            return AVERROR_EXTERNAL;
        av_frame_move_ref(frame, g->pending);
        g->pending_valid = 0;
        *got_frame = 1;
    }
    if (src) {
        if ((ret = av_frame_ref(g->pending, src)) < 0)
            return ret;
        if (pkt) {
            g->pending->pts     = pkt->pts;
            g->pending->pkt_dts = pkt->dts;
        }
        /* tickets complete in order: the newest one covers the copy-out of any older picture */
        g->pending_ticket = g->last_ticket;
        g->pending_valid = 1;
    }
    return 0;
}
