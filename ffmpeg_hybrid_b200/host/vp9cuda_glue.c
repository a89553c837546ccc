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
    int  (*wait)(Vp9CudaCtx *, uint64_t);
    void *(*host_alloc)(size_t);             /* pinned host memory (NULL in the dump sink) */
    void (*host_free)(void *);
} SinkApi;

struct VP9CudaGlue {
    SinkApi api;
    void *dl;
    Vp9CudaCtx *ctx;
    Vp9CudaFrameHeader hdr;
    Vp9CudaFrameRing fring;
    Vp9CudaLaneRing ring[VP9CUDA_MAX_LANES];
    Vp9CudaLaneCounts cnt[VP9CUDA_MAX_LANES];
    int n_lanes;
    int in_frame;
    int overflow;
    unsigned sb_cols;
    /* intra dependency levels: per plane one (tag << 16 | level) word per 4x4 cell; a word whose
     * tag is not the current frame's means "no intra block of this frame here" */
    uint32_t *lvl[3];
    int lvl_w[3], lvl_h[3];
    size_t lvl_cap[3];
    uint32_t lvl_tag;
    /* one frame in flight: the picture handed out by the NEXT ff_vp9cuda_deliver call */
    int pipeline;
    /* output pictures in pinned memory, so that the copy-out is a real asynchronous DMA */
    AVBufferPool *frame_pool;
    size_t frame_pool_size;
    AVFrame *pending;
    int pending_valid;
    uint64_t pending_ticket, last_ticket;
};

/* ------------------------------------------------------------------------------------------
 * dump sink: same calls, written to a file (format read by ffmpeg_hybrid_b200/replay.py)
 * ---------------------------------------------------------------------------------------- */
typedef struct DumpCtx {
    FILE *fp;
    Vp9CudaFrameHeader hdr;
    Vp9CudaLaneRing ring[VP9CUDA_MAX_LANES];
    Vp9CudaLaneCounts cnt[VP9CUDA_MAX_LANES];
    unsigned lane_sbs[VP9CUDA_MAX_LANES];
    Vp9CudaSbIndex *sb_index;
    Vp9CudaLoopFilterSb *lf;
    unsigned n_sb, n_sb_cap;
    uint32_t frame_id;
} DumpCtx;

static int dump_create(Vp9CudaCtx **out, int dev, int w, int h, int bpp, int nfl)
{
    DumpCtx *d = av_mallocz(sizeof(*d));
    const char *path = getenv("VP9CUDA_DUMP");
    if (!d)
        return -ENOMEM;
    d->fp = fopen(path, "wb");
    if (!d->fp) {
        av_free(d);
        return -EIO;
    }
    fwrite("V9R2", 1, 4, d->fp);
    {
        uint32_t sz[8] = { sizeof(Vp9CudaFrameHeader), sizeof(Vp9CudaLaneCounts),
                           sizeof(Vp9CudaInterBlock), sizeof(Vp9CudaTx), sizeof(Vp9CudaIntraTx),
                           sizeof(Vp9CudaSbIndex), sizeof(Vp9CudaLoopFilterSb), VP9CUDA_ABI_VERSION };
        fwrite(sz, 4, 8, d->fp);
    }
    *out = (Vp9CudaCtx *)d;
    return 0;
}

static void dump_free_rings(DumpCtx *d)
{
    for (int l = 0; l < VP9CUDA_MAX_LANES; l++) {
        av_freep(&d->ring[l].blk);
        for (int t = 0; t < 4; t++)
            av_freep(&d->ring[l].tx[t]);
        av_freep(&d->ring[l].itx);
        av_freep(&d->ring[l].coef);
        d->lane_sbs[l] = 0;
    }
}

static void dump_destroy(Vp9CudaCtx *c)
{
    DumpCtx *d = (DumpCtx *)c;
    dump_free_rings(d);
    av_freep(&d->sb_index);
    av_freep(&d->lf);
    fclose(d->fp);
    av_free(d);
}

static int dump_flush(Vp9CudaCtx *c) { return 0; }

static int dump_begin_frame(Vp9CudaCtx *c, const Vp9CudaFrameHeader *h, Vp9CudaFrameRing *fr)
{
    DumpCtx *d = (DumpCtx *)c;
    unsigned n_sb = ((h->w + 63) >> 6) * ((h->h + 63) >> 6);
    int cb = h->bpp > 8 ? 4 : 2;
    d->hdr = *h;
    if (n_sb > d->n_sb_cap) {
        av_freep(&d->sb_index);
        av_freep(&d->lf);
        d->sb_index = av_malloc_array(n_sb, sizeof(*d->sb_index));
        d->lf = av_malloc_array(n_sb, sizeof(*d->lf));
        d->n_sb_cap = n_sb;
    }
    memset(d->sb_index, 0, n_sb * sizeof(*d->sb_index));
    memset(d->lf, 0, n_sb * sizeof(*d->lf));
    d->n_sb = n_sb;
    uint32_t coef_base = 0;
    for (int l = 0; l < h->n_lanes; l++) {
        unsigned sbs = h->lane_sbs[l];
        Vp9CudaLaneRing *r = &d->ring[l];
        if (sbs > d->lane_sbs[l]) {
            av_freep(&r->blk);
            for (int t = 0; t < 4; t++)
                av_freep(&r->tx[t]);
            av_freep(&r->itx);
            av_freep(&r->coef);
            r->blk_cap = sbs * 64;
            r->tx_cap[0] = sbs * 384; r->tx_cap[1] = sbs * 96;
            r->tx_cap[2] = sbs * 24;  r->tx_cap[3] = sbs * 6;
            r->itx_cap = sbs * 384;
            r->coef_cap = sbs * 6144;
            r->blk = av_malloc_array(r->blk_cap, sizeof(*r->blk));
            for (int t = 0; t < 4; t++)
                r->tx[t] = av_malloc_array(r->tx_cap[t], sizeof(Vp9CudaTx));
            r->itx = av_malloc_array(r->itx_cap, sizeof(*r->itx));
            r->coef = av_malloc_array(r->coef_cap, 4);
            d->lane_sbs[l] = sbs;
        }
        (void)cb;
        r->coef_base = coef_base;
        coef_base += r->coef_cap;
        memset(&d->cnt[l], 0, sizeof(d->cnt[l]));
    }
    fr->sb_index = d->sb_index;
    fr->lf = d->lf;
    fr->n_sb = n_sb;
    fr->frame_id = ++d->frame_id;
    return 0;
}

static int dump_tile_ring(Vp9CudaCtx *c, uint32_t id, int lane, Vp9CudaLaneRing *r)
{
    *r = ((DumpCtx *)c)->ring[lane];
    return 0;
}

static int dump_submit_tile(Vp9CudaCtx *c, uint32_t id, int lane, const Vp9CudaLaneCounts *n)
{
    ((DumpCtx *)c)->cnt[lane] = *n;
    return 0;
}

static int dump_finish_frame(Vp9CudaCtx *c, uint32_t id, uint8_t *const out[3],
                             const ptrdiff_t st[3], uint64_t *ticket)
{
    DumpCtx *d = (DumpCtx *)c;
    int cb = d->hdr.bpp > 8 ? 4 : 2;
    fwrite("FRAM", 1, 4, d->fp);
    fwrite(&d->hdr, sizeof(d->hdr), 1, d->fp);
    fwrite(&d->n_sb, 4, 1, d->fp);
    for (int l = 0; l < d->hdr.n_lanes; l++) {
        const Vp9CudaLaneRing *r = &d->ring[l];
        const Vp9CudaLaneCounts *n = &d->cnt[l];
        fwrite(n, sizeof(*n), 1, d->fp);
        fwrite(&r->coef_base, 4, 1, d->fp);
        fwrite(r->blk, sizeof(*r->blk), n->blk, d->fp);
        for (int t = 0; t < 4; t++)
            fwrite(r->tx[t], sizeof(Vp9CudaTx), n->tx[t], d->fp);
        fwrite(r->itx, sizeof(*r->itx), n->itx, d->fp);
        fwrite(r->coef, cb, n->coef, d->fp);
    }
    fwrite(d->sb_index, sizeof(*d->sb_index), d->n_sb, d->fp);
    fwrite(d->lf, sizeof(*d->lf), d->n_sb, d->fp);
    *ticket = id;
    return 0;
}

static int dump_wait(Vp9CudaCtx *c, uint64_t t) { return 0; }

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
    VP9CudaGlue *g = opaque;
    AVBufferRef *ref;
    uint8_t *p = g->api.host_alloc(size);
    if (!p)
        return NULL;
    ref = av_buffer_create(p, size, pinned_buffer_free, NULL, 0);
    if (!ref)
        g->api.host_free(p);
    return ref;
}

static int vp9cuda_get_buffer2(AVCodecContext *avctx, AVFrame *frame, int flags)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g = s->cuda;
    int w = frame->width, h = frame->height, ret, i;
    int align[AV_NUM_DATA_POINTERS], linesize[4];
    ptrdiff_t ls[4];
    size_t plane[4], total = 0, off = 0;

    if (!g || !g->api.host_alloc)
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
    if (!g->frame_pool || g->frame_pool_size < total) {
        /* one pool, sized for the largest picture seen (streams with resize frames alternate) */
        av_buffer_pool_uninit(&g->frame_pool);
        g->frame_pool = av_buffer_pool_init2(total, g, pinned_buffer_alloc, NULL);
        if (!g->frame_pool)
            return AVERROR(ENOMEM);
        g->frame_pool_size = total;
    }
    frame->buf[0] = av_buffer_pool_get(g->frame_pool);
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

int ff_vp9cuda_init(AVCodecContext *avctx)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g;
    int ret, dev = 0;

    if (avctx->active_thread_type & FF_THREAD_FRAME) {
        av_log(avctx, AV_LOG_ERROR, "vp9cuda: frame threading is not wired up yet; "
               "use -threads 1 or -thread_type slice\n");
        return AVERROR(ENOSYS);
    }
    g = av_mallocz(sizeof(*g));
    if (!g)
        return AVERROR(ENOMEM);
    if (getenv("VP9CUDA_DUMP")) {
        g->api = (SinkApi){ dump_create, dump_destroy, dump_flush, dump_begin_frame,
                            dump_tile_ring, dump_submit_tile, dump_finish_frame, dump_wait };
    } else {
        const char *lib = getenv("VP9CUDA_LIB");
        g->dl = dlopen(lib ? lib : "libvp9cuda.so", RTLD_NOW | RTLD_LOCAL);
        if (!g->dl) {
            av_log(avctx, AV_LOG_ERROR, "vp9cuda: cannot load back end: %s\n", dlerror());
            av_free(g);
            return AVERROR(ENODEV);
        }
#define SYM(f) do { *(void **)&g->api.f = dlsym(g->dl, "vp9cuda_" #f); \
                    if (!g->api.f) { av_log(avctx, AV_LOG_ERROR, "vp9cuda: missing symbol %s\n", #f); \
                                     dlclose(g->dl); av_free(g); return AVERROR(ENODEV); } } while (0)
        SYM(create); SYM(destroy); SYM(flush); SYM(begin_frame); SYM(tile_ring);
        SYM(submit_tile); SYM(finish_frame); SYM(wait); SYM(host_alloc); SYM(host_free);
#undef SYM
    }
    if (getenv("VP9CUDA_DEVICE"))
        dev = atoi(getenv("VP9CUDA_DEVICE"));
    ret = g->api.create(&g->ctx, dev, avctx->coded_width, avctx->coded_height, 12, 2);
    if (ret < 0) {
        av_log(avctx, AV_LOG_ERROR, "vp9cuda: back end refused to start (%d)\n", ret);
        if (g->dl)
            dlclose(g->dl);
        av_free(g);
        return AVERROR(ENODEV);
    }
    g->pipeline = !getenv("VP9CUDA_PIPELINE") || atoi(getenv("VP9CUDA_PIPELINE"));
    if (g->api.host_alloc && avctx->get_buffer2 == avcodec_default_get_buffer2 &&
        (!getenv("VP9CUDA_PINNED_FRAMES") || atoi(getenv("VP9CUDA_PINNED_FRAMES")))) {
        pinned_host_free = g->api.host_free;
        avctx->get_buffer2 = vp9cuda_get_buffer2;
    }
    g->pending = av_frame_alloc();
    if (!g->pending) {
        g->api.destroy(g->ctx);
        if (g->dl)
            dlclose(g->dl);
        av_free(g);
        return AVERROR(ENOMEM);
    }
    s->cuda = g;
    return 0;
}

void ff_vp9cuda_uninit(AVCodecContext *avctx)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g = s->cuda;
    if (!g)
        return;
    av_frame_free(&g->pending);
    av_buffer_pool_uninit(&g->frame_pool);
    g->api.destroy(g->ctx);
    if (g->dl && !pinned_host_free)    /* pinned pictures handed to the application may still be alive */
        dlclose(g->dl);
    for (int p = 0; p < 3; p++)
        av_freep(&g->lvl[p]);
    av_free(g);
    s->cuda = NULL;
}

void ff_vp9cuda_flush(AVCodecContext *avctx)
{
    VP9Context *s = avctx->priv_data;
    if (s->cuda) {
        s->cuda->api.flush(s->cuda->ctx);
        if (s->cuda->pending_valid)
            av_frame_unref(s->cuda->pending);
        s->cuda->pending_valid = 0;
        s->cuda->last_ticket = 0;
    }
}

int ff_vp9cuda_begin_frame(AVCodecContext *avctx)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g = s->cuda;
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
    for (int i = 0; i < 3; i++) {
        h->ref_slot[i] = s->s.h.refidx[i];
        h->mvscale[i][0] = s->mvscale[i][0];
        h->mvscale[i][1] = s->mvscale[i][1];
        h->mvstep[i][0] = s->mvstep[i][0];
        h->mvstep[i][1] = s->mvstep[i][1];
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
    if (h->ss_h != 1 || h->ss_v != 1) {
        av_log(avctx, AV_LOG_ERROR, "vp9cuda: only 4:2:0 is implemented\n");
        return AVERROR_PATCHWELCOME;
    }
    ret = g->api.begin_frame(g->ctx, h, &g->fring);
    if (ret < 0)
        return AVERROR_EXTERNAL;
    for (int i = 0; i < g->n_lanes; i++) {
        ret = g->api.tile_ring(g->ctx, g->fring.frame_id, i, &g->ring[i]);
        if (ret < 0)
            return AVERROR_EXTERNAL;
        memset(&g->cnt[i], 0, sizeof(g->cnt[i]));
    }
    for (int p = 0; p < 3; p++) {
        int w = (s->cols * 2) >> !!p, hh = (s->rows * 2) >> !!p;
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
    g->in_frame = 1;
    g->overflow = 0;
    return 0;
}

/* move one transform block's dequantised coefficients into the lane ring and re-zero the
 * source, which is what every itxfm_add slot does (vp9dsp_template.c:1168,1182) */
static av_always_inline uint32_t take_coefs(VP9CudaGlue *g, int lane, int16_t *src, int n_coefs,
                                            int bytesperpixel)
{
    Vp9CudaLaneRing *r = &g->ring[lane];
    Vp9CudaLaneCounts *c = &g->cnt[lane];
    uint32_t off = c->coef;
    size_t bytes = (size_t)n_coefs * 2 * bytesperpixel;
    if (off + n_coefs > r->coef_cap) {
        g->overflow = 1;
        return 0;
    }
    memcpy((uint8_t *)r->coef + (size_t)off * 2 * bytesperpixel, src, bytes);
    memset(src, 0, bytes);
    c->coef = off + n_coefs;
    return r->coef_base + off;
}

static av_always_inline int lvl_get(const VP9CudaGlue *g, int p, int x, int y)
{
    uint32_t v = g->lvl[p][(size_t)y * g->lvl_w[p] + x];
    return (v >> 16) == g->lvl_tag ? (int)(v & 0xffff) : -1;
}

/* Dependency depth of one intra transform block (n x n cells at cell x,y of plane p) and
 * registration of its cells.  The neighbour set is the superset of what check_intra_mode()
 * may read (vp9recon.c:37-221): left column incl. top-left when have_left, top row when
 * have_top, one more cell to the right for 4x4 blocks with have_right.  All of them precede
 * the block in coding order and lie in the same tile column, so tile threads never race. */
static av_always_inline int intra_level(VP9CudaGlue *g, int p, int x, int y, int n, int have_left,
                                        int have_top, int topright)
{
    const int w = g->lvl_w[p], h = g->lvl_h[p];
    int lv = 0, i, j, xe, ye;
    if (have_left) {
        ye = FFMIN(y + n, h);
        for (j = have_top ? y - 1 : y; j < ye; j++)
            lv = FFMAX(lv, lvl_get(g, p, x - 1, j) + 1);
    }
    if (have_top) {
        xe = FFMIN(x + (topright ? 2 : n), w);
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
    const int bytesperpixel = s->bytesperpixel;
    const int row = td->row, col = td->col;
    const int lossless = s->s.h.lossless ? 0x80 : 0;
    int w4 = ff_vp9_bwh_tab[1][b->bs][0] << 1, h4 = ff_vp9_bwh_tab[1][b->bs][1] << 1;
    int end_x = FFMIN(2 * (s->cols - col), w4);
    int end_y = FFMIN(2 * (s->rows - row), h4);
    int step1d = 1 << b->tx, step = 1 << (b->tx * 2);
    int uvstep1d = 1 << b->uvtx, uvstep = 1 << (b->uvtx * 2);
    int x, y, n, p;

    if (b->intra) {
        Vp9CudaSbIndex *sbi = &g->fring.sb_index[(row >> 3) * g->sb_cols + (col >> 3)];
        const int have_left_blk = col > td->tile_col_start;
        unsigned first = c->itx;
        unsigned n_tx = ((end_x + step1d - 1) >> b->tx) * ((end_y + step1d - 1) >> b->tx) +
                        2 * (((end_x >> 1) + uvstep1d - 1) >> b->uvtx) *
                            (((end_y >> 1) + uvstep1d - 1) >> b->uvtx);
        Vp9CudaIntraTx *o;
        if (c->itx + n_tx > r->itx_cap) {
            g->overflow = 1;
            return;
        }
        o = &r->itx[c->itx];
        for (n = 0, y = 0; y < end_y; y += step1d) {
            for (x = 0; x < end_x; x += step1d, n += step, o++) {
                int mode = b->mode[b->bs > BS_8x8 && b->tx == TX_4X4 ? y * 2 + x : 0];
                int eob = b->skip ? 0 : b->tx > TX_8X8 ? AV_RN16A(&td->eob[n]) : td->eob[n];
                o->x4 = col * 2 + x;
                o->y4 = row * 2 + y;
                o->eob = eob;
                o->plane = 0;
                o->tx = b->tx;
                o->mode = mode;
                o->txtp = ff_vp9_intra_txfm_type[mode] | lossless;
                o->flags = ((have_left_blk || x > 0) ? VP9CUDA_ITX_HAVE_LEFT : 0) |
                           (x < w4 - 1 ? VP9CUDA_ITX_HAVE_RIGHT : 0);
                memset(o->pad, 0, sizeof(o->pad));
                o->level = intra_level(g, 0, o->x4, o->y4, step1d, o->flags & VP9CUDA_ITX_HAVE_LEFT,
                                       o->y4 > 0, b->tx == TX_4X4 && (o->flags & VP9CUDA_ITX_HAVE_RIGHT));
                o->coef_off = eob ? take_coefs(g, lane, td->block + 16 * n * bytesperpixel,
                                               16 << (2 * b->tx), bytesperpixel) : 0;
            }
        }
        w4 >>= 1;
        end_x >>= 1;
        end_y >>= 1;
        for (p = 0; p < 2; p++) {
            for (n = 0, y = 0; y < end_y; y += uvstep1d) {
                for (x = 0; x < end_x; x += uvstep1d, n += uvstep, o++) {
                    int eob = b->skip ? 0 : b->uvtx > TX_8X8 ? AV_RN16A(&td->uveob[p][n]) : td->uveob[p][n];
                    o->x4 = col + x;
                    o->y4 = row + y;
                    o->eob = eob;
                    o->plane = 1 + p;
                    o->tx = b->uvtx;
                    o->mode = b->uvmode;
                    o->txtp = DCT_DCT | lossless;
                    o->flags = ((have_left_blk || x > 0) ? VP9CUDA_ITX_HAVE_LEFT : 0) |
                               (x < w4 - 1 ? VP9CUDA_ITX_HAVE_RIGHT : 0);
                    memset(o->pad, 0, sizeof(o->pad));
                    o->level = intra_level(g, 1 + p, o->x4, o->y4, uvstep1d, o->flags & VP9CUDA_ITX_HAVE_LEFT,
                                           o->y4 > 0, b->uvtx == TX_4X4 && (o->flags & VP9CUDA_ITX_HAVE_RIGHT));
                    o->coef_off = eob ? take_coefs(g, lane, td->uvblock[p] + 16 * n * bytesperpixel,
                                                   16 << (2 * b->uvtx), bytesperpixel) : 0;
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
            g->overflow = 1;
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
        for (n = 0; n < 4; n++)
            for (p = 0; p < 2; p++) {
                o->mv[n][p][0] = b->mv[n][p].x;
                o->mv[n][p][1] = b->mv[n][p].y;
            }
        o->reserved = 0;
        if (!b->skip) {
            Vp9CudaTx *t;
            for (n = 0, y = 0; y < end_y; y += step1d) {
                for (x = 0; x < end_x; x += step1d, n += step) {
                    int eob = b->tx > TX_8X8 ? AV_RN16A(&td->eob[n]) : td->eob[n];
                    if (!eob)
                        continue;
                    if (c->tx[b->tx] >= r->tx_cap[b->tx]) {
                        g->overflow = 1;
                        return;
                    }
                    t = &r->tx[b->tx][c->tx[b->tx]++];
                    t->x4 = col * 2 + x;
                    t->y4 = row * 2 + y;
                    t->eob = eob;
                    t->plane = 0;
                    t->txtp = DCT_DCT | lossless;
                    t->coef_off = take_coefs(g, lane, td->block + 16 * n * bytesperpixel,
                                             16 << (2 * b->tx), bytesperpixel);
                }
            }
            end_x >>= 1;
            end_y >>= 1;
            for (p = 0; p < 2; p++) {
                for (n = 0, y = 0; y < end_y; y += uvstep1d) {
                    for (x = 0; x < end_x; x += uvstep1d, n += uvstep) {
                        int eob = b->uvtx > TX_8X8 ? AV_RN16A(&td->uveob[p][n]) : td->uveob[p][n];
                        if (!eob)
                            continue;
                        if (c->tx[b->uvtx] >= r->tx_cap[b->uvtx]) {
                            g->overflow = 1;
                            return;
                        }
                        t = &r->tx[b->uvtx][c->tx[b->uvtx]++];
                        t->x4 = col + x;
                        t->y4 = row + y;
                        t->eob = eob;
                        t->plane = 1 + p;
                        t->txtp = DCT_DCT | lossless;
                        t->coef_off = take_coefs(g, lane, td->uvblock[p] + 16 * n * bytesperpixel,
                                                 16 << (2 * b->uvtx), bytesperpixel);
                    }
                }
            }
        }
    }
}

void ff_vp9cuda_emit_lf(VP9Context *s, const VP9Filter *lflvl, int row, int col)
{
    VP9CudaGlue *g = s->cuda;
    memcpy(&g->fring.lf[(row >> 3) * g->sb_cols + (col >> 3)], lflvl, sizeof(Vp9CudaLoopFilterSb));
}

int ff_vp9cuda_end_frame(AVCodecContext *avctx)
{
    VP9Context *s = avctx->priv_data;
    VP9CudaGlue *g = s->cuda;
    AVFrame *f = s->s.frames[CUR_FRAME].tf.f;
    uint8_t *planes[3] = { f->data[0], f->data[1], f->data[2] };
    ptrdiff_t strides[3] = { f->linesize[0], f->linesize[1], f->linesize[2] };
    uint64_t ticket;
    int ret;

    if (!g->in_frame)
        return 0;
    g->in_frame = 0;
    if (g->overflow) {
        av_log(avctx, AV_LOG_ERROR, "vp9cuda: lane ring overflow\n");
        return AVERROR_BUG;
    }
    for (int i = 0; i < g->n_lanes; i++) {
        ret = g->api.submit_tile(g->ctx, g->fring.frame_id, i, &g->cnt[i]);
        if (ret < 0)
            return AVERROR_EXTERNAL;
    }
    ret = g->api.finish_frame(g->ctx, g->fring.frame_id, planes, strides, &ticket);
    if (ret < 0)
        return AVERROR_EXTERNAL;
    g->last_ticket = ticket;
    if (g->pipeline)
        return 0;                    /* ff_vp9cuda_deliver waits when it hands the picture out */
    ret = g->api.wait(g->ctx, ticket);
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
        if (g->pending_ticket && g->api.wait(g->ctx, g->pending_ticket) < 0)
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
