/*
 * vp9cuda.cu -- libvp9cuda.so: context, HBM frame / reference pool, record upload, kernel
 * sequencing and the C ABI of include/vp9cuda.h.
 *
 * Frame pipeline on the context's stream (all asynchronous to the host):
 *   H2D records -> k_cellmap -> k_inter (K2) -> k_itx x4 (K1) -> k_intra_sched + k_intra (K3)
 *   -> k_lf (K4) -> reference-slot refresh -> optional D2H of the visible crop.
 * The order is exact because nothing in reconstruction reads post-loop-filter pixels of the
 * current frame (vp9.c:1404-1416) and intra blocks read their neighbours after those got
 * their residual (vp9recon.c:109-120).
 *
 * There is no CPU implementation of anything in here: without a CUDA device every entry point
 * fails with -ENODEV.
 */
#include <cuda_runtime.h>

#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "vp9cuda.h"
#include "dev_common.cuh"
#include "k_inter.cuh"
#include "k_itx.cuh"
#include "k_intra.cuh"
#include "k_lf.cuh"

using namespace vp9;

#ifdef VP9CUDA_WATCHDOG
static int *g_wd_host;
static void wd_dump(void)
{
    if (!g_wd_host || !g_wd_host[0])
        return;
    fprintf(stderr, "vp9cuda watchdog: %d stuck waits\n", g_wd_host[0]);
    for (int i = 0; i < g_wd_host[0] && i < 255; i++) {
        const int *q = g_wd_host + 4 + 4 * i;
        fprintf(stderr, "  line %d id %d at %d block %d\n", q[0], q[1], q[2], q[3]);
    }
    g_wd_host[0] = 0;
}
#elif defined(VP9CUDA_HEARTBEAT)
static int *g_wd_host;
static void wd_dump(void)
{
    if (!g_wd_host)
        return;
    for (int b = 0; b < 64; b++)
        for (int w = 0; w < 32; w++) {
            const int *q = g_wd_host + (b * 32 + w) * 4;
            if (q[0] > 0)
                fprintf(stderr, "  block %d warp %d waits at line %d id %d at %d\n", b, w, q[0], q[1], q[2]);
        }
}
#else
static void wd_dump(void) {}
#endif

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
    fprintf(stderr, "vp9cuda: %s failed at %s:%d: %s\n", #call, __FILE__, __LINE__, cudaGetErrorString(e_)); \
    wd_dump(); return -EIO; } } while (0)

namespace {

struct Surface {
    void *pl[3] = { nullptr, nullptr, nullptr };
    int stride[3] = { 0, 0, 0 };       /* pixels */
    int w = 0, h = 0, bpp = 0, aw = 0, ah = 0;
    int refs = 0;                      /* reference slots + "current" holding this surface */
    cudaEvent_t ready = nullptr;       /* recorded on the compute stream when the picture is final */
    cudaEvent_t d2h_done = nullptr;    /* recorded on the download stream after the last copy-out */
    bool has_d2h = false;
};

/* growable device buffer */
struct DBuf {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t n)
    {
        if (n <= cap)
            return 0;
        if (p)
            cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = n + n / 4 + 256;
        if (cudaMalloc(&p, want) != cudaSuccess) {
            fprintf(stderr, "vp9cuda: cudaMalloc(%zu) failed\n", want);
            return -ENOMEM;
        }
        cap = want;
        return 0;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

/* growable pinned host buffer */
struct HBuf {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t n)
    {
        if (n <= cap)
            return 0;
        if (p)
            cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        if (cudaMallocHost(&p, n) != cudaSuccess)
            return -ENOMEM;
        cap = n;
        return 0;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

/* device-resident records + scratch of one frame */
struct DevRecords {
    DBuf blk, tx[4], itx, coef, sb_index, lf;
    DBuf cellmap, isb, irec, ires, itab, iprogress, lfops, progress, misc;
    uint32_t lane_itx_base[VP9CUDA_MAX_LANES] = { 0 };
    Vp9CudaFrameHeader hdr;
    uint32_t n_blk = 0, n_tx[4] = { 0, 0, 0, 0 }, n_itx = 0;
    bool used = false;
    cudaEvent_t uploaded = nullptr;    /* H2D of the records finished (upload stream) */
    cudaEvent_t dev_free = nullptr;    /* the kernels reading these records finished (compute stream) */
    cudaEvent_t prepped = nullptr;     /* the pixel-independent pre-passes finished (prep stream) */
    cudaEvent_t cellmapped = nullptr;  /* the cell map is built (prep stream) */
    cudaEvent_t inter_done = nullptr;  /* inter prediction finished (inter stream) */
    bool wait_upload = false, has_free = false;
    void release()
    {
        if (uploaded) cudaEventDestroy(uploaded);
        if (dev_free) cudaEventDestroy(dev_free);
        if (prepped) cudaEventDestroy(prepped);
        if (cellmapped) cudaEventDestroy(cellmapped);
        if (inter_done) cudaEventDestroy(inter_done);
        uploaded = dev_free = prepped = cellmapped = inter_done = nullptr;
        blk.release(); itx.release(); coef.release(); sb_index.release(); lf.release();
        for (auto &t : tx) t.release();
        misc.release(); cellmap.release(); isb.release(); irec.release(); ires.release(); itab.release(); iprogress.release(); lfops.release();
        progress.release();
    }
};

/* pinned rings of one frame in flight (begin_frame / tile_ring / submit_tile / finish_frame) */
struct RingSet {
    HBuf mem[VP9CUDA_MAX_LANES];           /* one slab per lane */
    Vp9CudaLaneRing ring[VP9CUDA_MAX_LANES];
    Vp9CudaLaneCounts counts[VP9CUDA_MAX_LANES];
    HBuf sb_index, lf;
    Vp9CudaFrameHeader hdr;
    uint32_t frame_id = 0;
    uint32_t n_sb = 0;
    cudaEvent_t consumed = nullptr;        /* H2D of this set finished */
    bool busy = false;
    DevRecords dev;
};

struct KStat {
    const char *name;
    double ms = 0;
    uint64_t launches = 0, bytes = 0;
};
enum { KS_CELLMAP, KS_INTER, KS_ITX, KS_ISCHED, KS_INTRA, KS_LFOPS, KS_LF, KS_WAVE, KS_H2D, KS_D2H, KS_N };

} /* namespace */

struct Vp9CudaCtx {
    int device = 0;
    cudaStream_t stream = nullptr;         /* compute */
    cudaStream_t s_up = nullptr;           /* record upload (H2D), overlaps the previous frame's kernels */
    cudaStream_t s_down = nullptr;         /* picture download (D2H), overlaps the next frame's kernels */
    cudaStream_t s_prep = nullptr;         /* pixel-independent pre-passes, overlap the previous frame / inter stage */
    cudaStream_t s_inter = nullptr;        /* inter prediction started behind the previous frame's loop filter */
    cudaEvent_t pre_wave = nullptr;        /* recorded on the compute stream right before a frame's wavefront launch */
    /* the frame whose loop filter may still be running when the next frame is submitted */
    DevRecords *prev_dev = nullptr;
    Surface *prev_surf = nullptr;
    bool prev_lf = false;
    int prev_sb_rows = 0, prev_w = 0, prev_h = 0;
    int inter_bands = 6;                   /* set_option "inter_bands" */
    int n_sm = 0;
    size_t ring_rr = 0;
    bool overlap_inter = false;            /* set_option "overlap_inter": measured +3 % (C3) .. -6 % (C2), off by default */
    bool fresh_surface = false;            /* surface_get just allocated (its clearing memset is on the compute stream) */
    int follow_mask = 3;                   /* debugging: 1 = wait for intra write-back, 2 = wait for the intra row below */
    bool trace = false;                    /* VP9CUDA_TRACE: one stderr line per submitted / finished frame */
    bool fuse_wave = true;                 /* set_option "fuse_wave": loop filter beside the intra wavefront */
    std::vector<Surface *> pool;           /* every surface ever allocated */
    Surface *ref[8] = { nullptr };
    Surface *last = nullptr;               /* most recently reconstructed frame */
    std::vector<RingSet *> rings;
    std::vector<DevRecords *> staged;
    uint32_t next_frame_id = 1;
    uint64_t next_ticket = 1;
    std::vector<cudaEvent_t> ticket_ev;    /* ring of events, ticket % size */
    std::mutex mu;
    KStat ks[KS_N];
    bool timing = false;
    /* per-stage timing without host syncs: event pairs from a pool, resolved in kernel_stats() */
    std::vector<cudaEvent_t> tpool;        /* 2 events per timed stage */
    std::vector<int> tstage;               /* stage id of pair i */
    size_t tused = 0;
    cudaEvent_t marks[16] = { nullptr };
    cudaEvent_t join_ev[4] = { nullptr, nullptr, nullptr, nullptr };
    DBuf sum_buf;
    DBuf phase_clk;                        /* optional profiling counters (set_option "phase_clk") */
    bool phase_on = false;
};

static Surface *surface_get(Vp9CudaCtx *c, int w, int h, int bpp)
{
    for (Surface *s : c->pool)
        if (!s->refs && s->w == w && s->h == h && s->bpp == bpp) {
            s->refs = 1;
            if (s->has_d2h) {               /* a copy-out of its previous picture may still be in flight */
                cudaStreamWaitEvent(c->stream, s->d2h_done, 0);
                cudaStreamWaitEvent(c->s_inter, s->d2h_done, 0);
                s->has_d2h = false;
            }
            c->fresh_surface = false;
            return s;
        }
    /* Unused surfaces of other sizes stay cached (streams with resize frames alternate between a
     * few sizes, and cudaFree / cudaMalloc would drain the pipeline every time); only when more
     * than POOL_IDLE_MAX sit idle are the idle ones of other sizes released. */
    enum { POOL_IDLE_MAX = 12 };
    int idle = 0;
    for (Surface *s : c->pool) idle += !s->refs;
    for (size_t i = 0; idle > POOL_IDLE_MAX && i < c->pool.size();) {
        Surface *s = c->pool[i];
        if (!s->refs && (s->w != w || s->h != h || s->bpp != bpp)) {
            if (s->has_d2h) cudaEventSynchronize(s->d2h_done);
            for (int p = 0; p < 3; p++) cudaFree(s->pl[p]);
            cudaEventDestroy(s->ready); cudaEventDestroy(s->d2h_done);
            delete s;
            c->pool.erase(c->pool.begin() + i);
            idle--;
        } else {
            i++;
        }
    }
    Surface *s = new Surface;
    s->w = w; s->h = h; s->bpp = bpp;
    s->aw = (w + 63) & ~63; s->ah = (h + 63) & ~63;
    const int bytes = bpp > 8 ? 2 : 1;
    for (int p = 0; p < 3; p++) {
        int sh = p ? 1 : 0;
        s->stride[p] = s->aw >> sh;
        size_t n = (size_t)s->stride[p] * (s->ah >> sh) * bytes;
        if (cudaMalloc(&s->pl[p], n) != cudaSuccess) {
            for (int q = 0; q < p; q++) cudaFree(s->pl[q]);
            delete s;
            return nullptr;
        }
        cudaMemsetAsync(s->pl[p], 0, n, c->stream);
    }
    cudaEventCreateWithFlags(&s->ready, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&s->d2h_done, cudaEventDisableTiming);
    s->refs = 1;
    c->fresh_surface = true;
    c->pool.push_back(s);
    return s;
}

static void sync_all(Vp9CudaCtx *c)
{
    cudaStreamSynchronize(c->s_up);
    cudaStreamSynchronize(c->s_prep);
    cudaStreamSynchronize(c->s_inter);
    cudaStreamSynchronize(c->stream);
    cudaStreamSynchronize(c->s_down);
}

static void surface_unref(Surface *s)
{
    if (s && s->refs > 0)
        s->refs--;
}

static SurfDesc surf_desc(const Surface *s)
{
    SurfDesc d;
    memset(&d, 0, sizeof(d));
    if (s) {
        for (int p = 0; p < 3; p++) { d.pl[p] = s->pl[p]; d.stride[p] = s->stride[p]; }
        d.w = s->w; d.h = s->h;
    }
    return d;
}

static void kstat_begin(Vp9CudaCtx *c, cudaStream_t st = nullptr)
{
    if (!st) st = c->stream;
    if (!c->timing)
        return;
    if (c->tused * 2 + 2 > c->tpool.size()) {
        cudaEvent_t a = nullptr, b = nullptr;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        c->tpool.push_back(a);
        c->tpool.push_back(b);
        c->tstage.push_back(0);
    }
    cudaEventRecord(c->tpool[c->tused * 2], st);
}
static void kstat_end(Vp9CudaCtx *c, int k, uint64_t launches, uint64_t bytes, cudaStream_t st = nullptr)
{
    if (!st) st = c->stream;
    c->ks[k].launches += launches;
    c->ks[k].bytes += bytes;
    if (c->timing) {
        cudaEventRecord(c->tpool[c->tused * 2 + 1], st);
        c->tstage[c->tused] = k;
        c->tused++;
    }
}
static void kstat_resolve(Vp9CudaCtx *c)
{
    if (!c->tused)
        return;
    cudaStreamSynchronize(c->stream);
    cudaStreamSynchronize(c->s_up);
    cudaStreamSynchronize(c->s_prep);
    cudaStreamSynchronize(c->s_inter);
    cudaStreamSynchronize(c->s_down);
    for (size_t i = 0; i < c->tused; i++) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, c->tpool[i * 2], c->tpool[i * 2 + 1]) == cudaSuccess)
            c->ks[c->tstage[i]].ms += ms;
    }
    c->tused = 0;
}

/* ------------------------------------------------------------------------------------------
 * record upload: host (pinned or pageable) -> DevRecords, lanes concatenated per record type,
 * coefficients placed at their lane's coef_base
 * ---------------------------------------------------------------------------------------- */
static int upload_records(Vp9CudaCtx *c, DevRecords &d, const Vp9CudaFrameHeader *hdr,
                          const Vp9CudaFrameRecords *recs)
{
    const int cb = hdr->bpp > 8 ? 4 : 2;
    const int nl = hdr->n_lanes;
    if (nl < 1 || nl > VP9CUDA_MAX_LANES)
        return -EINVAL;
    const uint32_t n_sb = ((hdr->w + 63) >> 6) * ((hdr->h + 63) >> 6);
    const int cols = (hdr->w + 7) >> 3, rows = (hdr->h + 7) >> 3, sb_rows = (hdr->h + 63) >> 6;
    uint64_t n_blk = 0, n_tx[4] = { 0, 0, 0, 0 }, n_itx = 0, coef_extent = 0;
    for (int l = 0; l < nl; l++) {
        const Vp9CudaLaneCounts &n = recs->counts[l];
        n_blk += n.blk; n_itx += n.itx;
        for (int t = 0; t < 4; t++) n_tx[t] += n.tx[t];
        uint64_t e = (uint64_t)recs->lanes[l].coef_base + n.coef;
        if (n.coef && e > coef_extent) coef_extent = e;
    }
    if (coef_extent > 0xffffffffull)
        return -EINVAL;
    int r = 0;
    r |= d.blk.reserve(n_blk * sizeof(Vp9CudaInterBlock));
    for (int t = 0; t < 4; t++) r |= d.tx[t].reserve(n_tx[t] * sizeof(Vp9CudaTx));
    r |= d.itx.reserve(n_itx * sizeof(Vp9CudaIntraTx));
    r |= d.coef.reserve(coef_extent * cb);
    r |= d.sb_index.reserve(n_sb * sizeof(Vp9CudaSbIndex));
    r |= d.lf.reserve(n_sb * sizeof(Vp9CudaLoopFilterSb));
    r |= d.cellmap.reserve((size_t)cols * rows * 4);
    r |= d.isb.reserve((size_t)n_sb * sizeof(IntraSb));
    r |= d.irec.reserve(n_itx * sizeof(IntraRec));
    if (n_itx) r |= d.ires.reserve((size_t)n_sb * INTRA_SB_RES * 2);
    if (n_itx) r |= d.itab.reserve((size_t)n_sb * INTRA_SB_TAB * 2);
    r |= d.iprogress.reserve((size_t)2 * sb_rows * 4);
    r |= d.lfops.reserve((size_t)n_sb * 3 * sizeof(LfSched));
    r |= d.progress.reserve((size_t)6 * sb_rows * 4);
    r |= d.misc.reserve(64);
    if (r)
        return -ENOMEM;
    if (!d.uploaded) {
        CK(cudaEventCreateWithFlags(&d.uploaded, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&d.dev_free, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&d.prepped, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&d.cellmapped, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&d.inter_done, cudaEventDisableTiming));
    }
    if (d.has_free) {                       /* the previous user of this landing zone must be done */
        CK(cudaStreamWaitEvent(c->s_up, d.dev_free, 0));
        d.has_free = false;                 /* the prep stream follows the upload, hence dev_free */
    }
    kstat_begin(c, c->s_up);
    uint64_t o_blk = 0, o_tx[4] = { 0, 0, 0, 0 }, o_itx = 0, bytes = 0;
    for (int l = 0; l < nl; l++) {
        const Vp9CudaLaneCounts &n = recs->counts[l];
        const Vp9CudaLaneRing &g = recs->lanes[l];
        if (n.blk)
            CK(cudaMemcpyAsync((char *)d.blk.p + o_blk * sizeof(Vp9CudaInterBlock), g.blk,
                               n.blk * sizeof(Vp9CudaInterBlock), cudaMemcpyHostToDevice, c->s_up));
        for (int t = 0; t < 4; t++)
            if (n.tx[t])
                CK(cudaMemcpyAsync((char *)d.tx[t].p + o_tx[t] * sizeof(Vp9CudaTx), g.tx[t],
                                   n.tx[t] * sizeof(Vp9CudaTx), cudaMemcpyHostToDevice, c->s_up));
        if (n.itx)
            CK(cudaMemcpyAsync((char *)d.itx.p + o_itx * sizeof(Vp9CudaIntraTx), g.itx,
                               n.itx * sizeof(Vp9CudaIntraTx), cudaMemcpyHostToDevice, c->s_up));
        if (n.coef)
            CK(cudaMemcpyAsync((char *)d.coef.p + (size_t)g.coef_base * cb, g.coef, (size_t)n.coef * cb,
                               cudaMemcpyHostToDevice, c->s_up));
        d.lane_itx_base[l] = (uint32_t)o_itx;
        bytes += n.blk * sizeof(Vp9CudaInterBlock) + n.itx * sizeof(Vp9CudaIntraTx) + (uint64_t)n.coef * cb;
        for (int t = 0; t < 4; t++) bytes += n.tx[t] * sizeof(Vp9CudaTx);
        o_blk += n.blk; o_itx += n.itx;
        for (int t = 0; t < 4; t++) o_tx[t] += n.tx[t];
    }
    if (n_itx) {
        CK(cudaMemcpyAsync(d.sb_index.p, recs->sb_index, n_sb * sizeof(Vp9CudaSbIndex), cudaMemcpyHostToDevice, c->s_up));
        bytes += n_sb * sizeof(Vp9CudaSbIndex);
    }
    if (hdr->lf_level) {
        CK(cudaMemcpyAsync(d.lf.p, recs->lf, n_sb * sizeof(Vp9CudaLoopFilterSb), cudaMemcpyHostToDevice, c->s_up));
        bytes += n_sb * sizeof(Vp9CudaLoopFilterSb);
    }
    kstat_end(c, KS_H2D, 0, bytes, c->s_up);
    CK(cudaEventRecord(d.uploaded, c->s_up));
    d.wait_upload = true;
    d.hdr = *hdr;
    d.n_blk = (uint32_t)n_blk; d.n_itx = (uint32_t)n_itx;
    for (int t = 0; t < 4; t++) d.n_tx[t] = (uint32_t)n_tx[t];
    d.used = true;
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * reconstruction of one frame from device-resident records
 * ---------------------------------------------------------------------------------------- */
template <class P, bool HBD>
static int launch_frame(Vp9CudaCtx *c, FrameDesc &f, DevRecords &d)
{
    cudaStream_t s = c->stream, sp = c->s_prep;
    const uint64_t P_bytes = (uint64_t)f.w * f.h * 3 / 2 * sizeof(P);
    const bool fused = c->fuse_wave && f.n_itx && f.lf_level;
    f.lf_follow = fused ? c->follow_mask : 0;
    /* ---- prep stream: everything that depends on the records only (sorting the intra blocks
     * into dependency steps, their residuals, the loop-filter edge schedule).  Runs beside the
     * inter stage of this frame and the wavefronts of the previous one. ---- */
    if (f.n_itx || f.lf_level) {
        if (d.has_free)                        /* staged records: scratch still read by their previous run */
            CK(cudaStreamWaitEvent(sp, d.dev_free, 0));
        if (d.wait_upload)
            CK(cudaStreamWaitEvent(sp, d.uploaded, 0));
        if (f.n_itx && f.n_blk) {
            /* the intra pre-pass asks the cell map which neighbours are intra */
            kstat_begin(c, sp);
            CK(cudaMemsetAsync(f.cellmap, 0xff, (size_t)f.cols * f.rows * 4, sp));
            CK(cudaMemsetAsync(f.mv_reach, 0, 4, sp));
            k_cellmap<<<(f.n_blk + 255) / 256, 256, 0, sp>>>(f.blk, f.n_blk, f.cellmap, f.cols, f.rows, f.mv_reach);
            kstat_end(c, KS_CELLMAP, 1, (uint64_t)f.n_blk * sizeof(Vp9CudaInterBlock) + (uint64_t)f.cols * f.rows * 4, sp);
            CK(cudaEventRecord(d.cellmapped, sp));
        }
        if (f.n_itx) {
            if (f.sb_cols > INTRA_MAX_SB_COLS)
                return -EINVAL;
            kstat_begin(c, sp);
            k_intra_prep<<<f.sb_cols * f.sb_rows, 128, 0, sp>>>(f);
            k_intra_res<HBD><<<f.sb_cols * f.sb_rows, IRES_WARPS * 32, 0, sp>>>(f);
            CK(cudaMemsetAsync(f.iprogress, 0, (size_t)2 * f.sb_rows * 4, sp));
            kstat_end(c, KS_ISCHED, 2, (uint64_t)f.n_itx * (sizeof(Vp9CudaIntraTx) + sizeof(IntraRec)), sp);
        }
        if (f.lf_level) {
            kstat_begin(c, sp);
            k_lf_sched<<<f.sb_cols * f.sb_rows, LF_SCHED_THREADS, 0, sp>>>(f);
            CK(cudaMemsetAsync(f.progress, 0, (size_t)6 * f.sb_rows * 4, sp));
            kstat_end(c, KS_LFOPS, 1, (uint64_t)f.sb_cols * f.sb_rows * (sizeof(Vp9CudaLoopFilterSb) + 3 * sizeof(LfSched)), sp);
        }
        CK(cudaEventRecord(d.prepped, sp));
    }
    const bool had_upload = d.wait_upload;
    if (d.wait_upload) {
        CK(cudaStreamWaitEvent(s, d.uploaded, 0));
        d.wait_upload = false;
    }
    /* ---- inter prediction: on the compute stream, or -- when the previous frame's loop filter is
     * still in flight and all of its CTAs are resident -- on the inter stream right behind that
     * wavefront, each tile waiting for the superblock rows it can reach (k_inter, dep_done) ---- */
    if (f.n_blk) {
        const bool overlap = c->overlap_inter && c->prev_dev && c->prev_dev != &d && c->prev_lf && !c->fresh_surface;
        bool depends = false;
        for (int i = 0; i < 3; i++)
            depends |= c->prev_surf && f.ref[i].pl[0] == c->prev_surf->pl[0];
        cudaStream_t si = overlap ? c->s_inter : s;
        if (overlap) {
            CK(cudaStreamWaitEvent(si, c->pre_wave, 0));     /* everything older than that wavefront is final */
            if (had_upload)
                CK(cudaStreamWaitEvent(si, d.uploaded, 0));
        }
        if (f.n_itx) {
            CK(cudaStreamWaitEvent(si, d.cellmapped, 0));
        } else {
            kstat_begin(c, si);
            CK(cudaMemsetAsync(f.cellmap, 0xff, (size_t)f.cols * f.rows * 4, si));
            CK(cudaMemsetAsync(f.mv_reach, 0, 4, si));
            k_cellmap<<<(f.n_blk + 255) / 256, 256, 0, si>>>(f.blk, f.n_blk, f.cellmap, f.cols, f.rows, f.mv_reach);
            kstat_end(c, KS_CELLMAP, 1, (uint64_t)f.n_blk * sizeof(Vp9CudaInterBlock) + (uint64_t)f.cols * f.rows * 4, si);
        }
        kstat_begin(c, si);
        const int gx = (f.cols * 8 + 63) / 64, gy = (f.rows * 8 + 31) / 32;
        if (overlap && depends) {
            /* in bands of tile rows, each behind a one-warp gate on the previous frame's loop-filter rows */
            const int dep_rows = c->prev_sb_rows, dep_whole = c->prev_w != f.w || c->prev_h != f.h;
            const int *dep_done = (const int *)c->prev_dev->progress.p + 3 * dep_rows;
            const int band = dep_whole ? gy : std::max(4, (gy + c->inter_bands - 1) / c->inter_bands);
            for (int y0 = 0; y0 < gy; y0 += band) {
                const int n = std::min(band, gy - y0);
                k_inter_gate<<<1, 32, 0, si>>>(dep_done, dep_rows, dep_whole, f.mv_reach, (y0 + n) * 32 - 1);
                f.tile_y0 = y0;
                k_inter<P><<<dim3(gx, n), 256, 0, si>>>(f);
            }
            f.tile_y0 = 0;
            kstat_end(c, KS_INTER, 2 * ((gy + band - 1) / band), 2 * P_bytes, si);
        } else {
            k_inter<P><<<dim3(gx, gy), 256, 0, si>>>(f);
            kstat_end(c, KS_INTER, 1, 2 * P_bytes, si);
        }
        if (overlap) {
            CK(cudaEventRecord(d.inter_done, si));
            CK(cudaStreamWaitEvent(s, d.inter_done, 0));
            /* the previous frame's flags were read until here: its records stay untouched that long */
            CK(cudaEventRecord(c->prev_dev->dev_free, s));
        }
    }
    kstat_begin(c);
    uint64_t itx_launches = 0, itx_bytes = 0;
    if (f.n_tx[0]) { k_itx<P, HBD, 4><<<(f.n_tx[0] + 31) / 32, 128, 0, s>>>(f); itx_launches++; }
    if (f.n_tx[1]) { k_itx<P, HBD, 8><<<(f.n_tx[1] + 15) / 16, 128, 0, s>>>(f); itx_launches++; }
    if (f.n_tx[2]) { k_itx<P, HBD, 16><<<(f.n_tx[2] + 7) / 8, 128, 0, s>>>(f); itx_launches++; }
    if (f.n_tx[3]) { k_itx<P, HBD, 32><<<(f.n_tx[3] + 3) / 4, 128, 0, s>>>(f); itx_launches++; }
    for (int t = 0; t < 4; t++)
        itx_bytes += (uint64_t)f.n_tx[t] * (16u << (2 * t)) * ((HBD ? 4 : 2) + 2 * sizeof(P));
    if (itx_launches)
        kstat_end(c, KS_ITX, itx_launches, itx_bytes);
    if (f.n_itx || f.lf_level)
        CK(cudaStreamWaitEvent(s, d.prepped, 0));
    CK(cudaEventRecord(c->pre_wave, s));
    c->prev_dev = &d;
    c->prev_lf = f.lf_level != 0;
    c->prev_sb_rows = f.sb_rows; c->prev_w = f.w; c->prev_h = f.h;
    if (fused) {
        kstat_begin(c);
        k_wave<P><<<4 * f.sb_rows, INTRA_THREADS, WaveSmem<P>::BYTES, s>>>(f);
        kstat_end(c, KS_WAVE, 1, 3 * P_bytes);
    } else {
        if (f.n_itx) {
            kstat_begin(c);
            k_intra<P><<<f.sb_rows, INTRA_THREADS, IntraSmem<P>::BYTES, s>>>(f);
            kstat_end(c, KS_INTRA, 1, P_bytes);
        }
        if (f.lf_level) {
            kstat_begin(c);
            k_lf<P><<<3 * f.sb_rows, LF_THREADS, 0, s>>>(f);
            kstat_end(c, KS_LF, 1, 2 * P_bytes);
        }
    }
    CK(cudaGetLastError());
    return 0;
}

static int copy_out(Vp9CudaCtx *c, Surface *s, uint8_t *const out_planes[3], const ptrdiff_t out_strides[3])
{
    const int bytes = s->bpp > 8 ? 2 : 1;
    uint64_t total = 0;
    CK(cudaStreamWaitEvent(c->s_down, s->ready, 0));
    kstat_begin(c, c->s_down);
    for (int p = 0; p < 3; p++) {
        int sh = p ? 1 : 0;
        int w = (s->w + sh) >> sh, h = (s->h + sh) >> sh;
        CK(cudaMemcpy2DAsync(out_planes[p], out_strides[p], s->pl[p], (size_t)s->stride[p] * bytes,
                             (size_t)w * bytes, h, cudaMemcpyDeviceToHost, c->s_down));
        total += (uint64_t)w * h * bytes;
    }
    kstat_end(c, KS_D2H, 0, total, c->s_down);
    CK(cudaEventRecord(s->d2h_done, c->s_down));
    s->has_d2h = true;
    return 0;
}

static int make_ticket(Vp9CudaCtx *c, uint64_t *ticket, cudaStream_t st)
{
    uint64_t t = c->next_ticket++;
    cudaEvent_t ev = c->ticket_ev[t % c->ticket_ev.size()];
    CK(cudaEventRecord(ev, st));
    if (ticket)
        *ticket = t;
    return 0;
}

/* VP9CUDA_TRACE: wait for `ev`; a frame that does not finish within 3 s gets its wavefront
 * counters printed (read over the idle upload stream) and the process ends. */
static void trace_wait(Vp9CudaCtx *c, const DevRecords &d, cudaEvent_t ev)
{
    for (int i = 0; cudaEventQuery(ev) == cudaErrorNotReady; i++) {
        struct timespec ts = { 0, 1000000 };
        nanosleep(&ts, nullptr);
        if (i < 3000)
            continue;
        const int sb_rows = (d.hdr.h + 63) >> 6;
        std::vector<int> ip(2 * sb_rows), lp(3 * sb_rows);
        cudaMemcpyAsync(ip.data(), d.iprogress.p, ip.size() * 4, cudaMemcpyDeviceToHost, c->s_up);
        cudaMemcpyAsync(lp.data(), d.progress.p, lp.size() * 4, cudaMemcpyDeviceToHost, c->s_up);
        cudaStreamSynchronize(c->s_up);
        fprintf(stderr, "vp9cuda: frame stuck; per superblock row: intra border / intra stored | lf Y / lf U / lf V\n");
        for (int r = 0; r < sb_rows; r++)
            fprintf(stderr, "  row %d: %d %d | %d %d %d\n", r, ip[r], ip[sb_rows + r], lp[r], lp[sb_rows + r], lp[2 * sb_rows + r]);
        fflush(stderr);
        _exit(3);
    }
}

static int reconstruct(Vp9CudaCtx *c, DevRecords &d, uint8_t *const out_planes[3],
                       const ptrdiff_t out_strides[3], uint64_t *ticket)
{
    const Vp9CudaFrameHeader &h = d.hdr;
    if (h.ss_h != 1 || h.ss_v != 1 || (h.bpp != 8 && h.bpp != 10 && h.bpp != 12) || !h.w || !h.h)
        return -EINVAL;
    Surface *cur = surface_get(c, h.w, h.h, h.bpp);
    if (!cur)
        return -ENOMEM;
    FrameDesc f;
    memset(&f, 0, sizeof(f));
    f.cur = surf_desc(cur);
    for (int i = 0; i < 3; i++) {
        Surface *r = c->ref[h.ref_slot[i] & 7];
        if (d.n_blk && (!r || r->bpp != h.bpp)) {
            surface_unref(cur);
            return -EINVAL;
        }
        f.ref[i] = surf_desc(r);
        f.mvscale[i][0] = h.mvscale[i][0]; f.mvscale[i][1] = h.mvscale[i][1];
        f.mvstep[i][0] = h.mvstep[i][0];   f.mvstep[i][1] = h.mvstep[i][1];
    }
    f.w = h.w; f.h = h.h; f.bpp = h.bpp;
    f.cols = (h.w + 7) >> 3; f.rows = (h.h + 7) >> 3;
    f.sb_cols = (h.w + 63) >> 6; f.sb_rows = (h.h + 63) >> 6;
    f.lossless = h.lossless; f.lf_level = h.lf_level;
    f.blk = (const Vp9CudaInterBlock *)d.blk.p; f.n_blk = d.n_blk;
    for (int t = 0; t < 4; t++) { f.tx[t] = (const Vp9CudaTx *)d.tx[t].p; f.n_tx[t] = d.n_tx[t]; }
    f.itx = (const Vp9CudaIntraTx *)d.itx.p; f.n_itx = d.n_itx;
    f.coef = d.coef.p;
    f.sb_index = (const Vp9CudaSbIndex *)d.sb_index.p;
    f.lf = (const Vp9CudaLoopFilterSb *)d.lf.p;
    f.cellmap = (int32_t *)d.cellmap.p;
    f.isb = (IntraSb *)d.isb.p;
    f.irec = (IntraRec *)d.irec.p;
    f.iprogress = (int *)d.iprogress.p;
    f.ires = (int16_t *)d.ires.p;
    f.itab = (uint16_t *)d.itab.p;
    memcpy(f.lane_itx_base, d.lane_itx_base, sizeof(f.lane_itx_base));
    f.lfsched = (LfSched *)d.lfops.p;
    f.progress = (int *)d.progress.p;
    f.mv_reach = (int *)d.misc.p;
    f.phase_clk = c->phase_on ? (unsigned long long *)c->phase_clk.p : nullptr;
    memcpy(f.lim_lut, h.lim_lut, 64);
    memcpy(f.mblim_lut, h.mblim_lut, 64);

    if (c->trace)
        fprintf(stderr, "vp9cuda: frame %dx%d bpp %d blk %u itx %u lf %d fuse %d\n", h.w, h.h, h.bpp, d.n_blk, d.n_itx,
                h.lf_level, (int)(c->fuse_wave && d.n_itx && h.lf_level));
    int r = h.bpp > 8 ? launch_frame<uint16_t, true>(c, f, d) : launch_frame<uint8_t, false>(c, f, d);
    if (r == 0 && d.dev_free) {
        CK(cudaEventRecord(d.dev_free, c->stream));
        d.has_free = true;
    }
    c->prev_surf = r == 0 ? cur : nullptr;
    if (r < 0)
        c->prev_dev = nullptr;
    if (r == 0)
        CK(cudaEventRecord(cur->ready, c->stream));
    if (r == 0 && c->trace)
        trace_wait(c, d, cur->ready);
    if (r < 0) {
        surface_unref(cur);
        return r;
    }
    /* reference refresh (vp9.c:1687-1692, 1848-1851) */
    for (int i = 0; i < 8; i++)
        if (h.refreshrefmask >> i & 1) {
            surface_unref(c->ref[i]);
            c->ref[i] = cur;
            cur->refs++;
        }
    surface_unref(c->last);
    c->last = cur;             /* keeps the "current" hold until the next frame */
    if (out_planes && out_planes[0]) {
        r = copy_out(c, cur, out_planes, out_strides);
        if (r < 0)
            return r;
        return make_ticket(c, ticket, c->s_down);
    }
    return make_ticket(c, ticket, c->stream);
}

/* ------------------------------------------------------------------------------------------
 * C ABI
 * ---------------------------------------------------------------------------------------- */
static const char *const ks_names[KS_N] = { "cellmap", "inter_pred", "itxfm_add", "intra_sort",
                                            "intra_wavefront", "lf_ops", "loop_filter", "intra_lf_wavefront", "h2d_records",
                                            "d2h_frame" };

extern "C" const char *vp9cuda_version(void)
{
    return "vp9cuda 0.1 (sm_100a) abi " "1";
}

extern "C" int vp9cuda_create(Vp9CudaCtx **out, int device, int max_w, int max_h, int max_bpp,
                              int frames_in_flight)
{
    (void)max_w; (void)max_h; (void)max_bpp;
    int n = 0;
    if (!out)
        return -EINVAL;
    *out = nullptr;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) {
        fprintf(stderr, "vp9cuda: no CUDA device; this back end has no CPU path\n");
        return -ENODEV;
    }
    if (device < 0 || device >= n)
        return -ENODEV;
    CK(cudaSetDevice(device));
    Vp9CudaCtx *c = new Vp9CudaCtx;
    c->device = device;
    CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&c->s_up, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&c->s_down, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&c->s_prep, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&c->s_inter, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&c->pre_wave, cudaEventDisableTiming));
    CK(cudaDeviceGetAttribute(&c->n_sm, cudaDevAttrMultiProcessorCount, device));
    if (const char *e = getenv("VP9CUDA_OVERLAP_INTER"))
        c->overlap_inter = atoi(e) != 0;
    if (const char *e = getenv("VP9CUDA_INTER_BANDS"))
        c->inter_bands = std::max(1, std::min(64, atoi(e)));
    CK(cudaFuncSetAttribute(k_wave<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, WaveSmem<uint8_t>::BYTES));
    CK(cudaFuncSetAttribute(k_wave<uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, WaveSmem<uint16_t>::BYTES));
    CK(cudaFuncSetAttribute(k_intra<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, IntraSmem<uint8_t>::BYTES));
    CK(cudaFuncSetAttribute(k_intra<uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, IntraSmem<uint16_t>::BYTES));
    c->ticket_ev.resize(256);
    for (auto &e : c->ticket_ev)
        CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto &e : c->marks)
        CK(cudaEventCreate(&e));
    for (auto &e : c->join_ev)
        CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
#if defined(VP9CUDA_WATCHDOG) || defined(VP9CUDA_HEARTBEAT)
    if (!g_wd_host) {
        int *dp = nullptr;
        CK(cudaHostAlloc((void **)&g_wd_host, 64 * 32 * 16, cudaHostAllocMapped));
        memset(g_wd_host, 0, 64 * 32 * 16);
        CK(cudaHostGetDevicePointer((void **)&dp, g_wd_host, 0));
        CK(cudaMemcpyToSymbol(vp9::g_wd_buf, &dp, sizeof(dp)));
    }
#endif
    if (const char *e = getenv("VP9CUDA_TRACE"))
        c->trace = atoi(e) != 0;
    if (const char *e = getenv("VP9CUDA_FOLLOW_MASK"))
        c->follow_mask = atoi(e);
    if (const char *e = getenv("VP9CUDA_FUSE_WAVE"))
        c->fuse_wave = atoi(e) != 0;
    if (frames_in_flight < 1) frames_in_flight = 1;
    if (frames_in_flight > 16) frames_in_flight = 16;
    for (int i = 0; i < frames_in_flight; i++) {
        RingSet *r = new RingSet;
        memset(r->ring, 0, sizeof(r->ring));
        memset(r->counts, 0, sizeof(r->counts));
        CK(cudaEventCreateWithFlags(&r->consumed, cudaEventDisableTiming));
        c->rings.push_back(r);
    }
    for (int k = 0; k < KS_N; k++)
        c->ks[k].name = ks_names[k];
    *out = c;
    return 0;
}

extern "C" int vp9cuda_flush(Vp9CudaCtx *c)
{
    if (!c)
        return -EINVAL;
    cudaSetDevice(c->device);
    sync_all(c);
    for (int i = 0; i < 8; i++) {
        surface_unref(c->ref[i]);
        c->ref[i] = nullptr;
    }
    surface_unref(c->last);
    c->last = nullptr;
    c->prev_dev = nullptr;
    c->prev_surf = nullptr;
    for (RingSet *r : c->rings)
        r->busy = false;
    return 0;
}

extern "C" void vp9cuda_destroy(Vp9CudaCtx *c)
{
    if (!c)
        return;
    cudaSetDevice(c->device);
    sync_all(c);
    for (Surface *s : c->pool) {
        for (int p = 0; p < 3; p++) cudaFree(s->pl[p]);
        cudaEventDestroy(s->ready); cudaEventDestroy(s->d2h_done);
        delete s;
    }
    for (RingSet *r : c->rings) {
        for (auto &m : r->mem) m.release();
        r->sb_index.release(); r->lf.release();
        r->dev.release();
        cudaEventDestroy(r->consumed);
        delete r;
    }
    for (DevRecords *d : c->staged)
        if (d) { d->release(); delete d; }
    for (auto &e : c->ticket_ev) cudaEventDestroy(e);
    for (auto &e : c->marks) cudaEventDestroy(e);
    for (auto &e : c->join_ev) cudaEventDestroy(e);
    for (auto &e : c->tpool) cudaEventDestroy(e);
    c->sum_buf.release();
    c->phase_clk.release();
    cudaStreamDestroy(c->stream);
    cudaStreamDestroy(c->s_up);
    cudaStreamDestroy(c->s_down);
    cudaStreamDestroy(c->s_prep);
    cudaStreamDestroy(c->s_inter);
    cudaEventDestroy(c->pre_wave);
    delete c;
}

extern "C" void *vp9cuda_host_alloc(size_t bytes)
{
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess)
        return nullptr;
    return p;
}

extern "C" void vp9cuda_host_free(void *p)
{
    if (p)
        cudaFreeHost(p);
}

static RingSet *ring_of(Vp9CudaCtx *c, uint32_t frame_id)
{
    for (RingSet *r : c->rings)
        if (r->busy && r->frame_id == frame_id)
            return r;
    return nullptr;
}

extern "C" int vp9cuda_begin_frame(Vp9CudaCtx *c, const Vp9CudaFrameHeader *hdr, Vp9CudaFrameRing *ring)
{
    if (!c || !hdr || !ring || hdr->n_lanes < 1 || hdr->n_lanes > VP9CUDA_MAX_LANES)
        return -EINVAL;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    /* round robin, so that consecutive frames land in different device scratch and the next
     * frame's upload / pre-passes / inter prediction can run beside this frame's wavefronts */
    RingSet *r = nullptr;
    for (size_t i = 0; i < c->rings.size(); i++) {
        RingSet *q = c->rings[(c->ring_rr + i) % c->rings.size()];
        if (!q->busy) { r = q; c->ring_rr = (c->ring_rr + i + 1) % c->rings.size(); break; }
    }
    if (!r)
        return -EBUSY;      /* caller must finish/wait a frame first */
    /* the previous user's H2D must be over before the producer scribbles on the pinned slab */
    CK(cudaEventSynchronize(r->consumed));
    const int cb = hdr->bpp > 8 ? 4 : 2;
    const uint32_t n_sb = ((hdr->w + 63) >> 6) * ((hdr->h + 63) >> 6);
    uint32_t coef_base = 0;
    for (int l = 0; l < hdr->n_lanes; l++) {
        const size_t sbs = hdr->lane_sbs[l] ? hdr->lane_sbs[l] : 1;
        Vp9CudaLaneRing &g = r->ring[l];
        g.blk_cap = (uint32_t)(sbs * 64);
        g.tx_cap[0] = (uint32_t)(sbs * 384); g.tx_cap[1] = (uint32_t)(sbs * 96);
        g.tx_cap[2] = (uint32_t)(sbs * 24);  g.tx_cap[3] = (uint32_t)(sbs * 6);
        g.itx_cap = (uint32_t)(sbs * 384);
        g.coef_cap = (uint32_t)(sbs * 6144);
        size_t off = 0, o_blk, o_tx[4], o_itx, o_coef;
        o_blk = off;  off += (size_t)g.blk_cap * sizeof(Vp9CudaInterBlock);
        for (int t = 0; t < 4; t++) { o_tx[t] = off; off += (size_t)g.tx_cap[t] * sizeof(Vp9CudaTx); }
        o_itx = off;  off += (size_t)g.itx_cap * sizeof(Vp9CudaIntraTx);
        off = (off + 63) & ~(size_t)63;
        o_coef = off; off += (size_t)g.coef_cap * cb;
        if (r->mem[l].reserve(off))
            return -ENOMEM;
        char *base = (char *)r->mem[l].p;
        g.blk = (Vp9CudaInterBlock *)(base + o_blk);
        for (int t = 0; t < 4; t++) g.tx[t] = (Vp9CudaTx *)(base + o_tx[t]);
        g.itx = (Vp9CudaIntraTx *)(base + o_itx);
        g.coef = base + o_coef;
        g.coef_base = coef_base;
        g.reserved = 0;
        coef_base += g.coef_cap;
        memset(&r->counts[l], 0, sizeof(r->counts[l]));
    }
    if (r->sb_index.reserve(n_sb * sizeof(Vp9CudaSbIndex)) || r->lf.reserve(n_sb * sizeof(Vp9CudaLoopFilterSb)))
        return -ENOMEM;
    memset(r->sb_index.p, 0, n_sb * sizeof(Vp9CudaSbIndex));
    memset(r->lf.p, 0, n_sb * sizeof(Vp9CudaLoopFilterSb));
    r->hdr = *hdr;
    r->n_sb = n_sb;
    r->frame_id = c->next_frame_id++;
    r->busy = true;
    ring->sb_index = (Vp9CudaSbIndex *)r->sb_index.p;
    ring->lf = (Vp9CudaLoopFilterSb *)r->lf.p;
    ring->n_sb = n_sb;
    ring->frame_id = r->frame_id;
    return 0;
}

extern "C" int vp9cuda_tile_ring(Vp9CudaCtx *c, uint32_t frame_id, int lane, Vp9CudaLaneRing *ring)
{
    if (!c || !ring)
        return -EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    RingSet *r = ring_of(c, frame_id);
    if (!r || lane < 0 || lane >= r->hdr.n_lanes)
        return -EINVAL;
    *ring = r->ring[lane];
    return 0;
}

extern "C" int vp9cuda_submit_tile(Vp9CudaCtx *c, uint32_t frame_id, int lane, const Vp9CudaLaneCounts *counts)
{
    if (!c || !counts)
        return -EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    RingSet *r = ring_of(c, frame_id);
    if (!r || lane < 0 || lane >= r->hdr.n_lanes)
        return -EINVAL;
    const Vp9CudaLaneRing &g = r->ring[lane];
    if (counts->blk > g.blk_cap || counts->itx > g.itx_cap || counts->coef > g.coef_cap)
        return -EINVAL;
    for (int t = 0; t < 4; t++)
        if (counts->tx[t] > g.tx_cap[t])
            return -EINVAL;
    r->counts[lane] = *counts;
    return 0;
}

extern "C" int vp9cuda_finish_frame(Vp9CudaCtx *c, uint32_t frame_id, uint8_t *const out_planes[3],
                                    const ptrdiff_t out_strides[3], uint64_t *ticket)
{
    if (!c)
        return -EINVAL;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    RingSet *r = ring_of(c, frame_id);
    if (!r)
        return -EINVAL;
    Vp9CudaFrameRecords recs;
    recs.lanes = r->ring;
    recs.counts = r->counts;
    recs.sb_index = (const Vp9CudaSbIndex *)r->sb_index.p;
    recs.lf = (const Vp9CudaLoopFilterSb *)r->lf.p;
    int ret = upload_records(c, r->dev, &r->hdr, &recs);
    if (ret == 0)
        CK(cudaEventRecord(r->consumed, c->s_up));
    r->busy = false;
    if (ret < 0)
        return ret;
    return reconstruct(c, r->dev, out_planes, out_strides, ticket);
}

extern "C" int vp9cuda_wait(Vp9CudaCtx *c, uint64_t ticket)
{
    if (!c || !ticket || ticket >= c->next_ticket)
        return -EINVAL;
    cudaSetDevice(c->device);
    if (c->next_ticket - ticket > c->ticket_ev.size()) {
        CK(cudaStreamSynchronize(c->stream));     /* event slot already recycled: everything older is done */
        CK(cudaStreamSynchronize(c->s_down));
        return 0;
    }
#ifdef VP9CUDA_HEARTBEAT
    for (int i = 0; cudaEventQuery(c->ticket_ev[ticket % c->ticket_ev.size()]) == cudaErrorNotReady; i++) {
        struct timespec ts = { 0, 1000000 };
        nanosleep(&ts, nullptr);
        if (i == 3000) {
            fprintf(stderr, "vp9cuda heartbeat: frame not finished after 3 s\n");
            wd_dump();
            fflush(stderr);
            _exit(3);
        }
    }
#endif
    CK(cudaEventSynchronize(c->ticket_ev[ticket % c->ticket_ev.size()]));
    CK(cudaGetLastError());
    if (c->trace)
        fprintf(stderr, "vp9cuda: ticket %llu done\n", (unsigned long long)ticket);
    return 0;
}

extern "C" int vp9cuda_show_existing(Vp9CudaCtx *c, int slot, uint8_t *const out_planes[3],
                                     const ptrdiff_t out_strides[3], uint64_t *ticket)
{
    if (!c || slot < 0 || slot > 7 || !c->ref[slot] || !out_planes)
        return -EINVAL;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    int r = copy_out(c, c->ref[slot], out_planes, out_strides);
    return r < 0 ? r : make_ticket(c, ticket, c->s_down);
}

extern "C" int vp9cuda_map_frame(Vp9CudaCtx *c, int slot, Vp9CudaMappedFrame *out)
{
    if (!c || !out || slot > 7)
        return -EINVAL;
    cudaSetDevice(c->device);
    Surface *s;
    {
        std::lock_guard<std::mutex> lk(c->mu);
        s = slot < 0 ? c->last : c->ref[slot];
        if (!s)
            return -EINVAL;
        s->refs++;                 /* keeps the surface out of the pool until unmap */
    }
    CK(cudaEventSynchronize(s->ready));
    const int bytes = s->bpp > 8 ? 2 : 1;
    for (int p = 0; p < 3; p++) {
        out->planes[p] = s->pl[p];
        out->strides[p] = (ptrdiff_t)s->stride[p] * bytes;
    }
    out->w = s->w; out->h = s->h; out->bpp = s->bpp;
    out->ss_h = out->ss_v = 1;
    return 0;
}

extern "C" int vp9cuda_unmap_frame(Vp9CudaCtx *c, const void *plane0)
{
    if (!c || !plane0)
        return -EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    for (Surface *s : c->pool)
        if (s->pl[0] == plane0 && s->refs > 0) {
            s->refs--;
            return 0;
        }
    return -EINVAL;
}

extern "C" int vp9cuda_decode_frame(Vp9CudaCtx *c, const Vp9CudaFrameHeader *hdr, const Vp9CudaFrameRecords *recs,
                                    uint8_t *const out_planes[3], const ptrdiff_t out_strides[3], uint64_t *ticket)
{
    if (!c || !hdr || !recs)
        return -EINVAL;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    /* use the device records of ring set (frame_id round robin) as the landing zone */
    RingSet *r = c->rings[c->next_frame_id++ % c->rings.size()];
    int ret = upload_records(c, r->dev, hdr, recs);
    if (ret < 0)
        return ret;
    return reconstruct(c, r->dev, out_planes, out_strides, ticket);
}

extern "C" int vp9cuda_stage_frame(Vp9CudaCtx *c, const Vp9CudaFrameHeader *hdr, const Vp9CudaFrameRecords *recs,
                                   int *handle)
{
    if (!c || !hdr || !recs || !handle)
        return -EINVAL;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    DevRecords *d = new DevRecords;
    int ret = upload_records(c, *d, hdr, recs);
    if (ret < 0) {
        d->release();
        delete d;
        return ret;
    }
    CK(cudaStreamSynchronize(c->s_up));
    d->wait_upload = false;
    c->staged.push_back(d);
    *handle = (int)c->staged.size() - 1;
    return 0;
}

extern "C" int vp9cuda_run_staged(Vp9CudaCtx *c, int handle, uint8_t *const out_planes[3],
                                  const ptrdiff_t out_strides[3], uint64_t *ticket)
{
    if (!c || handle < 0 || handle >= (int)c->staged.size() || !c->staged[handle])
        return -EINVAL;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    return reconstruct(c, *c->staged[handle], out_planes, out_strides, ticket);
}

extern "C" int vp9cuda_free_staged(Vp9CudaCtx *c, int handle)
{
    if (!c || handle < 0 || handle >= (int)c->staged.size() || !c->staged[handle])
        return -EINVAL;
    cudaSetDevice(c->device);
    sync_all(c);
    if (c->prev_dev == c->staged[handle])
        c->prev_dev = nullptr;
    c->staged[handle]->release();
    delete c->staged[handle];
    c->staged[handle] = nullptr;
    return 0;
}

extern "C" int vp9cuda_kernel_stats(Vp9CudaCtx *c, const char **names, double *ms, uint64_t *launches,
                                    uint64_t *bytes, int cap)
{
    if (!c)
        return -EINVAL;
    cudaSetDevice(c->device);
    kstat_resolve(c);
    int n = cap < KS_N ? cap : KS_N;
    for (int k = 0; k < n; k++) {
        if (names) names[k] = c->ks[k].name;
        if (ms) ms[k] = c->ks[k].ms;
        if (launches) launches[k] = c->ks[k].launches;
        if (bytes) bytes[k] = c->ks[k].bytes;
    }
    return n;
}

extern "C" int vp9cuda_set_option(Vp9CudaCtx *c, const char *key, int value)
{
    if (!c || !key)
        return -EINVAL;
    if (!strcmp(key, "timing")) {          /* per-stage CUDA-event timing on the ctx stream */
        c->timing = value != 0;
        return 0;
    }
    if (!strcmp(key, "inter_bands")) {
        if (value < 1 || value > 64)
            return -EINVAL;
        c->inter_bands = value;
        return 0;
    }
    if (!strcmp(key, "overlap_inter")) {   /* 1: inter prediction starts behind the previous frame's loop filter (default 0) */
        c->overlap_inter = value != 0;
        return 0;
    }
    if (!strcmp(key, "fuse_wave")) {       /* 1 (default): k_wave; 0: separate k_intra / k_lf launches */
        c->fuse_wave = value != 0;
        return 0;
    }
    if (!strcmp(key, "mark")) {            /* record timing mark `value` (0..15) on the ctx stream */
        if (value < 0 || value > 15)
            return -EINVAL;
        cudaSetDevice(c->device);
        /* join the upload and download streams so the mark brackets everything submitted so far */
        CK(cudaEventRecord(c->join_ev[0], c->s_up));
        CK(cudaEventRecord(c->join_ev[1], c->s_down));
        CK(cudaEventRecord(c->join_ev[2], c->s_prep));
        CK(cudaStreamWaitEvent(c->stream, c->join_ev[2], 0));
        CK(cudaEventRecord(c->join_ev[3], c->s_inter));
        CK(cudaStreamWaitEvent(c->stream, c->join_ev[3], 0));
        CK(cudaStreamWaitEvent(c->stream, c->join_ev[0], 0));
        CK(cudaStreamWaitEvent(c->stream, c->join_ev[1], 0));
        CK(cudaEventRecord(c->marks[value], c->stream));
        return 0;
    }
    if (!strcmp(key, "phase_clk")) {       /* profiling: accumulate per-phase clocks inside k_lf */
        cudaSetDevice(c->device);
        if (c->phase_clk.reserve(64 * 8))
            return -ENOMEM;
        CK(cudaStreamSynchronize(c->stream));
        CK(cudaMemset(c->phase_clk.p, 0, 64 * 8));
        c->phase_on = value != 0;
        return 0;
    }
    if (!strcmp(key, "sync")) {
        cudaSetDevice(c->device);
        sync_all(c);
        return 0;
    }
    if (!strcmp(key, "reset_stats")) {
        cudaSetDevice(c->device);
        kstat_resolve(c);
        for (int k = 0; k < KS_N; k++) { c->ks[k].ms = 0; c->ks[k].launches = 0; c->ks[k].bytes = 0; }
        return 0;
    }
    return -EINVAL;
}

extern "C" int vp9cuda_elapsed_ms(Vp9CudaCtx *c, int mark_a, int mark_b, double *ms)
{
    if (!c || !ms || mark_a < 0 || mark_a > 15 || mark_b < 0 || mark_b > 15)
        return -EINVAL;
    cudaSetDevice(c->device);
    float f = 0;
    CK(cudaEventSynchronize(c->marks[mark_b]));
    CK(cudaEventElapsedTime(&f, c->marks[mark_a], c->marks[mark_b]));
    *ms = f;
    return 0;
}

extern "C" int vp9cuda_phase_clocks(Vp9CudaCtx *c, uint64_t *out, int n)
{
    if (!c || !out || n < 1 || n > 64 || !c->phase_clk.p)
        return -EINVAL;
    cudaSetDevice(c->device);
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaMemcpy(out, c->phase_clk.p, (size_t)n * 8, cudaMemcpyDeviceToHost));
    return n;
}

/* ---- checksum of a frame in HBM ----------------------------------------------------------------- */
template <class P>
__global__ void k_checksum(const P *pl, int stride, int w, int h, unsigned long long *sum, unsigned salt)
{
    unsigned long long acc = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < w * h; i += gridDim.x * blockDim.x) {
        int x = i % w, y = i / w;
        unsigned long long v = pl[(size_t)y * stride + x];
        acc += (v + 1) * (0x9E3779B97F4A7C15ull ^ ((unsigned long long)(i + salt) * 0xD6E8FEB86659FD93ull));
    }
    for (int o = 16; o; o >>= 1)
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0)
        atomicAdd(sum, acc);
}

extern "C" int vp9cuda_frame_checksum(Vp9CudaCtx *c, int slot, uint64_t *sum)
{
    if (!c || !sum || slot > 7)
        return -EINVAL;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    Surface *s = slot < 0 ? c->last : c->ref[slot];
    if (!s)
        return -EINVAL;
    if (c->sum_buf.reserve(8))
        return -ENOMEM;
    CK(cudaMemsetAsync(c->sum_buf.p, 0, 8, c->stream));
    for (int p = 0; p < 3; p++) {
        int sh = p ? 1 : 0, w = (s->w + sh) >> sh, h = (s->h + sh) >> sh;
        if (s->bpp > 8)
            k_checksum<uint16_t><<<296, 256, 0, c->stream>>>((const uint16_t *)s->pl[p], s->stride[p], w, h,
                                                             (unsigned long long *)c->sum_buf.p, p * 0x01000000u);
        else
            k_checksum<uint8_t><<<296, 256, 0, c->stream>>>((const uint8_t *)s->pl[p], s->stride[p], w, h,
                                                            (unsigned long long *)c->sum_buf.p, p * 0x01000000u);
    }
    CK(cudaMemcpyAsync(sum, c->sum_buf.p, 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * slot-level entry points (tests): the batched equivalent of one VP9DSPContext slot.
 * Same device functions as the frame kernels.
 * ---------------------------------------------------------------------------------------- */
template <class P, bool HBD>
__global__ void k_slot_itx(int n, const uint8_t *tx, const uint8_t *txtp, const uint16_t *eob, P *pix,
                           const typename ItxTypes<HBD>::C *coef, int bpp)
{
    typedef typename ItxTypes<HBD>::C C;
    __shared__ C tmp[4][32 * 33];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = blockIdx.x * 4 + warp;
    if (i >= n)
        return;
    GlobalPix<P> g = { pix + (size_t)i * 4096, 64 };
    const C *cf = coef + (size_t)i * 1024;
    switch (tx[i]) {
    case 0: if (lane < 4)  itx_block<HBD, 4>(cf, tmp[warp], lane, txtp[i], eob[i], bpp, g, 0xfu); break;
    case 1: if (lane < 8)  itx_block<HBD, 8>(cf, tmp[warp], lane, txtp[i], eob[i], bpp, g, 0xffu); break;
    case 2: if (lane < 16) itx_block<HBD, 16>(cf, tmp[warp], lane, txtp[i], eob[i], bpp, g, 0xffffu); break;
    default: itx_block<HBD, 32>(cf, tmp[warp], lane, txtp[i], eob[i], bpp, g, 0xffffffffu); break;
    }
}

template <class P>
__global__ void k_slot_intra(int n, const uint8_t *tx, const uint8_t *mode, P *pix, const P *left,
                             const P *top, int bpp)
{
    const int i = blockIdx.x;
    if (i >= n)
        return;
    __shared__ int T[65], L[32];
    const int N = 4 << tx[i], m = mode[i];
    const P *tp = top + (size_t)i * 81, *lp = left + (size_t)i * 64;
    if (threadIdx.x == 0) {
        T[0] = tp[0];
        for (int k = 0; k < 2 * N; k++) {
            int src = k;
            if (N == 4) src = k;                     /* 4x4 may read top[4..7] */
            else if (k >= N) src = N - 1;
            T[1 + k] = tp[1 + src];
        }
        for (int y = 0; y < N; y++)
            L[y] = m == HOR_UP_PRED ? lp[y] : lp[N - 1 - y];
    }
    __syncthreads();
    int dc = intra_dc(m, N, T, L, bpp);
    for (int k = threadIdx.x; k < N * N; k += blockDim.x)
        pix[(size_t)i * 4096 + (k / N) * 64 + k % N] = (P)intra_pixel(m, N, k % N, k / N, T, L, dc, bpp);
}

template <class P>
__global__ void k_slot_mc(int n, const uint8_t *w, const uint8_t *h, const uint8_t *filter, const uint8_t *avg,
                          const uint8_t *mx, const uint8_t *my, const uint8_t *dx, const uint8_t *dy, P *pix,
                          const P *ref, int bpp)
{
    const int i = blockIdx.x;
    if (i >= n)
        return;
    RefPlane<P> rp = { ref + (size_t)i * 160 * 160, 160, 160, 160 };
    McPos pos = { 8, 8, mx[i], my[i], dx[i] ? dx[i] : 16, dy[i] ? dy[i] : 16 };
    for (int k = threadIdx.x; k < w[i] * h[i]; k += blockDim.x) {
        int ox = k % w[i], oy = k / w[i];
        int v = mc_pixel(rp, pos, ox, oy, filter[i], bpp);
        P *d = pix + (size_t)i * 4096 + oy * 64 + ox;
        *d = (P)(avg[i] ? (*d + v + 1) >> 1 : v);
    }
}

template <class P>
__global__ void k_slot_lf(int n, const uint8_t *wd, const uint8_t *dir, const uint8_t *len, const int32_t *E,
                          const int32_t *I, const int32_t *H, P *pix, int bpp)
{
    const int i = blockIdx.x;
    if (i >= n || threadIdx.x >= len[i])
        return;
    P *e = pix + (size_t)i * 4096 + 32 * 64 + 32;
    const int sa = dir[i] ? 1 : 64, sb = dir[i] ? 64 : 1;    /* dir 0: column edge, lines go down */
    e += threadIdx.x * sa;
    int px[16];
    for (int k = 0; k < 16; k++) px[k] = e[(k - 8) * sb];
    unsigned m = lf_line(px, wd[i], E[i], I[i], H[i], bpp);
    for (int k = 0; k < 16; k++)
        if (m >> k & 1) e[(k - 8) * sb] = (P)px[k];
}

namespace {
struct SlotBufs {
    std::vector<void *> d;
    ~SlotBufs() { for (void *p : d) cudaFree(p); }
    template <class T> T *up(const T *h, size_t n)
    {
        void *p = nullptr;
        if (cudaMalloc(&p, n * sizeof(T) + 16) != cudaSuccess) return nullptr;
        d.push_back(p);
        if (h) cudaMemcpy(p, h, n * sizeof(T), cudaMemcpyHostToDevice);
        return (T *)p;
    }
};
int slot_dev(int device)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n)
        return -ENODEV;
    return cudaSetDevice(device) == cudaSuccess ? 0 : -EIO;
}
}

extern "C" int vp9cuda_slot_itxfm_add(int device, int bpp, int n, const uint8_t *tx, const uint8_t *txtp,
                                      const uint16_t *eob, void *pix, const void *coef)
{
    if (slot_dev(device)) return -ENODEV;
    SlotBufs b;
    const int pb = bpp > 8 ? 2 : 1, cb = bpp > 8 ? 4 : 2;
    uint8_t *dtx = b.up(tx, n), *dtp = b.up(txtp, n);
    uint16_t *de = b.up(eob, n);
    uint8_t *dp = b.up((const uint8_t *)pix, (size_t)n * 4096 * pb);
    uint8_t *dc = b.up((const uint8_t *)coef, (size_t)n * 1024 * cb);
    if (!dtx || !dtp || !de || !dp || !dc) return -ENOMEM;
    if (bpp > 8) k_slot_itx<uint16_t, true><<<(n + 3) / 4, 128>>>(n, dtx, dtp, de, (uint16_t *)dp, (const int32_t *)dc, bpp);
    else         k_slot_itx<uint8_t, false><<<(n + 3) / 4, 128>>>(n, dtx, dtp, de, dp, (const int16_t *)dc, bpp);
    CK(cudaMemcpy(pix, dp, (size_t)n * 4096 * pb, cudaMemcpyDeviceToHost));
    CK(cudaGetLastError());
    return 0;
}

extern "C" int vp9cuda_slot_intra_pred(int device, int bpp, int n, const uint8_t *tx, const uint8_t *mode,
                                       void *pix, const void *left, const void *top)
{
    if (slot_dev(device)) return -ENODEV;
    SlotBufs b;
    const int pb = bpp > 8 ? 2 : 1;
    uint8_t *dtx = b.up(tx, n), *dm = b.up(mode, n);
    uint8_t *dp = b.up((const uint8_t *)pix, (size_t)n * 4096 * pb);
    uint8_t *dl = b.up((const uint8_t *)left, (size_t)n * 64 * pb), *dt = b.up((const uint8_t *)top, (size_t)n * 81 * pb);
    if (!dtx || !dm || !dp || !dl || !dt) return -ENOMEM;
    if (bpp > 8) k_slot_intra<uint16_t><<<n, 128>>>(n, dtx, dm, (uint16_t *)dp, (const uint16_t *)dl, (const uint16_t *)dt, bpp);
    else         k_slot_intra<uint8_t><<<n, 128>>>(n, dtx, dm, dp, dl, dt, bpp);
    CK(cudaMemcpy(pix, dp, (size_t)n * 4096 * pb, cudaMemcpyDeviceToHost));
    CK(cudaGetLastError());
    return 0;
}

extern "C" int vp9cuda_slot_mc(int device, int bpp, int n, const uint8_t *w, const uint8_t *h, const uint8_t *filter,
                               const uint8_t *avg, const uint8_t *mx, const uint8_t *my, const uint8_t *dx,
                               const uint8_t *dy, void *pix, const void *ref)
{
    if (slot_dev(device)) return -ENODEV;
    SlotBufs b;
    const int pb = bpp > 8 ? 2 : 1;
    uint8_t *dw = b.up(w, n), *dh = b.up(h, n), *df = b.up(filter, n), *da = b.up(avg, n), *dmx = b.up(mx, n),
            *dmy = b.up(my, n), *ddx = b.up(dx, n), *ddy = b.up(dy, n);
    uint8_t *dp = b.up((const uint8_t *)pix, (size_t)n * 4096 * pb);
    uint8_t *dr = b.up((const uint8_t *)ref, (size_t)n * 160 * 160 * pb);
    if (!dw || !dh || !df || !da || !dmx || !dmy || !ddx || !ddy || !dp || !dr) return -ENOMEM;
    if (bpp > 8) k_slot_mc<uint16_t><<<n, 256>>>(n, dw, dh, df, da, dmx, dmy, ddx, ddy, (uint16_t *)dp, (const uint16_t *)dr, bpp);
    else         k_slot_mc<uint8_t><<<n, 256>>>(n, dw, dh, df, da, dmx, dmy, ddx, ddy, dp, dr, bpp);
    CK(cudaMemcpy(pix, dp, (size_t)n * 4096 * pb, cudaMemcpyDeviceToHost));
    CK(cudaGetLastError());
    return 0;
}

extern "C" int vp9cuda_slot_loop_filter(int device, int bpp, int n, const uint8_t *wd, const uint8_t *dir,
                                        const uint8_t *len, const int32_t *E, const int32_t *I, const int32_t *H,
                                        void *pix)
{
    if (slot_dev(device)) return -ENODEV;
    SlotBufs b;
    const int pb = bpp > 8 ? 2 : 1;
    uint8_t *dwd = b.up(wd, n), *dd = b.up(dir, n), *dl = b.up(len, n);
    int32_t *dE = b.up(E, n), *dI = b.up(I, n), *dH = b.up(H, n);
    uint8_t *dp = b.up((const uint8_t *)pix, (size_t)n * 4096 * pb);
    if (!dwd || !dd || !dl || !dE || !dI || !dH || !dp) return -ENOMEM;
    if (bpp > 8) k_slot_lf<uint16_t><<<n, 32>>>(n, dwd, dd, dl, dE, dI, dH, (uint16_t *)dp, bpp);
    else         k_slot_lf<uint8_t><<<n, 32>>>(n, dwd, dd, dl, dE, dI, dH, dp, bpp);
    CK(cudaMemcpy(pix, dp, (size_t)n * 4096 * pb, cudaMemcpyDeviceToHost));
    CK(cudaGetLastError());
    return 0;
}
