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
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <set>
#include <vector>

#include "vp9cuda.h"
#include "dev_common.cuh"
#include "k_inter.cuh"
#include "k_itx.cuh"
#include "k_inter_sb.cuh"
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
    int ss = 1;                        /* chroma subsampling in both directions: 1 = 4:2:0, 0 = 4:4:4 */
    int refs = 0;                      /* reference slots + "current" holding this surface */
    cudaEvent_t ready = nullptr;       /* recorded on the compute stream when the picture is final */
    cudaEvent_t d2h_done = nullptr;    /* recorded on the download stream after the last copy-out */
    bool has_d2h = false;
};

/* per-superblock record capacities of a lane ring (include/vp9cuda.h, Vp9CudaLaneRing) */
enum { CAP_BLK = 64, CAP_TX = 384, CAP_ITX = 384, CAP_COEF = 6144 };
inline int coef_entry_bytes(int bpp) { return bpp > 8 ? (int)sizeof(Vp9CudaCoef32) : (int)sizeof(Vp9CudaCoef16); }
/* k = cap_scale(): a 4:4:4 superblock carries twice the samples of a 4:2:0 one */
inline int cap_scale(const Vp9CudaFrameHeader *h) { return (h->ss_h && h->ss_v) ? 1 : 2; }
inline size_t ring_bytes_per_sb(int ceb, int k)
{
    return (size_t)CAP_BLK * sizeof(Vp9CudaInterBlock) + (size_t)k * CAP_TX * sizeof(Vp9CudaTx) +
           (size_t)k * CAP_ITX * sizeof(Vp9CudaIntraTx) + (size_t)k * CAP_COEF * ceb;     /* a multiple of 64 */
}

/* device buffer that only ever grows; every (re)allocation is counted (vp9cuda_get_stat "allocs") */
struct DBuf {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t n, std::atomic<uint64_t> &allocs)
    {
        if (n <= cap)
            return 0;
        if (p)
            cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = n + 256;
        if (cudaMalloc(&p, want) != cudaSuccess) {
            fprintf(stderr, "vp9cuda: cudaMalloc(%zu) failed\n", want);
            return -ENOMEM;
        }
        allocs++;
        cap = want;
        return 0;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

/* pinned host buffer, same rules */
struct HBuf {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t n, std::atomic<uint64_t> &allocs)
    {
        if (n <= cap)
            return 0;
        if (p)
            cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        if (cudaMallocHost(&p, n) != cudaSuccess)
            return -ENOMEM;
        allocs++;
        cap = n;
        return 0;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

/* device-resident records + scratch of one frame.  Every record type is one array; lane l owns
 * the elements from lane_*_base[l] on (ring path: capacity-based, so that submit_tile can copy a
 * lane's new records to a fixed place while other lanes are still being produced; one-shot /
 * staged path: lanes packed back to back). */
struct DevRecords {
    DBuf blk, tx, itx, coef, sb_index, lf;
    DBuf cellmap, isb, irec, iblk, ires, itab, iprogress, lfops, progress, misc, ll, ill;
    uint32_t lane_blk_base[VP9CUDA_MAX_LANES] = { 0 }, lane_tx_base[VP9CUDA_MAX_LANES] = { 0 },
             lane_itx_base[VP9CUDA_MAX_LANES] = { 0 };
    uint32_t lane_blk_cnt[VP9CUDA_MAX_LANES] = { 0 }, lane_tx_cnt[VP9CUDA_MAX_LANES] = { 0 };
    Vp9CudaFrameHeader hdr;
    uint32_t n_blk = 0, n_tx = 0, n_itx = 0, max_blk = 0, max_tx = 0;
    uint64_t n_coef = 0;
    bool used = false;
    cudaEvent_t uploaded = nullptr;    /* H2D of the records finished (upload stream) */
    cudaEvent_t dev_free = nullptr;    /* the kernels reading these records finished (compute stream) */
    cudaEvent_t prepped = nullptr;     /* the pixel-independent pre-passes finished (prep stream) */
    cudaEvent_t cellmapped = nullptr;  /* the cell map is built (prep stream) */
    cudaEvent_t inter_done = nullptr;  /* inter prediction finished (inter stream) */
    bool wait_upload = false, has_free = false;
    int make_events()
    {
        if (uploaded)
            return 0;
        if (cudaEventCreateWithFlags(&uploaded, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&dev_free, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&prepped, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&cellmapped, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&inter_done, cudaEventDisableTiming) != cudaSuccess)
            return -EIO;
        return 0;
    }
    /* room for a frame of n_sb superblocks (cols x rows 8x8 cells) with the given record counts */
    int reserve(std::atomic<uint64_t> &allocs, uint32_t n_sb, int cols, int rows, int sb_rows, uint64_t cap_blk,
                uint64_t cap_tx, uint64_t cap_itx, uint64_t cap_coef, int ceb, bool intra)
    {
        int r = 0;
        r |= blk.reserve(cap_blk * sizeof(Vp9CudaInterBlock), allocs);
        r |= tx.reserve(cap_tx * sizeof(Vp9CudaTx), allocs);
        r |= itx.reserve(cap_itx * sizeof(Vp9CudaIntraTx), allocs);
        r |= coef.reserve(cap_coef * ceb, allocs);
        r |= sb_index.reserve((size_t)n_sb * sizeof(Vp9CudaSbIndex), allocs);
        r |= lf.reserve((size_t)n_sb * sizeof(Vp9CudaLoopFilterSb), allocs);
        r |= cellmap.reserve((size_t)cols * rows * 4, allocs);
        r |= isb.reserve((size_t)n_sb * sizeof(IntraSb), allocs);
        r |= irec.reserve(cap_itx * sizeof(IntraRec), allocs);
        r |= iblk.reserve(cap_itx * sizeof(IntraBlk), allocs);
        if (intra) r |= ires.reserve((size_t)n_sb * INTRA_SB_RES * 2, allocs);
        if (intra) r |= itab.reserve((size_t)n_sb * INTRA_SB_TAB * 2, allocs);
        r |= iprogress.reserve((size_t)2 * sb_rows * 4, allocs);
        r |= lfops.reserve((size_t)n_sb * 3 * sizeof(LfSched), allocs);
        r |= progress.reserve((size_t)6 * sb_rows * 4, allocs);
        {   /* hand-down words of the loop-filter rows: must not hold a tag of this context when they are new */
            const void *was = ll.p;
            r |= ll.reserve((size_t)3 * n_sb * LF_LL_WORDS * 8, allocs);
            if (!r && ll.p != was && (cudaMemset(ll.p, 0, ll.cap) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess))
                r = -1;                     /* (the context's streams do not wait for the default stream by themselves) */
            const void *iwas = ill.p;
            r |= ill.reserve((size_t)n_sb * INTRA_LL_WORDS * 8, allocs);
            if (!r && ill.p != iwas && (cudaMemset(ill.p, 0, ill.cap) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess))
                r = -1;
        }
        r |= misc.reserve(64, allocs);
        return r ? -ENOMEM : make_events();
    }
    void release()
    {
        if (uploaded) cudaEventDestroy(uploaded);
        if (dev_free) cudaEventDestroy(dev_free);
        if (prepped) cudaEventDestroy(prepped);
        if (cellmapped) cudaEventDestroy(cellmapped);
        if (inter_done) cudaEventDestroy(inter_done);
        uploaded = dev_free = prepped = cellmapped = inter_done = nullptr;
        blk.release(); tx.release(); itx.release(); coef.release(); sb_index.release(); lf.release();
        misc.release(); cellmap.release(); isb.release(); irec.release(); iblk.release(); ires.release(); itab.release(); iprogress.release(); lfops.release();
        progress.release(); ll.release(); ill.release();
    }
};

/* pinned rings of one frame in flight (begin_frame / tile_ring / submit_tile / finish_frame):
 * one slab, lane l's part starts sb_prefix[l] superblocks in */
struct RingSet {
    HBuf mem, sb_index, lf;
    Vp9CudaLaneRing ring[VP9CUDA_MAX_LANES];
    Vp9CudaLaneCounts counts[VP9CUDA_MAX_LANES];   /* submitted so far (cumulative) */
    Vp9CudaLaneCounts sent[VP9CUDA_MAX_LANES];     /* host->device copy enqueued so far */
    Vp9CudaFrameHeader hdr;
    uint32_t frame_id = 0;
    uint64_t seq = 0;                      /* position in decode order */
    uint32_t n_sb = 0;
    cudaEvent_t consumed = nullptr;        /* H2D of this set finished */
    bool busy = false;
    std::mutex mu;                         /* counts / sent */
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
    bool overlap_inter = false;            /* set_option "overlap_inter": measured +2 % (C3, 3 bands) .. -30 % (C2), off by default */
    bool fresh_surface = false;            /* surface_get just allocated (its clearing memset is on the compute stream) */
    unsigned ll_tag = 0;                   /* last tag given to a launch's loop-filter hand-down words */
    int follow_mask = 3;                   /* debugging: bit 0 = the loop filter waits for the intra write-back (bit 1 is unused since the tagged-word hand-over) */
    bool trace = false;                    /* VP9CUDA_TRACE: one stderr line per submitted / finished frame */
    bool fuse_wave = true;                 /* set_option "fuse_wave": loop filter beside the intra wavefront */
    bool inject_stall = false;             /* set_option "inject_stall": the next wavefront launch never satisfies one wait (test) */
    bool inter_sb = true;                  /* set_option "inter_sb": k_inter_sb (shared-memory windows + fused residual);
                                              0 = the strip kernel k_inter + k_itx of round 1 */
    std::vector<Surface *> pool;           /* every surface ever allocated */
    Surface *ref[8] = { nullptr };
    Surface *last = nullptr;               /* most recently reconstructed frame */
    std::vector<RingSet *> rings;
    std::vector<DevRecords *> staged;
    std::vector<DevRecords *> oneshot;     /* landing zones of vp9cuda_decode_frame, round robin */
    size_t oneshot_rr = 0;
    uint32_t next_frame_id = 1;
    std::atomic<uint64_t> next_ticket{1};
    std::vector<cudaEvent_t> ticket_ev;    /* ring of events, ticket % size */
    std::mutex mu;                         /* ring-set bookkeeping and the decode order */
    std::mutex gpu_mu;                     /* surface pool, reference slots, kernel sequencing */
    std::condition_variable cv;            /* a ring set became free / it is the next frame's turn */
    uint64_t next_begin_seq = 0;           /* decode order: handed out by begin_frame ... */
    uint64_t next_finish_seq = 0;          /* ... and honoured by finish_frame */
    std::set<uint64_t> aborted;            /* frames given back before it was their turn */
    int max_w = 0, max_h = 0, max_bpp = 0;
    std::atomic<uint64_t> n_allocs{0}, n_frames{0}, n_aborted{0}, h2d_bytes{0}, early_h2d_bytes{0}, d2h_bytes{0};
    /* host time spent inside the entry points a decoder calls per frame (ns, all calling threads summed) */
    std::atomic<uint64_t> ns_begin{0}, ns_submit{0}, ns_finish{0}, ns_wait{0}, n_submit{0}, n_submit_copies{0};
    KStat ks[KS_N];
    bool timing = false;
    /* per-stage timing without host syncs: event pairs from a pool, resolved in kernel_stats() */
    std::vector<cudaEvent_t> tpool;        /* 2 events per timed stage */
    std::vector<int> tstage;               /* stage id of pair i */
    size_t tused = 0;
    cudaEvent_t marks[16] = { nullptr };
    cudaEvent_t join_ev[4] = { nullptr, nullptr, nullptr, nullptr };
    DBuf sum_buf;
    DBuf err_dev;                          /* error word of the wavefront waits (dev_common.cuh: VP9_SPIN) */
    int *err_host = nullptr, *err_host_dev = nullptr;   /* its host-mapped twin */
    DBuf phase_clk;                        /* optional profiling counters (set_option "phase_clk") */
    bool phase_on = false;
};

static Surface *surface_get(Vp9CudaCtx *c, int w, int h, int bpp, int ss)
{
    for (Surface *s : c->pool)
        if (!s->refs && s->w == w && s->h == h && s->bpp == bpp && s->ss == ss) {
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
        if (!s->refs && (s->w != w || s->h != h || s->bpp != bpp || s->ss != ss)) {
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
    s->w = w; s->h = h; s->bpp = bpp; s->ss = ss;
    s->aw = (w + 63) & ~63; s->ah = (h + 63) & ~63;
    const int bytes = bpp > 8 ? 2 : 1;
    for (int p = 0; p < 3; p++) {
        int sh = p ? ss : 0;
        s->stride[p] = s->aw >> sh;
        size_t n = (size_t)s->stride[p] * (s->ah >> sh) * bytes;
        if (cudaMalloc(&s->pl[p], n) != cudaSuccess) {
            for (int q = 0; q < p; q++) cudaFree(s->pl[q]);
            delete s;
            return nullptr;
        }
        c->n_allocs++;
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
 * record upload: host (pinned or pageable) -> DevRecords
 * ---------------------------------------------------------------------------------------- */
static int check_header(const Vp9CudaFrameHeader *h, uint32_t *n_sb_out)
{
    if (!h->w || !h->h || h->w > 64 * INTRA_MAX_SB_COLS || h->h > 65535 ||
        (h->bpp != 8 && h->bpp != 10 && h->bpp != 12) || h->ss_h > 1 || h->ss_h != h->ss_v ||   /* 4:2:0 and 4:4:4 */
        h->n_lanes < 1 || h->n_lanes > VP9CUDA_MAX_LANES)
        return -EINVAL;
    const uint32_t n_sb = ((h->w + 63) >> 6) * ((h->h + 63) >> 6);
    uint64_t sum = 0;
    for (int l = 0; l < h->n_lanes; l++)
        sum += h->lane_sbs[l];
    if (sum > n_sb)
        return -EINVAL;
    *n_sb_out = n_sb;
    return 0;
}

/* copy records [from, to) of one lane from its ring to the lane's place in the device arrays */
static int copy_lane_range(Vp9CudaCtx *c, DevRecords &d, const Vp9CudaLaneRing &g, int lane,
                           const Vp9CudaLaneCounts &from, const Vp9CudaLaneCounts &to, int ceb, uint64_t *bytes)
{
    if (to.blk > from.blk) {
        const size_t n = (size_t)(to.blk - from.blk) * sizeof(Vp9CudaInterBlock);
        CK(cudaMemcpyAsync((char *)d.blk.p + ((size_t)d.lane_blk_base[lane] + from.blk) * sizeof(Vp9CudaInterBlock),
                           g.blk + from.blk, n, cudaMemcpyHostToDevice, c->s_up));
        *bytes += n;
    }
    if (to.tx > from.tx) {
        const size_t n = (size_t)(to.tx - from.tx) * sizeof(Vp9CudaTx);
        CK(cudaMemcpyAsync((char *)d.tx.p + ((size_t)d.lane_tx_base[lane] + from.tx) * sizeof(Vp9CudaTx),
                           g.tx + from.tx, n, cudaMemcpyHostToDevice, c->s_up));
        *bytes += n;
    }
    if (to.itx > from.itx) {
        const size_t n = (size_t)(to.itx - from.itx) * sizeof(Vp9CudaIntraTx);
        CK(cudaMemcpyAsync((char *)d.itx.p + ((size_t)d.lane_itx_base[lane] + from.itx) * sizeof(Vp9CudaIntraTx),
                           g.itx + from.itx, n, cudaMemcpyHostToDevice, c->s_up));
        *bytes += n;
    }
    if (to.coef > from.coef) {
        const size_t n = (size_t)(to.coef - from.coef) * ceb;
        CK(cudaMemcpyAsync((char *)d.coef.p + ((size_t)g.coef_base + from.coef) * ceb,
                           (const char *)g.coef + (size_t)from.coef * ceb, n, cudaMemcpyHostToDevice, c->s_up));
        *bytes += n;
    }
    return 0;
}

/* the per-superblock arrays + bookkeeping once every lane's records are on their way */
static int upload_frame_tail(Vp9CudaCtx *c, DevRecords &d, const Vp9CudaFrameHeader *hdr, uint32_t n_sb,
                             const Vp9CudaLaneCounts *counts, const Vp9CudaSbIndex *sb_index,
                             const Vp9CudaLoopFilterSb *lf, uint64_t *bytes)
{
    uint64_t n_blk = 0, n_tx = 0, n_itx = 0, n_coef = 0;
    uint32_t max_blk = 0, max_tx = 0;
    for (int l = 0; l < hdr->n_lanes; l++) {
        d.lane_blk_cnt[l] = counts[l].blk;
        d.lane_tx_cnt[l] = counts[l].tx;
        n_blk += counts[l].blk; n_tx += counts[l].tx; n_itx += counts[l].itx; n_coef += counts[l].coef;
        max_blk = std::max(max_blk, counts[l].blk);
        max_tx = std::max(max_tx, counts[l].tx);
    }
    if (n_itx || n_tx) {
        CK(cudaMemcpyAsync(d.sb_index.p, sb_index, (size_t)n_sb * sizeof(Vp9CudaSbIndex), cudaMemcpyHostToDevice, c->s_up));
        *bytes += (uint64_t)n_sb * sizeof(Vp9CudaSbIndex);
    }
    if (hdr->lf_level) {
        CK(cudaMemcpyAsync(d.lf.p, lf, (size_t)n_sb * sizeof(Vp9CudaLoopFilterSb), cudaMemcpyHostToDevice, c->s_up));
        *bytes += (uint64_t)n_sb * sizeof(Vp9CudaLoopFilterSb);
    }
    CK(cudaEventRecord(d.uploaded, c->s_up));
    d.wait_upload = true;
    d.hdr = *hdr;
    d.n_blk = (uint32_t)n_blk; d.n_tx = (uint32_t)n_tx; d.n_itx = (uint32_t)n_itx; d.n_coef = n_coef;
    d.max_blk = max_blk; d.max_tx = max_tx;
    d.used = true;
    return 0;
}

/* one-shot / staged path: all records of a frame from caller memory, lanes packed back to back */
static int upload_records(Vp9CudaCtx *c, DevRecords &d, const Vp9CudaFrameHeader *hdr,
                          const Vp9CudaFrameRecords *recs)
{
    uint32_t n_sb = 0;
    int r = check_header(hdr, &n_sb);
    if (r < 0)
        return r;
    const int ceb = coef_entry_bytes(hdr->bpp);
    const int nl = hdr->n_lanes;
    const int cols = (hdr->w + 7) >> 3, rows = (hdr->h + 7) >> 3, sb_rows = (hdr->h + 63) >> 6;
    uint64_t n_blk = 0, n_tx = 0, n_itx = 0, coef_extent = 0;
    for (int l = 0; l < nl; l++) {
        const Vp9CudaLaneCounts &n = recs->counts[l];
        d.lane_blk_base[l] = (uint32_t)n_blk; d.lane_tx_base[l] = (uint32_t)n_tx; d.lane_itx_base[l] = (uint32_t)n_itx;
        n_blk += n.blk; n_tx += n.tx; n_itx += n.itx;
        uint64_t e = (uint64_t)recs->lanes[l].coef_base + n.coef;
        if (n.coef && e > coef_extent) coef_extent = e;
    }
    if (coef_extent > 0xffffffffull || n_blk > 0x7fffffffull || n_tx > 0x7fffffffull || n_itx > 0x7fffffffull)
        return -EINVAL;
    r = d.reserve(c->n_allocs, n_sb, cols, rows, sb_rows, n_blk, n_tx, n_itx, coef_extent, ceb, n_itx != 0);
    if (r < 0)
        return r;
    if (d.has_free) {                       /* the previous user of this landing zone must be done */
        CK(cudaStreamWaitEvent(c->s_up, d.dev_free, 0));
        d.has_free = false;                 /* the prep stream follows the upload, hence dev_free */
    }
    kstat_begin(c, c->s_up);
    uint64_t bytes = 0;
    const Vp9CudaLaneCounts zero = { 0, 0, 0, 0 };
    for (int l = 0; l < nl; l++)
        if ((r = copy_lane_range(c, d, recs->lanes[l], l, zero, recs->counts[l], ceb, &bytes)) < 0)
            return r;
    r = upload_frame_tail(c, d, hdr, n_sb, recs->counts, recs->sb_index, recs->lf, &bytes);
    kstat_end(c, KS_H2D, 0, bytes, c->s_up);
    c->h2d_bytes += bytes;
    return r;
}

/* ring path: size one ring set (pinned slab + device records) for n_sb superblocks */
static int ringset_reserve(Vp9CudaCtx *c, RingSet *r, uint32_t n_sb, int cols, int rows, int sb_rows, int ceb, int k = 1)
{
    if (r->mem.reserve((size_t)n_sb * ring_bytes_per_sb(ceb, k), c->n_allocs) ||
        r->sb_index.reserve((size_t)n_sb * sizeof(Vp9CudaSbIndex), c->n_allocs) ||
        r->lf.reserve((size_t)n_sb * sizeof(Vp9CudaLoopFilterSb), c->n_allocs))
        return -ENOMEM;
    return r->dev.reserve(c->n_allocs, n_sb, cols, rows, sb_rows, (uint64_t)n_sb * CAP_BLK, (uint64_t)n_sb * k * CAP_TX,
                          (uint64_t)n_sb * k * CAP_ITX, (uint64_t)n_sb * k * CAP_COEF, ceb, true);
}

/* ------------------------------------------------------------------------------------------
 * reconstruction of one frame from device-resident records
 * ---------------------------------------------------------------------------------------- */
template <class P, bool HBD>
static int launch_frame(Vp9CudaCtx *c, FrameDesc &f, DevRecords &d)
{
    cudaStream_t s = c->stream, sp = c->s_prep;
    const uint64_t P_bytes = (uint64_t)f.w * f.h * (f.mono >= 0 ? 2 : 3) / 2 * sizeof(P);
    const bool fused = c->fuse_wave && f.n_itx && f.lf_level;
    f.lf_follow = fused ? c->follow_mask : 0;
    if (++c->ll_tag == 0)               /* the hand-down words of this launch carry a tag no earlier launch of the context used */
        c->ll_tag = 1;
    f.ll_tag = c->ll_tag;
    if (fused && c->inject_stall) {
        f.lf_follow |= 4;
        c->inject_stall = false;
    }
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
            k_cellmap<<<dim3((d.max_blk + 255) / 256, f.n_lanes), 256, 0, sp>>>(f);
            kstat_end(c, KS_CELLMAP, 1, (uint64_t)f.n_blk * sizeof(Vp9CudaInterBlock) + (uint64_t)f.cols * f.rows * 4, sp);
            CK(cudaEventRecord(d.cellmapped, sp));
        }
        if (f.n_itx) {
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
    bool residual_done = false;
    if (d.wait_upload) {
        CK(cudaStreamWaitEvent(s, d.uploaded, 0));
        d.wait_upload = false;
    }
    /* ---- inter prediction: on the compute stream, or -- when the previous frame's loop filter is
     * still in flight and all of its CTAs are resident -- on the inter stream right behind that
     * wavefront, each tile waiting for the superblock rows it can reach (k_inter, dep_done) ---- */
    if (f.n_blk) {
        const bool overlap = c->overlap_inter && f.mono < 0 && c->prev_dev && c->prev_dev != &d && c->prev_lf && !c->fresh_surface;
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
            k_cellmap<<<dim3((d.max_blk + 255) / 256, f.n_lanes), 256, 0, si>>>(f);
            kstat_end(c, KS_CELLMAP, 1, (uint64_t)f.n_blk * sizeof(Vp9CudaInterBlock) + (uint64_t)f.cols * f.rows * 4, si);
        }
        kstat_begin(c, si);
        const int gx = (f.cols * 8 + 63) / 64, gy = (f.rows * 8 + 31) / 32;
        if (overlap && depends && c->inter_sb) {
            /* in bands of superblock rows, each behind a one-warp gate on the previous frame's loop-filter rows */
            const int dep_rows = c->prev_sb_rows, dep_whole = c->prev_w != f.w || c->prev_h != f.h;
            const int *dep_done = (const int *)c->prev_dev->progress.p + 3 * dep_rows;
            const int band = dep_whole ? f.sb_rows : std::max(2, (f.sb_rows + c->inter_bands - 1) / c->inter_bands);
            for (int y0 = 0; y0 < f.sb_rows; y0 += band) {
                const int n = std::min(band, f.sb_rows - y0);
                k_inter_gate<<<1, 32, 0, si>>>(dep_done, dep_rows, dep_whole, f.mv_reach, (y0 + n) * 64 - 1, f.err, f.err_host);
                f.tile_y0 = y0;
                k_inter_sb<P, HBD><<<dim3(f.sb_cols, n), ISB_THREADS, IsbSmem<P>::BYTES, si>>>(f, 1);
            }
            f.tile_y0 = 0;
            kstat_end(c, KS_INTER, 2 * ((f.sb_rows + band - 1) / band), 2 * P_bytes + d.n_coef * sizeof(typename ItxTypes<HBD>::E), si);
            residual_done = true;
        } else if (overlap && depends) {
            /* in bands of tile rows, each behind a one-warp gate on the previous frame's loop-filter rows */
            const int dep_rows = c->prev_sb_rows, dep_whole = c->prev_w != f.w || c->prev_h != f.h;
            const int *dep_done = (const int *)c->prev_dev->progress.p + 3 * dep_rows;
            const int band = dep_whole ? gy : std::max(4, (gy + c->inter_bands - 1) / c->inter_bands);
            for (int y0 = 0; y0 < gy; y0 += band) {
                const int n = std::min(band, gy - y0);
                k_inter_gate<<<1, 32, 0, si>>>(dep_done, dep_rows, dep_whole, f.mv_reach, (y0 + n) * 32 - 1, f.err, f.err_host);
                f.tile_y0 = y0;
                k_inter<P><<<dim3(gx, n), 256, 0, si>>>(f);
            }
            f.tile_y0 = 0;
            kstat_end(c, KS_INTER, 2 * ((gy + band - 1) / band), 2 * P_bytes, si);
        } else if (c->inter_sb || f.mono >= 0) {
            /* one CTA per superblock: windows staged in shared memory, residual added in the tile */
            k_inter_sb<P, HBD><<<dim3(f.sb_cols, f.sb_rows), ISB_THREADS, IsbSmem<P>::BYTES, si>>>(f, 1);
            kstat_end(c, KS_INTER, 1, 2 * P_bytes + d.n_coef * sizeof(typename ItxTypes<HBD>::E), si);
            residual_done = true;
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
    if (f.n_tx && !residual_done) {
        kstat_begin(c);
        k_itx<P, HBD><<<dim3((d.max_tx + 3) / 4, f.n_lanes), 128, 0, s>>>(f);
        /* bytes: the coefficient entries of the frame (intra blocks included: the stream is shared) */
        kstat_end(c, KS_ITX, 1, d.n_coef * sizeof(typename ItxTypes<HBD>::E));
    }
    if (f.n_itx || f.lf_level)
        CK(cudaStreamWaitEvent(s, d.prepped, 0));
    CK(cudaEventRecord(c->pre_wave, s));
    c->prev_dev = &d;
    c->prev_lf = f.lf_level != 0;
    c->prev_sb_rows = f.sb_rows; c->prev_w = f.w; c->prev_h = f.h;
    if (fused) {
        kstat_begin(c);
        k_wave<P><<<(f.mono >= 0 ? 2 : 4) * f.sb_rows, INTRA_THREADS, WaveSmem<P>::BYTES, s>>>(f);
        kstat_end(c, KS_WAVE, 1, 3 * P_bytes);
    } else {
        if (f.n_itx) {
            kstat_begin(c);
            k_intra<P><<<f.sb_rows, INTRA_THREADS, IntraSmem<P>::BYTES, s>>>(f);
            kstat_end(c, KS_INTRA, 1, P_bytes);
        }
        if (f.lf_level) {
            kstat_begin(c);
            k_lf<P><<<(f.mono >= 0 ? 1 : 3) * f.sb_rows, LF_THREADS, 0, s>>>(f);
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
        int sh = p ? s->ss : 0;
        int w = (s->w + sh) >> sh, h = (s->h + sh) >> sh;
        CK(cudaMemcpy2DAsync(out_planes[p], out_strides[p], s->pl[p], (size_t)s->stride[p] * bytes,
                             (size_t)w * bytes, h, cudaMemcpyDeviceToHost, c->s_down));
        total += (uint64_t)w * h * bytes;
    }
    kstat_end(c, KS_D2H, 0, total, c->s_down);
    c->d2h_bytes += total;
    CK(cudaEventRecord(s->d2h_done, c->s_down));
    s->has_d2h = true;
    return 0;
}

static int make_ticket(Vp9CudaCtx *c, uint64_t *ticket, cudaStream_t st)
{
    uint64_t t = c->next_ticket.fetch_add(1);
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
    uint32_t n_sb_chk = 0;
    if (check_header(&h, &n_sb_chk) < 0)
        return -EINVAL;
    Surface *cur = surface_get(c, h.w, h.h, h.bpp, h.ss_h);
    if (!cur)
        return -ENOMEM;
    FrameDesc f;
    memset(&f, 0, sizeof(f));
    f.cur = surf_desc(cur);
    for (int i = 0; i < 3; i++) {
        Surface *r = c->ref[h.ref_slot[i] & 7];
        if (d.n_blk && (!r || r->bpp != h.bpp || r->ss != h.ss_h)) {
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
    f.tx = (const Vp9CudaTx *)d.tx.p; f.n_tx = d.n_tx;
    f.itx = (const Vp9CudaIntraTx *)d.itx.p; f.n_itx = d.n_itx;
    f.n_lanes = h.n_lanes;
    memcpy(f.lane_blk_base, d.lane_blk_base, sizeof(f.lane_blk_base));
    memcpy(f.lane_blk_cnt, d.lane_blk_cnt, sizeof(f.lane_blk_cnt));
    memcpy(f.lane_tx_base, d.lane_tx_base, sizeof(f.lane_tx_base));
    memcpy(f.lane_tx_cnt, d.lane_tx_cnt, sizeof(f.lane_tx_cnt));
    f.coef = d.coef.p;
    f.sb_index = (const Vp9CudaSbIndex *)d.sb_index.p;
    f.lf = (const Vp9CudaLoopFilterSb *)d.lf.p;
    f.cellmap = (int32_t *)d.cellmap.p;
    f.isb = (IntraSb *)d.isb.p;
    f.irec = (IntraRec *)d.irec.p;
    f.iblk = (IntraBlk *)d.iblk.p;
    f.iprogress = (int *)d.iprogress.p;
    f.ires = (int16_t *)d.ires.p;
    f.itab = (uint16_t *)d.itab.p;
    memcpy(f.lane_itx_base, d.lane_itx_base, sizeof(f.lane_itx_base));
    f.lfsched = (LfSched *)d.lfops.p;
    f.progress = (int *)d.progress.p;
    f.lf_ll = (unsigned long long *)d.ll.p;
    f.intra_ll = (unsigned long long *)d.ill.p;
    f.mv_reach = (int *)d.misc.p;
    f.phase_clk = c->phase_on ? (unsigned long long *)c->phase_clk.p : nullptr;
    f.err = (int *)c->err_dev.p;
    f.err_host = c->err_host_dev;
    memcpy(f.lim_lut, h.lim_lut, 64);
    memcpy(f.mblim_lut, h.mblim_lut, 64);

    if (c->trace)
        fprintf(stderr, "vp9cuda: frame %dx%d bpp %d blk %u itx %u lf %d fuse %d\n", h.w, h.h, h.bpp, d.n_blk, d.n_itx,
                h.lf_level, (int)(c->fuse_wave && d.n_itx && h.lf_level));
    int r = 0;
    if (h.ss_h) {
        f.mono = -1;
        r = h.bpp > 8 ? launch_frame<uint16_t, true>(c, f, d) : launch_frame<uint8_t, false>(c, f, d);
        if (r == 0 && d.dev_free) {
            CK(cudaEventRecord(d.dev_free, c->stream));
            d.has_free = true;
        }
    } else {
        /* 4:4:4: every plane is a luma-like picture of its own (FrameDesc.mono): three passes over the same records, each
         * with its plane in the place of plane 0.  The passes share the frame's scratch, so one starts its pre-passes when
         * the kernels of the one before are done (dev_free). */
        for (int p = 0; p < 3 && r == 0; p++) {
            FrameDesc fp = f;
            fp.mono = p;
            fp.cur.pl[0] = f.cur.pl[p]; fp.cur.stride[0] = f.cur.stride[p];
            for (int i = 0; i < 3; i++) { fp.ref[i].pl[0] = f.ref[i].pl[p]; fp.ref[i].stride[0] = f.ref[i].stride[p]; }
            r = h.bpp > 8 ? launch_frame<uint16_t, true>(c, fp, d) : launch_frame<uint8_t, false>(c, fp, d);
            if (r == 0 && d.dev_free) {
                CK(cudaEventRecord(d.dev_free, c->stream));
                d.has_free = true;
            }
        }
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
    c->n_frames++;
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

#define VP9CUDA_STR2(x) #x
#define VP9CUDA_STR(x) VP9CUDA_STR2(x)
extern "C" const char *vp9cuda_version(void)
{
    return "vp9cuda 0.2 (sm_100a) abi " VP9CUDA_STR(VP9CUDA_ABI_VERSION);
}

static int create_impl(Vp9CudaCtx *c, int device, int max_w, int max_h, int max_bpp, int frames_in_flight)
{
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
    CK(cudaFuncSetAttribute(k_inter_sb<uint8_t, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, IsbSmem<uint8_t>::BYTES));
    CK(cudaFuncSetAttribute(k_inter_sb<uint16_t, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, IsbSmem<uint16_t>::BYTES));
    if (const char *e = getenv("VP9CUDA_INTER_SB"))
        c->inter_sb = atoi(e) != 0;
    if (c->err_dev.reserve(64, c->n_allocs))
        return -ENOMEM;
    CK(cudaMemset(c->err_dev.p, 0, 64));
    CK(cudaHostAlloc((void **)&c->err_host, 64, cudaHostAllocMapped));
    memset(c->err_host, 0, 64);
    CK(cudaHostGetDevicePointer((void **)&c->err_host_dev, c->err_host, 0));
    c->ticket_ev.assign(256, nullptr);
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
    frames_in_flight = std::max(1, std::min(32, frames_in_flight));
    c->max_w = max_w; c->max_h = max_h; c->max_bpp = max_bpp;
    for (int i = 0; i < frames_in_flight; i++) {
        RingSet *r = new RingSet;
        memset(r->ring, 0, sizeof(r->ring));
        memset(r->counts, 0, sizeof(r->counts));
        memset(r->sent, 0, sizeof(r->sent));
        c->rings.push_back(r);
        CK(cudaEventCreateWithFlags(&r->consumed, cudaEventDisableTiming));
        if (max_w > 0 && max_h > 0) {
            /* everything a frame of the largest size can need, now, so that the frame path never allocates */
            const uint32_t n_sb = ((max_w + 63) >> 6) * ((max_h + 63) >> 6);
            if (ringset_reserve(c, r, n_sb, (max_w + 7) >> 3, (max_h + 7) >> 3, (max_h + 63) >> 6,
                                coef_entry_bytes(max_bpp)) < 0)
                return -ENOMEM;
        }
    }
    for (int k = 0; k < KS_N; k++)
        c->ks[k].name = ks_names[k];
    return 0;
}

extern "C" int vp9cuda_create(Vp9CudaCtx **out, int device, int max_w, int max_h, int max_bpp,
                              int frames_in_flight)
{
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
    if (max_w < 0 || max_h < 0 || max_w > 64 * INTRA_MAX_SB_COLS || max_h > 65535 ||
        (max_w && max_h && max_bpp != 8 && max_bpp != 10 && max_bpp != 12))
        return -EINVAL;
    CK(cudaSetDevice(device));
    Vp9CudaCtx *c = new Vp9CudaCtx;
    int r = create_impl(c, device, max_w, max_h, max_bpp, frames_in_flight);
    if (r < 0) {
        vp9cuda_destroy(c);      /* releases whatever the half-built context holds */
        return r;
    }
    *out = c;
    return 0;
}

static void advance_turn(Vp9CudaCtx *c)      /* c->mu held */
{
    c->next_finish_seq++;
    while (c->aborted.erase(c->next_finish_seq))
        c->next_finish_seq++;
    c->cv.notify_all();
}

/* give a begun frame back: c->mu held */
static void abort_set(Vp9CudaCtx *c, RingSet *r)
{
    /* copies submit_tile may have enqueued still read the slab: begin_frame waits for this event */
    cudaEventRecord(r->consumed, c->s_up);
    r->busy = false;
    c->n_aborted++;
    if (r->seq == c->next_finish_seq)
        advance_turn(c);
    else
        c->aborted.insert(r->seq);
    c->cv.notify_all();
}

extern "C" int vp9cuda_flush(Vp9CudaCtx *c)
{
    if (!c)
        return -EINVAL;
    cudaSetDevice(c->device);
    {
        std::lock_guard<std::mutex> lk(c->mu);
        for (RingSet *r : c->rings)
            if (r->busy)
                abort_set(c, r);
    }
    std::lock_guard<std::mutex> gk(c->gpu_mu);
    sync_all(c);
    if (c->err_host && *(volatile int *)c->err_host) {
        cudaMemset(c->err_dev.p, 0, 64);
        memset(c->err_host, 0, 64);
    }
    for (int i = 0; i < 8; i++) {
        surface_unref(c->ref[i]);
        c->ref[i] = nullptr;
    }
    surface_unref(c->last);
    c->last = nullptr;
    c->prev_dev = nullptr;
    c->prev_surf = nullptr;
    return 0;
}

namespace {
struct HostTimer {          /* adds the wall time of the enclosing scope to a counter */
    std::atomic<uint64_t> &acc;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    explicit HostTimer(std::atomic<uint64_t> &a) : acc(a) {}
    ~HostTimer() { acc += (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count(); }
};
}

extern "C" void vp9cuda_destroy(Vp9CudaCtx *c)
{
    if (!c)
        return;
    if (getenv("VP9CUDA_HOSTPROF"))
        fprintf(stderr, "vp9cuda host time inside the library: frames %llu, begin_frame %.3f ms, submit_tile %.3f ms (%llu calls, %llu with a copy), "
                        "finish_frame %.3f ms, wait %.3f ms (all calling threads summed)\n", (unsigned long long)c->n_frames.load(),
                c->ns_begin.load() * 1e-6, c->ns_submit.load() * 1e-6, (unsigned long long)c->n_submit.load(),
                (unsigned long long)c->n_submit_copies.load(), c->ns_finish.load() * 1e-6, c->ns_wait.load() * 1e-6);
    cudaSetDevice(c->device);
    if (c->stream)
        sync_all(c);
    for (Surface *s : c->pool) {
        for (int p = 0; p < 3; p++) cudaFree(s->pl[p]);
        cudaEventDestroy(s->ready); cudaEventDestroy(s->d2h_done);
        delete s;
    }
    for (RingSet *r : c->rings) {
        r->mem.release();
        r->sb_index.release(); r->lf.release();
        r->dev.release();
        if (r->consumed) cudaEventDestroy(r->consumed);
        delete r;
    }
    for (DevRecords *d : c->staged)
        if (d) { d->release(); delete d; }
    for (DevRecords *d : c->oneshot)
        if (d) { d->release(); delete d; }
    for (auto &e : c->ticket_ev) if (e) cudaEventDestroy(e);
    for (auto &e : c->marks) if (e) cudaEventDestroy(e);
    for (auto &e : c->join_ev) if (e) cudaEventDestroy(e);
    for (auto &e : c->tpool) cudaEventDestroy(e);
    c->sum_buf.release();
    c->err_dev.release();
    if (c->err_host) cudaFreeHost(c->err_host);
    c->phase_clk.release();
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->s_up) cudaStreamDestroy(c->s_up);
    if (c->s_down) cudaStreamDestroy(c->s_down);
    if (c->s_prep) cudaStreamDestroy(c->s_prep);
    if (c->s_inter) cudaStreamDestroy(c->s_inter);
    if (c->pre_wave) cudaEventDestroy(c->pre_wave);
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
    uint32_t n_sb = 0;
    if (!c || !hdr || !ring || check_header(hdr, &n_sb) < 0)
        return -EINVAL;
    HostTimer ht(c->ns_begin);
    cudaSetDevice(c->device);
    std::unique_lock<std::mutex> lk(c->mu);
    /* round robin, so that consecutive frames land in different device scratch and the next
     * frame's upload / pre-passes / inter prediction can run beside this frame's wavefronts */
    RingSet *r = nullptr;
    auto pick = [&]() {
        for (size_t i = 0; i < c->rings.size(); i++) {
            RingSet *q = c->rings[(c->ring_rr + i) % c->rings.size()];
            if (!q->busy) { r = q; c->ring_rr = (c->ring_rr + i + 1) % c->rings.size(); return true; }
        }
        return false;
    };
    /* all sets in use: another thread's finish_frame / abort_frame frees one (frame threading);
     * a caller that never finishes its frames gets -EBUSY instead of a hang */
    if (!c->cv.wait_for(lk, std::chrono::seconds(20), pick))
        return -EBUSY;
    r->busy = true;
    r->frame_id = c->next_frame_id++;
    if (!c->next_frame_id) c->next_frame_id = 1;
    r->seq = c->next_begin_seq++;
    r->hdr = *hdr;
    r->n_sb = n_sb;
    auto fail = [&](int e) { abort_set(c, r); return e; };
    /* the previous user's H2D must be over before the producer scribbles on the pinned slab */
    if (cudaEventSynchronize(r->consumed) != cudaSuccess)
        return fail(-EIO);
    const int ceb = coef_entry_bytes(hdr->bpp), k = cap_scale(hdr);
    const int cols = (hdr->w + 7) >> 3, rows = (hdr->h + 7) >> 3, sb_rows = (hdr->h + 63) >> 6;
    if (ringset_reserve(c, r, n_sb, cols, rows, sb_rows, ceb, k) < 0)
        return fail(-ENOMEM);
    DevRecords &d = r->dev;
    if (d.has_free) {                       /* the kernels of this landing zone's previous frame */
        if (cudaStreamWaitEvent(c->s_up, d.dev_free, 0) != cudaSuccess)
            return fail(-EIO);
        d.has_free = false;
    }
    uint32_t sb0 = 0;
    for (int l = 0; l < hdr->n_lanes; l++) {
        const uint32_t sbs = hdr->lane_sbs[l];
        Vp9CudaLaneRing &g = r->ring[l];
        char *base = (char *)r->mem.p + (size_t)sb0 * ring_bytes_per_sb(ceb, k);
        g.blk_cap = sbs * CAP_BLK; g.tx_cap = sbs * k * CAP_TX; g.itx_cap = sbs * k * CAP_ITX; g.coef_cap = sbs * k * CAP_COEF;
        g.blk = (Vp9CudaInterBlock *)base;              base += (size_t)g.blk_cap * sizeof(Vp9CudaInterBlock);
        g.tx = (Vp9CudaTx *)base;                       base += (size_t)g.tx_cap * sizeof(Vp9CudaTx);
        g.itx = (Vp9CudaIntraTx *)base;                 base += (size_t)g.itx_cap * sizeof(Vp9CudaIntraTx);
        g.coef = base;
        g.coef_base = sb0 * k * CAP_COEF;
        g.reserved = 0;
        d.lane_blk_base[l] = sb0 * CAP_BLK; d.lane_tx_base[l] = sb0 * k * CAP_TX; d.lane_itx_base[l] = sb0 * k * CAP_ITX;
        memset(&r->counts[l], 0, sizeof(r->counts[l]));
        memset(&r->sent[l], 0, sizeof(r->sent[l]));
        sb0 += sbs;
    }
    memset(r->sb_index.p, 0, (size_t)n_sb * sizeof(Vp9CudaSbIndex));
    memset(r->lf.p, 0, (size_t)n_sb * sizeof(Vp9CudaLoopFilterSb));
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
    HostTimer ht(c->ns_submit);
    c->n_submit++;
    RingSet *r;
    {
        std::lock_guard<std::mutex> lk(c->mu);
        r = ring_of(c, frame_id);
    }
    if (!r || lane < 0 || lane >= r->hdr.n_lanes)
        return -EINVAL;
    std::lock_guard<std::mutex> lk(r->mu);
    const Vp9CudaLaneRing &g = r->ring[lane];
    const Vp9CudaLaneCounts &was = r->counts[lane];
    if (counts->blk > g.blk_cap || counts->tx > g.tx_cap || counts->itx > g.itx_cap || counts->coef > g.coef_cap ||
        counts->blk < was.blk || counts->tx < was.tx || counts->itx < was.itx || counts->coef < was.coef)
        return -EINVAL;
    r->counts[lane] = *counts;
    /* start the copy of what is new; small increments are left to ride with the next call */
    const uint64_t fresh = (uint64_t)(counts->blk - r->sent[lane].blk) * sizeof(Vp9CudaInterBlock) +
                           (uint64_t)(counts->tx - r->sent[lane].tx) * sizeof(Vp9CudaTx) +
                           (uint64_t)(counts->itx - r->sent[lane].itx) * sizeof(Vp9CudaIntraTx) +
                           (uint64_t)(counts->coef - r->sent[lane].coef) * coef_entry_bytes(r->hdr.bpp);
    if (fresh < 64 * 1024)
        return 0;
    c->n_submit_copies++;
    cudaSetDevice(c->device);
    uint64_t bytes = 0;
    int ret = copy_lane_range(c, r->dev, g, lane, r->sent[lane], *counts, coef_entry_bytes(r->hdr.bpp), &bytes);
    if (ret < 0)
        return ret;
    r->sent[lane] = *counts;
    c->h2d_bytes += bytes;
    c->early_h2d_bytes += bytes;
    return 0;
}

extern "C" int vp9cuda_abort_frame(Vp9CudaCtx *c, uint32_t frame_id)
{
    if (!c)
        return -EINVAL;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    RingSet *r = ring_of(c, frame_id);
    if (!r)
        return -EINVAL;
    abort_set(c, r);
    return 0;
}

extern "C" int vp9cuda_finish_frame(Vp9CudaCtx *c, uint32_t frame_id, uint8_t *const out_planes[3],
                                    const ptrdiff_t out_strides[3], uint64_t *ticket)
{
    if (!c)
        return -EINVAL;
    HostTimer ht(c->ns_finish);
    cudaSetDevice(c->device);
    RingSet *r;
    {
        std::unique_lock<std::mutex> lk(c->mu);
        r = ring_of(c, frame_id);
        if (!r)
            return -EINVAL;
        /* decode order: every frame begun before this one has been finished or aborted */
        const uint64_t seq = r->seq;
        c->cv.wait(lk, [&]() { return c->next_finish_seq == seq || !r->busy || r->seq != seq; });
        if (!r->busy || r->seq != seq)
            return -EINVAL;      /* flushed / aborted while it waited for its turn */
    }
    int ret;
    {
        std::lock_guard<std::mutex> gk(c->gpu_mu);
        {
            std::lock_guard<std::mutex> rk(r->mu);
            const int ceb = coef_entry_bytes(r->hdr.bpp);
            uint64_t bytes = 0;
            ret = 0;
            kstat_begin(c, c->s_up);
            for (int l = 0; l < r->hdr.n_lanes && ret == 0; l++) {
                ret = copy_lane_range(c, r->dev, r->ring[l], l, r->sent[l], r->counts[l], ceb, &bytes);
                r->sent[l] = r->counts[l];
            }
            if (ret == 0)
                ret = upload_frame_tail(c, r->dev, &r->hdr, r->n_sb, r->counts, (const Vp9CudaSbIndex *)r->sb_index.p,
                                        (const Vp9CudaLoopFilterSb *)r->lf.p, &bytes);
            kstat_end(c, KS_H2D, 0, bytes, c->s_up);
            c->h2d_bytes += bytes;
            cudaEventRecord(r->consumed, c->s_up);
        }
        if (ret == 0)
            ret = reconstruct(c, r->dev, out_planes, out_strides, ticket);
    }
    std::lock_guard<std::mutex> lk(c->mu);
    r->busy = false;
    advance_turn(c);
    return ret;
}

extern "C" int vp9cuda_wait(Vp9CudaCtx *c, uint64_t ticket)
{
    if (!c || !ticket || ticket >= c->next_ticket)
        return -EINVAL;
    HostTimer ht(c->ns_wait);
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
    if (c->err_host && *(volatile int *)c->err_host) {
        /* a wavefront wait ran out of its budget (dev_common.cuh: VP9_SPIN): the pictures from here on are not
         * valid; the context stays in error until vp9cuda_flush */
        fprintf(stderr, "vp9cuda: a wavefront dependency was never satisfied (the wait at source line %d gave up)\n",
                *(volatile int *)c->err_host);
        return -EIO;
    }
    if (c->trace)
        fprintf(stderr, "vp9cuda: ticket %llu done\n", (unsigned long long)ticket);
    return 0;
}

extern "C" int vp9cuda_show_existing(Vp9CudaCtx *c, int slot, uint8_t *const out_planes[3],
                                     const ptrdiff_t out_strides[3], uint64_t *ticket)
{
    if (!c || slot < 0 || slot > 7 || !out_planes)
        return -EINVAL;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lk(c->gpu_mu);
    if (!c->ref[slot])
        return -EINVAL;
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
        std::lock_guard<std::mutex> lk(c->gpu_mu);
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
    out->ss_h = out->ss_v = s->ss;
    return 0;
}

extern "C" int vp9cuda_unmap_frame(Vp9CudaCtx *c, const void *plane0)
{
    if (!c || !plane0)
        return -EINVAL;
    std::lock_guard<std::mutex> lk(c->gpu_mu);
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
    {
        /* one-shot frames take their place in decode order at once: no begun frame may be open */
        std::lock_guard<std::mutex> lk(c->mu);
        for (RingSet *r : c->rings)
            if (r->busy)
                return -EBUSY;
    }
    std::lock_guard<std::mutex> gk(c->gpu_mu);
    /* landing zones of their own (three, round robin), so that the upload of one frame runs beside
     * the kernels of the one before */
    if (c->oneshot.empty())
        c->oneshot.assign(3, nullptr);
    DevRecords *&d = c->oneshot[c->oneshot_rr++ % c->oneshot.size()];
    if (!d)
        d = new DevRecords;
    int ret = upload_records(c, *d, hdr, recs);
    if (ret < 0)
        return ret;
    return reconstruct(c, *d, out_planes, out_strides, ticket);
}

extern "C" int vp9cuda_stage_frame(Vp9CudaCtx *c, const Vp9CudaFrameHeader *hdr, const Vp9CudaFrameRecords *recs,
                                   int *handle)
{
    if (!c || !hdr || !recs || !handle)
        return -EINVAL;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> gk(c->gpu_mu);
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
    std::lock_guard<std::mutex> gk(c->gpu_mu);
    return reconstruct(c, *c->staged[handle], out_planes, out_strides, ticket);
}

extern "C" int vp9cuda_free_staged(Vp9CudaCtx *c, int handle)
{
    if (!c || handle < 0 || handle >= (int)c->staged.size() || !c->staged[handle])
        return -EINVAL;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> gk(c->gpu_mu);
    sync_all(c);
    if (c->prev_dev == c->staged[handle])
        c->prev_dev = nullptr;
    c->staged[handle]->release();
    delete c->staged[handle];
    c->staged[handle] = nullptr;
    return 0;
}

extern "C" int vp9cuda_get_stat(Vp9CudaCtx *c, const char *key, uint64_t *value)
{
    if (!c || !key || !value)
        return -EINVAL;
    if (!strcmp(key, "allocs")) *value = c->n_allocs;
    else if (!strcmp(key, "frames")) *value = c->n_frames;
    else if (!strcmp(key, "aborted")) *value = c->n_aborted;
    else if (!strcmp(key, "ns_begin_frame")) *value = c->ns_begin;
    else if (!strcmp(key, "ns_submit_tile")) *value = c->ns_submit;
    else if (!strcmp(key, "ns_finish_frame")) *value = c->ns_finish;
    else if (!strcmp(key, "ns_wait")) *value = c->ns_wait;
    else if (!strcmp(key, "h2d_bytes")) *value = c->h2d_bytes;
    else if (!strcmp(key, "early_h2d_bytes")) *value = c->early_h2d_bytes;
    else if (!strcmp(key, "d2h_bytes")) *value = c->d2h_bytes;
    else return -EINVAL;
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
    if (!strcmp(key, "inject_stall")) {
        c->inject_stall = value != 0;
        return 0;
    }
    if (!strcmp(key, "inter_sb")) {        /* 1 (default): k_inter_sb; 0: k_inter + k_itx */
        c->inter_sb = value != 0;
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
        if (c->phase_clk.reserve(64 * 8, c->n_allocs))
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
    std::lock_guard<std::mutex> lk(c->gpu_mu);
    Surface *s = slot < 0 ? c->last : c->ref[slot];
    if (!s)
        return -EINVAL;
    if (c->sum_buf.reserve(8, c->n_allocs))
        return -ENOMEM;
    CK(cudaMemsetAsync(c->sum_buf.p, 0, 8, c->stream));
    for (int p = 0; p < 3; p++) {
        int sh = p ? s->ss : 0, w = (s->w + sh) >> sh, h = (s->h + sh) >> sh;
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

/* the frame kernel's own staging / filtering code (mc_batch, mc_batch_scaled of k_inter_sb.cuh), one warp per item */
template <class P>
__global__ void k_slot_mc(int n, const uint8_t *w, const uint8_t *h, const uint8_t *filter, const uint8_t *avg,
                          const uint8_t *mx, const uint8_t *my, const uint8_t *dx, const uint8_t *dy, P *pix,
                          const P *ref, int bpp)
{
    __shared__ __align__(16) unsigned char arena[ISB_ARENA];
    __shared__ UGeom geom[1];
    const int i = blockIdx.x;
    if (i >= n)
        return;
    const int W = w[i], H = h[i], pw = min(W, 32), ph = min(H, 16);
    const int wl = pw == 4 ? 0 : pw == 8 ? 1 : pw == 16 ? 2 : 3, hl = ph == 4 ? 0 : ph == 8 ? 1 : 2;
    const bool scaled = dx[i] != 0;
    for (int py = 0; py < H; py += ph)
        for (int px = 0; px < W; px += pw) {
            if (threadIdx.x == 0) {
                UGeom &g = geom[0];
                g.src = ref + (size_t)i * 160 * 160; g.stride = 160; g.rw = g.rh = 160;
                g.filter = filter[i];
                g.dst = (uint16_t)(py * 64 + px); g.dpitch = 64;
                g.pos.x0 = 8; g.pos.y0 = 8; g.pos.fx = mx[i]; g.pos.fy = my[i];
                g.pos.dx = scaled ? dx[i] : 16; g.pos.dy = scaled ? dy[i] : 16;
                g.sox = px; g.soy = py;
                g.fx = mx[i]; g.fy = my[i];
                g.wx = 8 + px - 3; g.wy = 8 + py - 3;
                g.interior = 1;
                g.a = (uint8_t)(g.wx & (IsbGeo<P>::APX - 1));
            }
            __syncwarp();
            if (scaled)
                mc_batch_scaled<P>(geom, 1, wl, hl, pix + (size_t)i * 4096, avg[i] != 0, bpp);
            else
                mc_batch<P>(geom, 1, wl, hl, arena, pix + (size_t)i * 4096, avg[i] != 0, bpp);
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
    if (bpp > 8) k_slot_mc<uint16_t><<<n, 32>>>(n, dw, dh, df, da, dmx, dmy, ddx, ddy, (uint16_t *)dp, (const uint16_t *)dr, bpp);
    else         k_slot_mc<uint8_t><<<n, 32>>>(n, dw, dh, df, da, dmx, dmy, ddx, ddy, dp, dr, bpp);
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

/* ------------------------------------------------------------------------------------------
 * vp9cuda_checksum_noalloc: the frame checksum of vp9cuda_frame_checksum for callers that must not
 * allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in
 * the reference.
 *
 * The kernel's scratch -- one partial sum per picture row and a ticket the CTAs take when they are
 * done -- lives in a file-scope __device__ array of VP9CUDA_NOALLOC_MAX_ROWS rows (128 KiB + 256 B
 * on every GPU that loads the library, until exit).  g_noalloc_scratch is a host-side stand-in in
 * host code, not a device address: the address the kernel gets comes from cudaGetSymbolAddress, once
 * per device, with that device current.  Every call shares the array: calls that overlap on the GPU
 * (different streams, different host threads) race; calls on one stream do not.
 * ---------------------------------------------------------------------------------------- */
#define NOALLOC_TICKET VP9CUDA_NOALLOC_MAX_ROWS       /* index of the ticket word behind the row sums */
static __device__ __align__(256) unsigned long long g_noalloc_scratch[VP9CUDA_NOALLOC_MAX_ROWS + 32];

struct NoAllocPlanes {
    const void *pl[3];
    long long stride[3];               /* pixels */
    int w[3], h[3];
};

__device__ __forceinline__ unsigned long long noalloc_block_sum(unsigned long long acc, unsigned long long *part)
{
    for (int o = 16; o; o >>= 1)
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __syncthreads();                                   /* part[] of the previous round has been read */
    if ((threadIdx.x & 31) == 0)
        part[threadIdx.x >> 5] = acc;
    __syncthreads();
    unsigned long long s = 0;
    for (int i = 0; i < (int)(blockDim.x >> 5); i++)
        s += part[i];
    return s;
}

/* same sum as k_checksum over the three planes (salt = plane << 24; the additions commute), taken row by row:
 * a CTA leaves the sum of each of its rows in scratch[row]; the CTA that takes the last ticket adds the rows up */
template <class P>
__global__ void k_checksum_rows(NoAllocPlanes a, unsigned long long *scratch, unsigned long long *out)
{
    __shared__ unsigned long long part[32];
    __shared__ int last;
    const int total = a.h[0] + a.h[1] + a.h[2];
    for (int r = blockIdx.x; r < total; r += gridDim.x) {
        const int p = r < a.h[0] ? 0 : r < a.h[0] + a.h[1] ? 1 : 2;
        const int y = r - (p ? a.h[0] : 0) - (p > 1 ? a.h[1] : 0);
        const P *row = (const P *)a.pl[p] + (size_t)y * a.stride[p];
        const int w = a.w[p];
        const unsigned salt = (unsigned)p * 0x01000000u;
        unsigned long long acc = 0;
        for (int x = threadIdx.x; x < w; x += blockDim.x) {
            const int i = y * w + x;
            const unsigned long long v = row[x];
            acc += (v + 1) * (0x9E3779B97F4A7C15ull ^ ((unsigned long long)(i + salt) * 0xD6E8FEB86659FD93ull));
        }
        const unsigned long long s = noalloc_block_sum(acc, part);
        if (threadIdx.x == 0)
            scratch[r] = s;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0)
        last = atomicAdd((unsigned *)&scratch[NOALLOC_TICKET], 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last)
        return;
    __threadfence();
    unsigned long long acc = 0;
    for (int r = threadIdx.x; r < total; r += blockDim.x)
        acc += __ldcg(&scratch[r]);
    const unsigned long long s = noalloc_block_sum(acc, part);
    if (threadIdx.x == 0) {
        *out = s;
        scratch[NOALLOC_TICKET] = 0;                   /* the next call starts from zero even without its own clear */
    }
}

#define NOALLOC_MAX_DEVICES 64
static std::mutex g_noalloc_mu;
static unsigned long long *g_noalloc_ptr[NOALLOC_MAX_DEVICES];      /* device address of g_noalloc_scratch, per device */

extern "C" int vp9cuda_checksum_noalloc(int device, const void *const planes[3], const ptrdiff_t strides[3],
                                        int w, int h, int bpp, int ss_h, int ss_v, uint64_t *sum_dev, void *stream)
{
    if (!planes || !strides || !sum_dev || !planes[0] || !planes[1] || !planes[2] || w < 1 || h < 1 || w > 65536 ||
        (bpp != 8 && bpp != 10 && bpp != 12) || (ss_h & ~1) || (ss_v & ~1) || device < 0 || device >= NOALLOC_MAX_DEVICES)
        return -EINVAL;
    const int pb = bpp > 8 ? 2 : 1;
    NoAllocPlanes a;
    long long rows = 0;
    for (int p = 0; p < 3; p++) {
        a.pl[p] = planes[p];
        if (strides[p] % pb)
            return -EINVAL;
        a.stride[p] = strides[p] / pb;
        a.w[p] = p ? (w + ss_h) >> ss_h : w;
        a.h[p] = p ? (h + ss_v) >> ss_v : h;
        rows += a.h[p];
    }
    if (rows > VP9CUDA_NOALLOC_MAX_ROWS) {
        fprintf(stderr, "vp9cuda_checksum_noalloc: a %dx%d picture has %lld rows in its three planes, the static scratch "
                        "holds VP9CUDA_NOALLOC_MAX_ROWS = %d\n", w, h, rows, VP9CUDA_NOALLOC_MAX_ROWS);
        return -E2BIG;
    }
    CK(cudaSetDevice(device));
    unsigned long long *scratch;
    {
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_ptr[device])                    /* once per device, with that device current */
            CK(cudaGetSymbolAddress((void **)&g_noalloc_ptr[device], g_noalloc_scratch));
        scratch = g_noalloc_ptr[device];
    }
    cudaStream_t st = (cudaStream_t)stream;
    CK(cudaMemsetAsync(scratch + NOALLOC_TICKET, 0, 256, st));      /* the ticket must start at zero */
    const int grid = (int)(rows < 592 ? rows : 592);
    if (bpp > 8) k_checksum_rows<uint16_t><<<grid, 256, 0, st>>>(a, scratch, (unsigned long long *)sum_dev);
    else         k_checksum_rows<uint8_t><<<grid, 256, 0, st>>>(a, scratch, (unsigned long long *)sum_dev);
    CK(cudaGetLastError());
    return 0;
}
