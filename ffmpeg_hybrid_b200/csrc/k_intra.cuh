/*
 * k_intra.cuh -- K3: intra prediction + residual as a superblock-row wavefront whose dependent
 * chains run entirely in shared memory.
 *
 * Every intra transform block depends on the reconstructed pixels left of, above and (4x4 blocks
 * inside their coding block only) above-right of it (check_intra_mode, vp9recon.c:37-221).  All
 * three stay inside the block's own superblock, its left neighbour and the superblock row above
 * up to the same column, so superblock (r, c) may run once (r, c-1) and (r-1, c) are done: a
 * wavefront with a lag of one superblock per row.
 *
 *   k_intra_prep : one CTA per superblock.  Sorts the superblock's records (all three planes)
 *       by the host-computed dependency depth (Vp9CudaIntraTx.level; equal depth = mutually
 *       independent) and writes them as compact 16-byte IntraRec with run lengths.
 *   k_intra_res : the inverse transforms do not depend on any neighbour, so they leave the
 *       dependency chain: one warp per coded block writes the rounded residual (int16, clamped
 *       to [-2^bpp, 2^bpp], which keeps clip(pred + residual) exact) into a per-superblock
 *       residual tile.  Fully parallel.
 *   k_intra : one CTA per superblock ROW.  The row's pixels slide through a shared-memory ring
 *       of four superblock columns per plane (row 0 of each ring = the pixel row above).  A
 *       body-loader warp streams tiles, records and residual tiles in with cp.async up to two
 *       turns ahead; a top-loader warp brings in the border line above -- as tagged words
 *       (pixels | launch tag, FrameDesc.intra_ll: data and "ready" in one 8-byte load) when the
 *       superblocks above went through the wavefront, else from the picture behind the row
 *       above's progress counter; twenty compute warps run the turn's dependency steps (edges ->
 *       prediction + residual, all in shared memory, one barrier per step); a storer warp sends
 *       the bottom line out as tagged words, then writes the finished tile back, bottom pixel
 *       row first, then fence + progress counter.  Superblocks without intra blocks cost one
 *       flag update.
 *
 * Replaces intra_recon + check_intra_mode (vp9recon.c:37-364) and the intra_pred slots
 * (vp9dsp_template.c:33-1153).  Reads the pre-loop-filter picture everywhere, which is what
 * the reference gets from intra_pred_data at superblock-row tops (vp9.c:1404-1416).
 * Algorithmic bytes: P_intra (write) + edges + coefficients.
 */
#ifndef VP9CUDA_K_INTRA_CUH
#define VP9CUDA_K_INTRA_CUH

#include "dev_common.cuh"
#include "k_itx.cuh"

namespace vp9 {

#ifndef INTRA_CW
#define INTRA_CW 20                         /* compute warps */
#endif
#define INTRA_CT (INTRA_CW * 32)
#define INTRA_THREADS (INTRA_CT + 3 * 32)   /* + body loader, top loader, storer */
#define INTRA_MAX_SB_COLS 256
#define INTRA_SB_RECS 384                   /* 256 luma + 2 x 64 chroma 4x4 blocks */
#define INTRA_SB_RES 6144                   /* residual tile: 64*64 + 2 * 32*32 int16 */
#define INTRA_SB_TAB 768                    /* work table: <= 384 chunk entries + <= 384 step headers */
#define INTRA_EDGE_N 104                    /* per compute warp: top[0..64] (0 = top-left) at 0, left[0..31] at 72 */

/* residual tile of one superblock: luma 64x64 row-major, then U 32x32, then V 32x32 (int16) */
__device__ __forceinline__ int res_plane_off(int pl) { return pl == 0 ? 0 : pl == 1 ? 4096 : 5120; }

/* 32-pixel chunks a block is dealt out in (a 4x4 block is one half-filled chunk) */
__device__ __forceinline__ int intra_chunks(int tx) { return tx == 0 ? 1 : 2 << (2 * (tx - 1)); }

/* ---------------------------------------------------------------------------------------------- */
/* Pre-pass, one CTA per superblock.
 *  1. sort the records by the host's frame-wide depth (a topological order);
 *  2. walk that order and compute each block's depth INSIDE the superblock, counting only the
 *     neighbours its (edge-converted) mode really reads -- left column / top-left / top row /
 *     top-right cell, the read set of check_intra_mode (vp9recon.c:72-221) -- and only cells that
 *     hold intra blocks of this superblock: everything else is final before the turn starts;
 *  3. sort by that local depth, emit IntraRec in that order and the work table: per step a header
 *     (0x8000 | chunk count) followed by one entry per 32-pixel chunk (record | chunk << 9). */
__global__ void __launch_bounds__(128) k_intra_prep(const __grid_constant__ FrameDesc f)
{
    __shared__ uint32_t keys[1024];
    __shared__ uint32_t inf[768];
    __shared__ uint8_t loc[768];
    __shared__ int n_match;
    __shared__ uint8_t lvl[3 * 256];
    __shared__ unsigned any_res, need_top;
    const int sb = blockIdx.x, tid = threadIdx.x;
    const Vp9CudaSbIndex si = f.sb_index[sb];
    /* a pass over one plane of a 4:4:4 frame (f.mono) sees the records of all three planes (up to 768) and keeps
     * its own (up to 256), as plane 0 */
    const int ntot = si.itx_count;
    int n = ntot;
    const uint32_t g0 = n ? f.lane_itx_base[si.lane] + si.itx_start : 0;
    if (n == 0 || n > (f.mono >= 0 ? 2 * INTRA_SB_RECS : INTRA_SB_RECS)) {
        if (tid == 0) {
            IntraSb o = { g0, 0, 0, 0, 0, 0 };
            f.isb[sb] = o;
        }
        return;
    }
    if (tid == 0) { any_res = 0; need_top = 0; n_match = 0; }
    int m = 32;
    while (m < n) m <<= 1;
    const int sbx = sb % f.sb_cols, sby = sb / f.sb_cols;
    for (int i = tid; i < 3 * 256; i += 128) lvl[i] = 0xff;
    __syncthreads();
    for (int i = tid; i < m; i += 128) {
        uint32_t key = 0xffffffffu;
        if (i < n && (f.mono < 0 || f.itx[g0 + i].plane == f.mono)) {
            Vp9CudaIntraTx r = f.itx[g0 + i];
            if (f.mono >= 0)
                r.plane = 0;
            atomicAdd(&n_match, 1);
            key = (uint32_t)r.level << 10 | i;
            const int ch = r.plane > 0;
            const int lx = r.x4 - sbx * (16 >> ch), ly = r.y4 - sby * (16 >> ch);
            const bool have_top = r.y4 > 0, have_left = r.flags & VP9CUDA_ITX_HAVE_LEFT;
            const int cmode = intra_mode_conv(r.mode, have_left, have_top);
            inf[i] = lx | ly << 4 | r.plane << 8 | r.tx << 10 | cmode << 12 | (r.flags & 3) << 16 |
                     (r.eob ? 1 : 0) << 18 | (have_top ? 1 : 0) << 19;
            if (r.eob)
                any_res = 1;
            /* Does this block read pixels of an INTRA block of the superblock row above?  Only then
             * must the row's wavefront wait for the row above; inter pixels are final already.
             * Cells read: from the top-left sample to the end of a possible top-right extension. */
            if (ly == 0 && have_top && (intra_mode_needs(cmode) & 14)) {
                if (f.n_blk == 0) {
                    need_top = 1;
                } else {
                    const int px0 = r.x4 * 4 - 1, px1 = r.x4 * 4 + (8 << r.tx) - 1;
                    const int c0 = max(ch ? px0 >> 2 : px0 >> 3, 0), c1 = min(ch ? px1 >> 2 : px1 >> 3, f.cols - 1);
                    const int32_t *above = f.cellmap + (size_t)(sby * 8 - 1) * f.cols;
                    for (int cx = c0; cx <= c1; cx++)
                        if (above[cx] < 0)
                            need_top = 1;
                }
            }
        }
        keys[i] = key;
    }
    __syncthreads();
    auto sort_keys = [&]() {
        for (int k = 2; k <= m; k <<= 1)
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int i = tid; i < m; i += 128) {
                    const int p = i ^ j;
                    if (p > i) {
                        const uint32_t a = keys[i], b = keys[p];
                        if ((a > b) == ((i & k) == 0)) { keys[i] = b; keys[p] = a; }
                    }
                }
                __syncthreads();
            }
    };
    sort_keys();
    n = n_match;                               /* the records that are not this pass's sort behind them */
    if (n == 0) {
        if (tid == 0) {
            IntraSb o = { g0, 0, 0, 0, 0, 0 };
            f.isb[sb] = o;
        }
        return;
    }
    /* local depth, one frame-wide depth value (= set of mutually independent blocks) at a time */
    for (int pos = 0; pos < n;) {
        const uint32_t L = keys[pos] >> 10;
        int end = pos + 1;
        while (end < n && (keys[end] >> 10) == L) end++;
        int mine[3], nm = 0;
        for (int k = pos + tid; k < end && nm < 3; k += 128) {
            const int i = keys[k] & 1023;
            const uint32_t v = inf[i];
            const int lx = v & 15, ly = v >> 4 & 15, pl = v >> 8 & 3, tx = v >> 10 & 3, cm = v >> 12 & 15;
            const bool have_left = v >> 16 & 1, have_right = v >> 17 & 1, have_top = v >> 19 & 1;
            const int needs = intra_mode_needs(cm), nb = 1 << tx, W = pl ? 8 : 16;
            int mx = -1;
            auto dep = [&](int x, int y) {
                if (x >= 0 && y >= 0 && x < W && y < W) {
                    const int d = lvl[pl * 256 + y * 16 + x];
                    if (d != 0xff && d > mx) mx = d;
                }
            };
            if (have_left) {
                if (needs & 1)
                    for (int j = 0; j < nb; j++) dep(lx - 1, ly + j);
                if ((needs & 4) && have_top)
                    dep(lx - 1, ly - 1);
            }
            if (have_top) {
                if (needs & 2)
                    for (int j = 0; j < nb; j++) dep(lx + j, ly - 1);
                if ((needs & 8) && tx == 0 && have_right)
                    dep(lx + 1, ly - 1);
            }
            loc[i] = (uint8_t)(mx + 1);
            mine[nm++] = i;
        }
        __syncthreads();
        for (int q = 0; q < nm; q++) {
            const uint32_t v = inf[mine[q]];
            const int lx = v & 15, ly = v >> 4 & 15, pl = v >> 8 & 3, nb = 1 << (v >> 10 & 3);
            for (int y = 0; y < nb; y++)
                for (int x = 0; x < nb; x++)
                    lvl[pl * 256 + (ly + y) * 16 + lx + x] = loc[mine[q]];
        }
        __syncthreads();
        pos = end;
    }
    for (int k = tid; k < n; k += 128) {
        const int i = keys[k] & 1023;
        keys[k] = (uint32_t)loc[i] << 10 | i;
    }
    __syncthreads();
    sort_keys();
    for (int k = tid; k < n; k += 128) {
        const int i = keys[k] & 1023;
        const uint32_t v = inf[i];
        Vp9CudaIntraTx r = f.itx[g0 + i];
        if (f.mono >= 0)
            r.plane = 0;
        IntraRec o;
        o.lx = v & 15; o.ly = v >> 4 & 15;
        o.plane = r.plane; o.tx = r.tx; o.mode = v >> 12 & 15; o.txtp = r.txtp;
        o.flags = (v >> 16 & 3) | (v >> 19 & 1) << 2;
        o.pad = 0;
        o.eob = r.eob;
        o.ncoef = r.ncoef;
        o.coef_off = r.coef_off;
        f.irec[g0 + k] = o;
        /* the wavefront's view of the block: addresses in its shared-memory rings (layout: IntraSmem) */
        {
            const int ps = f.bpp > 8 ? 2 : 1, rsby = 256 * ps + 16, rsbc = 128 * ps + 16;
            const int pl = r.plane, ch = pl > 0, S = 64 >> ch, N = 4 << r.tx, rsb = ch ? rsbc : rsby;
            const int o_pl = pl == 0 ? 0 : pl == 1 ? 65 * rsby : 65 * rsby + 33 * rsbc;
            const int bx = (sbx & 3) * S + o.lx * 4, by = 1 + o.ly * 4, xmask = 4 * S - 1;
            const bool have_left = o.flags & 1, have_right = o.flags & 2, have_top = o.flags & 4;
            const int avail_w = ((f.cols * 8) >> ch) - (sbx * S + o.lx * 4);
            const int avail_h = ((f.rows * 8) >> ch) - (sby * S + o.ly * 4);
            const int needs = intra_mode_needs(o.mode);
            IntraBlk b;
            b.dst = o_pl + by * rsb + bx * ps;
            b.top = o_pl + (by - 1) * rsb + bx * ps;
            b.tl = o_pl + (by - 1) * rsb + ((bx - 1) & xmask) * ps;
            b.left = o_pl + by * rsb + ((bx - 1) & xmask) * ps;
            b.res = (uint16_t)(res_plane_off(pl) + o.ly * 4 * S + o.lx * 4);
            b.mode = o.mode;
            b.lg = (uint8_t)(2 + r.tx);
            /* beyond the block: the real top-right only for 4x4 blocks that may read it, else replicate */
            b.klim = (uint8_t)min((N == 4 && (needs & 8) && have_right && avail_w >= 8) ? 7 : N - 1, avail_w - 1);
            b.hlim = (uint8_t)min(N - 1, avail_h - 1);
            b.flags = (uint8_t)((have_left ? 1 : 0) | (have_top ? 2 : 0) | (r.eob ? 4 : 0) | ((have_left && have_top) ? 8 : 0) |
                                (ch ? 16 : 0));
            for (int q = 0; q < 9; q++) b.pad[q] = 0;
            f.iblk[g0 + k] = b;
        }
    }
    /* work table: thread t owns sorted records 4t .. 4t+3; two block-wide exclusive scans give every
     * record the chunks and the step headers in front of it */
    {
        __shared__ int wsum[2][4];
        __shared__ int scnt[INTRA_SB_RECS];
        for (int i = tid; i < INTRA_SB_RECS; i += 128) scnt[i] = 0;
        int nch[4], first[4], tc = 0, ts = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int k = 4 * tid + j;
            nch[j] = 0; first[j] = 0;
            if (k < n) {
                nch[j] = intra_chunks(inf[keys[k] & 1023] >> 10 & 3);
                first[j] = k == 0 || (keys[k - 1] >> 10) != (keys[k] >> 10);
            }
            tc += nch[j]; ts += first[j];
        }
        int ic = tc, is = ts;
        const int lane = tid & 31, w = tid >> 5;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int a = __shfl_up_sync(0xffffffffu, ic, o), b2 = __shfl_up_sync(0xffffffffu, is, o);
            if (lane >= o) { ic += a; is += b2; }
        }
        if (lane == 31) { wsum[0][w] = ic; wsum[1][w] = is; }
        __syncthreads();
        int bc = 0, bs = 0, totc = 0, tots = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            if (q < w) { bc += wsum[0][q]; bs += wsum[1][q]; }
            totc += wsum[0][q]; tots += wsum[1][q];
        }
        int cpos = bc + ic - tc, sidx = bs + is - ts;       /* exclusive prefixes at record 4t */
        uint16_t *tab = f.itab + (size_t)sb * INTRA_SB_TAB;
        int hdr_pos[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int k = 4 * tid + j;
            hdr_pos[j] = -1;
            if (k < n) {
                sidx += first[j];
                if (first[j]) hdr_pos[j] = cpos + sidx - 1;
                for (int q = 0; q < nch[j]; q++)
                    tab[cpos + sidx + q] = (uint16_t)(k | q << 9);
                atomicAdd(&scnt[sidx - 1], nch[j]);
                cpos += nch[j];
            }
        }
        __syncthreads();
        sidx = bs + is - ts;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            sidx += first[j];
            if (hdr_pos[j] >= 0)
                tab[hdr_pos[j]] = (uint16_t)(0x8000 | scnt[sidx - 1]);
        }
        if (tid == 0) {
            IntraSb o = { g0, (uint16_t)n, (uint16_t)(totc + tots), (uint16_t)any_res, (uint16_t)need_top, 0 };
            f.isb[sb] = o;
        }
    }
}

/* ---------------------------------------------------------------------------------------------- */

struct ResBlk {
    int16_t *p; int stride;
    __device__ __forceinline__ int get(int, int) const { return 0; }
    __device__ __forceinline__ void put(int x, int y, int v) { p[y * stride + x] = (int16_t)v; }
};

#define IRES_WARPS 4
template <bool HBD>
__global__ void __launch_bounds__(IRES_WARPS * 32) k_intra_res(const __grid_constant__ FrameDesc f)
{
    typedef typename ItxTypes<HBD>::C C;
    typedef typename ItxTypes<HBD>::E E;
    __shared__ __align__(16) C dense_s[IRES_WARPS][32 * 32];
    __shared__ C tmp_s[IRES_WARPS][32 * 33];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    C *tmp = tmp_s[warp], *dense = dense_s[warp];
    /* one warp per (superblock, record); superblock = blockIdx.y keeps the residual tile address cheap */
    const int sb = blockIdx.x;
    const IntraSb si = f.isb[sb];
    if (!si.has_res)
        return;
    int16_t *tile = f.ires + (size_t)sb * INTRA_SB_RES;
    for (int k = warp; k < si.count; k += IRES_WARPS) {
        const IntraRec r = f.irec[si.start + k];
        if (!r.eob)
            continue;
        const int S = r.plane ? 32 : 64;
        ResBlk blk = { tile + res_plane_off(r.plane) + r.ly * 4 * S + r.lx * 4, S };
        const E *ent = (const E *)f.coef + r.coef_off;
        switch (r.tx) {
        case 0: if (lane < 4)  itx_entries<HBD, 4, ResBlk, true>(ent, r.ncoef, dense, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xfu); break;
        case 1: if (lane < 8)  itx_entries<HBD, 8, ResBlk, true>(ent, r.ncoef, dense, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xffu); break;
        case 2: if (lane < 16) itx_entries<HBD, 16, ResBlk, true>(ent, r.ncoef, dense, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xffffu); break;
        default: itx_entries<HBD, 32, ResBlk, true>(ent, r.ncoef, dense, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xffffffffu); break;
        }
        __syncwarp();
    }
}

/* ---------------------------------------------------------------------------------------------- */
struct IntraSig {
    volatile int body;     /* tiles / records / residuals of superblocks < body are in shared memory */
    volatile int top;      /* border rows of superblocks < top are in shared memory */
    volatile int done;     /* turns < done are complete */
    volatile int stored;   /* superblocks < stored have been written back */
};

/* a plane's ring in shared memory, addressed with ring-column / ring-row coordinates */
template <class P> struct TilePx {
    unsigned char *t; int rsb, mask;
    __device__ __forceinline__ int operator()(int x, int y) const { return *(const P *)(t + y * rsb + (x & mask) * (int)sizeof(P)); }
};

template <class P> struct IntraSmem {
    static constexpr int RSBY = 4 * 64 * (int)sizeof(P) + 16;      /* luma ring row stride, bytes */
    static constexpr int RSBC = 4 * 32 * (int)sizeof(P) + 16;
    static constexpr int O_Y = 0;
    static constexpr int O_U = O_Y + 65 * RSBY;
    static constexpr int O_V = O_U + 33 * RSBC;
    static constexpr int O_RES = O_V + 33 * RSBC;                  /* [4][INTRA_SB_RES] int16 */
    static constexpr int O_REC = O_RES + 4 * INTRA_SB_RES * 2;     /* [4][INTRA_SB_RECS] IntraBlk */
    static constexpr int O_TAB = O_REC + 4 * INTRA_SB_RECS * (int)sizeof(IntraBlk);   /* [4][INTRA_SB_TAB] u16 */
    static constexpr int O_INFO = O_TAB + 4 * INTRA_SB_TAB * 2;
    static constexpr int O_EDGE = O_INFO + INTRA_MAX_SB_COLS * (int)sizeof(IntraSb);   /* [INTRA_CW][INTRA_EDGE_N] int */
    static constexpr int BYTES = O_EDGE + INTRA_CW * INTRA_EDGE_N * 4;
};

#define ICLK(k) do { if (f.phase_clk) { long long c_ = clock64(); \
    atomicAdd(&f.phase_clk[k], (unsigned long long)(c_ - clk_)); clk_ = c_; } } while (0)

/* one superblock row of the intra wavefront; runs on all INTRA_THREADS threads of the CTA */
template <class P>
__device__ __forceinline__ void intra_row(const FrameDesc &f, const int row, unsigned char *sm, IntraSig &sig)
{
    typedef IntraSmem<P> L;
    auto tile = [&](int pl) -> unsigned char * { return sm + (pl == 0 ? L::O_Y : pl == 1 ? L::O_U : L::O_V); };
    int16_t *const resbuf = (int16_t *)(sm + L::O_RES);
    IntraBlk *const recbuf = (IntraBlk *)(sm + L::O_REC);
    uint16_t *const tabbuf = (uint16_t *)(sm + L::O_TAB);
    const IntraSb *const info = (const IntraSb *)(sm + L::O_INFO);
    const int ncol = f.sb_cols;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool has_top = row > 0;
    int *const ifull = f.iprogress + f.sb_rows;      /* [row]: superblocks < ifull are completely written back */

    if (threadIdx.x == 0) { sig.body = 0; sig.top = 0; sig.done = 0; sig.stored = 0; }
    for (int i = threadIdx.x; i < ncol * 4; i += INTRA_THREADS)
        ((uint32_t *)(sm + L::O_INFO))[i] = ((const uint32_t *)(f.isb + (size_t)row * ncol))[i];
    __syncthreads();

    if (warp < INTRA_CW) {
        /* ------------------------------- compute warps ------------------------------- */
        for (int c = 0; c < ncol; c++) {
            const int n = info[c].count;
            if (n == 0) {
                if (threadIdx.x == 0)
                    sig.done = c + 1;
                continue;
            }
            long long clk_ = 0;
            if (f.phase_clk && threadIdx.x == 0) clk_ = clock64();
            if (threadIdx.x == 0) {
                VP9_SPIN(sig.body < c + 1, VP9_TIGHT_NS, "intra compute: body", row, c);
                ICLK(10);
                if (has_top)
                    VP9_SPIN(sig.top < c + 1, VP9_TIGHT_NS, "intra compute: top", row, c);
                ICLK(11);
                __threadfence_block();
            }
            bar_named<1, INTRA_CT>();
#ifdef VP9CUDA_STEPSTATS   /* counters for tools/prof_frames.py; the chain is so tight that they cost ~4 % of the wavefront */
            int nsteps = 0, nrounds = 0, nchunks = 0, ntop = info[c].needs_top;
#endif
            const IntraBlk *rec = recbuf + (c & 3) * INTRA_SB_RECS;
            const int16_t *res = resbuf + (c & 3) * INTRA_SB_RES;
            const uint16_t *tab = tabbuf + (c & 3) * INTRA_SB_TAB;
            const int ntab = info[c].ntab;
            const int base = 128 << (f.bpp - 8);
            int *const T = (int *)(sm + L::O_EDGE) + warp * INTRA_EDGE_N, *const Lf = T + 72;
            for (int pos = 0; pos < ntab;) {
                const int cnt = tab[pos] & 0x7fff;
                for (int i = warp; i < cnt; i += INTRA_CW) {
                    const unsigned ent = tab[pos + 1 + i];
                    const uint4 w0 = *(const uint4 *)&rec[ent & 511];                       /* dst, top, tl, left */
                    const uint2 w1 = *(const uint2 *)((const unsigned char *)&rec[ent & 511] + 16);
                    const int q = ent >> 9 & 31;
                    const int mode = w1.x >> 16 & 0xff, lg = w1.x >> 24, N = 1 << lg;
                    const int klim = w1.y & 0xff, hlim = w1.y >> 8 & 0xff, flags = w1.y >> 16 & 0xff;
                    const int rsb = (flags & 16) ? L::RSBC : L::RSBY;
                    /* Edge samples once per chunk (rules of check_intra_mode, vp9recon.c:72-221, already turned
                     * into limits by the pre-pass; same values as intra_edges, vp9_pred.h): lane l fetches top[l]
                     * (0 = top-left, 1 + k = sample k above) and left[l] into the warp's own arrays, then the
                     * predictor indexes them freely. */
                    int tv = base - 1, lv = base + 1;
                    if (flags & 2)
                        tv = *(const P *)(sm + w0.y + min(lane - 1, klim) * (int)sizeof(P));
                    if (lane == 0)
                        tv = (flags & 8) ? (int)*(const P *)(sm + w0.z) : base + ((flags & 2) ? 1 : -1);
                    if (flags & 1)
                        lv = *(const P *)(sm + w0.w + min(lane, hlim) * rsb);
                    T[lane] = tv;
                    Lf[lane] = lv;
                    if (N >= 16) {                       /* top[32 ..]: all of them the last sample the block may read */
                        const int rep = (flags & 2) ? (int)*(const P *)(sm + w0.y + klim * (int)sizeof(P)) : base - 1;
                        if (N == 32) T[32 + lane] = rep;
                        if (lane == 0) T[N == 32 ? 64 : 32] = rep;
                    }
                    int dc = 0;
                    if (mode == DC_PRED || mode == LEFT_DC_PRED || mode == TOP_DC_PRED) {
                        /* intra_dc's sums (vp9_pred.h), lane-parallel: lane l holds top[l] and left[l] */
                        int sum = 0;
                        if (mode != LEFT_DC_PRED && lane >= 1 && lane <= min(N, 31)) sum += tv;
                        if (mode != TOP_DC_PRED && lane < N) sum += lv;
                        sum = __reduce_add_sync(0xffffffffu, sum);
                        if (N == 32 && mode != LEFT_DC_PRED)
                            sum += (flags & 2) ? (int)*(const P *)(sm + w0.y + min(31, klim) * (int)sizeof(P)) : base - 1;
                        dc = mode == DC_PRED ? (sum + N) >> (lg + 1) : (sum + (N >> 1)) >> lg;
                    } else if (mode >= DC_128_PRED) {
                        dc = base + (mode == DC_127_PRED ? -1 : mode == DC_129_PRED ? 1 : 0);
                    }
                    __syncwarp();
                    const int p = q * 32 + lane;
                    if (p < N * N) {
                        const int x = p & (N - 1), y = p >> lg;
                        int v = intra_pixel(mode, N, x, y, T, Lf, dc, f.bpp);
                        if (flags & 4)
                            v = clip_pixel(v + res[(w1.x & 0xffff) + (y << ((flags & 16) ? 5 : 6)) + x], f.bpp);
                        *(P *)(sm + w0.x + y * rsb + x * (int)sizeof(P)) = (P)v;
                    }
                    __syncwarp();                        /* the arrays are rewritten by this warp's next chunk */
                }
                bar_named<1, INTRA_CT>();
                pos += cnt + 1;
#ifdef VP9CUDA_STEPSTATS
                nsteps++;
                nrounds += (cnt + INTRA_CW - 1) / INTRA_CW;
                nchunks += cnt;
#endif
            }
            if (threadIdx.x == 0) {
                __threadfence_block();
                sig.done = c + 1;
                ICLK(12);
                if (f.phase_clk) {
                    atomicAdd(&f.phase_clk[14], 1ull);
                    atomicAdd(&f.phase_clk[15], (unsigned long long)n);
#ifdef VP9CUDA_STEPSTATS
                    atomicAdd(&f.phase_clk[13], (unsigned long long)nsteps);
                    atomicAdd(&f.phase_clk[16], (unsigned long long)nrounds);
                    atomicAdd(&f.phase_clk[17], (unsigned long long)nchunks);
                    atomicAdd(&f.phase_clk[18], (unsigned long long)ntop);
#endif
                }
            }
        }
    } else if (warp == INTRA_CW) {
        /* ------------------------------- body loader ------------------------------- */
        for (int j = 0; j < ncol; j++) {
            const IntraSb sj = info[j];
            const bool need_tile = sj.count > 0 || (j + 1 < ncol && info[j + 1].count > 0);
            /* ring slot j & 3 last served superblock j-4 (tile: also the left border of turn j-3) */
            /* lane 0 decides for the warp (see lf_row's body loader) */
            int wait = 0;
            if (lane == 0)
                wait = need_tile && (sig.done < j - 2 || sig.stored < j - 3);
            wait = __shfl_sync(0xffffffffu, wait, 0);
            if (wait) {
                cp_wait<0>();
                __syncwarp();
                if (lane == 0) {
                    __threadfence_block();
                    sig.body = j;
                    VP9_SPIN(sig.done < j - 2 || sig.stored < j - 3, 20, "intra body loader: ring", row, j);
                    __threadfence_block();
                }
                __syncwarp();
            }
            if (need_tile) {
#pragma unroll
                for (int pl = 0; pl < 3; pl++) {
                    if (pl && f.mono >= 0)
                        break;                       /* one plane of a 4:4:4 frame */
                    const int S = pl ? 32 : 64, rsb = pl ? L::RSBC : L::RSBY;
                    const int cpr = S * (int)sizeof(P) / 16;
                    const int stride = f.cur.stride[pl];
                    const unsigned char *src = (const unsigned char *)((const P *)f.cur.pl[pl] + (size_t)row * S * stride + j * S);
                    unsigned char *dst = tile(pl) + rsb + (j & 3) * S * (int)sizeof(P);
                    for (int i = lane; i < S * cpr; i += 32) {
                        const int r = i / cpr, k = i % cpr;
                        cp16(dst + r * rsb + 16 * k, src + (size_t)r * stride * sizeof(P) + 16 * k);
                    }
                }
            }
            if (sj.count > 0) {
                const unsigned char *rs = (const unsigned char *)(f.iblk + sj.start);
                unsigned char *rd = (unsigned char *)(recbuf + (j & 3) * INTRA_SB_RECS);
                for (int i = lane; i < 2 * sj.count; i += 32)
                    cp16(rd + 16 * i, rs + 16 * i);
                const unsigned char *ts = (const unsigned char *)(f.itab + (size_t)(row * ncol + j) * INTRA_SB_TAB);
                unsigned char *td = (unsigned char *)(tabbuf + (j & 3) * INTRA_SB_TAB);
                for (int i = lane; i < (sj.ntab * 2 + 15) / 16; i += 32)
                    cp16(td + 16 * i, ts + 16 * i);
                if (sj.has_res) {
                    const unsigned char *cs = (const unsigned char *)(f.ires + (size_t)(row * ncol + j) * INTRA_SB_RES);
                    unsigned char *cd = (unsigned char *)(resbuf + (j & 3) * INTRA_SB_RES);
                    for (int i = lane; i < INTRA_SB_RES * 2 / 16; i += 32)
                        cp16(cd + 16 * i, cs + 16 * i);
                }
            }
            cp_commit();
            cp_wait<1>();
            __syncwarp();
            if (lane == 0) { __threadfence_block(); sig.body = j; }
        }
        cp_wait<0>();
        __syncwarp();
        if (lane == 0) { __threadfence_block(); sig.body = ncol; }
    } else if (warp == INTRA_CW + 1) {
        /* ------------------------------- top loader ------------------------------- */
        if (has_top) {
            const volatile int *above = f.iprogress + row - 1;
            int prev_fast = 0;                      /* the previous turn took its border from the tagged words */
            for (int c = 0; c < ncol; c++) {
                if (info[c].count == 0) {
                    __syncwarp();
                    if (lane == 0) sig.top = c + 1;
                    prev_fast = 0;
                    continue;
                }
                /* Fast path: the superblock above (and the one above-left, for the top-left sample) went through the intra
                 * wavefront itself, so its bottom line arrives as tagged words (intra_ll) -- pixels and "ready" in one load. */
                int fast = 0;
                if (lane == 0) {
                    VP9_SPIN(sig.done < c - 2 || sig.stored < c - 3, 20, "intra top loader: ring", row, c);
                    const IntraSb *up = f.isb + (size_t)(row - 1) * ncol + c;
                    fast = info[c].needs_top && up[0].count > 0 && (c == 0 || up[-1].count > 0);
                }
                fast = __shfl_sync(0xffffffffu, fast, 0);
                if (fast) {
                    constexpr int WY = 64 * (int)sizeof(P) / 4, WC = 32 * (int)sizeof(P) / 4;     /* words per line */
                    const int k0 = c > 0 ? -4 : 0;                                             /* 16 bytes to the left: the top-left sample */
                    const int ny = WY - k0, nc = f.mono >= 0 ? 0 : WC - k0, total = ny + 2 * nc;
                    const unsigned long long *in = f.intra_ll + ((size_t)(row - 1) * ncol + c) * INTRA_LL_WORDS;
                    const unsigned tag = f.ll_tag;
                    const unsigned long long *src[3];
                    unsigned char *dst[3];
#pragma unroll
                    for (int j = 0; j < 3; j++) {
                        const int t = lane + 32 * j;
                        src[j] = nullptr; dst[j] = nullptr;
                        if (t < total) {
                            const int pl = t < ny ? 0 : t < ny + nc ? 1 : 2;
                            const int k = (pl == 0 ? t : pl == 1 ? t - ny : t - ny - nc) + k0;
                            const int W = pl ? WC : WY, S = pl ? 32 : 64, base = pl == 0 ? 0 : pl == 1 ? 32 : 48;
                            src[j] = k >= 0 ? in + base + k : in - INTRA_LL_WORDS + base + W + k;
                            dst[j] = tile(pl) + (((c & 3) * S * (int)sizeof(P) + 4 * k) & (4 * S * (int)sizeof(P) - 1));
                        }
                    }
                    unsigned long long w[3];
                    auto fetch = [&]() {
                        bool ready = true;
#pragma unroll
                        for (int j = 0; j < 3; j++) w[j] = src[j] ? ll_load(src[j]) : (unsigned long long)tag << 32;
#pragma unroll
                        for (int j = 0; j < 3; j++) ready &= (unsigned)(w[j] >> 32) == tag;
                        return ready;
                    };
                    VP9_SPIN(!fetch(), 0, "intra top loader: row above (words)", row, c);
#pragma unroll
                    for (int j = 0; j < 3; j++)
                        if (dst[j]) *(unsigned *)dst[j] = (unsigned)w[j];
                    __syncwarp();
                    if (lane == 0) { __threadfence_block(); sig.top = c + 1; }
                    prev_fast = 1;
                    continue;
                }
                if (lane == 0) {
                    if (info[c].needs_top)
                        VP9_SPIN(*above < c + 1, 20, "intra top loader: row above", row, c);
                    __threadfence();
                }
                __syncwarp();
#pragma unroll
                for (int pl = 0; pl < 3; pl++) {
                    if (pl && f.mono >= 0)
                        break;
                    const int S = pl ? 32 : 64;
                    const int cpr = S * (int)sizeof(P) / 16;
                    const int stride = f.cur.stride[pl];
                    const unsigned char *src = (const unsigned char *)((const P *)f.cur.pl[pl] + ((size_t)row * S - 1) * stride + c * S);
                    /* One extra chunk to the left: the top-left pixel lives in the previous ring slot.  Not after a turn
                     * that took its border from the tagged words: that slot then holds the final line already, while the
                     * picture may not show it yet (the words leave before the picture store), and a turn that does not
                     * wait for the row above would put stale pixels under the previous turn's blocks. */
                    const int k0 = (c > 0 && !prev_fast) ? -1 : 0;
                    if (lane - 1 >= k0 && lane - 1 < cpr) {
                        const int k = lane - 1;
                        const int col = ((c & 3) * S * (int)sizeof(P) + 16 * k) & (4 * S * (int)sizeof(P) - 1);
                        cp16(tile(pl) + col, src + 16 * k);
                    }
                }
                cp_commit();
                cp_wait<0>();
                __syncwarp();
                if (lane == 0) { __threadfence_block(); sig.top = c + 1; }
                prev_fast = 0;
            }
        }
    } else {
        /* ------------------------------- storer ------------------------------- */
        for (int c = 0; c < ncol; c++) {
            if (lane == 0) {
                VP9_SPIN(sig.done < c + 1, 20, "intra storer: done", row, c);
                __threadfence_block();
            }
            __syncwarp();
            if (info[c].count > 0) {
                auto store_rows = [&](int pl, int r0, int r1) {
                    if (pl && f.mono >= 0)
                        return;
                    const int S = pl ? 32 : 64, rsb = pl ? L::RSBC : L::RSBY;
                    const int cpr = S * (int)sizeof(P) / 16;
                    const int stride = f.cur.stride[pl];
                    unsigned char *dst = (unsigned char *)((P *)f.cur.pl[pl] + ((size_t)row * S - 1) * stride + c * S);
                    const unsigned char *src = tile(pl) + (c & 3) * S * (int)sizeof(P);
                    for (int i = lane + r0 * cpr; i < r1 * cpr; i += 32) {
                        const int r = i / cpr, k = i % cpr;
                        *(uint4 *)(dst + (ptrdiff_t)r * stride * (int)sizeof(P) + 16 * k) = *(const uint4 *)(src + r * rsb + 16 * k);
                    }
                };
                if (row + 1 < f.sb_rows) {      /* the bottom line as tagged words for the intra row below (its fast path) */
                    constexpr int WY = 64 * (int)sizeof(P) / 4, WC = 32 * (int)sizeof(P) / 4;
                    unsigned long long *out = f.intra_ll + ((size_t)row * ncol + c) * INTRA_LL_WORDS;
                    if (lane < WY)
                        ll_store(out + lane, *(const unsigned *)(tile(0) + 64 * L::RSBY + (c & 3) * 64 * (int)sizeof(P) + 4 * lane), f.ll_tag);
                    if (f.mono < 0 && lane < WC) {
                        ll_store(out + 32 + lane, *(const unsigned *)(tile(1) + 32 * L::RSBC + (c & 3) * 32 * (int)sizeof(P) + 4 * lane), f.ll_tag);
                        ll_store(out + 48 + lane, *(const unsigned *)(tile(2) + 32 * L::RSBC + (c & 3) * 32 * (int)sizeof(P) + 4 * lane), f.ll_tag);
                    }
                }
                store_rows(0, 64, 65); store_rows(1, 32, 33); store_rows(2, 32, 33);
                __syncwarp();
                if (lane == 0) {
                    __threadfence();
                    *(volatile int *)(f.iprogress + row) = c + 1;
                }
                store_rows(0, 1, 64); store_rows(1, 1, 32); store_rows(2, 1, 32);
                __syncwarp();
                if (lane == 0) {
                    __threadfence();
                    *(volatile int *)(ifull + row) = c + 1;
                }
            } else if (lane == 0) {
                *(volatile int *)(f.iprogress + row) = c + 1;
                *(volatile int *)(ifull + row) = c + 1;
            }
            __syncwarp();
            if (lane == 0)
                sig.stored = c + 1;
        }
    }
}

template <class P>
__global__ void __launch_bounds__(INTRA_THREADS) k_intra(const __grid_constant__ FrameDesc f)
{
    VP9_DYN_SMEM(sm);
    __shared__ IntraSig sig;
    intra_row<P>(f, blockIdx.x, sm, sig);
}

} /* namespace vp9 */
#endif
