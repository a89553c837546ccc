/*
 * k_intra.cuh -- K3: intra prediction + residual as a dependency-driven wavefront.
 *
 * Every intra transform block depends on the reconstructed pixels left of, above (and for some
 * 4x4 blocks above-right of) it (check_intra_mode, vp9recon.c:37-221).  The host records carry
 * each block's depth in that dependency graph (Vp9CudaIntraTx.level).  Here:
 *
 *   k_ilvl_hist / k_ilvl_scan / k_ilvl_scatter : counting sort of the frame's records by level
 *       (a topological order in which equal levels are mutually independent).
 *   k_intra : persistent warps claim blocks in that order from a global counter.  A warp first
 *       waits until the 4x4 cells it may read are final -- one lane polls one cell of the
 *       per-plane "done" map, and only cells that hold intra blocks of this frame are waited
 *       for -- then builds the edge arrays lane-parallel straight from the picture in L2,
 *       predicts into shared memory, adds the residual there (the same itx_block as K1),
 *       writes the block out once, fences, and marks its cells done.  Because claims are made
 *       in level order, every block a claimed block waits for has already been claimed by a
 *       running warp, so the scheme cannot deadlock whatever the grid size.
 *
 * The frame thus advances along its true dependency wavefront at 4x4-cell granularity instead
 * of superblock by superblock; in inter frames the scattered intra clusters are independent.
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

#define ILVL_MAX 65536

__global__ void k_ilvl_hist(const Vp9CudaIntraTx *__restrict__ itx, uint32_t n, uint32_t *hist)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
        atomicAdd(&hist[itx[i].level], 1u);
}

/* in-place exclusive scan of hist[0..ILVL_MAX) by one CTA of 1024 threads */
__global__ void __launch_bounds__(1024) k_ilvl_scan(uint32_t *hist)
{
    __shared__ uint32_t part[1024];
    constexpr int PER = ILVL_MAX / 1024;
    uint32_t loc[PER], s = 0;
    for (int k = 0; k < PER; k++) {
        loc[k] = hist[threadIdx.x * PER + k];
        s += loc[k];
    }
    part[threadIdx.x] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        uint32_t v = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    uint32_t base = part[threadIdx.x] - s;
    for (int k = 0; k < PER; k++) {
        hist[threadIdx.x * PER + k] = base;
        base += loc[k];
    }
}

__global__ void k_ilvl_scatter(const Vp9CudaIntraTx *__restrict__ itx, uint32_t n, uint32_t *cursor,
                               uint32_t *order)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
        order[atomicAdd(&cursor[itx[i].level], 1u)] = i;
}

/* current picture read through L2 (neighbours are written by other SMs during this kernel) */
template <class P> struct L2PlanePx {
    const P *p; int stride;
    __device__ __forceinline__ int operator()(int x, int y) const { return ld_cg(p + (size_t)y * stride + x); }
};
/* a block being built in shared memory */
template <class P> struct SmemBlk {
    P *p; int stride;
    __device__ __forceinline__ int get(int x, int y) const { return p[y * stride + x]; }
    __device__ __forceinline__ void put(int x, int y, int v) { p[y * stride + x] = (P)v; }
};

#define INTRA_WARPS 4

template <class P, bool HBD>
__global__ void __launch_bounds__(INTRA_WARPS * 32) k_intra(const __grid_constant__ FrameDesc f)
{
    typedef typename ItxTypes<HBD>::C C;
    __shared__ int edge_s[INTRA_WARPS][100];
    __shared__ P pred_s[INTRA_WARPS][32 * 32];
    __shared__ C tmp_s[INTRA_WARPS][32 * 33];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int *top = edge_s[warp], *left = top + 66;
    P *pred = pred_s[warp];
    C *tmp = tmp_s[warp];

    for (;;) {
        uint32_t i = 0;
        if (lane == 0)
            i = atomicAdd(f.icounter, 1u);
        i = __shfl_sync(0xffffffffu, i, 0);
        if (i >= f.n_itx)
            break;
        const Vp9CudaIntraTx r = f.itx[f.iorder[i]];
        const int pl = r.plane, ch = pl > 0;
        const int n = 1 << r.tx, N = 4 << r.tx;
        const int cx = r.x4, cy = r.y4;
        const bool have_left = r.flags & VP9CUDA_ITX_HAVE_LEFT, have_right = r.flags & VP9CUDA_ITX_HAVE_RIGHT;
        const bool have_top = cy > 0;
        const int W = f.done_w[pl], H = f.done_h[pl];
        uint32_t *done = f.done[pl];

        /* ---- wait for the cells this block may read (same set the host used for `level`) ---- */
        {
            int px = -1, py = -1;
            if (lane <= n) {                       /* left column, lane 0 = top-left */
                if (have_left && (lane > 0 || have_top) && cy - 1 + lane < H) { px = cx - 1; py = cy - 1 + lane; }
            } else if (lane >= 16 && have_top) {   /* top row (+ top-right for 4x4) */
                int k = lane - 16, xe = (r.tx == 0 && have_right) ? 2 : n;
                if (k < xe && cx + k < W) { px = cx + k; py = cy - 1; }
            }
            if (px >= 0 && f.n_blk) {
                /* only intra blocks of this frame are ever marked done: skip inter cells */
                int ci = ch ? py * f.cols + px : (py >> 1) * f.cols + (px >> 1);
                if (f.cellmap[ci] >= 0)
                    px = -1;
            }
            if (px >= 0) {
                const volatile uint32_t *flag = done + (size_t)py * W + px;
                unsigned ns = 32;
                while (*flag != f.gen) {
                    __nanosleep(ns);
                    if (ns < 256) ns <<= 1;
                }
            }
            __syncwarp();
            __threadfence();
        }

        /* ---- edges, prediction, residual, all in shared memory ---- */
        const int bx = cx * 4, by = cy * 4;
        const int mode = intra_mode_conv(r.mode, have_left, have_top);
        const int needs = intra_mode_needs(mode);
        const int stride = f.cur.stride[pl];
        P *plane = (P *)f.cur.pl[pl];
        L2PlanePx<P> px = { plane, stride };
        const int avail_w = ((f.cols * 8) >> ch) - bx, avail_h = ((f.rows * 8) >> ch) - by;
        for (int e = lane; e < 3 * N + 1; e += 32) {
            if (e <= 2 * N) {
                if (needs & 2)
                    top[e] = intra_edge_top(px, e, bx, by, N, needs, have_top, have_left, have_right, avail_w, f.bpp);
            } else if (needs & 1) {
                left[e - 2 * N - 1] = intra_edge_left(px, e - 2 * N - 1, bx, by, have_left, avail_h, f.bpp);
            }
        }
        __syncwarp();
        const int dc = intra_dc(mode, N, top, left, f.bpp);
        for (int k = lane; k < N * N; k += 32)
            pred[k] = (P)intra_pixel(mode, N, k & (N - 1), k >> (2 + r.tx), top, left, dc, f.bpp);
        __syncwarp();
        if (r.eob) {
            SmemBlk<P> blk = { pred, N };
            const C *coef = (const C *)f.coef + r.coef_off;
            switch (r.tx) {
            case 0: if (lane < 4)  itx_block<HBD, 4>(coef, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xfu); break;
            case 1: if (lane < 8)  itx_block<HBD, 8>(coef, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xffu); break;
            case 2: if (lane < 16) itx_block<HBD, 16>(coef, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xffffu); break;
            default: itx_block<HBD, 32>(coef, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xffffffffu); break;
            }
            __syncwarp();
        }
        for (int k = lane; k < N * N; k += 32)
            plane[(size_t)(by + (k >> (2 + r.tx))) * stride + bx + (k & (N - 1))] = pred[k];
        __threadfence();
        __syncwarp();
        /* ---- publish ---- */
        for (int k = lane; k < n * n; k += 32) {
            int x = cx + (k & (n - 1)), y = cy + (k >> r.tx);
            if (x < W && y < H)
                *(volatile uint32_t *)(done + (size_t)y * W + x) = f.gen;
        }
    }
}

} /* namespace vp9 */
#endif
