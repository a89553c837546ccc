/*
 * k_lf.cuh -- K4: the VP9 loop filter as a superblock wavefront, in the reference's exact
 * edge order.
 *
 * The reference filters superblock by superblock in raster order and, inside one, plane by
 * plane: every column edge, then every row edge (vp9lpf.c:222-229).  A filter reaches up to 8
 * pixels into the left and upper neighbours, so superblock (r, c) must run after (r, c-1)
 * [same CTA] and after (r-1, c+1) [whose column pass touches our upper strip]: a (left,
 * top-right) wavefront.  Planes never interact, so each plane is its own wavefront.
 *
 *   k_lf : one CTA per (superblock row, plane), walking the row left to right behind the row
 *          above (progress counters in global memory).  Per superblock: stage the (S+8)x(S+8)
 *          pixel window in shared memory, decode the VP9Filter masks into an edge-op table
 *          (vp9_lf.h, a 1:1 walk of filter_plane_cols/rows, vp9lpf.c:31-181), run the column
 *          pass with one thread per pixel row (edges left to right), the row pass with one
 *          thread per pixel column (edges top to bottom), and write the window back.
 *
 * Replaces ff_vp9_loopfilter_sb (vp9lpf.c:183-230) + loop_filter_8/16/mix2 slots
 * (vp9dsp_template.c:1780-1967).  Algorithmic bytes: 2*P (read + write of the picture).
 */
#ifndef VP9CUDA_K_LF_CUH
#define VP9CUDA_K_LF_CUH

#include "dev_common.cuh"

namespace vp9 {

#define LF_THREADS 128

/* pre-pass, fully parallel: one thread per (superblock, plane, direction) turns the VP9Filter
 * masks into the edge-op table the wavefront kernel consumes */
__global__ void k_lf_ops(const __grid_constant__ FrameDesc f)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int n_sb = f.sb_cols * f.sb_rows;
    if (t >= n_sb * 6)
        return;
    const int sb = t / 6, pl = (t % 6) >> 1, dir = t & 1;
    const LfSb &m = *(const LfSb *)&f.lf[sb];
    LfOps &o = f.lfops[sb * 3 + pl];
    if (dir == 0)
        lf_decode_cols(m, pl > 0, sb % f.sb_cols == 0, o);
    else
        lf_decode_rows(m, pl > 0, sb / f.sb_cols == 0, o);
}

#define LF_CLK(k) do { if (f.phase_clk && threadIdx.x == 0) { long long c_ = clock64(); \
    atomicAdd(&f.phase_clk[k], (unsigned long long)(c_ - clk_)); clk_ = c_; } } while (0)

template <class P, int S>
__device__ void lf_superblock(const FrameDesc &f, int pl, int sr, int sc, P *win /* (S+8)*(S+10) */, LfOps *ops,
                              long long &clk_)
{
    constexpr int W = S + 10;               /* odd word stride: conflict-free line-per-thread access */
    const int ch = pl > 0;
    const int x0 = sc * S, y0 = sr * S;
    const int stride = f.cur.stride[pl];
    P *plane = (P *)f.cur.pl[pl];
    (void)ch;
    for (int i = threadIdx.x; i < (S + 8) * (S + 8); i += blockDim.x) {
        int x = i % (S + 8) - 8, y = i / (S + 8) - 8;
        int gx = x0 + x, gy = y0 + y;
        P v = 0;
        if (gx >= 0 && gy >= 0 && !(x < 0 && y < 0))
            v = ld_cg(plane + (size_t)gy * stride + gx);
        win[(y + 8) * W + x + 8] = v;
    }
    {
        const uint32_t *src = (const uint32_t *)&f.lfops[(sr * f.sb_cols + sc) * 3 + pl];
        uint32_t *dst = (uint32_t *)ops;
        for (int i = threadIdx.x; i < (int)(sizeof(LfOps) / 4); i += blockDim.x)
            dst[i] = src[i];
    }
    __syncthreads();
    LF_CLK(1);

    /* column pass: thread = pixel row */
    if (threadIdx.x < S) {
        const int ln = threadIdx.x;
        P *row = win + (ln + 8) * W + 8;
#pragma unroll 1
        for (int p = 0; p < S / 4; p++) {
            LfOp op = ops->col[ln >> 3][p];
            if (!op.wd)
                continue;
            int px[16];
            P *e = row + 4 * p;
            if (op.wd == 16) {
#pragma unroll
                for (int i = 0; i < 16; i++) px[i] = e[i - 8];
            } else {
#pragma unroll
                for (int i = 4; i < 12; i++) px[i] = e[i - 8];
            }
            unsigned w = lf_line(px, op.wd, f.mblim_lut[op.L], f.lim_lut[op.L], op.L >> 4, f.bpp);
            if (w) {
#pragma unroll
                for (int i = 1; i < 15; i++)
                    if (w >> i & 1) e[i - 8] = (P)px[i];
            }
        }
    }
    __syncthreads();
    LF_CLK(2);
    /* row pass: thread = pixel column */
    if (threadIdx.x < S) {
        const int ln = threadIdx.x;
        P *colp = win + 8 * W + 8 + ln;
#pragma unroll 1
        for (int p = 0; p < S / 4; p++) {
            LfOp op = ops->row[p][ln >> 3];
            if (!op.wd)
                continue;
            int px[16];
            P *e = colp + 4 * p * W;
            if (op.wd == 16) {
#pragma unroll
                for (int i = 0; i < 16; i++) px[i] = e[(i - 8) * W];
            } else {
#pragma unroll
                for (int i = 4; i < 12; i++) px[i] = e[(i - 8) * W];
            }
            unsigned w = lf_line(px, op.wd, f.mblim_lut[op.L], f.lim_lut[op.L], op.L >> 4, f.bpp);
            if (w) {
#pragma unroll
                for (int i = 1; i < 15; i++)
                    if (w >> i & 1) e[(i - 8) * W] = (P)px[i];
            }
        }
    }
    __syncthreads();
    LF_CLK(3);
    for (int i = threadIdx.x; i < (S + 8) * (S + 8); i += blockDim.x) {
        int x = i % (S + 8) - 8, y = i / (S + 8) - 8;
        int gx = x0 + x, gy = y0 + y;
        if (gx >= 0 && gy >= 0 && !(x < 0 && y < 0))
            plane[(size_t)gy * stride + gx] = win[(y + 8) * W + x + 8];
    }
}

template <class P>
__global__ void __launch_bounds__(LF_THREADS) k_lf(const __grid_constant__ FrameDesc f)
{
    __shared__ P win[72 * 74];
    __shared__ LfOps ops;
    const int pl = blockIdx.x % 3, row = blockIdx.x / 3;
    int *prog = f.progress + pl * f.sb_rows;
    long long clk_ = clock64();
    for (int c = 0; c < f.sb_cols; c++) {
        if (row > 0 && threadIdx.x == 0) {
            const int need = min(c + 2, f.sb_cols);
            volatile int *above = prog + row - 1;
            while (*above < need)
                __nanosleep(100);
            __threadfence();
        }
        __syncthreads();
        LF_CLK(0);
        if (pl == 0)
            lf_superblock<P, 64>(f, pl, row, c, win, &ops, clk_);
        else
            lf_superblock<P, 32>(f, pl, row, c, win, &ops, clk_);
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0)
            atomicExch(prog + row, c + 1);
        LF_CLK(4);
    }
}

} /* namespace vp9 */
#endif
