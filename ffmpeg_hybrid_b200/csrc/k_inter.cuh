/*
 * k_inter.cuh -- K2: inter prediction for a whole frame (frame-centric, no dependencies).
 *
 *   k_cellmap : one thread per inter block record (grid.y = producer lane), writes the record index into every 8x8
 *               cell the block covers (the per-superblock mode / motion-vector map) and folds
 *               the frame's largest downward motion vector into mv_reach.
 *   k_inter_gate : optional (set_option "overlap_inter"): one warp in front of a band of k_inter
 *               tile rows, waits for the previous frame's loop-filter rows the band can reach.
 *   k_inter   : one CTA per 64x32 luma tile (+ the two 32x16 chroma tiles under it).  A thread
 *               owns a vertical strip (8 luma or 4 chroma rows of one column): it filters the
 *               rows it needs horizontally once, then slides the vertical 8-tap over them.
 *               Reference samples are read with coordinate clamping (= the reference's
 *               emulated_edge_mc, vp9recon.c:402-412).  Compound blocks average the two
 *               predictions in registers (vp9dsp_template.c:2051: (dst + pred + 1) >> 1).
 *               Scaled references use the same strip with a sliding window of
 *               horizontally filtered rows (mc_strip_scaled).
 *
 * Replaces inter_pred (vp9_mc_template.c:30-464) + mc_luma/chroma_{unscaled,scaled}
 * (vp9recon.c:376-628) + the mc/smc slots (vp9dsp_template.c:1969-2569).
 * Algorithmic bytes per frame: P (write) + r*P (reference reads), r = 1 or 2.
 */
#ifndef VP9CUDA_K_INTER_CUH
#define VP9CUDA_K_INTER_CUH

#include "dev_common.cuh"

namespace vp9 {

__global__ void k_cellmap(const __grid_constant__ FrameDesc f)
{
    /* blockIdx.y = producer lane; its records sit at lane_blk_base[lane] */
    const int lane = blockIdx.y;
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x, n = f.lane_blk_cnt[lane];
    const uint32_t i = f.lane_blk_base[lane] + k;
    const Vp9CudaInterBlock *__restrict__ blk = f.blk;
    int reach = 0;
    if (k < n) {
        /* how far below its own position any prediction of this frame may read (vertical MV, 1/8 pel) */
        const int nref = blk[i].comp ? 2 : 1, nsub = blk[i].bs > 9 ? 4 : 1;
        for (int q = 0; q < nsub; q++)
            for (int r = 0; r < nref; r++)
                reach = max(reach, (int)blk[i].mv[q][r][1]);
    }
    reach = __reduce_max_sync(0xffffffffu, reach);
    if ((threadIdx.x & 31) == 0 && reach > 0)
        atomicMax(f.mv_reach, reach);
    if (k >= n)
        return;
    int col = blk[i].col, row = blk[i].row, bs = blk[i].bs;
    int w = min(bs_w8(bs), f.cols - col), h = min(bs_h8(bs), f.rows - row);
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++)
            f.cellmap[(row + y) * f.cols + col + x] = (int32_t)i;
}

/* horizontal stage for one sample; CLAMP = false when the caller knows every tap is inside the plane */
template <class P, bool CLAMP>
__device__ __forceinline__ int mc_hrow(const RefPlane<P> &r, int x, int y, int filter, int phx, int bpp)
{
    if (CLAMP) {
        if (phx == 0)
            return r.at(x, y);
        int s[8];
#pragma unroll
        for (int i = 0; i < 8; i++)
            s[i] = r.at(x + i - 3, y);
        return filt8(s, filter, phx, bpp);
    } else {
        const P *row = r.pix + (size_t)y * r.stride + x;
        if (phx == 0)
            return row[0];
        int s[8];
#pragma unroll
        for (int i = 0; i < 8; i++)
            s[i] = row[i - 3];
        return filt8(s, filter, phx, bpp);
    }
}

/* NR vertically adjacent outputs at integer reference position (x, y), phases (phx, phy) */
template <class P, int NR, bool CLAMP>
__device__ __forceinline__ void mc_strip_c(const RefPlane<P> &r, int x, int y, int phx, int phy,
                                           int filter, int bpp, int *out)
{
    if (phy == 0) {
#pragma unroll
        for (int j = 0; j < NR; j++)
            out[j] = mc_hrow<P, CLAMP>(r, x, y + j, filter, phx, bpp);
    } else {
        int hbuf[NR + 7];
#pragma unroll
        for (int j = 0; j < NR + 7; j++)
            hbuf[j] = mc_hrow<P, CLAMP>(r, x, y - 3 + j, filter, phx, bpp);
#pragma unroll
        for (int j = 0; j < NR; j++)
            out[j] = filt8(&hbuf[j], filter, phy, bpp);
    }
}

template <class P, int NR>
__device__ __forceinline__ void mc_strip(const RefPlane<P> &r, int x, int y, int phx, int phy,
                                         int filter, int bpp, int *out)
{
    /* interior strips (nearly all of them) skip the coordinate clamps of the edge emulation */
    if (x >= 3 && x + 4 < r.w && y >= 3 && y + NR + 3 < r.h)
        mc_strip_c<P, NR, false>(r, x, y, phx, phy, filter, bpp, out);
    else
        mc_strip_c<P, NR, true>(r, x, y, phx, phy, filter, bpp, out);
}

/* Scaled reference (smc slots, vp9dsp_template.c:2363-2470): NR vertically adjacent outputs at
 * offset (ox, oy .. oy+NR-1) of a prediction call at position p.  The column's integer position
 * and phase are fixed; the row position advances by p.dy/16 per output, so the eight
 * horizontally filtered rows under the vertical taps slide down by 0..2 rows per output.  Same
 * values as mc_pixel (vp9_pred.h): horizontal pass rounded and clipped to a pixel, then vertical. */
template <class P, int NR>
__device__ __forceinline__ void mc_strip_scaled(const RefPlane<P> &r, const McPos &p, int ox, int oy,
                                                int filter, int bpp, int *out)
{
    const int ax = p.fx + ox * p.dx;
    const int x = p.x0 + (ax >> 4), phx = ax & 15;
    int ay = p.fy + oy * p.dy;
    int cur = (ay >> 4) - 8;             /* the first output pulls in all eight rows */
    int win[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    /* one call site for the row filter keeps the kernel small: rolled loops, outputs collected in
     * a 128-bit shift register (pixels are at most 12 bits wide) */
    unsigned long long lo = 0, hi = 0;
#pragma unroll 1
    for (int j = 0; j < NR; j++) {
        const int tgt = ay >> 4;
#pragma unroll 1
        while (cur < tgt) {
#pragma unroll
            for (int t = 0; t < 7; t++)
                win[t] = win[t + 1];
            cur++;
            win[7] = mc_hrow<P, true>(r, x, p.y0 + cur + 4, filter, phx, bpp);
        }
        const unsigned long long v = (unsigned long long)filt8(win, filter, ay & 15, bpp);
        lo = (lo >> 16) | (hi << 48);
        hi = (hi >> 16) | (v << 48);
        ay += p.dy;
    }
#pragma unroll
    for (int j = 0; j < NR; j++) {
        const int k = 8 - NR + j;        /* position in the 8-entry register */
        out[j] = (int)(((k < 4 ? lo >> (16 * k) : hi >> (16 * (k - 4)))) & 0xffff);
    }
}

/* Geometry of the prediction call of inter_pred() (vp9_mc_template.c:30-464) that covers pixel
 * (X, Y) of plane pl for reference i of block b: origin of the call in the plane, its offset and
 * the size of the coding block (the scaled path's MV clamp needs them), and the motion vector. */
struct InterCall {
    int ox0, oy0, px, py, pw, ph, mvx, mvy;
};
__device__ __forceinline__ InterCall inter_call(const Vp9CudaInterBlock &b, int i, int pl, int X, int Y)
{
    InterCall c;
    const int ch = pl > 0;
    c.px = c.py = 0;
    if (b.bs > 9) {
        if (!ch) {
            const int k = ((Y >> 2) & 1) * 2 + ((X >> 2) & 1);
            c.px = 4 * (k & 1); c.py = 4 * (k >> 1); c.pw = c.ph = 8;
            c.ox0 = b.col * 8 + c.px; c.oy0 = b.row * 8 + c.py;
            c.mvx = b.mv[k][i][0]; c.mvy = b.mv[k][i][1];
        } else {
            c.pw = c.ph = 4;
            c.ox0 = b.col * 4; c.oy0 = b.row * 4;
            c.mvx = rounded_div(b.mv[0][i][0] + b.mv[1][i][0] + b.mv[2][i][0] + b.mv[3][i][0], 4);
            c.mvy = rounded_div(b.mv[0][i][1] + b.mv[1][i][1] + b.mv[2][i][1] + b.mv[3][i][1], 4);
        }
    } else {
        c.pw = (bs_w8(b.bs) * 8) >> ch; c.ph = (bs_h8(b.bs) * 8) >> ch;
        c.ox0 = (b.col * 8) >> ch; c.oy0 = (b.row * 8) >> ch;
        c.mvx = b.mv[0][i][0]; c.mvy = b.mv[0][i][1];
    }
    return c;
}

/* One strip of NR rows at column X starting at row Y of plane pl, all inside one 8x8 cell
 * (NR = 8 luma, 4 chroma) and, for sub-8x8 luma, inside one 4x4 (NR = 4). */
template <class P, int NR>
__device__ __forceinline__ void inter_strip(const FrameDesc &f, const Vp9CudaInterBlock &b, int pl, int X,
                                            int Y, int *out)
{
    const int ch = pl > 0;
    const int nref = b.comp ? 2 : 1;
    int acc[NR];
    for (int i = 0; i < nref; i++) {
        const int ri = b.ref[i];
        const InterCall c = inter_call(b, i, pl, X, Y);
        const SurfDesc &rs = f.ref[ri];
        RefPlane<P> rp;
        rp.pix = (const P *)rs.pl[pl]; rp.stride = rs.stride[pl];
        rp.w = (rs.w + ch) >> ch; rp.h = (rs.h + ch) >> ch;
        int v[NR];
        if (f.mvscale[ri][0]) {
            const McPos pos = mc_pos_scaled(c.ox0, c.oy0, c.mvx, c.mvy, ch, c.px, c.py, c.pw, c.ph, f.cols, f.rows,
                                            f.mvscale[ri][0], f.mvscale[ri][1], f.mvstep[ri][0], f.mvstep[ri][1]);
            mc_strip_scaled<P, NR>(rp, pos, X - c.ox0, Y - c.oy0, b.filter, f.bpp, v);
        } else {
            const McPos pos = mc_pos_unscaled(X, Y, c.mvx, c.mvy, ch);
            mc_strip<P, NR>(rp, pos.x0, pos.y0, pos.fx, pos.fy, b.filter, f.bpp, v);
        }
        if (i == 0) {
#pragma unroll
            for (int j = 0; j < NR; j++) acc[j] = v[j];
        } else {
#pragma unroll
            for (int j = 0; j < NR; j++) acc[j] = (acc[j] + v[j] + 1) >> 1;
        }
    }
#pragma unroll
    for (int j = 0; j < NR; j++) out[j] = acc[j];
}

/* Gate in front of a band of k_inter tile rows that is launched while the PREVIOUS frame's loop
 * filter may still be running (inter stream): one warp waits until the superblock rows the band
 * can reach are final, then the band's k_inter launch follows in stream order.  A single warp
 * cannot keep the wavefront's CTAs off the machine, so nothing can deadlock.
 * Loop-filter row r finishes pixel rows [64r - 8, 64r + 56) (chroma: [32r - 8, 32r + 24)), so
 * luma-scale row Y is final with loop-filter row (Y + 16) >> 6 at the latest; rows finish top
 * down.  Reach below the band's last row: integer motion (k_cellmap's maximum) + 4 taps, with
 * slack for the rounding of the chroma / sub-8x8 average vectors. */
__global__ void k_inter_gate(const int *dep_done, int dep_rows, int dep_whole, const int *mv_reach, int y_last,
                             int *err, int *err_host)
{
    struct { int *err, *err_host; } f = { err, err_host };      /* what VP9_SPIN reports to */
    (void)f;
    if (threadIdx.x == 0) {
        int need = dep_rows - 1;
        if (!dep_whole)
            need = min(need, (y_last + (*(volatile const int *)mv_reach >> 3) + 12 + 16) >> 6);
        const volatile int *done = dep_done;
        for (int pl = 0; pl < 3; pl++)
            VP9_SPIN(done[pl * dep_rows + need] == 0, 500, "inter gate: previous frame's loop filter", need, pl);
        __threadfence();
    }
}

template <class P>
__global__ void __launch_bounds__(256, sizeof(P) == 2 ? 3 : 2) k_inter(const __grid_constant__ FrameDesc f)
{
    const int tid = threadIdx.x;
    const int X0 = blockIdx.x * 64, Y0 = (blockIdx.y + f.tile_y0) * 32;
    /* ---- luma: strip = column tid&63, rows (tid>>6)*8 .. +7 ---- */
    {
        const int X = X0 + (tid & 63), Y = Y0 + (tid >> 6) * 8;
        if (X < f.cols * 8 && Y < f.rows * 8) {
            int ci = f.cellmap[(Y >> 3) * f.cols + (X >> 3)];
            if (ci >= 0) {
                const Vp9CudaInterBlock &b = f.blk[ci];
                P *dst = (P *)f.cur.pl[0] + (size_t)Y * f.cur.stride[0] + X;
                int out[8];
                if (b.bs > 9) {
                    inter_strip<P, 4>(f, b, 0, X, Y, out);
                    inter_strip<P, 4>(f, b, 0, X, Y + 4, out + 4);
                } else {
                    inter_strip<P, 8>(f, b, 0, X, Y, out);
                }
#pragma unroll
                for (int j = 0; j < 8; j++)
                    dst[(size_t)j * f.cur.stride[0]] = (P)out[j];
            }
        }
    }
    /* ---- chroma: plane 1 + (tid>>7), strip = column tid&31, rows ((tid>>5)&3)*4 .. +3 ---- */
    {
        const int pl = 1 + (tid >> 7);
        const int X = (X0 >> 1) + (tid & 31), Y = (Y0 >> 1) + ((tid >> 5) & 3) * 4;
        if (X < f.cols * 4 && Y < f.rows * 4) {
            int ci = f.cellmap[(Y >> 2) * f.cols + (X >> 2)];
            if (ci >= 0) {
                const Vp9CudaInterBlock &b = f.blk[ci];
                P *dst = (P *)f.cur.pl[pl] + (size_t)Y * f.cur.stride[pl] + X;
                int out[4];
                inter_strip<P, 4>(f, b, pl, X, Y, out);
#pragma unroll
                for (int j = 0; j < 4; j++)
                    dst[(size_t)j * f.cur.stride[pl]] = (P)out[j];
            }
        }
    }
}

} /* namespace vp9 */
#endif
