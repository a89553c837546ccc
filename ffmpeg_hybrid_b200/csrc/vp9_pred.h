/*
 * vp9_pred.h -- per-pixel formulations of VP9 inter and intra prediction.
 *
 * Everything here is a pure function of edge / reference samples, written so that one GPU
 * thread (or one iteration of the host model's loop) can produce any single output pixel.
 * Reference behaviour reproduced (paths relative to the reference tree):
 *   sub-pel filters             libavcodec/vp9dsp.c:27-85, vp9dsp_template.c:2026-2113 (8-tap),
 *                               :2150-2256 (bilinear), :2363-2470 (scaled)
 *   motion-vector addressing    libavcodec/vp9recon.c:376-467 (unscaled), :492-628 (scaled)
 *   edge emulation              libavcodec/videodsp_template.c:27-105 (== coordinate clamp)
 *   intra edge preparation      libavcodec/vp9recon.c:37-221 (check_intra_mode)
 *   intra predictors            libavcodec/vp9dsp_template.c:33-1153
 */
#ifndef VP9_PRED_H
#define VP9_PRED_H

#include "vp9_itx.h"

namespace vp9 {

enum { VERT_PRED, HOR_PRED, DC_PRED, DIAG_DOWN_LEFT_PRED, DIAG_DOWN_RIGHT_PRED, VERT_RIGHT_PRED,
       HOR_DOWN_PRED, VERT_LEFT_PRED, HOR_UP_PRED, TM_VP8_PRED, LEFT_DC_PRED, TOP_DC_PRED,
       DC_128_PRED, DC_127_PRED, DC_129_PRED };
enum { FILTER_8TAP_SMOOTH, FILTER_8TAP_REGULAR, FILTER_8TAP_SHARP, FILTER_BILINEAR };

/* 8-tap kernels indexed [FilterMode][1/16 phase][tap] (VP9 bitstream specification tables) */
#ifdef __CUDACC__
#define VP9_TABLE static __device__ __constant__ const
#else
#define VP9_TABLE static const
#endif
#define VP9_SUBPEL_TABLE_INIT { \
  { /* smooth */ \
    {0,0,0,128,0,0,0,0},{-3,-1,32,64,38,1,-3,0},{-2,-2,29,63,41,2,-3,0},{-2,-2,26,63,43,4,-4,0}, \
    {-2,-3,24,62,46,5,-4,0},{-2,-3,21,60,49,7,-4,0},{-1,-4,18,59,51,9,-4,0},{-1,-4,16,57,53,12,-4,-1}, \
    {-1,-4,14,55,55,14,-4,-1},{-1,-4,12,53,57,16,-4,-1},{0,-4,9,51,59,18,-4,-1},{0,-4,7,49,60,21,-3,-2}, \
    {0,-4,5,46,62,24,-3,-2},{0,-4,4,43,63,26,-2,-2},{0,-3,2,41,63,29,-2,-2},{0,-3,1,38,64,32,-1,-3} }, \
  { /* regular */ \
    {0,0,0,128,0,0,0,0},{0,1,-5,126,8,-3,1,0},{-1,3,-10,122,18,-6,2,0},{-1,4,-13,118,27,-9,3,-1}, \
    {-1,4,-16,112,37,-11,4,-1},{-1,5,-18,105,48,-14,4,-1},{-1,5,-19,97,58,-16,5,-1},{-1,6,-19,88,68,-18,5,-1}, \
    {-1,6,-19,78,78,-19,6,-1},{-1,5,-18,68,88,-19,6,-1},{-1,5,-16,58,97,-19,5,-1},{-1,4,-14,48,105,-18,5,-1}, \
    {-1,4,-11,37,112,-16,4,-1},{-1,3,-9,27,118,-13,4,-1},{0,2,-6,18,122,-10,3,-1},{0,1,-3,8,126,-5,1,0} }, \
  { /* sharp */ \
    {0,0,0,128,0,0,0,0},{-1,3,-7,127,8,-3,1,0},{-2,5,-13,125,17,-6,3,-1},{-3,7,-17,121,27,-10,5,-2}, \
    {-4,9,-20,115,37,-13,6,-2},{-4,10,-23,108,48,-16,8,-3},{-4,10,-24,100,59,-19,9,-3},{-4,11,-24,90,70,-21,10,-4}, \
    {-4,11,-23,80,80,-23,11,-4},{-4,10,-21,70,90,-24,11,-4},{-3,9,-19,59,100,-24,10,-4},{-3,8,-16,48,108,-23,10,-4}, \
    {-2,6,-13,37,115,-20,9,-4},{-2,5,-10,27,121,-17,7,-3},{-1,3,-6,17,125,-13,5,-2},{0,1,-3,8,127,-7,3,-1} } }

#ifdef __CUDACC__
static __device__ __constant__ const int16_t d_subpel[3][16][8] = VP9_SUBPEL_TABLE_INIT;
#endif
static const int16_t h_subpel[3][16][8] = VP9_SUBPEL_TABLE_INIT;

VP9_HD const int16_t *subpel_taps(int filter, int phase)
{
#ifdef __CUDA_ARCH__
    return d_subpel[filter][phase];
#else
    return h_subpel[filter][phase];
#endif
}

VP9_HD int iclamp(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* One 1-D filter tap set applied to 8 samples s[0..7] (s[3] is the centre, phase toward s[4]).
 * filter 0..2: (sum F[i]*s[i] + 64) >> 7 clipped; filter 3 (bilinear): s3 + ((ph*(s4-s3)+8)>>4). */
VP9_HD int filt8(const int *s, int filter, int phase, int bpp)
{
    if (filter == FILTER_BILINEAR)
        return s[3] + ((phase * (s[4] - s[3]) + 8) >> 4);
    const int16_t *F = subpel_taps(filter, phase);
    int acc = 64;
#pragma unroll
    for (int i = 0; i < 8; i++)
        acc += F[i] * s[i];
    return clip_pixel(acc >> 7, bpp);
}

/* A reference plane with edge replication: sample(x, y) = pix[clamp(y)][clamp(x)], the exact
 * effect of emulated_edge_mc + direct reads (vp9recon.c:402-412). w,h = the reference FRAME's
 * own plane size in pixels (not the 8-aligned coded size). */
template <class P> struct RefPlane {
    const P *pix;
    int stride;   /* in pixels */
    int w, h;
    VP9_HDM int at(int x, int y) const
    {
        return pix[(size_t)iclamp(y, 0, h - 1) * stride + iclamp(x, 0, w - 1)];
    }
};

/* Position of one prediction call in the reference plane, in 1/16 pel, plus per-pixel steps.
 * Unscaled: step = 16.  (x0,y0) integer sample of output pixel (0,0); (fx,fy) its phase. */
struct McPos {
    int x0, y0, fx, fy, dx, dy;
};

/* mc_luma_unscaled / mc_chroma_unscaled (vp9recon.c:376-467).
 * bx,by: block origin in this plane (pixels); mvx,mvy: VP9mv in 1/8 luma pel;
 * chroma420: plane is subsampled 2x2 (MV unit becomes 1/16 chroma pel). */
VP9_HD McPos mc_pos_unscaled(int bx, int by, int mvx, int mvy, bool chroma420)
{
    McPos p;
    if (chroma420) {
        p.x0 = bx + (mvx >> 4); p.y0 = by + (mvy >> 4);
        p.fx = mvx & 15;        p.fy = mvy & 15;
    } else {
        p.x0 = bx + (mvx >> 3); p.y0 = by + (mvy >> 3);
        p.fx = (mvx & 7) << 1;  p.fy = (mvy & 7) << 1;
    }
    p.dx = p.dy = 16;
    return p;
}

VP9_HD int64_t scale_mv64(int64_t n, int scale) { return (n * scale) >> 14; }

/* mc_luma_scaled / mc_chroma_scaled (vp9recon.c:492-628) incl. libvpx's separate-component
 * rounding.  bx,by: origin of THIS prediction call in the plane; px,py: its offset inside the
 * coding block; pw,ph: coding block size in this plane; cols8/rows8 = s->cols*8, s->rows*8
 * (luma) -- for 4:2:0 chroma the limits use cols*4, rows*4 (vp9recon.c:577,585). */
VP9_HD McPos mc_pos_scaled(int bx, int by, int mvx, int mvy, bool chroma420, int px, int py,
                           int pw, int ph, int cols, int rows, int scale_x, int scale_y,
                           int step_x, int step_y)
{
    McPos p;
    int mx, my;
    if (chroma420) {
        int cx = iclamp(mvx, -(bx + pw - px + 4) * 16, (cols * 4 - bx + px + 3) * 16);
        int cy = iclamp(mvy, -(by + ph - py + 4) * 16, (rows * 4 - by + py + 3) * 16);
        mx = (int)(scale_mv64(cx, scale_x) + (scale_mv64(bx * 16, scale_x) & ~15) +
                   (scale_mv64(bx * 32, scale_x) & 15));
        my = (int)(scale_mv64(cy, scale_y) + (scale_mv64(by * 16, scale_y) & ~15) +
                   (scale_mv64(by * 32, scale_y) & 15));
    } else {
        int cx = iclamp(mvx, -(bx + pw - px + 4) * 8, (cols * 8 - bx + px + 3) * 8);
        int cy = iclamp(mvy, -(by + ph - py + 4) * 8, (rows * 8 - by + py + 3) * 8);
        mx = (int)(scale_mv64(cx * 2, scale_x) + scale_mv64(bx * 16, scale_x));
        my = (int)(scale_mv64(cy * 2, scale_y) + scale_mv64(by * 16, scale_y));
    }
    p.x0 = mx >> 4; p.y0 = my >> 4;
    p.fx = mx & 15; p.fy = my & 15;
    p.dx = step_x;  p.dy = step_y;
    return p;
}

/* Output pixel (ox, oy) of a prediction call: horizontal 8-tap at 8 rows (rounded + clipped to
 * pixel range, like the reference's intermediate buffer), then vertical 8-tap.  Phase-0 taps
 * are the identity, so this also equals the reference's 1-D and copy special cases. */
template <class P>
VP9_HD int mc_pixel(const RefPlane<P> &r, const McPos &p, int ox, int oy, int filter, int bpp)
{
    int ax = p.fx + ox * p.dx, ay = p.fy + oy * p.dy;
    int x = p.x0 + (ax >> 4), y = p.y0 + (ay >> 4);
    int phx = ax & 15, phy = ay & 15;
    int col[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        int s[8];
        if (phx == 0) {
            col[j] = r.at(x, y + j - 3);
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++)
                s[i] = r.at(x + i - 3, y + j - 3);
            col[j] = filt8(s, filter, phx, bpp);
        }
    }
    if (phy == 0)
        return col[3];
    return filt8(col, filter, phy, bpp);
}

/* ------------------------------------------------------------------------------------------ */
/* Intra prediction                                                                           */
/* ------------------------------------------------------------------------------------------ */

/* mode_conv of check_intra_mode (vp9recon.c:49-70): [mode][have_left][have_top] */
VP9_HD int intra_mode_conv(int mode, bool have_left, bool have_top)
{
    switch (mode) {
    case VERT_PRED:
    case DIAG_DOWN_LEFT_PRED:
    case VERT_LEFT_PRED:
        return have_top ? mode : DC_127_PRED;
    case HOR_PRED:
    case HOR_UP_PRED:
        return have_left ? mode : DC_129_PRED;
    case DC_PRED:
        return have_left ? (have_top ? DC_PRED : LEFT_DC_PRED) : (have_top ? TOP_DC_PRED : DC_128_PRED);
    case TM_VP8_PRED:
        return have_left ? (have_top ? TM_VP8_PRED : HOR_PRED) : (have_top ? VERT_PRED : DC_129_PRED);
    default:
        return mode;   /* D135, D117, D153 keep their mode and get synthetic edges */
    }
}

/* which edges a (converted) mode reads: bit0 left, bit1 top, bit2 top-left, bit3 top-right */
VP9_HD int intra_mode_needs(int mode)
{
    switch (mode) {
    case VERT_PRED: case TOP_DC_PRED:              return 2;
    case HOR_PRED: case LEFT_DC_PRED: case HOR_UP_PRED: return 1;
    case DC_PRED:                                  return 3;
    case DIAG_DOWN_LEFT_PRED: case VERT_LEFT_PRED: return 2 | 8;
    case DIAG_DOWN_RIGHT_PRED: case VERT_RIGHT_PRED: case HOR_DOWN_PRED: case TM_VP8_PRED:
        return 1 | 2 | 4;
    default:                                       return 0;
    }
}

/* Edge samples of one transform block in natural orientation:
 *   top[0] = top-left, top[1 + i] = above pixel i (i < 2N for 4x4 with top-right, else N,
 *   replicated up to 2N so every predictor can index freely), left[y] = left pixel of row y.
 * `Px(x, y)` reads the current (pre-loop-filter) plane.  N = 4 << tx.
 *   have_top  : block's y > 0 in the plane               (vp9recon.c:45)
 *   have_left : not on the tile's left edge              (vp9recon.c:46)
 *   have_right: not the last tx column of its coding block (vp9recon.c:47)
 *   avail_w/h : pixels from (bx,by) to the 8-aligned frame edge (n_px_have, :103,192)
 * Values that the chosen mode does not read are still filled deterministically. */
/* top[i] of the edge arrays: i = 0 is the top-left sample, i = 1 + k the k-th sample above */
template <class PX>
VP9_HD int intra_edge_top(const PX &Px, int i, int bx, int by, int N, int needs, bool have_top,
                          bool have_left, bool have_right, int avail_w, int bpp)
{
    const int base = 128 << (bpp - 8);
    if (i == 0)
        return (have_left && have_top) ? Px(bx - 1, by - 1) : base + (have_top ? 1 : -1);
    int k = i - 1;
    if (!have_top)
        return base - 1;
    if (k >= N) {
        /* beyond the block: real top-right only for 4x4 blocks that may read it, else replicate */
        if (!(N == 4 && (needs & 8) && have_right && 8 <= avail_w))
            k = N - 1;
    }
    return Px(bx + (k < avail_w ? k : avail_w - 1), by - 1);
}

/* left[y]: the sample left of row y */
template <class PX>
VP9_HD int intra_edge_left(const PX &Px, int y, int bx, int by, bool have_left, int avail_h, int bpp)
{
    if (!have_left)
        return (128 << (bpp - 8)) + 1;
    return Px(bx - 1, by + (y < avail_h ? y : avail_h - 1));
}

template <class PX>
VP9_HD void intra_edges(const PX &Px, int bx, int by, int N, int mode /* converted */,
                        bool have_top, bool have_left, bool have_right, int avail_w, int avail_h,
                        int bpp, int *top /* [1 + 2N] */, int *left /* [N] */)
{
    int needs = intra_mode_needs(mode);
    if (needs & 2)
        for (int i = 0; i <= 2 * N; i++)
            top[i] = intra_edge_top(Px, i, bx, by, N, needs, have_top, have_left, have_right, avail_w, bpp);
    if (needs & 1)
        for (int y = 0; y < N; y++)
            left[y] = intra_edge_left(Px, y, bx, by, have_left, avail_h, bpp);
}

VP9_HD int avg2(int a, int b) { return (a + b + 1) >> 1; }
VP9_HD int avg3(int a, int b, int c) { return (a + 2 * b + c + 2) >> 2; }

/* DC value for the DC family, computed once per block */
VP9_HD int intra_dc(int mode, int N, const int *top, const int *left, int bpp)
{
    int lg = N == 4 ? 2 : N == 8 ? 3 : N == 16 ? 4 : 5;
    int s = 0;
    switch (mode) {
    case DC_PRED:
        for (int i = 0; i < N; i++) s += top[1 + i] + left[i];
        return (s + N) >> (lg + 1);
    case LEFT_DC_PRED:
        for (int i = 0; i < N; i++) s += left[i];
        return (s + (N >> 1)) >> lg;
    case TOP_DC_PRED:
        for (int i = 0; i < N; i++) s += top[1 + i];
        return (s + (N >> 1)) >> lg;
    case DC_128_PRED: return 128 << (bpp - 8);
    case DC_127_PRED: return (128 << (bpp - 8)) - 1;
    case DC_129_PRED: return (128 << (bpp - 8)) + 1;
    }
    return 0;
}

/* Edge access for the predictors: top(i) = top[i] of intra_edges (i = 0 top-left), left(y).
 * ArrayEdges reads prepared arrays; kernels may plug in an accessor that evaluates
 * intra_edge_top/left on the fly. */
struct ArrayEdges {
    const int *t, *l;
    VP9_HDM int top(int i) const { return t[i]; }
    VP9_HDM int left(int y) const { return l[y]; }
};

/* S(k): the edge walked from bottom-left up the left column, through the top-left corner and
 * along the top row: S(k) = left[N-1-k] (k < N), S(N) = top-left, S(N+1+i) = above[i]. */
template <class E>
VP9_HD int edge_S(int k, int N, const E &e)
{
    return k < N ? e.left(N - 1 - k) : e.top(k - N);
}

/* Predicted sample (x, y) of an N x N block for a converted mode.  Edges as in intra_edges,
 * dc from intra_dc.  Closed forms of vp9dsp_template.c:33-1100. */
template <class E>
VP9_HD int intra_pixel_e(int mode, int N, int x, int y, const E &e, int dc, int bpp)
{
#define A_(i) e.top(1 + (i))                 /* A_(i) = above pixel i, A_(-1) = top-left */
    switch (mode) {
    case VERT_PRED: return A_(x);
    case HOR_PRED:  return e.left(y);
    case TM_VP8_PRED: return clip_pixel(A_(x) + e.left(y) - A_(-1), bpp);
    case DIAG_DOWN_LEFT_PRED: {
        int i = x + y;
        if (N == 4 && i == 6)
            return A_(7);
        return avg3(A_(i), A_(i + 1), A_(i + 2 < 2 * N ? i + 2 : 2 * N - 1));
    }
    case DIAG_DOWN_RIGHT_PRED: {
        int c = N + x - y;
        return avg3(edge_S(c - 1, N, e), edge_S(c, N, e), edge_S(c + 1, N, e));
    }
    case VERT_RIGHT_PRED: {
        int j = y >> 1, k = (N >> 1) - 1 - j + x;
        if (x >= j) {
            if (!(y & 1))
                return avg2(A_(x - j - 1), A_(x - j));
            int c = N + x - j;
            return avg3(edge_S(c - 1, N, e), edge_S(c, N, e), edge_S(c + 1, N, e));
        } else {
            int c = (y & 1) ? 2 * k + 2 : 2 * k + 3;
            return avg3(edge_S(c - 1, N, e), edge_S(c, N, e), edge_S(c + 1, N, e));
        }
    }
    case HOR_DOWN_PRED: {
        int m = 2 * N - 2 - 2 * y + x;
        if (m >= 2 * N) {
            int c = m - N + 1;
            return avg3(edge_S(c - 1, N, e), edge_S(c, N, e), edge_S(c + 1, N, e));
        }
        int i = m >> 1;
        if (!(m & 1))
            return avg2(edge_S(i, N, e), edge_S(i + 1, N, e));
        return avg3(edge_S(i, N, e), edge_S(i + 1, N, e), edge_S(i + 2, N, e));
    }
    case VERT_LEFT_PRED: {
        int k = (y >> 1) + x;
        if (!(y & 1))
            return avg2(A_(k), A_(k + 1));
        return avg3(A_(k), A_(k + 1), A_(k + 2));
    }
    case HOR_UP_PRED: {
        int m = 2 * y + x, i = m >> 1;
        int l0 = e.left(i < N ? i : N - 1), l1 = e.left(i + 1 < N ? i + 1 : N - 1);
        if (!(m & 1))
            return avg2(l0, l1);
        return avg3(l0, l1, e.left(i + 2 < N ? i + 2 : N - 1));
    }
    default:
        return dc;
    }
#undef A_
}

VP9_HD int intra_pixel(int mode, int N, int x, int y, const int *top, const int *left, int dc,
                       int bpp)
{
    ArrayEdges e = { top, left };
    return intra_pixel_e(mode, N, x, y, e, dc, bpp);
}

} /* namespace vp9 */
#endif /* VP9_PRED_H */
