/*
 * k_inter_sb.cuh -- K2 + K1 fused: inter prediction from shared-memory-staged reference windows
 * and the residual of the inter blocks, one CTA per 64x64 superblock.
 *
 *   1. The superblock's inter blocks (found through the 8x8 cell map) are cut into prediction
 *      UNITS of at most 32x16 samples per plane and reference, and bucketed by shape.
 *   2. Warps take batches of same-shaped units.  For each unit the (w+7) x (h+7) reference window
 *      (the reference's own footprint: vp9recon.c:402-412, vp9dsp_template.c:2076-2113) is staged
 *      in the warp's shared-memory arena with 128-bit cp.async copies of whole 16-byte chunks;
 *      windows that cross the reference's edge are gathered sample by sample with clamped
 *      coordinates instead (= emulated_edge_mc, videodsp_template.c:27-105).  Integer-pel motion
 *      vectors stage only the rows / chunks they read.
 *   3. Horizontal 8-tap pass window -> intermediate (rounded and clipped to a pixel exactly like the
 *      reference's temporary buffer), vertical pass intermediate -> the superblock's output tile in
 *      shared memory.  The second reference of a compound block is averaged into the tile in a
 *      second pass ((a + b + 1) >> 1, vp9dsp_template.c:2051).  Scaled references keep the
 *      sliding-window strip of k_inter.cuh (their windows depend on the step) but write the same tile.
 *   4. The inverse transforms of the superblock's inter residual blocks (Vp9CudaSbIndex.tx_*)
 *      are added in the tile (itx_entries, k_itx.cuh), which replaces the separate read-modify-write
 *      pass of k_itx over the picture (vp9recon.c:707-763).
 *   5. The tile goes to the picture with 16-byte stores, inter cells only (intra cells are written
 *      by the wavefront that follows).
 *
 * Replaces inter_pred (vp9_mc_template.c:30-464) + mc_luma/chroma_unscaled / _scaled
 * (vp9recon.c:376-628) + the mc / smc slots (vp9dsp_template.c:1969-2569) + inter_recon's residual
 * loop.  Algorithmic bytes per frame: P (write) + r*P (reference reads) + coefficients.
 */
#ifndef VP9CUDA_K_INTER_SB_CUH
#define VP9CUDA_K_INTER_SB_CUH

#include "dev_common.cuh"
#include "k_inter.cuh"
#include "k_itx.cuh"

namespace vp9 {

#define ISB_WARPS 8
#define ISB_THREADS (ISB_WARPS * 32)
#define ISB_MAX_UNITS 768                  /* 64 blocks of 4x4s: (4 luma + 2 chroma) x 2 references each */
#define ISB_WIN_BYTES 4352                 /* per warp: staged reference windows of the batch in flight */
#define ISB_MID_BYTES 1536                 /* per warp: horizontally filtered rows of the batch */
#define ISB_ARENA (ISB_WIN_BYTES + ISB_MID_BYTES)   /* also holds one 32x33 coefficient block (<= 4224 B) */
#define ISB_NCLS 24                        /* (scaled, log2 w/4, log2 h/4) */
#define ISB_MAX_BATCH 8

struct IUnit {
    uint32_t blk;                          /* index into FrameDesc.blk */
    uint16_t info;                         /* plane | ref << 2 | log2(w/4) << 3 | log2(h/4) << 5 */
    uint8_t ox, oy;                        /* origin inside the superblock's plane, pixels */
};

/* everything the phases of a batch need to know about one unit (per warp, shared memory) */
struct UGeom {
    const void *src;                       /* reference plane */
    int stride, rw, rh;                    /* its stride and size, pixels */
    int wx, wy;                            /* plane position of window sample (0, 0): tap 0 of output (0, 0) */
    McPos pos;                             /* scaled references: position of the whole prediction call ... */
    int sox, soy;                          /* ... and this unit's offset inside it */
    uint16_t dst;                          /* output (0, 0) in the tile, pixels */
    uint8_t dpitch;                        /* tile pitch, pixels */
    uint8_t a;                             /* column of wx in the staged window row (alignment slack) */
    uint8_t fx, fy, filter, interior;
};

__device__ __forceinline__ void isb_cp16(void *smem, const void *gmem) { cp16(smem, gmem); }
__device__ __forceinline__ void isb_cp_commit() { cp_commit(); }
__device__ __forceinline__ void isb_cp_wait_all() { cp_wait<0>(); }

template <class P> struct IsbGeo {
    static constexpr int PS = (int)sizeof(P);
    static constexpr int APX = 16 / PS;                               /* pixels per 16-byte chunk */
    /* 16-byte chunks per window row: (w + 7) samples at the worst alignment */
    static __device__ __forceinline__ int chunks(int w) { return ((w + 6 + APX) * PS + 15) >> 4; }
    /* units of shape (w, h) one warp handles at a time */
    static __device__ __forceinline__ int batch(int w, int h, bool scaled)
    {
        int k = 32 / w;
        if (!scaled) {
            const int nch = chunks(w), rows = h + 7;
            k = min(k, ISB_WIN_BYTES / (rows * nch * 16));
            k = min(k, ISB_MID_BYTES / (rows * w * PS));
            k = min(k, 32 / nch);
        }
        return min(k, ISB_MAX_BATCH);
    }
};

/* ---- 16-bit pixels: two outputs per lane, 8-tap filters as four packed dot products ------------------------
 * The 8-tap kernels fit signed bytes for every phase but 0 (which is never filtered: fx / fy == 0 copies), so an
 * output is 64 + sum over four (pixel pair) x (tap pair) dp2a, with pixel pairs taken as 32-bit words from shared
 * memory.  Same arithmetic as filt8 (vp9_pred.h): (sum + 64) >> 7, clipped. */
__device__ __forceinline__ void taps_packed(int filter, int phase, int &t0, int &t1)
{
    const int16_t *F = subpel_taps(filter, phase);
    t0 = (int)((unsigned)(F[0] & 0xff) | (unsigned)(F[1] & 0xff) << 8 | (unsigned)(F[2] & 0xff) << 16 | (unsigned)(F[3] & 0xff) << 24);
    t1 = (int)((unsigned)(F[4] & 0xff) | (unsigned)(F[5] & 0xff) << 8 | (unsigned)(F[6] & 0xff) << 16 | (unsigned)(F[7] & 0xff) << 24);
}
__device__ __forceinline__ int filt8_pairs(unsigned p0, unsigned p1, unsigned p2, unsigned p3, int t0, int t1, int bpp)
{
    int acc = 64;
    acc = __dp2a_lo((int)p0, t0, acc);
    acc = __dp2a_hi((int)p1, t0, acc);
    acc = __dp2a_lo((int)p2, t1, acc);
    acc = __dp2a_hi((int)p3, t1, acc);
    return clip_pixel(acc >> 7, bpp);
}

/* horizontal pass of one unit for the lane (row parity rp, output pair j): rows r0 + rp, r0 + rp + 2, ... < r1 */
__device__ __forceinline__ void mc_h_pairs(const unsigned char *wb, int pitch, int a, int j, int rp, int r0, int r1, int w,
                                           uint16_t *m, int filter, int fx, int bpp)
{
    const int q = a + 2 * j, wi = q >> 1, odd = q & 1;
    unsigned *m32 = (unsigned *)m + j;
    const int hw = w >> 1;
    if (fx) {
        int t0, t1;
        taps_packed(filter, fx, t0, t1);
        for (int r = r0 + rp; r < r1; r += 2) {
            const unsigned *row = (const unsigned *)(wb + r * pitch) + wi;
            const unsigned w0 = row[0], w1 = row[1], w2 = row[2], w3 = row[3], w4 = row[4];
            const unsigned f0 = __funnelshift_r(w0, w1, 16), f1 = __funnelshift_r(w1, w2, 16),
                           f2 = __funnelshift_r(w2, w3, 16), f3 = __funnelshift_r(w3, w4, 16);
            /* output 2j starts at pixel q, output 2j + 1 at q + 1; an odd q sits in the high half of its word */
            const int v0 = odd ? filt8_pairs(f0, f1, f2, f3, t0, t1, bpp) : filt8_pairs(w0, w1, w2, w3, t0, t1, bpp);
            const int v1 = odd ? filt8_pairs(w1, w2, w3, w4, t0, t1, bpp) : filt8_pairs(f0, f1, f2, f3, t0, t1, bpp);
            m32[r * hw] = (unsigned)v0 | (unsigned)v1 << 16;
        }
    } else {
        for (int r = r0 + rp; r < r1; r += 2) {
            const uint16_t *s = (const uint16_t *)(wb + r * pitch) + q;
            m32[r * hw] = (unsigned)s[3] | (unsigned)s[4] << 16;
        }
    }
}

/* vertical pass: output rows rp, rp + 2, ... < h of columns 2j, 2j + 1, written (or averaged) into the tile */
__device__ __forceinline__ void mc_v_pairs(const uint16_t *m, int j, int rp, int w, int h, uint16_t *dst, int dp, int filter,
                                           int fy, bool avg, int bpp)
{
    const unsigned *m32 = (const unsigned *)m + j;
    const int hw = w >> 1, dw = dp >> 1;
    unsigned *d32 = (unsigned *)dst + j;
    auto put = [&](int y, int v0, int v1) {
        if (avg) {
            const unsigned o = d32[y * dw];
            v0 = ((int)(o & 0xffff) + v0 + 1) >> 1;
            v1 = ((int)(o >> 16) + v1 + 1) >> 1;
        }
        d32[y * dw] = (unsigned)v0 | (unsigned)v1 << 16;
    };
    if (fy) {
        int t0, t1;
        taps_packed(filter, fy, t0, t1);
        /* vertically adjacent samples packed into words: (row, row + 1) of the low column and of the high column */
        int y = rp;
        unsigned a0 = m32[(y + 0) * hw], a1 = m32[(y + 1) * hw], a2 = m32[(y + 2) * hw], a3 = m32[(y + 3) * hw],
                 a4 = m32[(y + 4) * hw], a5 = m32[(y + 5) * hw];
        unsigned l0 = __byte_perm(a0, a1, 0x5410), h0 = __byte_perm(a0, a1, 0x7632);
        unsigned l1 = __byte_perm(a2, a3, 0x5410), h1 = __byte_perm(a2, a3, 0x7632);
        unsigned l2 = __byte_perm(a4, a5, 0x5410), h2 = __byte_perm(a4, a5, 0x7632);
        for (; y < h; y += 2) {
            const unsigned b0 = m32[(y + 6) * hw], b1 = m32[(y + 7) * hw];
            const unsigned l3 = __byte_perm(b0, b1, 0x5410), h3 = __byte_perm(b0, b1, 0x7632);
            put(y, filt8_pairs(l0, l1, l2, l3, t0, t1, bpp), filt8_pairs(h0, h1, h2, h3, t0, t1, bpp));
            l0 = l1; l1 = l2; l2 = l3;
            h0 = h1; h1 = h2; h2 = h3;
        }
    } else {
        for (int y = rp; y < h; y += 2) {
            const unsigned v = m32[(y + 3) * hw];
            put(y, (int)(v & 0xffff), (int)(v >> 16));
        }
    }
}

/* Unscaled prediction of `cnt` units of w x h samples (w = 4 << wl <= 32, h = 4 << hl <= 16) by one
 * warp.  geom[0 .. cnt): the units; arena: ISB_ARENA bytes, 16-byte aligned, owned by the warp;
 * tile: base of the output (any address space); avg: average into the tile instead of storing. */
template <class P>
__device__ __forceinline__ void mc_batch(const UGeom *geom, int cnt, int wl, int hl, unsigned char *arena, P *tile,
                                         bool avg, int bpp)
{
    typedef IsbGeo<P> G;
    const int lane = threadIdx.x & 31;
    const int w = 4 << wl, h = 4 << hl, rows = h + 7;
    const int nch = G::chunks(w), pitch = nch * 16, wbytes = rows * pitch;
    unsigned char *win = arena;
    P *mid = (P *)(arena + ISB_WIN_BYTES);
    /* ---- stage the windows: lanes = (row group q, unit us, chunk j) ---- */
    {
        const int per = cnt * nch, rg = 32 / per;
        if (lane < rg * per) {
            const int q = lane / per, rem = lane - q * per, us = rem / nch, j = rem - us * nch;
            const UGeom &g = geom[us];
            if (g.interior) {
                /* integer-pel vectors read one row / column per output: stage only those */
                const int r0 = g.fy ? 0 : 3, r1 = g.fy ? rows : h + 3;
                const int j0 = g.fx ? 0 : (g.a + 3) / G::APX;
                const int j1 = g.fx ? (g.a + w + 6) / G::APX : (g.a + w + 2) / G::APX;
                if (j >= j0 && j <= j1) {
                    const P *src = (const P *)g.src + (size_t)g.wy * g.stride + (g.wx - g.a) + j * G::APX;
                    unsigned char *dst = win + us * wbytes + 16 * j;
                    for (int r = r0 + q; r < r1; r += rg)
                        isb_cp16(dst + r * pitch, src + (size_t)r * g.stride);
                }
            }
        }
        isb_cp_commit();
        for (int us = 0; us < cnt; us++) {
            const UGeom &g = geom[us];
            if (g.interior)
                continue;
            RefPlane<P> rp;
            rp.pix = (const P *)g.src; rp.stride = g.stride; rp.w = g.rw; rp.h = g.rh;
            const int ww = w + 7;
            for (int i = lane; i < rows * ww; i += 32) {
                const int r = i / ww, c = i - r * ww;
                ((P *)(win + us * wbytes + r * pitch))[c] = (P)rp.at(g.wx + c, g.wy + r);
            }
        }
        isb_cp_wait_all();
        __syncwarp();
    }
    /* ---- lanes = (unit u, output column c); 16-bit pixels under an 8-tap filter: (unit u, row parity, column pair) ---- */
    const int u = lane >> (2 + wl), c = lane & (w - 1);
    const bool act = u < cnt;
    const bool pairs = sizeof(P) == 2 && act && geom[u].filter != FILTER_BILINEAR;
    const int prp = c >> (1 + wl), pj = c & ((w >> 1) - 1);
    if (pairs) {
        const UGeom &g = geom[u];
        mc_h_pairs(win + u * wbytes, pitch, g.a, pj, prp, g.fy ? 0 : 3, g.fy ? rows : h + 3, w,
                   (uint16_t *)(mid + u * rows * w), g.filter, g.fx, bpp);
    } else if (act) {
        const UGeom &g = geom[u];
        const unsigned char *wb = win + u * wbytes;
        P *m = mid + u * rows * w;
        const int r0 = g.fy ? 0 : 3, r1 = g.fy ? rows : h + 3;
        if (g.fx) {
            for (int r = r0; r < r1; r++) {
                const P *s = (const P *)(wb + r * pitch) + g.a + c;
                int t[8];
#pragma unroll
                for (int k = 0; k < 8; k++) t[k] = s[k];
                m[r * w + c] = (P)filt8(t, g.filter, g.fx, bpp);
            }
        } else {
            for (int r = r0; r < r1; r++)
                m[r * w + c] = ((const P *)(wb + r * pitch))[g.a + c + 3];
        }
    }
    __syncwarp();
    if (pairs) {
        const UGeom &g = geom[u];
        mc_v_pairs((const uint16_t *)(mid + u * rows * w), pj, prp, w, h, (uint16_t *)(tile + g.dst), g.dpitch, g.filter, g.fy, avg, bpp);
    } else if (act) {
        const UGeom &g = geom[u];
        const P *m = mid + u * rows * w + c;
        P *dst = tile + g.dst + c;
        const int dp = g.dpitch;
        if (g.fy) {
            int t[8];
#pragma unroll
            for (int k = 0; k < 7; k++) t[k] = m[k * w];
            for (int y = 0; y < h; y++) {
                t[7] = m[(y + 7) * w];
                const int v = filt8(t, g.filter, g.fy, bpp);
                dst[y * dp] = (P)(avg ? (dst[y * dp] + v + 1) >> 1 : v);
#pragma unroll
                for (int k = 0; k < 7; k++) t[k] = t[k + 1];
            }
        } else {
            for (int y = 0; y < h; y++) {
                const int v = m[(y + 3) * w];
                dst[y * dp] = (P)(avg ? (dst[y * dp] + v + 1) >> 1 : v);
            }
        }
    }
    __syncwarp();                              /* the arena is rewritten by the warp's next batch */
}

/* Scaled references: the same lane mapping, each lane runs the sliding-window strip of k_inter.cuh
 * down its column, four rows at a time. */
template <class P>
__device__ __forceinline__ void mc_batch_scaled(const UGeom *geom, int cnt, int wl, int hl, P *tile, bool avg, int bpp)
{
    const int lane = threadIdx.x & 31;
    const int w = 4 << wl, h = 4 << hl;
    const int u = lane >> (2 + wl), c = lane & (w - 1);
    if (u < cnt) {
        const UGeom &g = geom[u];
        RefPlane<P> rp;
        rp.pix = (const P *)g.src; rp.stride = g.stride; rp.w = g.rw; rp.h = g.rh;
        P *dst = tile + g.dst + c;
        const int dp = g.dpitch;
        for (int y0 = 0; y0 < h; y0 += 4) {
            int out[4];
            mc_strip_scaled<P, 4>(rp, g.pos, g.sox + c, g.soy + y0, g.filter, bpp, out);
#pragma unroll
            for (int j = 0; j < 4; j++) {
                P *d = dst + (y0 + j) * dp;
                *d = (P)(avg ? (*d + out[j] + 1) >> 1 : out[j]);
            }
        }
    }
    __syncwarp();
}

/* geometry of one unit of superblock (sbx, sby) */
template <class P>
__device__ __forceinline__ void isb_unit_geom(const FrameDesc &f, const IUnit &un, int sbx, int sby, UGeom &g)
{
    typedef IsbGeo<P> G;
    const Vp9CudaInterBlock &b = f.blk[un.blk];
    const int pl = un.info & 3, i = un.info >> 2 & 1, wl = un.info >> 3 & 3, hl = un.info >> 5 & 3;
    const int ch = pl > 0, S = 64 >> ch;
    const int X = sbx * S + un.ox, Y = sby * S + un.oy;
    const InterCall c = inter_call(b, i, pl, X, Y);
    const int ri = b.ref[i];
    const SurfDesc &rs = f.ref[ri];
    g.src = rs.pl[pl]; g.stride = rs.stride[pl];
    g.rw = (rs.w + ch) >> ch; g.rh = (rs.h + ch) >> ch;
    g.filter = b.filter;
    g.dst = (uint16_t)((pl == 0 ? 0 : pl == 1 ? 4096 : 5120) + un.oy * S + un.ox);
    g.dpitch = (uint8_t)S;
    if (f.mvscale[ri][0]) {
        g.pos = mc_pos_scaled(c.ox0, c.oy0, c.mvx, c.mvy, ch, c.px, c.py, c.pw, c.ph, f.cols, f.rows,
                              f.mvscale[ri][0], f.mvscale[ri][1], f.mvstep[ri][0], f.mvstep[ri][1]);
        g.sox = X - c.ox0; g.soy = Y - c.oy0;
        g.wx = g.wy = 0; g.a = g.fx = g.fy = g.interior = 0;
    } else {
        const McPos p = mc_pos_unscaled(X, Y, c.mvx, c.mvy, ch);
        const int w = 4 << wl, h = 4 << hl;
        g.fx = (uint8_t)p.fx; g.fy = (uint8_t)p.fy;
        g.wx = p.x0 - 3; g.wy = p.y0 - 3;
        g.interior = g.wx >= 0 && g.wy >= 0 && g.wx + w + 7 <= g.rw && g.wy + h + 7 <= g.rh;
        g.a = (uint8_t)(g.interior ? g.wx & (G::APX - 1) : 0);
        g.pos = p; g.sox = g.soy = 0;
    }
}

template <class P> struct IsbSmem {
    static constexpr int O_TILE = 0;                                       /* Y 64x64, U 32x32, V 32x32 */
    static constexpr int O_ARENA = 6144 * (int)sizeof(P);
    static constexpr int O_UNITS = O_ARENA + ISB_WARPS * ISB_ARENA;
    static constexpr int O_GEOM = O_UNITS + ISB_MAX_UNITS * (int)sizeof(IUnit);
    static constexpr int BYTES = O_GEOM + ISB_WARPS * ISB_MAX_BATCH * (int)sizeof(UGeom);
};

/* units of the inter block whose top-left 8x8 cell is (cx, cy) of the superblock; emit(pass, cls, unit) */
template <class F>
__device__ __forceinline__ void isb_block_units(const FrameDesc &f, uint32_t bi, int cx, int cy, int sbx, int sby, F emit)
{
    const Vp9CudaInterBlock &b = f.blk[bi];
    const int nref = b.comp ? 2 : 1;
    for (int i = 0; i < nref; i++) {
        const int sc = f.mvscale[b.ref[i]][0] ? 12 : 0;
        for (int pl = 0; pl < (f.mono >= 0 ? 1 : 3); pl++) {
            const int ch = pl > 0, S = 64 >> ch;
            const int lx = (cx * 8) >> ch, ly = (cy * 8) >> ch;
            IUnit un;
            un.blk = bi;
            if (b.bs > 9) {                          /* sub-8x8: four luma 4x4 with their own vectors, one chroma 4x4 */
                un.info = (uint16_t)(pl | i << 2);
                const int n = ch ? 1 : 4;
                for (int k = 0; k < n; k++) {
                    un.ox = (uint8_t)(lx + 4 * (k & 1)); un.oy = (uint8_t)(ly + 4 * (k >> 1));
                    emit(i, sc, un);
                }
                continue;
            }
            const int W = (bs_w8(b.bs) * 8) >> ch, H = (bs_h8(b.bs) * 8) >> ch;
            const int pw = min(W, 32), ph = min(H, 16);
            const int wl = pw == 4 ? 0 : pw == 8 ? 1 : pw == 16 ? 2 : 3, hl = ph == 4 ? 0 : ph == 8 ? 1 : 2;
            /* pieces that start beyond the frame's 8-aligned area are not predicted (k_inter.cuh does the same) */
            const int vis_w = ((f.cols * 8) >> ch) - sbx * S, vis_h = ((f.rows * 8) >> ch) - sby * S;
            un.info = (uint16_t)(pl | i << 2 | wl << 3 | hl << 5);
            for (int py = 0; py < H; py += ph)
                for (int px = 0; px < W; px += pw) {
                    if (lx + px >= vis_w || ly + py >= vis_h)
                        continue;
                    un.ox = (uint8_t)(lx + px); un.oy = (uint8_t)(ly + py);
                    emit(i, sc + wl * 3 + hl, un);
                }
        }
    }
}

template <class P, bool HBD>
__global__ void __launch_bounds__(ISB_THREADS, 3) k_inter_sb(const __grid_constant__ FrameDesc f, const int fuse_residual)
{
    typedef IsbSmem<P> L;
    typedef typename ItxTypes<HBD>::C C;
    typedef typename ItxTypes<HBD>::E E;
    VP9_DYN_SMEM(sm);
    __shared__ int s_cell[64];
    __shared__ int s_cnt[2][ISB_NCLS];
    __shared__ int s_base[2][ISB_NCLS + 1];
    P *tile = (P *)(sm + L::O_TILE);
    IUnit *units = (IUnit *)(sm + L::O_UNITS);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int sbx = blockIdx.x, sby = blockIdx.y + f.tile_y0;       /* tile_y0: first superblock row of a banded launch */
    unsigned char *arena = sm + L::O_ARENA + warp * ISB_ARENA;
    UGeom *geom = (UGeom *)(sm + L::O_GEOM) + warp * ISB_MAX_BATCH;

    if (tid < 64) {
        const int cx = sbx * 8 + (tid & 7), cy = sby * 8 + (tid >> 3);
        s_cell[tid] = (cx < f.cols && cy < f.rows) ? f.cellmap[cy * f.cols + cx] : -1;
    }
    if (tid < 2 * ISB_NCLS)
        (&s_cnt[0][0])[tid] = 0;
    __syncthreads();
    /* ---- the superblock's units, bucketed by (pass = reference index, shape class) ---- */
    bool head = false;
    int ci = -1;
    if (tid < 64) {
        ci = s_cell[tid];
        if (ci >= 0)
            head = f.blk[ci].col == sbx * 8 + (tid & 7) && f.blk[ci].row == sby * 8 + (tid >> 3);
    }
    if (head)
        isb_block_units(f, (uint32_t)ci, tid & 7, tid >> 3, sbx, sby,
                        [&](int pass, int cls, const IUnit &) { atomicAdd(&s_cnt[pass][cls], 1); });
    __syncthreads();
    if (tid == 0) {
        int acc = 0;
        for (int p = 0; p < 2; p++) {
            for (int k = 0; k < ISB_NCLS; k++) { s_base[p][k] = acc; acc += s_cnt[p][k]; }
            s_base[p][ISB_NCLS] = acc;
        }
    }
    __syncthreads();
    const int total = s_base[1][ISB_NCLS];
    if (total == 0 || total > ISB_MAX_UNITS)
        return;                                   /* no inter block in this superblock */
    if (tid < 2 * ISB_NCLS)
        (&s_cnt[0][0])[tid] = 0;
    __syncthreads();
    if (head)
        isb_block_units(f, (uint32_t)ci, tid & 7, tid >> 3, sbx, sby,
                        [&](int pass, int cls, const IUnit &un) { units[s_base[pass][cls] + atomicAdd(&s_cnt[pass][cls], 1)] = un; });
    __syncthreads();

    /* ---- prediction: pass 0 stores, pass 1 (second reference of compound blocks) averages ---- */
    for (int pass = 0; pass < 2; pass++) {
        int g0 = 0;                               /* batches handed out so far: batch g goes to warp g % ISB_WARPS */
        for (int cls = 0; cls < ISB_NCLS; cls++) {
            const int first = s_base[pass][cls], n = (cls == ISB_NCLS - 1 ? s_base[pass][ISB_NCLS] : s_base[pass][cls + 1]) - first;
            if (n == 0)
                continue;
            const bool scaled = cls >= 12;
            const int wl = (cls % 12) / 3, hl = cls % 3;
            const int K = IsbGeo<P>::batch(4 << wl, 4 << hl, scaled);
            const int nb = (n + K - 1) / K;
            for (int b = (warp - g0) & (ISB_WARPS - 1); b < nb; b += ISB_WARPS) {
                const int cnt = min(K, n - b * K);
                if (lane < cnt)
                    isb_unit_geom<P>(f, units[first + b * K + lane], sbx, sby, geom[lane]);
                __syncwarp();
                if (scaled)
                    mc_batch_scaled<P>(geom, cnt, wl, hl, tile, pass != 0, f.bpp);
                else
                    mc_batch<P>(geom, cnt, wl, hl, arena, tile, pass != 0, f.bpp);
            }
            g0 += nb;
        }
        __syncthreads();
    }

    /* ---- residual of the inter blocks, added in the tile ---- */
    if (fuse_residual && f.n_tx) {
        const Vp9CudaSbIndex si = f.sb_index[sby * f.sb_cols + sbx];
        if (si.tx_count) {
            const Vp9CudaTx *tx = f.tx + f.lane_tx_base[si.lane] + si.tx_start;
            C *buf = (C *)arena;
            for (int k = warp; k < si.tx_count; k += ISB_WARPS) {
                Vp9CudaTx r = tx[k];
                if (f.mono >= 0) {                 /* one plane of a 4:4:4 frame: its blocks only, as plane 0 */
                    if (r.plane != f.mono)
                        continue;
                    r.plane = 0;
                }
                const int S = r.plane ? 32 : 64;
                GlobalPix<P> pix;
                pix.stride = S;
                pix.p = tile + (r.plane == 0 ? 0 : r.plane == 1 ? 4096 : 5120) + (r.y4 * 4 - sby * S) * S + (r.x4 * 4 - sbx * S);
                const E *ent = (const E *)f.coef + r.coef_off;
                switch (r.tx) {
                case 0: if (lane < 4)  itx_entries<HBD, 4>(ent, r.ncoef, buf, buf, lane, r.txtp, r.eob, f.bpp, pix, 0xfu); break;
                case 1: if (lane < 8)  itx_entries<HBD, 8>(ent, r.ncoef, buf, buf, lane, r.txtp, r.eob, f.bpp, pix, 0xffu); break;
                case 2: if (lane < 16) itx_entries<HBD, 16>(ent, r.ncoef, buf, buf, lane, r.txtp, r.eob, f.bpp, pix, 0xffffu); break;
                default: itx_entries<HBD, 32>(ent, r.ncoef, buf, buf, lane, r.txtp, r.eob, f.bpp, pix, 0xffffffffu); break;
                }
                __syncwarp();
            }
        }
        __syncthreads();
    }

    /* ---- tile -> picture, inter cells only ---- */
    {
        constexpr int PS = (int)sizeof(P);
        P *const py = (P *)f.cur.pl[0] + (size_t)sby * 64 * f.cur.stride[0] + sbx * 64;
        for (int i = tid; i < 512; i += ISB_THREADS) {            /* luma: 64 rows x 8 cells of 8 samples */
            const int r = i >> 3, cxl = i & 7;
            if (s_cell[(r >> 3) * 8 + cxl] < 0)
                continue;
            const P *s = tile + r * 64 + cxl * 8;
            P *d = py + (size_t)r * f.cur.stride[0] + cxl * 8;
            if (PS == 2) *(uint4 *)d = *(const uint4 *)s;
            else         *(uint2 *)d = *(const uint2 *)s;
        }
        for (int i = tid; i < (f.mono >= 0 ? 0 : 512); i += ISB_THREADS) {   /* chroma: 2 planes x 32 rows x 8 cells of 4 samples */
            const int pl = 1 + (i >> 8), r = i >> 3 & 31, cxl = i & 7;
            if (s_cell[(r >> 2) * 8 + cxl] < 0)
                continue;
            const P *s = tile + (pl == 1 ? 4096 : 5120) + r * 32 + cxl * 4;
            P *d = (P *)f.cur.pl[pl] + ((size_t)sby * 32 + r) * f.cur.stride[pl] + sbx * 32 + cxl * 4;
            if (PS == 2) *(uint2 *)d = *(const uint2 *)s;
            else         *(uint32_t *)d = *(const uint32_t *)s;
        }
    }
}

} /* namespace vp9 */
#endif
