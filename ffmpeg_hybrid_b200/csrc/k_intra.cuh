/*
 * k_intra.cuh -- K3: intra prediction + residual as a superblock wavefront.
 *
 *   k_intra_sched : one thread per (superblock, plane).  Walks the superblock's intra
 *                   transform-block records in coding order and gives each the depth of its
 *                   dependency chain INSIDE the superblock (left column, top-left, top row and,
 *                   for 4x4 blocks that may read it, top-right: vp9recon.c:45-47,96-98), then
 *                   counting-sorts the plane's records by that level.
 *   k_intra       : one CTA per (superblock row, plane), walking the row left to right.  A
 *                   superblock starts when the one above it is finished (its left neighbour
 *                   was done by this same CTA; tile-column starts and top-right never reach
 *                   outside the superblock, vp9recon.c:46-47), which makes the frame a
 *                   (left, top) wavefront across CTAs.  The superblock's pixels live in shared
 *                   memory while its levels are processed: all blocks of one level run in
 *                   parallel, one warp per transform block (edge build, per-pixel predictor,
 *                   then the same itx_block as K1), with a CTA barrier between levels.
 *
 * Replaces intra_recon + check_intra_mode (vp9recon.c:37-364) and the intra_pred slots
 * (vp9dsp_template.c:33-1153).  Reads the pre-loop-filter picture everywhere, which is what
 * the reference gets from intra_pred_data at superblock-row tops (vp9.c:1404-1416).
 * Algorithmic bytes: P (write) + coefficients.
 */
#ifndef VP9CUDA_K_INTRA_CUH
#define VP9CUDA_K_INTRA_CUH

#include "dev_common.cuh"
#include "k_itx.cuh"

namespace vp9 {

#define ISCHED_MAX_LVL 64

__global__ void k_intra_sched(const __grid_constant__ FrameDesc f)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int n_sb = f.sb_cols * f.sb_rows;
    if (t >= n_sb * 3)
        return;
    const int sb = t / 3, pl = t % 3;
    const int ch = pl > 0;
    const Vp9CudaSbIndex si = f.sb_index[sb];
    const Vp9CudaIntraTx *rec = f.itx + f.itx_lane_base[si.lane] + si.itx_start;
    const int sbx4 = ((sb % f.sb_cols) * 16) >> ch, sby4 = ((sb / f.sb_cols) * 16) >> ch;
    const int S4 = 16 >> ch;                       /* superblock size in 4-pixel cells */
    int8_t cell[16][16];
    for (int y = 0; y < S4; y++)
        for (int x = 0; x < S4; x++)
            cell[y][x] = -1;
    uint16_t cnt[ISCHED_MAX_LVL];
    for (int i = 0; i < ISCHED_MAX_LVL; i++) cnt[i] = 0;
    /* segment of this plane = after the records of the lower planes of this superblock */
    uint32_t seg = f.itx_lane_base[si.lane] + si.itx_start;
    for (int k = 0; k < si.itx_count; k++)
        if (rec[k].plane < pl)
            seg++;
    /* pass 1: levels */
    uint8_t lv_local[256];
    int nl = 0, j = 0;
    for (int k = 0; k < si.itx_count; k++) {
        const Vp9CudaIntraTx r = rec[k];
        if (r.plane != pl)
            continue;
        const int n = 1 << r.tx;
        const int cx = r.x4 - sbx4, cy = r.y4 - sby4;
        int lv = 0;
        if (cx > 0) {
            for (int y = max(cy - 1, 0); y < min(cy + n, S4); y++)
                lv = max(lv, cell[y][cx - 1] + 1);
        }
        if (cy > 0) {
            int xe = cx + n;
            if (r.tx == 0 && (r.flags & VP9CUDA_ITX_HAVE_RIGHT))
                xe = cx + 2;
            for (int x = cx; x < min(xe, S4); x++)
                lv = max(lv, cell[cy - 1][x] + 1);
        }
        for (int y = cy; y < min(cy + n, S4); y++)
            for (int x = cx; x < min(cx + n, S4); x++)
                cell[y][x] = (int8_t)lv;
        cnt[lv]++;
        nl = max(nl, lv + 1);
        lv_local[j++] = (uint8_t)lv;
    }
    /* pass 2: counting sort */
    uint16_t start[ISCHED_MAX_LVL];
    uint16_t acc = 0;
    uint16_t *lvl_end = f.isched_lvl_end + (size_t)t * ISCHED_MAX_LVL;
    for (int i = 0; i < ISCHED_MAX_LVL; i++) {
        start[i] = acc;
        acc += cnt[i];
        lvl_end[i] = acc;
    }
    j = 0;
    for (int k = 0; k < si.itx_count; k++) {
        if (rec[k].plane != pl)
            continue;
        int lv = lv_local[j++];
        f.isched_order[seg + start[lv]++] = (uint16_t)k;
    }
    f.isched_seg[t] = seg;
    f.isched_nlvl[t] = (uint8_t)nl;
}

/* superblock tile in shared memory, with one row above and one column to the left */
template <class P, int S> struct IntraTile {
    P px[(S + 1) * (S + 1)];
    __device__ __forceinline__ P &at(int x, int y) { return px[(y + 1) * (S + 1) + x + 1]; }   /* x,y >= -1 */
};

template <class P, int S> struct TilePx {
    IntraTile<P, S> *t; int ox, oy;          /* plane coords of tile (0,0) */
    __device__ __forceinline__ int operator()(int x, int y) const { return t->at(x - ox, y - oy); }
};
template <class P, int S> struct TileBlk {
    IntraTile<P, S> *t; int bx, by;          /* tile coords of the block origin */
    __device__ __forceinline__ int get(int x, int y) const { return t->at(bx + x, by + y); }
    __device__ __forceinline__ void put(int x, int y, int v) { t->at(bx + x, by + y) = (P)v; }
};

#define INTRA_WARPS 8

template <class P, bool HBD, int S>
__device__ void intra_superblock(const FrameDesc &f, int pl, int sb, IntraTile<P, S> *tile,
                                 typename ItxTypes<HBD>::C *tmp_w, int *edge_w)
{
    typedef typename ItxTypes<HBD>::C C;
    const int ch = pl > 0;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t = sb * 3 + pl;
    const int nl = f.isched_nlvl[t];
    const uint32_t seg = f.isched_seg[t];
    const uint16_t *lvl_end = f.isched_lvl_end + (size_t)t * ISCHED_MAX_LVL;
    const Vp9CudaSbIndex si = f.sb_index[sb];
    const Vp9CudaIntraTx *rec = f.itx + f.itx_lane_base[si.lane] + si.itx_start;
    const int ox = ((sb % f.sb_cols) * 64) >> ch, oy = ((sb / f.sb_cols) * 64) >> ch;
    const int plane_w = (f.cols * 8) >> ch, plane_h = (f.rows * 8) >> ch;
    const int stride = f.cur.stride[pl];
    P *plane = (P *)f.cur.pl[pl];

    /* load the tile + its top row and left column (L2 reads: neighbours come from other SMs) */
    for (int i = threadIdx.x; i < (S + 1) * (S + 1); i += blockDim.x) {
        int x = i % (S + 1) - 1, y = i / (S + 1) - 1;
        int gx = ox + x, gy = oy + y;
        P v = 0;
        if (gx >= 0 && gy >= 0)
            v = ld_cg(plane + (size_t)gy * stride + gx);
        tile->at(x, y) = v;
    }
    __syncthreads();

    int *top = edge_w + warp * 100, *left = top + 66;
    C *tmp = tmp_w + warp * (32 * 33);
    int begin = 0;
    for (int L = 0; L < nl; L++) {
        const int end = lvl_end[L];
        for (int k = begin + warp; k < end; k += INTRA_WARPS) {
            const Vp9CudaIntraTx r = rec[f.isched_order[seg + k]];
            const int N = 4 << r.tx;
            const int bx = r.x4 * 4, by = r.y4 * 4;
            const bool have_top = by > 0, have_left = r.flags & VP9CUDA_ITX_HAVE_LEFT;
            const int mode = intra_mode_conv(r.mode, have_left, have_top);
            TilePx<P, S> px = { tile, ox, oy };
            /* edges: built redundantly by lane 0 into shared memory (small, keeps semantics 1:1) */
            if (lane == 0)
                intra_edges(px, bx, by, N, mode, have_top, have_left, r.flags & VP9CUDA_ITX_HAVE_RIGHT,
                            plane_w - bx, plane_h - by, f.bpp, top, left);
            __syncwarp();
            int dc = intra_dc(mode, N, top, left, f.bpp);
            TileBlk<P, S> blk = { tile, bx - ox, by - oy };
            for (int i = lane; i < N * N; i += 32)
                blk.put(i % N, i / N, intra_pixel(mode, N, i % N, i / N, top, left, dc, f.bpp));
            __syncwarp();
            if (r.eob) {
                const C *coef = (const C *)f.coef + r.coef_off;
                switch (r.tx) {
                case 0: if (lane < 4)  itx_block<HBD, 4>(coef, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xfu); break;
                case 1: if (lane < 8)  itx_block<HBD, 8>(coef, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xffu); break;
                case 2: if (lane < 16) itx_block<HBD, 16>(coef, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xffffu); break;
                default: itx_block<HBD, 32>(coef, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xffffffffu); break;
                }
            }
            __syncwarp();
        }
        begin = end;
        __syncthreads();
    }
    /* write the superblock back (interior only) */
    for (int i = threadIdx.x; i < S * S; i += blockDim.x) {
        int x = i % S, y = i / S;
        plane[(size_t)(oy + y) * stride + ox + x] = tile->at(x, y);
    }
}

template <class P, bool HBD>
__global__ void __launch_bounds__(INTRA_WARPS * 32) k_intra(const __grid_constant__ FrameDesc f)
{
    typedef typename ItxTypes<HBD>::C C;
    __shared__ IntraTile<P, 64> tile_mem;
    __shared__ C tmp_w[INTRA_WARPS * 32 * 33];
    __shared__ int edge_w[INTRA_WARPS * 100];
    const int pl = blockIdx.x % 3, row = blockIdx.x / 3;
    int *prog = f.progress + pl * f.sb_rows;           /* stage 0 */
    for (int c = 0; c < f.sb_cols; c++) {
        const int sb = row * f.sb_cols + c;
        if (f.isched_nlvl[sb * 3 + pl]) {
            if (row > 0 && threadIdx.x == 0) {
                volatile int *above = prog + row - 1;
                while (*above < c + 1)
                    __nanosleep(100);
                __threadfence();
            }
            __syncthreads();
            if (pl == 0)
                intra_superblock<P, HBD, 64>(f, pl, sb, &tile_mem, tmp_w, edge_w);
            else
                intra_superblock<P, HBD, 32>(f, pl, sb, (IntraTile<P, 32> *)&tile_mem, tmp_w, edge_w);
            __threadfence();
            __syncthreads();
        }
        /* a superblock without intra blocks is final already (inter kernels ran before), so
         * the row below never has to wait for anything through it */
        if (threadIdx.x == 0)
            atomicExch(prog + row, c + 1);
    }
}

} /* namespace vp9 */
#endif
