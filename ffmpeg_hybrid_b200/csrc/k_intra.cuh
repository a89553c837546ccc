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
 *       independent) and writes them as compact 16-byte IntraRec with run lengths, plus the
 *       superblock's coefficient range.
 *   k_intra : one CTA per superblock ROW.  The row's pixels slide through a shared-memory ring
 *       of four superblock columns per plane (row 0 of each ring = the pixel row above).  A
 *       body-loader warp streams tiles, records and coefficients in with cp.async one turn
 *       ahead; a top-loader warp polls the row above's progress counter and fetches the border
 *       row; eight compute warps run the turn's dependency steps (one block per warp and step,
 *       edges -> prediction -> inverse transform, all in shared memory, one barrier per step); a
 *       storer warp writes the finished tile back, bottom pixel row first, then fence + progress
 *       counter for the row below.  Superblocks without intra blocks cost one flag update.
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

#define INTRA_CW 8                          /* compute warps */
#define INTRA_CT (INTRA_CW * 32)
#define INTRA_THREADS (INTRA_CT + 3 * 32)   /* + body loader, top loader, storer */
#define INTRA_MAX_SB_COLS 256
#define INTRA_SB_RECS 384                   /* 256 luma + 2 x 64 chroma 4x4 blocks */
#define INTRA_SB_COEFS 6144                 /* 64*64 + 2 * 32*32 */

__device__ __forceinline__ void cp16(void *smem, const void *gmem)
{
    unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

/* ---------------------------------------------------------------------------------------------- */
__global__ void __launch_bounds__(128) k_intra_prep(const __grid_constant__ FrameDesc f)
{
    __shared__ uint32_t keys[512];
    __shared__ unsigned lo, hi;
    const int sb = blockIdx.x, tid = threadIdx.x;
    const Vp9CudaSbIndex si = f.sb_index[sb];
    const int n = si.itx_count;
    const uint32_t g0 = n ? f.lane_itx_base[si.lane] + si.itx_start : 0;
    if (n == 0 || n > 512) {
        if (tid == 0) {
            IntraSb o = { g0, 0, 0, 0, 0 };
            f.isb[sb] = o;
        }
        return;
    }
    if (tid == 0) { lo = 0xffffffffu; hi = 0; }
    int m = 32;
    while (m < n) m <<= 1;
    __syncthreads();
    for (int i = tid; i < m; i += 128) {
        uint32_t key = 0xffffffffu;
        if (i < n) {
            const Vp9CudaIntraTx &r = f.itx[g0 + i];
            key = (uint32_t)r.level << 10 | i;
            if (r.eob) {
                atomicMin(&lo, r.coef_off);
                atomicMax(&hi, r.coef_off + (16u << (2 * r.tx)));
            }
        }
        keys[i] = key;
    }
    __syncthreads();
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
    const int sbx = sb % f.sb_cols, sby = sb / f.sb_cols;
    for (int k = tid; k < n; k += 128) {
        const uint32_t key = keys[k], lvl = key >> 10;
        int run = 0;
        if (k == 0 || (keys[k - 1] >> 10) != lvl) {
            int e = k + 1;
            while (e < n && (keys[e] >> 10) == lvl) e++;
            run = e - k;
        }
        const Vp9CudaIntraTx r = f.itx[g0 + (key & 1023)];
        const int ch = r.plane > 0;
        IntraRec o;
        o.lx = (uint8_t)(r.x4 - sbx * (16 >> ch));
        o.ly = (uint8_t)(r.y4 - sby * (16 >> ch));
        o.plane = r.plane; o.tx = r.tx; o.mode = r.mode; o.txtp = r.txtp; o.flags = r.flags; o.pad = 0;
        o.eob = r.eob;
        o.run = (uint16_t)run;
        o.coef_rel = r.eob ? r.coef_off - lo : 0;
        f.irec[g0 + k] = o;
    }
    if (tid == 0) {
        IntraSb o = { g0, (uint16_t)n, 0, hi ? lo : 0, hi ? hi - lo : 0 };
        f.isb[sb] = o;
    }
}

/* ---------------------------------------------------------------------------------------------- */
struct IntraSig {
    volatile int body;     /* tiles / records / coefficients of superblocks < body are in shared memory */
    volatile int top;      /* border rows of superblocks < top are in shared memory */
    volatile int done;     /* turns < done are complete */
    volatile int stored;   /* superblocks < stored have been written back */
};

/* a plane's ring in shared memory, addressed with ring-column / ring-row coordinates */
template <class P> struct TilePx {
    unsigned char *t; int rsb, mask;
    __device__ __forceinline__ int operator()(int x, int y) const { return *(const P *)(t + y * rsb + (x & mask) * (int)sizeof(P)); }
};
template <class P> struct SmemBlk {
    P *p; int stride;
    __device__ __forceinline__ int get(int x, int y) const { return p[y * stride + x]; }
    __device__ __forceinline__ void put(int x, int y, int v) { p[y * stride + x] = (P)v; }
};

template <class P, bool HBD> struct IntraSmem {
    typedef typename ItxTypes<HBD>::C C;
    static constexpr int RSBY = 4 * 64 * (int)sizeof(P) + 16;      /* luma ring row stride, bytes */
    static constexpr int RSBC = 4 * 32 * (int)sizeof(P) + 16;
    static constexpr int O_Y = 0;
    static constexpr int O_U = O_Y + 65 * RSBY;
    static constexpr int O_V = O_U + 33 * RSBC;
    static constexpr int O_COEF = O_V + 33 * RSBC;                 /* [2][INTRA_SB_COEFS] C */
    static constexpr int O_REC = O_COEF + 2 * INTRA_SB_COEFS * (int)sizeof(C);
    static constexpr int O_TMP = O_REC + 2 * INTRA_SB_RECS * (int)sizeof(IntraRec);
    static constexpr int O_EDGE = O_TMP + INTRA_CW * 32 * 33 * (int)sizeof(C);
    static constexpr int O_INFO = O_EDGE + INTRA_CW * 100 * 4;
    static constexpr int BYTES = O_INFO + INTRA_MAX_SB_COLS * (int)sizeof(IntraSb);
};

template <class P, bool HBD>
__global__ void __launch_bounds__(INTRA_THREADS) k_intra(const __grid_constant__ FrameDesc f)
{
    typedef typename ItxTypes<HBD>::C C;
    typedef IntraSmem<P, HBD> L;
    extern __shared__ __align__(16) unsigned char sm[];
    __shared__ IntraSig sig;
    auto tile = [&](int pl) -> unsigned char * { return sm + (pl == 0 ? L::O_Y : pl == 1 ? L::O_U : L::O_V); };
    C *const coefbuf = (C *)(sm + L::O_COEF);
    IntraRec *const recbuf = (IntraRec *)(sm + L::O_REC);
    const IntraSb *const info = (const IntraSb *)(sm + L::O_INFO);
    const int row = blockIdx.x, ncol = f.sb_cols;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool has_top = row > 0;

    if (threadIdx.x == 0) { sig.body = 0; sig.top = 0; sig.done = 0; sig.stored = 0; }
    for (int i = threadIdx.x; i < ncol * 4; i += INTRA_THREADS)
        ((uint32_t *)(sm + L::O_INFO))[i] = ((const uint32_t *)(f.isb + (size_t)row * ncol))[i];
    __syncthreads();

    if (warp < INTRA_CW) {
        /* ------------------------------- compute warps ------------------------------- */
        int *top = (int *)(sm + L::O_EDGE) + warp * 100, *left = top + 66;
        C *tmp = (C *)(sm + L::O_TMP) + warp * 32 * 33;
        for (int c = 0; c < ncol; c++) {
            const int n = info[c].count;
            if (n == 0) {
                if (threadIdx.x == 0)
                    sig.done = c + 1;
                continue;
            }
            if (threadIdx.x == 0) {
                while (sig.body < c + 1) { }
                if (has_top)
                    while (sig.top < c + 1) { }
                __threadfence_block();
            }
            asm volatile("bar.sync 1, %0;\n" ::"n"(INTRA_CT) : "memory");
            const IntraRec *rec = recbuf + (c & 1) * INTRA_SB_RECS;
            const C *coefs = coefbuf + (c & 1) * INTRA_SB_COEFS;
            for (int a = 0; a < n;) {
                const int run = rec[a].run;
                for (int t = a + warp; t < a + run; t += INTRA_CW) {
                    const IntraRec r = rec[t];
                    const int pl = r.plane, ch = pl > 0;
                    const int N = 4 << r.tx, S = 64 >> ch;
                    const int rsb = ch ? L::RSBC : L::RSBY;
                    unsigned char *tl = tile(pl);
                    const int bx = (c & 3) * S + r.lx * 4, by = 1 + r.ly * 4;      /* ring coordinates */
                    const int abx = c * S + r.lx * 4, aby = row * S + r.ly * 4;    /* plane coordinates */
                    const bool have_left = r.flags & VP9CUDA_ITX_HAVE_LEFT, have_right = r.flags & VP9CUDA_ITX_HAVE_RIGHT;
                    const bool have_top = aby > 0;
                    const int mode = intra_mode_conv(r.mode, have_left, have_top);
                    const int needs = intra_mode_needs(mode);
                    const int avail_w = ((f.cols * 8) >> ch) - abx, avail_h = ((f.rows * 8) >> ch) - aby;
                    TilePx<P> px = { tl, rsb, 4 * S - 1 };
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
                    P *blkp = (P *)(tl + by * rsb) + bx;
                    const int bstride = rsb / (int)sizeof(P);
                    for (int k = lane; k < N * N; k += 32)
                        blkp[(k >> (2 + r.tx)) * bstride + (k & (N - 1))] =
                            (P)intra_pixel(mode, N, k & (N - 1), k >> (2 + r.tx), top, left, dc, f.bpp);
                    __syncwarp();
                    if (r.eob) {
                        SmemBlk<P> blk = { blkp, bstride };
                        const C *coef = coefs + r.coef_rel;
                        switch (r.tx) {
                        case 0: if (lane < 4)  itx_block<HBD, 4>(coef, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xfu); break;
                        case 1: if (lane < 8)  itx_block<HBD, 8>(coef, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xffu); break;
                        case 2: if (lane < 16) itx_block<HBD, 16>(coef, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xffffu); break;
                        default: itx_block<HBD, 32>(coef, tmp, lane, r.txtp, r.eob, f.bpp, blk, 0xffffffffu); break;
                        }
                        __syncwarp();
                    }
                }
                asm volatile("bar.sync 1, %0;\n" ::"n"(INTRA_CT) : "memory");
                a += run;
            }
            if (threadIdx.x == 0) {
                __threadfence_block();
                sig.done = c + 1;
            }
        }
    } else if (warp == INTRA_CW) {
        /* ------------------------------- body loader ------------------------------- */
        for (int j = 0; j < ncol; j++) {
            const IntraSb sj = info[j];
            const bool need_tile = sj.count > 0 || (j + 1 < ncol && info[j + 1].count > 0);
            if (need_tile && (sig.done < j - 1 || sig.stored < j - 3)) {
                cp_wait<0>();
                __syncwarp();
                if (lane == 0) { __threadfence_block(); sig.body = j; }
                while (sig.done < j - 1 || sig.stored < j - 3) __nanosleep(20);
            }
            if (need_tile) {
#pragma unroll
                for (int pl = 0; pl < 3; pl++) {
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
                const unsigned char *rs = (const unsigned char *)(f.irec + sj.start);
                unsigned char *rd = (unsigned char *)(recbuf + (j & 1) * INTRA_SB_RECS);
                for (int i = lane; i < sj.count; i += 32)
                    cp16(rd + 16 * i, rs + 16 * i);
                const unsigned char *cs = (const unsigned char *)((const C *)f.coef + sj.coef_lo);
                unsigned char *cd = (unsigned char *)(coefbuf + (j & 1) * INTRA_SB_COEFS);
                const int nch = (int)(sj.coef_n * sizeof(C) / 16);
                for (int i = lane; i < nch; i += 32)
                    cp16(cd + 16 * i, cs + 16 * i);
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
            for (int c = 0; c < ncol; c++) {
                if (info[c].count == 0) {
                    __syncwarp();
                    if (lane == 0) sig.top = c + 1;
                    continue;
                }
                while (sig.done < c - 2 || sig.stored < c - 3) __nanosleep(20);
                if (lane == 0) {
                    while (*above < c + 1) __nanosleep(20);
                    __threadfence();
                }
                __syncwarp();
#pragma unroll
                for (int pl = 0; pl < 3; pl++) {
                    const int S = pl ? 32 : 64;
                    const int cpr = S * (int)sizeof(P) / 16;
                    const int stride = f.cur.stride[pl];
                    const unsigned char *src = (const unsigned char *)((const P *)f.cur.pl[pl] + ((size_t)row * S - 1) * stride + c * S);
                    /* one extra chunk to the left: the top-left pixel lives in the previous ring slot */
                    const int k0 = c > 0 ? -1 : 0;
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
            }
        }
    } else {
        /* ------------------------------- storer ------------------------------- */
        for (int c = 0; c < ncol; c++) {
            while (sig.done < c + 1) __nanosleep(20);
            __threadfence_block();
            if (info[c].count > 0) {
                auto store_rows = [&](int pl, int r0, int r1) {
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
                store_rows(0, 64, 65); store_rows(1, 32, 33); store_rows(2, 32, 33);
                __syncwarp();
                if (lane == 0) {
                    __threadfence();
                    *(volatile int *)(f.iprogress + row) = c + 1;
                }
                store_rows(0, 1, 64); store_rows(1, 1, 32); store_rows(2, 1, 32);
            } else if (lane == 0) {
                *(volatile int *)(f.iprogress + row) = c + 1;
            }
            __syncwarp();
            if (lane == 0)
                sig.stored = c + 1;
        }
    }
}

} /* namespace vp9 */
#endif
