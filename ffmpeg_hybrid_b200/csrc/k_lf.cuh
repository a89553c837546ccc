/*
 * k_lf.cuh -- K4: the VP9 loop filter as a superblock-row wavefront, in the reference's exact
 * edge order, with the edges of one superblock scheduled by their true data dependencies.
 *
 * The reference filters superblock by superblock in raster order and, inside one, plane by
 * plane: every column edge C(c), then every row edge R(c) (vp9lpf.c:222-229).  Footprints:
 * C(c) = columns [x0-8, x0+S) x rows [y0, y0+S); R(c) = columns [x0, x0+S) x rows [y0-8, y0+S).
 * Only the superblock's left edge reaches into the left neighbour, so C(c) = Ce(c) (edge
 * position 0) followed by Cr(c) (the rest), and Ce(c+1) commutes with everything of superblock
 * c+1's upper neighbours.  One CTA owns one (superblock row, plane) and runs the turns
 *
 *        U(c) = Cr(c) -> R(c) -> Ce(c+1)                       c = 0 .. sb_cols-1
 *
 * after which columns [x0, x0+S) are final for rows [y0-8, y0+S-8) and "handed down" for the
 * last 8 rows.  U(c) of row r needs the hand-down of row r-1's U(c): a lag of ONE superblock
 * between rows (the plain raster dependency needs two), so the frame's critical path is
 * sb_cols + sb_rows turns.
 *
 * Inside a turn the ops are not run edge by edge: k_lf_sched (pre-pass, fully parallel) decodes
 * the VP9Filter masks (vp9_lf.h, 1:1 walk of filter_plane_cols/rows, vp9lpf.c:31-181) and sorts
 * each pass's ops by dependency step (LfSched); all (op, line) pairs of one step run
 * concurrently on 256 threads, one barrier per step.
 *
 * Pixels live in a shared-memory ring of 4 superblock columns x (S+8) rows.  Three helper warps
 * keep the turn free of global-memory latency: a body loader (cp.async of pre-filter pixels and
 * the schedule, up to 3 superblocks ahead), a top loader (polls the 8 handed-down rows, which
 * travel as tagged 8-byte words -- 4 bytes of pixels | launch tag, FrameDesc.lf_ll -- so that
 * pixels and "ready" arrive in one load), and a storer (sends the last 8 rows down as tagged
 * words, then writes rows [y0-8, y0+S-8) of the column block back: the 8 rows above included,
 * its own last 8 rows left to the row below, the last row writes its own).
 *
 * Replaces ff_vp9_loopfilter_sb (vp9lpf.c:183-230) + loop_filter_8/16/mix2 slots
 * (vp9dsp_template.c:1780-1967).  Algorithmic bytes: 2*P (read + write of the picture).
 */
#ifndef VP9CUDA_K_LF_CUH
#define VP9CUDA_K_LF_CUH

#include "dev_common.cuh"
#include "k_intra.cuh"

namespace vp9 {

#ifndef LF_CW
#define LF_CW 8                       /* compute warps */
#endif
#define LF_CT (LF_CW * 32)
#define LF_THREADS (LF_CT + 3 * 32)   /* + body loader, top loader, storer */

/* pre-pass, fully parallel: one CTA per superblock, one warp per (plane, direction).  Four lanes
 * decode the VP9Filter masks (two 8-line groups each), eight lanes compute the dependency steps
 * of their group (same rule as lf_schedule_dir, vp9_lf.h), then the warp sorts the ops by
 * (step, width) with a shared-memory histogram.  Equal-width ops sit together inside a step so
 * the warps of the wavefront kernel diverge less. */
#define LF_SCHED_THREADS 192
__global__ void __launch_bounds__(LF_SCHED_THREADS) k_lf_sched(const __grid_constant__ FrameDesc f)
{
    __shared__ LfSb m;
    __shared__ uint16_t ent[6][128];
    __shared__ uint8_t key[6][128];
    __shared__ int hist[6][72];
    const int sb = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int pl = warp >> 1, dir = warp & 1;
    if (threadIdx.x < 48)
        ((uint32_t *)&m)[threadIdx.x] = ((const uint32_t *)&f.lf[sb])[threadIdx.x];
    for (int i = threadIdx.x; i < 6 * 72; i += LF_SCHED_THREADS)
        (&hist[0][0])[i] = 0;
    __syncthreads();
    if (lane < 4) {
        LfOp o1[16], o2[16];
        if (dir == 0)
            lf_decode_cols_it(m, pl > 0, sb % f.sb_cols == 0, lane, o1, o2);
        else
            lf_decode_rows_j(m, pl > 0, sb / f.sb_cols == 0, lane, o1, o2);
#pragma unroll
        for (int p = 0; p < 16; p++) {
            ent[warp][(2 * lane) * 16 + p] = o1[p].wd ? lf_ent(2 * lane, p, o1[p]) : 0;
            ent[warp][(2 * lane + 1) * 16 + p] = o2[p].wd ? lf_ent(2 * lane + 1, p, o2[p]) : 0;
        }
    }
    __syncwarp();
    if (lane < 8) {
        int st[16];
#pragma unroll
        for (int p = 0; p < 16; p++) {
            const unsigned e = ent[warp][lane * 16 + p];
            st[p] = -1;
            int k = 0xff;
            if (e) {
                const int wc = e >> 7 & 3, lo = 4 * p - (wc == 3 ? 8 : 4);
                int s = (dir == 0 && p > 0) ? 1 : 0;
#pragma unroll
                for (int q = (p > 3 ? p - 3 : 0); q < p; q++) {
                    const unsigned eq = ent[warp][lane * 16 + q];
                    if (eq && 4 * q + ((eq >> 7 & 3) == 3 ? 8 : 4) > lo && st[q] + 1 > s)
                        s = st[q] + 1;
                }
                st[p] = s;
                k = s * 4 + wc;
            }
            key[warp][lane * 16 + p] = (uint8_t)k;
        }
    }
    __syncwarp();
    int pos[4], kk[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        kk[i] = key[warp][4 * lane + i];
        pos[i] = kk[i] != 0xff ? atomicAdd(&hist[warp][kk[i]], 1) : 0;
    }
    __syncwarp();
    {
        /* exclusive scan of hist[warp][0..71], 3 keys per lane (lanes 0..23) */
        int c0 = 0, c1 = 0, c2 = 0;
        if (lane < 24) { c0 = hist[warp][3 * lane]; c1 = hist[warp][3 * lane + 1]; c2 = hist[warp][3 * lane + 2]; }
        int tot = c0 + c1 + c2, inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
        __syncwarp();
        if (lane < 24) {
            int base = inc - tot;
            hist[warp][3 * lane] = base;
            hist[warp][3 * lane + 1] = base + c0;
            hist[warp][3 * lane + 2] = base + c0 + c1;
        }
    }
    __syncwarp();
    LfSched &out = f.lfsched[sb * 3 + pl];
    uint16_t *oe = dir ? out.row : out.col;
    uint8_t *os = dir ? out.rstart : out.cstart;
#pragma unroll
    for (int i = 0; i < 4; i++)
        if (kk[i] != 0xff)
            oe[hist[warp][kk[i]] + pos[i]] = ent[warp][4 * lane + i];
    if (lane <= LF_NSTEP)
        os[lane] = (uint8_t)hist[warp][4 * lane];
    /* steps are contiguous (an op of step s has a predecessor of step s-1): store the range */
    int top = -1;
#pragma unroll
    for (int i = 0; i < 4; i++)
        if (kk[i] != 0xff && (kk[i] >> 2) > top) top = kk[i] >> 2;
    top = __reduce_max_sync(0xffffffffu, top);
    if (lane == 0) {
        if (dir) out.rmax = (uint8_t)(top + 1);
        else     out.cmax = (uint8_t)(top < 0 ? 0 : top);
    }
}

__device__ __forceinline__ void lf_cp16(void *smem, const void *gmem) { cp16(smem, gmem); }
__device__ __forceinline__ void lf_cp_commit() { cp_commit(); }
template <int N> __device__ __forceinline__ void lf_cp_wait() { cp_wait<N>(); }
__device__ __forceinline__ void lf_bar_compute() { bar_named<1, LF_CT>(); }

/* signals between the warps of one CTA (shared memory, monotonically increasing counts) */
struct LfSig {
    volatile int body;     /* body blocks L(0..body-1) and their schedules are in the ring */
    volatile int top;      /* top strips T(0..top-1) are in the ring */
    volatile int done;     /* turns U(0..done-1) are complete in the ring */
    volatile int stored;   /* column blocks 0..stored-1 have been written back (ring slot reusable) */
};

template <class P> struct LfPix;
template <> struct LfPix<uint16_t> {
    typedef uint2 V;      /* 4 pixels */
    static __device__ __forceinline__ void unpack(V v, int *px) { px[0] = v.x & 0xffff; px[1] = v.x >> 16; px[2] = v.y & 0xffff; px[3] = v.y >> 16; }
    static __device__ __forceinline__ V pack(const int *px) { V v; v.x = px[0] | px[1] << 16; v.y = px[2] | px[3] << 16; return v; }
};
template <> struct LfPix<uint8_t> {
    typedef uint32_t V;
    static __device__ __forceinline__ void unpack(V v, int *px) { px[0] = v & 0xff; px[1] = v >> 8 & 0xff; px[2] = v >> 16 & 0xff; px[3] = v >> 24; }
    static __device__ __forceinline__ V pack(const int *px) { return px[0] | px[1] << 8 | px[2] << 16 | px[3] << 24; }
};

#define LF_CLK(k) do { if (f.phase_clk && ctid == 0) { long long c_ = clock64(); \
    atomicAdd(&f.phase_clk[(k) + (S == 64 ? 0 : 5)], (unsigned long long)(c_ - clk_)); clk_ = c_; } } while (0)

template <class P, int S>
__device__ void lf_row(const FrameDesc &f, int pl, int row, unsigned char *ring, LfSched *sched, LfSig *sig,
                       const uint8_t *lut /* lim[64], mblim[64] */)
{
    typedef typename LfPix<P>::V V;
    constexpr int RW = 4 * S;                           /* ring width in pixels */
    constexpr int RSB = RW * (int)sizeof(P) + 16;       /* ring row stride in bytes (16-B aligned, 4 words mod 32) */
    constexpr int CPR = S * (int)sizeof(P) / 16;        /* 16-byte chunks per block row */
    constexpr int SLOT = S * (int)sizeof(P);            /* bytes per ring slot row */
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ncol = f.sb_cols;
    const bool has_top = row > 0;
    const int stride = f.cur.stride[pl];
    P *plane = (P *)f.cur.pl[pl];
    const int y0 = row * S;
    /* Running beside the intra wavefront (k_wave): superblock (row, j) may be read once the intra
     * pass has written it back completely (which also means the intra row has fetched its border
     * line above columns j and j + 1, so that this row may overwrite it). */
    const volatile int *const intra_full = (f.lf_follow & 1) ? f.iprogress + f.sb_rows + row : nullptr;

    if (warp < LF_CW) {
        /* ------------------------------- compute warps ------------------------------- */
        const int ctid = threadIdx.x;
        long long clk_ = clock64();
        for (int c = 0; c < ncol; c++) {
            if (ctid == 0) {
                const int need = min(c + 2, ncol);
                VP9_SPIN(sig->body < need, VP9_TIGHT_NS, "lf compute: body", row * 3 + pl, c);
                __threadfence_block();
            }
            lf_bar_compute();
            LF_CLK(0);
            const LfSched &sc = sched[c & 3];
            /* one dependency step of column ops of superblock sbi */
            auto col_step = [&](const uint16_t *ent, int a, int b, int sbi) {
                for (int t = ctid; t < (b - a) * 8; t += LF_CT) {
                    const unsigned e = ent[a + (t >> 3)];
                    const int g = e & 7, p = e >> 3 & 15, wd = 2 << (e >> 7 & 3), L = e >> 9;
                    unsigned char *rp = ring + (8 + 8 * g + (t & 7)) * RSB;
                    const int xb = sbi * S + 4 * p;
                    int px[16];
                    if (wd == 16) {
                        LfPix<P>::unpack(*(const V *)(rp + ((xb - 8) & (RW - 1)) * sizeof(P)), px);
                        LfPix<P>::unpack(*(const V *)(rp + ((xb + 4) & (RW - 1)) * sizeof(P)), px + 12);
                    }
                    LfPix<P>::unpack(*(const V *)(rp + ((xb - 4) & (RW - 1)) * sizeof(P)), px + 4);
                    LfPix<P>::unpack(*(const V *)(rp + (xb & (RW - 1)) * sizeof(P)), px + 8);
                    const unsigned w = lf_line(px, wd, lut[64 + L], lut[L], L >> 4, f.bpp);
                    if (w & 0x000f) *(V *)(rp + ((xb - 8) & (RW - 1)) * sizeof(P)) = LfPix<P>::pack(px);
                    if (w & 0x00f0) *(V *)(rp + ((xb - 4) & (RW - 1)) * sizeof(P)) = LfPix<P>::pack(px + 4);
                    if (w & 0x0f00) *(V *)(rp + (xb & (RW - 1)) * sizeof(P)) = LfPix<P>::pack(px + 8);
                    if (w & 0xf000) *(V *)(rp + ((xb + 4) & (RW - 1)) * sizeof(P)) = LfPix<P>::pack(px + 12);
                }
            };
            /* Cr(c) */
#pragma unroll 1
            const int cmax = sc.cmax, rmax = sc.rmax;
            for (int s = 1; s <= cmax; s++) {
                const int a = sc.cstart[s], b = sc.cstart[s + 1];
                col_step(sc.col, a, b, c);
                lf_bar_compute();
            }
            LF_CLK(1);
            if (has_top) {
                if (ctid == 0) {
                    VP9_SPIN(sig->top < c + 1, VP9_TIGHT_NS, "lf compute: top", row * 3 + pl, c);
                    __threadfence_block();
                }
                lf_bar_compute();
            }
            LF_CLK(2);
            /* R(c) */
#pragma unroll 1
            for (int s = 0; s < rmax; s++) {
                const int a = sc.rstart[s], b = sc.rstart[s + 1];
                for (int t = ctid; t < (b - a) * 8; t += LF_CT) {
                    const unsigned e = sc.row[a + (t >> 3)];
                    const int cg = e & 7, ei = e >> 3 & 15, wd = 2 << (e >> 7 & 3), L = e >> 9;
                    const int col = (c * S + 8 * cg + (t & 7)) & (RW - 1);
                    unsigned char *cp = ring + 4 * ei * RSB + col * sizeof(P);
                    int px[16];
                    if (wd == 16) {
#pragma unroll
                        for (int i = 0; i < 16; i++) px[i] = *(const P *)(cp + i * RSB);
                    } else {
#pragma unroll
                        for (int i = 4; i < 12; i++) px[i] = *(const P *)(cp + i * RSB);
                    }
                    const unsigned w = lf_line(px, wd, lut[64 + L], lut[L], L >> 4, f.bpp);
                    if (w) {
#pragma unroll
                        for (int i = 1; i < 15; i++)
                            if (w >> i & 1) *(P *)(cp + i * RSB) = (P)px[i];
                    }
                }
                lf_bar_compute();
            }
            LF_CLK(3);
            /* Ce(c+1) */
            if (c + 1 < ncol) {
                const LfSched &sn = sched[(c + 1) & 3];
                const int a = sn.cstart[0], b = sn.cstart[1];
                if (a != b) {
                    col_step(sn.col, a, b, c + 1);
                    lf_bar_compute();
                }
            }
            if (ctid == 0) {
                __threadfence_block();
                sig->done = c + 1;
            }
            LF_CLK(4);
        }
    } else if (warp == LF_CW) {
        /* ------------------------------- body loader ------------------------------- */
        for (int j = 0; j < ncol; j++) {
            /* Lane 0 alone looks at the flags and the warp follows its verdict: a per-lane read of a
             * changing flag could send part of the warp into the hand-over branch and leave the
             * __syncwarp()s below unpaired. */
            int wait = 0;
            if (lane == 0)
                wait = (j >= 4 && sig->stored < j - 3) || (intra_full && *intra_full < j + 1);
            wait = __shfl_sync(0xffffffffu, wait, 0);
            if (wait) {
                /* hand over what is already in flight before sleeping */
                lf_cp_wait<0>();
                __syncwarp();
                if (lane == 0) {
                    __threadfence_block();
                    sig->body = j;
                    VP9_SPIN(j >= 4 && sig->stored < j - 3, 20, "lf body loader: ring", row * 3 + pl, j);
                    if (intra_full)
                        VP9_SPIN(*intra_full < j + 1, 40, "lf body loader: intra written back", row * 3 + pl, j);
                }
            }
            if (lane == 0) __threadfence();
            __syncwarp();
            const unsigned char *src = (const unsigned char *)(plane + (size_t)y0 * stride + j * S);
            unsigned char *dst = ring + 8 * RSB + (j & 3) * SLOT;
#pragma unroll 4
            for (int i = lane; i < S * CPR; i += 32) {
                const int r = i / CPR, k = i % CPR;
                lf_cp16(dst + r * RSB + 16 * k, src + (size_t)r * stride * sizeof(P) + 16 * k);
            }
            const unsigned char *ss = (const unsigned char *)&f.lfsched[(row * ncol + j) * 3 + pl];
            unsigned char *sd = (unsigned char *)&sched[j & 3];
            for (int i = lane; i < (int)sizeof(LfSched) / 16; i += 32)
                lf_cp16(sd + 16 * i, ss + 16 * i);
            lf_cp_commit();
            lf_cp_wait<1>();
            __syncwarp();
            if (lane == 0) { __threadfence_block(); sig->body = j; }
        }
        lf_cp_wait<0>();
        __syncwarp();
        if (lane == 0) { __threadfence_block(); sig->body = ncol; }
    } else if (warp == LF_CW + 1) {
        /* ------------------------------- top loader ------------------------------- */
        if (has_top) {
            /* the 8 rows above come as tagged words from the row above's storer (lf_ll): data and "ready" in one load */
            constexpr int LLW = 8 * S * (int)sizeof(P) / 4, WPR = S * (int)sizeof(P) / 4, NW = LLW / 32;
            const unsigned long long *in = f.lf_ll + ((size_t)pl * f.sb_rows + row - 1) * ncol * LF_LL_WORDS;
            const unsigned tag = f.ll_tag;
            for (int c = 0; c < ncol; c++) {
                if (lane == 0)
                    VP9_SPIN(c >= 4 && sig->stored < c - 3, 20, "lf top loader: ring", row * 3 + pl, c);
                __syncwarp();
                const unsigned long long *src = in + (size_t)c * LF_LL_WORDS + lane;
                unsigned long long w[NW];
                auto fetch = [&]() {
                    bool ready = true;
#pragma unroll
                    for (int j = 0; j < NW; j++) w[j] = ll_load(src + 32 * j);
#pragma unroll
                    for (int j = 0; j < NW; j++) ready &= (unsigned)(w[j] >> 32) == tag;
                    return ready;
                };
                VP9_SPIN(!fetch(), 0, "lf top loader: row above", row * 3 + pl, c);
                unsigned char *dst = ring + (c & 3) * SLOT;
#pragma unroll
                for (int j = 0; j < NW; j++) {
                    const int i = lane + 32 * j;
                    *(unsigned *)(dst + (i / WPR) * RSB + 4 * (i % WPR)) = (unsigned)w[j];
                }
                __syncwarp();
                if (lane == 0) { __threadfence_block(); sig->top = c + 1; }
            }
        }
    } else {
        /* ------------------------------- storer ------------------------------- */
        constexpr int LLW = 8 * S * (int)sizeof(P) / 4, WPR = S * (int)sizeof(P) / 4;
        const bool last_row = row + 1 == f.sb_rows;
        unsigned long long *out = f.lf_ll + ((size_t)pl * f.sb_rows + row) * ncol * LF_LL_WORDS;
        for (int c = 0; c < ncol; c++) {
            if (lane == 0) {
                VP9_SPIN(sig->done < c + 1, 20, "lf storer: done", row * 3 + pl, c);
                __threadfence_block();
            }
            __syncwarp();
            unsigned char *dst = (unsigned char *)(plane + (size_t)(y0 - 8) * stride + c * S);
            const unsigned char *src = ring + (c & 3) * SLOT;
            auto store_rows = [&](int r0, int r1) {
#pragma unroll 4
                for (int i = lane + r0 * CPR; i < r1 * CPR; i += 32) {
                    const int r = i / CPR, k = i % CPR;
                    *(uint4 *)(dst + (ptrdiff_t)r * stride * (int)sizeof(P) + 16 * k) = *(const uint4 *)(src + r * RSB + 16 * k);
                }
            };
            /* The last 8 rows go to the row below as tagged words: it filters across them and writes them to the
             * picture with its own column block c, after the intra row below has long fetched their pre-filter
             * bottom line -- so this row never overwrites what the intra wavefront still reads (the job
             * intra_pred_data does in the reference, vp9.c:1404-1416).  The last row writes them itself. */
            if (!last_row) {
#pragma unroll
                for (int i = lane; i < LLW; i += 32)
                    ll_store(out + (size_t)c * LF_LL_WORDS + i, *(const unsigned *)(src + (S + i / WPR) * RSB + 4 * (i % WPR)), f.ll_tag);
            } else {
                store_rows(S, S + 8);
            }
            store_rows(has_top ? 0 : 8, S);
            __syncwarp();
            if (lane == 0)
                sig->stored = c + 1;
        }
        /* every pixel of this (row, plane) is written: the next frame's inter prediction may read
         * rows [y0 - 8, y0 + S - 8) of the plane, and the rest once the row below says the same */
        __syncwarp();
        if (lane == 0) {
            __threadfence();
            *(volatile int *)(f.progress + (3 + pl) * f.sb_rows + row) = 1;
        }
    }
}

/* shared memory of one loop-filter CTA */
template <class P> struct LfSmem {
    static constexpr int O_RING = 0;                                      /* 72 ring rows */
    static constexpr int O_SCHED = 72 * (256 * (int)sizeof(P) + 16);       /* LfSched[4] */
    static constexpr int O_LUT = O_SCHED + 4 * (int)sizeof(LfSched);       /* lim[64], mblim[64] */
    static constexpr int BYTES = O_LUT + 128;
};

/* loop-filter CTA `idx` = (superblock row idx / 3, plane idx % 3); threads >= LF_THREADS idle */
template <class P>
__device__ __forceinline__ void lf_cta(const FrameDesc &f, int idx, unsigned char *sm, LfSig &sig)
{
    typedef LfSmem<P> L;
    unsigned char *ring = sm + L::O_RING;
    LfSched *sched = (LfSched *)(sm + L::O_SCHED);
    uint8_t *lut = sm + L::O_LUT;
    const int pl = f.mono >= 0 ? 0 : idx % 3, row = f.mono >= 0 ? idx : idx / 3;     /* 4:4:4: one plane per launch */
    if (threadIdx.x == 0) {
        sig.body = 0; sig.top = 0; sig.done = 0; sig.stored = 0;
    }
    if (threadIdx.x < 64) {
        lut[threadIdx.x] = f.lim_lut[threadIdx.x];
        lut[64 + threadIdx.x] = f.mblim_lut[threadIdx.x];
    }
    __syncthreads();
    if (threadIdx.x >= LF_THREADS)
        return;
    if (pl == 0)
        lf_row<P, 64>(f, pl, row, ring, sched, &sig, lut);
    else
        lf_row<P, 32>(f, pl, row, ring, sched, &sig, lut);
}

template <class P>
__global__ void __launch_bounds__(LF_THREADS) k_lf(const __grid_constant__ FrameDesc f)
{
    __shared__ __align__(16) unsigned char sm[LfSmem<P>::BYTES];
    __shared__ LfSig sig;
    lf_cta<P>(f, blockIdx.x, sm, sig);
}

/* ---------------------------------------------------------------------------------------------- */
/* k_wave: the intra wavefront and the loop-filter wavefront of one frame in ONE launch.
 *
 * CTAs [0, sb_rows) run intra_row, CTAs [sb_rows, 4*sb_rows) run lf_cta.  The loop filter of
 * superblock row r follows the intra pass at a distance of about three superblocks: it reads
 * block (r, j) once the intra pass has written it back (ifull[r] > j) and writes rows
 * [y0 - 8, y0 + 56) of column block c: never its own last 8 rows, whose pre-filter bottom line the
 * intra row below still reads -- they go down as tagged words and the row below writes them, which
 * is the role intra_pred_data plays in the reference (vp9.c:1404-1416).
 * Every wait points at a CTA with a lower block index, so in-order CTA dispatch rules out
 * deadlock even when the grid is larger than the machine (8K frames). */
template <class P> struct WaveSmem {
    static constexpr int BYTES = IntraSmem<P>::BYTES > LfSmem<P>::BYTES ? IntraSmem<P>::BYTES : LfSmem<P>::BYTES;
};

#ifndef VP9_WAVE_CTAS_PER_SM
#define VP9_WAVE_CTAS_PER_SM 1
#endif
template <class P>
__global__ void __launch_bounds__(INTRA_THREADS, VP9_WAVE_CTAS_PER_SM) k_wave(const __grid_constant__ FrameDesc f)
{
    VP9_DYN_SMEM(sm);
    __shared__ IntraSig isig;
    __shared__ LfSig lsig;
    /* test hook (set_option "inject_stall"): CTA 0 waits for something that never happens, so that the
     * budget of VP9_SPIN, the error word and the draining of the launch can be exercised */
    if ((f.lf_follow & 4) && blockIdx.x == 0 && threadIdx.x == 0)
        VP9_SPIN(true, 100, "injected stall", 0, 0);
    if ((int)blockIdx.x < f.sb_rows)
        intra_row<P>(f, blockIdx.x, sm, isig);
    else
        lf_cta<P>(f, blockIdx.x - f.sb_rows, sm, lsig);
}

} /* namespace vp9 */
#endif
