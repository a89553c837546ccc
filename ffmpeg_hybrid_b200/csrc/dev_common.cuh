/*
 * dev_common.cuh -- device-side view of one frame: surfaces, record arrays, frame scalars.
 * Shared by every kernel of libvp9cuda.so.
 */
#ifndef VP9CUDA_DEV_COMMON_CUH
#define VP9CUDA_DEV_COMMON_CUH

#ifdef VP9_EMU                      /* host emulation of a kernel inside the CPU tests (oracle/cuda_emu.h) */
#include "cuda_emu.h"
#else
#include <cuda_runtime.h>
#define VP9_DYN_SMEM(name) extern __shared__ __align__(16) unsigned char name[]
#endif
#include <stdint.h>
#include <stdio.h>

#include "vp9cuda.h"
#include "vp9_itx.h"
#include "vp9_pred.h"
#include "vp9_lf.h"

namespace vp9 {

/* One picture in HBM: three planes, 64-aligned padded size, stride in pixels.
 * w,h are the frame's own luma size (what motion compensation clamps to). */
struct SurfDesc {
    void *pl[3];
    int stride[3];
    int w, h;
};

/* Intra pre-pass output (k_intra_prep): one per superblock ... */
struct IntraSb {
    uint32_t start;                  /* first record of this superblock in irec[] */
    uint16_t count;                  /* intra transform blocks (all planes); 0 = nothing to do */
    uint16_t ntab;                   /* entries of its work table (itab) */
    uint16_t has_res;                /* any of them has coefficients */
    uint16_t needs_top;              /* some block reads intra pixels of the superblock row above */
    uint32_t pad2;
};
/* ... and one per intra transform block, sorted by dependency depth inside the superblock */
struct IntraRec {
    uint8_t lx, ly;                  /* position inside the superblock plane, 4-pixel units */
    uint8_t plane, tx, mode /* edge-converted */, txtp, flags /* 1 left, 2 right, 4 top */, pad;
    uint16_t eob;
    uint16_t ncoef;                  /* coefficient entries (non-zero coefficients) */
    uint32_t coef_off;               /* first entry in the frame's coefficient stream */
};

/* What the intra wavefront needs per transform block, with every address that does not depend on
 * pixel values worked out by the pre-pass (k_intra_prep): byte offsets into the wavefront CTA's
 * shared memory (pixel rings, IntraSmem) and limits of the edge rules of check_intra_mode
 * (vp9recon.c:72-221).  The dependent chain of the wavefront then is: this record -> edge samples ->
 * predictor taps -> store. */
struct IntraBlk {
    uint32_t dst;                    /* pixel (0,0) of the block */
    uint32_t top;                    /* sample 0 of the row above it */
    uint32_t tl;                     /* the top-left sample (may wrap to the ring's last column) */
    uint32_t left;                   /* the sample left of row 0 */
    uint16_t res;                    /* the block's first residual in the superblock's residual tile */
    uint8_t mode;                    /* edge-converted prediction mode */
    uint8_t lg;                      /* log2 of the block size: 2..5 */
    uint8_t klim;                    /* last sample of the row above the block may read (replicated beyond) */
    uint8_t hlim;                    /* last row whose left sample exists */
    uint8_t flags;                   /* 1 left, 2 top, 4 has residual, 8 real top-left, 16 chroma */
    uint8_t pad[9];
};                                   /* 32 bytes */

/* Everything a kernel needs to know about the frame being reconstructed. */
struct FrameDesc {
    SurfDesc cur;
    SurfDesc ref[3];                 /* by b->ref index (LAST, GOLDEN, ALTREF) */
    int w, h, cols, rows, sb_cols, sb_rows, bpp;
    int mono;                        /* -1: a 4:2:0 frame, all three planes in one go.  0..2: one plane of a 4:4:4 frame.  There
                                        every plane is predicted, transformed and filtered exactly like luma (full size, luma
                                        vector units, own vectors per 4x4, luma loop-filter masks: vp9_mc_template.c:78-121,
                                        vp9lpf.c:214 with ss_h == ss_v == 0), so the frame runs as three passes in which
                                        cur / ref plane 0 are that plane and only its records count, as plane 0 */
    uint16_t mvscale[3][2];
    uint8_t mvstep[3][2];
    uint8_t lossless, lf_level;
    /* records (device pointers); lane l owns elements [lane_*_base[l], lane_*_base[l] + lane_*_cnt[l]) */
    const Vp9CudaInterBlock *blk;  uint32_t n_blk;     /* totals over the lanes */
    const Vp9CudaTx *tx;           uint32_t n_tx;
    const Vp9CudaIntraTx *itx;     uint32_t n_itx;
    const void *coef;              /* Vp9CudaCoef16 (8-bit) or Vp9CudaCoef32 entries, indexed by coef_off */
    const Vp9CudaSbIndex *sb_index;
    const Vp9CudaLoopFilterSb *lf;
    int n_lanes;
    uint32_t lane_blk_base[VP9CUDA_MAX_LANES], lane_blk_cnt[VP9CUDA_MAX_LANES];
    uint32_t lane_tx_base[VP9CUDA_MAX_LANES], lane_tx_cnt[VP9CUDA_MAX_LANES];
    uint32_t lane_itx_base[VP9CUDA_MAX_LANES]; /* first itx record of each lane */
    /* scratch */
    int32_t *cellmap;              /* cols*rows: index into blk, -1 = not inter */
    IntraSb *isb;                  /* [sb] intra work of each superblock */
    IntraRec *irec;                /* n_itx records regrouped per superblock, sorted by depth */
    IntraBlk *iblk;                /* the same blocks as the wavefront reads them */
    int *iprogress;                /* [2][sb_rows] intra wavefront counters: border row stored, superblock stored */
    int lf_follow;                 /* the loop filter runs beside the intra wavefront (k_wave) */
    uint16_t *itab;                /* [sb][768] per-step chunk tables (k_intra_prep) */
    int16_t *ires;                 /* [sb][6144] intra residual tiles (k_intra_res -> k_intra) */
    LfSched *lfsched;              /* [sb][plane] loop-filter edge ops sorted by dependency step */
    int *progress;                 /* [3 planes][sb_rows] unused since the tagged-word hand-over, then [3][sb_rows] loop-filter "row done" */
    unsigned long long *lf_ll;     /* [3 planes][sb_rows][sb_cols][LF_LL_WORDS]: the 8 filtered rows a loop-filter row hands down,
                                      as (4 bytes of pixels | launch tag << 32) words: data and "ready" arrive in one 8-byte store */
    unsigned long long *intra_ll;  /* [sb_rows][sb_cols][INTRA_LL_WORDS]: the bottom pixel line of every superblock the intra wavefront
                                      reconstructed (luma at 0, U at 32, V at 48), as tagged words for the intra row below */
    unsigned ll_tag;               /* tag of this launch's words (never 0, unique per context and launch) */
    int *mv_reach;                 /* max vertical motion vector component of this frame, 1/8 pel (k_cellmap) */
    int tile_y0;                   /* first row of this launch when the frame is launched in bands: k_inter 32-row tile rows, k_inter_sb superblock rows */
    unsigned long long *phase_clk; /* optional: per-phase clock accumulation (profiling), or NULL */
    int *err;                      /* context error word (device memory): non-zero = a wavefront wait gave up */
    int *err_host;                 /* the same, host-mapped: what vp9cuda_wait reads */
    uint8_t lim_lut[64], mblim_lut[64];
};

__device__ __forceinline__ int bs_w8(int bs)
{
    /* ff_vp9_bwh_tab[1][bs][0] (vp9data.c:25-33): block width in 8-pixel units */
    return (int)((0x1111122244488ull >> (4 * bs)) & 15);   /* 8,8,4,4,4,2,2,2,1,1,1,1,1 */
}
__device__ __forceinline__ int bs_h8(int bs)
{
    /* ff_vp9_bwh_tab[1][bs][1]: 8,4,8,4,2,4,2,1,2,1,1,1,1 */
    return (int)((0x1111212424848ull >> (4 * bs)) & 15);
}

__device__ __forceinline__ int rounded_div(int a, int b)
{
    return a >= 0 ? (a + (b >> 1)) / b : (a - (b >> 1)) / b;
}

/* Spin-wait on `cond`; `ns` > 0 sleeps between polls.  Built with -DVP9CUDA_WATCHDOG a wait that
 * lasts longer than ~1 s reports where it is stuck into a host-mapped buffer and traps, so a
 * dependency bug shows up as an error instead of a hung GPU. */
#ifndef VP9_TIGHT_NS
#define VP9_TIGHT_NS 0      /* sleep (ns) inside the compute warps' waits for their loader warps */
#endif
#ifdef VP9_EMU
#define VP9_SPIN(cond, ns, tag, a, b) do { for (int i_ = 0; (cond); i_++) { if (i_ < 8) std::this_thread::yield(); else usleep(20); } } while (0)
#elif defined(VP9CUDA_WATCHDOG)
__device__ int *g_wd_buf;            /* host-mapped: [0] = count, then {line, id, at, block} per report */
#define VP9_SPIN(cond, ns, tag, a, b) do { long long t0_ = clock64(); bool rep_ = false; while (cond) { if ((ns) > 0) __nanosleep(ns); \
    const long long dt_ = clock64() - t0_; \
    if (dt_ > (1ll << 31) && !rep_) { rep_ = true; const int i_ = atomicAdd(g_wd_buf, 1); \
        if (i_ < 255) { int *q_ = g_wd_buf + 4 + 4 * i_; q_[0] = __LINE__; q_[1] = (int)(a); q_[2] = (int)(b); q_[3] = (int)blockIdx.x; } \
        __threadfence_system(); } \
    if (dt_ > (3ll << 30)) __trap(); } } while (0)
#elif defined(VP9CUDA_HEARTBEAT)
/* least-intrusive tracing: every wait notes {line, id, at} in a host-mapped table slot of its
 * (block, warp) on entry and negates the line on exit; the host dumps the table when a frame
 * does not finish (vp9cuda_wait). */
__device__ int *g_wd_buf;
#define VP9_SPIN(cond, ns, tag, a, b) do { int *q_ = g_wd_buf + ((blockIdx.x & 63) * 32 + (threadIdx.x >> 5)) * 4; \
    q_[1] = (int)(a); q_[2] = (int)(b); q_[0] = __LINE__; \
    while (cond) { if ((ns) > 0) __nanosleep(ns); } q_[0] = -__LINE__; } while (0)
#else
/* Production build: a wait that lasts longer than VP9_SPIN_BUDGET clock cycles (~4 s) is a broken
 * dependency (the wavefronts rely on CTAs being dispatched in block-index order, which CUDA does not
 * promise).  The waiter then sets the context's error word -- f.err[0] in device memory, polled by
 * every other wait of the context once per 1024 polls, and f.err_host[0] in host-mapped memory, read
 * by vp9cuda_wait -- and gives up; every other wait gives up as soon as it sees the word, the launch
 * drains, and the host reports -EIO for the frame instead of sitting on a hung GPU. */
#ifndef VP9_SPIN_BUDGET
#define VP9_SPIN_BUDGET (8ll << 30)
#endif
#define VP9_SPIN(cond, ns, tag, a, b) do { unsigned n_ = 0; long long t0_ = 0; \
    while (cond) { if ((ns) > 0) __nanosleep(ns); \
        if ((++n_ & 1023u) == 0 && f.err) { \
            if (*(volatile int *)f.err) break; \
            const long long now_ = clock64(); \
            if (!t0_) t0_ = now_; \
            else if (now_ - t0_ > VP9_SPIN_BUDGET) { \
                atomicExch(f.err, __LINE__); *(volatile int *)f.err_host = __LINE__; __threadfence_system(); break; } } } } while (0)
#endif

#define LF_LL_WORDS 256                 /* 8 rows x 64 pixels x 2 bytes / 4 */
#define INTRA_LL_WORDS 64               /* one line: 64 luma + 2 x 32 chroma pixels x 2 bytes / 4 */
__device__ __forceinline__ void ll_store(unsigned long long *p, unsigned data, unsigned tag)
{
    *(volatile unsigned long long *)p = (unsigned long long)tag << 32 | data;
}
__device__ __forceinline__ unsigned long long ll_load(const unsigned long long *p) { return *(const volatile unsigned long long *)p; }

/* 16-byte asynchronous global -> shared copies (LDGSTS.128) and named barriers, with host stand-ins for
 * the emulation build (oracle/cuda_emu.h) */
#ifdef VP9_EMU
__device__ __forceinline__ void cp16(void *smem, const void *gmem) { memcpy(smem, gmem, 16); }
__device__ __forceinline__ void cp_commit() {}
template <int N> __device__ __forceinline__ void cp_wait() {}
template <int ID, int N> __device__ __forceinline__ void bar_named() { emu::named_barrier(ID, N); }
#else
__device__ __forceinline__ void cp16(void *smem, const void *gmem)
{
    unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }
template <int ID, int N> __device__ __forceinline__ void bar_named() { asm volatile("bar.sync %0, %1;\n" ::"n"(ID), "n"(N) : "memory"); }
#endif

template <class P> __device__ __forceinline__ P ld_cg(const P *p);
template <> __device__ __forceinline__ uint8_t ld_cg<uint8_t>(const uint8_t *p)
{
    return __ldcg(p);
}
template <> __device__ __forceinline__ uint16_t ld_cg<uint16_t>(const uint16_t *p)
{
    return __ldcg(p);
}

} /* namespace vp9 */
#endif
