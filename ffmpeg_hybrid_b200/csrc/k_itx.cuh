/*
 * k_itx.cuh -- K1: batched inverse transform + add for the residual of inter blocks.
 *
 * One warp per record of a lane's transform-block list.  The block's non-zero coefficients are
 * scattered into a dense block in shared memory (coef_expand); N = 4 << tx lanes then work on it:
 * lane i runs the first-pass 1-D transform over block[i + x*N], parks the int16_t/int32_t-truncated
 * result in shared memory, then runs the second pass over tmp[x*N + i] and adds column i to the
 * picture (row j of all N lanes is one contiguous store).  DCT_DCT blocks with eob == 1 take the
 * DC-only shortcut exactly like the reference (vp9dsp_template.c:1165-1178).
 *
 * Replaces inter_recon's residual loop (vp9recon.c:707-763) + itxfm_add slots
 * (vp9dsp_template.c:1155-1778).  The same device function (itx_block) is reused by the intra
 * wavefront kernel.  Algorithmic bytes per block: ncoef entries (4 or 8 B each) + 2*N*N*Bpp.
 */
#ifndef VP9CUDA_K_ITX_CUH
#define VP9CUDA_K_ITX_CUH

#include "dev_common.cuh"

namespace vp9 {

template <bool HBD> struct ItxTypes;
template <> struct ItxTypes<false> { typedef W32 T; typedef int16_t C; typedef Vp9CudaCoef16 E; };
template <> struct ItxTypes<true>  { typedef W64 T; typedef int32_t C; typedef Vp9CudaCoef32 E; };

/* The coefficient stream carries only a block's non-zero coefficients (Vp9CudaCoef16/32: value +
 * position in the column-major block).  N cooperating lanes rebuild the dense N x N block the
 * transform reads in `dense` (shared memory, 16-byte aligned): clear, scatter. */
template <bool HBD, int N>
__device__ __forceinline__ void coef_expand(const typename ItxTypes<HBD>::E *__restrict__ ent, int ncoef,
                                            typename ItxTypes<HBD>::C *dense, int lane, unsigned syncmask)
{
    typedef typename ItxTypes<HBD>::E E;
    constexpr int VEC = N * N * (int)sizeof(typename ItxTypes<HBD>::C) / 16;
    for (int i = lane; i < VEC; i += N)
        ((uint4 *)dense)[i] = make_uint4(0, 0, 0, 0);
    __syncwarp(syncmask);
    for (int i = lane; i < ncoef; i += N) {
        const E e = ent[i];
        dense[e.pos & (N * N - 1)] = e.val;
    }
    __syncwarp(syncmask);
}

/* itx_block fed from the coefficient stream; DC-only blocks skip the scatter */
template <bool HBD, int N, class Pix, bool RES = false>
__device__ __forceinline__ void itx_entries(const typename ItxTypes<HBD>::E *__restrict__ ent, int ncoef,
                                            typename ItxTypes<HBD>::C *dense, typename ItxTypes<HBD>::C *tmp,
                                            int lane, int txtp, int eob, int bpp, Pix &pix, unsigned syncmask);

/* Full 2-D transform of one block by N cooperating threads (lane = 0..N-1, all N must call).
 * blk: coefficients (global); tmp: N*(N+1) elements of shared memory owned by this block;
 * Pix: accessor with int get(x,y) / void put(x,y,v) relative to the block origin.
 * `sync` synchronises the N threads (warp-level; N <= 32 and the threads sit in one warp). */
/* RES = true: instead of adding to the picture, put() receives the rounded residual clamped to
 * [-2^bpp, 2^bpp] -- clip_pixel(pred + r) is the same for every pred in [0, 2^bpp) whether or not
 * r was clamped to that range, so a later `clip_pixel(pred + residual)` is bit-exact. */
/* `tmp` may be the same memory as `blk` (the fused inter kernel has room for one block per warp):
 * every lane has its inputs in registers before any lane writes an intermediate. */
template <bool HBD, int N, class Pix, bool RES = false>
__device__ __forceinline__ void itx_block(const typename ItxTypes<HBD>::C *blk,
                                          typename ItxTypes<HBD>::C *tmp, int lane, int txtp, int eob,
                                          int bpp, Pix &pix, unsigned syncmask)
{
    typedef typename ItxTypes<HBD>::T T;
    typedef typename ItxTypes<HBD>::C C;
    constexpr int tx = N == 4 ? 0 : N == 8 ? 1 : N == 16 ? 2 : 3;
    const bool lossless = txtp & 0x80;
    txtp &= 3;
    const int bits = itx_bits(tx, lossless);
    if (!lossless && eob == 1 && (txtp == 0 || tx == 3)) {
        int r = itx_round(itx_dc_only<HBD>(blk[0]), bits);
#pragma unroll 4
        for (int j = 0; j < N; j++)
            pix.put(lane, j, RES ? iclamp(r, -(1 << bpp), 1 << bpp) : clip_pixel(pix.get(lane, j) + r, bpp));
        return;
    }
    if (N == 4 && lossless) {
        W32 in[4], out[4];
#pragma unroll
        for (int x = 0; x < 4; x++) in[x] = W32((int64_t)blk[lane + x * 4]);
        iwht4(in, 1, out, 0);
        __syncwarp(syncmask);
#pragma unroll
        for (int x = 0; x < 4; x++) tmp[lane * 5 + x] = (C)out[x].v;
        __syncwarp(syncmask);
#pragma unroll
        for (int x = 0; x < 4; x++) in[x] = W32((int64_t)tmp[x * 5 + lane]);
        iwht4(in, 1, out, 1);
#pragma unroll
        for (int j = 0; j < 4; j++)
            pix.put(lane, j, RES ? iclamp((int32_t)(C)out[j].v, -(1 << bpp), 1 << bpp)
                                 : clip_pixel(pix.get(lane, j) + (int32_t)(C)out[j].v, bpp));
        return;
    }
    {
        T in[N], out[N];
#pragma unroll
        for (int x = 0; x < N; x++) in[x] = T((int64_t)blk[lane + x * N]);
        itx_1d<T, N>(in, out, (txtp & 1) && N < 32);
        __syncwarp(syncmask);
#pragma unroll
        for (int x = 0; x < N; x++) tmp[lane * (N + 1) + x] = (C)out[x].v;
        __syncwarp(syncmask);
#pragma unroll
        for (int x = 0; x < N; x++) in[x] = T((int64_t)tmp[x * (N + 1) + lane]);
        itx_1d<T, N>(in, out, (txtp & 2) && N < 32);
#pragma unroll
        for (int j = 0; j < N; j++)
            pix.put(lane, j, RES ? iclamp(itx_round((int32_t)(C)out[j].v, bits), -(1 << bpp), 1 << bpp)
                                 : clip_pixel(pix.get(lane, j) + itx_round((int32_t)(C)out[j].v, bits), bpp));
    }
}

template <bool HBD, int N, class Pix, bool RES>
__device__ __forceinline__ void itx_entries(const typename ItxTypes<HBD>::E *__restrict__ ent, int ncoef,
                                            typename ItxTypes<HBD>::C *dense, typename ItxTypes<HBD>::C *tmp,
                                            int lane, int txtp, int eob, int bpp, Pix &pix, unsigned syncmask)
{
    typedef typename ItxTypes<HBD>::C C;
    constexpr int tx = N == 4 ? 0 : N == 8 ? 1 : N == 16 ? 2 : 3;
    if (!(txtp & 0x80) && eob == 1 && ((txtp & 3) == 0 || tx == 3)) {
        /* the one coefficient of an eob == 1 block is the DC (scan position 0) */
        C dc = 0;
        if (ncoef > 0 && ent[0].pos == 0)
            dc = ent[0].val;
        itx_block<HBD, N, Pix, RES>(&dc, tmp, lane, txtp, eob, bpp, pix, syncmask);
        return;
    }
    coef_expand<HBD, N>(ent, ncoef, dense, lane, syncmask);
    itx_block<HBD, N, Pix, RES>(dense, tmp, lane, txtp, eob, bpp, pix, syncmask);
}

/* picture accessor on global memory */
template <class P> struct GlobalPix {
    P *p; int stride;
    __device__ __forceinline__ int get(int x, int y) const { return p[(size_t)y * stride + x]; }
    __device__ __forceinline__ void put(int x, int y, int v) { p[(size_t)y * stride + x] = (P)v; }
};

/* one warp per transform block of the lane's list (grid.y = producer lane), N = 4 << tx lanes active */
template <class P, bool HBD>
__global__ void __launch_bounds__(128) k_itx(const __grid_constant__ FrameDesc f)
{
    typedef typename ItxTypes<HBD>::C C;
    typedef typename ItxTypes<HBD>::E E;
    __shared__ __align__(16) C dense_all[4][32 * 32];
    __shared__ C tmp_all[4][32 * 33];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, ln = blockIdx.y;
    const uint32_t k = blockIdx.x * 4 + warp;
    if (k >= f.lane_tx_cnt[ln])
        return;
    const Vp9CudaTx r = f.tx[f.lane_tx_base[ln] + k];
    GlobalPix<P> pix;
    pix.stride = f.cur.stride[r.plane];
    pix.p = (P *)f.cur.pl[r.plane] + (size_t)r.y4 * 4 * pix.stride + r.x4 * 4;
    const E *ent = (const E *)f.coef + r.coef_off;
    C *dense = dense_all[warp], *tmp = tmp_all[warp];
    switch (r.tx) {
    case 0: if (lane < 4)  itx_entries<HBD, 4>(ent, r.ncoef, dense, tmp, lane, r.txtp, r.eob, f.bpp, pix, 0xfu); break;
    case 1: if (lane < 8)  itx_entries<HBD, 8>(ent, r.ncoef, dense, tmp, lane, r.txtp, r.eob, f.bpp, pix, 0xffu); break;
    case 2: if (lane < 16) itx_entries<HBD, 16>(ent, r.ncoef, dense, tmp, lane, r.txtp, r.eob, f.bpp, pix, 0xffffu); break;
    default: itx_entries<HBD, 32>(ent, r.ncoef, dense, tmp, lane, r.txtp, r.eob, f.bpp, pix, 0xffffffffu); break;
    }
}

} /* namespace vp9 */
#endif
