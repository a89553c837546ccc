/*
 * k_itx.cuh -- K1: batched inverse transform + add for the residual of inter blocks.
 *
 * One launch per transform size over that size's record list.  N = 4 << tx threads work on one
 * transform block: thread i runs the first-pass 1-D transform over block[i + x*N] (a coalesced
 * read of the column-major coefficient block), parks the int16_t/int32_t-truncated result in
 * shared memory, then runs the second pass over tmp[x*N + i] and adds column i to the picture
 * (row j of all N threads is one contiguous store).  DCT_DCT blocks with eob == 1 take the
 * DC-only shortcut exactly like the reference (vp9dsp_template.c:1165-1178).
 *
 * Replaces inter_recon's residual loop (vp9recon.c:707-763) + itxfm_add slots
 * (vp9dsp_template.c:1155-1778).  The same device function (itx_block) is reused by the intra
 * wavefront kernel.  Algorithmic bytes per block: N*N coefficients (2 or 4 B each) + 2*N*N*Bpp.
 */
#ifndef VP9CUDA_K_ITX_CUH
#define VP9CUDA_K_ITX_CUH

#include "dev_common.cuh"

namespace vp9 {

template <bool HBD> struct ItxTypes;
template <> struct ItxTypes<false> { typedef W32 T; typedef int16_t C; };
template <> struct ItxTypes<true>  { typedef W64 T; typedef int32_t C; };

/* Full 2-D transform of one block by N cooperating threads (lane = 0..N-1, all N must call).
 * blk: coefficients (global); tmp: N*(N+1) elements of shared memory owned by this block;
 * Pix: accessor with int get(x,y) / void put(x,y,v) relative to the block origin.
 * `sync` synchronises the N threads (warp-level; N <= 32 and the threads sit in one warp). */
/* RES = true: instead of adding to the picture, put() receives the rounded residual clamped to
 * [-2^bpp, 2^bpp] -- clip_pixel(pred + r) is the same for every pred in [0, 2^bpp) whether or not
 * r was clamped to that range, so a later `clip_pixel(pred + residual)` is bit-exact. */
template <bool HBD, int N, class Pix, bool RES = false>
__device__ __forceinline__ void itx_block(const typename ItxTypes<HBD>::C *__restrict__ blk,
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

/* picture accessor on global memory */
template <class P> struct GlobalPix {
    P *p; int stride;
    __device__ __forceinline__ int get(int x, int y) const { return p[(size_t)y * stride + x]; }
    __device__ __forceinline__ void put(int x, int y, int v) { p[(size_t)y * stride + x] = (P)v; }
};

template <class P, bool HBD, int N>
__global__ void __launch_bounds__(128) k_itx(const __grid_constant__ FrameDesc f)
{
    typedef typename ItxTypes<HBD>::C C;
    constexpr int tx = N == 4 ? 0 : N == 8 ? 1 : N == 16 ? 2 : 3;
    constexpr int BPC = 128 / N;                    /* transform blocks per CTA */
    __shared__ C tmp_all[BPC * N * (N + 1)];
    const int sub = threadIdx.x / N, lane = threadIdx.x % N;
    const uint32_t idx = blockIdx.x * BPC + sub;
    if (idx >= f.n_tx[tx])
        return;
    const Vp9CudaTx r = f.tx[tx][idx];
    GlobalPix<P> pix;
    pix.stride = f.cur.stride[r.plane];
    pix.p = (P *)f.cur.pl[r.plane] + (size_t)r.y4 * 4 * pix.stride + r.x4 * 4;
    /* the N threads of one block are contiguous lanes of one warp */
    const unsigned wl = (threadIdx.x & 31) & ~(N - 1);
    const unsigned mask = N == 32 ? 0xffffffffu : (((1u << N) - 1u) << wl);
    itx_block<HBD, N>((const C *)f.coef + r.coef_off, tmp_all + sub * N * (N + 1), lane, r.txtp, r.eob,
                      f.bpp, pix, mask);
}

} /* namespace vp9 */
#endif
