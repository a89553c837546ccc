/*
 * vp9cuda.h -- C ABI of libvp9cuda.so, the B200 (sm_100a) VP9 reconstruction back end.
 *
 * This is the drop-in boundary for the reference's per-block reconstruction path
 * (SURVEY.md section 8b).  The host keeps libavcodec's serial entropy decode
 * (vp9.c decode_frame / vp9block.c decode_coeffs) and, instead of calling the
 * VP9DSPContext slots, writes records into pinned "lane rings" and hands them over here.
 * Every entry point names the reference interface it replaces (paths relative to the
 * reference tree).  Plain pointers and sizes only; no CUDA or torch types.
 *
 * Error model: 0 on success, negative errno-style code otherwise (AVERROR-compatible:
 * -EINVAL, -ENOMEM, -EIO for a CUDA failure, -ENODEV when no sm_100 GPU is present).
 * Nothing here ever falls back to a CPU implementation.
 *
 * Threading: one Vp9CudaCtx per decoder instance (= per stream), bound to one GPU and its own
 * CUDA streams.  begin_frame comes from one thread at a time in decode order; everything that
 * names a frame_id may come from any thread (libavcodec frame threading, pthread_frame.c, runs
 * one frame per worker).  A lane ring has exactly one producer thread; submit_tile may be called
 * concurrently for DIFFERENT lanes (slice threading, vp9.c:1442-1520: one lane per tile column).
 */
#ifndef VP9CUDA_H
#define VP9CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)   /* the library is built with -fvisibility=hidden */
#endif

#define VP9CUDA_ABI_VERSION 3
#define VP9CUDA_MAX_LANES   64   /* max tile columns, vp9.c:800-810 */

typedef struct Vp9CudaCtx Vp9CudaCtx;

/* ---- records (host writes, GPU reads) -------------------------------------------------- */

/* One per inter-coded block: what inter_pred() reads from VP9Block (vp9_mc_template.c:30-464,
 * vp9dec.h:89-97).  Sub-block / chroma motion vectors are derived on the GPU. */
typedef struct Vp9CudaInterBlock {
    uint16_t col, row;       /* td->col, td->row: position in 8x8 units */
    uint8_t  bs;             /* enum BlockSize, BS_64x64=0 .. BS_4x4=12 (vp9shared.h:86-101) */
    uint8_t  filter;         /* enum FilterMode 0 smooth, 1 regular, 2 sharp, 3 bilinear */
    uint8_t  comp;           /* compound prediction */
    uint8_t  ref[2];         /* b->ref[]: 0..2, index into Vp9CudaFrameHeader.ref_slot */
    uint8_t  pad[3];
    int16_t  mv[4][2][2];    /* b->mv[sub-block][ref] as {x, y}, 1/8 luma pel */
    uint32_t reserved;
} Vp9CudaInterBlock;         /* 48 bytes */

/* One per transform block with eob > 0 of an INTER block: the itxfm_add call of
 * inter_recon() (vp9recon.c:707-763).  One list per lane, in coding order, so that the blocks of
 * one superblock are contiguous (Vp9CudaSbIndex.tx_start / tx_count). */
typedef struct Vp9CudaTx {
    uint16_t x4, y4;         /* position inside the plane, in 4-pixel units */
    uint16_t eob;
    uint16_t ncoef;          /* coefficient entries of this block (non-zero values, <= eob) */
    uint8_t  plane;          /* 0 Y, 1 U, 2 V */
    uint8_t  tx;             /* enum TxfmMode 0..3 */
    uint8_t  txtp;           /* enum TxfmType | 0x80 when lossless (WHT, slot index 4) */
    uint8_t  pad;
    uint32_t coef_off;       /* first entry in the frame's coefficient stream
                                (= lane ring's coef_base + offset inside the lane) */
} Vp9CudaTx;                 /* 16 bytes */

/* One per transform block of an INTRA block, in the order intra_recon() walks them
 * (vp9recon.c:269-358): prediction (check_intra_mode, vp9recon.c:37-221) then itxfm_add. */
typedef struct Vp9CudaIntraTx {
    uint16_t x4, y4;
    uint16_t eob;            /* 0 = prediction only */
    uint16_t level;          /* depth of this block's dependency chain in the frame: 0 when no
                                neighbour its predictor reads is an intra block of this frame,
                                else 1 + the largest level among those neighbours.  The
                                neighbours are taken by mode, after check_intra_mode's
                                conversion for missing edges (vp9recon.c:49-70): left column,
                                top-left cell, top row, and the top-right cell for 4x4 blocks
                                with HAVE_RIGHT whose mode reads it.  Blocks of equal level are
                                independent; the GPU dispatches in level order. */
    uint8_t  plane;          /* 0 Y, 1 U, 2 V */
    uint8_t  tx;             /* enum TxfmMode 0..3 */
    uint8_t  mode;           /* signalled IntraPredMode 0..9 (edge remap happens on the GPU) */
    uint8_t  txtp;           /* enum TxfmType | 0x80 when lossless */
    uint8_t  flags;          /* VP9CUDA_ITX_* */
    uint8_t  pad;
    uint16_t ncoef;          /* coefficient entries of this block */
    uint32_t coef_off;
} Vp9CudaIntraTx;            /* 20 bytes */
#define VP9CUDA_ITX_HAVE_LEFT  1   /* col > tile_col_start || x > 0  (vp9recon.c:46) */
#define VP9CUDA_ITX_HAVE_RIGHT 2   /* x < w - 1 inside the coding block (vp9recon.c:47) */

/* Coefficient stream: only the NON-ZERO dequantised coefficients of a coded transform block
 * travel, each with its position in the column-major N x N block decode_coeffs fills
 * (`coef[rc]`, vp9block.c:912-917; rc = x * N + y).  The reference ships nothing -- it keeps the
 * dense block in td->block -- and checkasm bounds the block by eob (copy_subcoefs,
 * tests/checkasm/vp9dsp.c:253-296); here the eob-long scan prefix is walked once on the host.
 * 8-bit streams use Vp9CudaCoef16 (dctcoef = int16_t), 10/12-bit Vp9CudaCoef32 (int32_t,
 * bit_depth_template.c:82). */
typedef struct Vp9CudaCoef16 { int16_t val; uint16_t pos; } Vp9CudaCoef16;   /* 4 bytes */
typedef struct Vp9CudaCoef32 { int32_t val; uint32_t pos; } Vp9CudaCoef32;   /* 8 bytes */

/* Per 64x64 superblock: where its transform blocks sit in the producing lane's lists. */
typedef struct Vp9CudaSbIndex {
    uint32_t itx_start;      /* intra transform blocks (Vp9CudaIntraTx) */
    uint32_t tx_start;       /* residual blocks of inter blocks (Vp9CudaTx) */
    uint16_t itx_count;
    uint16_t tx_count;
    uint8_t  lane;
    uint8_t  pad[3];
} Vp9CudaSbIndex;            /* 16 bytes */

/* Byte-for-byte VP9Filter (vp9dec.h:83-87): level[64] then mask[2][2][8][4]. */
typedef struct Vp9CudaLoopFilterSb {
    uint8_t level[64];
    uint8_t mask[2][2][8][4];
} Vp9CudaLoopFilterSb;       /* 192 bytes */

/* ---- per-frame header ------------------------------------------------------------------- */

/* What the GPU needs from VP9BitstreamHeader / VP9Context after decode_frame_header()
 * (vp9.c:519-1113).  Replaces the fork's ff_vp9_webgpu_begin_frame (vp9_webgpu.h:477). */
typedef struct Vp9CudaFrameHeader {
    uint32_t w, h;               /* coded frame size in pixels */
    uint8_t  bpp;                /* 8, 10 or 12 */
    uint8_t  ss_h, ss_v;         /* 4:2:0 (1,1) or 4:4:4 (0,0); ss_h != ss_v is refused */
    uint8_t  keyframe, intraonly, lossless, invisible;
    uint8_t  lf_level;           /* s->s.h.filter.level; 0 = loop filter off */
    uint8_t  refreshrefmask;     /* vp9.c:1687-1692 */
    uint8_t  ref_slot[3];        /* s->s.h.refidx[0..2]: reference slot 0..7 */
    uint16_t mvscale[3][2];      /* s->mvscale, 0 = unscaled (vp9.c:857-872) */
    uint8_t  mvstep[3][2];       /* s->mvstep */
    uint8_t  n_lanes;            /* producer lanes this frame (1, or tile columns when slice-threaded) */
    uint8_t  pad[3];
    uint32_t lane_sbs[VP9CUDA_MAX_LANES]; /* superblocks each lane will produce (sizes its ring);
                                             their sum must not exceed the frame's superblock count */
    uint8_t  lim_lut[64];        /* s->filter_lut (vp9.c:673-686) */
    uint8_t  mblim_lut[64];
} Vp9CudaFrameHeader;

/* Pinned host memory of one producer lane for the current frame.  Capacities per superblock
 * of the lane: 64 blocks, 384 + 384 transform blocks, 6144 coefficient entries (= every
 * coefficient of every block non-zero), so a legal stream can never overflow a ring. */
typedef struct Vp9CudaLaneRing {
    Vp9CudaInterBlock *blk;  uint32_t blk_cap;
    Vp9CudaTx *tx;           uint32_t tx_cap;
    Vp9CudaIntraTx *itx;     uint32_t itx_cap;
    void *coef;              uint32_t coef_cap;    /* Vp9CudaCoef16 (8-bit) or Vp9CudaCoef32 entries */
    uint32_t coef_base;      /* index of coef[0] in the frame's coefficient stream: every
                                coef_off the producer writes is coef_base + offset-in-lane */
    uint32_t reserved;
} Vp9CudaLaneRing;

typedef struct Vp9CudaLaneCounts {
    uint32_t blk, tx, itx, coef;
} Vp9CudaLaneCounts;

/* Frame-level pinned arrays, one entry per superblock in raster order. */
typedef struct Vp9CudaFrameRing {
    Vp9CudaSbIndex *sb_index;
    Vp9CudaLoopFilterSb *lf;
    uint32_t n_sb;
    uint32_t frame_id;       /* names this frame in the calls below (frames may overlap in time
                                under libavcodec frame threading, pthread_frame.c) */
} Vp9CudaFrameRing;

/* ---- life cycle -------------------------------------------------------------------------- */

/* Replaces ff_vp9_webgpu_init (vp9_webgpu.h:371; call site vp9.c:1905-1917).
 * max_w x max_h at max_bpp (8, 10 or 12) bound every frame of the stream: all pinned rings and
 * device record buffers of the frames_in_flight (1..32) ring sets are allocated here, so that no
 * allocation happens on the frame path (vp9cuda_get_stat "allocs" stays constant).  With
 * max_w == 0 or max_h == 0 the buffers grow on demand instead (a frame larger than any before
 * reallocates; never in the steady state of a stream). */
int  vp9cuda_create(Vp9CudaCtx **out, int device, int max_w, int max_h, int max_bpp,
                    int frames_in_flight);
/* Replaces ff_vp9_webgpu_uninit (vp9_webgpu.h:374; call site vp9.c:1281-1292). */
void vp9cuda_destroy(Vp9CudaCtx *ctx);
/* Replaces vp9_decode_flush's reference drop (vp9.c:1865-1883).  Frames that were begun and
 * not finished are aborted. */
int  vp9cuda_flush(Vp9CudaCtx *ctx);

/* ---- per frame -----------------------------------------------------------------------------
 * Frames may overlap in time (libavcodec frame threading, pthread_frame.c: one vp9_decode_frame
 * per worker, only the serial entropy decode runs on the host): begin_frame fixes the DECODE
 * ORDER, every call after it names its frame by frame_id and may come from any thread, and
 * finish_frame hands the frames to the GPU in begin_frame order whatever order the workers
 * arrive in.  Every begun frame must end in finish_frame or abort_frame. */

/* Call after decode_frame_header + vp9_frame_alloc (vp9.c:1657-1692), before the tile loop, in
 * DECODE ORDER (one thread at a time; under frame threading: before ff_thread_finish_setup).
 * Blocks while all frames_in_flight ring sets are in use.  Fills *ring.
 * -EINVAL for a frame the back end cannot take (size 0, more than 16384 pixels wide, a bit
 * depth other than 8/10/12, chroma other than 4:2:0 / 4:4:4). */
int  vp9cuda_begin_frame(Vp9CudaCtx *ctx, const Vp9CudaFrameHeader *hdr, Vp9CudaFrameRing *ring);
/* Pinned ring of one lane of frame frame_id. */
int  vp9cuda_tile_ring(Vp9CudaCtx *ctx, uint32_t frame_id, int lane, Vp9CudaLaneRing *ring);
/* Records [previous counts, counts) of this lane are final: their host->device copy is enqueued
 * now, beside the entropy decode of the rest of the frame.  Called by the lane's producer at
 * tile / superblock-row granularity (decode_tiles, vp9.c:1325-1395; decode_tiles_mt,
 * vp9.c:1465-1518), any number of times.  Counts are cumulative for the frame and never
 * decrease; records below a submitted count must not be written again. */
int  vp9cuda_submit_tile(Vp9CudaCtx *ctx, uint32_t frame_id, int lane,
                         const Vp9CudaLaneCounts *counts);
/* All lanes submitted: copy what submit_tile has not yet sent plus the per-superblock arrays,
 * then enqueue inter prediction + residual -> intra wavefront -> loop filter -> reference
 * refresh and, when out_planes != NULL, the async copy of the visible w x h crop (framemd5
 * layout: packed rows, 10-bit as little-endian 16-bit).  Replaces the reconstruction half of
 * decode_tiles + ff_vp9_webgpu_end_frame (vp9.c:1839-1857).  Blocks until every frame begun
 * before this one has been finished or aborted.  Returns a ticket. */
int  vp9cuda_finish_frame(Vp9CudaCtx *ctx, uint32_t frame_id, uint8_t *const out_planes[3],
                          const ptrdiff_t out_strides[3], uint64_t *ticket);
/* The frame will not be finished (the host found the packet damaged between begin_frame and
 * finish_frame: vp9.c:1340-1346, 1389-1391, 1827-1832): give its ring set back and let the
 * frames behind it proceed.  Reference slots stay as they were (vp9.c:1860-1862 `fail:`). */
int  vp9cuda_abort_frame(Vp9CudaCtx *ctx, uint32_t frame_id);
/* Block until the frame behind `ticket` is complete (pixels in out_planes valid). */
int  vp9cuda_wait(Vp9CudaCtx *ctx, uint64_t ticket);
/* Device-resident output (the output site is vp9.c:1853-1857): hand the finished
 * picture to a GPU consumer without the device->host copy.  slot 0..7 = reference slot, slot < 0
 * = the most recently reconstructed frame.  Blocks until that picture is final, then returns
 * device pointers to the three planes (visible w x h at the top left of a 64-aligned surface),
 * their strides in bytes, and the format.  The surface stays untouched until
 * vp9cuda_unmap_frame(planes[0]); every map needs one unmap.  0 or negative errno. */
typedef struct Vp9CudaMappedFrame {
    void     *planes[3];
    ptrdiff_t strides[3];    /* bytes */
    int       w, h, bpp;     /* luma size; 8 -> uint8_t samples, 10/12 -> uint16_t */
    int       ss_h, ss_v;    /* chroma subsampling (1, 1 = 4:2:0) */
} Vp9CudaMappedFrame;
int  vp9cuda_map_frame(Vp9CudaCtx *ctx, int slot, Vp9CudaMappedFrame *out);
int  vp9cuda_unmap_frame(Vp9CudaCtx *ctx, const void *plane0);
/* show_existing_frame (vp9.c:1636-1653): copy reference slot `slot` to host. */
int  vp9cuda_show_existing(Vp9CudaCtx *ctx, int slot, uint8_t *const out_planes[3],
                           const ptrdiff_t out_strides[3], uint64_t *ticket);

/* ---- one-shot form: the caller owns the record memory ----------------------------------------- */

/* All records of one frame, in caller memory laid out exactly like the lane rings above
 * (lanes[i].coef_base places lane i in the coefficient arena; capacities are ignored). */
typedef struct Vp9CudaFrameRecords {
    const Vp9CudaLaneRing *lanes;        /* hdr->n_lanes entries */
    const Vp9CudaLaneCounts *counts;     /* hdr->n_lanes entries */
    const Vp9CudaSbIndex *sb_index;      /* one per superblock, raster order */
    const Vp9CudaLoopFilterSb *lf;       /* one per superblock, raster order */
} Vp9CudaFrameRecords;

/* Page-locked host memory for records and output planes (pageable memory also works, slower). */
void *vp9cuda_host_alloc(size_t bytes);
void  vp9cuda_host_free(void *p);

/* begin_frame + submit_tile (all lanes) + finish_frame in one call, from caller memory:
 * host->device copy of the records, reconstruction, optional device->host copy of the
 * picture.  The record memory may be reused once the ticket has been waited for. */
int  vp9cuda_decode_frame(Vp9CudaCtx *ctx, const Vp9CudaFrameHeader *hdr,
                          const Vp9CudaFrameRecords *recs, uint8_t *const out_planes[3],
                          const ptrdiff_t out_strides[3], uint64_t *ticket);

/* ---- measurement / test support ------------------------------------------------------------- */

/* decode_frame split in two so a bench can keep records resident in HBM:
 * stage = copy one frame's records host->device and keep them under `handle`;
 * run   = reconstruct a staged frame (any number of times, in stream order). */
int  vp9cuda_stage_frame(Vp9CudaCtx *ctx, const Vp9CudaFrameHeader *hdr,
                         const Vp9CudaFrameRecords *recs, int *handle);
int  vp9cuda_run_staged(Vp9CudaCtx *ctx, int handle, uint8_t *const out_planes[3],
                        const ptrdiff_t out_strides[3], uint64_t *ticket);
int  vp9cuda_free_staged(Vp9CudaCtx *ctx, int handle);
/* MD5-free frame check: 64-bit FNV-style checksum of the visible planes of reference slot
 * `slot` (or of the last reconstructed frame when slot < 0), computed on the GPU. */
int  vp9cuda_frame_checksum(Vp9CudaCtx *ctx, int slot, uint64_t *sum);
/* Cumulative device time (CUDA events on the ctx's stream) and launch count per kernel family.
 * names/ms/launches hold up to `cap` entries; returns the number filled. */
int  vp9cuda_kernel_stats(Vp9CudaCtx *ctx, const char **names, double *ms, uint64_t *launches,
                          uint64_t *bytes, int cap);
/* keys: "timing" 0/1 (per-stage CUDA events on the ctx stream, no host syncs),
 * "reset_stats", "mark" = record timing mark 0..15 on the ctx stream, "sync" = drain the stream */
int  vp9cuda_set_option(Vp9CudaCtx *ctx, const char *key, int value);
/* profiling aid: after set_option("phase_clk", 1), SM clock cycles accumulated per phase of the
 * loop-filter wavefront CTAs (0 wait, 1 load, 2 column pass, 3 row pass, 4 store+publish) */
int  vp9cuda_phase_clocks(Vp9CudaCtx *ctx, uint64_t *out, int n);
/* device time between two marks (waits for mark_b) */
int  vp9cuda_elapsed_ms(Vp9CudaCtx *ctx, int mark_a, int mark_b, double *ms);
/* counters: "allocs" = device / pinned allocations made so far (constant once the stream is in
 * its steady state), "frames" = frames reconstructed, "aborted" = frames given back unfinished,
 * "h2d_bytes" / "d2h_bytes" = record bytes copied in / picture bytes copied out so far,
 * "early_h2d_bytes" = the part of h2d_bytes that submit_tile sent before finish_frame,
 * "ns_begin_frame" / "ns_submit_tile" / "ns_finish_frame" / "ns_wait" = host time spent inside
 * those entry points so far, nanoseconds, all calling threads summed ($VP9CUDA_HOSTPROF=1 prints
 * them when the context is destroyed) */
int  vp9cuda_get_stat(Vp9CudaCtx *ctx, const char *key, uint64_t *value);
const char *vp9cuda_version(void);

/* ---- slot-level entry points: the batched equivalent of one VP9DSPContext slot -------------
 * (vp9dsp.h:40-122).  Host buffers in and out; used by the checkasm-style parity tests
 * (reference tests/checkasm/vp9dsp.c).  Item i works on its own 64x64-pixel tile at
 * pix + i*64*64*bytes_per_pixel with a stride of 64 pixels. */
int  vp9cuda_slot_itxfm_add(int device, int bpp, int n, const uint8_t *tx, const uint8_t *txtp,
                            const uint16_t *eob, void *pix, const void *coef /* n*1024 */);
int  vp9cuda_slot_intra_pred(int device, int bpp, int n, const uint8_t *tx, const uint8_t *mode,
                             void *pix, const void *left /* n*64 */, const void *top /* n*(1+64+16) */);
int  vp9cuda_slot_mc(int device, int bpp, int n, const uint8_t *w, const uint8_t *h,
                     const uint8_t *filter, const uint8_t *avg, const uint8_t *mx,
                     const uint8_t *my, const uint8_t *dx, const uint8_t *dy /* 0 = unscaled */,
                     void *pix, const void *ref /* n tiles of 160x160 pixels, block origin at (8,8) */);
int  vp9cuda_slot_loop_filter(int device, int bpp, int n, const uint8_t *wd /* 4,8,16 */,
                              const uint8_t *dir, const uint8_t *len /* 8 or 16 */,
                              const int32_t *E, const int32_t *I, const int32_t *H,
                              void *pix /* edge at (32,32) of each tile */);

/* ---- allocation-free entry point --------------------------------------------------------------
 * No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Computes the checksum vp9cuda_frame_checksum defines (same value, bit for bit) of a picture
 * the CALLER holds in device memory -- e.g. the planes vp9cuda_map_frame returned -- and leaves it in
 * *sum_dev (DEVICE memory, 8 bytes, 8-aligned), in stream order on `stream` (a cudaStream_t, NULL =
 * the default stream); the call does not wait for the GPU.  It makes no cudaMalloc / cudaFree and
 * calls no allocating library routine: the kernel's scratch (one partial sum per picture row + a
 * ticket word) is a static __device__ array of VP9CUDA_NOALLOC_MAX_ROWS rows, about 128 KiB that
 * every GPU which loads the library holds until exit.  Its device address is looked up once per
 * device on the first call for that device (make that call outside a capture).
 *   planes / strides : device pointers and strides in BYTES of the three planes; w, h luma size;
 *                      bpp 8 (uint8_t samples) or 10 / 12 (uint16_t); ss_h, ss_v chroma shifts (0 / 1).
 * Returns 0, -EINVAL (arguments), -E2BIG when h + 2 * ((h + ss_v) >> ss_v) exceeds
 * VP9CUDA_NOALLOC_MAX_ROWS (also reported on stderr), -EIO (CUDA error).
 * Every call on a device shares the one scratch array: calls that overlap on the GPU (different
 * streams or host threads without ordering between them) race; calls ordered on one stream do not.
 * The call leaves `device` current on the calling thread, like the other entry points. */
#define VP9CUDA_NOALLOC_MAX_ROWS 16384      /* 8192x4320 4:2:0 has 8640, 4:4:4 12960 */
int  vp9cuda_checksum_noalloc(int device, const void *const planes[3], const ptrdiff_t strides[3],
                              int w, int h, int bpp, int ss_h, int ss_v, uint64_t *sum_dev, void *stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* VP9CUDA_H */
