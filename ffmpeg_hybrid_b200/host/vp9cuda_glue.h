/*
 * vp9cuda_glue.h -- host side of the vp9cuda back end, compiled INTO libavcodec next to vp9.c
 * (it takes the place of the fork's vp9_webgpu.[ch]; same seam).
 *
 * The hooks below are called from the sites where the reference's fork attached its WebGPU
 * prototype (vp9.c:1319-1324, 1839-1845, 1905-1917, 1281-1292; vp9block.c:1389-1405;
 * vp9lpf.c:188-208).  With the back end active no pixel is touched on the host: the hooks only
 * write records into the pinned lane rings of include/vp9cuda.h.
 */
#ifndef AVCODEC_VP9CUDA_GLUE_H
#define AVCODEC_VP9CUDA_GLUE_H

#include <stddef.h>

struct AVCodecContext;
struct AVFrame;
struct VP9Context;
struct VP9TileData;
struct VP9Filter;
typedef struct VP9CudaGlue VP9CudaGlue;

/* vp9_decode_init: attach the back end.  Fails (negative AVERROR) when libvp9cuda.so or a
 * B200 is missing -- there is deliberately no software fallback. */
int  ff_vp9cuda_init(struct AVCodecContext *avctx);
/* vp9_decode_free: first wait for every copy-out that may still write into a picture about to be
 * unreferenced (quiesce, before the frames are dropped), then detach (uninit) */
void ff_vp9cuda_quiesce(struct AVCodecContext *avctx);
void ff_vp9cuda_uninit(struct AVCodecContext *avctx);
/* vp9_decode_update_thread_context (frame threading, vp9.c:1923-1962): all workers share one back end */
int  ff_vp9cuda_update_thread_context(struct AVCodecContext *dst, const struct AVCodecContext *src);
/* vp9_decode_flush */
void ff_vp9cuda_flush(struct AVCodecContext *avctx);
/* vp9_decode_frame, after decode_frame_header and before vp9_frame_alloc: back end exists, CUR_FRAME
 * will come from the pinned pool */
int  ff_vp9cuda_prepare_frame(struct AVCodecContext *avctx);
/* vp9_decode_frame, after the header is parsed and CUR_FRAME allocated, before the tile loop and
 * before ff_thread_finish_setup: fixes the frame's place in decode order */
int  ff_vp9cuda_begin_frame(struct AVCodecContext *avctx);
/* vp9_decode_frame returned an error after begin_frame: give the frame's ring set back */
void ff_vp9cuda_abort_frame(struct AVCodecContext *avctx);
/* decode_tiles / decode_tiles_mt, after each superblock row of a tile column: its records may
 * start their way to the GPU */
void ff_vp9cuda_row_done(struct VP9TileData *td);
/* ff_vp9_decode_block, in place of ff_vp9_{intra,inter}_recon_{8,16}bpp + overhang copies */
void ff_vp9cuda_emit_block(struct VP9TileData *td);
/* ff_vp9_loopfilter_sb, in place of filter_plane_cols/rows */
void ff_vp9cuda_emit_lf(struct VP9Context *s, const struct VP9Filter *lflvl, int row, int col);
/* vp9_decode_frame, after the tile loop: submit, reconstruct, copy CUR_FRAME back (asynchronously
 * when pipelined) */
int  ff_vp9cuda_end_frame(struct AVCodecContext *avctx);
/* vp9_decode_frame's output sites (vp9.c:1636-1653 show_existing_frame, :1853-1857 end of frame)
 * and the drain call: `src` is the picture this call would normally return (NULL: none -- hidden
 * frame or drain).  Pipelined (default, $VP9CUDA_PIPELINE=0 turns it off), the picture is held back
 * until the next call so that its reconstruction and copy-out run beside the next frame's
 * entropy decode; the decoder therefore advertises AV_CODEC_CAP_DELAY.  `pkt` != NULL stamps the
 * picture with the packet's timestamps (show_existing_frame). */
struct AVPacket;
int  ff_vp9cuda_deliver(struct AVCodecContext *avctx, struct AVFrame *frame, int *got_frame,
                        struct AVFrame *src, const struct AVPacket *pkt);

#endif /* AVCODEC_VP9CUDA_GLUE_H */
