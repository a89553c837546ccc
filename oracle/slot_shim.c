/*
 * slot_shim.c -- TEST INFRASTRUCTURE (oracle/): makes the reference's own VP9DSPContext slots
 * (libavcodec/vp9dsp.h:40-122, filled by ff_vp9dsp_init, vp9dsp.c:88-112 ->
 * vp9dsp_template.c:2571-2578) callable from ctypes, one plain-C entry per slot family.
 * Built by oracle/Makefile into oracle/_ref/libvp9dsp_ref.so against the unmodified
 * reference objects.  This is what the checkasm-style slot tests compare against
 * (reference tests/checkasm/vp9dsp.c:50-625 does the same C-vs-candidate memcmp).
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include "libavcodec/vp9dsp.h"
#include "libavcodec/videodsp.h"

static VP9DSPContext dsp[3];
static VideoDSPContext vdsp[3];
static int inited;

static int idx(int bpp)
{
    if (!inited) {
        for (int i = 0; i < 3; i++) {
            memset(&dsp[i], 0, sizeof(dsp[i]));
            ff_vp9dsp_init(&dsp[i], 8 + 2 * i, 1);
            ff_videodsp_init(&vdsp[i], 8 + 2 * i);
        }
        inited = 1;
    }
    return (bpp - 8) >> 1;
}

/* itxfm_add[tx 0..4 (4 = lossless WHT)][txtp 0..3]; zeroes `block` like the slot does */
void ref_itxfm_add(int bpp, int tx, int txtp, uint8_t *dst, ptrdiff_t stride, int16_t *block, int eob)
{
    dsp[idx(bpp)].itxfm_add[tx][txtp](dst, stride, block, eob);
}

/* intra_pred[tx 0..3][mode 0..14]; left is bottom-to-top, top[-1] is the top-left pixel */
void ref_intra_pred(int bpp, int tx, int mode, uint8_t *dst, ptrdiff_t stride,
                    const uint8_t *left, const uint8_t *top)
{
    dsp[idx(bpp)].intra_pred[tx][mode](dst, stride, left, top);
}

/* loop_filter_8[wd 0..2 = 4/8/16-wide][dir 0 = col edge (h), 1 = row edge (v)] */
void ref_loop_filter_8(int bpp, int wd, int dir, uint8_t *dst, ptrdiff_t stride, int E, int I, int H)
{
    dsp[idx(bpp)].loop_filter_8[wd][dir](dst, stride, E, I, H);
}

void ref_loop_filter_16(int bpp, int dir, uint8_t *dst, ptrdiff_t stride, int E, int I, int H)
{
    dsp[idx(bpp)].loop_filter_16[dir](dst, stride, E, I, H);
}

void ref_loop_filter_mix2(int bpp, int wd1, int wd2, int dir, uint8_t *dst, ptrdiff_t stride,
                          int E, int I, int H)
{
    dsp[idx(bpp)].loop_filter_mix2[wd1][wd2][dir](dst, stride, E, I, H);
}

/* mc[hsize 0..4 = 64..4][filter 0..3 = smooth/regular/sharp/bilinear][avg][!!mx][!!my] */
void ref_mc(int bpp, int hsize, int filter, int avg, uint8_t *dst, ptrdiff_t dst_stride,
            const uint8_t *ref, ptrdiff_t ref_stride, int h, int mx, int my)
{
    dsp[idx(bpp)].mc[hsize][filter][avg][!!mx][!!my](dst, dst_stride, ref, ref_stride, h, mx, my);
}

/* smc[hsize][filter][avg] */
void ref_smc(int bpp, int hsize, int filter, int avg, uint8_t *dst, ptrdiff_t dst_stride,
             const uint8_t *ref, ptrdiff_t ref_stride, int h, int mx, int my, int dx, int dy)
{
    dsp[idx(bpp)].smc[hsize][filter][avg](dst, dst_stride, ref, ref_stride, h, mx, my, dx, dy);
}

/* VideoDSPContext.emulated_edge_mc (videodsp_template.c:27-105) */
void ref_emulated_edge_mc(int bpp, uint8_t *buf, const uint8_t *src, ptrdiff_t buf_linesize,
                          ptrdiff_t src_linesize, int block_w, int block_h,
                          int src_x, int src_y, int w, int h)
{
    vdsp[idx(bpp)].emulated_edge_mc(buf, src, buf_linesize, src_linesize, block_w, block_h,
                                    src_x, src_y, w, h);
}
