/*
 * vp9_lf.h -- VP9 loop filter: the per-line filter and the decoding of one superblock's
 * VP9Filter masks into an explicit table of edge operations.
 *
 * Reference behaviour reproduced:
 *   filter arithmetic   libavcodec/vp9dsp_template.c:1780-1888 (loop_filter), wrappers :1890-1967
 *   mask / level walk   libavcodec/vp9lpf.c:31-181 (filter_plane_cols / filter_plane_rows),
 *                       order per superblock and plane: all column edges, then all row edges
 *                       (vp9lpf.c:222-229)
 *   E/I/H from level    vp9lpf.c:47-48 with s->filter_lut (vp9.c:673-686)
 *
 * An "op" says: on these 8 lines, at this edge position, run a filter of width wd (4, 8 or 16)
 * with level L.  Ops on the same line are applied in increasing position, exactly the order
 * in which the reference reaches them.
 */
#ifndef VP9_LF_H
#define VP9_LF_H

#include "vp9_itx.h"

namespace vp9 {

/* px[0..15] = p7..p0, q0..q7 across the edge (only px[4..11] are read for wd < 16).
 * Returns a bit mask of modified entries (bit i = px[i] written).
 * E, I, H are the 8-bit-domain thresholds (mblim_lut[L], lim_lut[L], L >> 4).
 * Same decisions and outputs as loop_filter() (vp9dsp_template.c:1780-1888); the 15- and 7-tap
 * smoothers are evaluated as running window sums (each output's window is the previous one
 * shifted by a pixel, ends replicated), which is the same integer sum the reference spells out. */
VP9_HD unsigned lf_line(int *px, int wd, int E, int I, int H, int bpp)
{
    const int sh = bpp - 8, F = 1 << sh;
    E <<= sh; I <<= sh; H <<= sh;
    const int p3 = px[4], p2 = px[5], p1 = px[6], p0 = px[7];
    const int q0 = px[8], q1 = px[9], q2 = px[10], q3 = px[11];
#define VP9_ABS(v) ((v) < 0 ? -(v) : (v))
#define VP9_MAX(a, b) ((a) > (b) ? (a) : (b))
    const int a_p1p0 = VP9_ABS(p1 - p0), a_q1q0 = VP9_ABS(q1 - q0);
    int m = VP9_MAX(VP9_ABS(p3 - p2), VP9_ABS(p2 - p1));
    int m2 = VP9_MAX(VP9_ABS(q2 - q1), VP9_ABS(q3 - q2));
    m = VP9_MAX(m, a_p1p0); m2 = VP9_MAX(m2, a_q1q0);
    m = VP9_MAX(m, m2);
    if (m > I || VP9_ABS(p0 - q0) * 2 + (VP9_ABS(p1 - q1) >> 1) > E)
        return 0;
    if (wd >= 8) {
        int fi = VP9_MAX(VP9_ABS(p3 - p0), VP9_ABS(p2 - p0));
        int fi2 = VP9_MAX(VP9_ABS(q2 - q0), VP9_ABS(q3 - q0));
        fi = VP9_MAX(fi, a_p1p0); fi2 = VP9_MAX(fi2, a_q1q0);
        if (VP9_MAX(fi, fi2) <= F) {
            if (wd >= 16) {
                int fo = VP9_MAX(VP9_ABS(px[0] - p0), VP9_ABS(px[1] - p0));
                int fo2 = VP9_MAX(VP9_ABS(px[2] - p0), VP9_ABS(px[3] - p0));
                int fo3 = VP9_MAX(VP9_ABS(px[12] - q0), VP9_ABS(px[13] - q0));
                int fo4 = VP9_MAX(VP9_ABS(px[14] - q0), VP9_ABS(px[15] - q0));
                fo = VP9_MAX(fo, fo2); fo3 = VP9_MAX(fo3, fo4);
                if (VP9_MAX(fo, fo3) <= F) {
                    /* 15-tap: out[i] = (sum_{k=-7..7} v[clamp(i+k, 0, 15)] + v[i] + 8) >> 4, i = 1..14 */
                    const int v0 = px[0], v1 = px[1], v2 = px[2], v3 = px[3], v12 = px[12], v13 = px[13],
                              v14 = px[14], v15 = px[15];
                    int s = 7 * v0 + 2 * v1 + v2 + v3 + p3 + p2 + p1 + p0 + q0 + 8;
                    px[1] = s >> 4;  s += v2 - v1 + q1 - v0;
                    px[2] = s >> 4;  s += v3 - v2 + q2 - v0;
                    px[3] = s >> 4;  s += p3 - v3 + q3 - v0;
                    px[4] = s >> 4;  s += p2 - p3 + v12 - v0;
                    px[5] = s >> 4;  s += p1 - p2 + v13 - v0;
                    px[6] = s >> 4;  s += p0 - p1 + v14 - v0;
                    px[7] = s >> 4;  s += q0 - p0 + v15 - v0;
                    px[8] = s >> 4;  s += q1 - q0 + v15 - v1;
                    px[9] = s >> 4;  s += q2 - q1 + v15 - v2;
                    px[10] = s >> 4; s += q3 - q2 + v15 - v3;
                    px[11] = s >> 4; s += v12 - q3 + v15 - p3;
                    px[12] = s >> 4; s += v13 - v12 + v15 - p2;
                    px[13] = s >> 4; s += v14 - v13 + v15 - p1;
                    px[14] = s >> 4;
                    return 0x7FFE;
                }
            }
            /* 7-tap: out[i] = (sum_{k=-3..3} w[clamp(i+k, 0, 7)] + w[i] + 4) >> 3 over w = p3..q3 */
            int s = 3 * p3 + 2 * p2 + p1 + p0 + q0 + 4;
            px[5] = s >> 3;  s += p1 - p2 + q1 - p3;
            px[6] = s >> 3;  s += p0 - p1 + q2 - p3;
            px[7] = s >> 3;  s += q0 - p0 + q3 - p3;
            px[8] = s >> 3;  s += q1 - q0 + q3 - p2;
            px[9] = s >> 3;  s += q2 - q1 + q3 - p1;
            px[10] = s >> 3;
            return 0x07E0;
        }
    }
    {
        const int lim_hi = (1 << (bpp - 1)) - 1, lim_lo = -(1 << (bpp - 1));
        const bool hev = a_p1p0 > H || a_q1q0 > H;
        int f = hev ? iclamp(p1 - q1, lim_lo, lim_hi) : 0;
        f = iclamp(3 * (q0 - p0) + f, lim_lo, lim_hi);
        const int f1 = (f + 4 < lim_hi ? f + 4 : lim_hi) >> 3;
        const int f2 = (f + 3 < lim_hi ? f + 3 : lim_hi) >> 3;
        px[7] = clip_pixel(p0 + f2, bpp);
        px[8] = clip_pixel(q0 - f1, bpp);
        if (hev)
            return 0x0180;
        f = (f1 + 1) >> 1;
        px[6] = clip_pixel(p1 + f, bpp);
        px[9] = clip_pixel(q1 - f, bpp);
        return 0x03C0;
    }
#undef VP9_MAX
#undef VP9_ABS
}

/* byte-for-byte VP9Filter (vp9dec.h:83-87) */
struct LfSb {
    uint8_t level[64];
    uint8_t mask[2][2][8][4];
};

/* One edge operation: wd = 0 (none), 4, 8, 16; L = filter level */
struct LfOp {
    uint8_t wd, L;
};

/* Edge-op tables of one superblock and one plane.
 *   luma  : col[g][p]  g = 8-row group 0..7,  p = edge at x = 4p (p = 0..15)
 *           row[e][c]  e = edge at y = 4e (0..15), c = 8-column group 0..7
 *   chroma (4:2:0): col[g][p] g = 0..3 (8 chroma rows), p = edge at x = 4p (0..7)
 *           row[e][c]  e = edge at y = 4e (0..7), c = 8-column group 0..3
 * Unused entries have wd = 0. */
struct LfOps {
    LfOp col[8][16];
    LfOp row[16][8];
};

VP9_HD void lf_set(LfOp &o, int wd, int L) { o.wd = (uint8_t)wd; o.L = (uint8_t)L; }

/* One iteration `it` of filter_plane_cols' outer loop (vp9lpf.c:31-107): the column ops of the two
 * 8-line groups 2*it (o1) and 2*it+1 (o2), indexed by edge position / 4.
 * first_col: superblock is at the frame's left edge. */
VP9_HD void lf_decode_cols_it(const LfSb &f, bool chroma, bool first_col, int it, LfOp *o1, LfOp *o2)
{
    const int ss = chroma ? 1 : 0;
    const uint8_t (*mask)[4] = f.mask[ss][0];
    for (int p = 0; p < 16; p++) {
        lf_set(o1[p], 0, 0);
        lf_set(o2[p], 0, 0);
    }
    const int y = it * (2 << ss);
    if (y >= 8)
        return;
    const uint8_t *h1 = mask[y], *h2 = mask[y + 1 + ss];
    const uint8_t *lvl = f.level + it * (16 << ss);
    unsigned hm1 = h1[0] | h1[1] | h1[2], hm13 = h1[3];
    unsigned hm2 = h2[1] | h2[2], hm23 = h2[3];
    unsigned hm = hm1 | hm2 | hm13 | hm23;
    int loff = 0;
    for (int b = 0; b < 8 && (hm & ~((1u << b) - 1)); b++) {
        unsigned x = 1u << b;
        const uint8_t *l = lvl + loff;
        int pm = chroma ? b : 2 * b;     /* position index of the main edge */
        if (!first_col || b > 0) {
            if (hm1 & x) {
                int L = l[0];
                if (h1[0] & x) {
                    lf_set(o1[pm], 16, L);
                    if (h2[0] & x)
                        lf_set(o2[pm], 16, L);
                } else if (hm2 & x) {
                    lf_set(o1[pm], (h1[1] & x) ? 8 : 4, L);
                    lf_set(o2[pm], (h2[1] & x) ? 8 : 4, l[8 << ss]);
                } else {
                    lf_set(o1[pm], (h1[1] & x) ? 8 : 4, L);
                }
            } else if (hm2 & x) {
                lf_set(o2[pm], (h2[1] & x) ? 8 : 4, l[8 << ss]);
            }
        }
        if (chroma) {
            if (x & 0xAA)
                loff += 2;
        } else {
            if (hm13 & x)
                lf_set(o1[pm + 1], 4, l[0]);
            if (hm23 & x)
                lf_set(o2[pm + 1], 4, l[8]);
            loff++;
        }
    }
}

VP9_HD void lf_decode_cols(const LfSb &f, bool chroma, bool first_col, LfOps &ops)
{
    for (int it = 0; it < 4; it++)
        lf_decode_cols_it(f, chroma, first_col, it, ops.col[2 * it], ops.col[2 * it + 1]);
}

/* Iteration `j` of filter_plane_rows' inner loop for every y (vp9lpf.c:109-181): the row ops of
 * the two 8-column groups 2*j (o1) and 2*j+1 (o2), indexed by edge position / 4.
 * first_row: superblock is at the frame's top edge. */
VP9_HD void lf_decode_rows_j(const LfSb &f, bool chroma, bool first_row, int j, LfOp *o1, LfOp *o2)
{
    const int ss = chroma ? 1 : 0;
    const uint8_t (*mask)[4] = f.mask[ss][1];
    for (int e = 0; e < 16; e++) {
        lf_set(o1[e], 0, 0);
        lf_set(o2[e], 0, 0);
    }
    const int b = j * (2 << ss);
    if (b >= 8)
        return;
    const unsigned x = 1u << b, x2 = x << (1 + ss);
    for (int y = 0; y < 8; y++) {
        const uint8_t *vmask = mask[y];
        unsigned vm = vmask[0] | vmask[1] | vmask[2], vm3 = vmask[3];
        if (!(vm & ~(x - 1)))
            continue;                         /* the reference's x loop has ended before this pair */
        int em = chroma ? y : 2 * y;          /* edge index of the main edge (units of 4 rows) */
        const uint8_t *l = f.level + (chroma ? (y >> 1) * 16 : y * 8) + j * (2 << ss);
        if (!first_row || y) {
            if (vm & x) {
                int L = l[0];
                if (vmask[0] & x) {
                    lf_set(o1[em], 16, L);
                    if (vmask[0] & x2)
                        lf_set(o2[em], 16, L);
                } else if (vm & x2) {
                    lf_set(o1[em], (vmask[1] & x) ? 8 : 4, L);
                    lf_set(o2[em], (vmask[1] & x2) ? 8 : 4, l[1 + ss]);
                } else {
                    lf_set(o1[em], (vmask[1] & x) ? 8 : 4, L);
                }
            } else if (vm & x2) {
                lf_set(o2[em], (vmask[1] & x2) ? 8 : 4, l[1 + ss]);
            }
        }
        if (!chroma) {
            if (vm3 & x)
                lf_set(o1[em + 1], 4, l[0]);
            if (vm3 & x2)
                lf_set(o2[em + 1], 4, l[1]);
        }
    }
}

VP9_HD void lf_decode_rows(const LfSb &f, bool chroma, bool first_row, LfOps &ops)
{
    for (int j = 0; j < 4; j++) {
        LfOp o1[16], o2[16];
        lf_decode_rows_j(f, chroma, first_row, j, o1, o2);
        for (int e = 0; e < 16; e++) {
            ops.row[e][2 * j] = o1[e];
            ops.row[e][2 * j + 1] = o2[e];
        }
    }
}

/*
 * Dependency schedule of one superblock plane's edge ops.
 *
 * The reference applies the ops of one 8-line group strictly in increasing edge position
 * (vp9lpf.c:59-106 / :131-180).  Two ops of the same group commute exactly when the pixel
 * ranges they READ are disjoint (a filter never writes outside what it reads): read range of an
 * op at position X is [X-8, X+8) for wd 16 and [X-4, X+4) otherwise.  `step` of an op is the
 * length of the longest chain of non-commuting earlier ops in its group, so all ops of equal
 * step (any group) may run concurrently and steps run in increasing order.  Typical content
 * (uniform 8x8 / 16x16 / 32x32 transform areas) needs 1-3 steps per pass instead of 16.
 *
 * Column ops at position 0 (the superblock's left edge, reaching 8 pixels into the left
 * neighbour) always get step 0 and every other column op a step >= 1, so the wavefront kernel
 * can run "step 0 of superblock c+1" as the tail of superblock c's turn.
 *
 * Entry: bits 0-2 group (col: 8-row group, row: 8-column group), 3-6 edge index (position / 4),
 * 7-8 width code (1,2,3 = 4,8,16), 9-14 level.  Entries are sorted by step;
 * cstart/rstart[s] is the first entry of step s, [LF_NSTEP] the total.
 */
#define LF_NSTEP 17
struct LfSched {
    uint16_t col[128], row[128];
    uint8_t cstart[LF_NSTEP + 1], rstart[LF_NSTEP + 1];
    uint8_t cmax, rmax;          /* highest non-empty column step; number of row steps (device pre-pass) */
    uint8_t pad[26];
};

VP9_HD uint16_t lf_ent(int grp, int e, const LfOp &o)
{
    return (uint16_t)(grp | e << 3 | (o.wd == 16 ? 3 : o.wd == 8 ? 2 : 1) << 7 | (o.L & 63) << 9);
}

/* dir 0: column ops (o.col[g][p]); dir 1: row ops (o.row[e][c]) */
VP9_HD void lf_schedule_dir(const LfOps &o, int dir, uint16_t *ent, uint8_t *start)
{
    uint8_t step[8][16];
    uint8_t cnt[LF_NSTEP + 1];
    for (int s = 0; s <= LF_NSTEP; s++) cnt[s] = 0;
    for (int g = 0; g < 8; g++)
        for (int p = 0; p < 16; p++) {
            const LfOp &op = dir ? o.row[p][g] : o.col[g][p];
            step[g][p] = 0xff;
            if (!op.wd)
                continue;
            const int lo = 4 * p - (op.wd == 16 ? 8 : 4);
            int s = (dir == 0 && p > 0) ? 1 : 0;
            for (int q = p > 3 ? p - 3 : 0; q < p; q++) {
                const LfOp &pr = dir ? o.row[q][g] : o.col[g][q];
                if (pr.wd && 4 * q + (pr.wd == 16 ? 8 : 4) > lo && step[g][q] + 1 > s)
                    s = step[g][q] + 1;
            }
            step[g][p] = (uint8_t)s;
            cnt[s]++;
        }
    int acc = 0;
    for (int s = 0; s <= LF_NSTEP; s++) {
        int n = cnt[s];
        start[s] = (uint8_t)acc;
        cnt[s] = (uint8_t)acc;
        acc += n;
    }
    for (int g = 0; g < 8; g++)
        for (int p = 0; p < 16; p++)
            if (step[g][p] != 0xff)
                ent[cnt[step[g][p]]++] = lf_ent(g, p, dir ? o.row[p][g] : o.col[g][p]);
}

} /* namespace vp9 */
#endif /* VP9_LF_H */
