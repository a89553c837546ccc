/*
 * replay_model.cc -- TEST INFRASTRUCTURE (oracle/).  Serial CPU restatement of the VP9
 * reconstruction path, driven by the record stream (.v9r) that the libavcodec glue emits.
 *
 * It restates, one block at a time and in the reference's own order, what
 *   ff_vp9_inter_recon_*  (libavcodec/vp9recon.c:655-764, vp9_mc_template.c:30-464),
 *   ff_vp9_intra_recon_*  (libavcodec/vp9recon.c:235-364, check_intra_mode :37-221),
 *   ff_vp9_loopfilter_sb  (libavcodec/vp9lpf.c:183-230)
 * and the VP9DSPContext slots behind them (libavcodec/vp9dsp_template.c) compute, using the
 * per-pixel / per-line arithmetic of ffmpeg_hybrid_b200/csrc/vp9_{itx,pred,lf}.h compiled for
 * the host.  Its output is pinned against the UNMODIFIED reference decoder
 * (oracle/_ref/ffvp9_ref -raw, built by oracle/Makefile from /root/reference) frame by frame in
 * tests/test_oracle_model.py, and the arithmetic headers are pinned slot by slot against
 * oracle/_ref/libvp9dsp_ref.so in tests/test_oracle_slots.py.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may execute this.
 * The product (libvp9cuda.so) never links or calls it.
 *
 * usage: replay_model in.v9r out.yuv        (planar, visible crop, 16-bit LE when bpp > 8)
 *   also exported (libreplay_model.so): model_* slot entry points used by the slot tests.
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <vector>
#include <algorithm>
#include <memory>

#include "vp9cuda.h"
#include "vp9_itx.h"
#include "vp9_pred.h"
#include "vp9_lf.h"

using namespace vp9;

struct Surf {
    int w = 0, h = 0, bpp = 8;
    int pw[3], ph[3], stride[3];          /* visible plane size, stride in pixels */
    std::vector<uint16_t> p[3];
    void alloc(int w_, int h_, int bpp_, int ss = 1)     /* ss: chroma subsampling in both directions (4:2:0) or none (4:4:4) */
    {
        w = w_; h = h_; bpp = bpp_;
        int aw = (w + 63) & ~63, ah = (h + 63) & ~63;
        for (int i = 0; i < 3; i++) {
            int s = i ? ss : 0;
            pw[i] = (w + s) >> s; ph[i] = (h + s) >> s;
            stride[i] = aw >> s;
            p[i].assign((size_t)stride[i] * (ah >> s), 0);
        }
    }
};

static const uint8_t bwh8[13][2] = { {8,8},{8,4},{4,8},{4,4},{4,2},{2,4},{2,2},{2,1},{1,2},{1,1},{1,1},{1,1},{1,1} };

static int rdiv(int a, int b) { return a >= 0 ? (a + (b >> 1)) / b : (a - (b >> 1)) / b; }

/* ---- inverse transform + add (the itxfm_add slot) ------------------------------------------ */
template <bool HBD>
static void itx_add_t(uint16_t *dst, int stride, const void *coefs, int tx, int txtp, int eob, int bpp)
{
    typedef typename std::conditional<HBD, W64, W32>::type T;
    typedef typename std::conditional<HBD, int32_t, int16_t>::type C;
    const C *blk = (const C *)coefs;
    const bool lossless = txtp & 0x80;
    txtp &= 3;
    const int N = 4 << tx, bits = itx_bits(tx, lossless);
    if (!lossless && eob == 1 && (txtp == 0 || tx == 3)) {
        int32_t t = itx_dc_only<HBD>(blk[0]);
        int r = itx_round(t, bits);
        for (int y = 0; y < N; y++)
            for (int x = 0; x < N; x++)
                dst[y * stride + x] = clip_pixel(dst[y * stride + x] + r, bpp);
        return;
    }
    std::vector<C> tmp(N * N);
    for (int i = 0; i < N; i++) {
        if (lossless) {
            W32 in[4], out[4];
            for (int x = 0; x < 4; x++) in[x] = W32((int64_t)blk[i + x * 4]);
            iwht4(in, 1, out, 0);
            for (int x = 0; x < 4; x++) tmp[i * 4 + x] = (C)out[x].v;
        } else {
            T in[32], out[32];
            for (int x = 0; x < N; x++) in[x] = T((int64_t)blk[i + x * N]);
            bool adst = (txtp & 1) && tx < 3;
            switch (tx) {
            case 0: itx_1d<T, 4>(in, out, adst); break;
            case 1: itx_1d<T, 8>(in, out, adst); break;
            case 2: itx_1d<T, 16>(in, out, adst); break;
            default: itx_1d<T, 32>(in, out, false); break;
            }
            for (int x = 0; x < N; x++) tmp[i * N + x] = (C)out[x].v;
        }
    }
    for (int i = 0; i < N; i++) {
        C res[32];
        if (lossless) {
            W32 in[4], out[4];
            for (int x = 0; x < 4; x++) in[x] = W32((int64_t)tmp[i + x * 4]);
            iwht4(in, 1, out, 1);
            for (int x = 0; x < 4; x++) res[x] = (C)out[x].v;
        } else {
            T in[32], out[32];
            for (int x = 0; x < N; x++) in[x] = T((int64_t)tmp[i + x * N]);
            bool adst = (txtp & 2) && tx < 3;
            switch (tx) {
            case 0: itx_1d<T, 4>(in, out, adst); break;
            case 1: itx_1d<T, 8>(in, out, adst); break;
            case 2: itx_1d<T, 16>(in, out, adst); break;
            default: itx_1d<T, 32>(in, out, false); break;
            }
            for (int x = 0; x < N; x++) res[x] = (C)out[x].v;
        }
        for (int j = 0; j < N; j++)
            dst[j * stride + i] = clip_pixel(dst[j * stride + i] + itx_round(res[j], bits), bpp);
    }
}

static void itx_add(uint16_t *dst, int stride, const void *coefs, int tx, int txtp, int eob, int bpp)
{
    if (bpp > 8) itx_add_t<true>(dst, stride, coefs, tx, txtp, eob, bpp);
    else         itx_add_t<false>(dst, stride, coefs, tx, txtp, eob, bpp);
}

/* ---- intra prediction of one transform block ------------------------------------------------- */
struct PlanePx {
    const uint16_t *p; int stride;
    int operator()(int x, int y) const { return p[(size_t)y * stride + x]; }
};

static void intra_block(uint16_t *plane, int stride, int bx, int by, int tx, int mode, bool have_top,
                        bool have_left, bool have_right, int avail_w, int avail_h, int bpp)
{
    const int N = 4 << tx;
    int top[65], left[32];
    int m = intra_mode_conv(mode, have_left, have_top);
    PlanePx px = { plane, stride };
    intra_edges(px, bx, by, N, m, have_top, have_left, have_right, avail_w, avail_h, bpp, top, left);
    int dc = intra_dc(m, N, top, left, bpp);
    for (int y = 0; y < N; y++)
        for (int x = 0; x < N; x++)
            plane[(size_t)(by + y) * stride + bx + x] = intra_pixel(m, N, x, y, top, left, dc, bpp);
}

/* ---- loop filter of one superblock plane ------------------------------------------------------ */
static void lf_apply(uint16_t *plane, int stride, int x0, int y0, bool chroma, const LfOps &ops,
                     bool rows, const uint8_t *lim_lut, const uint8_t *mblim_lut, int bpp)
{
    const int n_lines = chroma ? 32 : 64, n_pos = chroma ? 8 : 16;
    for (int ln = 0; ln < n_lines; ln++) {
        for (int p = 0; p < n_pos; p++) {
            LfOp op = rows ? ops.row[p][ln >> 3] : ops.col[ln >> 3][p];
            if (!op.wd)
                continue;
            int px[16];
            int ex = rows ? x0 + ln : x0 + 4 * p, ey = rows ? y0 + 4 * p : y0 + ln;
            int sx = rows ? 0 : 1, sy = rows ? 1 : 0;
            int lo = op.wd == 16 ? 0 : 4, hi = op.wd == 16 ? 16 : 12;
            for (int i = lo; i < hi; i++)
                px[i] = plane[(size_t)(ey + (i - 8) * sy) * stride + ex + (i - 8) * sx];
            unsigned m = lf_line(px, op.wd, mblim_lut[op.L], lim_lut[op.L], op.L >> 4, bpp);
            for (int i = lo; i < hi; i++)
                if (m >> i & 1)
                    plane[(size_t)(ey + (i - 8) * sy) * stride + ex + (i - 8) * sx] = px[i];
        }
    }
}

/* ---- one frame ----------------------------------------------------------------------------------- */
struct Lane {
    Vp9CudaLaneCounts n;
    uint32_t coef_base;
    std::vector<Vp9CudaInterBlock> blk;
    std::vector<Vp9CudaTx> tx;
    std::vector<Vp9CudaIntraTx> itx;
    std::vector<uint8_t> coef;            /* Vp9CudaCoef16 / Vp9CudaCoef32 entries */
};

struct Model {
    std::shared_ptr<Surf> ref[8];
    std::shared_ptr<Surf> cur;
};

/* The coefficient stream carries the non-zero coefficients of a block with their positions in
 * the column-major N x N block (include/vp9cuda.h: Vp9CudaCoef16/32).  Rebuilds the dense block
 * the itxfm_add slot takes; checks what the producer promises (positions inside the block, values
 * non-zero, at most eob of them, an eob == 1 block holds the DC). */
static const void *coef_block(const std::vector<Lane> &lanes, uint32_t off, int ncoef, int eob, int tx, bool hbd,
                              std::vector<int32_t> &dense)
{
    const int nn = 16 << (2 * tx);
    dense.assign(nn, 0);
    if (ncoef > eob || eob > nn) { fprintf(stderr, "ncoef %d / eob %d out of range\n", ncoef, eob); exit(3); }
    if (!ncoef)
        return dense.data();
    for (const Lane &l : lanes) {
        if (off < l.coef_base || (uint64_t)off + ncoef > (uint64_t)l.coef_base + l.n.coef)
            continue;
        for (int i = 0; i < ncoef; i++) {
            int32_t val; uint32_t pos;
            if (hbd) { const Vp9CudaCoef32 &e = ((const Vp9CudaCoef32 *)l.coef.data())[off - l.coef_base + i]; val = e.val; pos = e.pos; }
            else     { const Vp9CudaCoef16 &e = ((const Vp9CudaCoef16 *)l.coef.data())[off - l.coef_base + i]; val = e.val; pos = e.pos; }
            if (pos >= (uint32_t)nn || !val || (eob == 1 && pos)) { fprintf(stderr, "bad coefficient entry pos %u val %d\n", pos, val); exit(3); }
            if (hbd) dense[pos] = val; else ((int16_t *)dense.data())[pos] = (int16_t)val;
        }
        return dense.data();
    }
    fprintf(stderr, "coef_off %u (+%d) outside every lane\n", off, ncoef);
    exit(3);
}

static void inter_block(Model &m, const Vp9CudaFrameHeader &h, const Vp9CudaInterBlock &b)
{
    Surf &c = *m.cur;
    const int cols = (h.w + 7) >> 3, rows = (h.h + 7) >> 3;
    const int sub8 = b.bs > 9;
    for (int pl = 0; pl < 3; pl++) {
        /* 4:4:4 chroma is predicted exactly like luma: full-size planes, luma vector units, every 4x4 of a sub-8x8
         * block with its own vector (vp9_mc_template.c:78-121 with ss_h == ss_v == 0) */
        const int ch = pl > 0 && h.ss_h;
        const int vis_w = (cols * 8) >> ch, vis_h = (rows * 8) >> ch;
        const int obx = (b.col * 8) >> ch, oby = (b.row * 8) >> ch;
        int ncall = (sub8 && !ch) ? 4 : 1;
        for (int k = 0; k < ncall; k++) {
            int cw, chh, px = 0, py = 0, pw, ph;
            if (sub8) {
                cw = chh = 4;
                if (!ch) { px = 4 * (k & 1); py = 4 * (k >> 1); pw = ph = 8; } else { pw = ph = 4; }
            } else {
                cw = (bwh8[b.bs][0] * 8) >> ch; chh = (bwh8[b.bs][1] * 8) >> ch;
                pw = cw; ph = chh;
            }
            McPos pos[2];
            RefPlane<uint16_t> rp[2];
            int nref = b.comp ? 2 : 1;
            for (int i = 0; i < nref; i++) {
                int mvx, mvy;
                if (sub8 && ch) {
                    mvx = rdiv(b.mv[0][i][0] + b.mv[1][i][0] + b.mv[2][i][0] + b.mv[3][i][0], 4);
                    mvy = rdiv(b.mv[0][i][1] + b.mv[1][i][1] + b.mv[2][i][1] + b.mv[3][i][1], 4);
                } else {
                    int kk = sub8 ? k : 0;
                    mvx = b.mv[kk][i][0]; mvy = b.mv[kk][i][1];
                }
                Surf *r = m.ref[h.ref_slot[b.ref[i]]].get();
                if (!r) { fprintf(stderr, "missing reference\n"); exit(3); }
                rp[i].pix = r->p[pl].data(); rp[i].stride = r->stride[pl];
                rp[i].w = r->pw[pl]; rp[i].h = r->ph[pl];
                int sc = h.mvscale[b.ref[i]][0];
                if (sc)
                    pos[i] = mc_pos_scaled(obx + px, oby + py, mvx, mvy, ch, px, py, pw, ph, cols, rows,
                                           h.mvscale[b.ref[i]][0], h.mvscale[b.ref[i]][1],
                                           h.mvstep[b.ref[i]][0], h.mvstep[b.ref[i]][1]);
                else
                    pos[i] = mc_pos_unscaled(obx + px, oby + py, mvx, mvy, ch);
            }
            for (int oy = 0; oy < chh; oy++) {
                for (int ox = 0; ox < cw; ox++) {
                    int X = obx + px + ox, Y = oby + py + oy;
                    if (X >= vis_w || Y >= vis_h)
                        continue;
                    int v = mc_pixel(rp[0], pos[0], ox, oy, b.filter, h.bpp);
                    if (b.comp)
                        v = (v + mc_pixel(rp[1], pos[1], ox, oy, b.filter, h.bpp) + 1) >> 1;
                    c.p[pl][(size_t)Y * c.stride[pl] + X] = v;
                }
            }
        }
    }
}

static void decode_frame(Model &m, const Vp9CudaFrameHeader &h, const std::vector<Lane> &lanes,
                         const std::vector<Vp9CudaSbIndex> &sbi, const std::vector<Vp9CudaLoopFilterSb> &lf)
{
    const bool hbd = h.bpp > 8;
    std::vector<int32_t> dense;
    const int cols = (h.w + 7) >> 3, rows = (h.h + 7) >> 3;
    const int sb_cols = (h.w + 63) >> 6, sb_rows = (h.h + 63) >> 6;
    m.cur = std::make_shared<Surf>();
    m.cur->alloc(h.w, h.h, h.bpp, h.ss_h);
    Surf &c = *m.cur;
    if (h.ss_h != h.ss_v) { fprintf(stderr, "4:2:2 / 4:4:0 are not restated\n"); exit(3); }
    /* phase 1: inter prediction; phase 2: inter residual */
    for (const Lane &l : lanes)
        for (const auto &b : l.blk)
            inter_block(m, h, b);
    for (const Lane &l : lanes)
        for (const auto &r : l.tx)
            itx_add(&c.p[r.plane][(size_t)r.y4 * 4 * c.stride[r.plane] + r.x4 * 4], c.stride[r.plane],
                    coef_block(lanes, r.coef_off, r.ncoef, r.eob, r.tx, hbd, dense), r.tx, r.txtp, r.eob, h.bpp);
    /* the per-superblock index must name exactly the lane's residual blocks that lie inside it */
    {
        std::vector<size_t> seen(lanes.size(), 0);
        for (int s = 0; s < sb_rows * sb_cols; s++) {
            if (!sbi[s].tx_count) continue;
            const Lane &l = lanes[sbi[s].lane];
            for (unsigned k = 0; k < sbi[s].tx_count; k++) {
                const Vp9CudaTx &r = l.tx[sbi[s].tx_start + k];
                const int ch = r.plane > 0 && h.ss_h, sx = (r.x4 * 4) >> (ch ? 5 : 6), sy = (r.y4 * 4) >> (ch ? 5 : 6);
                if (sy * sb_cols + sx != s) { fprintf(stderr, "superblock %d lists a residual block of superblock %d\n", s, sy * sb_cols + sx); exit(4); }
            }
            seen[sbi[s].lane] += sbi[s].tx_count;
        }
        for (size_t i = 0; i < lanes.size(); i++)
            if (seen[i] != lanes[i].tx.size()) { fprintf(stderr, "lane %zu: %zu residual blocks indexed, %zu recorded\n", i, seen[i], lanes[i].tx.size()); exit(4); }
    }
    /* phase 3: intra blocks, superblocks in raster order, coding order inside.  Along the way
     * the records' dependency levels are re-derived and checked (exactly 1 + max over the intra
     * neighbours the block's predictor reads, by mode), which is what the GPU's level-ordered dispatch relies on. */
    std::vector<int> lvmap[3];
    int lvw[3], lvh[3];
    for (int p = 0; p < 3; p++) {
        lvw[p] = (cols * 2) >> (p && h.ss_h); lvh[p] = (rows * 2) >> (p && h.ss_v);
        lvmap[p].assign((size_t)lvw[p] * lvh[p], -1);
    }
    for (int s = 0; s < sb_rows * sb_cols; s++) {
        const Lane &l = lanes[sbi[s].lane];
        for (unsigned k = 0; k < sbi[s].itx_count; k++) {
            const Vp9CudaIntraTx &r = l.itx[sbi[s].itx_start + k];
            int pl = r.plane, ch = pl > 0 && h.ss_h;
            int bx = r.x4 * 4, by = r.y4 * 4;
            int avail_w = ((cols * 8) >> ch) - bx, avail_h = ((rows * 8) >> ch) - by;
            {
                const int n = 1 << r.tx, x = r.x4, y = r.y4, w = lvw[pl], hh = lvh[pl];
                const bool hl = r.flags & VP9CUDA_ITX_HAVE_LEFT, ht = y > 0;
                const int needs = intra_mode_needs(intra_mode_conv(r.mode, hl, ht));      /* edges this block's predictor reads */
                int lv = 0;
                if (hl && (needs & 1))
                    for (int j = y; j < std::min(y + n, hh); j++)
                        lv = std::max(lv, lvmap[pl][(size_t)j * w + x - 1] + 1);
                if (hl && ht && (needs & 4))
                    lv = std::max(lv, lvmap[pl][(size_t)(y - 1) * w + x - 1] + 1);
                if (ht && (needs & 2))
                    for (int i = x; i < std::min(x + ((r.tx == 0 && (needs & 8) && (r.flags & VP9CUDA_ITX_HAVE_RIGHT)) ? 2 : n), w); i++)
                        lv = std::max(lv, lvmap[pl][(size_t)(y - 1) * w + i] + 1);
                if (lv != r.level) {
                    fprintf(stderr, "level mismatch plane %d cell %d,%d: record %d, derived %d\n", pl, x, y, r.level, lv);
                    exit(4);
                }
                for (int j = y; j < std::min(y + n, hh); j++)
                    for (int i = x; i < std::min(x + n, w); i++)
                        lvmap[pl][(size_t)j * w + i] = lv;
            }
            intra_block(c.p[pl].data(), c.stride[pl], bx, by, r.tx, r.mode, by > 0,
                        r.flags & VP9CUDA_ITX_HAVE_LEFT, r.flags & VP9CUDA_ITX_HAVE_RIGHT,
                        avail_w, avail_h, h.bpp);
            if (r.eob)
                itx_add(&c.p[pl][(size_t)by * c.stride[pl] + bx], c.stride[pl],
                        coef_block(lanes, r.coef_off, r.ncoef, r.eob, r.tx, hbd, dense), r.tx, r.txtp, r.eob, h.bpp);
        }
    }
    /* phase 4: loop filter, superblocks in raster order */
    if (h.lf_level) {
        LfOps ops;
        for (int sr = 0; sr < sb_rows; sr++)
            for (int sc = 0; sc < sb_cols; sc++) {
                const LfSb &f = *(const LfSb *)&lf[sr * sb_cols + sc];
                for (int pl = 0; pl < 3; pl++) {
                    int ch = pl > 0 && h.ss_h;       /* 4:4:4: the chroma planes take the luma masks (vp9lpf.c:214) */
                    lf_decode_cols(f, ch, sc == 0, ops);
                    lf_apply(c.p[pl].data(), c.stride[pl], (sc * 64) >> ch, (sr * 64) >> ch, ch, ops, false,
                             h.lim_lut, h.mblim_lut, h.bpp);
                    lf_decode_rows(f, ch, sr == 0, ops);
                    lf_apply(c.p[pl].data(), c.stride[pl], (sc * 64) >> ch, (sr * 64) >> ch, ch, ops, true,
                             h.lim_lut, h.mblim_lut, h.bpp);
                }
            }
    }
    for (int i = 0; i < 8; i++)
        if (h.refreshrefmask >> i & 1)
            m.ref[i] = m.cur;
}

static void rd(FILE *fp, void *p, size_t n)
{
    if (n && fread(p, 1, n, fp) != n) { fprintf(stderr, "truncated .v9r\n"); exit(2); }
}

#ifndef MODEL_NO_MAIN
int main(int argc, char **argv)
{
    if (argc < 3) { fprintf(stderr, "usage: %s in.v9r out.yuv\n", argv[0]); return 2; }
    FILE *fp = fopen(argv[1], "rb"), *out = fopen(argv[2], "wb");
    if (!fp || !out) { perror("open"); return 2; }
    char magic[4]; uint32_t sz[8];
    rd(fp, magic, 4); rd(fp, sz, 32);
    if (memcmp(magic, "V9R3", 4) || sz[0] != sizeof(Vp9CudaFrameHeader) || sz[2] != sizeof(Vp9CudaInterBlock) ||
        sz[3] != sizeof(Vp9CudaTx) || sz[4] != sizeof(Vp9CudaIntraTx) || sz[7] != VP9CUDA_ABI_VERSION) {
        fprintf(stderr, "bad .v9r header\n"); return 2;
    }
    Model m;
    int nf = 0;
    while (fread(magic, 1, 4, fp) == 4) {
        if (memcmp(magic, "FRAM", 4)) { fprintf(stderr, "bad frame tag\n"); return 2; }
        Vp9CudaFrameHeader h; uint32_t n_sb;
        rd(fp, &h, sizeof(h)); rd(fp, &n_sb, 4);
        const int ceb = h.bpp > 8 ? sizeof(Vp9CudaCoef32) : sizeof(Vp9CudaCoef16);
        std::vector<Lane> lanes(h.n_lanes);
        for (Lane &l : lanes) {
            rd(fp, &l.n, sizeof(l.n)); rd(fp, &l.coef_base, 4);
            l.blk.resize(l.n.blk); rd(fp, l.blk.data(), l.n.blk * sizeof(Vp9CudaInterBlock));
            l.tx.resize(l.n.tx); rd(fp, l.tx.data(), l.n.tx * sizeof(Vp9CudaTx));
            l.itx.resize(l.n.itx); rd(fp, l.itx.data(), l.n.itx * sizeof(Vp9CudaIntraTx));
            l.coef.resize((size_t)l.n.coef * ceb); rd(fp, l.coef.data(), l.coef.size());
        }
        std::vector<Vp9CudaSbIndex> sbi(n_sb); rd(fp, sbi.data(), n_sb * sizeof(Vp9CudaSbIndex));
        std::vector<Vp9CudaLoopFilterSb> lf(n_sb); rd(fp, lf.data(), n_sb * sizeof(Vp9CudaLoopFilterSb));
        decode_frame(m, h, lanes, sbi, lf);
        /* like the decoder, only shown frames are output (vp9.c:1853-1857) */
        Surf &c = *m.cur;
        for (int pl = 0; pl < 3 && !h.invisible; pl++)
            for (int y = 0; y < c.ph[pl]; y++)
                for (int x = 0; x < c.pw[pl]; x++) {
                    uint16_t v = c.p[pl][(size_t)y * c.stride[pl] + x];
                    if (h.bpp > 8) fwrite(&v, 2, 1, out); else fputc(v, out);
                }
        nf++;
    }
    fprintf(stderr, "model: %d frames\n", nf);
    fclose(out);
    return 0;
}
#endif

/* ------------------------------------------------------------------------------------------
 * Slot-level entry points of the restatement (libreplay_model.so), same calling convention as
 * oracle/slot_shim.c's ref_* functions so the tests can memcmp them against the real slots.
 * Pixels are uint8_t (bpp 8) or uint16_t; strides in BYTES like the reference slots.
 * ---------------------------------------------------------------------------------------- */
template <class P> static void load_px(std::vector<uint16_t> &v, const uint8_t *src, ptrdiff_t stride, int w, int h)
{
    v.resize((size_t)w * h);
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++)
            v[(size_t)y * w + x] = ((const P *)(src + y * stride))[x];
}
template <class P> static void store_px(const std::vector<uint16_t> &v, uint8_t *dst, ptrdiff_t stride, int w, int h)
{
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++)
            ((P *)(dst + y * stride))[x] = (P)v[(size_t)y * w + x];
}

extern "C" {

void model_itxfm_add(int bpp, int tx, int txtp, uint8_t *dst, ptrdiff_t stride, void *block, int eob)
{
    const int lossless = tx == 4, t = lossless ? 0 : tx, N = 4 << t;
    std::vector<uint16_t> v;
    if (bpp > 8) load_px<uint16_t>(v, dst, stride, N, N); else load_px<uint8_t>(v, dst, stride, N, N);
    itx_add(v.data(), N, block, t, (txtp & 3) | (lossless ? 0x80 : 0), eob, bpp);
    if (bpp > 8) store_px<uint16_t>(v, dst, stride, N, N); else store_px<uint8_t>(v, dst, stride, N, N);
    memset(block, 0, (size_t)N * N * (bpp > 8 ? 4 : 2));     /* the slot zeroes its input */
}

/* left: bottom-to-top (top-to-bottom for HOR_UP), top[-1] = top-left: the slot's convention */
void model_intra_pred(int bpp, int tx, int mode, uint8_t *dst, ptrdiff_t stride, const uint8_t *left, const uint8_t *top)
{
    const int N = 4 << tx;
    int T[65], L[32];
    auto rd = [&](const uint8_t *p, int i) -> int { return bpp > 8 ? ((const uint16_t *)p)[i] : p[i]; };
    T[0] = rd(top, -1);
    for (int k = 0; k < 2 * N; k++)
        T[1 + k] = rd(top, N == 4 ? k : (k < N ? k : N - 1));
    for (int y = 0; y < N; y++)
        L[y] = mode == HOR_UP_PRED ? rd(left, y) : rd(left, N - 1 - y);
    int dc = intra_dc(mode, N, T, L, bpp);
    for (int y = 0; y < N; y++)
        for (int x = 0; x < N; x++) {
            int v = intra_pixel(mode, N, x, y, T, L, dc, bpp);
            if (bpp > 8) ((uint16_t *)(dst + y * stride))[x] = v; else dst[y * stride + x] = v;
        }
}

/* mc / smc: ref points at the block's integer position inside a padded buffer, like the slots */
void model_mc(int bpp, int w, int filter, int avg, uint8_t *dst, ptrdiff_t dst_stride, const uint8_t *ref,
              ptrdiff_t ref_stride, int h, int mx, int my, int dx, int dy)
{
    const int bps = bpp > 8 ? 2 : 1;
    McPos pos = { 0, 0, mx, my, dx ? dx : 16, dy ? dy : 16 };
    for (int oy = 0; oy < h; oy++)
        for (int ox = 0; ox < w; ox++) {
            /* unclamped plane: reads reach up to 3 left/above and (step-scaled) beyond, all inside
             * the caller's padding, so give RefPlane a window that never clamps */
            int v;
            if (bpp > 8) {
                RefPlane<uint16_t> rp = { (const uint16_t *)ref - 64 * (ref_stride / bps) - 64, (int)(ref_stride / bps), 1 << 20, 1 << 20 };
                McPos p2 = pos; p2.x0 = 64; p2.y0 = 64;
                v = mc_pixel(rp, p2, ox, oy, filter, bpp);
                uint16_t *d = (uint16_t *)(dst + oy * dst_stride) + ox;
                *d = avg ? (*d + v + 1) >> 1 : v;
            } else {
                RefPlane<uint8_t> rp = { ref - 64 * ref_stride - 64, (int)ref_stride, 1 << 20, 1 << 20 };
                McPos p2 = pos; p2.x0 = 64; p2.y0 = 64;
                v = mc_pixel(rp, p2, ox, oy, filter, bpp);
                uint8_t *d = dst + oy * dst_stride + ox;
                *d = avg ? (*d + v + 1) >> 1 : v;
            }
        }
}

/* loop filter over `len` lines (8 or 16); dir 0 = column edge ("h" slots), 1 = row edge ("v") */
void model_loop_filter(int bpp, int wd, int dir, int len, uint8_t *dst, ptrdiff_t stride, int E, int I, int H)
{
    const int bps = bpp > 8 ? 2 : 1;
    for (int ln = 0; ln < len; ln++) {
        int px[16];
        auto at = [&](int k) -> uint8_t * {
            return dir ? dst + (ptrdiff_t)(k - 8) * stride + ln * bps : dst + (ptrdiff_t)ln * stride + (k - 8) * bps;
        };
        for (int k = 0; k < 16; k++)
            px[k] = bpp > 8 ? *(uint16_t *)at(k) : *at(k);
        unsigned m = lf_line(px, wd, E, I, H, bpp);
        for (int k = 0; k < 16; k++)
            if (m >> k & 1) { if (bpp > 8) *(uint16_t *)at(k) = px[k]; else *at(k) = px[k]; }
    }
}

/* mask walk of one superblock plane: fills wd/L tables (col[8][16], row[16][8]) */
void model_lf_decode(const uint8_t *vp9filter /* 192 bytes */, int chroma, int first_col, int first_row,
                     uint8_t *col_wd, uint8_t *col_L, uint8_t *row_wd, uint8_t *row_L)
{
    LfOps ops;
    lf_decode_cols(*(const LfSb *)vp9filter, chroma, first_col, ops);
    lf_decode_rows(*(const LfSb *)vp9filter, chroma, first_row, ops);
    for (int g = 0; g < 8; g++)
        for (int p = 0; p < 16; p++) { col_wd[g * 16 + p] = ops.col[g][p].wd; col_L[g * 16 + p] = ops.col[g][p].L; }
    for (int e = 0; e < 16; e++)
        for (int c = 0; c < 8; c++) { row_wd[e * 8 + c] = ops.row[e][c].wd; row_L[e * 8 + c] = ops.row[e][c].L; }
}

/* Test hook for the dependency schedule (vp9_lf.h: lf_schedule_dir): filters one superblock plane
 * of seeded pixels (a) in the reference's sequential edge order and (b) step by step in schedule
 * order, within a step in REVERSE list order; returns the number of differing pixels, or -1 if the
 * schedule is malformed (entry count, step-0 rule for column position 0). */
int model_lf_sched_check(const uint8_t *vp9filter, int chroma, int first_col, int first_row, unsigned seed,
                         int bpp, int *n_steps_out)
{
    LfSb sb = *(const LfSb *)vp9filter;
    for (int i = 0; i < 64; i++) sb.level[i] &= 63;
    LfOps ops;
    lf_decode_cols(sb, chroma, first_col, ops);
    lf_decode_rows(sb, chroma, first_row, ops);
    LfSched sc;
    lf_schedule_dir(ops, 0, sc.col, sc.cstart);
    lf_schedule_dir(ops, 1, sc.row, sc.rstart);
    uint8_t lim[64], mblim[64];
    for (int L = 0; L < 64; L++) {          /* sharpness 0 luts, vp9.c:673-686 */
        int l = L < 1 ? 1 : L;
        lim[L] = l; mblim[L] = 2 * (L + 2) + l;
    }
    const int W = 96, S = chroma ? 32 : 64;
    std::vector<uint16_t> a(W * W), b;
    unsigned st = seed * 2654435761u + 12345u;
    int v = 1 << (bpp - 1);
    for (int i = 0; i < W * W; i++) {
        st = st * 1664525u + 1013904223u;
        int r = st >> 24;
        v += (r & 0x80) ? ((r & 0x7f) % 3) - 1 : ((r & 0x7f) % 41) - 20;   /* flat runs + jumps */
        if (v < 0) v = 0;
        if (v > (1 << bpp) - 1) v = (1 << bpp) - 1;
        a[i] = (uint16_t)v;
    }
    b = a;
    lf_apply(a.data(), W, 16, 16, chroma, ops, false, lim, mblim, bpp);
    lf_apply(a.data(), W, 16, 16, chroma, ops, true, lim, mblim, bpp);
    int n_ops[2] = { 0, 0 };
    for (int g = 0; g < 8; g++)
        for (int p = 0; p < 16; p++) { n_ops[0] += ops.col[g][p].wd != 0; n_ops[1] += ops.row[p][g].wd != 0; }
    if (sc.cstart[LF_NSTEP] != n_ops[0] || sc.rstart[LF_NSTEP] != n_ops[1])
        return -1;
    int steps = 0;
    for (int dir = 0; dir < 2; dir++) {
        const uint16_t *ent = dir ? sc.row : sc.col;
        const uint8_t *start = dir ? sc.rstart : sc.cstart;
        for (int s = 0; s < LF_NSTEP; s++) {
            if (start[s + 1] > start[s]) steps++;
            for (int k = start[s + 1] - 1; k >= start[s]; k--) {
                int e = ent[k], g = e & 7, p = e >> 3 & 15, wd = 2 << (e >> 7 & 3), L = e >> 9;
                const LfOp &ref = dir ? ops.row[p][g] : ops.col[g][p];
                if (ref.wd != wd || ref.L != L) return -1;
                if (!dir && ((p == 0) != (s == 0))) return -1;
                for (int ln = 0; ln < 8; ln++) {
                    int ex = dir ? 16 + 8 * g + ln : 16 + 4 * p, ey = dir ? 16 + 4 * p : 16 + 8 * g + ln;
                    int sx = dir ? 0 : 1, sy = dir ? 1 : 0, px[16];
                    for (int i = 0; i < 16; i++) px[i] = b[(ey + (i - 8) * sy) * W + ex + (i - 8) * sx];
                    unsigned m = lf_line(px, wd, mblim[L], lim[L], L >> 4, bpp);
                    for (int i = 0; i < 16; i++)
                        if (m >> i & 1) b[(ey + (i - 8) * sy) * W + ex + (i - 8) * sx] = px[i];
                }
            }
        }
    }
    (void)S;
    if (n_steps_out) *n_steps_out = steps;
    int diff = 0;
    for (int i = 0; i < W * W; i++) diff += a[i] != b[i];
    return diff;
}

} /* extern "C" */
