/*
 * emu_inter.cc -- TEST INFRASTRUCTURE (oracle/).  Runs the CUDA kernel k_inter_sb
 * (ffmpeg_hybrid_b200/csrc/k_inter_sb.cuh: shared-memory staged inter prediction + fused inter
 * residual) ON THE HOST through oracle/cuda_emu.h -- same source, every CUDA thread an OS thread --
 * and compares the pixels of every inter cell of every frame with the oracle restatement
 * (replay_model.cc phases 1 and 2: inter_block + itx_add, which are pinned against the unmodified
 * reference decoder).  This checks the kernel's indexing, staging, batching and synchronisation
 * before it ever runs on a GPU; the GPU parity tests then check the real thing.
 *
 * usage: emu_inter [-full] in.v9r [max_frames]      exit 0 = identical
 *   -full: also k_intra_prep, k_intra_res, k_lf_sched and the intra / loop-filter wavefronts (k_wave), CTAs in
 *          block-index order; the whole reconstructed frame is compared with the restatement
 */
#define VP9_EMU_IMPL
#include "cuda_emu.h"
#define MODEL_NO_MAIN
#include "replay_model.cc"
#include "k_inter_sb.cuh"
#include "k_intra.cuh"
#include "k_lf.cuh"

static bool g_full;          /* -full: also the pre-passes and the intra / loop-filter wavefronts, whole frame compared */

static long g_compared;
template <class P, bool HBD>
static long run_wave(FrameDesc &f, const Vp9CudaFrameHeader &h, const std::vector<Lane> &lanes,
                     const std::vector<Vp9CudaSbIndex> &sbi, const std::vector<Vp9CudaLoopFilterSb> &lf,
                     std::vector<P> *cur, Model &m)
{
    /* the rest of launch_frame (csrc/vp9cuda.cu): pre-passes, then the wavefronts.  CTAs run one after the
     * other in block-index order, which satisfies every cross-CTA wait of k_wave (they all point at lower
     * block indices) */
    const int n_sb = f.sb_cols * f.sb_rows;
    std::vector<Vp9CudaIntraTx> itx;
    for (size_t l = 0; l < lanes.size(); l++) {
        f.lane_itx_base[l] = (uint32_t)itx.size();
        itx.insert(itx.end(), lanes[l].itx.begin(), lanes[l].itx.end());
    }
    f.itx = itx.data(); f.n_itx = (uint32_t)itx.size();
    f.lf = lf.data();
    f.lossless = h.lossless; f.lf_level = h.lf_level;
    memcpy(f.lim_lut, h.lim_lut, 64); memcpy(f.mblim_lut, h.mblim_lut, 64);
    std::vector<IntraSb> isb(n_sb);
    std::vector<IntraRec> irec(itx.size() + 1);
    std::vector<IntraBlk> iblk(itx.size() + 1);
    std::vector<int> iprogress(2 * f.sb_rows, 0), progress(6 * f.sb_rows, 0);
    std::vector<uint16_t> itab((size_t)n_sb * INTRA_SB_TAB);
    std::vector<int16_t> ires((size_t)n_sb * INTRA_SB_RES);
    std::vector<LfSched> lfs((size_t)n_sb * 3);
    f.isb = isb.data(); f.irec = irec.data(); f.iblk = iblk.data(); f.iprogress = iprogress.data();
    f.itab = itab.data(); f.ires = ires.data(); f.lfsched = lfs.data(); f.progress = progress.data();
    std::vector<unsigned long long> ll((size_t)3 * n_sb * LF_LL_WORDS, 0);
    static unsigned ll_tag = 0;
    std::vector<unsigned long long> ill((size_t)n_sb * INTRA_LL_WORDS, 0);
    f.lf_ll = ll.data(); f.intra_ll = ill.data(); f.ll_tag = ++ll_tag;
    const bool fused = f.n_itx && f.lf_level && !getenv("EMU_NOFUSE");
    f.lf_follow = fused ? 3 : 0;
    if (f.n_itx) {
        emu::launch(dim3(n_sb), 128, 0, [&] { k_intra_prep(f); });
        emu::launch(dim3(n_sb), IRES_WARPS * 32, 0, [&] { k_intra_res<HBD>(f); });
    }
    if (getenv("EMU_NOLF")) f.lf_level = 0;
    if (f.lf_level)
        emu::launch(dim3(n_sb), LF_SCHED_THREADS, 0, [&] { k_lf_sched(f); });
    if (fused) {
        emu::launch(dim3((f.mono >= 0 ? 2 : 4) * f.sb_rows), INTRA_THREADS, WaveSmem<P>::BYTES, [&] { k_wave<P>(f); });
    } else {
        if (f.n_itx)
            emu::launch(dim3(f.sb_rows), INTRA_THREADS, IntraSmem<P>::BYTES, [&] { k_intra<P>(f); });
        if (f.lf_level)
            emu::launch(dim3((f.mono >= 0 ? 1 : 3) * f.sb_rows), LF_THREADS, 0, [&] { k_lf<P>(f); });
    }
    (void)cur; (void)m;
    return 0;
}

/* the whole picture against the restatement's (all four phases) */
template <class P>
static long compare_full(const FrameDesc &f, const Vp9CudaFrameHeader &h, const std::vector<Lane> &lanes,
                         const std::vector<Vp9CudaSbIndex> &sbi, const std::vector<Vp9CudaLoopFilterSb> &lf,
                         std::vector<P> *cur, Model &m)
{
    Model mf = m;
    Vp9CudaFrameHeader h2 = h;
    if (getenv("EMU_NOLF")) h2.lf_level = 0;
    decode_frame(mf, h2, lanes, sbi, lf);
    long bad = 0;
    for (int pl = 0; pl < 3; pl++)
        for (int y = 0; y < mf.cur->ph[pl]; y++)
            for (int x = 0; x < mf.cur->pw[pl]; x++) {
                g_compared++;
                const int want = mf.cur->p[pl][(size_t)y * mf.cur->stride[pl] + x], got = cur[pl][(size_t)y * f.cur.stride[pl] + x];
                if (want != got && bad++ < 12)
                    fprintf(stderr, "  full frame, plane %d (%d, %d): kernels %d, restatement %d\n", pl, x, y, got, want);
            }
    return bad;
}

template <class P, bool HBD>
static long run_frame(Model &m, const Vp9CudaFrameHeader &h, const std::vector<Lane> &lanes,
                      const std::vector<Vp9CudaSbIndex> &sbi, const std::vector<Vp9CudaLoopFilterSb> &lf)
{
    const int cols = (h.w + 7) >> 3, rows = (h.h + 7) >> 3;
    const int sb_cols = (h.w + 63) >> 6, sb_rows = (h.h + 63) >> 6;
    /* expected: phases 1 and 2 of the restatement on a fresh picture */
    Model mx = m;
    mx.cur = std::make_shared<Surf>();
    mx.cur->alloc(h.w, h.h, h.bpp, h.ss_h);
    std::vector<int32_t> dense;
    for (const Lane &l : lanes)
        for (const auto &b : l.blk)
            inter_block(mx, h, b);
    for (const Lane &l : lanes)
        for (const auto &r : l.tx)
            itx_add(&mx.cur->p[r.plane][(size_t)r.y4 * 4 * mx.cur->stride[r.plane] + r.x4 * 4], mx.cur->stride[r.plane],
                    coef_block(lanes, r.coef_off, r.ncoef, r.eob, r.tx, HBD, dense), r.tx, r.txtp, r.eob, h.bpp);
    /* the frame as the kernel sees it */
    FrameDesc f;
    memset(&f, 0, sizeof(f));
    std::vector<P> cur[3], refp[3][3];
    for (int pl = 0; pl < 3; pl++) {
        cur[pl].assign(mx.cur->p[pl].size(), (P)0);
        f.cur.pl[pl] = cur[pl].data(); f.cur.stride[pl] = mx.cur->stride[pl];
    }
    f.cur.w = h.w; f.cur.h = h.h;
    for (int i = 0; i < 3; i++) {
        Surf *r = m.ref[h.ref_slot[i] & 7].get();
        if (!r) continue;
        for (int pl = 0; pl < 3; pl++) {
            refp[i][pl].assign(r->p[pl].begin(), r->p[pl].end());
            f.ref[i].pl[pl] = refp[i][pl].data(); f.ref[i].stride[pl] = r->stride[pl];
        }
        f.ref[i].w = r->w; f.ref[i].h = r->h;
        f.mvscale[i][0] = h.mvscale[i][0]; f.mvscale[i][1] = h.mvscale[i][1];
        f.mvstep[i][0] = h.mvstep[i][0]; f.mvstep[i][1] = h.mvstep[i][1];
    }
    f.w = h.w; f.h = h.h; f.cols = cols; f.rows = rows; f.sb_cols = sb_cols; f.sb_rows = sb_rows; f.bpp = h.bpp;
    std::vector<Vp9CudaInterBlock> blk;
    std::vector<Vp9CudaTx> tx;
    typedef typename ItxTypes<HBD>::E E;
    std::vector<E> coef;
    f.n_lanes = h.n_lanes;
    for (size_t l = 0; l < lanes.size(); l++) {
        f.lane_blk_base[l] = (uint32_t)blk.size(); f.lane_blk_cnt[l] = lanes[l].n.blk;
        f.lane_tx_base[l] = (uint32_t)tx.size(); f.lane_tx_cnt[l] = lanes[l].n.tx;
        blk.insert(blk.end(), lanes[l].blk.begin(), lanes[l].blk.end());
        tx.insert(tx.end(), lanes[l].tx.begin(), lanes[l].tx.end());
        if (coef.size() < (size_t)lanes[l].coef_base + lanes[l].n.coef)
            coef.resize((size_t)lanes[l].coef_base + lanes[l].n.coef);
        if (lanes[l].n.coef)
            memcpy(coef.data() + lanes[l].coef_base, lanes[l].coef.data(), lanes[l].n.coef * sizeof(E));
    }
    f.blk = blk.data(); f.n_blk = (uint32_t)blk.size();
    f.tx = tx.data(); f.n_tx = (uint32_t)tx.size();
    f.coef = coef.data();
    f.sb_index = sbi.data();
    std::vector<int32_t> cellmap((size_t)cols * rows, -1);
    for (size_t i = 0; i < blk.size(); i++) {         /* what k_cellmap writes */
        const int w = std::min<int>(bwh8[blk[i].bs][0], cols - blk[i].col), hh = std::min<int>(bwh8[blk[i].bs][1], rows - blk[i].row);
        for (int y = 0; y < hh; y++)
            for (int x = 0; x < w; x++)
                cellmap[(size_t)(blk[i].row + y) * cols + blk[i].col + x] = (int32_t)i;
    }
    f.cellmap = cellmap.data();
    if (blk.empty() && !g_full)
        return 0;
    long bad = 0;
    /* 4:2:0: one pass over all planes.  4:4:4: three passes, each with one plane in the place of plane 0
     * (FrameDesc.mono), exactly what csrc/vp9cuda.cu: reconstruct() launches */
    for (int pass = 0; pass < (h.ss_h ? 1 : 3); pass++) {
        FrameDesc fp = f;
        fp.mono = h.ss_h ? -1 : pass;
        if (!h.ss_h) {
            fp.cur.pl[0] = f.cur.pl[pass]; fp.cur.stride[0] = f.cur.stride[pass];
            for (int i = 0; i < 3; i++) { fp.ref[i].pl[0] = f.ref[i].pl[pass]; fp.ref[i].stride[0] = f.ref[i].stride[pass]; }
        }
        if (!blk.empty())
            emu::launch(dim3(sb_cols, sb_rows), ISB_THREADS, IsbSmem<P>::BYTES, [&] { k_inter_sb<P, HBD>(fp, 1); });
        if (g_full)
            run_wave<P, HBD>(fp, h, lanes, sbi, lf, cur, m);
    }
    if (g_full)
        bad += compare_full<P>(f, h, lanes, sbi, lf, cur, m);
    for (int pl = 0; pl < 3 && !g_full; pl++) {
        const int ch = pl > 0 && h.ss_h, cs = 8 >> ch;
        for (int y = 0; y < (rows * 8) >> ch; y++)
            for (int x = 0; x < (cols * 8) >> ch; x++) {
                if (cellmap[(size_t)(y / cs) * cols + x / cs] < 0)
                    continue;
                g_compared++;
                const int want = mx.cur->p[pl][(size_t)y * mx.cur->stride[pl] + x], got = cur[pl][(size_t)y * f.cur.stride[pl] + x];
                if (want != got && bad++ < 12)
                    fprintf(stderr, "  plane %d (%d, %d): kernel %d, restatement %d\n", pl, x, y, got, want);
            }
    }
    return bad;
}

int main(int argc, char **argv)
{
    if (argc < 2) { fprintf(stderr, "usage: %s in.v9r [max_frames]\n", argv[0]); return 2; }
    if (argc > 1 && !strcmp(argv[1], "-full")) { g_full = true; argv++; argc--; }
    const int max_frames = argc > 2 ? atoi(argv[2]) : 1 << 30;
    FILE *fp = fopen(argv[1], "rb");
    if (!fp) { perror("open"); return 2; }
    char magic[4]; uint32_t sz[8];
    rd(fp, magic, 4); rd(fp, sz, 32);
    if (memcmp(magic, "V9R3", 4) || sz[0] != sizeof(Vp9CudaFrameHeader) || sz[7] != VP9CUDA_ABI_VERSION) {
        fprintf(stderr, "bad .v9r header\n"); return 2;
    }
    Model m;
    int nf = 0, ninter = 0;
    long bad = 0;
    while (nf < max_frames && fread(magic, 1, 4, fp) == 4) {
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
        size_t nblk = 0;
        for (const Lane &l : lanes) nblk += l.blk.size();
        if (nblk || g_full) {
            const long b = h.bpp > 8 ? run_frame<uint16_t, true>(m, h, lanes, sbi, lf) : run_frame<uint8_t, false>(m, h, lanes, sbi, lf);
            if (b) fprintf(stderr, "frame %d: %ld differing pixels\n", nf, b);
            bad += b;
            ninter++;
        }
        decode_frame(m, h, lanes, sbi, lf);
        nf++;
    }
    fprintf(stderr, "emu_inter: %d frames, %d with inter blocks, %ld samples compared, %ld differing\n", nf, ninter, g_compared, bad);
    return bad ? 1 : 0;
}
