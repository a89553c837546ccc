/*
 * vp9author.c -- TEST TOOLING: a syntax-driven VP9 tile writer that uses the reference's own parser
 * as its grammar.
 *
 * The reference decoder (built from /root/reference by tools/vp9author.mk with three small hooks, see
 * tools/vp9author_overlay.py) is run over a seed stream (real headers from streamgen.py, any tile
 * payload).  Inside the tiles every bool the parser asks for -- vpx_rac_get_prob[_branchy],
 * vp89_rac_get_tree: decode_mode vp9block.c:80-802, decode_coeffs vp9block.c:805-923, decode_sb
 * vp9.c:1115-1242, find_ref_mvs / read_mv vp9mvs.c -- is not READ from the payload but CHOSEN here
 * and ENCODED with a boolean encoder (the inverse of vpx_rac.h:84-107) under the very probability the
 * parser passed, i.e. with the parser's own contexts, neighbour rules and adapted probabilities.  The
 * parser then continues with the chosen symbol, so what it walks through IS the decode of the
 * stream being written: contexts, motion-vector predictors and backward adaptation of later frames
 * follow the authored content by construction.  Output: the seed's headers + the encoded tiles.
 *
 * What is chosen:
 *   default          : the symbol is sampled from the parser's probability (like a random payload);
 *   motion vectors   : for NEWMV blocks the writer picks the FINAL vector -- uniform in +-mv_range
 *                      (1/8 pel; default +-64 pixels, so every sub-pel phase and, near the frame's
 *                      edges, off-frame windows on all four sides occur by construction) -- and
 *                      encodes joint / sign / class / bits / fraction / hp of (target - predictor),
 *                      the predictor being the one find_ref_mvs just computed (vp9mvs.c:297-360);
 *   inter mode       : newmv=<percent> of the inter-mode trees are forced to NEWMV;
 *   partition        : part_none=<percent> of the partition trees are forced to PARTITION_NONE (large blocks);
 *   tile marker      : 0 (vp9.c:1345).
 * Policy: $VP9AUTHOR = "seed=7,mv_range=512,newmv=70,part_none=50".
 *
 * usage: ffvp9_author seed.ivf out.ivf        (single-frame packets; decoder runs on one thread)
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "libavcodec/avcodec.h"
#include "libavcodec/vp9.h"
#include "libavcodec/vp9data.h"
#include "libavcodec/vp9dec.h"
#include "libavutil/frame.h"

/* ---- boolean encoder (inverse of vpx_rac_get_prob) ------------------------------------------------ */
typedef struct BoolEnc {
    uint8_t *buf;
    size_t pos, cap;
    unsigned low, range;
    int count;
} BoolEnc;

static void enc_start(BoolEnc *e)
{
    e->pos = 0; e->low = 0; e->range = 255; e->count = -24;
}
static void enc_byte(BoolEnc *e, unsigned v)
{
    if (e->pos == e->cap) {
        e->cap = e->cap ? 2 * e->cap : 4096;
        e->buf = realloc(e->buf, e->cap);
    }
    e->buf[e->pos++] = (uint8_t)v;
}
static void enc_put(BoolEnc *e, int bit, int prob)
{
    unsigned split = 1 + (((e->range - 1) * prob) >> 8);
    int shift = 0;
    if (bit) { e->low += split; e->range -= split; } else e->range = split;
    while ((e->range << shift) < 128) shift++;
    e->range <<= shift;
    e->count += shift;
    if (e->count >= 0) {
        int offset = shift - e->count;
        if ((e->low << (offset - 1)) & 0x80000000u) {
            long x = (long)e->pos - 1;
            while (x >= 0 && e->buf[x] == 0xff) e->buf[x--] = 0;
            if (x >= 0) e->buf[x]++;
        }
        enc_byte(e, e->low >> (24 - offset));
        e->low = (e->low << offset) & 0xffffff;
        shift = e->count;
        e->count -= 8;
    }
    e->low <<= shift;
}
static void enc_finish(BoolEnc *e)
{
    for (int i = 0; i < 32; i++) enc_put(e, 0, 128);
}

/* ---- state ------------------------------------------------------------------------------------------ */
#define MAX_TILES 256
typedef struct Tile {
    const VPXRangeCoder *c;
    BoolEnc enc;
    int fresh;                       /* the next bool is the tile marker */
} Tile;

static struct {
    Tile tile[MAX_TILES];
    int n_tiles;
    size_t tile_off;                 /* where the tile data starts in the current packet */
    int have_frame;
    uint8_t script[128];             /* bools to hand out before anything is sampled again */
    int script_len, script_pos;
    const VPXRangeCoder *script_c;
    uint64_t rng;
    int mv_range, newmv_pct, part_none_pct;
} A = { .rng = 0x9E3779B97F4A7C15ull, .mv_range = 512, .newmv_pct = 0, .part_none_pct = 0 };

static unsigned rnd(void)
{
    A.rng ^= A.rng << 13; A.rng ^= A.rng >> 7; A.rng ^= A.rng << 17;
    return (unsigned)(A.rng >> 24);
}

static Tile *tile_of(const VPXRangeCoder *c)
{
    for (int i = A.n_tiles - 1; i >= 0; i--)     /* the newest registration of a pointer wins: tile rows reuse c_b[] */
        if (A.tile[i].c == c)
            return &A.tile[i];
    return NULL;
}

/* ---- hooks called from the patched copies of vp9.c / vpx_rac.h / vp89_rac.h / vp9mvs.c -------------- */
void ff_vp9author_frame_begin(size_t tile_off)
{
    for (int i = 0; i < A.n_tiles; i++) A.tile[i].c = NULL;
    A.n_tiles = 0;
    A.tile_off = tile_off;
    A.have_frame = 1;
    A.script_len = A.script_pos = 0;
}

void ff_vp9author_register(const VPXRangeCoder *c)
{
    if (A.n_tiles == MAX_TILES) { fprintf(stderr, "vp9author: too many tiles\n"); exit(2); }
    Tile *t = &A.tile[A.n_tiles++];
    t->c = c;
    t->fresh = 1;
    enc_start(&t->enc);
}

int ff_vp9author_bit(const VPXRangeCoder *c, int prob)
{
    Tile *t = tile_of(c);
    int bit;
    if (!t)
        return -1;                               /* header coder: decode from the input as usual */
    if (t->fresh) {
        t->fresh = 0;
        bit = 0;                                 /* tile marker */
    } else if (A.script_pos < A.script_len && A.script_c == c) {
        bit = A.script[A.script_pos++];
    } else {
        bit = (int)(rnd() & 0xff) >= prob;       /* P(bit == 0) = prob / 256, the decoder's own model */
    }
    enc_put(&t->enc, bit, prob);
    return bit;
}

/* the bools that lead to leaf `value` of a tree (vp89_rac.h:49-61); returns the number written or -1 */
static int tree_path(const int8_t (*tree)[2], int node, int value, uint8_t *out, int depth)
{
    for (int b = 0; b < 2; b++) {
        int nx = tree[node][b];
        out[depth] = (uint8_t)b;
        if (nx <= 0) {
            if (-nx == value) return depth + 1;
        } else {
            int r = tree_path(tree, nx, value, out, depth + 1);
            if (r >= 0) return r;
        }
    }
    return -1;
}

static void script_tree(const int8_t (*tree)[2], int value)
{
    int n = tree_path(tree, 0, value, A.script + A.script_len, 0);
    if (n < 0) { fprintf(stderr, "vp9author: value %d not in tree\n", value); exit(2); }
    A.script_len += n;
}
static void script_bit(int b) { A.script[A.script_len++] = (uint8_t)b; }

/* vp89_rac_get_tree for an authored coder: -1 = not authored */
int ff_vp9author_tree(const VPXRangeCoder *c, const int8_t (*tree)[2], const uint8_t *probs)
{
    Tile *t = tile_of(c);
    int i = 0;
    if (!t)
        return -1;
    if (tree == ff_vp9_inter_mode_tree && !(A.script_pos < A.script_len) && (int)(rnd() % 100) < A.newmv_pct) {
        A.script_len = A.script_pos = 0;
        A.script_c = c;
        script_tree(tree, NEWMV);
    }
    if (tree == ff_vp9_partition_tree && !(A.script_pos < A.script_len) && (int)(rnd() % 100) < A.part_none_pct) {
        A.script_len = A.script_pos = 0;
        A.script_c = c;
        script_tree(tree, PARTITION_NONE);
    }
    do {
        i = tree[i][ff_vp9author_bit(c, probs[i])];
    } while (i > 0);
    return -i;
}

/* one motion-vector component difference d != 0 in the order read_mv_component reads it (vp9mvs.c:236-295) */
static void script_mv_component(int d, int hp)
{
    int n = abs(d) - 1, c = 0;
    script_bit(d < 0);
    while (c < 10 && n >= (16 << c)) c++;          /* class c holds n in [8 << c, 16 << c), class 0: n < 16 */
    script_tree(ff_vp9_mv_class_tree, c);
    if (c) {
        int r = n - (8 << c);
        for (int m = 0; m < c; m++) script_bit(r >> (3 + m) & 1);
        script_tree(ff_vp9_mv_fp_tree, r >> 1 & 3);
        if (hp) script_bit(r & 1);
    } else {
        script_bit(n >> 3);
        script_tree(ff_vp9_mv_fp_tree, n >> 1 & 3);
        if (hp) script_bit(n & 1);
    }
}

/* ff_vp9_fill_mv, NEWMV, right before the joint is read: pred = the predictor find_ref_mvs produced (already
 * rounded when !hp).  Picks the final vector and scripts joint + components of (target - pred). */
void ff_vp9author_mv(const VPXRangeCoder *c, const VP9mv *pred, int hp)
{
    if (!tile_of(c))
        return;
    int tx = (int)(rnd() % (2 * A.mv_range + 1)) - A.mv_range;
    int ty = (int)(rnd() % (2 * A.mv_range + 1)) - A.mv_range;
    if (!hp) {                                     /* without the hp bit a difference is even, the predictor too */
        tx &= ~1; ty &= ~1;
    }
    int dx = tx - pred->x, dy = ty - pred->y;
    A.script_len = A.script_pos = 0;
    A.script_c = c;
    script_tree(ff_vp9_mv_joint_tree, (dx ? MV_JOINT_H : 0) | (dy ? MV_JOINT_V : 0));
    if (dy) script_mv_component(dy, hp);
    if (dx) script_mv_component(dx, hp);
}

/* ---- the tool ------------------------------------------------------------------------------------- */
static unsigned rl32(const unsigned char *p) { return p[0] | p[1] << 8 | p[2] << 16 | (unsigned)p[3] << 24; }
static void wl32(unsigned char *p, unsigned v) { p[0] = v; p[1] = v >> 8; p[2] = v >> 16; p[3] = v >> 24; }

int main(int argc, char **argv)
{
    if (argc < 3) { fprintf(stderr, "usage: %s seed.ivf out.ivf   ($VP9AUTHOR=seed=..,mv_range=..,newmv=..)\n", argv[0]); return 2; }
    if (getenv("VP9AUTHOR")) {
        char *s = strdup(getenv("VP9AUTHOR"));
        for (char *tok = strtok(s, ","); tok; tok = strtok(NULL, ",")) {
            int v;
            if (sscanf(tok, "seed=%d", &v) == 1) A.rng ^= (uint64_t)v * 0xD6E8FEB86659FD93ull;
            else if (sscanf(tok, "mv_range=%d", &v) == 1) A.mv_range = v;
            else if (sscanf(tok, "newmv=%d", &v) == 1) A.newmv_pct = v;
            else if (sscanf(tok, "part_none=%d", &v) == 1) A.part_none_pct = v;
            else { fprintf(stderr, "vp9author: unknown policy item %s\n", tok); return 2; }
        }
        free(s);
    }
    FILE *fp = fopen(argv[1], "rb"), *out = fopen(argv[2], "wb");
    if (!fp || !out) { perror("open"); return 2; }
    unsigned char hdr[32], fh[12];
    if (fread(hdr, 1, 32, fp) != 32 || memcmp(hdr, "DKIF", 4)) { fprintf(stderr, "not IVF\n"); return 2; }
    fwrite(hdr, 1, 32, out);
    const AVCodec *codec = avcodec_find_decoder(AV_CODEC_ID_VP9);
    AVCodecContext *c = avcodec_alloc_context3(codec);
    c->thread_count = 1;
    c->flags |= AV_CODEC_FLAG_BITEXACT;
    c->err_recognition |= AV_EF_EXPLODE;
    if (avcodec_open2(c, codec, NULL) < 0) { fprintf(stderr, "open failed\n"); return 2; }
    AVPacket *pkt = av_packet_alloc();
    AVFrame *fr = av_frame_alloc();
    int n = 0;
    size_t total = 0;
    while (fread(fh, 1, 12, fp) == 12) {
        unsigned fsz = rl32(fh);
        if (av_new_packet(pkt, fsz) < 0 || fread(pkt->data, 1, fsz, fp) != fsz) { fprintf(stderr, "truncated\n"); return 2; }
        A.have_frame = 0;
        if (avcodec_send_packet(c, pkt) < 0) { fprintf(stderr, "vp9author: the parser rejected packet %d\n", n); return 1; }
        while (avcodec_receive_frame(c, fr) >= 0) av_frame_unref(fr);
        if (!A.have_frame || !A.n_tiles) {           /* show_existing_frame: nothing to author */
            fwrite(fh, 1, 12, out);
            fwrite(pkt->data, 1, fsz, out);
        } else {
            size_t sz = A.tile_off;
            for (int i = 0; i < A.n_tiles; i++) {
                enc_finish(&A.tile[i].enc);
                sz += A.tile[i].enc.pos + (i + 1 < A.n_tiles ? 4 : 1);
            }
            wl32(fh, (unsigned)sz);
            fwrite(fh, 1, 12, out);
            fwrite(pkt->data, 1, A.tile_off, out);
            for (int i = 0; i < A.n_tiles; i++) {
                BoolEnc *e = &A.tile[i].enc;
                if (i + 1 < A.n_tiles) {
                    unsigned char be[4] = { e->pos >> 24, e->pos >> 16, e->pos >> 8, e->pos };
                    fwrite(be, 1, 4, out);
                }
                fwrite(e->buf, 1, e->pos, out);
            }
            fputc(0, out);                          /* the packet must not end like a superframe index */
            total += sz;
        }
        av_packet_unref(pkt);
        n++;
    }
    fclose(out);
    fprintf(stderr, "vp9author: %d packets, %zu bytes authored\n", n, total);
    return 0;
}
