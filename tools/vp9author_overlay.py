#!/usr/bin/env python3
"""Patched copies of five reference files for the stream-authoring build (tools/vp9author.c):

    tools/vp9author_overlay.py <reference tree> <out dir>      -> <out dir>/libavcodec/{vp9.c, vp9mvs.c, vp9dec.h, vpx_rac.h, vp89_rac.h}

Three hooks, applied as exact-string edits on fresh copies (the script fails if the reference text differs):
  * vpx_rac.h : vpx_rac_get_prob / _branchy ask ff_vp9author_bit() first (tile coders answer, header coders decline);
  * vp89_rac.h: vp89_rac_get_tree asks ff_vp9author_tree() first (trees are recognised by address);
  * vp9.c     : vp9_decode_frame reports where the tile data starts, decode_tiles registers every tile's coder;
  * vp9mvs.c  : ff_vp9_fill_mv hands the motion-vector predictor over before a NEWMV difference is read.
Nothing is written into the reference tree and the copies are build artefacts (oracle/_ref/author/, git-ignored)."""
import sys
from pathlib import Path

PROTOS = '''
/* stream authoring hooks (tools/vp9author.c) */
struct VPXRangeCoder;
int ff_vp9author_bit(const struct VPXRangeCoder *c, int prob);
int ff_vp9author_tree(const struct VPXRangeCoder *c, const int8_t (*tree)[2], const uint8_t *probs);
void ff_vp9author_register(const struct VPXRangeCoder *c);
void ff_vp9author_frame_begin(size_t tile_off);
'''


def edit(text: str, old: str, new: str, count: int = 1) -> str:
    assert text.count(old) == count, f"expected {count} x {old!r}, found {text.count(old)}"
    return text.replace(old, new)


def main():
    ref, out = Path(sys.argv[1]) / "libavcodec", Path(sys.argv[2]) / "libavcodec"
    out.mkdir(parents=True, exist_ok=True)
    s = (ref / "vpx_rac.h").read_text()
    s = edit(s, "extern const uint8_t ff_vpx_norm_shift[256];", PROTOS + "\nextern const uint8_t ff_vpx_norm_shift[256];")
    s = edit(s, "static av_always_inline int vpx_rac_get_prob(VPXRangeCoder *c, uint8_t prob)\n{\n",
             "static av_always_inline int vpx_rac_get_prob(VPXRangeCoder *c, uint8_t prob)\n{\n"
             "    { int a_ = ff_vp9author_bit(c, prob); if (a_ >= 0) return a_; }\n")
    s = edit(s, "static av_always_inline int vpx_rac_get_prob_branchy(VPXRangeCoder *c, int prob)\n{\n",
             "static av_always_inline int vpx_rac_get_prob_branchy(VPXRangeCoder *c, int prob)\n{\n"
             "    { int a_ = ff_vp9author_bit(c, prob); if (a_ >= 0) return a_; }\n")
    (out / "vpx_rac.h").write_text(s)

    s = (ref / "vp89_rac.h").read_text()
    s = edit(s, "    int i = 0;\n\n    do {\n        i = tree[i][vpx_rac_get_prob(c, probs[i])];",
             "    int i = ff_vp9author_tree(c, tree, probs);\n\n    if (i >= 0)\n        return i;\n    i = 0;\n"
             "    do {\n        i = tree[i][vpx_rac_get_prob(c, probs[i])];")
    (out / "vp89_rac.h").write_text(s)

    (out / "vp9dec.h").write_text((ref / "vp9dec.h").read_text())

    s = (ref / "vp9.c").read_text()
    s = edit(s, "            ret = ff_vpx_init_range_decoder(&td->c_b[tile_col], data, tile_size);\n            if (ret < 0)\n                return ret;\n",
             "            ret = ff_vpx_init_range_decoder(&td->c_b[tile_col], data, tile_size);\n            if (ret < 0)\n                return ret;\n"
             "            ff_vp9author_register(&td->c_b[tile_col]);\n")
    s = edit(s, "    data += ret;\n    size -= ret;\n", "    data += ret;\n    size -= ret;\n    ff_vp9author_frame_begin(data - pkt->data);\n")
    (out / "vp9.c").write_text(s)

    s = (ref / "vp9mvs.c").read_text()
    s = edit(s, "            enum MVJoint j = vp89_rac_get_tree(td->c, ff_vp9_mv_joint_tree,\n                                               s->prob.p.mv_joint);\n\n"
                "            td->counts.mv_joint[j]++;\n            if (j >= MV_JOINT_V)\n                mv[0].y += read_mv_component(td, 0, hp);",
             "            enum MVJoint j;\n            ff_vp9author_mv(td->c, &mv[0], hp);\n"
             "            j = vp89_rac_get_tree(td->c, ff_vp9_mv_joint_tree,\n                                               s->prob.p.mv_joint);\n\n"
             "            td->counts.mv_joint[j]++;\n            if (j >= MV_JOINT_V)\n                mv[0].y += read_mv_component(td, 0, hp);")
    s = edit(s, "                enum MVJoint j = vp89_rac_get_tree(td->c, ff_vp9_mv_joint_tree,\n                                                   s->prob.p.mv_joint);\n\n"
                "                td->counts.mv_joint[j]++;\n                if (j >= MV_JOINT_V)\n                    mv[1].y += read_mv_component(td, 0, hp);",
             "                enum MVJoint j;\n                ff_vp9author_mv(td->c, &mv[1], hp);\n"
             "                j = vp89_rac_get_tree(td->c, ff_vp9_mv_joint_tree,\n                                                   s->prob.p.mv_joint);\n\n"
             "                td->counts.mv_joint[j]++;\n                if (j >= MV_JOINT_V)\n                    mv[1].y += read_mv_component(td, 0, hp);")
    s = edit(s, '#include "vpx_rac.h"\n', '#include "vpx_rac.h"\n\nvoid ff_vp9author_mv(const VPXRangeCoder *c, const VP9mv *pred, int hp);\n')
    (out / "vp9mvs.c").write_text(s)
    (Path(sys.argv[2]) / ".patched").write_text("ok\n")


if __name__ == "__main__":
    main()
