"""The five BASELINE.json workloads (and small variants for the tests) as StreamSpec objects.

Feature mix per config follows SURVEY.md section 8(d).  All streams are produced by
streamgen.encode_stream (valid headers + seeded model-distributed tile payloads).
"""
from __future__ import annotations

from typing import Dict

from .streamgen import (FILTER_SWITCHABLE, PRED_SWITCHABLE, TX_4X4, TX_8X8, TX_16X16, TX_32X32, TX_SWITCHABLE,
                        FrameSpec, StreamSpec)


def _ippp(n: int, tile_cols: int = 0, compound: bool = False, **kw) -> list:
    """key frame + (n-1) inter frames cycling through the three reference slots."""
    frames = [FrameSpec(keyframe=True, refreshrefmask=0xFF, log2_tile_cols=tile_cols, **kw)]
    for i in range(1, n):
        frames.append(FrameSpec(
            refreshrefmask=1 << (i % 3), refidx=(0, 1, 2),
            signbias=(0, 0, 1) if compound else (0, 0, 0),
            comppredmode=PRED_SWITCHABLE, filtermode=FILTER_SWITCHABLE,
            txfmmode=TX_SWITCHABLE, log2_tile_cols=tile_cols, **kw))
    return frames


def c1(frames: int = 30, seed: int = 1) -> StreamSpec:
    """352x288 8-bit profile 0, 1 key + inter, single tile."""
    st = StreamSpec(352, 288, seed=seed)
    st.frames = _ippp(frames)
    return st


def c2(frames: int = 60, seed: int = 2, w: int = 1920, h: int = 1080) -> StreamSpec:
    """1080p 8-bit IPPP, 4 tile columns, compound prediction, switchable filters."""
    st = StreamSpec(w, h, seed=seed)
    st.frames = _ippp(frames, tile_cols=2, compound=True)
    return st


def c3(frames: int = 30, seed: int = 3, w: int = 3840, h: int = 2160) -> StreamSpec:
    """4K 10-bit profile 2, 32x32 transforms, scaled-reference (resize) frames.

    Every shown frame is w x h.  Every 6th coded frame is a HIDDEN half-size inter frame
    (predicted from full-size references => down-scaled motion compensation) that refreshes
    slot 3; the shown frames after it predict from slot 3 as their third reference
    (=> up-scaled motion compensation, ratio 2x, within vp9.c:862)."""
    st = StreamSpec(w, h, profile=2, bitdepth=10, seed=seed)
    fr = [FrameSpec(keyframe=True, refreshrefmask=0xFF, log2_tile_cols=2)]
    shown = 1
    i = 1
    while shown < frames:
        if i % 6 == 3:
            fr.append(FrameSpec(show=False, width=w // 2, height=h // 2, refreshrefmask=1 << 3,
                                refidx=(0, 1, 2), log2_tile_cols=2))
        else:
            fr.append(FrameSpec(refreshrefmask=1 << (shown % 3), refidx=(0, 1, 3) if i > 3 else (0, 1, 2),
                                log2_tile_cols=2))
            shown += 1
        i += 1
    st.frames = fr
    return st


def c3p(frames: int = 30, seed: int = 3, w: int = 3840, h: int = 2160) -> StreamSpec:
    """C3 authored for a parallel host (SURVEY.md section 7 "hard parts" (b)): the same 4K 10-bit
    content model, but every frame sets frame_parallel_decoding_mode (no backward probability
    adaptation, so the entropy decode of frame N+1 does not wait for frame N, vp9.c:1731-1756) and
    the frames carry 8 tile columns instead of 4."""
    st = c3(frames=frames, seed=seed, w=w, h=h)
    for f in st.frames:
        f.parallelmode = True
        f.log2_tile_cols = 3
    return st


def natural_stats(st: StreamSpec) -> StreamSpec:
    """Give a stream the statistics of encoded video instead of the default model's: every frame's header raises
    the end-of-block / zero-token / skip probabilities (sparse coefficients) and, on inter frames, lowers the
    intra probability and raises the no-split partition probabilities (large blocks, a few per cent of intra
    blocks).  The updates are differential, coded against the context the generator tracks (streamgen.ProbTracker).  The payload is still a seeded random sample -- of this model."""
    part = [0] * 48
    none = {0: (235, 210, 210, 180), 1: (225, 200, 200, 160), 2: (215, 180, 180, 140), 3: (230, 200, 200, 170)}
    from .streamgen import default_probs
    d = default_probs()["partition"]
    for lvl in range(4):
        for ctx in range(4):
            part[(lvl * 4 + ctx) * 3:(lvl * 4 + ctx) * 3 + 3] = [none[lvl][ctx], d[(lvl * 4 + ctx) * 3 + 1], d[(lvl * 4 + ctx) * 3 + 2]]
    for f in st.frames:
        if f.show_existing >= 0:
            continue
        f.prob_update = {"coef": (1.25, 1.1, 1.0), "skip": [200, 140, 80]}
        if not f.keyframe and not f.intraonly:
            f.prob_update.update({"intra": [3, 8, 24, 60], "partition": part})
    return st


def c3n(frames: int = 30, seed: int = 3, w: int = 3840, h: int = 2160) -> StreamSpec:
    """C3 with natural statistics (natural_stats): the same 4K 10-bit profile 2 structure -- 32x32 transforms,
    hidden half-size frames, scaled references, 4 tile columns, backward adaptation -- but inter frames made
    of large, mostly skipped inter blocks with sparse residuals and only a few per cent of intra blocks."""
    return natural_stats(c3(frames=frames, seed=seed, w=w, h=h))


def c4(frames: int = 8, seed: int = 4, w: int = 7680, h: int = 4320) -> StreamSpec:
    """8K 8-bit, all key frames (intra + loop-filter wavefront bound); second half at max LF level."""
    st = StreamSpec(w, h, seed=seed)
    st.frames = [FrameSpec(keyframe=True, refreshrefmask=0xFF, log2_tile_cols=3,
                           lf_level=32 if i < frames // 2 else 63, lf_sharpness=0)
                 for i in range(frames)]
    return st


def c5(rank: int, frames: int = 30, w: int = 3840, h: int = 2160) -> StreamSpec:
    """one of eight independent 4K 8-bit C2-style streams (one per GPU)."""
    return c2(frames=frames, seed=50 + rank, w=w, h=h)


def small_variants() -> Dict[str, StreamSpec]:
    """Reduced-size relatives of the five configs + edge cases, decoded in seconds on a CPU."""
    out: Dict[str, StreamSpec] = {}
    out["c1_small"] = c1(frames=6)
    out["c2_small"] = c2(frames=5, w=456, h=264)           # non-multiple-of-8 size, 4 tile cols cap
    out["c3_small"] = c3(frames=8, w=400, h=240)
    out["c3p_small"] = c3p(frames=8, w=528, h=240)          # frame-parallel headers + 2 tile columns
    out["c4_small"] = c4(frames=2, w=520, h=392)
    st = StreamSpec(200, 136, profile=2, bitdepth=12, seed=7)  # 12-bit
    st.frames = _ippp(4, compound=True)
    out["hbd12"] = st
    st = StreamSpec(328, 200, seed=8)                        # lossless (WHT) key + inter
    st.frames = _ippp(3, yac_qi=0)
    out["lossless"] = st
    st = StreamSpec(296, 168, seed=9)                        # sharpness, deltas, high q
    st.frames = _ippp(4, lf_level=40, lf_sharpness=5, yac_qi=200, ydc_qdelta=-3, uvdc_qdelta=4,
                      uvac_qdelta=-5, lf_ref_deltas=(2, 0, -2, -4), lf_mode_deltas=(1, -1))
    out["lf_q"] = st
    st = StreamSpec(264, 264, seed=10)                       # loop filter off, fixed 8-tap sharp
    st.frames = _ippp(3, lf_level=0)
    for f in st.frames[1:]:
        f.filtermode = 2
    out["nolf"] = st
    st = StreamSpec(200, 120, seed=12)                       # hidden intra-only frame mid-stream
    st.frames = _ippp(5)
    st.frames[2] = FrameSpec(intraonly=True, show=False, refreshrefmask=0x06, resetctx=0)
    out["intraonly"] = st
    st = StreamSpec(328, 392, seed=13)                       # 2 tile rows x 2 tile columns
    st.frames = _ippp(4, tile_cols=1, compound=True, log2_tile_rows=1)
    out["tilerows"] = st
    st = StreamSpec(264, 200, seed=14)                       # error-resilient / frame-parallel headers
    st.frames = _ippp(4, errorres=True)
    for f in st.frames:
        f.refreshctx = False
        f.parallelmode = True
    out["parallel"] = st
    st = StreamSpec(296, 232, seed=15)                       # strongest loop filter
    st.frames = _ippp(3, lf_level=63, lf_sharpness=7)
    out["lf63"] = st
    st = StreamSpec(453, 267, seed=16)                       # size not a multiple of 8
    st.frames = _ippp(4, tile_cols=1, compound=True)
    out["odd"] = st
    st = StreamSpec(264, 200, seed=17)                       # hidden frame + shown frame in one superframe packet,
    st.frames = _ippp(5, compound=True)                      # later shown again with show_existing_frame
    st.frames[2].show = False
    st.frames[2].refreshrefmask = 0x04
    st.frames[2].superframe = True
    st.frames[4] = FrameSpec(show_existing=2)
    out["superframe"] = st
    st = StreamSpec(328, 264, seed=18)                       # segmentation: per-segment quantiser / loop-filter
    st.frames = _ippp(6, compound=True)                      # level / reference / skip, every map mode
    none = (None, None, None, False)
    deltas = [none, (40, -10, None, False), (-30, 12, None, False), (90, None, None, False),
              (None, 20, None, False), (-50, -16, 1, False), (10, 4, None, True), (25, None, 0, False)]
    absolute = [(60, 16, None, False), (0, 0, None, False), (255, 63, None, False), (120, 40, None, False),
                none, (30, 8, 2, False), (200, 50, None, True), (90, 24, None, False)]
    tree = (120, 100, 140, 128, 90, 160, 110)
    st.frames[0].segmentation = {"probs": tree, "feats": deltas}                      # key frame: map + data
    st.frames[1].segmentation = {"probs": tree, "pred_probs": (200, 128, 60)}          # temporally predicted map
    st.frames[2].segmentation = {}                                                      # map and data kept
    st.frames[3].segmentation = {"probs": (128, None, 128, 128, None, 128, 128), "abs": True, "feats": absolute}
    st.frames[4].segmentation = {"pred_probs": None, "probs": tree}                     # coded map, no prediction
    out["segment"] = st                                      # frame 5: segmentation off again
    st = StreamSpec(512, 320, seed=19)                       # reference scaling at its limits (vp9.c:854-877):
    kw = dict(filtermode=FILTER_SWITCHABLE, txfmmode=TX_SWITCHABLE)      # 2:1 down, 16:1 up, non-integer and
    st.frames = [                                                        # anisotropic ratios, scaled + unscaled
        FrameSpec(keyframe=True, refreshrefmask=0xFF),                   # references (and compound) in one frame
        FrameSpec(show=False, width=256, height=160, refreshrefmask=1 << 3, refidx=(0, 1, 2), **kw),
        FrameSpec(intraonly=True, show=False, width=32, height=20, refreshrefmask=1 << 4, resetctx=0),
        FrameSpec(refreshrefmask=1 << 5, refidx=(4, 4, 4), **kw),
        FrameSpec(width=352, height=224, refreshrefmask=1 << 6, refidx=(0, 5, 1), signbias=(0, 0, 1), **kw),
        FrameSpec(refreshrefmask=1 << 1, refidx=(0, 6, 3), signbias=(0, 1, 0), **kw),
        FrameSpec(width=416, height=260, refreshrefmask=1 << 2, refidx=(6, 0, 3), signbias=(0, 0, 1), **kw),
    ]
    out["scale_mix"] = st
    st = StreamSpec(4096, 136, seed=20)                      # 16 tile columns (the most a 4096-wide frame may
    st.frames = _ippp(3, tile_cols=4, compound=True, log2_tile_rows=1)   # have, vp9.c:800-810) x 2 tile rows
    out["manytiles"] = st
    st = StreamSpec(296, 200, seed=21)                       # every fixed interpolation filter (header literals
    st.frames = _ippp(6, compound=True)                      # 0..3 incl. bilinear) x every fixed transform mode,
    for i, f in enumerate(st.frames):                        # integer-precision motion vectors on odd frames
        f.txfmmode = (TX_32X32, TX_4X4, TX_8X8, TX_16X16, TX_32X32, TX_SWITCHABLE)[i]
        if i:
            f.filtermode = (0, 1, 2, 3, FILTER_SWITCHABLE)[i - 1]
            f.highprecisionmvs = i % 2 == 0
    out["filters_tx"] = st
    st = StreamSpec(264, 168, seed=22)                       # coarsest quantiser: largest coefficients the 8-bit
    st.frames = _ippp(3, compound=True, yac_qi=255, ydc_qdelta=15, uvdc_qdelta=-15, uvac_qdelta=15)
    out["maxq"] = st                                         # (int16, wrapping) transform path sees
    st = StreamSpec(200, 136, profile=2, bitdepth=12, seed=23)   # ... and the 12-bit (int32) one
    st.frames = _ippp(3, compound=True, yac_qi=255, ydc_qdelta=15, uvdc_qdelta=15, uvac_qdelta=-15)
    out["maxq12"] = st
    st = StreamSpec(200, 120, seed=24)                       # key frames that change the size mid-stream (surface
    kw = dict(filtermode=FILTER_SWITCHABLE, txfmmode=TX_SWITCHABLE)      # pool re-created), size taken from a
    st.frames = [                                                        # reference, non-reference frames, one slot
        FrameSpec(keyframe=True, refreshrefmask=0xFF),                   # named three times
        FrameSpec(refreshrefmask=0x00, refidx=(0, 0, 0), **kw),
        FrameSpec(refreshrefmask=0x02, refidx=(0, 1, 2), size_from_ref=1, **kw),
        FrameSpec(keyframe=True, refreshrefmask=0xFF, width=328, height=264),
        FrameSpec(refreshrefmask=0x05, refidx=(2, 2, 5), size_from_ref=0, signbias=(0, 0, 1), **kw),
        FrameSpec(refreshrefmask=0x00, refidx=(0, 2, 7), size_from_ref=2, **kw),
        FrameSpec(keyframe=True, refreshrefmask=0xFF, width=72, height=40),
        FrameSpec(refreshrefmask=0x80, refidx=(7, 3, 1), size_from_ref=2, signbias=(1, 0, 0), **kw),
    ]
    out["keyresize"] = st
    st = StreamSpec(640, 384, seed=25)                       # fixed bilinear filter at 1/16-pel precision with
    st.frames = _ippp(9, compound=True)                      # compound prediction on 64-wide blocks: the bilinear
    for f in st.frames[1:]:                                  # mc slots the switchable streams rarely reach
        f.filtermode = 3
    out["bilinear"] = st
    st = StreamSpec(640, 384, seed=26)                       # scaled references (2:1 down and up) under every fixed
    kw = dict(txfmmode=TX_SWITCHABLE)                        # filter, single and compound: the smc slots
    st.frames = [FrameSpec(keyframe=True, refreshrefmask=0xFF),
                 FrameSpec(show=False, width=320, height=192, refreshrefmask=1 << 3, refidx=(0, 1, 2),
                           signbias=(0, 0, 1), filtermode=3, **kw)]
    for i, fm in enumerate((3, 0, 2, 1, 3, 0, FILTER_SWITCHABLE, 0)):
        st.frames.append(FrameSpec(refreshrefmask=1 << (i % 3), refidx=(3, 0, 3) if i % 2 else (0, 3, 3),
                                   signbias=(0, 0, 1), filtermode=fm, **kw))
    out["scale_filters"] = st
    st = StreamSpec(1024, 576, seed=27)                      # many 64- and 32-wide compound blocks under each fixed
    st.frames = _ippp(11, tile_cols=1, compound=True)        # 8-tap filter with quarter-pel vectors (integer phase in
    for i, f in enumerate(st.frames[1:]):                    # one direction only is common then)
        f.filtermode = (0, 2, 0, 2, 1, 0, 2, 3, 0, 2)[i]
        f.highprecisionmvs = False
    out["wide64"] = st
    out["c3n_small"] = c3n(frames=8, w=784, h=464)           # natural statistics through forward probability updates
    # 4:4:4 (profiles 1 and 3): every plane predicted, transformed and filtered like luma
    st = StreamSpec(328, 200, profile=1, bitdepth=8, seed=61, ss_h=0, ss_v=0)     # 8-bit, compound, 2 tile columns
    st.frames = _ippp(5, tile_cols=1, compound=True)
    out["p1_444"] = st
    st = StreamSpec(400, 240, profile=3, bitdepth=10, seed=62, ss_h=0, ss_v=0)    # 10-bit with C3's hidden half-size
    st.frames = c3(frames=6, w=400, h=240).frames                                   # frames and scaled references
    out["p3_444_10"] = st
    st = StreamSpec(200, 136, profile=3, bitdepth=12, seed=63, ss_h=0, ss_v=0)    # 12-bit, intra-only frame, lossless tail
    st.frames = _ippp(5, compound=True)
    st.frames[2] = FrameSpec(intraonly=True, show=False, refreshrefmask=0x06, resetctx=0)
    st.frames[4].yac_qi = 0
    out["p3_444_12"] = st
    st = StreamSpec(72, 40, seed=11)                         # tiny, single superblock column
    st.frames = _ippp(4, compound=True)
    out["tiny"] = st
    return out


def damaged_variants() -> Dict[str, StreamSpec]:
    """Streams with one deliberately broken frame: the decoder must report the error for that packet,
    keep its reference slots as they were and decode the following packets like the reference does."""
    out: Dict[str, StreamSpec] = {}
    for name, frame, damage in (("bad_marker", 2, "marker:1"), ("bad_tilesize", 3, "tilesize"),
                                ("bad_first_tile", 1, "marker:0")):
        st = c2(frames=6, w=456, h=264, seed=30 + len(out))
        st.frames[frame].damage = damage
        out[name] = st
    st = c2(frames=14, w=456, h=264, seed=35)                # four broken frames in one stream, two of them back
    for frame, damage in ((2, "marker:1"), (5, "tilesize"), (6, "marker:0"), (10, "marker:1")):   # to back: every
        st.frames[frame].damage = damage                     # failed frame must give its ring set back
    out["bad_many"] = st
    st = c1(frames=5, seed=34)                               # second key frame with a broken sync code
    st.frames[2] = FrameSpec(keyframe=True, refreshrefmask=0xFF, damage="sync")
    out["bad_sync"] = st
    return out
