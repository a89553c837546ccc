"""Host logic of the synthetic stream generator: the boolean encoder is the exact inverse of the
decoder it targets (reference libavcodec/vpx_rac.h:84-107), headers parse, tiles split."""
import numpy as np

from ffmpeg_hybrid_b200 import configs
from ffmpeg_hybrid_b200.streamgen import BoolEncoder, _tile_bounds, encode_stream, resolve_sizes, tile_col_limits


class BoolDecoder:
    """plain restatement of vpx_rac.c:41-53 / vpx_rac.h:84-107 (bit-at-a-time form)"""

    def __init__(self, data: bytes):
        self.data = data + b"\0" * 8
        self.pos = 0
        self.value = 0
        self.count = -8
        self.range = 255
        self._fill()

    def _fill(self):
        while self.count < 0:
            self.value |= self.data[self.pos] << (-self.count + 16)   # keep 24+ bits of window
            self.pos += 1
            self.count += 8

    def get(self, prob: int) -> int:
        split = 1 + (((self.range - 1) * prob) >> 8)
        big = split << 24
        if self.value >= big:
            self.range -= split
            self.value -= big
            bit = 1
        else:
            self.range = split
            bit = 0
        while self.range < 128:
            self.range <<= 1
            self.value = (self.value << 1) & 0xFFFFFFFF
            self.count -= 1
            self._fill()
        return bit


def test_bool_encoder_roundtrip():
    rng = np.random.default_rng(5)
    probs = rng.integers(1, 256, size=4000)
    bits = [(int(rng.integers(0, 256)) >= p) * 1 for p in probs]
    enc = BoolEncoder()
    for b, p in zip(bits, probs):
        enc.put(int(b), int(p))
    data = enc.finish()
    dec = BoolDecoder(data)
    got = [dec.get(int(p)) for p in probs]
    assert got == bits


def test_tile_layout_rules():
    assert tile_col_limits(30) == (0, 2)       # 1080p: up to 4 tile columns (vp9.c:800-804)
    assert tile_col_limits(120) == (1, 4)      # 8K needs at least 2
    assert _tile_bounds(0, 2, 30) == (0, 7) and _tile_bounds(3, 2, 30) == (22, 30)


def test_resolve_sizes_tracks_reference_slots():
    st = configs.c3(frames=8, w=400, h=240)
    sizes = resolve_sizes(st)
    assert sizes[0] == (400, 240)
    assert (200, 120) in sizes and sizes[-1] == (400, 240)


def test_ivf_container():
    data = encode_stream(configs.small_variants()["tiny"])
    assert data[:4] == b"DKIF" and data[8:12] == b"VP90"
    n = int.from_bytes(data[24:28], "little")
    pos, k = 32, 0
    while pos < len(data):
        sz = int.from_bytes(data[pos:pos + 4], "little")
        assert data[pos + 12] >> 6 == 2            # VP9 frame marker
        pos += 12 + sz
        k += 1
    assert k == n and pos == len(data)
