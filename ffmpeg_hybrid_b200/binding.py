"""ctypes binding of libvp9cuda.so (include/vp9cuda.h) + helpers to feed it .v9r record dumps.

This is plumbing for the tests and bench.py: the product is the shared library and the
libavcodec glue.  Loading fails loudly when the library is missing -- there is no fallback.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path
from typing import List, Optional, Sequence

import numpy as np

from . import v9r

LIB_PATH = Path(__file__).resolve().parent / "libvp9cuda.so"
MAX_LANES = 64

API = [
    "vp9cuda_version", "vp9cuda_create", "vp9cuda_destroy", "vp9cuda_flush", "vp9cuda_begin_frame",
    "vp9cuda_tile_ring", "vp9cuda_submit_tile", "vp9cuda_finish_frame", "vp9cuda_abort_frame", "vp9cuda_wait", "vp9cuda_get_stat",
    "vp9cuda_show_existing", "vp9cuda_host_alloc", "vp9cuda_host_free", "vp9cuda_decode_frame",
    "vp9cuda_stage_frame", "vp9cuda_run_staged", "vp9cuda_free_staged", "vp9cuda_frame_checksum",
    "vp9cuda_kernel_stats", "vp9cuda_set_option", "vp9cuda_elapsed_ms", "vp9cuda_phase_clocks", "vp9cuda_slot_itxfm_add",
    "vp9cuda_slot_intra_pred", "vp9cuda_slot_mc", "vp9cuda_slot_loop_filter", "vp9cuda_map_frame", "vp9cuda_unmap_frame",
    "vp9cuda_checksum_noalloc",
]


class MappedFrame(C.Structure):
    """Vp9CudaMappedFrame: a finished picture left in HBM for a GPU consumer."""
    _fields_ = [("planes", C.c_void_p * 3), ("strides", C.c_ssize_t * 3),
                ("w", C.c_int), ("h", C.c_int), ("bpp", C.c_int), ("ss_h", C.c_int), ("ss_v", C.c_int)]


class LaneRing(C.Structure):
    _fields_ = [("blk", C.c_void_p), ("blk_cap", C.c_uint32),
                ("tx", C.c_void_p), ("tx_cap", C.c_uint32),
                ("itx", C.c_void_p), ("itx_cap", C.c_uint32),
                ("coef", C.c_void_p), ("coef_cap", C.c_uint32),
                ("coef_base", C.c_uint32), ("reserved", C.c_uint32)]


class LaneCounts(C.Structure):
    _fields_ = [("blk", C.c_uint32), ("tx", C.c_uint32), ("itx", C.c_uint32), ("coef", C.c_uint32)]


class FrameRecords(C.Structure):
    _fields_ = [("lanes", C.POINTER(LaneRing)), ("counts", C.POINTER(LaneCounts)),
                ("sb_index", C.c_void_p), ("lf", C.c_void_p)]


class FrameRing(C.Structure):
    _fields_ = [("sb_index", C.c_void_p), ("lf", C.c_void_p), ("n_sb", C.c_uint32), ("frame_id", C.c_uint32)]


_lib = None


def load() -> C.CDLL:
    """dlopen libvp9cuda.so; raises if it is not built (no CPU fallback exists)."""
    global _lib
    if _lib is None:
        import os
        path = Path(os.environ.get("VP9CUDA_LIB", LIB_PATH))      # instrumented builds for the profiling tools
        if not path.exists():
            raise RuntimeError(f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`")
        lib = C.CDLL(str(path))
        lib.vp9cuda_version.restype = C.c_char_p
        lib.vp9cuda_host_alloc.restype = C.c_void_p
        lib.vp9cuda_host_alloc.argtypes = [C.c_size_t]
        lib.vp9cuda_host_free.argtypes = [C.c_void_p]
        lib.vp9cuda_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        lib.vp9cuda_destroy.argtypes = [C.c_void_p]
        lib.vp9cuda_destroy.restype = None
        lib.vp9cuda_flush.argtypes = [C.c_void_p]
        lib.vp9cuda_wait.argtypes = [C.c_void_p, C.c_uint64]
        lib.vp9cuda_decode_frame.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(FrameRecords),
                                             C.POINTER(C.c_void_p), C.POINTER(C.c_ssize_t), C.POINTER(C.c_uint64)]
        lib.vp9cuda_stage_frame.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(FrameRecords), C.POINTER(C.c_int)]
        lib.vp9cuda_run_staged.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_ssize_t),
                                           C.POINTER(C.c_uint64)]
        lib.vp9cuda_free_staged.argtypes = [C.c_void_p, C.c_int]
        lib.vp9cuda_frame_checksum.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
        lib.vp9cuda_map_frame.argtypes = [C.c_void_p, C.c_int, C.POINTER(MappedFrame)]
        lib.vp9cuda_unmap_frame.argtypes = [C.c_void_p, C.c_void_p]
        lib.vp9cuda_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        lib.vp9cuda_elapsed_ms.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double)]
        lib.vp9cuda_phase_clocks.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.c_int]
        lib.vp9cuda_begin_frame.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(FrameRing)]
        lib.vp9cuda_tile_ring.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.POINTER(LaneRing)]
        lib.vp9cuda_submit_tile.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.POINTER(LaneCounts)]
        lib.vp9cuda_finish_frame.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p), C.POINTER(C.c_ssize_t),
                                             C.POINTER(C.c_uint64)]
        lib.vp9cuda_get_stat.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_uint64)]
        lib.vp9cuda_abort_frame.argtypes = [C.c_void_p, C.c_uint32]
        lib.vp9cuda_kernel_stats.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_double),
                                             C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_int]
        lib.vp9cuda_checksum_noalloc.argtypes = [C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_ssize_t), C.c_int, C.c_int,
                                                 C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        _lib = lib
    return _lib


def check(rc: int, what: str) -> None:
    if rc < 0:
        raise RuntimeError(f"{what} failed with {rc}")


class HostFrame:
    """One frame's records in page-locked host memory, laid out like the lane rings."""

    def __init__(self, lib: C.CDLL, fr: v9r.Frame, pinned: bool = True):
        self.lib = lib
        self.pinned = pinned
        self._allocs: List[int] = []
        self._keep: List[object] = []
        self.hdr_buf = (C.c_char * v9r.HDR_DTYPE.itemsize).from_buffer_copy(fr.hdr.tobytes())
        n = len(fr.lanes)
        self.lanes = (LaneRing * n)()
        self.counts = (LaneCounts * n)()
        self.h2d_bytes = fr.record_bytes
        for i, l in enumerate(fr.lanes):
            r = self.lanes[i]
            r.blk = self._put(l.blk)
            r.tx = self._put(l.tx)
            r.itx = self._put(l.itx)
            r.coef = self._put(l.coef)
            r.coef_base = l.coef_base
            c = self.counts[i]
            c.blk = int(l.counts["blk"])
            c.tx = int(l.counts["tx"])
            c.itx = int(l.counts["itx"])
            c.coef = int(l.counts["coef"])
        self.recs = FrameRecords()
        self.recs.lanes = C.cast(self.lanes, C.POINTER(LaneRing))
        self.recs.counts = C.cast(self.counts, C.POINTER(LaneCounts))
        self.recs.sb_index = self._put(fr.sb_index)
        self.recs.lf = self._put(fr.lf)
        self.w, self.h, self.bpp = int(fr.hdr["w"]), int(fr.hdr["h"]), int(fr.hdr["bpp"])
        self.invisible = bool(fr.hdr["invisible"])

    def _put(self, data: bytes) -> int:
        n = max(len(data), 16)
        if self.pinned:
            p = self.lib.vp9cuda_host_alloc(n)
            if not p:
                raise MemoryError("vp9cuda_host_alloc")
            self._allocs.append(p)
            C.memmove(p, data, len(data))
            return p
        buf = C.create_string_buffer(data, n)
        self._keep.append(buf)
        return C.addressof(buf)

    def free(self) -> None:
        for p in self._allocs:
            self.lib.vp9cuda_host_free(p)
        self._allocs = []


class OutFrame:
    """Pinned host planes for the visible crop of one picture (framemd5 layout)."""

    def __init__(self, lib: C.CDLL, w: int, h: int, bpp: int, ss: int = 1):
        self.lib = lib
        bps = 2 if bpp > 8 else 1
        ch = ((h + ss) >> ss, (w + ss) >> ss)             # ss = 1: 4:2:0, ss = 0: 4:4:4
        self.shape = [(h, w), ch, ch]
        self.bps = bps
        self.sizes = [r * c * bps for r, c in self.shape]
        self.base = lib.vp9cuda_host_alloc(sum(self.sizes))
        if not self.base:
            raise MemoryError("vp9cuda_host_alloc")
        self.planes = (C.c_void_p * 3)()
        self.strides = (C.c_ssize_t * 3)()
        off = 0
        for i in range(3):
            self.planes[i] = self.base + off
            self.strides[i] = self.shape[i][1] * bps
            off += self.sizes[i]
        self.nbytes = off

    def bytes(self) -> bytes:
        return C.string_at(self.base, self.nbytes)

    def free(self) -> None:
        if self.base:
            self.lib.vp9cuda_host_free(self.base)
            self.base = None


class Context:
    def __init__(self, device: int = 0, frames_in_flight: int = 2, max_w: int = 0, max_h: int = 0, max_bpp: int = 12):
        self.lib = load()
        self.ctx = C.c_void_p()
        check(self.lib.vp9cuda_create(C.byref(self.ctx), device, max_w, max_h, max_bpp, frames_in_flight), "vp9cuda_create")

    def stat(self, key: str) -> int:
        v = C.c_uint64()
        check(self.lib.vp9cuda_get_stat(self.ctx, key.encode(), C.byref(v)), f"get_stat({key})")
        return v.value

    def close(self) -> None:
        if self.ctx:
            self.lib.vp9cuda_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def decode(self, hf: HostFrame, out: Optional[OutFrame]) -> int:
        t = C.c_uint64()
        planes = out.planes if out is not None else None
        strides = out.strides if out is not None else None
        check(self.lib.vp9cuda_decode_frame(self.ctx, C.addressof(hf.hdr_buf), C.byref(hf.recs), planes, strides,
                                            C.byref(t)), "vp9cuda_decode_frame")
        return t.value

    def stage(self, hf: HostFrame) -> int:
        h = C.c_int()
        check(self.lib.vp9cuda_stage_frame(self.ctx, C.addressof(hf.hdr_buf), C.byref(hf.recs), C.byref(h)),
              "vp9cuda_stage_frame")
        return h.value

    def run_staged(self, handle: int, out: Optional[OutFrame] = None) -> int:
        t = C.c_uint64()
        planes = out.planes if out is not None else None
        strides = out.strides if out is not None else None
        check(self.lib.vp9cuda_run_staged(self.ctx, handle, planes, strides, C.byref(t)), "vp9cuda_run_staged")
        return t.value

    def begin(self, fr: "v9r.Frame") -> "FrameRing":
        """vp9cuda_begin_frame for a dumped frame; returns the frame ring (its frame_id names the frame)"""
        ring = FrameRing()
        hdr = (C.c_char * v9r.HDR_DTYPE.itemsize).from_buffer_copy(fr.hdr.tobytes())
        check(self.lib.vp9cuda_begin_frame(self.ctx, C.addressof(hdr), C.byref(ring)), "vp9cuda_begin_frame")
        return ring

    def produce(self, fr: "v9r.Frame", ring: "FrameRing", splits: int = 1) -> None:
        """what the libavcodec glue does between begin_frame and finish_frame: write the lanes' records into
        the pinned rings and submit them in `splits` growing portions per lane"""
        ceb = v9r.COEF32_DTYPE.itemsize if fr.hdr["bpp"] > 8 else v9r.COEF16_DTYPE.itemsize
        for i, l in enumerate(fr.lanes):
            lr = LaneRing()
            check(self.lib.vp9cuda_tile_ring(self.ctx, ring.frame_id, i, C.byref(lr)), "vp9cuda_tile_ring")
            assert lr.coef_base == l.coef_base, "the dump and the library place the lanes alike"
            tot = {k: int(l.counts[k]) for k in ("blk", "tx", "itx", "coef")}
            assert tot["blk"] <= lr.blk_cap and tot["tx"] <= lr.tx_cap and tot["itx"] <= lr.itx_cap and tot["coef"] <= lr.coef_cap
            done = dict.fromkeys(tot, 0)
            for k in range(1, splits + 1):
                upto = {n: tot[n] * k // splits for n in tot}
                for name, ptr_, data, sz in (("blk", lr.blk, l.blk, v9r.BLK_SIZE), ("tx", lr.tx, l.tx, v9r.TX_SIZE),
                                             ("itx", lr.itx, l.itx, v9r.ITX_SIZE), ("coef", lr.coef, l.coef, ceb)):
                    a, b = done[name] * sz, upto[name] * sz
                    if b > a:
                        C.memmove(ptr_ + a, data[a:b], b - a)
                done = upto
                c = LaneCounts(upto["blk"], upto["tx"], upto["itx"], upto["coef"])
                check(self.lib.vp9cuda_submit_tile(self.ctx, ring.frame_id, i, C.byref(c)), "vp9cuda_submit_tile")
        C.memmove(ring.sb_index, fr.sb_index, len(fr.sb_index))
        C.memmove(ring.lf, fr.lf, len(fr.lf))

    def finish(self, ring: "FrameRing", out: Optional[OutFrame]) -> int:
        t = C.c_uint64()
        planes = out.planes if out is not None else None
        strides = out.strides if out is not None else None
        check(self.lib.vp9cuda_finish_frame(self.ctx, ring.frame_id, planes, strides, C.byref(t)), "vp9cuda_finish_frame")
        return t.value

    def abort(self, ring: "FrameRing") -> None:
        check(self.lib.vp9cuda_abort_frame(self.ctx, ring.frame_id), "vp9cuda_abort_frame")

    def wait(self, ticket: int) -> None:
        check(self.lib.vp9cuda_wait(self.ctx, ticket), "vp9cuda_wait")

    def option(self, key: str, value: int = 1) -> None:
        check(self.lib.vp9cuda_set_option(self.ctx, key.encode(), value), f"set_option({key})")

    def elapsed_ms(self, a: int, b: int) -> float:
        ms = C.c_double()
        check(self.lib.vp9cuda_elapsed_ms(self.ctx, a, b, C.byref(ms)), "vp9cuda_elapsed_ms")
        return ms.value

    def phase_clocks(self, n: int = 8) -> list:
        out = (C.c_uint64 * n)()
        check(self.lib.vp9cuda_phase_clocks(self.ctx, out, n), "vp9cuda_phase_clocks")
        return list(out)

    def map_frame(self, slot: int = -1) -> "MappedFrame":
        m = MappedFrame()
        check(self.lib.vp9cuda_map_frame(self.ctx, slot, C.byref(m)), "vp9cuda_map_frame")
        return m

    def unmap_frame(self, m: "MappedFrame") -> None:
        check(self.lib.vp9cuda_unmap_frame(self.ctx, m.planes[0]), "vp9cuda_unmap_frame")

    def checksum(self, slot: int = -1) -> int:
        s = C.c_uint64()
        check(self.lib.vp9cuda_frame_checksum(self.ctx, slot, C.byref(s)), "vp9cuda_frame_checksum")
        return s.value

    def kernel_stats(self) -> dict:
        n = 16
        names = (C.c_char_p * n)()
        ms = (C.c_double * n)()
        launches = (C.c_uint64 * n)()
        nbytes = (C.c_uint64 * n)()
        k = self.lib.vp9cuda_kernel_stats(self.ctx, names, ms, launches, nbytes, n)
        check(k, "vp9cuda_kernel_stats")
        return {names[i].decode(): {"ms": ms[i], "launches": int(launches[i]), "bytes": int(nbytes[i])}
                for i in range(k)}


def load_stream(lib: C.CDLL, path: Path, pinned: bool = True) -> List[HostFrame]:
    return [HostFrame(lib, fr, pinned) for fr in v9r.read_v9r(path)]
