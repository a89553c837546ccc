"""vp9cuda_checksum_noalloc (include/vp9cuda.h): the allocation-free variant of the frame checksum, whose
scratch is a static __device__ array.  No counterpart in the reference, so the
check is against the existing entry point: the value must equal vp9cuda_frame_checksum's bit for bit, on calls
in a row that share the scratch array.  Collected last on purpose (pytest -x)."""
import ctypes as C
import errno

import pytest

from ffmpeg_hybrid_b200 import configs, parity

pytestmark = pytest.mark.gpu


def _noalloc(lib, m, out, w=None, h=None):
    planes = (C.c_void_p * 3)(*[m.planes[p] for p in range(3)])
    strides = (C.c_ssize_t * 3)(*[m.strides[p] for p in range(3)])
    return lib.vp9cuda_checksum_noalloc(0, planes, strides, w or m.w, h or m.h, m.bpp, m.ss_h, m.ss_v,
                                        C.c_void_p(out.data_ptr()), None)


@pytest.mark.parametrize("name", ["c3_small", "hbd12", "p1_444"])
def test_noalloc_checksum_equals_frame_checksum(name, workdir):
    import torch
    from ffmpeg_hybrid_b200 import binding
    lib = binding.load()
    ivf = parity.write_stream(configs.small_variants()[name], workdir / f"na_{name}.ivf")
    parity.dump_records(ivf, workdir / f"na_{name}.v9r")
    frames = binding.load_stream(lib, workdir / f"na_{name}.v9r", pinned=True)
    ctx = binding.Context(0)
    out = torch.full((2,), -1, dtype=torch.int64, device="cuda:0")      # device memory of the caller; [1] is a guard
    torch.cuda.synchronize()
    n = 0
    for f in frames:
        ctx.wait(ctx.decode(f, None))
        want = ctx.checksum(-1)
        m = ctx.map_frame(-1)
        allocs = ctx.stat("allocs")
        for _ in range(2):                                                # twice in a row on the shared scratch
            out.fill_(-1)
            assert _noalloc(lib, m, out) == 0
            torch.cuda.synchronize()
            got = out.cpu()
            assert int(got[0]) & (2**64 - 1) == want
            assert int(got[1]) == -1
            n += 1
        assert ctx.stat("allocs") == allocs                               # the library's own allocation counter did not move
        ctx.unmap_frame(m)
    assert n == 2 * len(frames) and n >= 4
    # beyond the static scratch: a clear error, nothing launched
    m = ctx.map_frame(-1)
    assert _noalloc(lib, m, out, h=2 * 16384) == -errno.E2BIG
    assert lib.vp9cuda_checksum_noalloc(0, None, None, 64, 64, 8, 1, 1, C.c_void_p(out.data_ptr()), None) == -errno.EINVAL
    # and the next call still works
    out.fill_(-1)
    assert _noalloc(lib, m, out) == 0
    torch.cuda.synchronize()
    assert int(out.cpu()[0]) & (2**64 - 1) == ctx.checksum(-1)
    ctx.unmap_frame(m)
    ctx.close()
    for f in frames:
        f.free()
