"""The allocation-free group-mode pack entry point (collected last on purpose).

`splitq_quant_pack_group_noalloc` exists for hosts that cannot hand the library a workspace:
a serving runtime that already owns all device memory, or a callback that is forbidden to
allocate while it enqueues work.  Its scratch is a fixed array inside the library, shared by all
calls and never reset between them, so the test makes TWO calls in a row -- the second on
smaller-magnitude data and a different batch size, which would inherit stale maxima if the
entry point forgot to clear its scratch -- and requires each payload to be byte-identical to
what the regular `splitq_quant_pack` (group mode, caller-supplied workspace) produces.
Everything goes through the C ABI via ctypes; buffers are created before the calls.
"""
import ctypes as C

import pytest
import torch

import splitq
from splitq import _native as N

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _case(seed, B, S, D, magnitude):
    gen = torch.Generator().manual_seed(seed)
    h = (torch.randn(B, S, D, generator=gen) * magnitude).to(DEV)
    imp = torch.rand(B, S, generator=gen).to(DEV)
    ladder = [(S // 4, 2), (S // 4, 4), (S // 8, 8)]
    bits, row_off = splitq.select_bits(imp, ladder, D)
    nbytes = splitq.payload_bytes(S, D, ladder, N.SCALE_GROUP)
    want = splitq.quant_pack(h, bits, row_off, N.SCALE_GROUP, nbytes)
    got = torch.zeros(B, nbytes, dtype=torch.uint8, device=DEV)
    return h, bits, row_off, want, got


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    lib = N.load()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    first = _case(11, 3, 200, 512, magnitude=50.0)
    second = _case(12, 2, 96, 896, magnitude=0.01)
    torch.cuda.synchronize()
    for h, bits, row_off, want, got in (first, second):
        B, S, D = h.shape
        rc = lib.splitq_quant_pack_group_noalloc(C.c_void_p(h.data_ptr()), B, S, D, C.c_void_p(bits.data_ptr()),
                                                 C.c_void_p(row_off.data_ptr()), C.c_void_p(got.data_ptr()),
                                                 got.stride(0), stream)
        assert rc == 0, lib.splitq_last_error().decode()
        torch.cuda.synchronize()
        assert torch.equal(got, want)  # byte-exact, like every other payload test
        assert torch.equal(splitq.dequant(got, S, D), splitq.dequant(want, S, D))


def test_noalloc_python_wrapper_and_generic_shapes():
    h, bits, row_off, want, got = _case(13, 2, 19, 5, magnitude=3.0)  # odd D -> scalar kernels
    assert torch.equal(splitq.quant_pack_group_noalloc(h, bits, row_off, got), want)


def test_noalloc_rejects_batches_beyond_its_scratch():
    lib = N.load()
    B = N.NOALLOC_MAX_WINDOWS + 1
    dummy = torch.zeros(16, device=DEV)
    rc = lib.splitq_quant_pack_group_noalloc(C.c_void_p(dummy.data_ptr()), B, 1, 4, C.c_void_p(dummy.data_ptr()),
                                             C.c_void_p(dummy.data_ptr()), C.c_void_p(dummy.data_ptr()), 64,
                                             C.c_void_p(0))
    assert rc == -1
    msg = lib.splitq_last_error().decode()
    assert "exceeds the built-in scratch" in msg and str(N.NOALLOC_MAX_WINDOWS) in msg
