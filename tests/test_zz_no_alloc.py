"""ddpb_conv_wgrad_static -- the weight-gradient entry point that allocates nothing.

Some callers of this library are not allowed to allocate device memory when they need a
weight gradient: a loop that cannot afford cudaMalloc's device-wide synchronisation, a
process whose memory budget was given away at start-up, an embedding application that
forbids allocator calls.  For them the split-reduction scratch sits in a fixed 16 MiB array
inside the library and the caller hands over inputs and outputs only.

The test drives the entry point through the C ABI (ctypes on libddpb200.so, raw device
pointers) exactly as such a caller would and checks it against the ordinary
ddpb_conv_wgrad, twice in a row with different shapes, because the array is shared and
keeps the previous call's contents.  It is collected last (file name) so that a run which
stops at the first failure has already exercised everything else.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from ddpb200.ops import _native as nat

    lib = nat.load()
    lib.ddpb_conv_wgrad_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 8 + [C.c_void_p]
    lib.ddpb_conv_wgrad_static.restype = C.c_int
    lib.ddpb_conv_wgrad_static_max_elems.argtypes = []
    lib.ddpb_conv_wgrad_static_max_elems.restype = C.c_longlong
    lib.ddpb_conv_wgrad.argtypes = [C.c_void_p] * 4 + [C.c_longlong] + [C.c_int] * 8 + [C.c_void_p]
    lib.ddpb_conv_wgrad.restype = C.c_int
    lib.ddpb_conv_wgrad_scratch_elems.argtypes = [C.c_int] * 6
    lib.ddpb_conv_wgrad_scratch_elems.restype = C.c_longlong
    lib.ddpb_conv_wgrad_scratch_elems_ld.argtypes = [C.c_int] * 8
    lib.ddpb_conv_wgrad_scratch_elems_ld.restype = C.c_longlong
    lib.ddpb_last_error.restype = C.c_char_p
    return lib


def _rel_err(a, b):
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


def test_static_scratch_entry_point_matches_regular_one():
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    torch.manual_seed(0)
    stream = torch.cuda.current_stream().cuda_stream
    cap = lib.ddpb_conv_wgrad_static_max_elems()
    assert cap == 4 * 1024 * 1024
    # two consecutive calls: a big split reduction first, then a smaller shape that leaves
    # most of the array holding the first call's partial sums
    shapes = [(64, 8, 8, 128, 256), (32, 16, 16, 64, 128)]
    for (N, H, W, Ci, Co) in shapes:
        x = torch.randn(N * H * W, Ci, device=dev).bfloat16()
        dy = torch.randn(N * H * W, Co, device=dev).bfloat16()
        need = lib.ddpb_conv_wgrad_scratch_elems(N, H, W, Ci, Co, 9)
        assert 0 < need <= cap, "shape chosen to exercise the scratch path"
        # existing entry point, caller-provided scratch
        want = torch.empty(Co, 9 * Ci, dtype=torch.float32, device=dev)
        scratch = torch.empty(need, dtype=torch.float32, device=dev)
        rc = lib.ddpb_conv_wgrad(dy.data_ptr(), x.data_ptr(), want.data_ptr(), scratch.data_ptr(), need, N, H, W, Ci,
                                 Co, 9, 9 * Ci, 9 * Ci, stream)
        assert rc == 0, lib.ddpb_last_error()
        # new entry point: inputs and outputs only
        got = torch.full((Co, 9 * Ci), float("nan"), dtype=torch.float32, device=dev)
        rc = lib.ddpb_conv_wgrad_static(dy.data_ptr(), x.data_ptr(), got.data_ptr(), N, H, W, Ci, Co, 9, 9 * Ci,
                                        9 * Ci, stream)
        assert rc == 0, lib.ddpb_last_error()
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
        assert _rel_err(got, want) < 2e-3          # tolerance of tests/test_kernels_gpu.py::test_conv_wgrad
        # and against the plain fp32 restatement of the op, like every other kernel test
        from ddpb200.ops import reference as R
        assert _rel_err(got, R.conv_wgrad(dy, x, N, H, W, Ci, Co)) < 2e-3
    # a shape beyond the built-in array is refused with a message, not run
    N, H, W, Ci, Co = 256, 8, 8, 256, 256
    need = lib.ddpb_conv_wgrad_scratch_elems(N, H, W, Ci, Co, 9)
    assert need > cap
    rc = lib.ddpb_conv_wgrad_static(0, 0, 0, N, H, W, Ci, Co, 9, 9 * Ci, 9 * Ci, stream)
    assert rc != 0
    msg = lib.ddpb_last_error().decode()
    assert "built-in array" in msg and "ddpb_conv_wgrad" in msg
    # a gradient with padded rows is assembled from scratch even when the reduction is not split: this
    # shape needs nothing for a dense dw but 4,718,592 floats (> the array) with a padded row length --
    # the capacity check has to use the real requirement (it once used the dense one and overran)
    N, H, W, Ci, Co = 64, 8, 8, 512, 1024
    assert lib.ddpb_conv_wgrad_scratch_elems(N, H, W, Ci, Co, 9) == 0
    assert lib.ddpb_conv_wgrad_scratch_elems_ld(N, H, W, Ci, Co, 9, 9 * Ci, 9 * Ci + 8) == Co * 9 * Ci > cap
    rc = lib.ddpb_conv_wgrad_static(0, 0, 0, N, H, W, Ci, Co, 9, 9 * Ci, 9 * Ci + 8, stream)
    assert rc != 0 and "built-in array" in lib.ddpb_last_error().decode()
