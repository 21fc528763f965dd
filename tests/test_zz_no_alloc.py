"""GPU: cb_chamfer_static, the Chamfer entry point that allocates nothing.

Who this is for: code that embeds the library but may not call cudaMalloc — a plugin inside an
application whose own memory pool already owns the card, a serving process with a fixed per-model
memory budget, a real-time loop that cannot take cudaMalloc's device-wide synchronisation.  Such a
caller cannot even make a cb_ctx (cb_create allocates the workspace arena), so this entry point takes
none and keeps its scratch in a fixed ~3 MB array inside the library's device image.

What is checked: called through the C ABI with caller-owned inputs and outputs only, it returns what
cb_chamfer returns, on two calls in a row with different inputs — the second call finds the first
call's keys and partial sums still in the shared array and must not be disturbed by them — in one and
two directions, with and without a pose; and it refuses clouds beyond CB_CHAMFER_STATIC_MAX_POINTS with
a readable message instead of overrunning the array.  The file name sorts last on purpose: under
`pytest -x` a failure here cannot hide another test."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

RTOL = 1e-5  # the Chamfer contract of tests/test_gpu_chamfer_retrieval.py; the two paths share kernels
MAX_POINTS = 65536  # CB_CHAMFER_STATIC_MAX_POINTS in include/corsair_b200.h


def _pose(rng):
    q, _ = np.linalg.qr(rng.standard_normal((3, 3)))
    if np.linalg.det(q) < 0:
        q[:, 0] = -q[:, 0]
    T = np.eye(4)
    T[:3, :3] = q
    T[:3, 3] = 0.1 * rng.standard_normal(3)
    return T


def _static_call(lib, a, b, T, two_dir, out, nn_idx, stream):
    """The call a cudaMalloc-less host would make: raw device pointers, no context object."""
    err = ctypes.create_string_buffer(256)
    rc = lib.cb_chamfer_static(ctypes.c_void_p(a.data_ptr()), a.shape[0], ctypes.c_void_p(b.data_ptr()), b.shape[0],
                               ctypes.c_void_p(T.data_ptr() if T is not None else 0), int(two_dir),
                               ctypes.c_void_p(out.data_ptr()),
                               ctypes.c_void_p(nn_idx.data_ptr() if nn_idx is not None else 0), err, len(err),
                               ctypes.c_void_p(stream))
    return rc, err.value.decode()


def test_static_scratch_chamfer_matches_arena_path_twice(ctx):
    from corsair_b200._capi import ptr
    rng = np.random.default_rng(2026)
    dev = ctx.device
    stream = torch.cuda.current_stream(dev).cuda_stream
    # two different problems back to back; the second is smaller than the first so stale keys and
    # partial sums of the first call lie beyond and inside the region the second one uses
    cases = [(9000, 7000, True, True), (3001, 4100, False, True), (513, 2049, True, False), (1, 1, False, False)]
    with torch.cuda.device(dev):
        for na, nb, with_pose, two_dir in cases:
            a = torch.from_numpy(rng.random((na, 3), dtype=np.float32) - 0.5).to(dev)
            b = torch.from_numpy(rng.random((nb, 3), dtype=np.float32) - 0.5).to(dev)
            T = torch.from_numpy(_pose(rng)).to(dev) if with_pose else None
            want = torch.full((1,), -1.0, dtype=torch.float64, device=dev)
            want_idx = torch.full((na,), -1, dtype=torch.int32, device=dev)
            ctx.call("cb_chamfer", ptr(a), na, ptr(b), nb, ptr(T), int(two_dir), ptr(want), ptr(want_idx), ctx.stream)
            got = torch.full((1,), -2.0, dtype=torch.float64, device=dev)
            got_idx = torch.full((na,), -2, dtype=torch.int32, device=dev)
            rc, msg = _static_call(ctx.lib, a, b, T, two_dir, got, got_idx, stream)
            assert rc == 0, msg
            torch.cuda.synchronize(dev)
            w, g = want.item(), got.item()
            assert w > 0 and abs(g - w) <= RTOL * w, (na, nb, with_pose, two_dir, g, w)
            assert torch.equal(got_idx, want_idx)


def test_static_scratch_chamfer_rejects_what_does_not_fit(ctx):
    dev = ctx.device
    with torch.cuda.device(dev):
        big = torch.zeros((MAX_POINTS + 1, 3), dtype=torch.float32, device=dev)
        small = torch.zeros((8, 3), dtype=torch.float32, device=dev)
        out = torch.full((1,), 7.0, dtype=torch.float64, device=dev)
        for a, b in ((big, small), (small, big)):
            rc, msg = _static_call(ctx.lib, a, b, None, 1, out, None, 0)
            assert rc == -4 and str(MAX_POINTS) in msg, (rc, msg)
        rc, msg = _static_call(ctx.lib, small[:0], small, None, 0, out, None, 0)
        assert rc == -1 and msg
        torch.cuda.synchronize(dev)
        assert out.item() == 7.0  # a refused call launches nothing
        # exactly at the limit it runs
        edge = torch.rand((MAX_POINTS, 3), dtype=torch.float32, device=dev)
        rc, msg = _static_call(ctx.lib, edge, edge, None, 1, out, None, 0)
        assert rc == 0, msg
        torch.cuda.synchronize(dev)
        assert out.item() == 0.0  # a cloud against itself
