"""ekf_pose_moments_noalloc (include/ekf_b200.h): the allocation-free variant of ekf_batch_pose_moments, whose per-CTA
partial sums live in a file-scope device array. No counterpart in the reference; collected last on purpose (the file
name sorts after every other test file), and it has to pass alone in a fresh process as well as after the others.

Checked through the C-ABI against the existing entry point (same kernels, same summation order: bit-exact, which is
inside the 1e-9 bar of conftest.TOL), on two calls in a row per size, with sizes that use 1, 274 and all 592 scratch
records in an order that leaves stale records behind from the previous call.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import TOL


def _filters(eb, torch, B):
    from slam_duckietown_b200 import workloads as wl
    w = wl.make_batched(B, 8, 2, steps=2)
    f = eb.BatchedEKF(B, 8)
    f.set_noise(w.motion_var, w.meas_var)
    f.init(0.1)
    for t in range(2):
        f.step(*(torch.from_numpy(np.ascontiguousarray(a[t])).to(f.dev) for a in (w.ang, w.lin, w.slot, w.rng, w.brg)))
    return f


@pytest.mark.gpu
def test_noalloc_pose_moments_matches_existing_entry_point():
    import torch
    assert torch.cuda.is_available()
    from slam_duckietown_b200 import capi, ekf_bindings as eb
    lib = capi.lib()                                   # raises when the CUDA library is missing: no fallback
    dev = torch.cuda.current_device()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for B in (70001, 37, 400000, 37):
        f = _filters(eb, torch, B)
        want = f.pose_moments()                        # existing entry point (handle-owned cudaMalloc'ed scratch)
        got = [torch.full((12,), float("nan"), dtype=torch.float64, device=f.dev) for _ in range(2)]
        stride = int(f.mu.stride(0))
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        for out in got:                                # two calls in a row, same stream
            capi.call("ekf_pose_moments_noalloc", C.c_void_p(f.mu.data_ptr()), stride, B,
                      C.c_void_p(out.data_ptr()), dev, stream)
        torch.cuda.synchronize()
        if B != 70001:                                 # the very first call may load the module lazily
            assert torch.cuda.mem_get_info()[0] == free0, "the no-alloc entry point changed free device memory"
        w = want.cpu().numpy()
        for out in got:
            g = out.cpu().numpy()
            assert np.all(np.isfinite(g))
            assert np.max(np.abs(g - w)) <= TOL * max(1.0, np.abs(w).max())
            assert np.array_equal(g, w), "same kernels and summation order: expected identical bits"
        # a compact [B][3] pose copy (another stride) gives the same sums
        pose = f.mu[:, :3].contiguous()
        out = torch.empty(12, dtype=torch.float64, device=f.dev)
        capi.call("ekf_pose_moments_noalloc", C.c_void_p(pose.data_ptr()), 3, B, C.c_void_p(out.data_ptr()), dev,
                  stream)
        assert np.array_equal(out.cpu().numpy(), w)
        del f


@pytest.mark.gpu
def test_noalloc_pose_moments_rejects_sizes_beyond_the_static_scratch():
    import torch
    assert torch.cuda.is_available()
    from slam_duckietown_b200 import capi
    lib = capi.lib()
    dev = torch.cuda.current_device()
    buf = torch.zeros(16, dtype=torch.float64, device="cuda")
    limit = 4194304                                    # EKF_POSE_MOMENTS_NOALLOC_MAX_FILTERS
    p = C.c_void_p(buf.data_ptr())
    for bad_batch in (limit + 1, 0, -5):
        rc = lib.ekf_pose_moments_noalloc(p, 3, bad_batch, p, dev, None)
        assert rc < 0
    rc = lib.ekf_pose_moments_noalloc(p, 3, limit + 1, p, dev, None)
    assert rc < 0 and b"EKF_POSE_MOMENTS_NOALLOC_MAX_FILTERS" in lib.ekf_last_error()
    assert lib.ekf_pose_moments_noalloc(None, 3, 4, p, dev, None) < 0
    assert lib.ekf_pose_moments_noalloc(p, 2, 4, p, dev, None) < 0
    torch.cuda.synchronize()                           # nothing was launched, nothing faulted
    assert float(buf.abs().sum()) == 0.0
