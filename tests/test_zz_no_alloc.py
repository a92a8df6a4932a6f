"""The allocation-free entry point isogs_depth_outlier_threshold_noalloc (no counterpart in the reference; added
for callers that cannot call cudaMalloc): its scratch is a file-scope device
array inside the library.  Called through the C ABI and compared with isogs_depth_outlier_threshold (caller
scratch) -- the two run the same kernels, so the threshold is required bit-equal -- on two calls in a row whose
image sizes differ, so that contents left in the shared array by the first call would show in the second.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(W, H, seed):
    rng = np.random.RandomState(seed)
    gt = rng.uniform(0.5, 4.0, (H, W)).astype(np.float32)
    gt[rng.rand(H, W) < 0.1] = 0.0                                  # invalid depth
    rd = (gt + rng.normal(0, 0.05, (H, W))).astype(np.float32)
    rd[rng.rand(H, W) < 0.02] += 1.5                                # outliers
    return torch.from_numpy(rd).cuda(), torch.from_numpy(gt).cuda()


def test_noalloc_outlier_threshold_matches_the_caller_scratch_entry_twice_in_a_row(pkg):
    L, lb = pkg.lib(), pkg._lib
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # inputs, outputs and the existing entry's scratch are all created BEFORE the calls under test
    cases = []
    for W, H, seed in ((640, 480, 1), (208, 144, 2)):
        rd, gt = _inputs(W, H, seed)
        P = W * H
        scratch = torch.empty((2 * P,), dtype=torch.float32, device="cuda")
        tmp_b = max(lb.temp_bytes(0, P)) + (1 << 20)
        tmp = torch.empty((tmp_b,), dtype=torch.uint8, device="cuda")
        cases.append((pkg.make_camera(W, H, 500.0, 500.0, W / 2, H / 2), rd, gt, scratch, tmp, tmp_b,
                      torch.zeros((1,), dtype=torch.float32, device="cuda"),
                      torch.zeros((1,), dtype=torch.float32, device="cuda")))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for cam, rd, gt, scratch, tmp, tmp_b, want, got in cases:
        lb.check(L.isogs_depth_outlier_threshold(C.byref(cam), rd.data_ptr(), gt.data_ptr(), 10.0, scratch.data_ptr(),
                                                 tmp.data_ptr(), tmp_b, want.data_ptr(), st))
        lb.check(L.isogs_depth_outlier_threshold_noalloc(C.byref(cam), rd.data_ptr(), gt.data_ptr(), 10.0,
                                                         got.data_ptr(), st))
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    for cam, rd, gt, scratch, tmp, tmp_b, want, got in cases:
        w, g = float(want[0]), float(got[0])
        err = (torch.abs(gt - rd) * (gt > 0)).flatten()
        ref = float(np.float32(10.0) * np.float32(float(torch.sort(err).values[(err.numel() - 1) // 2])))
        assert w > 0 and abs(w - ref) <= 1e-6 * ref, (w, ref)      # the existing entry: 10 x lower median
        assert g == w, (g, w)                                      # same kernels, different scratch: bit-equal


def test_noalloc_outlier_threshold_rejects_images_beyond_the_built_in_scratch(pkg):
    L, lb = pkg.lib(), pkg._lib
    cam = pkg.make_camera(2048, 1025, 500.0, 500.0, 1024.0, 512.0)   # 2 099 200 > ISOGS_NOALLOC_MAX_PIXELS = 2^21
    x = torch.zeros((4,), dtype=torch.float32, device="cuda")        # never dereferenced: rejected on the host
    rc = L.isogs_depth_outlier_threshold_noalloc(C.byref(cam), x.data_ptr(), x.data_ptr(), 10.0, x.data_ptr(), None)
    assert rc == lb.E_ARG
    assert b"exceed" in L.isogs_last_error()
