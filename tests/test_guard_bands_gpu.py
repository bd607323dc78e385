"""Out-of-bounds writes, checked by hand (compute-sanitizer is closed on this GPU pool: profiles/r02_sanitizer_unavailable.txt).

Every buffer the library WRITES — the per-view workspace, the binning scratch, the image outputs, the compositor-gradient
block, the flat leaf gradients, the deterministic backward's scratch — is carved out of a larger allocation with a 64 KB
band of a known bit pattern on either side, sized EXACTLY as the ABI's size functions say. After gs_view_fwd / gs_view_bwd
/ gs_view_bwd_det the bands must be untouched. Cases: ragged image sizes, one and several views per call, both binning
pipelines, a capacity that is too small (the overflow protocol: nothing may be written past I_cap) and a capacity of
zero."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
BAND = 64 * 1024
PATTERN = 0x5A


class Guarded:
    """`nbytes` usable bytes (256-byte aligned) between two bands of PATTERN."""

    def __init__(self, nbytes, dev):
        self.n = int(nbytes)
        pad = (256 - self.n % 256) % 256
        self.buf = torch.full((BAND + self.n + pad + BAND,), PATTERN, dtype=torch.uint8, device=dev)
        self.ptr = self.buf.data_ptr() + BAND
        assert self.ptr % 256 == 0

    def view(self, dtype):
        return self.buf[BAND:BAND + self.n].view(dtype)

    def intact(self):
        lo, hi = self.buf[:BAND], self.buf[BAND + self.n:]
        return bool((lo == PATTERN).all().item()) and bool((hi == PATTERN).all().item())


def _run(dev, n, W, H, views, cap, impl, det=False):
    from gaussctrl_exp_b200 import _lib, render, scene
    L = _lib.lib()
    p = {k: v.to(dev).contiguous() for k, v in scene.make_scene(n, 11).items()}
    cams = scene.make_cameras(views + 1, 5, intrinsics=dict(fx=0.9 * W, fy=0.9 * W, cx=W / 2 - 0.3, cy=H / 2 + 0.2, W=W, H=H))
    gcs = [render.gs_camera(c) for c in cams[:views]]
    cam_arr = (_lib.GsCamera * views)(*gcs)
    T_view = gcs[0].tiles_x * gcs[0].tiles_y
    lay = _lib.GsViewLayout()
    assert L.gs_view_layout(n, views, cap, T_view, ctypes.byref(lay)) == 0
    ws = Guarded(lay.bytes, dev)
    scratch = Guarded(L.gs_view_scratch_bytes(n, views, cap, T_view), dev)
    hw = views * H * W
    out, outc, alpha, depth = (Guarded(4 * k * hw, dev) for k in (3, 3, 1, 1))
    fT, fidx = Guarded(4 * hw, dev), Guarded(4 * hw, dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    pinned = torch.zeros(1, dtype=torch.int32).pin_memory()
    ev = torch.cuda.Event()
    ev.record()
    prev = L.gs_set_bin_impl(0 if impl == "bucket" else 1)
    try:
        rc = L.gs_view_fwd(n, views, p["means"].data_ptr(), p["scales"].data_ptr(), p["quats"].data_ptr(),
                           p["opacities"].data_ptr(), p["features_dc"].data_ptr(), p["features_rest"].data_ptr(), None, 1, 3,
                           ctypes.addressof(cam_arr), bg.data_ptr(), cap, ws.ptr, ws.n, scratch.ptr, scratch.n, out.ptr,
                           outc.ptr, alpha.ptr, depth.ptr, fT.ptr, fidx.ptr, pinned.data_ptr(), ev.cuda_event,
                           _lib.stream_ptr())
    finally:
        L.gs_set_bin_impl(prev)
    assert rc == 0
    torch.cuda.synchronize()
    I = int(pinned[0])
    guards = dict(ws=ws, scratch=scratch, out=out, out_clamped=outc, alpha=alpha, depth=depth, final_T=fT, final_idx=fidx)
    if views == 1 and I <= cap:
        Np = (n + 3) // 4 * 4
        gbuf, flat = Guarded(4 * (9 * Np + 64), dev), Guarded(4 * 59 * Np, dev)
        v_out = torch.randn(H, W, 3, device=dev) / (H * W)
        args = (n, p["means"].data_ptr(), p["scales"].data_ptr(), p["quats"].data_ptr(), p["opacities"].data_ptr(), 1, 3,
                ctypes.addressof(gcs[0]), bg.data_ptr(), ws.ptr, ws.n, cap, fT.ptr, fidx.ptr, v_out.data_ptr(), None, out.ptr,
                gbuf.ptr, flat.ptr)
        if det:
            ds = Guarded(L.gs_view_bwd_det_scratch_bytes(n, cap), dev)
            rc = L.gs_view_bwd_det(*args, ds.ptr, ds.n, _lib.stream_ptr())
            guards["det_scratch"] = ds
        else:
            rc = L.gs_view_bwd(*args, _lib.stream_ptr())
        assert rc == 0
        torch.cuda.synchronize()
        guards.update(gbuf=gbuf, flat=flat)
        assert torch.isfinite(flat.view(torch.float32)).all()
    bad = [k for k, g in guards.items() if not g.intact()]
    assert not bad, f"guard bands overwritten: {bad} (n={n} {W}x{H} views={views} cap={cap} I={I} impl={impl})"
    return I


@pytest.mark.parametrize("impl", ["lsd", "bucket"])
@pytest.mark.parametrize("n,W,H,views", [(20000, 160, 128, 1), (5003, 97, 61, 1), (3000, 33, 17, 3), (12000, 250, 3, 2),
                                          (40000, 512, 512, 1)])
def test_no_write_outside_the_declared_buffers(cuda_dev, n, W, H, views, impl):
    I = _run(cuda_dev, n, W, H, views, cap=max(8 * n * views, 1024), impl=impl)
    assert I > 0
    # exact fit, one entry short (overflow: nothing may be written past the capacity), and no capacity at all
    assert _run(cuda_dev, n, W, H, views, cap=I, impl=impl) == I
    assert _run(cuda_dev, n, W, H, views, cap=I - 1, impl=impl) == I
    assert _run(cuda_dev, n, W, H, views, cap=max(I // 3, 1), impl=impl) == I
    assert _run(cuda_dev, n, W, H, views, cap=0, impl=impl) == I


def test_no_write_outside_the_declared_buffers_deterministic_backward(cuda_dev):
    I = _run(cuda_dev, 20000, 160, 128, 1, cap=200000, impl="lsd", det=True)
    assert _run(cuda_dev, 20000, 160, 128, 1, cap=I, impl="lsd", det=True) == I
