"""BASELINE.json configs[4] / SURVEY.md §8d(5): stress sweep — N in {1, 2, 5} M Gaussians at 1920x1080, a 64-camera
batch; tile-binning/sort and raster throughput vs Gaussian count. One GPU, CUDA events, synthetic scene.

Per N it reports: intersections I, the all-view eval render (RGB+depth+alpha, `render_views`) in views/s, the
training step (fwd+bwd) in views/s, and the binning / compositing stages in M intersections/s.
Prints one JSON line per N (and writes them to gpurun_out/stress_sweep.jsonl when that directory exists).
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussctrl_exp_b200 import _lib, render, scene  # noqa: E402
from gaussctrl_exp_b200.model import GaussianModel  # noqa: E402


def timed(fn, reps):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps):
        fn(i)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=float, nargs="+", default=[1e6, 2e6, 5e6])
    ap.add_argument("--cams", type=int, default=64)
    ap.add_argument("--train-steps", type=int, default=32)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    cams = scene.make_cameras(args.cams, 0, intrinsics=scene.HD_INTRINSICS)
    H, W = cams[0].H, cams[0].W
    v = (torch.randn(H, W, 3, generator=torch.Generator().manual_seed(1234)) / (H * W)).to(dev)
    L = _lib.lib()
    lines = []
    for n in args.n:
        N = int(n)
        torch.cuda.empty_cache()
        m = GaussianModel(scene.make_scene(N, 0), background_color=torch.tensor([0.1, 0.2, 0.3])).to(dev)
        # ---- all-view eval render ----
        m.eval()
        m.render_views(cams[:4])                       # warm-up (capacity hints, streams, allocator)
        torch.cuda.synchronize()
        ms_eval = timed(lambda i: m.render_views(cams), 2) / len(cams)
        out = m.get_outputs_for_camera(cams[0])
        I = int(m.view_state.num_intersects) if m.view_state.num_intersects else None
        # ---- training step ----
        m.train()

        def step(i):
            for q in m.parameters():
                q.grad = None
            (m(cams[i % len(cams)])["rgb"] * v).sum().backward()
        for i in range(4):
            step(i)
        ms_train = timed(step, args.train_steps)
        I_train = int(m.view_state.num_intersects)
        # ---- stages on camera 0 (device-resident) ----
        cam = render.gs_camera(cams[0])
        p = [t.detach().float().contiguous() for t in (m.means, m.scales, m.quats, m.opacities, m.features_dc,
                                                       m.features_rest)]
        bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
        with torch.no_grad(), torch.cuda.device(dev):
            res = render._forward_impl(*p, cam, True, 3, bg, strict=True)
            xys, depths, conics, rgbs, opac, radii, hit, ids_sorted, bins, out_img, fT, fidx, I0 = res[:13]
            T = cam.tiles_x * cam.tiles_y
            cap = ids_sorted.numel()
            scratch = torch.empty(L.gs_bin_fast_scratch_bytes(N, cap, T), dtype=torch.uint8, device=dev)
            total = torch.empty(1, dtype=torch.int32, device=dev)
            st = _lib.stream_ptr()

            def bin_(i):
                _lib.check(L.gs_bin_fast(N, cap, xys.data_ptr(), depths.data_ptr(), radii.data_ptr(), hit.data_ptr(),
                                         cam.tiles_x, cam.tiles_y, ids_sorted.data_ptr(), bins.data_ptr(),
                                         total.data_ptr(), scratch.data_ptr(), scratch.numel(), st), "bin")

            def rast(i):
                _lib.check(L.gs_raster_fwd(3, H, W, cam.tiles_x, cam.tiles_y, ids_sorted.data_ptr(), bins.data_ptr(),
                                           xys.data_ptr(), conics.data_ptr(), rgbs.data_ptr(), opac.data_ptr(),
                                           bg.data_ptr(), out_img.data_ptr(), fT.data_ptr(), fidx.data_ptr(), st), "rf")
            bin_(0); rast(0)
            ms_bin = timed(bin_, 10)
            ms_rast = timed(rast, 10)
        line = {"workload": f"stress: {N} Gaussians, {W}x{H}, {len(cams)} cameras, SH deg 3", "n_gaussians": N,
                "intersections_cam0": int(I0), "intersections_last_train_view": I_train,
                "eval_render_views_per_s": 1e3 / ms_eval, "train_fwd_bwd_views_per_s": 1e3 / ms_train,
                "bin_fast_us": ms_bin * 1e3, "bin_fast_Mintersections_per_s": I0 / ms_bin / 1e3,
                "raster_fwd_us": ms_rast * 1e3, "raster_fwd_Mintersections_per_s": I0 / ms_rast / 1e3,
                "peak_mem_GB": torch.cuda.max_memory_allocated() / 2**30}
        print(json.dumps(line), flush=True)
        lines.append(line)
        del m, res, xys, depths, conics, rgbs, opac, radii, hit, ids_sorted, bins, out_img, fT, fidx, scratch, p
    if os.path.isdir("gpurun_out"):
        with open("gpurun_out/stress_sweep.jsonl", "w") as f:
            for ln in lines:
                f.write(json.dumps(ln) + "\n")


if __name__ == "__main__":
    main()
