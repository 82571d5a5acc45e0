#!/usr/bin/env python3
"""Workload for compute-sanitizer (tools/sanitize.sh): every kernel variant at CI size -- the five BASELINE scenes at
64 x 36 x 4 spp with the shipped integrator settings, light_samples 3 (the padded lane group of k_nee), RecursivePT, the
physically corrected mode, work counters, traced paths, a closest-hit batch, tonemap (both entry points), and a two-shard
asynchronous render with a mid-render snapshot (k_combine)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import pathtrace_b200 as pt  # noqa: E402
import raygen  # noqa: E402

W, H, SPP = 64, 36, 4
SCENES = ["three_orbs", "cornell_box", "cornell_box_small_lights", "cornell_box_with_volume", "book2_final"]


def render(scene, **over):
    replace = {k: over.pop(k) for k in list(over) if k in ("batch_samples", "profile", "stats_clamp", "devices",
                                                            "async_render", "sample_offset")}
    cfg = pt.Config.from_json(pt.default_config_dict(width=W, height=H, samples=SPP,
                                                     scene=os.path.join(ROOT, "scenes", scene + ".json"), **over))
    sc = pt.Scene.from_file(cfg.scene_path, W, H, device=0)
    r = pt.renderer_from_config(sc, cfg.replace(**replace) if replace else cfg).start_render()
    rep = r.finalize()
    fb = r.framebuffer
    assert np.isfinite(fb).all() and rep["camera_rays"] == W * H * SPP, (scene, over)
    return sc, r, fb


def main():
    n = 0
    for scene in SCENES:
        sc, r, fb = render(scene, batch_samples=2)
        img, mx, _ = r.tonemap()
        assert img.shape == (H, W, 3) and mx > 0
        info = sc.bvh_info()
        hits = sc.trace_batch(raygen.random_rays(info["bounds"][:3], info["bounds"][3:], 2000, seed=3))
        assert (hits["instance"] >= 0).any()
        render(scene, light_samples=3, batch_samples=3)
        render(scene, integrator_type="recursive path tracing", max_bounces=4)
        render(scene, physically_corrected=True)
        render(scene, profile=2, stats_clamp=2.0)
        n += 5
    render("cornell_box", should_trace_paths=True, avg_number_of_paths=50)[1].paths()
    render("book2_final", light_samples=40)  # more than 32 light samples: the float-atomics path of k_nee
    # two shards on one device, asynchronous, with a snapshot while the workers run
    os.environ["B200PT_ALLOW_DUPLICATE_DEVICES"] = "1"
    cfg = pt.Config.from_json(pt.default_config_dict(width=W, height=H, samples=16,
                                                     scene=os.path.join(ROOT, "scenes", "book2_final.json")))
    sc = pt.Scene.from_file(cfg.scene_path, W, H, device=0)
    r = pt.renderer_from_config(sc, cfg.replace(devices=[0, 0], async_render=1, batch_samples=1)).start_render()
    time.sleep(0.01)
    mid = r.framebuffer
    r.tonemap(max(1, r.spp_done))
    r.finalize()
    assert np.isfinite(mid).all() and np.isfinite(r.framebuffer).all()
    # the allocation-free tonemap on a render's device film
    import ctypes as C
    import torch
    sc1, r1, _ = render("cornell_box")
    out = torch.zeros(W * H * 3, dtype=torch.uint8, device="cuda")
    d_film, _ = r1.device_buffer()
    mx = C.c_float()
    pt._capi.check(pt.lib().b200pt_tonemap_static(0, d_film, W, H, SPP, r1.config.exposure, out.data_ptr(), None,
                                                  C.byref(mx), None))
    assert mx.value > 0
    print(f"sanitize_run ok: {n + 4} renders")


if __name__ == "__main__":
    main()
