#!/usr/bin/env python3
"""Renders the five BASELINE.json configs at their FULL sizes (film and samples per pixel) on cuda:0 through the public
API, writes gpurun_out/full_configs.json (device time, rates, ray counts) and a 256-pixel-wide PNG thumbnail of every
tonemapped image (the reference's own tonemap, renderer.h:24-55) under gpurun_out/renders/."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pathtrace_b200 as pt  # noqa: E402

CONFIGS = [("three_orbs", 400, 225, 64), ("cornell_box", 1024, 1024, 1024), ("cornell_box_small_lights", 1024, 1024, 1024),
           ("cornell_box_with_volume", 1024, 1024, 2048), ("book2_final", 3840, 2160, 1024)]
names = sys.argv[1:]
out_dir = os.path.join(ROOT, "gpurun_out", "renders")
os.makedirs(out_dir, exist_ok=True)
rows = []
for name, w, h, spp in CONFIGS:
    if names and name not in names:
        continue
    cfgd = pt.default_config_dict(width=w, height=h, samples=spp, scene=os.path.join(ROOT, "scenes", name + ".json"))
    cfg = pt.Config.from_json(cfgd)
    t0 = time.perf_counter()
    scene = pt.Scene.from_file(cfg.scene_path, w, h, device=0)
    r = pt.renderer_from_config(scene, cfg).start_render()
    rep = r.finalize()
    img, max_lum, avg_lum = r.tonemap()
    wall = time.perf_counter() - t0
    row = {"scene": name, "width": w, "height": h, "spp": spp, "camera_paths": rep["camera_rays"], "rays": rep["rays"],
           "device_seconds": rep["seconds"], "wall_seconds": wall, "mpaths_per_s": rep["mpaths_per_s"],
           "mrays_per_s": rep["mrays_per_s"], "rays_per_path": rep["rays"] / rep["camera_rays"],
           "max_luminance": max_lum, "avg_luminance": avg_lum}
    try:
        from PIL import Image
        im = Image.fromarray(img, "RGB")
        im.thumbnail((256, 256))
        im.save(os.path.join(out_dir, name + ".png"))
    except Exception as e:  # thumbnails are a nicety
        row["thumbnail_error"] = str(e)
    rows.append(row)
    print(json.dumps(row), flush=True)
    r.close()
    scene.close()
with open(os.path.join(ROOT, "gpurun_out", "full_configs.json"), "w") as f:
    json.dump(rows, f, indent=1)
