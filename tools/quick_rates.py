#!/usr/bin/env python3
"""Development aid: render every BASELINE scene briefly on cuda:0 and print Mpaths/s / Mrays/s."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pathtrace_b200 as pt  # noqa: E402

CASES = [("three_orbs", 400, 225, 64), ("cornell_box", 1024, 1024, 32), ("cornell_box_small_lights", 1024, 1024, 32),
         ("cornell_box_with_volume", 1024, 1024, 32), ("book2_final", 3840, 2160, 4)]
if len(sys.argv) > 1:
    CASES = [c for c in CASES if c[0] in sys.argv[1:]]
out = []
for name, w, h, spp in CASES:
    cfgd = pt.default_config_dict(width=w, height=h, samples=spp, scene=os.path.join(ROOT, "scenes", name + ".json"))
    cfg = pt.Config.from_json(cfgd)
    t0 = time.time()
    scene = pt.Scene.from_file(cfg.scene_path, w, h, device=0)
    t1 = time.time()
    for rep in range(2):
        r = pt.renderer_from_config(scene, cfg).start_render()
        report = r.finalize()
        r.close()
    report.update(scene=name, w=w, h=h, spp=spp, scene_s=t1 - t0, rays_per_path=report["rays"] / report["camera_rays"])
    print(json.dumps(report), flush=True)
    out.append(report)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", "quick_rates.json"), "w") as f:
    json.dump(out, f, indent=1)
