#!/usr/bin/env python3
"""Development aid: per-kernel-class device time (config.profile = 1) and rates for the BASELINE scenes on cuda:0."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pathtrace_b200 as pt  # noqa: E402

CASES = {"three_orbs": (400, 225, 64), "cornell_box": (1024, 1024, 16), "cornell_box_small_lights": (1024, 1024, 16),
         "cornell_box_with_volume": (1024, 1024, 16), "book2_final": (3840, 2160, 4)}
names = sys.argv[1:] or ["cornell_box", "cornell_box_with_volume", "book2_final"]
for name in names:
    w, h, spp = CASES[name]
    cfgd = pt.default_config_dict(width=w, height=h, samples=spp, scene=os.path.join(ROOT, "scenes", name + ".json"))
    scene = pt.Scene.from_file(cfgd["scene"], w, h, device=0)
    out = {}
    for mode in (0, 0, 1, 2):
        cfg = pt.Config.from_json(cfgd).replace(profile=mode)
        r = pt.renderer_from_config(scene, cfg).start_render()
        rep = r.finalize()
        if mode == 0:
            out.update(scene=name, mpaths=round(rep["mpaths_per_s"], 1), mrays=round(rep["mrays_per_s"], 1),
                       ms=round(rep["seconds"] * 1e3, 3))
        elif mode == 1:
            out["kernel_ms"] = {k: round(v, 3) for k, v in r.profile()["ms"].items()}
        else:
            p = r.profile()
            out["ext"] = [round(p["ext_boxes"] / max(1, p["ext_rays"]), 2), round(p["ext_prims"] / max(1, p["ext_rays"]), 2)]
            out["nee"] = [round(p["nee_boxes"] / max(1, p["nee_rays"]), 2), round(p["nee_prims"] / max(1, p["nee_rays"]), 2)]
            out["grays_nee"] = round(p["nee_rays"] / (out["kernel_ms"]["nee"] * 1e-3) / 1e9, 3) if out["kernel_ms"]["nee"] else 0
            out["grays_ext"] = round(p["ext_rays"] / (out["kernel_ms"]["extend"] * 1e-3) / 1e9, 3)
        r.close()
    print(json.dumps(out), flush=True)
