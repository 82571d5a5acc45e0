#!/usr/bin/env python3
"""Development aid: wall-clock split of one end-to-end step (what bench.py's e2e times) on cuda:0."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pathtrace_b200 as pt  # noqa: E402  (B200PT_LIB selects another build of the library)

name = sys.argv[1] if len(sys.argv) > 1 else "cornell_box"
w, h, spp = (3840, 2160, 4) if name == "book2_final" else (1024, 1024, 16)
cfgd = pt.default_config_dict(width=w, height=h, samples=spp, scene=os.path.join(ROOT, "scenes", name + ".json"))
scene_json = open(cfgd["scene"]).read()
cfg_json = json.dumps(cfgd)
for rep in range(4):
    t = [time.perf_counter()]
    sc = pt.Scene.from_json(scene_json, w, h, device=0); t.append(time.perf_counter())
    c = pt.Config.from_json(cfg_json).replace(device=0); t.append(time.perf_counter())
    r = pt.renderer_from_config(sc, c).start_render(); t.append(time.perf_counter())
    rep_ = r.finalize(); t.append(time.perf_counter())
    fb = r.framebuffer; t.append(time.perf_counter())
    r.close(); sc.close(); t.append(time.perf_counter())
    names = ["scene_create", "config", "render_start", "render", "read_film", "close"]
    print(json.dumps({"rep": rep, **{n: round((b - a) * 1e3, 3) for n, a, b in zip(names, t, t[1:])},
                      "total_ms": round((t[-1] - t[0]) * 1e3, 3), "device_ms": round(rep_["seconds"] * 1e3, 3)}), flush=True)
