#!/usr/bin/env python3
"""Development aid: time of b200pt_scene_create_from_json for the 8192-instance scene on cuda:0 (parse, compile, BVH,
pack, upload); B200PT_LIB selects another build of the library.  Run from the repo root."""
import sys, time, os
sys.path.insert(0, os.getcwd())
import pathtrace_b200 as pt
js = open("scenes/book2_final.json").read()
ts = []
for i in range(12):
    t = time.perf_counter(); s = pt.Scene.from_json(js, 3840, 2160, device=0); ts.append((time.perf_counter() - t) * 1e3); s.close()
ts = sorted(ts[2:]); print(os.environ.get("B200PT_LIB"), "scene_create ms: min %.2f median %.2f" % (ts[0], ts[len(ts) // 2]))
