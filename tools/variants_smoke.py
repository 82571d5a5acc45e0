#!/usr/bin/env python3
"""Small renders through every kernel variant (small-scene walk, BVH with and without media, plain path tracing,
image textures, corrected estimator, counting kernels) plus a closest-hit batch.  With a library built with
`make VARIANT=_dbg EXTRA=-DPT_DEBUG_BOUNDS` (B200PT_LIB=pathtrace_b200/libb200pt_dbg.so) every node, instance and stack
index of the traversal is range-checked on the device and a violation traps."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pathtrace_b200 as pt  # noqa: E402

for name, over in [("cornell_box", {}), ("cornell_box_with_volume", {}), ("book2_final", {}),
                   ("book2_final", {"integrator_type": "recursive path tracing"}),
                   ("cornell_box_image_textures", {}), ("cornell_box", {"physically_corrected": True}),
                   ("book2_final", {"profile": 2})]:
    w, h = 64, 36
    cfgd = pt.default_config_dict(width=w, height=h, samples=2, scene=os.path.join(ROOT, "scenes", name + ".json"))
    cfgd.update({k: v for k, v in over.items() if k != "profile"})
    sc = pt.Scene.from_file(cfgd["scene"], w, h, device=0)
    cfg = pt.Config.from_json(json.dumps(cfgd)).replace(device=0, **({"profile": over["profile"]} if "profile" in over else {}))
    r = pt.renderer_from_config(sc, cfg).start_render()
    r.finalize()
    fb = r.framebuffer
    img, mx, avg = r.tonemap()
    b = sc.bvh_info()["bounds"]
    rays = np.random.default_rng(1).uniform(-1, 1, (4096, 8)).astype(np.float32)
    rays[:, :3] = b[:3] + (rays[:, :3] * 0.5 + 0.5) * (b[3:] - b[:3])
    rays[:, 6], rays[:, 7] = 0.001, 3.0e38
    hits = sc.trace_batch(rays)
    print(name, over, "film mean", float(fb.mean()), "hits", int((hits["instance"] >= 0).sum()), flush=True)
    r.close()
    sc.close()
print("done")
