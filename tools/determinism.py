#!/usr/bin/env python3
"""Race check by determinism: each of the five scenes rendered three times (4 batches each, sort on) must give bit-identical
films and identical ray counters; an unsynchronised read-modify-write between kernels or lanes shows up as a difference."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pathtrace_b200 as pt  # noqa: E402

for scene, over in [("three_orbs", {}), ("cornell_box", {}), ("cornell_box_small_lights", {"light_samples": 3}),
                    ("cornell_box_with_volume", {}), ("book2_final", {}), ("book2_final", {"light_samples": 5})]:
    w, h, spp = 192, 108, 8
    cfg = pt.Config.from_json(pt.default_config_dict(width=w, height=h, samples=spp,
                                                     scene=os.path.join(ROOT, "scenes", scene + ".json"), **over))
    sc = pt.Scene.from_file(cfg.scene_path, w, h, device=0)
    films, counts = [], []
    for rep in range(3):
        r = pt.renderer_from_config(sc, cfg.replace(batch_samples=2)).start_render()
        out = r.finalize()
        films.append(r.framebuffer)
        counts.append((out["rays_extension"], out["rays_shadow"], out["rays_reference"]))
        r.close()
    same = all(np.array_equal(films[0], f) for f in films[1:]) and len(set(counts)) == 1
    print(scene, over, "bit-identical over 3 runs:", same, counts[0], flush=True)
    assert same
print("determinism ok")
