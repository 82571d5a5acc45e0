"""The drop-in CLI (pathtrace_b200/b200pt_main: main.cpp:108-168 through the C ABI) and the traced-path debug dump
(should_trace_paths: config.h:122-131, renderer.h:57-96,652-678, integrator.h:195-197, camera.h:49-109)."""
import json
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, scene_path

pytestmark = pytest.mark.gpu

CLI = os.path.join(ROOT, "pathtrace_b200", "b200pt_main")


def _project(cam, p):
    """camera::project (camera.h:49-109) in numpy, for points in front of the film plane."""
    o, ll, hz, vt, w = (np.asarray(cam[k], dtype=np.float64) for k in ("origin", "lower_left", "horizontal", "vertical", "w"))
    d = o - p
    t = np.dot(ll - p, w) / np.dot(d, w)
    q = p + t * d - ll
    return np.dot(hz, q) / np.dot(hz, hz), np.dot(vt, q) / np.dot(vt, vt)


def test_traced_paths(tmp_path):
    import pathtrace_b200 as pt
    w, h, spp, avg = 64, 48, 16, 300
    cfgd = pt.default_config_dict(width=w, height=h, samples=spp, scene=scene_path("cornell_box"), should_trace_paths=True,
                                  avg_number_of_paths=avg, traced_paths_output_path=str(tmp_path / "o" / "paths.txt"),
                                  traced_paths_2d_output_path=str(tmp_path / "o" / "paths2d.txt"))
    cfg = pt.Config.from_json(cfgd)
    assert abs(cfg.trace_probability - avg / (w * h * spp)) < 1e-9
    scene = pt.Scene.from_file(cfg.scene_path, w, h, device=0)
    r = pt.renderer_from_config(scene, cfg.replace(batch_samples=4)).start_render()  # 4 batches share one path store
    r.finalize()
    paths = r.paths()
    assert abs(len(paths) - avg) < 6 * np.sqrt(avg), len(paths)  # binomial, mean avg
    keys = [(p["sample"], p["pixel"]) for p in paths]
    assert keys == sorted(keys) and len(set(keys)) == len(keys)
    assert {p["sample"] // 4 for p in paths} == {0, 1, 2, 3}  # every batch contributed
    flat = scene.describe()
    rays, expect = [], []
    for p in paths:
        pts = p["points"]
        assert len(pts) <= cfg.max_bounces + 1
        if len(pts) == 0:  # the camera ray passed beside the box: nothing was pushed (and the file skips the path)
            continue
        # the first vertex is seen through the pixel the sample belongs to (row 0 = image bottom)
        x, y = _project(flat["camera"], pts[0].astype(np.float64))
        i, j = p["pixel"] % w, p["pixel"] // w
        assert i - 1e-3 <= x * w <= i + 1 + 1e-3 and j - 1e-3 <= y * h <= j + 1 + 1e-3, (p["pixel"], x * w, y * h)
        for a, b in zip(pts[:-1], pts[1:]):  # consecutive vertices see each other
            rays.append(np.concatenate([a, b - a, [1e-3, 3.4e38]]))
            expect.append(b)
    rays = np.asarray(rays, dtype=np.float32)
    hits = scene.trace_batch(rays)
    err = np.linalg.norm(hits["p"] - np.asarray(expect), axis=1)
    assert (hits["instance"] >= 0).mean() > 0.99 and np.mean(err < 0.5) > 0.98, (np.mean(err < 0.5), err.max())
    # the same samples are recorded whatever the batching, and recording does not change the film
    r2 = pt.renderer_from_config(scene, cfg.replace(batch_samples=16)).start_render()
    r2.finalize()
    assert [(p["sample"], p["pixel"]) for p in r2.paths()] == keys
    r3 = pt.renderer_from_config(scene, cfg.replace(trace_probability=0.0)).start_render()
    r3.finalize()
    assert np.array_equal(r3.framebuffer, r2.framebuffer) and r3.paths() == []
    # the reference's two text files (renderer.h:57-96)
    n = r.write_paths()
    paths = [p for p in paths if len(p["points"])]
    assert n == len(paths) > 0.5 * avg
    blocks3 = open(cfgd["traced_paths_output_path"]).read().strip("\n").split("\n\n")
    assert len(blocks3) == n
    first = np.array([[float(v) for v in line.split(",")] for line in blocks3[0].split("\n")])
    np.testing.assert_allclose(first, paths[0]["points"], rtol=1e-5)
    blocks2 = open(cfgd["traced_paths_2d_output_path"]).read().split("\n\n")
    xy = [float(v) for v in blocks2[0].strip().split("\n")[0].rstrip("!").split(",")]
    x, y = _project(flat["camera"], paths[0]["points"][0].astype(np.float64))
    assert abs(xy[0] - x) < 1e-4 and abs(xy[1] - y) < 1e-4


def test_cli_is_a_drop_in_for_main(tmp_path):
    """b200pt_main reads ./config.json like the reference's main() (main.cpp:108-121), prints its rate lines
    (renderer.h:700-724, the lines bench.py parses from the reference) and writes the PPM."""
    import pathtrace_b200 as pt
    assert os.path.exists(CLI), "build first: make -C pathtrace_b200/csrc"
    cfgd = pt.default_config_dict(width=96, height=64, samples=8, scene=scene_path("three_orbs"),
                                  ppm_output_path="render.ppm", should_trace_paths=True, avg_number_of_paths=50,
                                  traced_paths_output_path="paths.txt", traced_paths_2d_output_path="paths2d.txt")
    with open(tmp_path / "config.json", "w") as f:
        json.dump(cfgd, f)
    out = subprocess.run([CLI], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-500:]
    m1 = re.search(r"computed (\d+) camera rays in ([0-9.eE+-]+)s", out.stdout)
    m2 = re.search(r"computed (\d+) rays, at ([0-9.eE+-]+) rays per second", out.stdout)
    assert m1 and m2 and int(m1.group(1)) == 96 * 64 * 8 and int(m2.group(1)) > int(m1.group(1))
    assert "max lum" in out.stdout and "time taken to build bvh" in out.stdout
    data = open(tmp_path / "render.ppm", "rb").read()
    header = b"P6\n96 64\n255\n"
    assert data.startswith(header) and len(data) == len(header) + 96 * 64 * 3
    img = np.frombuffer(data[len(header):], dtype=np.uint8)
    assert 5 < img.mean() < 250
    assert re.search(r"wrote (\d+) traced paths", out.stdout) and os.path.getsize(tmp_path / "paths.txt") > 0
