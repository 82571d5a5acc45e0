"""INTEGRATION.md section B, compiled and run: oracle/_ref/wavefront_main is `Wavefront : public Renderer` (renderer.h:114-150)
built against the UNMODIFIED reference headers + libb200pt.so and driven by the reference's own lifecycle loop
(main.cpp:153-167).  `"render_type": "wavefront"` selects it.  Its PPM is the REFERENCE's film path (calculate_luminance +
output_to_file, helpers.h:146-168, renderer.h:24-55) applied to the GPU's film; it must equal the PPM the library's own
tonemap writes through the drop-in CLI for the same config."""
import json
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, scene_path

pytestmark = pytest.mark.gpu

BINDING = os.path.join(ROOT, "oracle", "_ref", "wavefront_main")
CLI = os.path.join(ROOT, "pathtrace_b200", "b200pt_main")


def _ppm(path, w, h):
    data = open(path, "rb").read()
    header = f"P6\n{w} {h}\n255\n".encode()
    assert data.startswith(header) and len(data) == len(header) + w * h * 3, (len(data), len(header) + w * h * 3)
    return np.frombuffer(data[len(header):], dtype=np.uint8).astype(np.int32)


@pytest.mark.parametrize("scene,w,h,spp", [("cornell_box", 160, 120, 32), ("book2_final", 480, 270, 24)])
def test_reference_side_binding_renders_through_the_library(tmp_path, scene, w, h, spp):
    import pathtrace_b200 as pt
    if not os.path.exists(BINDING):
        pytest.skip("oracle/_ref/wavefront_main not built (needs /root/reference at build time)")
    cfgd = pt.default_config_dict(width=w, height=h, samples=spp, scene=scene_path(scene), render_type="wavefront",
                                  ppm_output_path="render.ppm", traced_paths_output_path="p.txt",
                                  traced_paths_2d_output_path="p2.txt", should_trace_paths=True, avg_number_of_paths=40,
                                  batch_samples=2)
    outs = {}
    for tag, exe, args in (("binding", BINDING, ["2"]), ("cli", CLI, [])):
        d = tmp_path / tag
        d.mkdir()
        with open(d / "config.json", "w") as f:
            json.dump(cfgd, f)
        out = subprocess.run([exe] + args, cwd=d, capture_output=True, text=True, timeout=600)
        assert out.returncode == 0, (out.stdout[-400:], out.stderr[-800:])
        outs[tag] = (out.stdout, _ppm(d / "render.ppm", w, h), d)
    text = outs["binding"][0]
    assert "selected and constructed Wavefront renderer" in text and " done" in text
    m1 = re.search(r"computed (\d+) camera rays in ([0-9.eE+-]+)s", text)
    m2 = re.search(r"computed (\d+) rays, at ([0-9.eE+-]+) rays per second", text)  # the lines bench.py parses
    assert m1 and m2 and int(m1.group(1)) == w * h * spp and int(m2.group(1)) > int(m1.group(1))
    assert int(re.search(r"wavefront previews (\d+)", text).group(1)) >= 1  # sync_progress ran (main.cpp:158-163)
    # the reference's output_to_file on the GPU film == the library's k_tonemap (an ulp of powf can move one level)
    diff = np.abs(outs["binding"][1] - outs["cli"][1])
    assert diff.max() <= 1 and (diff > 0).mean() < 0.01, (diff.max(), (diff > 0).mean())
    assert 5 < outs["binding"][1].mean() < 250
    # max lum as the reference's calculate_luminance prints it vs the library's device reduce
    mb = float(re.search(r"max lum ([0-9.eE+-]+)", text).group(1))
    mc = float(re.search(r"max lum ([0-9.eE+-]+)", outs["cli"][0]).group(1))
    assert abs(mb - mc) <= 1e-5 * mc
    # traced paths: the same device store, the same files
    added = int(re.search(r"added (\d+) paths", text).group(1))
    assert added > 5 and open(outs["binding"][2] / "p.txt").read() == open(outs["cli"][2] / "p.txt").read()
