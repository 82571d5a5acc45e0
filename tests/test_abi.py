"""The C-ABI surface without a GPU: the library loads, exports every symbol include/b200pt.h declares, parses
config.json like Config(json) (config.h:98-132, quirks included), and refuses to trace without a device."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

from conftest import ROOT, scene_path


def _header_symbols():
    with open(os.path.join(ROOT, "include", "b200pt.h")) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200pt_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import pathtrace_b200 as pt
    from pathtrace_b200 import _capi
    names = _header_symbols()
    assert len(names) >= 20
    handle = C.CDLL(pt.LIB_PATH)
    for n in names:
        assert hasattr(handle, n), n
    assert sorted(_capi.SIGNATURES) == names  # the ctypes table binds exactly the header
    assert b"sm_100a" in pt.lib().b200pt_version()


def test_struct_sizes_match_header():
    from pathtrace_b200 import _capi
    assert C.sizeof(_capi.Instance) == 48 + 4 * 4 + 24 + 4 + 4
    assert C.sizeof(_capi.Material) == 32 and C.sizeof(_capi.Texture) == 32
    assert _capi.HIT_DTYPE.itemsize == 40
    assert C.sizeof(_capi.Config) % 8 == 0


def test_config_defaults_and_quirks():
    import pathtrace_b200 as pt
    minimal = {"film": {}, "traced_paths_output_path": "a", "traced_paths_2d_output_path": "b"}
    c = pt.Config.from_json(minimal)
    assert (c.width, c.height) == (400, 300)                      # config.h:21-22
    assert abs(c.exposure - 2.2) < 1e-6 and c.gamma == 0.0        # Q1: "gamma" -> exposure, "exposure" -> gamma
    assert c.integrator == 0 and c.max_bounces == 10 and c.samples == 20 and c.light_samples == 1
    assert c.russian_roulette == 1 and c.only_direct_illumination == 0 and abs(c.normal_offset - 1e-4) < 1e-9
    assert c.scene_path == "scenes/scene.json" and c.ppm_output_path == "out.ppm"
    c = pt.Config.from_json(dict(minimal, film={"gamma": 1.5, "exposure": 3.0, "width": 7, "height": 5}))
    assert abs(c.exposure - 1.5) < 1e-6 and abs(c.gamma - 3.0) < 1e-6 and (c.width, c.height) == (7, 5)
    c = pt.Config.from_json(dict(minimal, integrator_type="iterative nee path tracing"))
    assert c.integrator == 4
    c = pt.Config.from_json(dict(minimal, integrator_type="metropolis light transport"))
    assert c.integrator == 0                                      # unknown strings end in RecursivePT (config.h:59-72)
    with pytest.raises(pt.B200ptError):
        pt.Config.from_json(dict(minimal, integrator_type="recursive nee path tracing"))
    with pytest.raises(pt.B200ptError):                           # config.h:104-105: required keys
        pt.Config.from_json({"film": {}})
    with pytest.raises(pt.B200ptError):
        pt.Config.from_json("{ broken")
    shipped = pt.Config.from_json(pt.default_config_dict())
    assert (shipped.width, shipped.height, shipped.light_samples, shipped.integrator) == (600, 600, 4, 4)


def test_no_cpu_fallback():
    import pathtrace_b200 as pt
    s = pt.Scene.from_file(scene_path("cornell_box"), 32, 32, device=-1)
    assert s.bvh_info()["n_nodes"] == 4  # 8 instances: 7 binary nodes collapsed to 4-wide ones
    with pytest.raises(pt.B200ptError) as e:
        s.trace_batch(np.zeros((1, 8), np.float32))
    assert e.value.code == 3
    cfg = pt.Config.from_json(pt.default_config_dict(width=32, height=32, samples=1)).replace(device=-1)
    with pytest.raises(pt.B200ptError) as e:
        pt.renderer_from_config(s, cfg).start_render()
    assert e.value.code == 3


def test_scene_errors():
    import pathtrace_b200 as pt
    with open(scene_path("cornell_box")) as f:
        sj = json.load(f)
    bad = dict(sj, instances=[{"type": "ref", "primitive": {"id": "does-not-exist"}}])
    with pytest.raises(pt.B200ptError):
        pt.Scene.from_json(bad, 32, 32, device=-1)
    with pytest.raises(pt.B200ptError):
        pt.Scene.from_json(dict(sj, instances=[]), 32, 32, device=-1)


def test_config_device_keys():
    """The scheduling keys this backend adds next to the reference's "threads": "devices": N, "device_ids": [...],
    "async_render"; like "threads" they do not change the image.  Absent: one GPU, the paced mode."""
    import pathtrace_b200 as pt
    base = pt.default_config_dict()
    c = pt.Config.from_json(base)
    assert c.device_count <= 1 and c.async_render == 0 and c.device == 0
    c = pt.Config.from_json(dict(base, devices=4))
    assert c.device_count == 4 and list(c.device_ids)[:4] == [0, 1, 2, 3]
    c = pt.Config.from_json(dict(base, device_ids=[2, 5, 7], async_render=True))
    assert c.device_count == 3 and list(c.device_ids)[:3] == [2, 5, 7] and c.device == 2 and c.async_render == 1
    for bad in (dict(devices=9), dict(device_ids=[]), dict(device_ids=list(range(9)))):
        with pytest.raises(pt.B200ptError):
            pt.Config.from_json(dict(base, **bad))
    # Config.replace(devices=[...]) is the Python spelling
    c = pt.Config.from_json(base).replace(devices=[1, 3])
    assert c.device_count == 2 and list(c.device_ids)[:2] == [1, 3] and c.device == 1
    assert pt.Config.from_json(base).replace(devices=[2]).device_count == 0


def test_render_start_refuses_bad_device_lists_without_a_gpu():
    """No device here: a multi-device render must fail with NO_DEVICE (never fall back), a one-device one likewise."""
    import pathtrace_b200 as pt
    scene = pt.Scene.from_file(scene_path("cornell_box"), 16, 16, device=-1)
    cfg = pt.Config.from_json(pt.default_config_dict(width=16, height=16, samples=4, scene=scene_path("cornell_box")))
    for c in (cfg.replace(device=-1), cfg.replace(devices=[0, 1])):
        with pytest.raises(pt.B200ptError):
            pt.renderer_from_config(scene, c).start_render()
