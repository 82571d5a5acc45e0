"""a20 (film: calculate_luminance, output_to_file, tonemap_uncharted, to_srgb) and f3 (camera::project,
handle_traced_paths) pinned to the UNMODIFIED reference: oracle/_ref/ref_driver `film` / `project` run the
reference's own functions (helpers.h:146-168, renderer.h:24-55,57-96, tonemap.h:4-24, camera.h:49-109) on inputs
supplied by the test, and the library's outputs are compared with theirs."""
import os

import numpy as np
import pytest

from conftest import scene_path
from oracle_lib import RefDriver, have_ref_driver

needs_ref = pytest.mark.skipif(not have_ref_driver(), reason="oracle/_ref/ref_driver not built")


def _points(scene, rng, n):
    """points all over: inside the scene, behind the camera, in the camera's plane, absurdly far (the failure return)"""
    b = scene.bvh_info()["bounds"]
    cam = scene.describe()["camera"]
    inside = rng.uniform(b[:3], b[3:], size=(n, 3))
    around = cam["origin"] + rng.normal(size=(n, 3)) * 50.0
    behind = cam["origin"] + cam["w"] * rng.uniform(1.0, 500.0, size=(n, 1)) + rng.normal(size=(n, 3))
    far = rng.normal(size=(n // 4, 3)) * 1e11
    # t < 0 means: between the camera and the focal plane; the failure return also needs a coordinate beyond 1e10
    focus = float(np.dot(cam["origin"] - cam["lower_left"], cam["w"]))
    far_behind = (cam["origin"] - cam["w"] * focus * rng.uniform(0.05, 0.95, size=(n // 4, 1))
                  + cam["u"] * 1e11 * rng.choice([-1.0, 1.0], size=(n // 4, 1)) * rng.uniform(1.0, 5.0, size=(n // 4, 1)))
    in_plane = cam["origin"] + cam["u"] * rng.normal(size=(8, 1)) + cam["v"] * rng.normal(size=(8, 1))
    return np.concatenate([inside, around, behind, far, far_behind, in_plane, cam["origin"][None]]).astype(np.float32)


@needs_ref
@pytest.mark.parametrize("name,w,h", [("cornell_box", 64, 64), ("three_orbs", 400, 225), ("book2_final", 96, 54)])
def test_camera_project_matches_reference(name, w, h):
    import pathtrace_b200 as pt
    scene = pt.Scene.from_file(scene_path(name), w, h, device=-1)  # host-side entry point: no GPU needed
    pts = _points(scene, np.random.default_rng(5), 400)
    xy, ok = scene.project(pts)
    ref = RefDriver.project(scene_path(name), w, h, [pts])["projected"]
    assert np.array_equal(ok, ref[:, 2] > 0.5)
    assert ok.sum() > 800
    if name == "cornell_box":  # axis-aligned camera: the far-away points keep their exact depth in fp32
        assert (~ok).sum() > 20  # the failure return is exercised
    fin = np.isfinite(ref[:, :2]).all(axis=1)
    assert np.array_equal(fin, np.isfinite(xy).all(axis=1))
    np.testing.assert_allclose(xy[fin], ref[fin, :2], rtol=2e-5, atol=1e-6)
    behind = (ref[:, 0] == 0) & (ref[:, 1] == 0)
    assert behind.sum() > 30 and np.array_equal(behind, (xy[:, 0] == 0) & (xy[:, 1] == 0))  # x, y left at 0 (t < 0)


def _compare_film(r, fb, n):
    img, mx, avg = r.tonemap(n)
    ref_img, ref_mx, ref_avg, _ = RefDriver.film(fb, n, r.config.exposure)
    assert abs(mx - ref_mx) <= 1e-6 * abs(ref_mx), (mx, ref_mx)
    # helpers.h:167 divides the luminance total by pixels * n a second time; the total is a float sum over all pixels
    assert abs(avg - ref_avg) <= 1e-4 * abs(ref_avg) + 1e-12, (avg, ref_avg)
    diff = np.abs(img.astype(np.int32) - ref_img.astype(np.int32))
    # truncation to bytes: an ulp in powf (CUDA libm vs glibc) can move a value across an integer
    assert diff.max() <= 1 and (diff > 0).mean() < 0.01, (diff.max(), (diff > 0).mean())
    return img


@pytest.mark.gpu
@needs_ref
def test_film_matches_reference_output_to_file(tmp_path):
    import pathtrace_b200 as pt
    w, h, spp = 56, 40, 8
    cfg = pt.Config.from_json(pt.default_config_dict(width=w, height=h, samples=spp, scene=scene_path("cornell_box")))
    scene = pt.Scene.from_file(cfg.scene_path, w, h, device=0)
    r = pt.renderer_from_config(scene, cfg).start_render()
    r.finalize()
    fb = r.framebuffer
    img = _compare_film(r, fb, spp)
    _compare_film(r, fb, 3)  # sync_progress's divisor 1 + done / pixels (renderer.h:611-616)
    # the file itself: header + payload byte for byte what the library's tonemap returned
    path = str(tmp_path / "render.ppm")
    r.write_ppm(path)
    data = open(path, "rb").read()
    header = f"P6\n{w} {h}\n255\n".encode()
    assert data.startswith(header) and np.array_equal(np.frombuffer(data[len(header):], dtype=np.uint8), img.ravel())
    # a synthetic film: NaNs (de_nan in calculate_luminance only), negatives, zeros, a huge value that sets the white point
    rng = np.random.default_rng(11)
    syn = (rng.gamma(0.5, 2.0, size=(h, w, 3)) * spp).astype(np.float32)
    syn[0, 0] = [np.nan, 1.0, 2.0]
    syn[1, 1] = [-3.0, 0.0, 5.0]
    syn[2, 2] = 0.0
    syn[3, 3] = [4000.0, 100.0, 1.0]
    r.write_film(np.concatenate([syn, np.zeros((h, w, 1), np.float32)], axis=-1), spp)
    img2, mx2, _ = r.tonemap(spp)
    ref_img2, ref_mx2, _, _ = RefDriver.film(syn, spp, r.config.exposure)
    assert abs(mx2 - ref_mx2) <= 1e-6 * ref_mx2
    ok = ~np.isnan(syn).any(axis=-1)[::-1]  # int(NaN) is undefined in C++: that pixel is not compared
    diff = np.abs(img2.astype(np.int32) - ref_img2.astype(np.int32))[ok]
    assert diff.max() <= 1 and (diff > 0).mean() < 0.01


@pytest.mark.gpu
@needs_ref
def test_traced_path_files_match_reference_handle_traced_paths(tmp_path):
    import pathtrace_b200 as pt
    w, h, spp, avg = 64, 48, 8, 200
    cfgd = pt.default_config_dict(width=w, height=h, samples=spp, scene=scene_path("three_orbs"), should_trace_paths=True,
                                  avg_number_of_paths=avg, traced_paths_output_path=str(tmp_path / "p3.txt"),
                                  traced_paths_2d_output_path=str(tmp_path / "p2.txt"))
    cfg = pt.Config.from_json(cfgd)
    scene = pt.Scene.from_file(cfg.scene_path, w, h, device=0)
    r = pt.renderer_from_config(scene, cfg).start_render()
    r.finalize()
    paths = r.paths()
    n = r.write_paths()
    ref = RefDriver.project(cfg.scene_path, w, h, [p["points"] for p in paths])
    assert n == ref["added_paths"] > 50
    assert open(cfgd["traced_paths_output_path"]).read() == ref["text3d"]  # same values, same formatting
    ours2, ref2 = open(cfgd["traced_paths_2d_output_path"]).read().split("\n"), ref["text2d"].split("\n")
    assert len(ours2) == len(ref2)
    for a, b in zip(ours2, ref2):
        assert (a == "") == (b == "") and a.endswith("!") == b.endswith("!"), (a, b)
        if a:
            va, vb = [float(v) for v in a.rstrip("!").split(",")], [float(v) for v in b.rstrip("!").split(",")]
            np.testing.assert_allclose(va, vb, rtol=2e-5, atol=1e-6)
