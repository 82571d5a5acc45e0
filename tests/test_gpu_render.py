"""Parity check 2, GPU side: renders through the C ABI against the CPU restatement, which draws the
same Philox variates -- so the comparison is per SAMPLE SUM, not statistical: it checks the wavefront plumbing
(queues, compaction, material sort, batching, accumulation order).  The restatement itself is pinned against the
unmodified reference in tests/test_oracle_vs_reference.py (CPU) and test_render_matches_reference_golden below."""
import json
import os

import numpy as np
import pytest

import parity_protocol
from conftest import ROOT, scene_path
from oracle_lib import PortOracle, make_cconfig

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _render(cfgd, **over):
    import pathtrace_b200 as pt
    cfg = pt.Config.from_json(cfgd)
    if over:
        cfg = cfg.replace(**over)
    scene = pt.Scene.from_file(cfg.scene_path, cfg.width, cfg.height, device=0)
    r = pt.renderer_from_config(scene, cfg).start_render()
    rep = r.finalize()
    return r, rep


def _cfg(scene, w, h, spp, **kw):
    import pathtrace_b200 as pt
    return pt.default_config_dict(width=w, height=h, samples=spp, scene=scene_path(scene), **kw)


def _close(a, b, what, frac_limit=5e-3, total_limit=1e-3, level="loose"):
    """Sample sums agree up to libm ulps (CUDA's sinf / cosf / logf / powf vs glibc's) amplified through a few bounces; a
    plumbing error -- a lost queue slot, a sample added twice -- shows up as O(1) on the pixels it touches.  Limits are a
    few times what the cases measure (gpurun_out/parity/close_stats.jsonl).  level:
      "bit"    no libm call on the path (direct lighting from rect lights, RecursivePT's products): bit-equal sums
      "tight"  trig / powf ulps only, no decision can flip at these sizes: every pixel within 1e-4 relative (measured <= 6e-6)
      "loose"  media (logf decides whether a scatter event happens), Russian roulette on long paths, sphere lights: a flipped
               decision changes a whole sample -- at most 5e-3 of the pixels beyond 1e-3 (measured 1.6e-3), totals 1e-3."""
    scale = np.maximum(np.abs(b), 1e-3 * max(1.0, float(np.abs(b).max())))
    rel = np.abs(a - b) / scale
    frac_bad = float((rel > 1e-3).mean())
    tot_a, tot_b = float(a.astype(np.float64).sum()), float(b.astype(np.float64).sum())
    try:
        os.makedirs(os.path.join(ROOT, "gpurun_out", "parity"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "parity", "close_stats.jsonl"), "a") as f:
            f.write(json.dumps({"what": str(what), "frac_gt_1e-3": frac_bad, "rel_max": float(rel.max()),
                                "bit_equal": bool(np.array_equal(a, b)),
                                "total_rel": abs(tot_a - tot_b) / max(abs(tot_b), 1e-30)}) + "\n")
    except OSError:
        pass
    if level == "bit":
        assert np.array_equal(a, b), (what, frac_bad, float(rel.max()))
    if level == "tight":
        assert float(rel.max()) <= 1e-4, (what, float(rel.max()))
    assert frac_bad < frac_limit, (what, frac_bad, float(rel.max()))
    assert abs(tot_a - tot_b) <= total_limit * abs(tot_b) + 1e-6, (what, tot_a, tot_b)


CASES = [
    ("three_orbs", dict(), "tight"),
    ("cornell_box", dict(), "tight"),
    ("cornell_box_small_lights", dict(light_samples=3), "tight"),
    ("cornell_box_with_volume", dict(), "tight"),
    ("cornell_box_with_volume2", dict(russian_roulette=False, max_bounces=4), "tight"),
    ("light_test", dict(light_samples=1), "loose"),
    ("book2_final", dict(), "loose"),
    ("cornell_box", dict(integrator_type="recursive path tracing"), "bit"),
    ("three_orbs", dict(integrator_type="recursive path tracing", max_bounces=5), "bit"),
    ("cornell_box", dict(only_direct_illumination=True), "bit"),
    ("cornell_box", dict(max_bounces=1), "bit"),
    ("cornell_box", dict(russian_roulette=False, max_bounces=3), "tight"),
    ("cornell_box_image_textures", dict(), "tight"),
    ("cornell_box_image_textures", dict(integrator_type="recursive path tracing", max_bounces=3), "bit"),
]


@pytest.mark.parametrize("scene,over,level", CASES)
def test_render_matches_cpu_restatement(scene, over, level):
    w, h, spp = 48, 40, 6
    cfgd = _cfg(scene, w, h, spp, **over)
    r, rep = _render(cfgd, batch_samples=4)  # 2 batches: 4 + 2 samples
    fb = r.framebuffer
    oracle = PortOracle.from_file(scene_path(scene), w, h)
    ref = oracle.render(make_cconfig(cfgd), stats=False)
    _close(fb, ref["sums"], scene, level=level)
    # ray accounting: exact where no decision can flip; else a differing libm ulp can flip a Russian-roulette or medium
    # decision on a handful of paths
    for k in ("rays_extension", "rays_shadow", "rays_reference"):
        if level == "bit":
            assert rep[k] == ref[k], (k, rep[k], ref[k])
        assert abs(rep[k] - ref[k]) <= 2e-3 * ref[k] + 8, (k, rep[k], ref[k])
    assert rep["camera_rays"] == w * h * spp


def test_image_textured_light_matches_cpu_restatement(tmp_path):
    """The variant the reference cannot render (scenes/make_scenes.py: its NEE indexes the image with int(NaN)): the
    light's emission AND alpha come from a png (material.h:209-228), shadow hits must then carry p, u, v
    (PtSceneView::emit_needs_p).  GPU against the CPU restatement, sample for sample."""
    with open(scene_path("cornell_box_image_textures")) as f:
        sj = json.load(f)
    for m in sj["materials"]:
        if m["id"] == "light":
            m["data"] = {"texture": "panel", "power": 3.0, "two_sided": True}
    path = str(tmp_path / "image_light.json")
    with open(path, "w") as f:
        json.dump(sj, f)
    import pathtrace_b200 as pt
    w, h, spp = 48, 40, 6
    cfgd = pt.default_config_dict(width=w, height=h, samples=spp, scene=path)
    here = os.getcwd()
    os.chdir(ROOT)  # "scenes/assets/panel.png" is relative to the repository root
    try:
        r, rep = _render(cfgd, batch_samples=4)
        oracle = PortOracle.from_file(path, w, h)
    finally:
        os.chdir(here)
    ref = oracle.render(make_cconfig(cfgd), stats=False)
    fb = r.framebuffer
    _close(fb, ref["sums"], "image light")
    assert fb.max() > 0
    plain = _render(_cfg("cornell_box_image_textures", w, h, spp), batch_samples=4)[0].framebuffer
    assert not np.allclose(plain, fb)  # the textured light does change the picture


def test_batching_and_sharding_invariance():
    """Philox is keyed on (pixel, GLOBAL sample): the film does not depend on the batch size, and two spp shards
    (sample_offset) add up to the unsharded film -- what multi-GPU sharding relies on."""
    w, h, spp = 64, 48, 8
    cfgd = _cfg("cornell_box", w, h, spp)
    a = _render(cfgd, batch_samples=1)[0].framebuffer
    b = _render(cfgd, batch_samples=8)[0].framebuffer
    assert np.array_equal(a, b)
    half = dict(cfgd, samples=spp // 2)
    s0 = _render(half, sample_offset=0)[0].framebuffer
    s1 = _render(half, sample_offset=spp // 2)[0].framebuffer
    np.testing.assert_allclose(s0 + s1, a, rtol=1e-5, atol=1e-6)
    c = _render(cfgd, seed=123)[0].framebuffer
    assert not np.array_equal(a, c)


def test_stats_buffers():
    w, h, spp = 32, 32, 16
    cfgd = _cfg("cornell_box", w, h, spp)
    r, _ = _render(cfgd, stats_clamp=2.0)
    s, q = r.stats()
    oracle = PortOracle.from_file(scene_path("cornell_box"), w, h)
    ref = oracle.render(make_cconfig(cfgd, stats_clamp=2.0))
    _close(s, ref["csum"], "csum")
    _close(q, ref["csq"], "csq")
    assert float(s.max()) <= 2.0 * spp + 1e-3


def _tonemap_numpy(fb, n, exposure):
    """output_to_file + calculate_luminance (renderer.h:24-55, helpers.h:78-93,146-168, tonemap.h:4-24) in numpy fp32."""
    f32 = np.float32
    inv_n = f32(1.0 / float(f32(n)))
    c = np.where(np.isnan(fb), f32(0), fb).astype(f32) * inv_n
    lum = np.sqrt((c.astype(np.float64) ** 2).sum(-1))
    max_lum = f32(lum.max())

    def u2(x):
        A, B, C, D, E, F = map(f32, (0.15, 0.50, 0.10, 0.20, 0.02, 0.30))
        return ((x * (A * x + C * B) + D * E) / (x * (A * x + B) + D * F)) - E / F

    x = fb.astype(f32) * inv_n * f32(16 + exposure)
    x = np.clip(u2(x) / u2(max_lum), 0, 1).astype(f32)
    srgb = np.where(x < f32(0.0031308), f32(323) * x / f32(25), (f32(211) * np.power(x, f32(5.0 / 12)) - f32(11)) / f32(200))
    img = (f32(255) * srgb.astype(f32)).astype(np.int32).astype(np.uint8)
    return img[::-1], float(max_lum)


def test_tonemap_and_ppm(tmp_path):
    w, h, spp = 40, 24, 4
    cfgd = _cfg("cornell_box", w, h, spp)
    r, _ = _render(cfgd)
    fb = r.framebuffer
    img, mx, avg = r.tonemap()
    exp_img, exp_max = _tonemap_numpy(fb, spp, r.config.exposure)
    assert abs(mx - exp_max) <= 1e-5 * exp_max
    diff = np.abs(img.astype(np.int32) - exp_img.astype(np.int32))
    assert diff.max() <= 1 and (diff > 0).mean() < 0.01  # truncation to bytes: an ulp in powf can move one level
    path = str(tmp_path / "out" / "render.ppm")
    r.write_ppm(path)
    with open(path, "rb") as f:
        data = f.read()
    header = f"P6\n{w} {h}\n255\n".encode()
    assert data.startswith(header) and len(data) == len(header) + w * h * 3
    assert np.array_equal(np.frombuffer(data[len(header):], dtype=np.uint8).reshape(h, w, 3), img)
    assert abs(r.config.exposure - 2.2) < 1e-6  # Q1: JSON "gamma" lands in film.exposure (config.h:24-25)


@pytest.mark.parametrize("tag", ["cornell_box_d1", "cornell_box_b3", "cornell_box_rpt", "three_orbs_full",
                                 "cornell_box_small_lights_d1", "cornell_box_with_volume_b2", "light_test_b2",
                                 "book2_final_b2", "cornell_box_image_textures_b2"])
def test_render_matches_reference_golden(tag):
    """Statistical parity with the UNMODIFIED reference: golden per-pixel statistics of integrator->color made by
    tests/golden/make_golden.py (ref_driver `color` mode).  Clamped samples on both sides (finite variance)."""
    path = os.path.join(GOLDEN, f"color_{tag}.npz")
    if not os.path.exists(path):
        pytest.skip("no golden file")
    g = np.load(path, allow_pickle=False)
    cfgd = json.loads(str(g["config"]))
    cfgd["scene"] = scene_path(os.path.basename(cfgd["scene"]).replace(".json", ""))
    n_ref, clamp = int(g["n"]), float(g["clamp"])
    n_gpu = 4096
    cfgd["samples"] = n_gpu
    r, _ = _render(cfgd, stats_clamp=clamp, seed=77)
    s, q = r.stats()
    mean_g = s.astype(np.float64) / n_gpu
    var_g = np.maximum(q.astype(np.float64) / n_gpu - mean_g ** 2, 0) * n_gpu / (n_gpu - 1)
    mean_r = g["cmean"]
    var_r = g["cm2"] / (n_ref - 1)
    se = np.sqrt(var_g / n_gpu + var_r / n_ref)
    ok = se > 0
    z = (mean_g[ok] - mean_r[ok]) / se[ok]
    assert abs(z.mean()) < 0.2, z.mean()  # skewed samples bias z by O(skew / sqrt(n)): not exactly 0
    assert (np.abs(z) > 3).mean() < 0.02, (np.abs(z) > 3).mean()
    assert (np.abs(z) > 6).mean() < 1e-3, np.abs(z).max()
    # deterministic pixels (zero variance on both sides) must agree outright
    np.testing.assert_allclose(mean_g[~ok], mean_r[~ok], rtol=1e-4, atol=1e-6)


REPORT_DIR = os.path.join(ROOT, "gpurun_out", "parity")


@pytest.mark.parametrize("tag", parity_protocol.FULL_TAGS + parity_protocol.LADDER_TAGS + ["cornell_box_4096"])
def test_full_depth_protocol_gpu_vs_reference(tag):
    """SURVEY 8c protocol on the GPU at the SHIPPED integrator settings (config.json:12-25: INEEPT, max_bounces 10,
    light_samples 4, RR on) for every BASELINE scene, the RR-off / direct-only rungs, and the north star's literal
    4096 spp on cornell_box: (2) clamped per-pixel z-scores, (3) 8x8-block rRMSE <= 1.2 x the reference-vs-reference
    floor, (4) unclamped global mean within 3 sigma -- against goldens of the UNMODIFIED reference
    (integrator.h:176-339 run by oracle/ref_driver `color`, tests/golden/make_golden.py).  The figures, including the
    literal per-pixel rRMSE next to its floor, are appended to gpurun_out/parity/gpu_vs_reference.jsonl."""
    g = parity_protocol.load_golden(tag)
    cfgd = dict(g["config"])
    cfgd["scene"] = scene_path(os.path.basename(cfgd["scene"]).replace(".json", ""))
    n1 = g["n"]  # step 2 at the reference's own sample count, steps 3-4 over all samples (parity_protocol.evaluate)
    n_gpu = max(4096, 2 * n1)
    parts = []
    for first, count in ((0, n1), (n1, n_gpu - n1)):  # two spp shards of one render (Philox is keyed on the global index)
        cfgd["samples"] = count
        r, _ = _render(cfgd, stats_clamp=g["clamp"], seed=77, sample_offset=first)
        parts.append(r.stats() + (r.framebuffer,))
        r.close()
    s, q, sums = (parts[0][k] + parts[1][k] for k in range(3))
    cfgd["samples"] = n_gpu
    r2, _ = _render(cfgd, stats_clamp=1e30, seed=77)  # the unclamped second moment, same samples
    _, q_un = r2.stats()
    np.testing.assert_allclose(r2.framebuffer, sums, rtol=2e-4, atol=1e-4)  # fp32 sums of 4096 samples, two orders
    r2.close()
    rep = parity_protocol.evaluate(g, (parts[0][0], parts[0][1], n1), (s, q, sums, n_gpu), q_un)
    parity_protocol.record(REPORT_DIR, "gpu_vs_reference", tag, rep)


def _many_lights_scene(n_lights, n_spheres):
    """A floor, `n_lights` small rect/sphere lights above it and `n_spheres` diffuse / glass / metal spheres: with more
    than 8 emitters NEE falls back to the plain closest-hit shadow query, with more than 16 instances to the BVH."""
    rng = np.random.Generator(np.random.PCG64(11))
    mats = [{"id": "white", "type": "lambertian", "data": {"color": [0.7, 0.7, 0.7]}},
            {"id": "blue", "type": "lambertian", "data": {"color": [0.2, 0.3, 0.8]}},
            {"id": "glass", "type": "dielectric", "data": {"ior": 1.5}},
            {"id": "steel", "type": "metal", "data": {"color": [0.8, 0.8, 0.9]}},
            {"id": "light", "type": "diffuse_light", "data": {"color": [12.0, 11.0, 9.0]}}]
    inst = [{"type": "direct", "primitive": {"type": "rect", "material": {"id": "white"}, "size": [60, 60]}}]
    for k in range(n_lights):
        x, z = (k % 4) * 10.0 - 15.0, (k // 4) * 10.0 - 10.0
        prim = ({"type": "rect", "material": {"id": "light"}, "size": [3, 3]} if k % 2 == 0 else
                {"type": "sphere", "material": {"id": "light"}, "radius": 1.0})
        inst.append({"type": "direct", "primitive": prim, "transform": {"translate": [x, 12.0 + (k % 3), z]}})
    for k in range(n_spheres):
        p = rng.uniform([-20, 1.5, -15], [20, 1.5, 15])
        inst.append({"type": "direct", "primitive": {"type": "sphere", "material": {"id": ["blue", "glass", "steel"][k % 3]},
                                                     "radius": 1.5},
                     "transform": {"translate": [float(p[0]), float(p[1]), float(p[2])]}})
    return {"camera": {"look_from": [0.0, 18.0, -45.0], "look_at": [0.0, 3.0, 0.0], "fov": 40.0, "aperture": 0.0,
                       "dist_to_focus": 40.0},
            "world": {"color": [0.02, 0.02, 0.03]}, "assets": [], "textures": [], "materials": mats, "primitives": [],
            "instances": inst}


@pytest.mark.parametrize("n_lights,n_spheres,over", [(12, 3, dict()), (12, 20, dict()), (3, 20, dict(light_samples=2)),
                                                     (9, 0, dict(light_samples=5, max_bounces=3))])
def test_many_lights_and_mid_size_scenes(tmp_path, n_lights, n_spheres, over):
    """The kernel variants the shipped scenes do not reach: > 8 emitters (closest-hit shadow query, `ANY` = false) on
    the small-scene table and on the BVH, and a BVH scene with few emitters; sphere AND rect lights; light_samples
    that are not a power of two (atomics instead of the butterfly).  GPU against the CPU restatement."""
    path = str(tmp_path / "many_lights.json")
    with open(path, "w") as f:
        json.dump(_many_lights_scene(n_lights, n_spheres), f)
    import pathtrace_b200 as pt
    w, h, spp = 48, 32, 6
    cfgd = pt.default_config_dict(width=w, height=h, samples=spp, scene=path, **over)
    r, rep = _render(cfgd, batch_samples=4)
    ref = PortOracle.from_file(path, w, h).render(make_cconfig(cfgd), stats=False)
    _close(r.framebuffer, ref["sums"], f"{n_lights} lights, {n_spheres} spheres")
    for k in ("rays_extension", "rays_shadow", "rays_reference"):
        assert abs(rep[k] - ref[k]) <= 2e-3 * ref[k] + 8, (k, rep[k], ref[k])
    assert r.framebuffer.max() > 0


def test_degenerate_films_and_settings():
    """1 x 1 and odd-sized films, one sample, light_samples = 1, max_bounces = 1, a sample offset beyond 2^31;
    light_samples = 0 is refused (the reference divides by it, integrator.h:265)."""
    import pathtrace_b200 as pt
    with pytest.raises(Exception, match="light_samples"):
        _render(_cfg("cornell_box", 3, 5, 2, light_samples=0))
    for (w, h, spp, over) in [(1, 1, 1, dict()), (3, 5, 2, dict(light_samples=2)), (7, 2, 3, dict(light_samples=1, max_bounces=1)),
                              (5, 3, 2, dict(russian_roulette=False, max_bounces=2))]:
        cfgd = _cfg("cornell_box", w, h, spp, **over)
        r, rep = _render(cfgd)
        ref = PortOracle.from_file(scene_path("cornell_box"), w, h).render(make_cconfig(cfgd), stats=False)
        assert rep["camera_rays"] == w * h * spp
        _close(r.framebuffer, ref["sums"], (w, h, spp, over))
    cfgd = _cfg("cornell_box", 8, 8, 2)
    far = _render(cfgd, sample_offset=(1 << 33) + 5)[0].framebuffer
    near = _render(cfgd, sample_offset=5)[0].framebuffer
    assert np.isfinite(far).all() and far.max() > 0 and far.shape == near.shape
    assert not np.array_equal(far, near)  # the high word of the global sample index is part of the Philox counter
    ref = PortOracle.from_file(scene_path("cornell_box"), 8, 8).render(make_cconfig(cfgd, sample_offset=(1 << 33) + 5),
                                                                       stats=False)
    _close(far, ref["sums"], "offset 2^33 + 5")
    with pytest.raises(Exception, match="sample_offset"):
        _render(cfgd, sample_offset=1 << 56)
    with pytest.raises(Exception, match="sample_offset"):
        _render(cfgd, sample_offset=-1)


@pytest.mark.parametrize("scene", ["book2_final", "cornell_box", "light_test"])
def test_horizon_skip_keeps_ray_accounting_and_film(scene):
    """k_shade leaves vertices whose lights all lie below their horizon out of the NEE queue (pt_shade.h
    lights_below_horizon).  Their shadow queries must still be counted -- in the report (`rays_shadow`, compared with
    the CPU restatement, which has no such shortcut) and in the work counters of the profiling build -- and the
    film must be the one the restatement computes without the shortcut."""
    cfgd = _cfg(scene, 96, 54, 8)
    r, rep = _render(cfgd, profile=2)
    prof = r.profile()
    assert prof["nee_rays"] == rep["rays_shadow"]
    assert prof["ext_rays"] == rep["rays_extension"]
    oracle = PortOracle.from_file(scene_path(scene), 96, 54)
    ref = oracle.render(make_cconfig(cfgd), stats=False)
    # a differing libm ulp can flip a Russian-roulette or medium decision on a handful of paths (as in
    # test_render_matches_cpu_restatement); a vertex dropped without being counted would be a few per cent
    for k in ("rays_extension", "rays_shadow", "rays_reference"):
        assert abs(rep[k] - ref[k]) <= 2e-3 * ref[k] + 8, (k, rep[k], ref[k])
    _close(r.framebuffer, ref["sums"], scene)
