"""BASELINE.json's FULL film sizes on the GPU -- all five configs, from three_orbs 400 x 225 to the north-star target book2_final
3840 x 2160, at the shipped integrator settings -- checked through properties that do not depend on the size, because the CPU restatement
cannot render 8.3 M pixels in seconds:

  * a band of pixel rows of the full-size film against the CPU restatement, sample sum for sample sum (the restatement renders
    only that band: the Philox key of a sample is (pixel, global sample), whatever else is in flight);
  * batching invariance: one sample per wavefront batch vs all of them in one batch, bit-equal films (32 Mi slots in flight
    against 8 Mi: different queue lengths, sort histograms and launch geometry);
  * sharding: two halves of the global sample range (sample_offset) add up to the whole -- the sum multi-GPU renders rely on --
    and so do their ray counters, exactly (a checksum of checksums);
  * ray accounting: every camera sample is counted once, extension rays >= camera rays, and the reference's count
    (integrator.h:192,247) covers extension + shadow.
"""
import numpy as np
import pytest

from conftest import scene_path
from oracle_lib import PortOracle, make_cconfig
from test_gpu_render import _cfg, _close, _render

pytestmark = pytest.mark.gpu

FULL = [
    # scene, width, height, samples, rows of the band checked against the restatement
    ("book2_final", 3840, 2160, 4, 2),      # BASELINE.json configs[4], the north-star target
    ("cornell_box", 1024, 1024, 8, 4),      # configs[1]
    ("cornell_box_small_lights", 1024, 1024, 4, 4),  # configs[2]
    ("cornell_box_with_volume", 1024, 1024, 4, 4),   # configs[3]
    ("three_orbs", 400, 225, 8, 8),         # configs[0]
]


@pytest.mark.parametrize("scene,w,h,spp,band_rows", FULL)
def test_full_size_film_properties(scene, w, h, spp, band_rows):
    cfgd = _cfg(scene, w, h, spp)
    r_all, rep_all = _render(cfgd, batch_samples=spp)
    whole = r_all.framebuffer.copy()
    assert whole.shape == (h, w, 3) and np.isfinite(whole).all()
    assert rep_all["camera_rays"] == w * h * spp
    assert rep_all["rays_extension"] >= rep_all["camera_rays"]
    # the reference counts light_samples per vertex it reaches, the shadow counter only vertices that ask for light samples
    assert rep_all["rays_reference"] >= rep_all["rays_extension"] + rep_all["rays_shadow"]

    # batching invariance at full size
    r_one, rep_one = _render(cfgd, batch_samples=1)
    assert np.array_equal(r_one.framebuffer, whole)
    for k in ("rays_extension", "rays_shadow", "rays_reference"):
        assert rep_one[k] == rep_all[k], (k, rep_one[k], rep_all[k])
    del r_one

    # two shards of the global sample range add up to the whole: films to fp32 summation order, counters exactly
    half = dict(cfgd, samples=spp // 2)
    r0, rep0 = _render(half, sample_offset=0)
    r1, rep1 = _render(half, sample_offset=spp // 2)
    np.testing.assert_allclose(r0.framebuffer + r1.framebuffer, whole, rtol=1e-5, atol=1e-6)
    for k in ("rays_extension", "rays_shadow", "rays_reference", "camera_rays"):
        assert rep0[k] + rep1[k] == rep_all[k], (k, rep0[k], rep1[k], rep_all[k])
    s_whole = whole.astype(np.float64).sum()
    s_parts = r0.framebuffer.astype(np.float64).sum() + r1.framebuffer.astype(np.float64).sum()
    assert abs(s_whole - s_parts) <= 1e-6 * abs(s_whole)
    del r0, r1

    # a band of rows in the middle of the film against the CPU restatement (same Philox streams)
    row0 = h // 2
    p0, p1 = row0 * w, (row0 + band_rows) * w
    oracle = PortOracle.from_file(scene_path(scene), w, h)
    ref = oracle.render(make_cconfig(cfgd), stats=False, pixel_begin=p0, pixel_end=p1)
    got = whole.reshape(-1, 3)[p0:p1]
    exp = ref["sums"].reshape(-1, 3)[p0:p1]
    assert float(np.abs(exp).max()) > 0
    _close(got, exp, f"{scene} {w}x{h} rows {row0}..{row0 + band_rows}", level="loose")
