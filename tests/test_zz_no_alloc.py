"""b200pt_tonemap_static: the allocation-free tonemap (no counterpart in the reference; added
for callers that cannot call cudaMalloc).  Its luminance-reduce scratch is a file-scope `static __device__` array
of the library, shared by every call and keeping the previous call's contents -- so the check is made on TWO calls in a
row, a large film and then a small one (the second call must not see the partials the first left behind), each against the
existing entry point b200pt_render_tonemap on the same film.

Collected last on purpose.  Passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from conftest import scene_path

pytestmark = pytest.mark.gpu

MAX_PIXELS = 1024 * 16384  # PT_STATIC_MAX_PIXELS (include/b200pt.h: 16 777 216)


def _rendered(w, h, spp, scene="cornell_box"):
    import pathtrace_b200 as pt
    cfg = pt.Config.from_json(pt.default_config_dict(width=w, height=h, samples=spp, scene=scene_path(scene)))
    sc = pt.Scene.from_file(cfg.scene_path, w, h, device=0)
    r = pt.renderer_from_config(sc, cfg).start_render()
    r.finalize()
    return sc, r


def _static_tonemap(r, n_samples, out, want_lum=True):
    """The new entry point on the render's own device film buffer; `out` is a torch uint8 CUDA tensor of H*W*3 bytes."""
    import torch
    import pathtrace_b200 as pt
    from pathtrace_b200._capi import check
    d_film, nbytes = r.device_buffer()
    w, h = r.config.width, r.config.height
    assert nbytes == w * h * 16 and out.numel() == w * h * 3
    mx, avg = C.c_float(-1.0), C.c_float(-1.0)
    check(pt.lib().b200pt_tonemap_static(0, d_film, w, h, n_samples, r.config.exposure, out.data_ptr(), None,
                                         C.byref(mx) if want_lum else None, C.byref(avg) if want_lum else None))
    torch.cuda.synchronize()
    return out.cpu().numpy().reshape(h, w, 3), mx.value, avg.value


def _compare(r, n_samples, out):
    img_ref, mx_ref, avg_ref = r.tonemap(n_samples)  # the existing entry point (allocates its scratch per render)
    img, mx, avg = _static_tonemap(r, n_samples, out)
    assert mx == mx_ref  # a maximum does not depend on the order
    assert abs(avg - avg_ref) <= 1e-4 * abs(avg_ref) + 1e-12, (avg, avg_ref)  # float sum, other order
    assert np.array_equal(img, img_ref)  # same white point, same arithmetic
    assert 5 < img.mean() < 250
    return img


def test_static_tonemap_matches_render_tonemap_on_two_calls_in_a_row():
    import torch
    big_scene, big = _rendered(640, 360, 6)        # 225 tiles of 1024 pixels
    small_scene, small = _rendered(96, 64, 4)      # 6 tiles: cells 6 .. 224 still hold the first call's partials
    out_big = torch.zeros(640 * 360 * 3, dtype=torch.uint8, device="cuda")
    out_small = torch.zeros(96 * 64 * 3, dtype=torch.uint8, device="cuda")
    first = _compare(big, 6, out_big)
    _compare(small, 4, out_small)
    # ... and back: the same answer as the first time, bit for bit (fixed summation order), with another divisor too
    again = _static_tonemap(big, 6, out_big)
    assert np.array_equal(again[0], first)
    _compare(big, 3, out_big)  # sync_progress's divisor 1 + done / pixels (renderer.h:611-616)
    # without the luminance outputs nothing is copied back; the image is the same
    out_big.zero_()
    img, mx, avg = _static_tonemap(big, 6, out_big, want_lum=False)
    assert np.array_equal(img, first) and mx == -1.0 and avg == -1.0
    for h in (big, small):
        h.close()
    for s in (big_scene, small_scene):
        s.close()


def test_static_tonemap_refuses_films_beyond_its_scratch():
    import pathtrace_b200 as pt
    lib = pt.lib()
    dummy = C.c_void_p(256)  # never dereferenced: the size is checked first
    rc = lib.b200pt_tonemap_static(0, dummy, 8192, 4096, 1, 2.2, dummy, None, None, None)  # 33.5 M pixels
    assert rc != 0
    msg = lib.b200pt_last_error().decode()
    assert "exceeds" in msg and str(MAX_PIXELS) in msg, msg
    assert lib.b200pt_tonemap_static(0, None, 16, 16, 1, 2.2, dummy, None, None, None) != 0
    assert lib.b200pt_tonemap_static(0, dummy, 16, 16, 0, 2.2, dummy, None, None, None) != 0
    # exactly at the limit the size check passes (4096 x 4096); a film that small in pixels is rendered for real above
    assert 4096 * 4096 == MAX_PIXELS
