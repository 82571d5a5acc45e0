"""Seeded ray batches for the closest-hit parity check (parity check 1)."""
import numpy as np

TMIN = 0.001
FLT_MAX = np.float32(3.402823466e+38)


def camera_rays(cam, w, h):
    """Pixel-centre camera rays (camera.h:38-47 with aperture 0)."""
    i, j = np.meshgrid(np.arange(w), np.arange(h))
    u = ((i + 0.5) / w).astype(np.float32).ravel()
    v = ((j + 0.5) / h).astype(np.float32).ravel()
    o = np.asarray(cam["origin"], dtype=np.float32)
    ll = np.asarray(cam["lower_left_corner"] if "lower_left_corner" in cam else cam["lower_left"], dtype=np.float32)
    hor = np.asarray(cam["horizontal"], dtype=np.float32)
    ver = np.asarray(cam["vertical"], dtype=np.float32)
    d = ll[None, :] + u[:, None] * hor[None, :] + v[:, None] * ver[None, :] - o[None, :]
    rays = np.zeros((w * h, 8), dtype=np.float32)
    rays[:, 0:3] = o
    rays[:, 3:6] = d
    rays[:, 6] = TMIN
    rays[:, 7] = FLT_MAX
    return rays


def random_rays(bounds_lo, bounds_hi, n, seed):
    """Origins uniform in the scene box inflated by 10 %, directions uniform on the sphere with log-uniform length
    in [0.1, 1000] (the reference keeps directions un-normalised, so t scales with 1/|d|)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    lo = np.asarray(bounds_lo, dtype=np.float64)
    hi = np.asarray(bounds_hi, dtype=np.float64)
    c, e = (lo + hi) / 2, (hi - lo) / 2 * 1.1
    o = c + (rng.random((n, 3)) * 2 - 1) * e
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d *= np.exp(rng.uniform(np.log(0.1), np.log(1000.0), size=(n, 1)))
    rays = np.zeros((n, 8), dtype=np.float32)
    rays[:, 0:3] = o
    rays[:, 3:6] = d
    rays[:, 6] = TMIN
    rays[:, 7] = FLT_MAX
    return rays


def secondary_rays(hits, n, seed, offset=1e-4):
    """Rays that start ON surfaces (previous hit points pushed along the normal), cosine-ish directions."""
    rng = np.random.Generator(np.random.PCG64(seed))
    ok = np.nonzero(hits["instance"] >= 0)[0]
    if len(ok) == 0:
        return np.zeros((0, 8), dtype=np.float32)
    pick = ok[rng.integers(0, len(ok), size=n)]
    p = hits["p"][pick].astype(np.float64)
    nrm = hits["n"][pick].astype(np.float64)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d + nrm * 1.001  # biased towards the normal side, still covers grazing directions
    sign = np.where(rng.random(n) < 0.2, -1.0, 1.0)[:, None]  # 20 % go through the surface (insides of boxes/spheres)
    rays = np.zeros((n, 8), dtype=np.float32)
    rays[:, 0:3] = p + offset * nrm * sign
    rays[:, 3:6] = d * sign
    rays[:, 6] = TMIN
    rays[:, 7] = FLT_MAX
    return rays


def shadow_rays(hits, light_points, n, seed):
    """Shadow-style rays: from hit points (un-offset, integrator.h:235) to sampled light points, un-normalised."""
    rng = np.random.Generator(np.random.PCG64(seed))
    ok = np.nonzero(hits["instance"] >= 0)[0]
    if len(ok) == 0 or len(light_points) == 0:
        return np.zeros((0, 8), dtype=np.float32)
    pick = ok[rng.integers(0, len(ok), size=n)]
    lp = light_points[rng.integers(0, len(light_points), size=n)]
    p = hits["p"][pick]
    rays = np.zeros((n, 8), dtype=np.float32)
    rays[:, 0:3] = p
    rays[:, 3:6] = lp - p
    rays[:, 6] = TMIN
    rays[:, 7] = FLT_MAX
    return rays


def compare_hits(ref, got, exclude_instances=()):
    """Agreement statistics between two HIT_DTYPE arrays whose instance ids are in the SAME numbering."""
    n = len(ref)
    keep = np.ones(n, dtype=bool)
    for e in exclude_instances:
        keep &= (ref["instance"] != e) & (got["instance"] != e)
    # A ray lying IN a rect's plane gives t = 0/0 = NaN, which passes `t < t0 || t > t1` (primitive.h:196) in the
    # reference and here alike.  With a NaN record in play the reference's winner depends on the SHAPE of its tree
    # (`left.t < right.t ? left : right`, bvh.h:39-48, is not associative over NaN), so such rays are left out of
    # the id statistics; they carry no radiance either way (cos = 0, pdf = NaN -> the sample is dropped,
    # integrator.h:255-262).  They only arise when origin and direction are exactly in the plane.
    nan_any = (np.isnan(ref["t"]) & (ref["instance"] >= 0)) | (np.isnan(got["t"]) & (got["instance"] >= 0))
    keep &= ~nan_any
    same = ref["instance"] == got["instance"]
    both = same & (ref["instance"] >= 0) & keep
    n_nan_mismatch = int(nan_any.sum())
    t_ref = ref["t"][both].astype(np.float64)
    dt = np.abs(got["t"][both].astype(np.float64) - t_ref) / np.maximum(np.abs(t_ref), 1e-30)
    dn = np.linalg.norm(got["n"][both].astype(np.float64) - ref["n"][both].astype(np.float64), axis=1)
    pr = ref["p"][both].astype(np.float64)
    dp = np.linalg.norm(got["p"][both].astype(np.float64) - pr, axis=1) / np.maximum(np.linalg.norm(pr, axis=1), 1e-30)
    return {
        "n": int(keep.sum()),
        "id_agree": float((same & keep).sum() / max(1, keep.sum())),
        "n_hit": int(both.sum()),
        "nan_rays": n_nan_mismatch,
        "t_rel_max": float(dt.max()) if len(dt) else 0.0,
        "t_rel_p9999": float(np.quantile(dt, 0.9999)) if len(dt) else 0.0,
        "t_bad": float((dt > 1e-5).mean()) if len(dt) else 0.0,
        "n_err_max": float(dn.max()) if len(dn) else 0.0,
        "n_bad": float((dn > 1e-5).mean()) if len(dn) else 0.0,
        "p_rel_max": float(dp.max()) if len(dp) else 0.0,
    }
