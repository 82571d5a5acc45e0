"""Test-side access to the oracles (TEST INFRASTRUCTURE: nothing in pathtrace_b200/ imports this).

  * RefDriver   -- oracle/_ref/ref_driver: the UNMODIFIED reference compiled from /root/reference (prebuilt binary
                   on the GPU box); modes scene / hit / color.
  * PortOracle  -- oracle/_build/libpt_oracle.so: the scalar CPU restatement sharing the Philox keying of the kernels.
"""
import ctypes as C
import json
import os
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
PORT_LIB = os.path.join(ROOT, "oracle", "_build", "libpt_oracle.so")
SCENES = os.path.join(ROOT, "scenes")

HIT_DTYPE = np.dtype([("instance", "<i4"), ("t", "<f4"), ("p", "<f4", 3), ("n", "<f4", 3), ("u", "<f4"), ("v", "<f4")])


def have_ref_driver():
    return os.path.exists(REF_DRIVER) and os.access(REF_DRIVER, os.X_OK)


def have_port():
    return os.path.exists(PORT_LIB)


class RefDriver:
    """Runs the reference harness as a subprocess (it is single-threaded and deterministic)."""

    @staticmethod
    def scene(scene_path, w, h):
        with tempfile.TemporaryDirectory() as d:
            out = os.path.join(d, "scene.json")
            subprocess.run([REF_DRIVER, "scene", scene_path, str(w), str(h), out], check=True, cwd=ROOT,
                           stdout=subprocess.DEVNULL)
            with open(out) as f:
                return json.load(f)

    @staticmethod
    def hit(scene_path, rays, burn=0):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        with tempfile.TemporaryDirectory() as d:
            rin, rout = os.path.join(d, "rays.f32"), os.path.join(d, "hits.bin")
            rays.tofile(rin)
            subprocess.run([REF_DRIVER, "hit", scene_path, rin, rout, str(burn)], check=True, cwd=ROOT,
                           stdout=subprocess.DEVNULL)
            return np.fromfile(rout, dtype=HIT_DTYPE)

    @staticmethod
    def color(config_dict, nsamples, clamp, burn=0, timeout=None):
        """Per-pixel statistics of integrator->color: returns dict(mean, cmean, cm2, total_rays), arrays H x W x 3
        with row 0 = image bottom."""
        with tempfile.TemporaryDirectory() as d:
            cfg, out = os.path.join(d, "config.json"), os.path.join(d, "color.f64")
            with open(cfg, "w") as f:
                json.dump(config_dict, f)
            subprocess.run([REF_DRIVER, "color", cfg, out, str(nsamples), repr(float(clamp)), str(burn)], check=True,
                           cwd=ROOT, stdout=subprocess.DEVNULL, timeout=timeout)
            raw = np.fromfile(out, dtype=np.float64)
        w, h, n = int(raw[0]), int(raw[1]), int(raw[2])
        body = raw[5:].reshape(h, w, 9)
        return {"w": w, "h": h, "n": n, "clamp": raw[3], "total_rays": int(raw[4]), "mean": body[..., 0:3],
                "cmean": body[..., 3:6], "cm2": body[..., 6:9]}


    @staticmethod
    def film(fb_sums, n_samples, exposure):
        """The reference's own film path on a supplied framebuffer of sums (H x W x 3, row 0 = bottom):
        calculate_luminance (helpers.h:146-168) + output_to_file (renderer.h:24-55).  Returns (image H x W x 3 uint8
        with the TOP row first, max_luminance, avg_luminance, the file's bytes)."""
        fb = np.ascontiguousarray(fb_sums, dtype=np.float32)
        h, w, _ = fb.shape
        with tempfile.TemporaryDirectory() as d:
            fin, ppm, js = os.path.join(d, "fb.f32"), os.path.join(d, "out.ppm"), os.path.join(d, "out.json")
            fb.tofile(fin)
            subprocess.run([REF_DRIVER, "film", fin, str(w), str(h), str(int(n_samples)), repr(float(exposure)), ppm, js],
                           check=True, cwd=ROOT, stdout=subprocess.DEVNULL)
            with open(ppm, "rb") as f:
                data = f.read()
            with open(js) as f:
                j = json.load(f)
        header = f"P6\n{w} {h}\n255\n".encode()
        assert data.startswith(header) and len(data) == len(header) + w * h * 3, (len(data), len(header) + w * h * 3)
        img = np.frombuffer(data[len(header):], dtype=np.uint8).reshape(h, w, 3)
        return img, j["max_luminance"], j["avg_luminance"], data

    @staticmethod
    def project(scene_path, w, h, paths):
        """camera::project (camera.h:49-109) per point and handle_traced_paths (renderer.h:57-96) on the supplied paths
        (a list of (k, 3) arrays).  Returns dict(projected = (n, 3) [x, y, ok], text3d, text2d, added_paths)."""
        with tempfile.TemporaryDirectory() as d:
            pin, o3, o2, js = (os.path.join(d, n) for n in ("paths.txt", "o3.txt", "o2.txt", "out.json"))
            with open(pin, "w") as f:
                for p in paths:
                    for v in np.asarray(p, dtype=np.float32).reshape(-1, 3):
                        f.write("%.9g %.9g %.9g\n" % (v[0], v[1], v[2]))
                    f.write("\n")
            subprocess.run([REF_DRIVER, "project", scene_path, str(w), str(h), pin, o3, o2, js], check=True, cwd=ROOT,
                           stdout=subprocess.DEVNULL)
            with open(js) as f:
                j = json.load(f)
            return {"projected": np.array(j["projected"], dtype=np.float64).reshape(-1, 3), "text3d": open(o3).read(),
                    "text2d": open(o2).read(), "added_paths": j["added_paths"]}


class _CConfig(C.Structure):  # mirrors b200pt_config (include/b200pt.h)
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("exposure", C.c_float), ("gamma", C.c_float),
                ("integrator", C.c_int32), ("max_bounces", C.c_int32), ("samples", C.c_int32),
                ("light_samples", C.c_int32), ("russian_roulette", C.c_int32),
                ("only_direct_illumination", C.c_int32), ("normal_offset", C.c_float),
                ("block_width", C.c_int32), ("block_height", C.c_int32), ("threads", C.c_int32),
                ("seed", C.c_uint64), ("sample_offset", C.c_int64), ("device", C.c_int32),
                ("batch_samples", C.c_int32), ("stats_clamp", C.c_float), ("profile", C.c_int32),
                ("trace_probability", C.c_float), ("trace_max_paths", C.c_int32),
                ("physically_corrected", C.c_int32), ("device_count", C.c_int32), ("device_ids", C.c_int32 * 8),
                ("async_render", C.c_int32)]


INTEGRATORS = {"iterative nee path tracing": 4, "recursive path tracing": 0}


def make_cconfig(config_dict, **over):
    """config.json dict -> the C struct, with the reference's defaults (config.h:98-132), without libb200pt."""
    film = config_dict.get("film", {})
    c = _CConfig()
    c.width = int(film.get("width", 400))
    c.height = int(film.get("height", 300))
    c.exposure = float(film.get("gamma", 2.2))
    c.gamma = float(film.get("exposure", 0.0))
    c.integrator = INTEGRATORS.get(config_dict.get("integrator_type", "recursive path tracing"), 0)
    c.max_bounces = int(config_dict.get("max_bounces", 10))
    c.samples = int(config_dict.get("samples", 20))
    c.light_samples = int(config_dict.get("light_samples", 1))
    c.russian_roulette = int(bool(config_dict.get("russian_roulette", True)))
    c.only_direct_illumination = int(bool(config_dict.get("only_direct_illumination", False)))
    c.normal_offset = float(config_dict.get("normal_offset", 0.0001))
    c.seed = int(config_dict.get("seed", 0))
    c.physically_corrected = int(bool(config_dict.get("physically_corrected", False)))
    for k, v in over.items():
        setattr(c, k, v)
    return c


class PortOracle:
    def __init__(self, scene_json_text, w, h, asset_dir=None):
        self.lib = C.CDLL(PORT_LIB)
        self.lib.pt_oracle_scene_create.restype = C.c_void_p
        self.lib.pt_oracle_scene_create.argtypes = [C.c_char_p, C.c_int, C.c_int]
        self.lib.pt_oracle_scene_destroy.argtypes = [C.c_void_p]
        self.lib.pt_oracle_last_error.restype = C.c_char_p
        self.lib.pt_oracle_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        self.lib.pt_oracle_trace_shadow.argtypes = self.lib.pt_oracle_trace.argtypes
        self.lib.pt_oracle_trace_linear.argtypes = self.lib.pt_oracle_trace.argtypes
        self.lib.pt_oracle_emitters.argtypes = [C.c_void_p, C.c_void_p]
        self.lib.pt_oracle_render.argtypes = [C.c_void_p, C.POINTER(_CConfig), C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.POINTER(C.c_int64), C.c_int, C.c_int64, C.c_int64]
        self.lib.pt_oracle_scene_create2.restype = C.c_void_p
        self.lib.pt_oracle_scene_create2.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_char_p]
        self.h = self.lib.pt_oracle_scene_create2(scene_json_text.encode(), w, h,
                                                  asset_dir.encode() if asset_dir else None)
        if not self.h:
            raise RuntimeError("pt_oracle: " + self.lib.pt_oracle_last_error().decode())
        self.w, self.hgt = w, h

    @classmethod
    def from_file(cls, path, w, h):
        with open(path) as f:
            return cls(f.read(), w, h, asset_dir=os.path.dirname(os.path.abspath(path)))

    def close(self):
        if self.h:
            self.lib.pt_oracle_scene_destroy(self.h)
            self.h = None

    def trace(self, rays, medium_xi=None):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
        xi = None if medium_xi is None else np.ascontiguousarray(medium_xi, dtype=np.float32)
        self.lib.pt_oracle_trace(self.h, rays.ctypes.data, rays.shape[0], None if xi is None else xi.ctypes.data,
                                 hits.ctypes.data)
        return hits

    def trace_shadow(self, rays, medium_xi=None):
        """The NEE shadow query (emitters first, then any-hit): a record only when an emitter is met first."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
        xi = None if medium_xi is None else np.ascontiguousarray(medium_xi, dtype=np.float32)
        self.lib.pt_oracle_trace_shadow(self.h, rays.ctypes.data, rays.shape[0],
                                        None if xi is None else xi.ctypes.data, hits.ctypes.data)
        return hits

    def trace_linear(self, rays, medium_xi=None):
        """World::hit without any tree: every instance tested in index order."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
        xi = None if medium_xi is None else np.ascontiguousarray(medium_xi, dtype=np.float32)
        self.lib.pt_oracle_trace_linear(self.h, rays.ctypes.data, rays.shape[0],
                                        None if xi is None else xi.ctypes.data, hits.ctypes.data)
        return hits

    def horizon_check(self, points, normals, samples=16):
        """(vertices the horizon test rejects, their samples that are not void [must be 0], kept vertices whose
        samples are all void) -- pt_shade.h lights_below_horizon against nee_prepare."""
        pn = np.ascontiguousarray(np.concatenate([points, normals], axis=1), dtype=np.float32)
        out = np.zeros(3, dtype=np.uint64)
        self.lib.pt_oracle_horizon_check.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
        self.lib.pt_oracle_horizon_check(self.h, pn.ctypes.data, pn.shape[0], samples, out.ctypes.data)
        return int(out[0]), int(out[1]), int(out[2])

    def emitters(self):
        n = self.lib.pt_oracle_emitters(self.h, None)
        flags = np.zeros(n, dtype=np.uint8)
        self.lib.pt_oracle_emitters(self.h, flags.ctypes.data)
        return flags.astype(bool)

    def render(self, cconfig, threads=None, pixel_begin=0, pixel_end=0, stats=True):
        h, w = cconfig.height, cconfig.width
        sums = np.zeros((h, w, 3), dtype=np.float32)
        cs = np.zeros((h, w, 3), dtype=np.float32)
        cq = np.zeros((h, w, 3), dtype=np.float32)
        counts = (C.c_int64 * 3)()
        threads = threads or os.cpu_count() or 1
        self.lib.pt_oracle_render(self.h, C.byref(cconfig), sums.ctypes.data, cs.ctypes.data if stats else None,
                                  cq.ctypes.data if stats else None, counts, threads, pixel_begin, pixel_end)
        return {"sums": sums, "csum": cs, "csq": cq, "rays_extension": counts[0], "rays_shadow": counts[1],
                "rays_reference": counts[2]}
