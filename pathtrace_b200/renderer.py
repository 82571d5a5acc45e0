"""Host-side mirror of the reference's interface for this path, over the C ABI.

    reference (C++)                               here
    -------------------------------------------  ------------------------------------------
    Config(json)                config.h:98-132   Config.from_json / Config.from_file
    build_scene + setup_camera  scene_parser.h:241, main.cpp:86   Scene.from_json / Scene.from_file
    World::hit                  world.h:17-20     Scene.trace_batch (n rays at once)
    renderer_from_config        main.cpp:57-84    Wavefront(scene, config)
    Renderer::start_render      renderer.h:553    Wavefront.start_render
    Renderer::sync_progress     renderer.h:605    Wavefront.sync_progress
    Renderer::is_done           renderer.h:622    Wavefront.is_done
    Renderer::finalize          renderer.h:693    Wavefront.finalize
    Renderer::framebuffer       renderer.h:141    Wavefront.framebuffer  (H x W x 3 sums, row 0 = bottom)

Same names, same argument meaning; errors are exceptions (B200ptError) instead of assert/terminate.
All compute happens in libb200pt.so on the GPU.
"""
import ctypes as C
import json
import os

import numpy as np

from . import _capi
from ._capi import B200ptError, Config as _CConfig, check, lib


class Config:
    """config.json -> render configuration (same keys and defaults as config.h:98-132)."""

    def __init__(self, cconfig, scene_path="", ppm_output_path="out.ppm", traced_paths_output_path="",
                 traced_paths_2d_output_path=""):
        self.c = cconfig
        self.scene_path = scene_path
        self.ppm_output_path = ppm_output_path
        self.traced_paths_output_path = traced_paths_output_path
        self.traced_paths_2d_output_path = traced_paths_2d_output_path

    @classmethod
    def from_json(cls, text):
        if isinstance(text, dict):
            text = json.dumps(text)
        c = _CConfig()
        scene = C.create_string_buffer(1024)
        ppm = C.create_string_buffer(1024)
        check(lib().b200pt_config_from_json(text.encode(), C.byref(c), scene, 1024, ppm, 1024))
        extra = []
        for key in ("traced_paths_output_path", "traced_paths_2d_output_path"):
            buf = C.create_string_buffer(1024)
            check(lib().b200pt_config_string(text.encode(), key.encode(), buf, 1024))
            extra.append(buf.value.decode())
        return cls(c, scene.value.decode(), ppm.value.decode(), *extra)

    @classmethod
    def from_file(cls, path="config.json"):
        with open(path) as f:
            return cls.from_json(f.read())

    def __getattr__(self, name):
        return getattr(self.c, name)

    def replace(self, **kw):
        """A copy with fields of b200pt_config replaced; `devices=[0, 1, ...]` sets device_count / device_ids / device."""
        c = _CConfig()
        C.memmove(C.byref(c), C.byref(self.c), C.sizeof(c))
        devices = kw.pop("devices", None)
        if devices is not None:
            devices = list(devices)
            if not 1 <= len(devices) <= 8:
                raise ValueError("1 to 8 devices")
            c.device = devices[0]
            c.device_count = len(devices) if len(devices) > 1 else 0
            for k, d in enumerate(devices):
                c.device_ids[k] = d
        for k, v in kw.items():
            setattr(c, k, v)
        return Config(c, self.scene_path, self.ppm_output_path, self.traced_paths_output_path,
                      self.traced_paths_2d_output_path)


def default_config_dict(**over):
    """The shipped config.json (config.json:1-26) with path tracing of debug paths off; override by keyword."""
    d = {
        "film": {"width": 600, "height": 600, "exposure": 0.0, "gamma": 2.2},
        "ppm_output_path": "output/render.ppm", "png_output_path": "output/render.png",
        "traced_paths_output_path": "output/out.txt", "traced_paths_2d_output_path": "output/out_2d.txt",
        "scene": "scenes/cornell_box.json", "render_type": "tiled",
        "integrator_type": "iterative nee path tracing", "should_trace_paths": False,
        "only_direct_illumination": False, "avg_number_of_paths": 100, "block_width": 128, "block_height": 128,
        "normal_offset": 0.0001, "max_bounces": 10, "samples": 20, "light_samples": 4, "russian_roulette": True,
        "threads": 10,
    }
    film = over.pop("film", None)
    if film:
        d["film"].update(film)
    for k in ("width", "height"):
        if k in over:
            d["film"][k] = over.pop(k)
    d.update(over)
    return d


class Scene:
    """A compiled, uploaded scene (immutable)."""

    def __init__(self, handle, device):
        self._h = handle
        self.device = device

    @classmethod
    def from_json(cls, text, film_width, film_height, device=0):
        if isinstance(text, dict):
            text = json.dumps(text)
        h = C.c_void_p()
        check(lib().b200pt_scene_create_from_json(text.encode(), film_width, film_height, device, C.byref(h)))
        return cls(h, device)

    @classmethod
    def from_file(cls, path, film_width, film_height, device=0):
        h = C.c_void_p()
        check(lib().b200pt_scene_create_from_file(os.fspath(path).encode(), film_width, film_height, device, C.byref(h)))
        return cls(h, device)

    @classmethod
    def from_desc(cls, desc, device=0):
        h = C.c_void_p()
        check(lib().b200pt_scene_create(C.byref(desc), device, C.byref(h)))
        return cls(h, device)

    def close(self):
        if self._h:
            lib().b200pt_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def desc(self):
        d = _capi.SceneDesc()
        check(lib().b200pt_scene_get_desc(self._h, C.byref(d)))
        return d

    def describe(self):
        """The flat description as plain Python data (copies)."""
        d = self.desc()
        inst = []
        for i in range(d.n_instances):
            s = d.instances[i]
            inst.append({"fwd": np.array(s.fwd[:], dtype=np.float32).reshape(3, 4), "prim_kind": s.prim_kind,
                         "plane": s.plane, "flags": s.flags, "material": s.material,
                         "params": np.array(s.params[:], dtype=np.float32), "density": s.density,
                         "light_index": s.light_index})
        mats = [{"kind": d.materials[i].kind, "texture": d.materials[i].texture,
                 "color": list(d.materials[i].color), "ior": d.materials[i].ior, "power": d.materials[i].power,
                 "two_sided": d.materials[i].two_sided} for i in range(d.n_materials)]
        tex = [{"kind": d.textures[i].kind, "color": list(d.textures[i].color), "alpha": d.textures[i].alpha,
                "scale": d.textures[i].scale, "odd": d.textures[i].odd, "even": d.textures[i].even}
               for i in range(d.n_textures)]
        cam = {k: np.array(getattr(d.camera, k)[:], dtype=np.float32)
               for k in ("origin", "lower_left", "horizontal", "vertical", "u", "v", "w")}
        cam["lens_radius"] = d.camera.lens_radius
        out = {"instances": inst, "materials": mats, "textures": tex, "lights": [d.lights[i] for i in range(d.n_lights)],
               "background": d.background, "camera": cam}
        out["images"] = [np.ctypeslib.as_array(d.images[i].rgba, shape=(d.images[i].height, d.images[i].width, 4)).copy()
                         for i in range(d.n_images)]
        if d.perlin_ranvec:
            out["perlin_ranvec"] = np.ctypeslib.as_array(d.perlin_ranvec, shape=(256, 3)).copy()
            out["perlin_perm"] = np.ctypeslib.as_array(d.perlin_perm, shape=(3, 256)).copy()
        return out

    def bvh_info(self):
        n, depth = C.c_int32(), C.c_int32()
        bounds = (C.c_float * 6)()
        check(lib().b200pt_scene_bvh_info(self._h, C.byref(n), C.byref(depth), bounds))
        return {"n_nodes": n.value, "max_depth": depth.value, "bounds": np.array(bounds[:], dtype=np.float32)}

    def trace_batch(self, rays, medium_xi=None):
        """World::hit for an (n, 8) float32 array {o, d, tmin, tmax}; returns a structured array (HIT_DTYPE)."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        n = rays.shape[0]
        hits = np.zeros(n, dtype=_capi.HIT_DTYPE)
        xi = None
        if medium_xi is not None:
            xi = np.ascontiguousarray(medium_xi, dtype=np.float32)
            assert xi.shape == (n,)
        check(lib().b200pt_trace_batch(self._h, rays.ctypes.data, n, xi.ctypes.data if xi is not None else None,
                                       hits.ctypes.data))
        return hits

    def project(self, points):
        """camera::project (camera.h:49-109): (n, 3) world points -> ((n, 2) film coordinates, (n,) bool return values)."""
        pts = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
        xy = np.zeros((pts.shape[0], 2), dtype=np.float32)
        ok = np.zeros(pts.shape[0], dtype=np.int32)
        check(lib().b200pt_camera_project(self._h, pts.ctypes.data, pts.shape[0], xy.ctypes.data, ok.ctypes.data))
        return xy, ok.astype(bool)

    def trace_count(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        counts = (C.c_uint64 * 3)()
        check(lib().b200pt_trace_count(self._h, rays.ctypes.data, rays.shape[0], counts))
        return {"boxes": counts[0], "prims": counts[1], "hits": counts[2], "rays": rays.shape[0]}

    def trace_batch_device(self, d_rays, n, d_hits, d_medium_xi=None, repeat=1, timed=True):
        ms = C.c_float(0)
        check(lib().b200pt_trace_batch_device(self._h, d_rays, n, d_medium_xi, d_hits, repeat,
                                              C.byref(ms) if timed else None))
        return ms.value


class Wavefront:
    """The wavefront renderer behind the reference's Renderer lifecycle (renderer.h:114-150)."""

    def __init__(self, scene, config):
        self.scene = scene
        self.config = config
        self._h = None
        self.progress = _capi.Progress()
        # samples per pixel the film's sums stand for: config.samples, or -- on rank 0 after a sharded render has
        # reduced every rank's share into this film (distributed.render_sharded) -- the samples of ALL ranks
        self.spp_total = config.samples

    # -- lifecycle ---------------------------------------------------------------------------------------------
    def preprocess(self):
        pass

    def start_render(self, program_start_time=None):
        h = C.c_void_p()
        check(lib().b200pt_render_start(self.scene._h, C.byref(self.config.c), C.byref(h)))
        self._h = h
        return self

    def sync_progress(self, max_batches=1):
        """Advance by up to `max_batches` wavefront batches (0 = all) and refresh `progress`."""
        check(lib().b200pt_render_poll(self._h, max_batches, C.byref(self.progress)))
        return self.progress

    def is_done(self):
        return bool(self.progress.done)

    def wait(self):
        """Block until every batch is rendered (both modes) and refresh `progress`."""
        check(lib().b200pt_render_wait(self._h, C.byref(self.progress)))
        return self.progress

    def finalize(self, ppm_path=None):
        """Run to completion; optionally write the PPM (renderer.h:693-727).  Returns the rate report."""
        self.wait()
        p = self.progress
        secs = p.seconds if p.seconds > 0 else float("nan")
        report = {
            "camera_rays": p.samples_done, "seconds": p.seconds,
            "rays": p.rays_extension + p.rays_shadow, "rays_extension": p.rays_extension,
            "rays_shadow": p.rays_shadow, "rays_reference": p.rays_reference,
            "mpaths_per_s": p.samples_done / secs * 1e-6, "mrays_per_s": (p.rays_extension + p.rays_shadow) / secs * 1e-6,
        }
        if ppm_path:
            self.write_ppm(ppm_path)
        return report

    def close(self):
        if self._h:
            lib().b200pt_render_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- film ----------------------------------------------------------------------------------------------------
    @property
    def framebuffer(self):
        h, w = self.config.height, self.config.width
        out = np.zeros((h, w, 3), dtype=np.float32)
        spp = C.c_int64()
        check(lib().b200pt_render_read(self._h, out.ctypes.data, C.byref(spp)))
        self.spp_done = spp.value
        return out

    def stats(self):
        h, w = self.config.height, self.config.width
        s = np.zeros((h, w, 3), dtype=np.float32)
        q = np.zeros((h, w, 3), dtype=np.float32)
        check(lib().b200pt_render_read_stats(self._h, s.ctypes.data, q.ctypes.data))
        return s, q

    def profile(self):
        """Per-kernel device time / launch counts / traversal work counters (config.profile)."""
        p = _capi.Profile()
        check(lib().b200pt_render_profile(self._h, C.byref(p)))
        out = {"ms": dict(zip(_capi.KERNEL_NAMES, p.ms[:])), "launches": dict(zip(_capi.KERNEL_NAMES, p.launches[:]))}
        for k in ("ext_rays", "ext_boxes", "ext_prims", "nee_rays", "nee_boxes", "nee_prims", "nee_traced", "nee_queued"):
            out[k] = getattr(p, k)
        return out

    def device_buffer(self):
        p, n = C.c_void_p(), C.c_size_t()
        check(lib().b200pt_render_device_buffer(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def write_film(self, rgba_sums, spp):
        a = np.ascontiguousarray(rgba_sums, dtype=np.float32)
        assert a.size == self.config.height * self.config.width * 4
        check(lib().b200pt_render_write(self._h, a.ctypes.data, spp))

    def tonemap(self, n_samples=None):
        """(H, W, 3) uint8 image, TOP row first, plus (max_luminance, avg_luminance)."""
        n = n_samples if n_samples is not None else self.spp_total
        h, w = self.config.height, self.config.width
        out = np.zeros((h, w, 3), dtype=np.uint8)
        mx, avg = C.c_float(), C.c_float()
        check(lib().b200pt_render_tonemap(self._h, n, out.ctypes.data, C.byref(mx), C.byref(avg)))
        return out, mx.value, avg.value

    def write_ppm(self, path, n_samples=None):
        n = n_samples if n_samples is not None else self.spp_total
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)  # the reference needs pre_render.py for this
        check(lib().b200pt_render_write_ppm(self._h, n, path.encode()))


    def paths(self):
        """Traced paths (should_trace_paths, renderer.h:652-678): a list of dicts {pixel, sample, points (k x 3)},
        ordered by (sample, pixel)."""
        n, stride = C.c_int32(), C.c_int32()
        check(lib().b200pt_render_paths(self._h, 0, C.byref(n), C.byref(stride), None, None, None, None))
        if n.value == 0:
            return []
        pixel = np.zeros(n.value, dtype=np.int32)
        sample = np.zeros(n.value, dtype=np.int64)
        length = np.zeros(n.value, dtype=np.int32)
        pts = np.zeros((n.value, max(1, stride.value), 3), dtype=np.float32)
        check(lib().b200pt_render_paths(self._h, n.value, C.byref(n), C.byref(stride), pixel.ctypes.data,
                                        sample.ctypes.data, length.ctypes.data, pts.ctypes.data))
        return [{"pixel": int(pixel[i]), "sample": int(sample[i]), "points": pts[i, :length[i]].copy()}
                for i in range(len(pixel))]

    def write_paths(self, path_3d=None, path_2d=None):
        """handle_traced_paths (renderer.h:57-96): the two text files; returns the number of paths written."""
        path_3d = path_3d or self.config.traced_paths_output_path
        path_2d = path_2d or self.config.traced_paths_2d_output_path
        for p in (path_3d, path_2d):
            if os.path.dirname(p):
                os.makedirs(os.path.dirname(p), exist_ok=True)
        n = C.c_int32()
        check(lib().b200pt_render_write_paths(self._h, path_3d.encode(), path_2d.encode(), C.byref(n)))
        return n.value


def renderer_from_config(scene, config):
    """main.cpp:57-84: every render_type of the reference maps onto the one wavefront scheduler."""
    return Wavefront(scene, config)


def render_file(config_path="config.json", device=0):
    """What the reference's main() does (main.cpp:108-168), returning (renderer, report)."""
    cfg = Config.from_file(config_path)
    cfg = cfg.replace(device=device)
    scene = Scene.from_file(cfg.scene_path, cfg.width, cfg.height, device)
    r = renderer_from_config(scene, cfg).start_render()
    report = r.finalize(cfg.ppm_output_path)
    if cfg.trace_probability > 0:
        report["traced_paths"] = r.write_paths()
    return r, report
