#!/usr/bin/env python3
"""bench.py -- throughput of the hot path (camera ray -> closest hit -> shade -> NEE -> accumulate) in Mpaths/s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--scene NAME] [--spp-per-step S]

Workload (BASELINE.json configs[1]): scenes/cornell_box.json at 1024x1024 with the shipped integrator settings
("iterative nee path tracing", max_bounces 10, light_samples 4, Russian roulette on).  One STEP renders S samples
per pixel of the whole film (S x 1024 x 1024 camera paths), as wavefront batches on the device; rates do not depend on
the number of samples, so K steps of S spp stand for the config's 1024 spp (K*S = 1024 reproduces it in full).
Multi-GPU (torchrun, one rank per GPU): every rank renders its own S-spp share of each step (spp sharding, global
sample indices keyed into Philox), no collective on the data path; ONE NCCL sum-reduce of the film to rank 0 closes
the timed region.  `value` = camera paths of all ranks / max-over-ranks device time -> "scaling": "weak".

Output: ONE JSON line (see the keys at the bottom).  `--impl reference` times the UNMODIFIED reference CPU renderer
(oracle/_ref/ref_driver main = its own main()) on the host cores for the same scene/film/settings.
"""
import argparse
import json
import os
import re
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SCENES = {  # name -> (width, height, spp of the BASELINE config)
    "three_orbs": (400, 225, 64),
    "cornell_box": (1024, 1024, 1024),
    "cornell_box_small_lights": (1024, 1024, 1024),
    "cornell_box_with_volume": (1024, 1024, 2048),
    "book2_final": (3840, 2160, 1024),
}
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # 74.5 TFLOP/s fp32 FMA peak at max clock


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scene", default="cornell_box", choices=sorted(SCENES))
    ap.add_argument("--spp-per-step", type=int, default=0, help="0 = 16 (4 for the 4K scene)")
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target length of the CPU baseline sample")
    return ap.parse_args()


def config_dict(args):
    import pathtrace_b200 as pt
    w, h, _ = SCENES[args.scene]
    w, h = args.width or w, args.height or h
    return pt.default_config_dict(width=w, height=h, scene=os.path.join(ROOT, "scenes", args.scene + ".json"))


# ---- clocks during the timed region (B200_PROFILING.md) ---------------------------------------------------------

class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.rows, self.stop_flag, self.thread = device, [], threading.Event(), None

    def _run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.FIELDS}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 7:
                    self.rows.append(parts)
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def __enter__(self):
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop_flag.set()
        self.thread.join(timeout=15)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[3 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


# ---- the reference's CPU renderer on the host cores ------------------------------------------------------------

REF_DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")


def run_reference_once(cfgd, spp, threads, block=32, timeout=3000):
    """One run of the UNMODIFIED reference main() (main.cpp:108-168 via oracle/_ref/ref_driver main): tiled
    renderer, `threads` workers, block x block tiles so that tiles >= threads.  Returns the two rates
    the program prints itself (renderer.h:703-706).  Its clock has the 0.5 s granularity of its polling loop."""
    if not os.path.exists(REF_DRIVER):
        raise FileNotFoundError(REF_DRIVER + " missing: run __graft_entry__.build() where /root/reference is mounted")
    with tempfile.TemporaryDirectory() as d:
        c = dict(cfgd, samples=int(spp), threads=int(threads), render_type="tiled", block_width=block,
                 block_height=block, should_trace_paths=False, ppm_output_path=os.path.join(d, "render.ppm"),
                 png_output_path=os.path.join(d, "render.png"), traced_paths_output_path=os.path.join(d, "p.txt"),
                 traced_paths_2d_output_path=os.path.join(d, "p2.txt"))
        with open(os.path.join(d, "config.json"), "w") as f:
            json.dump(c, f)
        t0 = time.perf_counter()
        out = subprocess.run([REF_DRIVER, "main"], cwd=d, capture_output=True, text=True, timeout=timeout).stdout
        wall = time.perf_counter() - t0
    m1 = re.search(r"computed (\d+) camera rays in ([0-9.eE+-]+)s", out)
    m2 = re.search(r"computed (\d+) rays, at ([0-9.eE+-]+) rays per second", out)
    if not m1 or not m2:
        raise RuntimeError("reference output not understood: " + out[-400:])
    paths, secs, rays = int(m1.group(1)), float(m1.group(2)), int(m2.group(1))
    return {"paths": paths, "rays": rays, "seconds": secs, "wall_seconds": wall,
            "mpaths_per_s": paths / secs * 1e-6, "mrays_per_s": rays / secs * 1e-6}


def cpu_sample_spp(cfgd, threads, target_seconds):
    """Samples per pixel for a run of about `target_seconds`: calibrated on a 1-spp run of the same film."""
    probe = run_reference_once(cfgd, 1, threads)
    rate = probe["paths"] / max(probe["wall_seconds"], 0.25)
    npix = cfgd["film"]["width"] * cfgd["film"]["height"]
    return max(1, min(64, int(round(rate * target_seconds / npix)))), probe


def reference_arm(args):
    """--impl reference: K timed runs (after W warm-up runs) of the reference on all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cfgd = config_dict(args)
    threads = os.cpu_count() or 1
    w, h = cfgd["film"]["width"], cfgd["film"]["height"]
    # a step = the whole film at `spp` samples, sized for ~10 s; K and W are capped so the arm ends within minutes
    spp, probe = cpu_sample_spp(cfgd, threads, 10.0)
    steps, warm = max(1, min(args.steps, 6)), max(0, min(args.warmup, 1))
    for _ in range(warm):
        run_reference_once(cfgd, spp, threads)
    runs = [run_reference_once(cfgd, spp, threads) for _ in range(steps)]
    paths = sum(r["paths"] for r in runs)
    secs = sum(r["seconds"] for r in runs)
    rays = sum(r["rays"] for r in runs)
    value = paths / secs * 1e-6
    sample = f"{args.scene} {w}x{h} at {spp} spp per step, tiled 32x32 blocks, unmodified reference main()"
    line = {
        "impl": "reference", "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": secs / steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.scene} {w}x{h}, iterative nee path tracing, max_bounces 10, light_samples 4, RR on",
                   "spp_per_step": spp, "steps_requested": args.steps},
        "mrays_per_s": rays / secs * 1e-6, "rays_per_path": rays / paths,
        "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": threads, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ---- our arm --------------------------------------------------------------------------------------------------------

def scene_upload_bytes(scene):
    """Bytes b200pt_scene_create_from_json copies host -> device (capi.cu upload_scene)."""
    d = scene.desc()
    n_nodes = scene.bvh_info()["n_nodes"]
    return (n_nodes * 128 + d.n_instances * (80 + 48) + d.n_materials * 32 + d.n_textures * 32 + d.n_lights * 4
            + 768 * 4 * 2)


def traversal_roofline(prof_time, prof_count, peaks):
    """bytes/ray = 32*boxes + 80*prims + 48, flops/ray = 20*boxes/... for the dominant kernel class."""
    ms = prof_time["ms"]
    top = max(("nee", "extend"), key=lambda k: ms[k])
    pre = "nee" if top == "nee" else "ext"
    rays, boxes, prims = prof_count[pre + "_rays"], prof_count[pre + "_boxes"], prof_count[pre + "_prims"]
    launches = max(1, prof_time["launches"][top])
    bytes_total = 32.0 * boxes + 80.0 * prims + 48.0 * rays
    flops_total = 20.0 * boxes + prims * (36.0 + 30.0) + 9.0 * rays
    secs = ms[top] * 1e-3
    achieved = bytes_total / secs / 1e9 if secs > 0 else 0.0
    peak = peaks["hbm_gbs"]
    share = ms[top] / max(1e-12, sum(ms.values()))

    def one(kernel, pre):  # the same figures for either traversal kernel
        r, b, p = prof_count[pre + "_rays"], prof_count[pre + "_boxes"], prof_count[pre + "_prims"]
        t = ms[kernel] * 1e-3
        by = 32.0 * b + 80.0 * p + 48.0 * r
        return {"rays": r, "boxes_per_ray": round(b / max(1, r), 3), "prims_per_ray": round(p / max(1, r), 3),
                "bytes_per_ray": round(by / max(1, r), 1), "grays_per_s": round(r / t / 1e9, 3) if t > 0 else 0.0,
                "achieved": round(by / t / 1e9, 1) if t > 0 else 0.0, "frac": round(by / t / 1e9 / peak, 4) if t > 0 else 0.0}

    return {
        "traversal_kernels": {"k_extend": one("extend", "ext"), "k_nee": one("nee", "nee")},
        "bound": "hbm", "kernel": "k_nee" if top == "nee" else "k_extend", "achieved": achieved, "peak": peak,
        "unit": "GB/s", "frac": achieved / peak, "peak_source": peaks["source"], "traffic": peaks.get("traffic"),
        "algorithmic_bytes_per_launch": bytes_total / launches, "avg_launch_ms": ms[top] / launches,
        "launches": launches, "rays": rays, "boxes_per_ray": boxes / max(1, rays), "prims_per_ray": prims / max(1, rays),
        "bytes_per_ray": bytes_total / max(1, rays), "grays_per_s": rays / secs / 1e9 if secs > 0 else 0.0,
        "fp32_tflops": flops_total / secs / 1e12 if secs > 0 else 0.0,
        "fp32_frac_of_nominal": (flops_total / secs / 1e12) / FP32_PEAK_TFLOPS if secs > 0 else 0.0,
        "share_of_step": share, "kernel_ms": {k: round(v, 4) for k, v in ms.items()},
        "note": "scene data (BVH + instance table) is L2-resident; the byte figure is the algorithmic node/primitive "
                "fetch traffic of SURVEY.md 8d, not DRAM traffic",
    }


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peaks = {"hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}
    if os.path.exists(path):
        with open(path) as f:
            j = json.load(f)
        peaks = {"hbm_gbs": float(j["hbm_gbs"]), "source": "measured (MEASURED_PEAKS.json, burst copy)"}
    tr = os.path.join(ROOT, "profiles", "traffic.json")  # dram bytes per launch of the top kernel from ncu --set full
    if os.path.exists(tr):
        with open(tr) as f:
            peaks["traffic"] = json.load(f).get("dram_bytes_per_launch")
    return peaks


def ours_arm(args):
    import torch
    import torch.distributed as dist

    import pathtrace_b200 as pt
    from pathtrace_b200 import distributed as ptd

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this backend has no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # the version banner goes to stdout, which carries the ONE JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    cfgd = config_dict(args)
    w, h = cfgd["film"]["width"], cfgd["film"]["height"]
    npix = w * h
    spp_step = args.spp_per_step or (4 if npix > 4_000_000 else 16)
    K, W = args.steps, max(args.warmup, 3)
    with open(cfgd["scene"]) as f:
        scene_json = f.read()
    cfg_json = json.dumps(cfgd)

    # ---- device-timed region: inputs (scene, wavefront state) resident in HBM ----------------------------------------
    scene = pt.Scene.from_json(scene_json, w, h, device=local)
    total_spp = (W + K) * spp_step
    base = pt.Config.from_json(cfg_json).replace(device=local)
    # rank r renders global samples [r * total_spp, (r+1) * total_spp): weak scaling, disjoint Philox streams
    cfg = base.replace(samples=total_spp, sample_offset=rank * total_spp, batch_samples=0)
    r = pt.renderer_from_config(scene, cfg).start_render()
    target_slots = int(os.environ.get("B200PT_BATCH_SLOTS", "0")) or (32 << 20)
    batch = max(1, min(target_slots // npix, total_spp))  # the library's automatic wavefront batch (capi.cu)
    batches_per_step = -(-spp_step // batch)
    assert spp_step % batch == 0 or batch > spp_step, "a step must be a whole number of wavefront batches"
    if batch > spp_step:
        r.close()
        cfg = cfg.replace(batch_samples=spp_step)
        r = pt.renderer_from_config(scene, cfg).start_render()
        batches_per_step = 1
    for _ in range(W):
        r.sync_progress(batches_per_step)
    film = ptd.film_tensor(r)
    if world > 1:  # warm-up of the one collective: same size, scratch buffer (the first NCCL reduce sets up channels)
        scratch = torch.zeros_like(film)
        for _ in range(2):
            ptd.reduce_film(scratch, 0)
        del scratch
    launches0 = sum(r.profile()["launches"].values())
    p0 = r.progress
    t_dev0, paths0, rays0 = p0.seconds, p0.samples_done, p0.rays_extension + p0.rays_shadow
    red0, red1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    with ClockSampler(local) as clocks:
        wall0 = time.perf_counter()
        for _ in range(K):
            r.sync_progress(batches_per_step)  # device time by CUDA events on the render stream (capi.cu poll)
        red0.record()
        if world > 1:
            ptd.reduce_film(film, 0)  # the one exchange step: NCCL sum-reduce of the film to rank 0
        red1.record()
        barrier()
        wall = time.perf_counter() - wall0
    p1 = r.progress
    dev_s = (p1.seconds - t_dev0) + red0.elapsed_time(red1) * 1e-3
    paths = p1.samples_done - paths0
    rays = p1.rays_extension + p1.rays_shadow - rays0
    launches = sum(r.profile()["launches"].values()) - launches0 + (1 if world > 1 else 0)
    stat = torch.tensor([dev_s, wall], dtype=torch.float64, device="cuda")
    cnt = torch.tensor([paths, rays, launches], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(stat, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    dev_s, wall = [float(x) for x in stat.tolist()]
    paths, rays, launches = [int(x) for x in cnt.tolist()]
    r.close()

    # ---- end to end through the public API: host JSON in, film in host memory out, every step ----------------------
    def e2e_step(step):
        sc = pt.Scene.from_json(scene_json, w, h, device=local)                     # parse + BVH + H2D upload
        c = pt.Config.from_json(cfg_json).replace(device=local, samples=spp_step * world,
                                                  sample_offset=step * spp_step * world)
        rr, rep = ptd.render_sharded(sc, c, rank, world)                            # render (+ NCCL reduce)
        fb = rr.framebuffer if rank == 0 else None                                  # D2H read of the film
        rr.close()
        sc.close()
        return rep, fb

    e2e_step(0)
    barrier()
    e0 = time.perf_counter()
    e_paths = 0
    for s in range(K):
        rep, fb = e2e_step(s + 1)
        e_paths += rep["camera_rays"] if rank == 0 else 0
    barrier()
    e_wall = torch.tensor([time.perf_counter() - e0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e_wall, op=dist.ReduceOp.MAX)
    e_wall = float(e_wall.item())
    h2d = scene_upload_bytes(scene)
    d2h = npix * 16
    return dict(rank=rank, world=world, local=local, cfgd=cfgd, w=w, h=h, spp_step=spp_step, K=K, W=W, dev_s=dev_s,
                wall=wall, paths=paths, rays=rays, launches=launches, clocks=clocks.summary(), e_wall=e_wall,
                e_paths=e_paths, h2d=h2d, d2h=d2h, scene=scene, base=base, batch=min(batch, spp_step))


def profile_passes(res):
    """Two extra single-step renders on this rank: per-kernel CUDA-event times, then traversal work counters of the
    same samples -- kept out of the timed region."""
    import pathtrace_b200 as pt
    out = []
    for mode in (1, 2):
        cfg = res["base"].replace(samples=res["spp_step"], sample_offset=0, batch_samples=res["batch"], profile=mode)
        r = pt.renderer_from_config(res["scene"], cfg).start_render()
        if mode == 1:  # warm the caches / clocks like the timed region
            warm = pt.renderer_from_config(res["scene"], cfg.replace(profile=0)).start_render()
            warm.finalize()
            warm.close()
        r.finalize()
        out.append(r.profile())
        r.close()
    return out


def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    res = ours_arm(args)
    rank, world = res["rank"], res["world"]
    if rank != 0:
        import torch.distributed as dist
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0
    prof_time, prof_count = profile_passes(res)
    roof = traversal_roofline(prof_time, prof_count, load_peaks())
    cpu = None
    if world == 1 and not args.no_cpu_baseline and os.path.exists(REF_DRIVER):
        threads = os.cpu_count() or 1
        try:
            spp, _ = cpu_sample_spp(res["cfgd"], threads, args.cpu_seconds)
            run = run_reference_once(res["cfgd"], spp, threads)
            cpu = {"value": run["mpaths_per_s"], "unit": "Mpaths/s", "cores": threads, "kind": "reference",
                   "mrays_per_s": run["mrays_per_s"], "seconds": run["seconds"],
                   "sample": f"{args.scene} {res['w']}x{res['h']} at {spp} spp, unmodified reference main() "
                             f"(tiled, 32x32 blocks, {threads} threads)"}
        except Exception as e:  # the baseline is reported, never required
            cpu = {"value": None, "unit": "Mpaths/s", "cores": threads, "kind": "reference", "sample": f"failed: {e}"}
    value = res["paths"] / res["dev_s"] * 1e-6
    line = {
        "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": res["K"],
        "warmup": res["W"], "ms_per_step": res["dev_s"] / res["K"] * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.scene} {res['w']}x{res['h']}, iterative nee path tracing, max_bounces 10, "
                               f"light_samples 4, RR on (BASELINE configs: cornell_box = configs[1])",
                   "spp_per_step_per_gpu": res["spp_step"], "wavefront_batch_spp": res["batch"],
                   "parallelism": f"spp-sharding x{world}, one NCCL reduce of the film",
                   "l2": "per-step wavefront state (>= 1.3 GB) exceeds the 126 MB L2; no flush needed"},
        "mrays_per_s": res["rays"] / res["dev_s"] * 1e-6, "rays_per_path": res["rays"] / max(1, res["paths"]),
        "wall_ms_per_step": res["wall"] / res["K"] * 1e3,
        "e2e": {"value": res["e_paths"] / res["e_wall"] * 1e-6, "unit": "Mpaths/s",
                "h2d_bytes_per_step": res["h2d"], "d2h_bytes_per_step": res["d2h"],
                "what": "per step: scene JSON (host) -> b200pt_scene_create_from_json -> render -> film read to host"},
        "gpu_launches": res["launches"], "clocks": res["clocks"], "roofline": roof, "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
