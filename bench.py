#!/usr/bin/env python3
"""bench.py -- throughput of the hot path (camera ray -> closest hit -> shade -> NEE -> accumulate) in Mpaths/s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--scene NAME] [--spp-per-step S]

Workload (default): the north star's TARGET config, BASELINE.json configs[4] -- scenes/book2_final.json (8192 instances,
media, BVH) at 3840x2160 with the shipped integrator settings ("iterative nee path tracing", max_bounces 10,
light_samples 4, Russian roulette on).  It fits one GPU.  `--scene` selects any of the other four configs.
One STEP renders a fixed TOTAL of S samples per pixel of the whole film (default 64 at 4K: 531 M camera paths), so 16
steps are the config's 1024-spp frame; rates do not depend on the number of samples.
Multi-GPU (torchrun, one rank per GPU): the S spp of a step are SPLIT over the ranks (spp sharding on global Philox
sample indices, the reference's own split: renderer.h:352-360), no collective on the data path; ONE NCCL sum-reduce of
the film to rank 0 closes the job inside the timed region.  Total work is fixed as N grows -> "scaling": "strong".
`value` = camera paths of all ranks / max-over-ranks device time.

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
CONFIG_INDEX = {"three_orbs": 0, "cornell_box": 1, "cornell_box_small_lights": 2, "cornell_box_with_volume": 3,
                "book2_final": 4}
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # 74.5 TFLOP/s fp32 FMA peak at max clock


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scene", default="book2_final", choices=sorted(SCENES))
    ap.add_argument("--spp-per-step", type=int, default=0,
                    help="TOTAL samples per pixel of one step, split over the ranks; 0 = 64 (256 for the 1024^2 films)")
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target length of the CPU baseline sample")
    ap.add_argument("--cpu-one-thread", action="store_true", help="also time the reference on ONE thread (SURVEY 8d)")
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
    the program prints itself (renderer.h:703-706).  Its clock has the 0.5 s granularity of its polling loop.

    The reference never joins its worker threads (main.cpp:153-167) and, with 16 of them, was seen not to EXIT after
    finalize() had printed everything (reproduced here: rates printed after 2 s, process still there after 120 s).  So the
    harness does not wait for the process: it reads the output as it appears and stops the process once the last line
    finalize() prints ("max lum", renderer.h:722) is there.  Raises subprocess.TimeoutExpired if that takes > `timeout` s."""
    if not os.path.exists(REF_DRIVER):
        raise FileNotFoundError(REF_DRIVER + " missing: run __graft_entry__.build() where /root/reference is mounted")
    with tempfile.TemporaryDirectory() as d:
        c = dict(cfgd, samples=int(spp), threads=int(threads), render_type="tiled", block_width=block,
                 block_height=block, should_trace_paths=False, ppm_output_path=os.path.join(d, "render.ppm"),
                 png_output_path=os.path.join(d, "render.png"), traced_paths_output_path=os.path.join(d, "p.txt"),
                 traced_paths_2d_output_path=os.path.join(d, "p2.txt"))
        with open(os.path.join(d, "config.json"), "w") as f:
            json.dump(c, f)
        log = os.path.join(d, "stdout.txt")
        t0 = time.perf_counter()
        with open(log, "wb") as sink:
            proc = subprocess.Popen([REF_DRIVER, "main"], cwd=d, stdout=sink, stderr=subprocess.DEVNULL)
        out, wall = "", None
        try:
            while True:
                with open(log, "r", errors="replace") as f:
                    out = f.read()
                if "max lum" in out or proc.poll() is not None:
                    wall = time.perf_counter() - t0
                    break
                if time.perf_counter() - t0 > timeout:
                    raise subprocess.TimeoutExpired([REF_DRIVER, "main"], timeout)
                time.sleep(0.05)
            try:
                proc.wait(timeout=2.0)  # a well-behaved exit; else stop it -- everything we need has been printed
            except subprocess.TimeoutExpired:
                pass
        finally:
            if proc.poll() is None:
                proc.kill()
                proc.wait()
        with open(log, "r", errors="replace") as f:
            out = f.read()
    m1 = re.search(r"computed (\d+) camera rays in ([0-9.eE+-]+)s", out)
    m2 = re.search(r"computed (\d+) rays, at ([0-9.eE+-]+) rays per second", out)
    if not m1 or not m2:
        raise RuntimeError("reference output not understood: " + out[-400:])
    paths, secs, rays = int(m1.group(1)), float(m1.group(2)), int(m2.group(1))
    return {"paths": paths, "rays": rays, "seconds": secs, "wall_seconds": wall,
            "mpaths_per_s": paths / secs * 1e-6, "mrays_per_s": rays / secs * 1e-6}


def usable_threads():
    """Host threads this process may run on (respects cpusets / affinity; os.cpu_count() reports the whole machine)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def with_film(cfgd, w, h):
    return dict(cfgd, film=dict(cfgd["film"], width=int(w), height=int(h)))


def cpu_sample(cfgd, threads, target_seconds):
    """One BOUNDED run of the reference on the bench scene, sized for about `target_seconds` of rendering.  The rates the
    reference prints (paths / s, rays / s) do not depend on the film size or the sample count, so the sample is the same
    scene, camera and settings on a film scaled down (same aspect) until 1 spp fits the budget -- the full film at several
    spp when that fits too.  Calibrated on a 1/8-size film and re-sized once more if the first sized run turns out
    shorter than a third of the target (the reference's clock ticks in 0.5 s steps: a short run gives a coarse rate);
    every run has a hard timeout, and if a sized run does not finish the last finished one is the sample.
    Returns (description "WxH at N spp", run)."""
    w, h = cfgd["film"]["width"], cfgd["film"]["height"]
    pw, ph = max(32, w // 8), max(18, h // 8)
    try:
        run = run_reference_once(with_film(cfgd, pw, ph), 1, threads, timeout=180)
    except subprocess.TimeoutExpired:  # a pathologically slow host: a thumbnail, or give up (the caller reports it)
        pw, ph = 64, max(18, 64 * h // max(w, 1))
        run = run_reference_once(with_film(cfgd, pw, ph), 1, threads, timeout=300)
        return f"{pw}x{ph} at 1 spp (the 1/8-size probe exceeded its time limit)", run
    desc = f"{pw}x{ph} at 1 spp"

    def shape(n_paths):  # the film / spp that holds about n_paths camera paths
        if n_paths >= w * h:
            return w, h, max(1, min(256, int(n_paths // (w * h))))
        f = max((n_paths / float(w * h)) ** 0.5, pw / float(w))
        return max(pw, int(w * f)), max(ph, int(h * f)), 1

    for _ in range(6):
        secs = max(run["seconds"], 0.25)  # its own render clock: it ticks in 0.5 s steps (the polling loop of main())
        if secs >= 0.6 * target_seconds:  # long enough that a 0.5 s tick is a few per cent
            break
        # below ~1 s the clock says nothing (one tick): grow geometrically; above, scale to the target
        grow = 8.0 if secs < 1.1 else min(8.0, target_seconds / secs)
        sw, sh, spp = shape(run["paths"] * grow)
        new_desc = f"{sw}x{sh} at {spp} spp"
        if new_desc == desc:
            break
        try:
            run = run_reference_once(with_film(cfgd, sw, sh), spp, threads, timeout=max(60.0, 5.0 * target_seconds))
            desc = new_desc
        except subprocess.TimeoutExpired:
            desc += f" (the sized run of {new_desc} exceeded its time limit)"
            break
    return desc, run


def workload_name(args, w, h):
    return (f"{args.scene} {w}x{h}, iterative nee path tracing, max_bounces 10, light_samples 4, RR on "
            f"(BASELINE.json configs[{CONFIG_INDEX[args.scene]}])")


def reference_arm(args):
    """--impl reference: timed runs of the UNMODIFIED reference main() on all host threads, each a bounded sample of the
    bench film (whole film, few spp), the whole arm bounded to a few minutes whatever --steps asks for."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cfgd = config_dict(args)
    threads = usable_threads()
    w, h = cfgd["film"]["width"], cfgd["film"]["height"]
    t_arm = time.perf_counter()
    try:
        # a step of about 10 s (its 0.5 s clock tick is then ~5 %), shorter only if K steps would not fit the budget
        step_target = max(3.0, min(10.0, 140.0 / max(1, args.steps)))
        desc, first = cpu_sample(cfgd, threads, step_target)  # calibration + the first sample; doubles as the warm-up run
    except Exception as e:  # noqa: BLE001 -- the arm must end with a line and exit 0 whatever the host does
        print(json.dumps({"impl": "reference", "unavailable": f"reference CPU run failed or timed out: {e}"[:300]}), flush=True)
        return 0
    per_step = max(first["wall_seconds"], 0.5)
    budget = 170.0  # seconds of timed runs, whatever --steps asks for
    steps = max(1, min(args.steps, int(budget / per_step)))
    warm = 1
    m = re.match(r"(\d+)x(\d+) at (\d+) spp", desc)
    sw, sh, spp = int(m.group(1)), int(m.group(2)), int(m.group(3))
    runs = []
    for _ in range(steps):
        if runs and time.perf_counter() - t_arm > 270.0:
            break  # never let the arm run away
        try:
            runs.append(run_reference_once(with_film(cfgd, sw, sh), spp, threads, timeout=max(60.0, 6.0 * per_step)))
        except subprocess.TimeoutExpired:
            break
    if not runs:
        runs = [first]
    steps = len(runs)
    paths = sum(r["paths"] for r in runs)
    secs = sum(r["seconds"] for r in runs)
    rays = sum(r["rays"] for r in runs)
    value = paths / secs * 1e-6
    sample = (f"{args.scene}, film {desc} per step x {steps} steps (bounded sample of the {w}x{h} workload: the same scene, "
              f"camera and settings, rates do not depend on film size; --steps {args.steps} asked), tiled 32x32 blocks, "
              f"{threads} threads, unmodified reference main(); rates as it prints them (renderer.h:705-706)")
    line = {
        "impl": "reference", "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": secs / steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args, w, h), "sample_film": f"{sw}x{sh}", "spp_per_step": spp,
                   "steps_requested": args.steps},
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


def traversal_roofline(prof_time, prof_count, peaks, scene_name, small_scene):
    """bytes/ray = 32*boxes + 80*prims + 48 for the dominant traversal kernel, divided by its CUDA-event
    time -- plus what makes the figure readable: whether it is a bound at all (`bound_valid`), the rays that were
    really traced next to the reference-accounted ones, and the issue-slot figure ncu says binds these kernels."""
    ms = prof_time["ms"]
    top = max(("nee", "extend"), key=lambda k: ms[k])
    pre = "nee" if top == "nee" else "ext"
    launches = max(1, prof_time["launches"][top])
    peak = peaks["hbm_gbs"]
    share = ms[top] / max(1e-12, sum(ms.values()))
    ncu = peaks.get("ncu", {}).get(scene_name, {})

    def one(kernel, pre):  # the same figures for either traversal kernel
        r, b, p = prof_count[pre + "_rays"], prof_count[pre + "_boxes"], prof_count[pre + "_prims"]
        traced = prof_count["nee_traced"] if pre == "nee" else r  # every extension ray is traced
        t = ms[kernel] * 1e-3
        by = 32.0 * b + 80.0 * p + 48.0 * traced  # 48 B of ray in / hit out only for queries that are traced
        fl = 20.0 * b + p * (36.0 + 30.0) + 9.0 * traced
        out = {"rays_reference_accounting": r, "rays_traced": traced,
               "rays_queued": prof_count["nee_queued"] if pre == "nee" else r,
               "boxes_per_traced_ray": round(b / max(1, traced), 3), "prims_per_traced_ray": round(p / max(1, traced), 3),
               "bytes_per_traced_ray": round(by / max(1, traced), 1),
               "grays_traced_per_s": round(traced / t / 1e9, 3) if t > 0 else 0.0,
               "grays_reference_per_s": round(r / t / 1e9, 3) if t > 0 else 0.0,
               "algorithmic_gbs": round(by / t / 1e9, 1) if t > 0 else 0.0,
               "frac_of_hbm": round(by / t / 1e9 / peak, 4) if t > 0 else 0.0,
               "fp32_tflops": round(fl / t / 1e12, 3) if t > 0 else 0.0, "ms": round(ms[kernel], 4)}
        n = ncu.get("k_" + kernel)
        if n:  # from the committed ncu pass of this build (tools/capture_profiles.sh -> profiles/ncu_metrics.json)
            out["ncu"] = n
            out["useful_issue_frac"] = round(n["issue_active_pct"] / 100.0 * n["lanes_per_warp_inst"] / 32.0, 4)
        return out

    kernels = {"k_extend": one("extend", "ext"), "k_nee": one("nee", "nee")}
    k = kernels["k_" + top]
    frac = k["frac_of_hbm"]
    valid = (not small_scene) and frac <= 1.0
    why = ("scene records are staged in shared memory: the 8d byte count never comes from HBM" if small_scene else
           ("frac > 1: the node / instance records are served from L1/L2 (the scene is 1.6 MB), not HBM -- the SURVEY 8d "
            "byte model is not a bound here; see issue_bound and traffic (measured DRAM bytes)" if frac > 1.0 else
            "scene is L2-resident (1.6 MB): an upper bound on fetch traffic, not on DRAM traffic"))
    return {
        "bound": "hbm", "kernel": "k_" + top, "achieved": k["algorithmic_gbs"], "peak": peak, "unit": "GB/s",
        "frac": frac, "bound_valid": valid, "bound_note": why, "peak_source": peaks["source"],
        "traffic": (k.get("ncu") or {}).get("dram_bytes_per_launch"),
        "traffic_source": (peaks.get("ncu_source") if k.get("ncu") else None),
        "algorithmic_bytes_per_launch": k["algorithmic_gbs"] * 1e9 * ms[top] * 1e-3 / launches,
        "avg_launch_ms": ms[top] / launches, "launches": launches, "share_of_step": share,
        "issue_bound": {"what": "smsp__issue_active x (active lanes / 32) of the same kernel: the fraction of the SM's "
                                "issue slots doing useful lane work -- the resource ncu shows these kernels are bound by",
                        "frac": k.get("useful_issue_frac")},
        "fp32_frac_of_nominal": k["fp32_tflops"] / FP32_PEAK_TFLOPS,
        "traversal_kernels": kernels, "kernel_ms": {kk: round(v, 4) for kk, v in ms.items()},
    }


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peaks = {"hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}
    if os.path.exists(path):
        with open(path) as f:
            j = json.load(f)
        peaks = {"hbm_gbs": float(j["hbm_gbs"]), "source": "measured (MEASURED_PEAKS.json, burst copy)"}
    nc = os.path.join(ROOT, "profiles", "ncu_metrics.json")  # ncu --set full figures of the top kernels, per scene
    if os.path.exists(nc):
        with open(nc) as f:
            j = json.load(f)
        peaks["ncu"] = j.get("scenes", {})
        peaks["ncu_source"] = (f"profiles/ncu_metrics.json (ncu metrics pass over every launch of one warm batch, "
                               f"tools/batch_profile.sh; commit {j.get('commit', '?')})")
    return peaks


def pick_batch(count, auto):
    """Samples per wavefront batch: the largest divisor of a rank's per-step share that fits the automatic batch."""
    best = 1
    for d in range(1, count + 1):
        if count % d == 0 and d <= auto:
            best = d
    return best


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
    spp_step = args.spp_per_step or (256 if 1_000_000 <= npix < 2_000_000 else 64)  # TOTAL over all ranks
    if spp_step < world:
        raise SystemExit("bench.py: --spp-per-step must be at least the number of GPUs")
    K, W = args.steps, max(args.warmup, 3)
    with open(cfgd["scene"]) as f:
        scene_json = f.read()
    cfg_json = json.dumps(cfgd)

    # ---- device-timed region: inputs (scene, wavefront state) resident in HBM ----------------------------------------
    # STRONG scaling: the job is (W + K) x spp_step samples of every pixel whatever the number of GPUs; rank r renders the
    # contiguous share shard_samples() gives it (global Philox sample indices), count_r spp of it per step.
    scene = pt.Scene.from_json(scene_json, w, h, device=local)
    first_r, count_r = ptd.shard_samples(spp_step, world, rank)
    base = pt.Config.from_json(cfg_json).replace(device=local)
    target_slots = int(os.environ.get("B200PT_BATCH_SLOTS", "0")) or (128 << 20)  # the library's default (capi.cu)
    batch = pick_batch(count_r, max(1, target_slots // npix))
    batches_per_step = count_r // batch
    cfg = base.replace(samples=(W + K) * count_r, sample_offset=(W + K) * first_r, batch_samples=batch)
    r = pt.renderer_from_config(scene, cfg).start_render()
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
    t_dev0, paths0 = p0.seconds, p0.samples_done
    rays0 = (p0.rays_extension, p0.rays_shadow)
    red0, red1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    with ClockSampler(local) as clocks:
        wall0 = time.perf_counter()
        for _ in range(K):
            r.sync_progress(batches_per_step)  # device time by CUDA events on the render stream (capi.cu poll)
        local_ck = film.double().sum(dim=(0, 1)) if world > 1 else None  # per-channel checksum of this rank's film
        red0.record()
        if world > 1:
            ptd.reduce_film(film, 0)  # the one exchange step: NCCL sum-reduce of the film to rank 0
        red1.record()
        barrier()
        wall = time.perf_counter() - wall0
    p1 = r.progress
    dev_s = (p1.seconds - t_dev0) + red0.elapsed_time(red1) * 1e-3
    paths = p1.samples_done - paths0
    rays_ext, rays_shadow = p1.rays_extension - rays0[0], p1.rays_shadow - rays0[1]
    launches = sum(r.profile()["launches"].values()) - launches0 + (1 if world > 1 else 0)
    film_check = None
    if world > 1:
        # a checksum of checksums: the reduced film on rank 0 must sum to the sum of the per-rank sums, channel by
        # channel, and its unused .w lane must still be zero (a wrong root, a stale stream or garbage lanes show here)
        dist.all_reduce(local_ck, op=dist.ReduceOp.SUM)
        if rank == 0:
            got = film.double().sum(dim=(0, 1))
            rel = ((got[:3] - local_ck[:3]).abs() / local_ck[:3].abs().clamp_min(1e-30)).max().item()
            film_check = {"what": "per-channel sum of the NCCL-reduced film vs the sum of the per-rank film sums (float64)",
                          "checksum_rel_err": rel, "w_lane_abs_sum": float(got[3].abs().item()),
                          "ok": bool(rel < 1e-5 and got[3].abs().item() == 0.0)}
    stat = torch.tensor([dev_s, wall], dtype=torch.float64, device="cuda")
    cnt = torch.tensor([paths, rays_ext, rays_shadow, launches], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(stat, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    dev_s, wall = [float(x) for x in stat.tolist()]
    paths, rays_ext, rays_shadow, launches = [int(x) for x in cnt.tolist()]
    r.close()

    # ---- end to end through the public API: host JSON in, film in host memory out, every step ----------------------
    def e2e_step(step):
        sc = pt.Scene.from_json(scene_json, w, h, device=local)                     # parse + BVH + H2D upload
        c = pt.Config.from_json(cfg_json).replace(device=local, samples=spp_step, sample_offset=step * spp_step)
        rr, rep = ptd.render_sharded(sc, c, rank, world)                            # render (+ NCCL reduce)
        fb = rr.framebuffer if rank == 0 else None                                  # D2H read of the film
        rr.close()
        sc.close()
        return rep, fb

    e2e_step(0)
    barrier()
    e0 = time.perf_counter()
    e_paths = 0
    for s_ in range(K):
        rep, fb = e2e_step(s_ + 1)
        e_paths += rep["camera_rays"] if rank == 0 else 0
    barrier()
    e_wall = torch.tensor([time.perf_counter() - e0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e_wall, op=dist.ReduceOp.MAX)
    e_wall = float(e_wall.item())
    h2d = scene_upload_bytes(scene)
    d2h = npix * 12
    return dict(rank=rank, world=world, local=local, cfgd=cfgd, w=w, h=h, spp_step=spp_step, spp_rank=count_r, K=K, W=W,
                dev_s=dev_s, wall=wall, paths=paths, rays_ext=rays_ext, rays_shadow=rays_shadow, launches=launches,
                clocks=clocks.summary(), e_wall=e_wall, e_paths=e_paths, h2d=h2d, d2h=d2h, scene=scene, base=base,
                batch=batch, film_check=film_check, small_scene=scene.desc().n_instances <= 16)


def profile_passes(res):
    """Two extra single-step renders on this rank: per-kernel CUDA-event times, then traversal work counters of the
    same samples -- kept out of the timed region."""
    import pathtrace_b200 as pt
    out = []
    for mode in (1, 2):
        cfg = res["base"].replace(samples=res["spp_rank"], sample_offset=0, batch_samples=res["batch"], profile=mode)
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
    roof = traversal_roofline(prof_time, prof_count, load_peaks(), args.scene, res["small_scene"])
    cpu = None
    if world == 1 and not args.no_cpu_baseline and os.path.exists(REF_DRIVER):
        threads = usable_threads()
        try:
            desc, run = cpu_sample(res["cfgd"], threads, args.cpu_seconds)
            cpu = {"value": run["mpaths_per_s"], "unit": "Mpaths/s", "cores": threads, "kind": "reference",
                   "mrays_per_s": run["mrays_per_s"], "seconds": run["seconds"],
                   "sample": f"{args.scene}, film {desc} (bounded sample of the {res['w']}x{res['h']} workload: same scene, "
                             f"camera and settings; rates do not depend on film size), unmodified reference main() "
                             f"(tiled, 32x32 blocks, {threads} threads)"}
            if args.cpu_one_thread:  # the deterministic single-thread rate (no RNG race), on a reduced film
                d1, r1 = cpu_sample(res["cfgd"], 1, min(8.0, args.cpu_seconds))
                cpu["one_thread"] = {"value": r1["mpaths_per_s"], "mrays_per_s": r1["mrays_per_s"],
                                     "sample": f"film {d1}, 1 thread"}
        except Exception as e:  # the baseline is reported, never required
            cpu = {"value": None, "unit": "Mpaths/s", "cores": threads, "kind": "reference", "sample": f"failed: {e}"}
    value = res["paths"] / res["dev_s"] * 1e-6
    rays = res["rays_ext"] + res["rays_shadow"]
    # shadow queries really traced = reference-accounted ones x the traced share measured by the counters-on pass
    nee = roof["traversal_kernels"]["k_nee"]
    traced_share = nee["rays_traced"] / max(1, nee["rays_reference_accounting"])
    rays_traced = res["rays_ext"] + res["rays_shadow"] * traced_share
    line = {
        "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": res["K"],
        "warmup": res["W"], "ms_per_step": res["dev_s"] / res["K"] * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args, res["w"], res["h"]),
                   "spp_per_step_total": res["spp_step"], "spp_per_step_this_rank": res["spp_rank"],
                   "wavefront_batch_spp": res["batch"],
                   "parallelism": f"spp-sharding x{world} (strong: the step's samples are split over the ranks), one NCCL "
                                  f"reduce of the film closes the job",
                   "l2": "per-step wavefront state (>= 1.3 GB) exceeds the 126 MB L2; no flush needed"},
        "mrays_per_s": rays / res["dev_s"] * 1e-6, "rays_per_path": rays / max(1, res["paths"]),
        "rays": {"rule": "Mrays/s counts what the reference counts (integrator.h:192,247): one per path segment + "
                         "light_samples per vertex, including light samples that need no ray (coefficient 0: light below the "
                         "horizon, pdf 0). mrays_traced_per_s counts only queries that are really intersected.",
                 "extension": res["rays_ext"], "shadow_reference_accounting": res["rays_shadow"],
                 "shadow_traced_share": round(traced_share, 4),
                 "mrays_traced_per_s": rays_traced / res["dev_s"] * 1e-6},
        "wall_ms_per_step": res["wall"] / res["K"] * 1e3,
        "e2e": {"value": res["e_paths"] / res["e_wall"] * 1e-6, "unit": "Mpaths/s",
                "h2d_bytes_per_step": res["h2d"], "d2h_bytes_per_step": res["d2h"],
                "what": "per step: scene JSON (host) -> b200pt_scene_create_from_json -> render (sharded) -> NCCL reduce -> "
                        "film read to host"},
        "gpu_launches": res["launches"], "clocks": res["clocks"], "roofline": roof, "cpu_baseline": cpu,
        "film_check": res["film_check"],
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
