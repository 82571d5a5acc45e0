"""The Renderer lifecycle behind the C ABI in its asynchronous form (renderer.h:553-603: start_render spawns the workers
and returns, main.cpp:158-163 polls every 0.5 s) and multi-GPU rendering INSIDE the library (config device_count /
device_ids: one host thread and stream per device, spp sharding on global sample indices as the reference's own worker
split renderer.h:352-360, films summed on the first device by k_combine over peer loads).

On a one-GPU box the multi-device path runs with several shards on the same device (B200PT_ALLOW_DUPLICATE_DEVICES); with
two or more GPUs visible the same tests also run across real devices, and torchrun checks the NCCL path."""
import json
import os
import re
import subprocess
import sys
import time

import numpy as np
import pytest

from conftest import ROOT, scene_path

pytestmark = pytest.mark.gpu

CLI = os.path.join(ROOT, "pathtrace_b200", "b200pt_main")


def _gpu_count():
    try:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=30).stdout
        return len([l for l in out.splitlines() if l.startswith("GPU ")])
    except Exception:
        return 0


def _cfg(scene, w, h, spp, **kw):
    import pathtrace_b200 as pt
    return pt.Config.from_json(pt.default_config_dict(width=w, height=h, samples=spp, scene=scene_path(scene), **kw))


def _paced(scene_name, cfg, **over):
    import pathtrace_b200 as pt
    scene = pt.Scene.from_file(cfg.scene_path, cfg.width, cfg.height, device=0)
    r = pt.renderer_from_config(scene, cfg.replace(**over) if over else cfg).start_render()
    rep = r.finalize()
    return scene, r, rep


def test_async_render_start_returns_and_poll_reports():
    """async_render = 1: start_render returns while the GPU works, poll never advances the render itself, a read in the
    middle returns a consistent (film, spp) pair, and the finished film is bit-equal to the paced render's."""
    import pathtrace_b200 as pt
    w, h, spp = 1280, 720, 64
    cfg = _cfg("book2_final", w, h, spp)
    scene, ref, ref_rep = _paced("book2_final", cfg, batch_samples=2)
    final = ref.framebuffer
    r = pt.renderer_from_config(scene, cfg.replace(batch_samples=2, async_render=1))
    r.start_render()
    seen, mid = [], None
    while True:
        p = r.sync_progress(0)
        seen.append(p.samples_done)
        if p.done:
            break
        if mid is None and 0 < p.samples_done < p.samples_total:
            fb = r.framebuffer  # a snapshot while the workers go on
            mid = (fb, r.spp_done)
        time.sleep(0.002)
    rep = r.finalize()
    assert seen[0] < w * h * spp, "start_render returned only when the render was over"
    assert seen == sorted(seen) and seen[-1] == w * h * spp
    assert len(set(seen)) > 2, "poll never saw the render in progress"
    assert np.array_equal(r.framebuffer, final)
    for k in ("rays_extension", "rays_shadow", "rays_reference", "camera_rays"):
        assert rep[k] == ref_rep[k], k
    assert mid is not None
    fb, n = mid
    assert 0 < n < spp and n % 2 == 0  # whole batches
    # the snapshot is the film of exactly its first n samples: the paced render of those samples, bit for bit
    part = _paced("book2_final", cfg.replace(samples=n), batch_samples=2)[1].framebuffer
    assert np.array_equal(fb, part)
    img, mx, _ = r.tonemap()
    img_ref, mx_ref, _ = ref.tonemap()
    assert np.array_equal(img, img_ref) and mx == mx_ref


def test_scene_destroyed_before_its_render():
    """A garbage collector may finalize the Scene object before the Wavefront that renders it (seen when a failing test left
    both to the cyclic collector): b200pt_scene_destroy then only drops the caller's reference, the render keeps working and
    its destruction releases the scene."""
    import pathtrace_b200 as pt
    cfg = _cfg("book2_final", 96, 54, 4)
    scene, r, _ = _paced("book2_final", cfg)
    before = r.framebuffer
    scene.close()                       # the handle is gone ...
    assert np.array_equal(r.framebuffer, before)   # ... the render is not
    img, mx, _ = r.tonemap()
    assert mx > 0 and img.mean() > 0
    r.close()
    # and a render destroyed twice / a scene closed twice are no-ops on the Python side
    r.close()
    scene.close()


@pytest.mark.parametrize("n_dev,async_render,no_peer", [(2, 0, False), (3, 1, False), (2, 1, True)])
def test_devices_inside_the_library(monkeypatch, n_dev, async_render, no_peer):
    """device_count shards behind ONE handle: the combined film equals the sum of the shards rendered one by one
    (fp32 sums in device order), counters add up, tonemap / PPM come from the combined film.  Real devices when the box has
    them, else several shards on device 0; no_peer forces the peer-copy fallback of k_combine's inputs."""
    import pathtrace_b200 as pt
    real = _gpu_count() >= n_dev
    if not real:
        monkeypatch.setenv("B200PT_ALLOW_DUPLICATE_DEVICES", "1")
    if no_peer:
        monkeypatch.setenv("B200PT_NO_PEER_LOADS", "1")
    devices = list(range(n_dev)) if real else [0] * n_dev
    w, h, spp = 96, 64, 11  # 11 samples over 2 or 3 devices: uneven shares
    cfg = _cfg("cornell_box_with_volume", w, h, spp, light_samples=3)
    scene = pt.Scene.from_file(cfg.scene_path, w, h, device=0)
    multi = pt.renderer_from_config(scene, cfg.replace(devices=devices, async_render=async_render, batch_samples=2,
                                                       stats_clamp=2.5)).start_render()
    rep = multi.finalize()
    fb = multi.framebuffer
    assert multi.spp_done == spp
    total = np.zeros((h, w, 3), np.float32)
    cs_tot, cq_tot = np.zeros_like(total), np.zeros_like(total)
    counts = dict(rays_extension=0, rays_shadow=0, rays_reference=0, camera_rays=0)
    base, extra, first = spp // n_dev, spp % n_dev, 0
    for i in range(n_dev):
        n_i = base + (1 if i < extra else 0)
        one = pt.renderer_from_config(scene, cfg.replace(samples=n_i, sample_offset=first, batch_samples=2,
                                                         stats_clamp=2.5)).start_render()
        rep_i = one.finalize()
        total = total + one.framebuffer  # device order, like k_combine
        cs, cq = one.stats()
        cs_tot, cq_tot = cs_tot + cs, cq_tot + cq
        for k in counts:
            counts[k] += rep_i[k]
        first += n_i
        one.close()
    assert np.array_equal(fb, total)
    for k in counts:
        assert rep[k] == counts[k], k
    cs, cq = multi.stats()
    assert np.array_equal(cs, cs_tot) and np.array_equal(cq, cq_tot)
    # the film of all devices in one paced single-device render: the same estimator, other summation order
    whole = pt.renderer_from_config(scene, cfg.replace(batch_samples=2)).start_render()
    whole.finalize()
    np.testing.assert_allclose(fb, whole.framebuffer, rtol=2e-5, atol=1e-5)
    img, mx, avg = multi.tonemap()
    whole.write_film(np.concatenate([fb, np.zeros((h, w, 1), np.float32)], axis=-1), spp)
    img1, mx1, avg1 = whole.tonemap()
    assert np.array_equal(img, img1) and mx == mx1 and abs(avg - avg1) <= 1e-6 * abs(avg1)


def test_cli_renders_on_devices_with_previews(tmp_path, monkeypatch):
    """b200pt_main with "devices": 2 in config.json (the key a user adds next to "threads"): asynchronous workers, the
    polling loop of main.cpp:158-163 rewriting the PPM from snapshots (sync_progress, renderer.h:605-620) -- kept here as
    preview files -- and a final image equal to the one-device run's."""
    import pathtrace_b200 as pt
    if _gpu_count() < 2:
        monkeypatch.setenv("B200PT_ALLOW_DUPLICATE_DEVICES", "1")
    w, h, spp = 640, 360, 48
    outs = {}
    for tag, extra in (("one", {}), ("two", {"devices": 2} if _gpu_count() >= 2 else {"device_ids": [0, 0]})):
        d = tmp_path / tag
        d.mkdir()
        cfgd = pt.default_config_dict(width=w, height=h, samples=spp, scene=scene_path("book2_final"),
                                      ppm_output_path="render.ppm", traced_paths_output_path="p.txt",
                                      traced_paths_2d_output_path="p2.txt", batch_samples=2, **extra)
        with open(d / "config.json", "w") as f:
            json.dump(cfgd, f)
        env = dict(os.environ, B200PT_CLI_POLL_MS="5", B200PT_CLI_KEEP_PREVIEWS=str(d / "preview_"))
        out = subprocess.run([CLI], cwd=d, capture_output=True, text=True, timeout=600, env=env)
        assert out.returncode == 0, out.stderr[-800:]
        header = f"P6\n{w} {h}\n255\n".encode()
        data = open(d / "render.ppm", "rb").read()
        assert data.startswith(header) and len(data) == len(header) + w * h * 3
        outs[tag] = (np.frombuffer(data[len(header):], dtype=np.uint8).astype(np.int32), out.stdout)
        previews = [(int(a), int(b)) for a, b in re.findall(r"preview (\d+) samples (\d+)", out.stdout)]
        assert len(previews) >= 2, "the render finished before the polling loop ever wrote a preview"
        ns = [n for _, n in previews]
        assert ns == sorted(ns) and 1 <= ns[0] and ns[-1] <= spp + 1
        final_mean = outs[tag][0].mean()
        for k, n in previews:
            pdata = open(d / f"preview_{k}.ppm", "rb").read()
            assert pdata.startswith(header) and len(pdata) == len(data)
            if n > 4:  # a few samples per pixel in: the preview already looks like the final picture
                pm = np.frombuffer(pdata[len(header):], dtype=np.uint8).mean()
                assert 0.6 * final_mean < pm < 1.6 * final_mean, (k, n, pm, final_mean)
        m = re.search(r"computed (\d+) camera rays", out.stdout)
        assert m and int(m.group(1)) == w * h * spp
    diff = np.abs(outs["one"][0] - outs["two"][0])
    assert diff.max() <= 1 and (diff > 0).mean() < 0.01  # the same film up to fp32 summation order
    assert "2 GPU(s)" in outs["two"][1]


def test_nccl_reduced_film_equals_single_gpu_film(tmp_path):
    """torchrun --nproc-per-node 2: the NCCL-reduced film of distributed.render_sharded on rank 0 is the one-GPU film of
    the same global samples (up to fp32 summation order), its tonemap divides by ALL samples, and every rank raises when
    there are fewer samples than ranks."""
    if _gpu_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "two_ranks.py"
    script.write_text('''
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, %r)
import pathtrace_b200 as pt
from pathtrace_b200 import distributed as ptd
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cfg = pt.Config.from_json(pt.default_config_dict(width=96, height=64, samples=9, scene=%r)).replace(device=local)
scene = pt.Scene.from_file(cfg.scene_path, 96, 64, device=local)
r, rep = ptd.render_sharded(scene, cfg, rank, world)
if rank == 0:
    img, mx, _ = r.tonemap()
    np.savez(%r, fb=r.framebuffer, img=img, mx=mx, rays=rep["rays"], paths=rep["camera_rays"])
try:
    ptd.render_sharded(scene, cfg.replace(samples=1), rank, world)
    raised = False
except ValueError:
    raised = True
assert raised
dist.barrier()
dist.destroy_process_group()
''' % (ROOT, scene_path("cornell_box"), str(tmp_path / "out.npz")))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-1500:]
    got = np.load(tmp_path / "out.npz")
    cfg = _cfg("cornell_box", 96, 64, 9)
    _, one, rep = _paced("cornell_box", cfg)
    np.testing.assert_allclose(got["fb"], one.framebuffer, rtol=2e-5, atol=1e-5)
    assert int(got["paths"]) == rep["camera_rays"] and int(got["rays"]) == rep["rays"]
    img, mx, _ = one.tonemap()
    assert abs(float(got["mx"]) - mx) <= 1e-5 * mx
    diff = np.abs(got["img"].astype(np.int32) - img.astype(np.int32))
    assert diff.max() <= 1 and (diff > 0).mean() < 0.01
