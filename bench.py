#!/usr/bin/env python
"""bench.py — headline benchmark of the volumetric hot path (BASELINE.json metric).

Metric: Msamples/s (camera paths completed per second) on the Disney-cloud-like 1280x720 volpath
workload (configs[1]); a "step" renders `--spp-per-step` samples per pixel of the full frame.

  python bench.py --gpus N --steps K --warmup W            # our arm (torchrun launches it for N > 1)
  python bench.py --impl reference ...                     # the reference's CPU path on host cores

One JSON line on rank 0.  `value` = device-resident throughput (kernel timed with CUDA events on the
launching stream, film reduce over NCCL included for N > 1); `e2e` = the same through the C-ABI
entry point with HOST film buffers (tables H2D + film D2H inside the timed region).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from acceleratedvolrenderer_b200 import scenes, sharding  # noqa: E402

WORKLOADS = {
    # Disney-cloud stand-in (the real wdas_cloud .nvdb is not in the reference repo and cannot be
    # fetched): fBm cloud at the quarter-res voxel count, 64^3 majorant grid as NanoVDBMedium uses
    # (media.cpp:521), high albedo, forward-peaked phase function, long paths.
    # sigma_t gives an optical depth of ~100 through the centre; camera framed so the
    # cloud fills about half of the frame as in the pbrt-v4-scenes disney-cloud view.
    "disney-cloud-standin-720p": dict(grid=(512, 352, 624), width=1280, height=720, sigma_t=200.0, albedo=0.999,
                                      g=0.877, max_depth=100, majorant_res=(64, 64, 64), eye=(0.0, 0.3, -2.4),
                                      fov=16.0),
    # small variant for smoke runs / CPU-box sanity
    "cloud-small": dict(grid=(96, 80, 104), width=320, height=180, sigma_t=200.0, albedo=0.999, g=0.877,
                        max_depth=100, majorant_res=(64, 64, 64), eye=(0.0, 0.3, -2.4), fov=16.0),
}


def fbm_density_torch(nx, ny, nz, device, seed=7, octaves=5, frequency=3.0, threshold=0.3, sharp=3.0):
    """Same procedural cloud as scenes.fbm_density, evaluated with torch (fast for 10^8 voxels)."""
    import torch
    rng = np.random.default_rng(seed)
    zs = (torch.arange(nz, device=device, dtype=torch.float32) + 0.5) / nz
    ys = (torch.arange(ny, device=device, dtype=torch.float32) + 0.5) / ny
    xs = (torch.arange(nx, device=device, dtype=torch.float32) + 0.5) / nx
    out = torch.zeros((nz, ny, nx), device=device, dtype=torch.float32)
    amp, freq, norm = 1.0, frequency, 0.0
    for _ in range(octaves):
        n = int(np.ceil(freq)) + 2
        lat = torch.from_numpy(rng.random((n, n, n), dtype=np.float32)).to(device)

        def axis(c):
            f = c * freq
            i = torch.floor(f).long()
            t = f - i
            return i, t * t * (3 - 2 * t)
        iz, tz = axis(zs); iy, ty = axis(ys); ix, tx = axis(xs)
        TX = tx[None, None, :]; TY = ty[None, :, None]; TZ = tz[:, None, None]

        def tap(dz, dy, dx):
            return lat[(iz + dz)[:, None, None], (iy + dy)[None, :, None], (ix + dx)[None, None, :]]
        lo = (tap(0, 0, 0) * (1 - TX) + tap(0, 0, 1) * TX) * (1 - TY) + (tap(0, 1, 0) * (1 - TX) + tap(0, 1, 1) * TX) * TY
        hi = (tap(1, 0, 0) * (1 - TX) + tap(1, 0, 1) * TX) * (1 - TY) + (tap(1, 1, 0) * (1 - TX) + tap(1, 1, 1) * TX) * TY
        out += amp * (lo * (1 - TZ) + hi * TZ)
        del lo, hi
        norm += amp
        amp *= 0.5
        freq *= 2.0
    out /= norm
    r = torch.sqrt((xs - 0.5)[None, None, :] ** 2 + (ys - 0.5)[None, :, None] ** 2 + (zs - 0.5)[:, None, None] ** 2) * 2
    out = out * torch.clamp(1.25 - r, 0, 1)
    return torch.clamp((out - threshold * 0.6) * sharp, 0, 1)


def make_density(w, device):
    nx, ny, nz = w["grid"]
    if device is not None:
        return fbm_density_torch(nx, ny, nz, device).cpu().numpy()
    return scenes.fbm_density(nx, ny, nz)


def make_scene(w, density, spp, exact):
    """The cloud as a "nanovdb" medium — the medium type of pbrt-v4-scenes' Disney cloud (media.cpp:628-665):
    index-space trilinear sampling, 64^3 majorant grid — inside a sphere "interface" boundary."""
    nx, ny, nz = w["grid"]
    return scenes.cloud_scene_nanovdb(density, voxel_size=1.0 / max(nx, ny, nz), width=w["width"], height=w["height"],
                                      spp=spp, sigma_t=w["sigma_t"], albedo=w["albedo"], g=w["g"],
                                      max_depth=w["max_depth"], exact_math=exact, eye=w.get("eye", (0.0, 0.35, -3.2)),
                                      fov=w.get("fov", 20.0))


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, smax, reasons = [], 0.0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); smax = max(smax, float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


def measured_traffic(workload, layout, exact, spp):
    """DRAM bytes per k_volpath launch from the committed ncu capture of this configuration (or None)."""
    p = os.path.join(ROOT, "profiles", "r1_volpath_dram_traffic.json")
    if not os.path.exists(p):
        return None, None
    for e in json.load(open(p))["entries"]:
        if e.get("superseded_by"):
            continue
        if (e["workload"], e["layout"], e["exact_math"], e["spp_per_launch"]) == (workload, layout, int(exact), spp):
            return e["dram_read_bytes"] + e["dram_write_bytes"], e["capture"]
    return None, None


def run_reference_arm(args, w, density):
    """--impl reference / cpu_baseline: the CPU path on all host threads, each step a bounded sample of the
    frame (a pixel lattice, 1 spp).  kind "reference": the reference's OWN NanoVDBMedium (over a real
    nanovdb grid built with its vendored GridBuilder) and VolPathIntegrator::Li, compiled from
    /root/reference into oracle/_ref/libavr_refprobe.so (shipped to the GPU box as a built artefact).
    kind "port": the oracle restatement, when that library is not there."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"], check=True)
    from cpu_checker import Checker, have_ref
    threads = os.cpu_count() or 1
    nd, rd = make_scene(w, density, 1, 1)
    W, H = w["width"], w["height"]
    kind = "reference" if have_ref() else "port"
    if kind == "port":
        orc = Checker("orc")
        mo = orc.medium_nanovdb(nd)

        def render(stride, seed):
            orc.lib.orc_set_pixel_stride(C.c_int(stride))
            mo.render(rd.with_(sampler_seed=seed), threads=threads)
            orc.lib.orc_set_pixel_stride(C.c_int(1))
    else:
        ref = Checker("ref")
        mo = ref.medium_nanovdb(nd)

        def render(stride, seed):
            xs, ys = np.arange(0, W, stride), np.arange(0, H, stride)
            smp = np.stack([np.tile(xs, ys.size), np.repeat(ys, xs.size), np.zeros(xs.size * ys.size)], axis=1).astype(np.int32)
            r = rd.with_(sampler_seed=seed)
            smp, rays = mo.camera_rays(r, smp)
            # interleaved shares: neighbouring pixels cost alike, so every thread gets the same mix
            parts = [(smp[k::threads * 4], rays[k::threads * 4]) for k in range(threads * 4)]
            it = iter(parts)
            lock = threading.Lock()

            def work():
                while True:
                    with lock:
                        job = next(it, None)
                    if job is None:
                        return
                    mo.volpath_li(r, job[0], job[1])  # ctypes releases the GIL inside the reference code
            pool = [threading.Thread(target=work) for _ in range(threads)]
            for t in pool:
                t.start()
            for t in pool:
                t.join()

    # bounded sample: every `stride`-th pixel in x and y of the whole frame, stride calibrated so that one
    # step is a few seconds of CPU work
    stride = 16
    t = time.perf_counter(); render(stride, 0); cal = time.perf_counter() - t
    while stride > 1 and cal * 4 < 4.0:
        stride //= 2; cal *= 4
    cw, ch = (W + stride - 1) // stride, (H + stride - 1) // stride
    for i in range(args.warmup):
        render(stride, 1 + i)
    t = time.perf_counter()
    for i in range(args.steps):
        render(stride, 100 + i)
    dt = time.perf_counter() - t
    mo.close()
    value = cw * ch * args.steps / dt / 1e6
    what = ("the reference's NanoVDBMedium + VolPathIntegrator::Li (oracle/_ref)" if kind == "reference"
            else "oracle port of the reference renderer")
    lattice = "every pixel" if stride == 1 else f"every {stride}th pixel in x and y"
    sample_desc = f"{lattice} of {W}x{H} ({cw}x{ch} paths), 1 spp per step; {what}"
    return value, dt / args.steps * 1e3, threads, sample_desc, kind


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="avr_b200")
    ap.add_argument("--workload", default=os.environ.get("AVR_BENCH_WORKLOAD", "disney-cloud-standin-720p"))
    ap.add_argument("--spp-per-step", type=int, default=16)
    ap.add_argument("--layout", default="octet32")
    ap.add_argument("--exact", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    w = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    base = {"metric": "Msamples/s on Disney cloud 720p volpath", "unit": "Msamples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "medium": "nanovdb (dense fBm stand-in flattened over its index box)",
                       "film": [w["width"], w["height"]], "grid": list(w["grid"]),
                       "majorant_res": list(w["majorant_res"]), "sigma_t": w["sigma_t"], "albedo": w["albedo"],
                       "g": w["g"], "max_depth": w["max_depth"], "sampler": "independent",
                       "l2": "density store (octets, 3.6 GB at 720p workload) is far larger than the 126 MB L2"}}

    if args.impl == "reference":
        if rank != 0:
            return 0
        import torch  # only for the density generator when a GPU is around
        dev = "cuda:0" if torch.cuda.is_available() else None
        density = make_density(w, dev) if dev else scenes.fbm_density(*w["grid"])
        value, ms, threads, sample_desc, kind = run_reference_arm(args, w, density)
        out = dict(base)
        out.update({"impl": "reference", "value": value, "ms_per_step": ms, "n_gpus": args.gpus,
                    "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": threads, "kind": kind,
                                     "sample": sample_desc},
                    "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                    "gpu_launches": 0})
        out["config"]["spp_per_step"] = 1
        print(json.dumps(out))
        return 0

    import torch
    import torch.distributed as dist
    from acceleratedvolrenderer_b200.api import Context
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device and no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    density = make_density(w, dev)
    spp = args.spp_per_step
    nd, rd = make_scene(w, density, spp, args.exact)
    ctx = Context(local_rank)
    t0 = time.perf_counter()
    med = ctx.create_medium("nanovdb", nd, layout=args.layout)
    build_s = time.perf_counter() - t0
    W, H = w["width"], w["height"]
    film = torch.zeros((H, W, 4), dtype=torch.float64, device=dev)
    # weak scaling by sample index: every rank renders its own `spp` samples of every pixel
    rdr = sharding.shard_samples(rd, rank, world, spp)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def device_step(step):
        film.zero_()
        r = rdr.with_(sampler_seed=step)
        st = ctx.render_volpath_device(med, r, film.data_ptr(), want_stats=True)  # syncs the library stream
        sharding.reduce_film(film, dst=0)
        return st

    for i in range(args.warmup):
        device_step(i)
    sync_all()
    clocks = ClockSampler(local_rank)
    clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stats_acc = {"camera_fetches": 0, "shadow_fetches": 0, "segments": 0, "camera_rays": 0, "kernel_ms": 0.0}
    t_wall = time.perf_counter()
    ev0.record()
    for i in range(args.steps):
        st = device_step(args.warmup + i)
        for k in stats_acc:
            stats_acc[k] += st[k]
    ev1.record()
    sync_all()
    wall = time.perf_counter() - t_wall
    clk = clocks.stop()
    # kernel time on the library's launching stream (CUDA events inside the library) + reduce, as wall
    # between synchronised points; take the max over ranks
    tt = torch.tensor([wall, stats_acc["kernel_ms"] / 1e3], dtype=torch.float64, device=dev)
    cnt = torch.tensor([stats_acc["camera_rays"], stats_acc["camera_fetches"], stats_acc["shadow_fetches"],
                        stats_acc["segments"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    wall_max, kern_max = float(tt[0]), float(tt[1])
    rays, cam_f, sh_f, segs = [float(x) for x in cnt]
    samples_per_step = W * H * spp * world
    value = samples_per_step * args.steps / wall_max / 1e6

    # e2e through the host-buffer C-ABI call (rank-local partition; host film accumulate)
    host_film = np.zeros((H, W, 4), dtype=np.float64)
    ctx.render_volpath(med, rdr, film=host_film)  # warm
    sync_all()
    t = time.perf_counter()
    e2e_steps = max(1, args.steps)
    for i in range(e2e_steps):
        # progressive accumulation: every step adds its samples to the same host film (the C entry ADDS)
        ctx.render_volpath(med, rdr.with_(sampler_seed=100 + i), film=host_film)
    if world > 1:
        dist.barrier()
    e2e_dt = time.perf_counter() - t
    te = torch.tensor([e2e_dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = samples_per_step * e2e_steps / float(te[0]) / 1e6

    if rank == 0:
        peaks, peak_kind = measured_peaks()
        fetch_bytes = 32 if args.layout != "octet16" else 16
        alg_bytes = (cam_f + sh_f) * fetch_bytes + segs * 4
        achieved = alg_bytes / max(kern_max, 1e-9) / 1e9 / max(world, 1)  # per-GPU GB/s of the dominant kernel
        traffic, traffic_src = measured_traffic(args.workload, args.layout, args.exact, spp)
        out = dict(base)
        out.update({"impl": "avr_b200", "value": value, "ms_per_step": wall_max / args.steps * 1e3,
                    "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": 4 * 471 * 4,
                            "d2h_bytes_per_step": W * H * 4 * 8},
                    "gpu_launches": args.steps,
                    "roofline": {"bound": "hbm", "kernel": "k_volpath", "achieved": achieved, "peak": peaks["hbm_gbs"],
                                 "peak_kind": peak_kind, "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                                 "traffic": traffic, "traffic_unit": "bytes per launch (ncu dram read+write)",
                                 "traffic_capture": traffic_src, "algorithmic_bytes": alg_bytes / args.steps / max(world, 1),
                                 "fetches_per_step": (cam_f + sh_f) / args.steps,
                                 "segments_per_step": segs / args.steps, "bytes_per_fetch": fetch_bytes,
                                 "kernel_ms_per_step": kern_max / args.steps * 1e3},
                    "clocks": clk, "medium_build_s": build_s, "medium_bytes": med.bytes()})
        out["config"].update({"spp_per_step": spp, "layout": args.layout, "exact_math": args.exact,
                              "parallelism": f"sample-index sharded x{world} ({spp} spp per rank per step), film ncclReduce" if world > 1 else "1 gpu"})
        if not args.no_cpu_baseline:
            try:
                v, ms, threads, sample_desc, kind = run_reference_arm(argparse.Namespace(steps=1, warmup=0), w, density)
                out["cpu_baseline"] = {"value": v, "unit": "Msamples/s", "cores": threads, "kind": kind,
                                       "sample": sample_desc}
            except Exception as e:  # the GPU numbers stand on their own
                out["cpu_baseline"] = {"value": None, "error": str(e)}
        print(json.dumps(out))
    med.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
