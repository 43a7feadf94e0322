#!/usr/bin/env python
"""bench.py — headline benchmark of the volumetric hot path (BASELINE.json metric).

Metric: Msamples/s (camera paths completed per second) on the Disney-cloud-like 1280x720 volpath
workload (configs[1]); a "step" renders `--spp-per-step` samples per pixel of the full frame.

  python bench.py --gpus N --steps K --warmup W            # our arm (torchrun launches it for N > 1)
  python bench.py --impl reference ...                     # the reference's CPU path on host cores
  python bench.py --workload graph-1e7 ...                 # config 4: radiance propagation (Mvertex-bounces/s)

One JSON line on rank 0.
  value      device-resident throughput of the arithmetic mode the adapter ships (FAST, exact_math = 0): render
             kernel + film ncclReduce for N > 1, timed with CUDA events on the launching stream, max over ranks
  exact      the same measurement with exact_math = 1 (the mode whose per-ray parity is pinned), same command
  e2e        the same metric through the C-ABI call a host makes: avr_render_volpath (N = 1) /
             avr_render_volpath_multi (N > 1) with a HOST film — tables H2D, render, reduce, film D2H + add
  parity     GPU film (FAST and EXACT, >= 256 spp) against the reference's own pixels of the SAME workload:
             z-scores of 16x16-pixel block means, from the reference's per-block sample variance
  roofline   algorithmic bytes of the dominant kernel / its CUDA-event time, against the measured HBM peak
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
    "disney-cloud-standin-720p": dict(kind="volpath", grid=(512, 352, 624), width=1280, height=720, sigma_t=200.0,
                                      albedo=0.999, g=0.877, max_depth=100, majorant_res=(64, 64, 64),
                                      eye=(0.0, 0.3, -2.4), fov=16.0, spp_per_step=16),
    # configs[2]: "Disney cloud HD, high-albedo multiple scattering (long paths)": 1920x1080, depth 1000
    "disney-cloud-standin-hd": dict(kind="volpath", grid=(512, 352, 624), width=1920, height=1080, sigma_t=200.0,
                                    albedo=0.999, g=0.877, max_depth=1000, majorant_res=(64, 64, 64),
                                    eye=(0.0, 0.3, -2.4), fov=16.0, spp_per_step=8),
    # small variant for smoke runs / CPU-box sanity
    "cloud-small": dict(kind="volpath", grid=(96, 80, 104), width=320, height=180, sigma_t=200.0, albedo=0.999, g=0.877,
                        max_depth=100, majorant_res=(64, 64, 64), eye=(0.0, 0.3, -2.4), fov=16.0, spp_per_step=16),
    # configs[4] (C5): synthetic cubes with the media the general / chromatic kernels serve: RGB grids
    # ({scale,c0,c1,c2} sigmoid-polynomial voxels for sigma_a and sigma_s), an emissive grid (blackbody from a
    # temperature grid) and a scalar grid with wavelength-dependent sigma_a / sigma_s.  Same parameters as
    # tools/config_sweep.py; the CPU arm of these workloads is the oracle port on a pixel crop.
    "c5-rgbgrid-512": dict(kind="volpath", medium="rgbgrid", grid=(512, 512, 512), width=1280, height=720, sigma_t=8.08,
                           albedo=1 / 1.01, g=0.0, max_depth=100, majorant_res=(16, 16, 16), fov=24.0, spp_per_step=4),
    "c5-emissive-512": dict(kind="volpath", medium="emissive", grid=(512, 512, 512), width=1280, height=720, sigma_t=8.08,
                            albedo=1 / 1.01, g=0.0, max_depth=100, majorant_res=(16, 16, 16), fov=24.0, spp_per_step=4),
    "c5-chromatic-512": dict(kind="volpath", medium="chromatic", grid=(512, 512, 512), width=1280, height=720,
                             sigma_t=64.64, albedo=1 / 1.01, g=0.877, max_depth=100, majorant_res=(16, 16, 16), fov=24.0,
                             spp_per_step=4),
    # configs[3]: src/graph propagation, 8 iterations, synthetic random-geometric transport matrix
    "graph-1e7": dict(kind="graph", n=10_000_000, avg_deg=6, bounces=8),
    "graph-1e5": dict(kind="graph", n=100_000, avg_deg=6, bounces=8),
}
BLOCK = 16  # parity statistic: block means over BLOCK x BLOCK pixels


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


def c5_scene(w, dens, spp):
    """(medium kind, medium desc, render desc, layout, algorithmic bytes per fetch) of a C5 workload; dens: numpy."""
    from acceleratedvolrenderer_b200._abi import GridDesc, RGBGridDesc
    kw = dict(width=w["width"], height=w["height"], spp=spp, sigma_t=w["sigma_t"], albedo=w["albedo"], g=w["g"],
              max_depth=w["max_depth"], exact_math=0, fov=w["fov"])
    if w["medium"] == "chromatic":
        gd, rd = scenes.cloud_scene(dens, chromatic=True, **kw)
        return "uniformgrid", gd, rd, "octet32", 32
    gd, rd = scenes.cloud_scene(dens, **kw)
    if w["medium"] == "emissive":   # temperature = 300 + 2000 d, Lescale = 1: density + temperature + Lescale taps
        ge = GridDesc(dens, gd.sigma_a, gd.sigma_s, render_from_medium=gd.m, g=w["g"], temperature=300 + 2000 * dens,
                      lescale=np.ones_like(dens), le_norm=1.0)
        return "uniformgrid", ge, rd, "dense", 96
    r2 = np.random.default_rng(7)   # sigma_s = d (1, .9, .8) k, sigma_a = d (.02, .05, .1) k as sigmoid-polynomial voxels
    ss = scenes.rgb_voxels(dens * 8.0, r2); sa = scenes.rgb_voxels(dens * 0.4, r2)
    return "rgbgrid", RGBGridDesc(sigma_a=sa, sigma_s=ss, render_from_medium=gd.m, g=w["g"]), rd, "dense", 256


def c5_cpu_arm(kind, desc, rd, seconds):
    """The oracle port (kind "port") on a centred crop of the frame, all host threads, ~`seconds` of work."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"], check=True)
    import cpu_checker
    ck = cpu_checker.Checker("orc")
    mo = ck.medium_rgbgrid(desc) if kind == "rgbgrid" else ck.medium_uniformgrid(desc)
    threads = os.cpu_count() or 1
    W, H = rd.c.width, rd.c.height
    cw, ch = 64, 36
    n, k, t = 0, 0, time.perf_counter()
    while k < 2 or (time.perf_counter() - t < seconds and k < 64):
        r = rd.with_(pixel_bounds=(W // 2 - cw // 2, H // 2 - ch // 2, W // 2 + cw // 2, H // 2 + ch // 2), sampler_seed=100 + k,
                     exact_math=1)
        mo.render(r, threads=threads)
        n += cw * ch * rd.c.spp
        k += 1
    dt = time.perf_counter() - t
    mo.close()
    return n / dt / 1e6, threads, f"{k} steps of a centred {cw}x{ch} pixel crop x {rd.c.spp} spp; oracle/ CPU restatement (port)"


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
    p = os.path.join(ROOT, "profiles", "volpath_dram_traffic.json")
    if not os.path.exists(p):
        return None, None
    for e in json.load(open(p))["entries"]:
        if e.get("superseded_by"):
            continue
        if (e["workload"], e["layout"], e["exact_math"], e["spp_per_launch"]) == (workload, layout, int(exact), spp):
            return e["dram_read_bytes"] + e["dram_write_bytes"], e["capture"]
    return None, None


# ---- the reference's CPU path (--impl reference, cpu_baseline, and the reference pixels of `parity`) ---------
def film_values(rd, out12):
    """Per-sample sensor RGB exactly as RGBFilm::AddSample forms it before the sum (film.h:95-100,239-255, weight 1):
    imagingRatio * mean_i(sensor_c(lambda_i) * L_i / pdf_i); NaN / Inf radiance -> black (cpu/integrators.cpp:286-296)."""
    L = out12[:, 0:4].astype(np.float64); lam = out12[:, 4:8]; pdf = out12[:, 8:12].astype(np.float64)
    bad = ~np.isfinite(L).all(axis=1)
    idx = np.clip(np.rint(lam).astype(np.int64) - 360, 0, 470)
    inside = (lam >= 360) & (lam <= 830) & (pdf != 0)
    q = np.where(inside, L / np.where(pdf != 0, pdf, 1.0), 0.0)
    sens = rd.sensor_rgb.astype(np.float64)
    rgb = np.stack([(sens[c][idx] * q).mean(axis=1) for c in range(3)], axis=1) * float(rd.c.imaging_ratio)
    rgb[bad] = 0.0
    return rgb


class ReferenceArm:
    """The reference's OWN NanoVDBMedium (over a real nanovdb grid built with its vendored GridBuilder) and
    VolPathIntegrator::Li, compiled from /root/reference into oracle/_ref (shipped to the GPU box as a built
    artefact): the timing flavour (-O3 -march=x86-64-v3, contraction on, as the reference's Release build) when it
    is there, else the parity flavour; kind "port" (the oracle restatement) only when neither exists.  All host
    threads, the medium and its grid built once, every sample goes through the film arithmetic (film_values)."""

    def __init__(self, w, density, spp):
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"], check=True)
        import cpu_checker
        self.threads = os.cpu_count() or 1
        self.w, self.W, self.H, self.spp = w, w["width"], w["height"], spp
        self.nd, self.rd = make_scene(w, density, spp, 1)
        timing = os.path.join(ROOT, "oracle", "_ref", "libavr_refprobe_timing.so")
        parity_path = cpu_checker.REF_PATH
        if os.path.exists(timing):
            cpu_checker.REF_PATH = timing
            self.kind, self.flags = "reference", "-O3 -march=x86-64-v3 -ffp-contract=fast"
        elif cpu_checker.have_ref():
            self.kind, self.flags = "reference", "-O2 -ffp-contract=off (parity build)"
        else:
            self.kind, self.flags = "port", "-O2 -ffp-contract=off (oracle restatement)"
        try:
            self.ck = cpu_checker.Checker("ref" if self.kind == "reference" else "orc")
        finally:
            cpu_checker.REF_PATH = parity_path
        self.mo = self.ck.medium_nanovdb(self.nd)
        nb = ((self.H + BLOCK - 1) // BLOCK) * ((self.W + BLOCK - 1) // BLOCK)
        self.bsum = np.zeros(nb); self.bsq = np.zeros(nb); self.bcnt = np.zeros(nb)
        self.lattice = None

    def close(self):
        self.mo.close()

    def render(self, stride, seed):
        """One step: `spp` samples of every `stride`-th pixel in x and y.  Returns the number of samples."""
        W, H, spp = self.W, self.H, self.spp
        xs, ys = np.arange(0, W, stride), np.arange(0, H, stride)
        px = np.tile(np.repeat(xs, spp), ys.size); py = np.repeat(ys, xs.size * spp)
        si = np.tile(np.arange(spp), xs.size * ys.size)
        smp = np.stack([px, py, si], axis=1).astype(np.int32)
        r = self.rd.with_(sampler_seed=seed)
        if self.kind == "port":
            # the restatement renders films, not sample lists: timing only, no pixels for `parity`
            self.ck.lib.orc_set_pixel_stride(C.c_int(stride))
            self.mo.render(r, threads=self.threads)
            self.ck.lib.orc_set_pixel_stride(C.c_int(1))
            return smp.shape[0]
        else:
            smp, rays = self.mo.camera_rays(r, smp)
            out = np.zeros((smp.shape[0], 12), dtype=np.float32)
            # interleaved shares: neighbouring pixels cost alike, so every thread gets the same mix
            parts = list(range(self.threads * 4))
            it = iter(parts)
            lock = threading.Lock()
            step = self.threads * 4

            def work():
                while True:
                    with lock:
                        k = next(it, None)
                    if k is None:
                        return
                    o, _ = self.mo.volpath_li(r, smp[k::step], rays[k::step])  # ctypes releases the GIL
                    out[k::step] = o
            pool = [threading.Thread(target=work) for _ in range(self.threads)]
            for t in pool:
                t.start()
            for t in pool:
                t.join()
        # the film: sensor RGB of every sample, binned into BLOCK x BLOCK pixel blocks (Y channel) for `parity`
        y = film_values(self.rd, out)[:, 1]
        nbx = (W + BLOCK - 1) // BLOCK
        b = (smp[:, 1] // BLOCK) * nbx + smp[:, 0] // BLOCK
        n = self.bsum.size
        self.bsum += np.bincount(b, weights=y, minlength=n)
        self.bsq += np.bincount(b, weights=y * y, minlength=n)
        self.bcnt += np.bincount(b, minlength=n)
        self.lattice = (stride if self.lattice in (None, stride) else -1)
        return smp.shape[0]

    def calibrate(self, target_s):
        """Lattice stride such that one step is about target_s seconds of CPU work."""
        stride = 32
        t = time.perf_counter(); self.render(stride, 900); cal = time.perf_counter() - t
        while stride > 1 and cal * 4 < target_s:
            stride //= 2; cal *= 4
        self.bsum[:] = 0; self.bsq[:] = 0; self.bcnt[:] = 0; self.lattice = None
        return stride

    def describe(self, stride):
        W, H = self.W, self.H
        cw, ch = (W + stride - 1) // stride, (H + stride - 1) // stride
        what = ("the reference's NanoVDBMedium + VolPathIntegrator::Li + film arithmetic (oracle/_ref, "
                if self.kind == "reference" else "oracle port of the reference renderer (")
        lattice = "every pixel" if stride == 1 else f"every {stride}th pixel in x and y"
        return f"{lattice} of {W}x{H} ({cw}x{ch} pixels) x {self.spp} spp per step; {what}{self.flags})"


def lattice_blocks(W, H, stride):
    """Block index of every lattice pixel (row-major over ys x xs) and the lattice pixel count per block."""
    nbx = (W + BLOCK - 1) // BLOCK
    nb = nbx * ((H + BLOCK - 1) // BLOCK)
    xs, ys = np.arange(0, W, stride), np.arange(0, H, stride)
    bidx = ((ys // BLOCK)[:, None] * nbx + (xs // BLOCK)[None, :]).ravel()
    return xs, ys, bidx, np.bincount(bidx, minlength=nb).astype(np.float64)


def parity_block(ref: ReferenceArm, gpu_steps, spp_step):
    """z-scores of the BLOCK x BLOCK block means (film Y) of the GPU renders against the reference's, over the pixel
    lattice the reference rendered.  gpu_steps[name] = array (k, blocks): block means of k independent GPU steps.
      z = (gpu - ref) / sqrt(var_ref / n_ref + SE_gpu^2): var_ref = the reference's own per-block sample variance,
      SE_gpu^2 = the variance of the k GPU step means / k.  Under "same distribution" the z are N(0, 1): rms_z ~ 1,
      max |z| over ~1500 lit blocks ~ 3.5; a relative bias b of the GPU estimator shows as global_mean_z = b / (the
      relative standard error of the reference's global mean, global_mean_rel_se)."""
    W, H, stride = ref.W, ref.H, ref.lattice
    if not stride or stride < 0 or ref.bcnt.sum() == 0:
        return None
    _, _, _, npix = lattice_blocks(W, H, stride)
    cnt = np.maximum(ref.bcnt, 1)
    m_ref = ref.bsum / cnt
    var_ref = np.maximum(ref.bsq / cnt - m_ref ** 2, 0) * cnt / np.maximum(cnt - 1, 1)
    out = {"block": BLOCK, "lattice_stride": stride, "reference_samples": int(ref.bcnt.sum()),
           "statistic": "film Y (sensor g-bar) block means over the reference's pixel lattice",
           "reference": f"{ref.kind} ({ref.flags})"}
    dump = {"m_ref": m_ref, "var_ref": var_ref, "cnt": ref.bcnt, "npix": npix}
    for name, steps in gpu_steps.items():
        k = steps.shape[0]
        m_gpu = steps.mean(axis=0)
        se2_gpu = steps.var(axis=0, ddof=1) / k
        lit = (npix > 0) & ((var_ref > 0) | (se2_gpu > 0))
        d = m_gpu[lit] - m_ref[lit]
        se = np.sqrt(var_ref[lit] / cnt[lit] + se2_gpu[lit])
        z = d / se
        wgt = npix[lit] / npix[lit].sum()
        gse = float(np.sqrt(((se * wgt) ** 2).sum()))
        gref = float((m_ref[lit] * wgt).sum())
        out[name] = {"gpu_spp": int(k * spp_step), "blocks": int(lit.sum()),
                     "blocks_black_in_both": int(((npix > 0) & ~lit).sum()),
                     "max_abs_z": float(np.abs(z).max()), "rms_z": float(np.sqrt((z * z).mean())),
                     "frac_abs_z_lt_4": float((np.abs(z) < 4).mean()), "frac_abs_z_gt_2": float((np.abs(z) > 2).mean()),
                     "global_mean_ratio": float((m_gpu[lit] * wgt).sum() / gref),
                     "global_mean_z": float((d * wgt).sum() / gse), "global_mean_rel_se": gse / gref}
        dump["m_" + name] = m_gpu; dump["se2_" + name] = se2_gpu
    if os.environ.get("AVR_BENCH_DUMP"):
        np.savez(os.environ["AVR_BENCH_DUMP"], **dump)
    return out


def run_reference_arm(args, w, density, spp):
    """--impl reference: W warm-up + K timed steps, each a bounded lattice sample of the repo arm's step
    (same spp per step) sized so that the run ends within a few minutes."""
    ref = ReferenceArm(w, density, spp)
    total = max(1, args.steps + args.warmup)
    stride = ref.calibrate(min(8.0, 150.0 / total))
    for i in range(args.warmup):
        ref.render(stride, 1 + i)
    t = time.perf_counter()
    n = 0
    for i in range(args.steps):
        n += ref.render(stride, 100 + i)
    dt = time.perf_counter() - t
    desc = ref.describe(stride)
    ref.close()
    return n / dt / 1e6, dt / max(args.steps, 1) * 1e3, ref.threads, desc, ref.kind


# ---- config 4: graph propagation ------------------------------------------------------------------------------
def run_graph(args, w, rank, world, local_rank, base):
    import torch
    import torch.distributed as dist
    from acceleratedvolrenderer_b200.api import Context
    n, bounces = w["n"], w["bounces"]
    rowptr, col, val = scenes.random_graph(n, avg_deg=w["avg_deg"], seed=5)
    L0h = np.random.default_rng(5).random(n, dtype=np.float32)
    nnz = int(col.size)
    base = dict(base)
    base.update({"metric": "Mvertex-bounces/s on src/graph radiance propagation", "unit": "Mvertex-bounces/s",
                 "scaling": "strong", "dtype": "f32",
                 "config": {"workload": args.workload, "vertices": n, "nnz": nnz, "bounces": bounces,
                            "matrix": "synthetic random-geometric transport graph (scenes.random_graph, seed 5)"}})
    if args.impl == "reference":
        if rank != 0:
            return 0
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"], check=True)
        import cpu_checker
        kind = "reference" if cpu_checker.have_ref() else "port"
        ck = cpu_checker.Checker("ref" if kind == "reference" else "orc")
        for _ in range(min(args.warmup, 1)):
            ck.graph_propagate(rowptr, col, val, L0h, bounces)
        steps = max(1, min(args.steps, 5))
        t = time.perf_counter()
        for _ in range(steps):
            ck.graph_propagate(rowptr, col, val, L0h, bounces)
        dt = time.perf_counter() - t
        v = n * bounces * steps / dt / 1e6
        out = dict(base)
        out.update({"impl": "reference", "value": v, "ms_per_step": dt / steps * 1e3, "steps": steps, "gpu_launches": 0,
                    "cpu_baseline": {"value": v, "unit": base["unit"], "cores": 1, "kind": kind,
                                     "sample": f"{steps} full propagations; LightingCalculator::ComputeFinalLight "
                                               "(Eigen sparse product, one core as in the reference)"},
                    "e2e": {"value": v, "unit": base["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
        print(json.dumps(out))
        return 0
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ctx = Context(local_rank)
    g = ctx.create_graph(rowptr, col, val)
    L0 = torch.from_numpy(L0h).to(dev)
    total = torch.zeros(n, dtype=torch.float32, device=dev)
    comm = sharding.init_comm(ctx) if world > 1 else None

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def step():
        if comm is not None:
            comm.propagate(g, L0.data_ptr(), bounces, total.data_ptr(), want_iters=False)
        else:
            g.propagate_device(L0.data_ptr(), bounces, total.data_ptr())

    for _ in range(max(args.warmup, 3)):
        step()
    sync_all()
    clocks = ClockSampler(local_rank)
    clocks.start()
    t = time.perf_counter()
    for _ in range(args.steps):
        step()
    sync_all()
    dt = time.perf_counter() - t
    clk = clocks.stop()
    tt = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dt = float(tt[0])
    value = n * bounces * args.steps / dt / 1e6
    # N > 1: what the step costs WITHOUT the collective — the same `bounces` SpMVs over this rank's rows, no
    # all-gather, no finiteness pass (max over ranks) — so that the line names the all-gather's share of the step
    spmv_only_ms = None
    if world > 1:
        chunk = (n + world - 1) // world
        r0, r1 = min(rank * chunk, n), min((rank + 1) * chunk, n)
        nxt = torch.empty(n, dtype=torch.float32, device=dev)
        flag = torch.tensor([0, 2 ** 31 - 1], dtype=torch.int32, device=dev)
        for _ in range(3):
            g.spmv_rows_device(L0.data_ptr(), nxt.data_ptr(), r0, r1, flag.data_ptr())
        sync_all()
        t = time.perf_counter()
        for _ in range(args.steps * bounces):
            g.spmv_rows_device(L0.data_ptr(), nxt.data_ptr(), r0, r1, flag.data_ptr())
        torch.cuda.synchronize()
        ts = torch.tensor([time.perf_counter() - t], dtype=torch.float64, device=dev)
        dist.all_reduce(ts, op=dist.ReduceOp.MAX)
        spmv_only_ms = float(ts[0]) / args.steps * 1e3
    # e2e: host vectors through avr_graph_propagate (rank 0 alone: the host-buffer entry is single-GPU)
    e2e = None
    if rank == 0:
        g.compute_final_light(L0h, bounces)
        t = time.perf_counter()
        k = max(1, min(args.steps, 5))
        for _ in range(k):
            g.compute_final_light(L0h, bounces)
        e2e = n * bounces * k / (time.perf_counter() - t) / 1e6
    if world > 1:
        dist.barrier()
    if rank == 0:
        peaks, peak_kind = measured_peaks()
        alg = bounces * (12.0 * nnz + 16.0 * n) / max(world, 1)   # per GPU: its rows' share
        achieved = alg / (dt / args.steps) / 1e9
        out = dict(base)
        out.update({"impl": "avr_b200", "value": value, "ms_per_step": dt / args.steps * 1e3,
                    "e2e": {"value": e2e, "unit": base["unit"], "h2d_bytes_per_step": 4 * n, "d2h_bytes_per_step": 4 * n,
                            "note": "avr_graph_propagate on rank 0 (host L0 / total)"},
                    "gpu_launches": args.steps * (bounces + 2),
                    "collective": None if spmv_only_ms is None else {
                        "spmv_only_ms_per_step": spmv_only_ms,
                        "exchange_share": max(0.0, 1.0 - spmv_only_ms / (dt / args.steps * 1e3)),
                        "bytes_gathered_per_bounce": 4 * n,
                        "exchange": {1: "peer memory: k_spmv_p2p stores every row into all ranks' vectors, "
                                        "k_p2p_barrier between bounces", -1: "ncclAllGather per bounce"}.get(comm.graph_exchange(), "?"),
                        "note": "share of the step that is not this rank's SpMV arithmetic (the exchange and its "
                                "synchronisation); spmv_only = the same SpMVs writing to local memory alone"},
                    "roofline": {"bound": "hbm", "kernel": "k_spmv", "achieved": achieved, "peak": peaks["hbm_gbs"],
                                 "peak_kind": peak_kind, "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"], "traffic": None,
                                 "algorithmic_bytes": alg,
                                 "note": "whole step (SpMV + exchange + checks) / (12 nnz + 16 N) bytes per bounce per GPU"},
                    "clocks": clk})
        how = "the exchange fused into the SpMV kernel over peer memory" if comm is not None and comm.graph_exchange() == 1 \
            else "ncclAllGather per bounce"
        out["config"]["parallelism"] = (f"vertex-range rows x{world}, {how}, one CUDA graph per step" if world > 1 else "1 gpu")
        print(json.dumps(out))
    if comm is not None:
        comm.close()
    g.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="avr_b200")
    ap.add_argument("--workload", default=os.environ.get("AVR_BENCH_WORKLOAD", "disney-cloud-standin-720p"))
    ap.add_argument("--spp-per-step", type=int, default=0)
    ap.add_argument("--layout", default="octet32")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-exact", action="store_true", help="skip the exact_math=1 measurement")
    ap.add_argument("--parity-spp", type=int, default=256)
    args = ap.parse_args()
    w = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    base = {"metric": "Msamples/s on Disney cloud 720p volpath", "unit": "Msamples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
    if w["kind"] == "graph":
        return run_graph(args, w, rank, world, local_rank, base)
    spp = args.spp_per_step or w["spp_per_step"]
    # `config` names the workload only (identical in both arms); what is specific to our kernels is `kernel_config`
    base["config"] = {"workload": args.workload, "medium": "nanovdb (dense fBm stand-in flattened over its index box)",
                      "film": [w["width"], w["height"]], "grid": list(w["grid"]),
                      "majorant_res": list(w["majorant_res"]), "sigma_t": w["sigma_t"], "albedo": w["albedo"],
                      "g": w["g"], "max_depth": w["max_depth"], "sampler": "independent", "spp_per_step": spp,
                      "l2": "inputs (3.6 GB octet store at this workload) are far larger than the 126 MB L2: no flush between steps"}

    c5 = w.get("medium", "nanovdb") != "nanovdb"
    if c5:
        base["metric"] = "Msamples/s on synthetic C5 cube volpath"
        base["config"]["medium"] = {"rgbgrid": "rgbgrid: sigma_a, sigma_s as RGBUnboundedSpectrum voxels (general kernel)",
                                    "emissive": "uniformgrid + temperature grid, blackbody Le (general kernel)",
                                    "chromatic": "uniformgrid, wavelength-dependent sigma_a / sigma_s (4-channel kernel)"}[w["medium"]]
        base["config"]["l2"] = "inputs are far larger than the 126 MB L2: no flush between steps"
    if args.impl == "reference" and c5:
        if rank != 0:
            return 0
        import torch
        kind, desc, rdc, _, _ = c5_scene(w, np.asarray(make_density(w, "cuda:0" if torch.cuda.is_available() else None)), spp)
        v, threads, sample = c5_cpu_arm(kind, desc, rdc, 25.0 * max(1, args.steps + args.warmup) / 11)
        out = dict(base)
        out.update({"impl": "reference", "value": v, "ms_per_step": None, "n_gpus": args.gpus,
                    "cpu_baseline": {"value": v, "unit": "Msamples/s", "cores": threads, "kind": "port", "sample": sample},
                    "e2e": {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                    "gpu_launches": 0})
        print(json.dumps(out))
        return 0
    if args.impl == "reference":
        if rank != 0:
            return 0
        import torch  # only for the density generator when a GPU is around
        dev = "cuda:0" if torch.cuda.is_available() else None
        density = make_density(w, dev) if dev else scenes.fbm_density(*w["grid"])
        value, ms, threads, sample_desc, kind = run_reference_arm(args, w, density, spp)
        out = dict(base)
        out.update({"impl": "reference", "value": value, "ms_per_step": ms, "n_gpus": args.gpus,
                    "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": threads, "kind": kind,
                                     "sample": sample_desc},
                    "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                    "gpu_launches": 0})
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
    ctx = Context(local_rank)
    c5_fetch_bytes = None
    if c5:
        c5_kind, c5_desc, rd, args.layout, c5_fetch_bytes = c5_scene(w, np.asarray(density), spp)
        t0 = time.perf_counter()
        med = ctx.create_medium(c5_kind, c5_desc) if c5_kind == "rgbgrid" else ctx.create_medium(c5_kind, c5_desc, layout=args.layout)
    else:
        nd, rd = make_scene(w, density, spp, 0)
        t0 = time.perf_counter()
        med = ctx.create_medium("nanovdb", nd, layout=args.layout)
    build_s = time.perf_counter() - t0
    comm = sharding.init_comm(ctx) if world > 1 else None
    W, H = w["width"], w["height"]
    film = torch.zeros((H, W, 4), dtype=torch.float64, device=dev)
    # weak scaling by sample index: every rank renders its own `spp` samples of every pixel
    rdr = sharding.shard_samples(rd, rank, world, spp)
    # a stream of our own: torch's default stream has handle 0, which the C ABI reads as "the context's stream"
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def device_step(step, exact, ev=None):
        # everything on ONE stream (torch's current one): film clear -> render kernel -> ncclReduce
        film.zero_()
        r = rdr.with_(sampler_seed=step, exact_math=exact)
        if ev:
            ev[0].record(stream)
        st = ctx.render_volpath_device(med, r, film.data_ptr(), stream=stream.cuda_stream, want_stats=True)
        if ev:
            ev[1].record(stream)
        sharding.reduce_film(film, dst=0)
        return st

    def timed(exact, steps, warmup):
        for i in range(warmup):
            device_step(i, exact)
        sync_all()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        acc = {"camera_fetches": 0, "shadow_fetches": 0, "segments": 0, "camera_rays": 0}
        ev0.record(stream)
        for i in range(steps):
            st = device_step(warmup + i, exact, kev[i])
            for k in acc:
                acc[k] += st[k]
        ev1.record(stream)
        sync_all()
        # device time of the whole loop and of the render kernel alone (events on the launching stream); max over ranks
        tt = torch.tensor([ev0.elapsed_time(ev1) / 1e3, sum(a.elapsed_time(b) for a, b in kev) / 1e3],
                          dtype=torch.float64, device=dev)
        cnt = torch.tensor([acc["camera_rays"], acc["camera_fetches"], acc["shadow_fetches"], acc["segments"]],
                           dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        return float(tt[0]), float(tt[1]), [float(x) for x in cnt]

    def e2e(exact, steps):
        """The call a host makes: avr_render_volpath (N = 1) or avr_render_volpath_multi (N > 1, ncclReduce inside the
        library), HOST film on rank 0; every step ADDS its samples to that film (progressive render)."""
        host_film = np.zeros((H, W, 4), dtype=np.float64) if rank == 0 else None
        full = rd.with_(spp=spp * world, exact_math=exact)

        def one(seed):
            if comm is not None:
                comm.render_volpath(med, full.with_(sampler_seed=seed), film=host_film)
            else:
                ctx.render_volpath(med, full.with_(sampler_seed=seed), film=host_film)
        one(99)
        sync_all()
        t = time.perf_counter()
        for i in range(steps):
            one(100 + i)
        if world > 1:
            dist.barrier()
        dt = time.perf_counter() - t
        te = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        if rank == 0:
            assert float(host_film[..., 3].min()) == spp * world * (steps + 1), "e2e film is incomplete"
        return samples_per_step * steps / float(te[0]) / 1e6

    samples_per_step = W * H * spp * world
    clocks = ClockSampler(local_rank)
    clocks.start()
    loop_s, kern_s, (rays, cam_f, sh_f, segs) = timed(0, args.steps, args.warmup)
    clk = clocks.stop()
    value = samples_per_step * args.steps / loop_s / 1e6
    e2e_value = e2e(0, max(1, args.steps))
    exact_out = None
    if not args.no_exact:
        xs = max(2, args.steps // 4)
        xl, xk, (_, xcam, xsh, xseg) = timed(1, xs, 1)
        exact_out = {"value": samples_per_step * xs / xl / 1e6, "ms_per_step": xl / xs * 1e3, "steps": xs,
                     "e2e": e2e(1, max(1, xs // 2)), "fetches_per_step": (xcam + xsh) / xs, "segments_per_step": xseg / xs,
                     "kernel_ms_per_step": xk / xs * 1e3}

    # parity renders (rank 0's GPU, full frame, >= parity_spp samples per pixel in both modes): k independent steps,
    # film Y of every step kept at full resolution until the reference's lattice is known
    gpu_y = {}
    if rank == 0 and not args.no_cpu_baseline and args.parity_spp > 0 and not c5:
        k = max(2, (args.parity_spp + spp - 1) // spp)
        for name, exact in (("fast", 0), ("exact", 1)):
            ys_ = []
            for i in range(k):
                film.zero_()
                ctx.render_volpath_device(med, rd.with_(sampler_seed=5000 + i, exact_math=exact), film.data_ptr(),
                                          stream=stream.cuda_stream, want_stats=False)
                ys_.append((film[..., 1] / film[..., 3].clamp(min=1)).float().cpu().numpy())
            gpu_y[name] = ys_

    if rank == 0:
        peaks, peak_kind = measured_peaks()
        # algorithmic bytes of one trilinear fetch: the 32-byte octet, the 16-byte fp16 octet, or (sparse brick pool)
        # two 8-byte tap rows + the 4-byte indirection entry
        fetch_bytes = {"octet16": 16, "brick16": 20}.get(args.layout, 32)
        if c5_fetch_bytes:
            fetch_bytes = c5_fetch_bytes  # emissive: density + temperature + Lescale taps; rgbgrid: 2 grids x 8 x 16 B
        # the L2 half of the metric ("density-fetch GB/s vs HBM/L2 roofline"): measured here, with the 64 MiB-resident
        # read kernel BASELINE.md promises; it is the denominator when the medium's store fits the L2
        l2_peak = ctx.measure_read_bandwidth(64 << 20, iters=40)
        hbm_read = ctx.measure_read_bandwidth(4 << 30, iters=2)
        L2_BYTES = 126 * 1000 * 1000
        fits_l2 = med.bytes() < L2_BYTES

        def roof(cam, sh, seg, ksec, steps, exact):
            alg = (cam + sh) * fetch_bytes + seg * 4
            ach = alg / max(ksec, 1e-9) / 1e9 / max(world, 1)  # per-GPU GB/s of the dominant kernel
            traffic, src = measured_traffic(args.workload, args.layout, exact, spp)
            peak = l2_peak if fits_l2 else peaks["hbm_gbs"]
            return {"bound": "l2" if fits_l2 else "hbm", "kernel": "k_volpath", "achieved": ach, "peak": peak,
                    "peak_kind": "measured here (64 MiB-resident read kernel)" if fits_l2 else peak_kind,
                    "l2_peak": l2_peak, "hbm_read_peak_here": hbm_read, "hbm_peak": peaks["hbm_gbs"],
                    "unit": "GB/s", "frac": ach / peak, "traffic": traffic,
                    "traffic_unit": "bytes per launch (ncu dram read+write)", "traffic_capture": src,
                    "algorithmic_bytes": alg / steps / max(world, 1), "fetches_per_step": (cam + sh) / steps,
                    "segments_per_step": seg / steps, "bytes_per_fetch": fetch_bytes,
                    "kernel_ms_per_step": ksec / steps * 1e3}
        out = dict(base)
        out.update({"impl": "avr_b200", "value": value, "ms_per_step": loop_s / args.steps * 1e3,
                    "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": 4 * 471 * 4,
                            "d2h_bytes_per_step": W * H * 4 * 8,
                            "call": "avr_render_volpath_multi (ncclReduce of the film inside the library, one D2H on rank 0)"
                                    if world > 1 else "avr_render_volpath"},
                    "gpu_launches": args.steps,
                    "roofline": roof(cam_f, sh_f, segs, kern_s, args.steps, 0),
                    "clocks": clk, "medium_build_s": build_s, "medium_bytes": med.bytes(),
                    "kernel_config": {"value_is": "exact_math=0 (FAST: the mode the adapter ships; per-ray parity is pinned "
                                                  "for exact_math=1, reported under `exact`)",
                                      "layout": args.layout, "exact_math": 0,
                                      "parallelism": f"sample-index sharded x{world} ({spp} spp per rank per step), film ncclReduce"
                                                     if world > 1 else "1 gpu"}})
        if exact_out:
            exact_out["roofline_frac"] = roof(xcam, xsh, xseg, xk, xs, 1)["frac"]
            out["exact"] = exact_out
        if not args.no_cpu_baseline and c5:
            try:
                v, threads, sample = c5_cpu_arm(c5_kind, c5_desc, rd, 20.0)
                out["cpu_baseline"] = {"value": v, "unit": "Msamples/s", "cores": threads, "kind": "port", "sample": sample}
            except Exception as e:
                out["cpu_baseline"] = {"value": None, "error": repr(e)}
        elif not args.no_cpu_baseline:
            try:
                ref = ReferenceArm(w, density, spp)
                stride = ref.calibrate(6.0)
                t = time.perf_counter()
                n, k = 0, 0
                while k < 3 or (time.perf_counter() - t < 18.0 and k < 64):   # bounded: ~20 s of CPU work
                    n += ref.render(stride, 100 + k)
                    k += 1
                dt = time.perf_counter() - t
                out["cpu_baseline"] = {"value": n / dt / 1e6, "unit": "Msamples/s", "cores": ref.threads, "kind": ref.kind,
                                       "sample": f"{k} steps of " + ref.describe(stride)}
                if gpu_y:
                    xs_, ys_, bidx, npix = lattice_blocks(W, H, stride)
                    steps_ = {name: np.stack([np.bincount(bidx, weights=y[ys_][:, xs_].ravel().astype(np.float64),
                                                          minlength=npix.size) / np.maximum(npix, 1) for y in lst])
                              for name, lst in gpu_y.items()}
                    out["parity"] = parity_block(ref, steps_, spp)
                ref.close()
            except Exception as e:  # the GPU numbers stand on their own
                out["cpu_baseline"] = {"value": None, "error": repr(e)}
        print(json.dumps(out))
    if comm is not None:
        comm.close()
    med.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
