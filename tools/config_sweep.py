#!/usr/bin/env python
"""Throughput of every kernel on the path over BASELINE.json's other configurations (GPU box only).

One JSON line per case: C1 cube, C2/C3 Disney-cloud stand-in (720p depth 100,
1080p depth 1000), C5 synthetic cubes 256^3..1024^3 (scalar thin/thick, g 0 / 0.877, RGB grids,
emissive), the SampleT_maj kernel on 2^20-ray batches, and C4 graph propagation on synthetic graphs
of 1e5..1e7 vertices.  Every line carries algorithmic GB/s and its fraction of the measured HBM peak.
Timings: k_volpath by the library's CUDA events around the launch (stats.kernel_ms); the batch
kernels by wall clock between synchronisations over several repetitions.  Nothing here touches the
oracle.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

import bench  # noqa: E402
from acceleratedvolrenderer_b200 import scenes  # noqa: E402
from acceleratedvolrenderer_b200._abi import GridDesc, RGBGridDesc  # noqa: E402
from acceleratedvolrenderer_b200.api import Context  # noqa: E402

LAM = np.array([450.0, 520.0, 610.0, 700.0], dtype=np.float32)


def emit(**kw):
    print(json.dumps(kw), flush=True)


def density_cube(n, dev):
    return bench.fbm_density_torch(n, n, n, dev).cpu().numpy()


def time_render(ctx, med, rd, film, reps):
    best = None
    for i in range(reps + 1):
        film.zero_()
        st = ctx.render_volpath_device(med, rd.with_(sampler_seed=i), film.data_ptr(), want_stats=True)
        if i > 0 and (best is None or st["kernel_ms"] < best["kernel_ms"]):
            best = st
    return best


def volpath_case(ctx, peak, name, med, rd, bytes_per_fetch, reps, **extra):
    dev = torch.device("cuda", 0)
    film = torch.zeros((rd.height, rd.width, 4), dtype=torch.float64, device=dev)
    st = time_render(ctx, med, rd, film, reps)
    f = st["camera_fetches"] + st["shadow_fetches"]
    gbs = (f * bytes_per_fetch + st["segments"] * 4) / st["kernel_ms"] / 1e6
    emit(case=name, kernel="k_volpath", kernel_ms=round(st["kernel_ms"], 3),
         Msamples_s=round(rd.width * rd.height * rd.spp / st["kernel_ms"] / 1e3, 2), fetches=f,
         segments=st["segments"], fetches_per_path=round(f / (rd.width * rd.height * rd.spp), 1),
         bytes_per_fetch=bytes_per_fetch, GBs=round(gbs, 1), hbm_frac=round(gbs / peak, 4),
         medium_MB=round(med.bytes() / 1e6, 1), **extra)
    del film


def run_volpath(ctx, peak, args):
    dev = torch.device("cuda", 0)
    if not args.c5_only:
        run_c1_c3(ctx, peak, args, dev)
    run_c5(ctx, peak, args, dev)


def run_c1_c3(ctx, peak, args, dev):
    # C1: cube.pbrt-like, 32^3 uniform-random density, 640x480
    rng = np.random.default_rng(1)
    dens = rng.random((32, 32, 32), dtype=np.float32)
    for exact in (1, 0):
        gd, rd = scenes.cloud_scene(dens, width=640, height=480, spp=16, sigma_t=8.0, albedo=0.9, g=0.0, max_depth=5,
                                    exact_math=exact, boundary="box")
        med = ctx.create_medium("uniformgrid", gd, layout="octet32")
        volpath_case(ctx, peak, "C1 cube 32^3 640x480 16spp depth5", med, rd, 32, args.reps, exact_math=exact)
        med.close()
    # C2 / C3: Disney-cloud stand-in
    w = dict(bench.WORKLOADS["disney-cloud-standin-720p"])
    density = bench.make_density(w, dev)
    for name, width, height, depth, spp in (("C2 cloud 720p depth100", 1280, 720, 100, 4),
                                            ("C3 cloud 1080p depth1000", 1920, 1080, 1000, 2)):
        w2 = dict(w, width=width, height=height, max_depth=depth)
        for layout, bpf in (("octet32", 32), ("octet16", 16), ("dense", 32)):
            for exact in ((0, 1) if layout == "octet32" else (0,)):
                gd, rd = bench.make_scene(w2, density, spp, exact)
                med = ctx.create_medium("nanovdb", gd, layout=layout)
                volpath_case(ctx, peak, f"{name} {spp}spp", med, rd, bpf, args.reps, layout=layout, exact_math=exact)
                med.close()
    del density


def run_c5(ctx, peak, args, dev):
    # C5: synthetic cubes
    for n in args.cubes:
        dens = density_cube(n, dev)
        for k, tag in ((8.0, "thin"), (64.0, "thick")):
            for g in (0.0, 0.877):
                gd, rd = scenes.cloud_scene(dens, width=1280, height=720, spp=2, sigma_t=1.01 * k, albedo=1 / 1.01, g=g,
                                            max_depth=100, exact_math=0, majorant_res=(16, 16, 16), fov=24.0)
                med = ctx.create_medium("uniformgrid", gd, layout="octet32")
                volpath_case(ctx, peak, f"C5 scalar {n}^3 {tag} g={g} 720p 2spp", med, rd, 32, args.reps, exact_math=0)
                med.close()
        # wavelength-dependent sigma_a / sigma_s on the scalar grid: the 4-channel (chromatic MIS) kernel
        gd, rd = scenes.cloud_scene(dens, width=1280, height=720, spp=2, sigma_t=64.64, albedo=1 / 1.01, g=0.877,
                                    max_depth=100, exact_math=0, fov=24.0, chromatic=True)
        med = ctx.create_medium("uniformgrid", gd, layout="octet32")
        for exact in (1, 0):
            volpath_case(ctx, peak, f"C5 chromatic scalar {n}^3 thick g=0.877 720p 2spp", med, rd.with_(exact_math=exact), 32,
                         args.reps, exact_math=exact)
        med.close()
        if n <= args.general_max:
            # emissive scalar grid: temperature = 300 + 2000 d, Lescale = 1 (general kernel: exact math, dense layout)
            gd, rd = scenes.cloud_scene(dens, width=1280, height=720, spp=1, sigma_t=8.08, albedo=1 / 1.01, g=0.0,
                                        max_depth=100, exact_math=1, fov=24.0)
            ge = GridDesc(dens, gd.sigma_a, gd.sigma_s, render_from_medium=gd.m, g=0.0, temperature=300 + 2000 * dens,
                          lescale=np.ones_like(dens), le_norm=1.0)
            med = ctx.create_medium("uniformgrid", ge, layout="dense")
            for exact in (1, 0):
                volpath_case(ctx, peak, f"C5 emissive {n}^3 thin 720p 1spp", med, rd.with_(exact_math=exact), 96, args.reps,
                             exact_math=exact)
            med.close()
            # RGB grids: sigma_s = d (1, .9, .8) k, sigma_a = d (.02, .05, .1) k as {scale,c0,c1,c2} voxels
            r2 = np.random.default_rng(7)
            ss = scenes.rgb_voxels(dens * 8.0, r2); sa = scenes.rgb_voxels(dens * 0.4, r2)
            gr = RGBGridDesc(sigma_a=sa, sigma_s=ss, render_from_medium=gd.m, g=0.0)
            med = ctx.create_medium("rgbgrid", gr)
            for exact in (1, 0):
                volpath_case(ctx, peak, f"C5 rgbgrid {n}^3 thin 720p 1spp", med, rd.with_(exact_math=exact), 256, args.reps,
                             exact_math=exact)
            med.close()
            del ss, sa
        del dens


def run_tmaj(ctx, peak, args):
    dev = torch.device("cuda", 0)
    n = 1 << 20
    rng = np.random.default_rng(7)
    for cube in (256, 512):
        dens = density_cube(cube, dev)
        gd, _ = scenes.cloud_scene(dens, width=8, height=8, spp=1, sigma_t=64.0, albedo=0.99, g=0.0)
        # origins on the bounding sphere r = 1, directions towards uniform points of the cube (render space)
        o = rng.normal(size=(n, 3)); o /= np.linalg.norm(o, axis=1, keepdims=True)
        tgt = rng.uniform(-0.5, 0.5, (n, 3))
        d = tgt - o
        rays = np.concatenate([o, d, np.full((n, 1), 2.0)], axis=1).astype(np.float32)
        seeds = rng.integers(0, 2 ** 63, (n, 2), dtype=np.uint64)
        u = rng.random(n, dtype=np.float32)
        t_rays = torch.from_numpy(rays).to(dev); t_seeds = torch.from_numpy(seeds.view(np.int64)).to(dev)
        t_u = torch.from_numpy(u).to(dev)
        out = torch.zeros((n, 20), dtype=torch.float32, device=dev)  # avr_tmaj_result = 80 bytes
        for layout, bpf in (("octet32", 32), ("dense", 32)):
            med = ctx.create_medium("uniformgrid", gd, layout=layout)
            for mode, mname in ((0, "ratio"), (1, "delta"), (2, "shadow")):
                for exact in (1, 0):
                    def call():
                        med.sample_t_maj_device(mode, exact, t_rays.data_ptr(), n, t_seeds.data_ptr(), t_u.data_ptr(), LAM,
                                                out.data_ptr())
                    call(); ctx_sync(ctx)
                    t = time.perf_counter()
                    for _ in range(args.reps * 3):
                        call()
                    ctx_sync(ctx)
                    ms = (time.perf_counter() - t) / (args.reps * 3) * 1e3
                    res = out.cpu().numpy().view(np.int32)
                    fetch = int(res[:, 18].sum()); seg = int(res[:, 19].sum())
                    gbs = (fetch * bpf + seg * 4 + n * (28 + 16 + 4 + 80)) / ms / 1e6
                    emit(case=f"SampleT_maj 2^20 rays {cube}^3 thick", kernel=f"k_sample_tmaj[{mname}]", layout=layout,
                         exact_math=exact, ms=round(ms, 3), Mrays_s=round(n / ms / 1e3, 1), fetches=fetch, segments=seg,
                         GBs=round(gbs, 1), hbm_frac=round(gbs / peak, 4))
            med.close()
        del dens


def ctx_sync(ctx):
    torch.cuda.synchronize()  # device-wide: covers the library's own stream


def run_graph(ctx, peak, args):
    dev = torch.device("cuda", 0)
    for n in args.graph_sizes:
        rowptr, col, val = scenes.random_graph(n, avg_deg=6, seed=3, albedo=0.95)
        g = ctx.create_graph(rowptr, col, val)
        nnz = int(col.size)
        L0 = torch.rand(n, dtype=torch.float32, device=dev)
        total = torch.empty_like(L0)
        for bounces in (1, 2, 4, 8):
            g.propagate_device(L0.data_ptr(), bounces, total.data_ptr()); ctx_sync(ctx)
            reps = max(3, args.reps * 3)
            t = time.perf_counter()
            for _ in range(reps):
                g.propagate_device(L0.data_ptr(), bounces, total.data_ptr())
            ctx_sync(ctx)
            ms = (time.perf_counter() - t) / reps * 1e3
            bytes_ = bounces * (12 * nnz + 16 * n)
            gbs = bytes_ / ms / 1e6
            emit(case=f"C4 graph propagate N={n} nnz={nnz}", kernel="k_spmv_* (graph of `bounces` iterations)", bounces=bounces,
                 ms=round(ms, 4), GBs=round(gbs, 1), hbm_frac=round(gbs / peak, 4),
                 Medges_s=round(bounces * nnz / ms / 1e3, 1))
        g.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", default="volpath,tmaj,graph")
    ap.add_argument("--cubes", default="256,512,1024")
    ap.add_argument("--general-max", type=int, default=512, help="largest cube for the RGB / emissive variants")
    ap.add_argument("--graph-sizes", default="100000,1000000,10000000")
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--c5-only", action="store_true", help="volpath: only the synthetic cubes (C5)")
    args = ap.parse_args()
    args.cubes = [int(x) for x in args.cubes.split(",") if x]
    args.graph_sizes = [int(x) for x in args.graph_sizes.split(",") if x]
    peaks, kind = bench.measured_peaks()
    peak = peaks["hbm_gbs"]
    emit(hbm_peak_GBs=peak, peak_kind=kind, device=torch.cuda.get_device_name(0))
    ctx = Context(0)
    for what in args.what.split(","):
        {"volpath": run_volpath, "tmaj": run_tmaj, "graph": run_graph}[what](ctx, peak, args)
    ctx.close()


if __name__ == "__main__":
    main()
