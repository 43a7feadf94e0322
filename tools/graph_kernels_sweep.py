#!/usr/bin/env python
"""Timings of the graph-side kernels on ONE device-built graph (GPU box only).

A fBm cloud is turned into a free graph by avr_graph_build_free, then every step of cmd/graph_maker.cpp's
pipeline and of GraphIntegrator runs on it: light vector, transport matrix from the edge list, propagation (on
this NON-LOCAL graph: neighbours are close in space, not in index), vertex upload, ConnectToGraph queries, graph
render.  A second SpMV case uses a skewed (power-law) degree distribution.  One JSON line per step; under
`ncu -k regex:<kernel> -c 1` the same script gives the per-kernel captures summarised in
profiles/r2_graph_kernels_ncu_summary.md.  Nothing here touches the oracle.
"""
import argparse
import json
import math
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

import bench  # noqa: E402
from acceleratedvolrenderer_b200 import scenes  # noqa: E402
from acceleratedvolrenderer_b200.api import Context  # noqa: E402

LAM = np.array([450.0, 520.0, 610.0, 700.0], dtype=np.float32)


def emit(**kw):
    print(json.dumps(kw), flush=True)


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    best = None
    for _ in range(reps):
        t = time.perf_counter()
        out = fn()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t) * 1e3
        best = dt if best is None or dt < best else best
    return best, out


def skewed_graph(n, seed=5, albedo=0.95):
    """CSR with Zipf-like out-degrees (a few vertices with thousands of neighbours, most with 1-3) and columns
    anywhere in the graph: the opposite of scenes.random_graph's index-local rows."""
    rng = np.random.default_rng(seed)
    deg = np.minimum(rng.zipf(1.8, n), 4096).astype(np.int64)
    rowptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(deg, out=rowptr[1:])
    nnz = int(rowptr[-1])
    col = rng.integers(0, n, nnz, dtype=np.int64)
    rows = np.repeat(np.arange(n, dtype=np.int64), deg)
    order = np.lexsort((col, rows))
    col = col[order].astype(np.int32)
    w = rng.random(nnz).astype(np.float32) + 0.05
    rowsum = np.bincount(rows, weights=w, minlength=n).astype(np.float32)
    val = (w / np.maximum(rowsum[rows], 1e-20) * albedo).astype(np.float32)
    return rowptr, col, val, int(deg.max())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=96)
    ap.add_argument("--dimension-steps", type=int, default=700)
    ap.add_argument("--iterations-per-step", type=int, default=4)
    ap.add_argument("--max-depth", type=int, default=60)
    ap.add_argument("--radius-modifier", type=float, default=4.0)
    ap.add_argument("--width", type=int, default=640)
    ap.add_argument("--height", type=int, default=360)
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--queries", type=int, default=1 << 20)
    ap.add_argument("--skewed-n", type=int, default=2_000_000)
    args = ap.parse_args()
    peaks, kind = bench.measured_peaks()
    peak = peaks["hbm_gbs"]
    emit(hbm_peak_GBs=peak, peak_kind=kind, device=torch.cuda.get_device_name(0))
    ctx = Context(0)
    n = args.grid
    dens = scenes.fbm_density(n, n, n, seed=4)
    gd, rd = scenes.cloud_scene(dens, width=args.width, height=args.height, spp=args.spp, sigma_t=20.0, albedo=0.95, g=0.0,
                                max_depth=8)
    md = ctx.create_medium("uniformgrid", gd, layout="octet32")
    radius = float(rd.c.boundary[3])
    max_dist = radius * math.sqrt(3.0)                     # util.h:56-57 for the sphere's bounding box
    vertex_radius = max_dist * 2 / 1000 * args.radius_modifier   # util.h:465-467 x radiusModifier
    center = [float(rd.c.boundary[i]) for i in range(3)]
    kw = dict(vertex_radius=vertex_radius, max_dist_to_center=max_dist, sampling_resolution=8192, lam=LAM)

    # n2: the builder
    t = time.perf_counter()
    built = md.build_free_graph(rd, bounds_center=center, dimension_steps=args.dimension_steps,
                                iterations_per_step=args.iterations_per_step, max_depth=args.max_depth,
                                search_range_neighbours=4, **kw)
    ms = (time.perf_counter() - t) * 1e3
    nv, ne = built["pos"].shape[0], built["efrom"].size
    emit(case="n2 avr_graph_build_free", ms=round(ms, 2), paths=int(built["n_paths"]), scatter_points=int(built["n_points"]),
         vertices=nv, edges=ne, cluster_rounds=int(built["cluster_rounds"]),
         Mpoints_s=round(built["n_points"] / ms / 1e3, 2))

    # a18: the light vector
    ms, L0 = timed(lambda: md.graph_light_vector(rd, built["pos"], light_iterations=8, points_on_radius=2, **kw), reps=2)
    rays = nv * 8 * (1 + 2 * 4)  # GetDiskPointsSize(2) is about 1 + 4 per ring; reported as an order of magnitude only
    emit(case="a18 avr_graph_light_vector", ms=round(ms, 2), vertices=nv, light_iterations=8, points_on_radius=2,
         Mvertices_s=round(nv / ms / 1e3, 3), approx_rays=rays)

    # a16: transport matrix from the edge list
    ms, g = timed(lambda: ctx.create_graph_from_edges(built["vsamples"], built["efrom"], built["eto"], built["esamples"]), reps=2)
    emit(case="a16 avr_graph_create_from_edges (incl. H2D of the edge list)", ms=round(ms, 2), edges=ne,
         Medges_s=round(ne / ms / 1e3, 2))

    # a17: propagation on the built (non-local) graph, device-resident vectors
    dev = torch.device("cuda", 0)
    rowptr, col, val = g.csr()
    nnz = int(col.size)
    L0d = torch.from_numpy(L0).to(dev)
    total = torch.empty_like(L0d)
    for bounces in (1, 8):
        ms, _ = timed(lambda: g.propagate_device(L0d.data_ptr(), bounces, total.data_ptr()), reps=5)
        gbs = bounces * (12 * nnz + 16 * nv) / ms / 1e6
        emit(case="a17 propagate, device-built graph (spatially local, index non-local)", vertices=nv, nnz=nnz, bounces=bounces,
             ms=round(ms, 4), GBs=round(gbs, 1), hbm_frac=round(gbs / peak, 4))
    tot, _ = g.compute_final_light(L0, 3)

    # a19: vertex upload, ConnectToGraph, graph render
    ms, _ = timed(lambda: g.set_vertices(built["pos"], tot, built["vrange"], vertex_radius), reps=2)
    emit(case="a19 avr_graph_set_vertices (hash grid build on the host)", ms=round(ms, 2), vertices=nv)
    rng = np.random.default_rng(1)
    q = (np.array(center, dtype=np.float32) + rng.normal(size=(args.queries, 3)).astype(np.float32) * radius * 0.4)
    ms, got = timed(lambda: g.connect(q), reps=2)
    emit(case="a19 avr_graph_connect (ConnectToGraph, host points)", ms=round(ms, 2), queries=args.queries,
         Mqueries_s=round(args.queries / ms / 1e3, 2), hit_fraction=round(float((np.asarray(got) > 0).mean()), 3))
    ms, (film, st) = timed(lambda: ctx.render_graph(md, g, rd), reps=2)
    emit(case="a19 avr_render_graph (host film)", ms=round(ms, 2), film=[args.width, args.height], spp=args.spp,
         Msamples_s=round(args.width * args.height * args.spp / ms / 1e3, 2),
         kernel_ms=round(st.get("kernel_ms", float("nan")), 3) if isinstance(st, dict) else None)
    g.close()

    # a17 on a skewed-degree graph
    rowptr, col, val, dmax = skewed_graph(args.skewed_n)
    g = ctx.create_graph(rowptr, col, val)
    nnz = int(col.size)
    x = torch.rand(args.skewed_n, dtype=torch.float32, device=dev)
    total = torch.empty_like(x)
    ms, _ = timed(lambda: g.propagate_device(x.data_ptr(), 8, total.data_ptr()), reps=5)
    gbs = 8 * (12 * nnz + 16 * args.skewed_n) / ms / 1e6
    emit(case="a17 propagate, skewed (Zipf 1.8) degrees, random columns", vertices=args.skewed_n, nnz=nnz, max_degree=dmax,
         bounces=8, ms=round(ms, 4), GBs=round(gbs, 1), hbm_frac=round(gbs / peak, 4))
    g.close()
    md.close()
    ctx.close()


if __name__ == "__main__":
    main()
