#!/usr/bin/env python
"""Sweep the scheduling knobs of k_volpath on the bench workload (GPU box only).

Prints one line per configuration: kernel ms for 1 spp at the bench workload, fetches, GB/s.
Knobs are environment variables read by the library at every render call:
AVR_OP_WEIGHTS (R,S,A,B,E1,E2), AVR_WALK_STEPS, AVR_COLLIDE_REPEAT, AVR_BLOCKS_PER_SM.
"""
import argparse
import itertools
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

import bench  # noqa: E402
from acceleratedvolrenderer_b200.api import Context  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="disney-cloud-standin-720p")
    ap.add_argument("--spp", type=int, default=1)
    ap.add_argument("--layouts", default="octet32")
    ap.add_argument("--exact", default="0")
    ap.add_argument("--env", action="append", default=[],
                    help="NAME=v1;v2;... (repeatable); the product of all lists is swept")
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--sampler", default="independent", help="independent | zsobol (fastowen)")
    ap.add_argument("--filter", default="box", help="box | gaussian (pbrt's default pixel filter: radius 1.5, sigma 0.5) | triangle")
    args = ap.parse_args()
    w = bench.WORKLOADS[args.workload]
    dev = torch.device("cuda", 0)
    density = bench.make_density(w, dev)
    W, H = w["width"], w["height"]
    film = torch.zeros((H, W, 4), dtype=torch.float64, device=dev)
    for layout in args.layouts.split(","):
        gd, rd = bench.make_scene(w, density, args.spp, 0)
        names = [e.split("=", 1)[0] for e in args.env]
        lists = [e.split("=", 1)[1].split(";") for e in args.env]
        for exact, *vals in itertools.product(args.exact.split(","), *lists):
            for n, v in zip(names, vals):
                os.environ[n] = v
            # the library reads its knobs when a context is created: one context (and medium) per configuration
            ctx = Context(0)
            med = ctx.create_medium("nanovdb", gd, layout=layout)
            r = rd.with_(exact_math=int(exact), sampler=1 if args.sampler == "zsobol" else 0, randomize=2)
            if args.filter == "gaussian":
                from acceleratedvolrenderer_b200._abi import gaussian_filter_table
                r = r.with_filter(gaussian_filter_table((1.5, 1.5), 0.5), (1.5, 1.5))
            elif args.filter == "triangle":
                r = r.with_filter("triangle", (2.0, 2.0))
            best = None
            for i in range(args.reps + 1):
                film.zero_()
                st = ctx.render_volpath_device(med, r.with_(sampler_seed=i), film.data_ptr(), want_stats=True)
                if i > 0 and (best is None or st["kernel_ms"] < best["kernel_ms"]):
                    best = st
            f = best["camera_fetches"] + best["shadow_fetches"]
            bpf = {"octet16": 16, "brick16": 20}.get(layout, 32)
            print(json.dumps({"layout": layout, "exact": int(exact), "sampler": args.sampler, "filter": args.filter, "env": dict(zip(names, vals)),
                              "kernel_ms": round(best["kernel_ms"], 2),
                              "Msamples_s": round(W * H * args.spp / best["kernel_ms"] / 1e3, 2),
                              "fetches": f, "GBs": round((f * bpf + best["segments"] * 4) / best["kernel_ms"] / 1e6, 1),
                              "medium_MB": round(med.bytes() / 1e6, 1),
                              "stats": {k: v for k, v in best.items() if k != "kernel_ms"}}),
                  flush=True)
            med.close()
            ctx.close()


if __name__ == "__main__":
    main()
