#!/usr/bin/env python
"""One-at-a-time sweep of k_volpath's scheduling knobs around their defaults on the bench workload (GPU box only).

The knobs order the work inside the path pool and never change a result; the library reads them when a context is
created, so every setting gets its own context and medium.  One JSON line per (mode, knob, value): kernel ms of the
best of `--reps` launches at `--spp` samples per pixel.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

import bench  # noqa: E402
from acceleratedvolrenderer_b200.api import Context  # noqa: E402

SWEEP = {
    "AVR_WALK_STEPS": [2, 3, 4, 5, 6],
    "AVR_REFILL_LANES": [8, 12, 20, 24],
    "AVR_TRACK_MIN": [8, 16, 24],
    "AVR_YIELD_LANES": [0, 4, 12, 16],
    "AVR_YIELD_LANES_FULL": [12, 16, 24, 28],
    "AVR_EVENT_SMSP_MASK": [0, 8, 14, 15],
    "AVR_COLLIDE_MIN": [4, 8],
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spp", type=int, default=8)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--exact", default="0,1")
    args = ap.parse_args()
    w = bench.WORKLOADS["disney-cloud-standin-720p"]
    dev = torch.device("cuda", 0)
    density = bench.make_density(w, dev)
    W, H = w["width"], w["height"]
    film = torch.zeros((H, W, 4), dtype=torch.float64, device=dev)
    gd, rd = bench.make_scene(w, density, args.spp, 0)

    def run(exact, name=None, value=None):
        for k in SWEEP:
            os.environ.pop(k, None)
        if name:
            os.environ[name] = str(value)
        ctx = Context(0)
        med = ctx.create_medium("nanovdb", gd, layout="octet32")
        best = None
        for i in range(args.reps + 1):
            film.zero_()
            torch.cuda.synchronize()
            st = ctx.render_volpath_device(med, rd.with_(exact_math=exact, sampler_seed=i), film.data_ptr(), want_stats=True)
            if i > 0 and (best is None or st["kernel_ms"] < best):
                best = st["kernel_ms"]
        med.close(); ctx.close()
        print(json.dumps({"exact": exact, "knob": name or "defaults", "value": value, "kernel_ms": round(best, 3),
                          "Msamples_s": round(W * H * args.spp / best / 1e3, 1)}), flush=True)

    for exact in [int(e) for e in args.exact.split(",")]:
        run(exact)
        for name, values in SWEEP.items():
            for v in values:
                run(exact, name, v)
        run(exact)


if __name__ == "__main__":
    main()
