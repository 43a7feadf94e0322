#!/usr/bin/env python
"""Sweep the scheduling knobs of k_volpath on the bench workload (GPU box only).

Prints one line per configuration: kernel ms for 1 spp at the bench workload, fetches, GB/s.
Knobs are environment variables read by the library at every render call:
AVR_REGEN_LANES, AVR_EVENT_LANES, AVR_BLOCKS_PER_SM.
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
    ap.add_argument("--regen", default="8")
    ap.add_argument("--event", default="8")
    ap.add_argument("--bpsm", default="0")
    ap.add_argument("--reps", type=int, default=2)
    args = ap.parse_args()
    w = bench.WORKLOADS[args.workload]
    dev = torch.device("cuda", 0)
    density = bench.make_density(w, dev)
    ctx = Context(0)
    W, H = w["width"], w["height"]
    film = torch.zeros((H, W, 4), dtype=torch.float64, device=dev)
    for layout in args.layouts.split(","):
        gd, rd = bench.make_scene(w, density, args.spp, 0)
        med = ctx.create_medium("uniformgrid", gd, layout=layout)
        for exact, rg, ev, bp in itertools.product(args.exact.split(","), args.regen.split(","),
                                                   args.event.split(","), args.bpsm.split(",")):
            os.environ["AVR_REGEN_LANES"] = rg
            os.environ["AVR_EVENT_LANES"] = ev
            os.environ["AVR_BLOCKS_PER_SM"] = bp
            r = rd.with_(exact_math=int(exact))
            best = None
            for i in range(args.reps + 1):
                film.zero_()
                st = ctx.render_volpath_device(med, r.with_(sampler_seed=i), film.data_ptr(), want_stats=True)
                if i > 0 and (best is None or st["kernel_ms"] < best["kernel_ms"]):
                    best = st
            f = best["camera_fetches"] + best["shadow_fetches"]
            bpf = 16 if layout == "octet16" else 32
            print(json.dumps({"layout": layout, "exact": int(exact), "regen": int(rg), "event": int(ev), "bpsm": int(bp),
                              "kernel_ms": round(best["kernel_ms"], 2),
                              "Msamples_s": round(W * H * args.spp / best["kernel_ms"] / 1e3, 2),
                              "fetches": f, "GBs": round((f * bpf + best["segments"] * 4) / best["kernel_ms"] / 1e6, 1)}),
                  flush=True)
        med.close()
    ctx.close()


if __name__ == "__main__":
    main()
