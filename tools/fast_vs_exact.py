#!/usr/bin/env python
"""FAST vs EXACT arithmetic of k_volpath on the same scene at high sample counts: image means and
per-pixel differences in units of the Monte Carlo standard error (a bias check for fast-math changes).
usage: python tools/fast_vs_exact.py [spp]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from acceleratedvolrenderer_b200 import scenes  # noqa: E402
from acceleratedvolrenderer_b200.api import Context  # noqa: E402


def main():
    spp = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    ctx = Context(0)
    dens = scenes.fbm_density(48, 40, 52, seed=4)
    d2 = dens.copy(); d2[d2 < 0.02] = 0
    cases = []
    gd, rd = scenes.cloud_scene(dens, width=32, height=24, spp=spp, sigma_t=14.0, albedo=0.95, g=0.4, max_depth=6)
    cases.append(("uniformgrid", "uniformgrid", gd, rd, "octet32"))
    gdc, rdc = scenes.cloud_scene(dens, width=32, height=24, spp=spp, sigma_t=14.0, albedo=0.95, g=0.4, max_depth=6, chromatic=True)
    cases.append(("chromatic", "uniformgrid", gdc, rdc, "octet32"))
    nd, rdn = scenes.cloud_scene_nanovdb(d2, width=32, height=24, spp=spp, sigma_t=18.0, albedo=0.95, g=0.4, max_depth=6)
    cases.append(("nanovdb", "nanovdb", nd, rdn, "octet32"))
    cases.append(("nanovdb-dense", "nanovdb", nd, rdn, "dense"))
    cases.append(("nanovdb-oct16", "nanovdb", nd, rdn, "octet16"))
    nd2, rdn2 = scenes.cloud_scene_nanovdb(d2, width=32, height=24, spp=spp, sigma_t=200.0, albedo=0.999, g=0.877, max_depth=100)
    cases.append(("nanovdb-thick", "nanovdb", nd2, rdn2, "octet32"))
    from acceleratedvolrenderer_b200._abi import GridDesc
    em = GridDesc(dens, gd.sigma_a, gd.sigma_s, render_from_medium=gd.m, g=0.4, temperature=dens * 2500 + 200,
                  lescale=dens * 0.5, le_norm=1.0)
    cases.append(("grid-emissive", "uniformgrid", em, rd, "dense"))
    nde, rde = scenes.cloud_scene_nanovdb(d2, temperature=(d2 * 2500 + 200).astype(np.float32), le_scale=0.4, width=32,
                                          height=24, spp=spp, sigma_t=18.0, albedo=0.95, g=0.4, max_depth=6)
    cases.append(("nvdb-emissive", "nanovdb", nde, rde, "dense"))
    for name, kind, desc, r, layout in cases:
        md = ctx.create_medium(kind, desc, layout=layout)
        imgs = {}
        for exact in (1, 0):
            halves = []
            for seed in (11, 12):
                film, _ = ctx.render_volpath(md, r.with_(exact_math=exact, sampler_seed=seed))
                halves.append(ctx.film_resolve(r, film).astype(np.float64))
            imgs[exact] = halves
        md.close()
        e = (imgs[1][0] + imgs[1][1]) / 2; f = (imgs[0][0] + imgs[0][1]) / 2
        # standard error of a 2-seed mean from the seed-to-seed difference
        se = np.sqrt(((imgs[1][0] - imgs[1][1]) ** 2 + (imgs[0][0] - imgs[0][1]) ** 2) / 4 + 1e-12)
        z = (f - e) / se
        lit = e > 0.05 * e.max()
        print(f"{name:14s} mean exact {e.mean():.5f} fast {f.mean():.5f} ratio {f.mean() / e.mean():.4f} "
              f"| z over lit pixels: mean {z[lit].mean():+.3f} rms {np.sqrt((z[lit] ** 2).mean()):.3f} (unbiased: ~0, ~1)")


if __name__ == "__main__":
    main()
