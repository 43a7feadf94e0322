#!/usr/bin/env python
"""EXACT renders must be reproducible sample for sample whatever the scheduling did: render the same scene
repeatedly (per-sample dump, bit comparison) and through different sample shards (GPU box only)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from acceleratedvolrenderer_b200 import scenes  # noqa: E402
from acceleratedvolrenderer_b200.api import Context  # noqa: E402


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
    ctx = Context(0)
    dens = scenes.fbm_density(48, 40, 52, seed=4)
    bad = 0
    for (w, h, spp) in ((64, 48, 12), (200, 150, 8)):
        gd, rd = scenes.cloud_scene(dens, width=w, height=h, spp=spp, sigma_t=14.0, albedo=0.95, g=0.4, max_depth=8)
        med = ctx.create_medium("uniformgrid", gd, layout="octet32")
        for exact in (1, 0):
            ref, _ = ctx.render_volpath_samples(med, rd.with_(exact_math=exact))
            film0, _ = ctx.render_volpath(med, rd.with_(exact_math=exact))
            for i in range(reps):
                got, _ = ctx.render_volpath_samples(med, rd.with_(exact_math=exact))
                d = (got.view(np.uint32) != ref.view(np.uint32)).any(axis=1)
                film, _ = ctx.render_volpath(med, rd.with_(exact_math=exact))
                fd = ~np.isclose(film, film0, rtol=1e-12, atol=1e-300)
                # sample shards into one film
                acc = np.zeros_like(film0)
                k = 3 if spp % 3 == 0 else 2
                for s in range(0, spp, spp // k):
                    ctx.render_volpath(med, rd.with_(exact_math=exact, sample_begin=s, sample_end=s + spp // k), film=acc)
                sd = ~np.isclose(acc, film0, rtol=1e-12, atol=1e-300)
                if d.any() or fd.any() or sd.any():
                    bad += 1
                    idx = np.nonzero(d)[0][:4]
                    print(f"{w}x{h}x{spp} exact={exact} rep {i}: {int(d.sum())} samples differ {idx.tolist()}, "
                          f"{int(fd.any(axis=2).sum())} film pixels, {int(sd.any(axis=2).sum())} sharded-film pixels", flush=True)
                    for j in idx[:2]:
                        print("   ", ref[j, :4], got[j, :4])
        med.close()
    print("DETERMINISM", "OK" if bad == 0 else f"FAILED ({bad} runs)")
    ctx.close()


if __name__ == "__main__":
    main()
