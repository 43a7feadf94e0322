import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from acceleratedvolrenderer_b200 import scenes
from acceleratedvolrenderer_b200.api import Context
from cpu_checker import Checker

def rel_mse(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    return float(np.mean((a - b) ** 2 / (b ** 2 + 1e-3)))

dens = scenes.fbm_density(48, 40, 52, seed=4)
gd, rd = scenes.cloud_scene(dens, width=32, height=24, spp=64, sigma_t=14.0, albedo=0.95, g=0.4, max_depth=6)
orc = Checker("orc")
mo = orc.medium_uniformgrid(gd)
ref_a, _ = mo.render(rd, threads=8)
ref_b, _ = mo.render(rd.with_(sampler_seed=1), threads=8)
ctx = Context(0)
img_a = ctx.film_resolve(rd, ref_a); img_b = ctx.film_resolve(rd, ref_b)
print("noise", rel_mse(img_a, img_b), "mean", img_a.mean())
md = ctx.create_medium("uniformgrid", gd, layout="dense")
for dbg in [0] + [int(x) for x in sys.argv[1:]]:
    os.environ["AVR_DEBUG"] = str(dbg)
    for exact in (0,):
        film, st = ctx.render_volpath(md, rd.with_(exact_math=exact, sampler_seed=2))
        img = ctx.film_resolve(rd, film)
        print("debug", dbg, "exact", exact, "relmse", rel_mse(img, img_a), "mean", img.mean(), {k: st[k] for k in ("camera_fetches", "shadow_fetches", "segments", "scatter_events")})
film, st = ctx.render_volpath(md, rd.with_(exact_math=1, sampler_seed=2))
img = ctx.film_resolve(rd, film)
print("exact relmse", rel_mse(img, img_a), "mean", img.mean(), {k: st[k] for k in ("camera_fetches", "shadow_fetches", "segments", "scatter_events")})
