"""Seeded input cases shared by tests/golden/make_golden.py (which runs them through the reference's own
compiled templates, oracle/_ref) and tests/test_oracle_golden.py (which runs them through the CPU
restatement, oracle/liboracle.so).  Every case returns {name: ndarray}; the bar is bit-exactness.

Case inputs are small on purpose: the fixtures travel with the repository.
"""
import os

import numpy as np

from acceleratedvolrenderer_b200 import scenes
from acceleratedvolrenderer_b200._abi import GridDesc, NanoVDBDesc, RGBGridDesc

LAMBDA = np.array([412.3, 533.9, 601.2, 688.8], dtype=np.float32)


def case_scalar(ck):
    """util/math.h FastExp, util/hash.h, util/rng.h, samplers.h IndependentSampler, util/sampling.h,
    util/scattering.h HG, util/spectrum.h wavelengths/blackbody, util/containers.h SampledGrid::Lookup."""
    rng = np.random.default_rng(1)
    out = {}
    x = np.concatenate([rng.uniform(-100, 5, 2000), [-1e30, 0, -0.0, 88, 100, -87.5, -126 * 0.6931]]).astype(np.float32)
    out["fast_exp"] = ck.fast_exp(x)
    out["mix_bits"] = ck.mix_bits(rng.integers(0, 2 ** 63, 256, dtype=np.uint64))
    out["hash_float"] = ck.hash_float(rng.random(256, dtype=np.float32))
    out["hash_point3"] = ck.hash_point3(rng.normal(size=(255, 3)).astype(np.float32))
    for i, args in enumerate([(5, 7, 1, 0), (123456789, 0, 0, 1000), (2 ** 40 + 3, 99, 1, 65536 * 3 + 5)]):
        u32, f = ck.rng_stream(*args, 64)
        out[f"rng_u32_{i}"] = u32
        out[f"rng_f_{i}"] = f
    out["sampler_a"] = ck.independent_sampler(3, 9, 0, 5, 0, 32)
    out["sampler_b"] = ck.independent_sampler(300, 200, 7, 127, 3, 32)
    w = rng.random((1000, 3), dtype=np.float32)
    w[:50, 2] = 0
    u = rng.random(1000, dtype=np.float32)
    u[:5] = [0, 0.99999994, 0.5, 1e-9, 0.999]
    out["sample_discrete3"] = ck.sample_discrete3(w, u)
    out["sample_exponential"] = ck.sample_exponential(u, rng.uniform(0.01, 100, 1000).astype(np.float32))
    wo = rng.normal(size=(1000, 3))
    wo /= np.linalg.norm(wo, axis=1, keepdims=True)
    u2 = rng.random((1000, 2), dtype=np.float32)
    for g in [0.0, 0.877, -0.5, 0.995, 5e-4]:
        wi, pdf = ck.hg_sample(wo, g, u2)
        out[f"hg_wi_{g}"] = wi
        out[f"hg_pdf_{g}"] = pdf
        out[f"hg_p_{g}"] = ck.hg_p(rng.uniform(-1, 1, 200), g)
    out["wavelengths"] = ck.sample_wavelengths(u)
    out["blackbody"] = ck.blackbody(rng.uniform(50, 6000, 200), [400., 500., 600., 700.])
    grid = rng.random((7, 9, 11), dtype=np.float32)
    p = rng.uniform(-0.1, 1.1, (3000, 3)).astype(np.float32)
    p[:10] = [[0, 0, 0], [1, 1, 1], [0.5, 0.5, 0.5], [1, 0, 0], [0, 1, 0], [0, 0, 1], [.9999999, .5, .5],
              [1e-8, .5, .5], [.5 / 11, .5 / 9, .5 / 7], [1 - .5 / 11, .5, .5]]
    out["grid_lookup"] = ck.grid_lookup(grid, p)
    return out


def _dda_rays(rng, bounds, n):
    b = np.asarray(bounds, dtype=np.float32)
    c = (b[:3] + b[3:]) / 2
    e = (b[3:] - b[:3])
    o_ = c + rng.normal(size=(n, 3)) * e * 1.5
    tgt = b[:3] + rng.random((n, 3)) * e
    d = tgt - o_
    d[:n // 10, 0] = 0
    d[n // 10:n // 5, 1] = -0.0
    d[n // 5:n // 4, 2] = 0
    rays = np.zeros((n, 8), dtype=np.float32)
    rays[:, :3] = o_
    rays[:, 3:6] = d
    rays[:, 6] = rng.uniform(0, 0.8, n)
    rays[:, 7] = rays[:, 6] + rng.uniform(0, 1.5, n)
    return rays


def case_dda(ck):
    """DDAMajorantIterator, media.h:136-214: per-ray segment lists (t range, sigma_maj, voxel)."""
    rng = np.random.default_rng(2)
    out = {}
    for k, (bounds, res, n, cap) in enumerate([((0, 0, 0, 1, 1, 1), (16, 16, 16), 600, 64),
                                               ((-1, -2, 0.5, 3, 1, 2.5), (64, 64, 64), 200, 200),
                                               ((0, 0, 0, 1, 1, 1), (5, 7, 3), 600, 32)]):
        vox = rng.random(res[0] * res[1] * res[2], dtype=np.float32)
        vox[rng.random(vox.size) < 0.3] = 0
        rays = _dda_rays(rng, bounds, n)
        cnt, t, sig, vx = ck.dda(bounds, res, vox, rays, sigma_t=1.7, max_per_ray=cap)
        out[f"dda{k}_count"] = cnt
        # keep the fixture small: only the valid prefix of every ray
        m = np.arange(cap)[None, :] < np.minimum(cnt, cap)[:, None]
        out[f"dda{k}_t"] = t.reshape(n, cap, 2)[m]
        out[f"dda{k}_sigma"] = sig.reshape(n, cap)[m]
        out[f"dda{k}_voxel"] = vx.reshape(n, cap)[m]
    return out


def _rot_scale():
    M = scenes.scale_translate((2, 0.5, 1.5), (0.3, -1, 2)).astype(np.float64)
    th = 0.7
    R = np.eye(4)
    R[0, 0] = np.cos(th); R[0, 2] = np.sin(th); R[2, 0] = -np.sin(th); R[2, 2] = np.cos(th)
    M2 = (M @ R).astype(np.float32)
    M2i = np.linalg.inv(M2.astype(np.float64)).astype(np.float32)
    return M2, M2i


def case_transform(ck):
    """Transform::operator()(Ray) / ApplyInverse(Ray), util/transform.h:391-433 (interval arithmetic)."""
    rng = np.random.default_rng(3)
    M2, M2i = _rot_scale()
    rays7 = np.zeros((1500, 7), dtype=np.float32)
    rays7[:, :3] = rng.normal(size=(1500, 3)) * 3
    rays7[:, 3:6] = rng.normal(size=(1500, 3))
    rays7[:, 6] = rng.uniform(0, 10, 1500)
    rays7[:10, 3:6] = 0
    eye = np.eye(4, dtype=np.float32)
    return {"apply_inverse_ray": ck.apply_inverse_ray(M2, M2i, rays7), "apply_ray": ck.apply_ray(M2, M2i, rays7),
            "apply_inverse_ray_identity": ck.apply_inverse_ray(eye, eye, rays7)}


def _tmaj_fields(prefix, res, out, approx_t_event=False):
    for f in res.dtype.names:
        if f == "n_seg":
            continue  # the probe counts segments differently (it cannot see inside the iterator)
        if f == "t_event" and approx_t_event:
            continue
        out[f"{prefix}_{f}"] = np.ascontiguousarray(res[f])


def case_uniformgrid(ck):
    """GridMedium ctor / SamplePoint / SampleRay (media.h:265-337, media.cpp:221-331) and SampleT_maj
    (media.h:741-806) under the three callbacks of the hot path."""
    rng = np.random.default_rng(4)
    out = {}
    dens = scenes.fbm_density(24, 20, 28, seed=3)
    M2, M2i = _rot_scale()
    variants = [dict(), dict(render_from_medium=M2, medium_from_render=M2i, p0=(-1, 0, 0.5), p1=(1, 2, 1.5), scale=3.0),
                dict(temperature=dens * 3000 + 50, lescale=dens * 2, le_norm=0.5, temperature_offset=10, temperature_scale=1.1),
                dict(le=scenes.ramp_spectrum(0.2, 1.0), le_norm=0.01)]
    for k, kw in enumerate(variants):
        gd = GridDesc(dens, scenes.ramp_spectrum(0.5, 1.5), scenes.ramp_spectrum(6, 9), g=0.3, **kw)
        m = ck.medium_uniformgrid(gd)
        out[f"ug{k}_majorants"] = m.majorants()
        out[f"ug{k}_emissive"] = np.array([m.is_emissive()], dtype=np.int32)
        b0, b1 = gd.p0, gd.p1
        npts = 1200
        pm = (b0 + rng.uniform(-0.05, 1.05, (npts, 3)) * (b1 - b0)).astype(np.float32)
        pr = (np.c_[pm, np.ones(npts)] @ gd.m.reshape(4, 4).T.astype(np.float64))[:, :3].astype(np.float32)
        out[f"ug{k}_sample_point"] = m.sample_point(pr, LAMBDA)
        n = 1500
        org = rng.normal(size=(n, 3))
        org /= np.linalg.norm(org, axis=1, keepdims=True)
        org *= 3
        ctr = (np.r_[(b0 + b1) / 2, 1] @ gd.m.reshape(4, 4).T)[:3]
        tgt = pr[rng.integers(0, npts, n)]
        rays7 = np.zeros((n, 7), dtype=np.float32)
        rays7[:, :3] = org + ctr
        rays7[:, 3:6] = (tgt - rays7[:, :3]) * rng.uniform(0.5, 2, (n, 1))
        rays7[:, 6] = rng.uniform(0.2, 3, n)
        out[f"ug{k}_ray_to_medium"] = m.ray_to_medium(rays7)
        seeds = rng.integers(0, 2 ** 63, (n, 2), dtype=np.uint64)
        u = rng.random(n, dtype=np.float32)
        for mode in (0, 1, 2):
            _tmaj_fields(f"ug{k}_tmaj{mode}", m.sample_tmaj(mode, rays7, seeds, u, LAMBDA), out, approx_t_event=(mode == 1))
        m.close()
    return out


def case_rgbgrid(ck):
    """RGBGridMedium (media.h:341-427, media.cpp:333-443)."""
    rng = np.random.default_rng(5)
    out = {}
    dens = scenes.fbm_density(20, 18, 22, seed=3)
    sa = scenes.rgb_voxels(dens * 0.5, rng)
    ss = scenes.rgb_voxels(dens * 6, rng)
    le = scenes.rgb_voxels(dens * 2, rng)
    for k, kw in enumerate([dict(sigma_a=sa, sigma_s=ss), dict(sigma_s=ss, scale=2.0),
                            dict(sigma_a=sa, sigma_s=ss, le=le, le_illuminant=scenes.ramp_spectrum(0.5, 1.5), le_scale=0.7, g=0.5)]):
        gd = RGBGridDesc(**kw)
        m = ck.medium_rgbgrid(gd)
        out[f"rgb{k}_majorants"] = m.majorants()
        out[f"rgb{k}_emissive"] = np.array([m.is_emissive()], dtype=np.int32)
        pr = rng.uniform(-0.05, 1.05, (1000, 3)).astype(np.float32)
        out[f"rgb{k}_sample_point"] = m.sample_point(pr, LAMBDA)
        n = 800
        org = rng.normal(size=(n, 3))
        org /= np.linalg.norm(org, axis=1, keepdims=True)
        org = org * 3 + 0.5
        tgt = pr[rng.integers(0, 1000, n)]
        rays7 = np.zeros((n, 7), dtype=np.float32)
        rays7[:, :3] = org
        rays7[:, 3:6] = (tgt - org) * rng.uniform(0.5, 2, (n, 1))
        rays7[:, 6] = rng.uniform(0.2, 3, n)
        seeds = rng.integers(0, 2 ** 63, (n, 2), dtype=np.uint64)
        u = rng.random(n, dtype=np.float32)
        for mode in (0, 1, 2):
            _tmaj_fields(f"rgb{k}_tmaj{mode}", m.sample_tmaj(mode, rays7, seeds, u, LAMBDA), out, approx_t_event=True)
        m.close()
    return out


def nanovdb_variants(ck):
    """(name, NanoVDBDesc) pairs shared with the GPU parity tests."""
    dens = scenes.fbm_density(40, 33, 37, seed=5)
    dens[dens < 0.05] = 0          # sparse: inactive voxels = background
    dens[:2] = 0; dens[:, :, -3:] = 0   # empty border slabs: index box != build box
    temp = (scenes.fbm_density(40, 33, 37, seed=9) * 2500 + 80).astype(np.float32)
    temp[:, :5] = 0
    M2, M2i = _rot_scale()
    g0 = ck.nanovdb_grid(dens, (-7, 3, 11), 0.07, (0.37, -1.21, 0.05))  # voxel size not a power of two
    g1 = ck.nanovdb_grid(dens, (100, -200, -30), 0.03125, (-4.0, 6.5, 1.25))
    t1 = ck.nanovdb_grid(temp, (101, -198, -33), 0.03125, (-4.0, 6.5, 1.25))
    return [("nv0", NanoVDBDesc(g0, scenes.ramp_spectrum(0.5, 1.5), scenes.ramp_spectrum(6, 9), g=0.3)),
            ("nv1", NanoVDBDesc(g1, scenes.ramp_spectrum(0.2, 0.4), scenes.ramp_spectrum(3, 5), temperature=t1,
                                render_from_medium=M2, medium_from_render=M2i, scale=2.5, g=0.877, le_scale=0.6,
                                temperature_offset=20, temperature_scale=1.05))]


def case_nanovdb(ck):
    """NanoVDBMedium (media.h:602-685, media.cpp:511-616) over real nanovdb::FloatGrids: the grid
    read-back the host adapter does, the 64^3 majorant grid, bounds, SamplePoint (SampleFromVoxels
    trilinear + blackbody emission), SampleRay prologue and SampleT_maj under the three callbacks."""
    rng = np.random.default_rng(8)
    out = {}
    for name, nd in nanovdb_variants(ck):
        for gname, gr in (("density", nd.density), ("temperature", nd.temperature)):
            if gr is None:
                continue
            out[f"{name}_{gname}_values"] = gr.values
            out[f"{name}_{gname}_index_min"] = gr.index_min
            out[f"{name}_{gname}_inv_mat"] = gr.inv_mat
            out[f"{name}_{gname}_vec"] = gr.vec
            out[f"{name}_{gname}_world_bbox"] = gr.world_bbox
        m = ck.medium_nanovdb(nd)
        out[f"{name}_bounds"] = m.bounds()
        out[f"{name}_majorants"] = m.majorants()
        out[f"{name}_emissive"] = np.array([m.is_emissive()], dtype=np.int32)
        b0, b1 = nd.p0, nd.p1
        npts = 1500
        pm = (b0 + rng.uniform(-0.05, 1.05, (npts, 3)) * (b1 - b0)).astype(np.float32)
        pr = (np.c_[pm, np.ones(npts)] @ nd.m.reshape(4, 4).T.astype(np.float64))[:, :3].astype(np.float32)
        out[f"{name}_sample_point"] = m.sample_point(pr, LAMBDA)
        n = 1200
        org = rng.normal(size=(n, 3))
        org /= np.linalg.norm(org, axis=1, keepdims=True)
        org *= 2.5 * float(np.linalg.norm(b1 - b0))
        ctr = (np.r_[(b0 + b1) / 2, 1] @ nd.m.reshape(4, 4).T)[:3]
        tgt = pr[rng.integers(0, npts, n)]
        rays7 = np.zeros((n, 7), dtype=np.float32)
        rays7[:, :3] = org + ctr
        rays7[:, 3:6] = (tgt - rays7[:, :3]) * rng.uniform(0.5, 2, (n, 1))
        rays7[:, 6] = rng.uniform(0.2, 3, n)
        out[f"{name}_ray_to_medium"] = m.ray_to_medium(rays7)
        seeds = rng.integers(0, 2 ** 63, (n, 2), dtype=np.uint64)
        u = rng.random(n, dtype=np.float32)
        for mode in (0, 1, 2):
            _tmaj_fields(f"{name}_tmaj{mode}", m.sample_tmaj(mode, rays7, seeds, u, LAMBDA), out, approx_t_event=True)
        m.close()
    return out


def case_graph(ck):
    """src/graph: Eigen SpMV propagation (free_graph.cpp:345-375) and the light-ray transmittance
    estimator (graph/util.h SampleTransmittance)."""
    rng = np.random.default_rng(6)
    out = {}
    for n in (50, 2000):
        rowptr, col, val = scenes.random_graph(n, avg_deg=4, seed=n)
        L0 = rng.random(n, dtype=np.float32)
        for b in (0, 1, 3, 7):
            total, iters = ck.graph_propagate(rowptr, col, val, L0, b)
            out[f"graph{n}_b{b}_total"] = total
            out[f"graph{n}_b{b}_iters"] = np.array([iters], dtype=np.int32)
        total, iters = ck.graph_propagate(rowptr, col, val * 4e7, L0, 20)  # overflows: early-out path
        out[f"graph{n}_overflow_total"] = total
        out[f"graph{n}_overflow_iters"] = np.array([iters], dtype=np.int32)
    dens = scenes.fbm_density(24, 20, 28, seed=3)
    gd = GridDesc(dens, scenes.const_spectrum(0.2), scenes.const_spectrum(9.0), g=0.3)
    m = ck.medium_uniformgrid(gd)
    n = 600
    org = np.c_[rng.uniform(0, 1, (n, 2)), np.full(n, -1.0)]
    d = np.tile([0.05, 0.02, 1.0], (n, 1))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.c_[org, d].astype(np.float32)
    pix = np.c_[rng.integers(0, 500, n), rng.integers(0, 500, n)].astype(np.int32)
    ts = np.c_[rng.uniform(1.0, 1.5, n), rng.uniform(0.05, 0.4, n)].astype(np.float32)
    out["graph_light_rays"] = m.graph_light_rays(rays, pix, ts, 8, 0, LAMBDA)
    m.close()
    return out


def case_graphfile(ck):
    """src/graph: FreeGraph containers (AddEdge merge / drop, graph.cpp:202-235), the text format
    (graph.cpp:284-416, :644-655; bytes of the written file), LightingCalculator::GetTransportMatrix
    (lighting_calculator.cpp:61-82) and ComputeFinalLight over that matrix (:23-59)."""
    import tempfile
    out = {}
    rng = np.random.default_rng(12)
    for n, ne in ((40, 150), (1500, 7000)):
        r = scenes.random_edge_graph(n, ne, seed=n)
        vid = np.arange(n, dtype=np.int32)
        eid = np.arange(r["efrom"].size, dtype=np.int32)
        pid = np.arange(r["poff"].size - 1, dtype=np.int32)
        g = ck.graph_build(0.0125, vid, r["pos"], vtype=r["vtype"], vlight=r["vlight"], vsamples=r["vsamples"],
                           vrange=r["vrange"], vcoors=r["vcoors"], eid=eid, efrom=r["efrom"], eto=r["eto"],
                           esamples=r["esamples"], pid=pid, poff=r["poff"], pverts=r["pverts"])
        for k, v in g.get().items():
            out[f"g{n}_{k}"] = v
        L0 = rng.random(n, dtype=np.float32)
        for b in (0, 3):
            rowptr, col, val, total, iters = g.transport(L0, b)
            out[f"g{n}_b{b}_total"] = total
            out[f"g{n}_b{b}_iters"] = np.array([iters], dtype=np.int32)
        out[f"g{n}_rowptr"], out[f"g{n}_col"], out[f"g{n}_val"] = rowptr, col, val
        with tempfile.TemporaryDirectory() as td:
            # what graph_maker writes (cmd/graph_maker.cpp:171-173): vertices only, lighting + search range
            for tag, flags, opts in (("maker", (0, 0, 0, 1, 1), (1, 0, 0)), ("full", (1, 1, 1, 1, 1), (1, 1, 1)),
                                     ("bare", (0, 0, 0, 0, 0), (1, 1, 0))):
                path = f"{td}/{tag}.txt"
                g.write(path, "basic", flags, opts)
                data = open(path, "rb").read()
                # records appear in the writer's container order (an unordered_map in the reference):
                # compare the file as a header + a sorted set of record lines
                lines = data.split(b"\n")
                out[f"g{n}_{tag}_header"] = np.frombuffer(b"\n".join(lines[:4]), dtype=np.uint8).copy()
                out[f"g{n}_{tag}_records"] = np.frombuffer(b"\n".join(sorted(lines[4:])), dtype=np.uint8).copy()
                g2 = ck.graph_read(path)
                c, radius = g2.counts()
                out[f"g{n}_{tag}_counts"] = c
                out[f"g{n}_{tag}_radius"] = np.array([radius], dtype=np.float32)
                for k, v in g2.get().items():
                    out[f"g{n}_{tag}_read_{k}"] = v
                g2.close()
        g.close()
    return out


# ---- PerspectiveCamera (cameras.h:343-370, cameras.cpp:428-490) + GetCameraSample (samplers.h:797-814) ----
def camera_configs():
    """(width, height, fov, eye, sampler seed): landscape and portrait films (fov spans the shorter axis)."""
    return [(48, 36, 20.0, (0.0, 0.35, -3.2), 0), (31, 57, 37.5, (1.5, -0.7, 2.2), 3)]


def camera_samples(cfg, n=400):
    W, H = cfg[0], cfg[1]
    rng = np.random.default_rng(2 + W)
    s = np.stack([rng.integers(0, W, n), rng.integers(0, H, n), rng.integers(0, 4, n)], axis=1).astype(np.int32)
    s[0] = (0, 0, 0); s[1] = (W - 1, H - 1, 3)
    return s


def case_camera_reference(ref):
    """Reference side only (fixture keys camera/*): matrices the real PerspectiveCamera constructor builds and
    the rays its GenerateRayDifferential returns for pixel samples drawn as EvaluatePixelSample draws them."""
    out = {}
    for i, (W, H, fov, eye, seed) in enumerate(camera_configs()):
        aspect = W / H
        sw = (-aspect, -1, aspect, 1) if aspect > 1 else (-1, -1 / aspect, 1, 1 / aspect)
        rays, mats = ref.camera_rays_reference(scenes.look_at(eye, (0, 0, 0), (0, 1, 0)), fov, sw, W, H, seed, 4,
                                               camera_samples((W, H)))
        out[f"rays{i}"] = rays
        out[f"matrices{i}"] = mats
    return out


def camera_rays_restated(orc, matrices, i):
    """The restatement's camera rays for configuration i, fed with the reference's two matrices."""
    W, H, fov, eye, seed = camera_configs()[i]
    dens = scenes.fbm_density(8, 8, 8, seed=1)
    gd, rd = scenes.cloud_scene(dens, width=W, height=H, spp=4, fov=fov, eye=eye, sampler_seed=seed)
    for k in range(16):
        rd.c.camera_from_raster[k] = float(matrices[0].ravel()[k])
        rd.c.render_from_camera[k] = float(matrices[1].ravel()[k])
    m = orc.medium_uniformgrid(gd)
    _, rays = m.camera_rays(rd, camera_samples((W, H)))
    m.close()
    own = np.stack([np.array(list(scenes.cloud_scene(dens, width=W, height=H, spp=4, fov=fov, eye=eye)[1].c.camera_from_raster),
                             dtype=np.float32).reshape(4, 4),
                    np.array(list(scenes.cloud_scene(dens, width=W, height=H, spp=4, fov=fov, eye=eye)[1].c.render_from_camera),
                             dtype=np.float32).reshape(4, 4)])
    return rays, own


# ---- PixelSensor::ToSensorRGB, RGBFilm::AddSample / GetPixelRGB (film.h:95-100, :239-275) ---------------
def film_case_inputs():
    """A 12x9 film, 4000 samples: radiance over 8 decades, zero pdfs (terminated secondary wavelengths), black
    samples, and a clamp (maxcomponentvalue) that bites on part of them; a non-trivial output matrix."""
    dens = scenes.fbm_density(8, 8, 8, seed=1)
    _, rd = scenes.cloud_scene(dens, width=12, height=9, spp=1)
    rd = rd.with_(max_component_value=2.5)
    M = np.array([1.2, -0.15, -0.05, -0.1, 1.3, -0.2, 0.02, -0.3, 1.28], dtype=np.float32)
    for i, v in enumerate(M):
        rd.c.output_rgb_from_sensor[i] = float(v)
    rng = np.random.default_rng(21)
    n = 4000
    pixels = np.stack([rng.integers(0, 12, n), rng.integers(0, 9, n)], axis=1).astype(np.int32)
    lam = rng.uniform(360, 830, (n, 4)).astype(np.float32)
    pdf = (0.0039398042 / np.cosh(0.0072 * (lam.astype(np.float64) - 538)) ** 2).astype(np.float32)
    L = (10.0 ** rng.uniform(-6, 2, (n, 4))).astype(np.float32) * pdf
    pdf[rng.random((n, 4)) < 0.1] = 0          # SampledWavelengths::TerminateSecondary
    L[rng.random(n) < 0.05] = 0
    return rd, pixels, np.concatenate([L, lam, pdf], axis=1).astype(np.float32)


def case_film(ck):
    rd, pixels, vals = film_case_inputs()
    film, rgb = ck.film_add_samples(rd, pixels, vals)
    return {"film": film, "rgb": rgb}  # the product's avr_film_resolve is checked against "rgb" too (tests/test_abi.py)


# ---- GraphIntegrator::Initialize + ConnectToGraph (graph/graph_integrator.cpp:24-80, :249-280) ----------
def case_graphconnect(ck):
    """ck "ref": the reference's own compiled GraphIntegrator methods over a graph file written and read by the
    reference (oracle/refprobe_graphint.cpp); ck "orc": the restatement over the same file contents (its
    writer is byte-identical, case "graphfile")."""
    import tempfile
    from cpu_checker import RefGraphIntegrator
    n, ne, radius = 1500, 7000, 0.05
    r = scenes.random_edge_graph(n, ne, seed=n)
    g = ck.graph_build(radius, np.arange(n), r["pos"], vtype=r["vtype"], vlight=r["vlight"], vsamples=r["vsamples"],
                       vrange=r["vrange"] * 3, vcoors=r["vcoors"])
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        os.makedirs(os.path.join(tmp, "files", "graphs"))
        for use_range in (1, 0):
            path = os.path.join(tmp, "files", "graphs", f"probe{use_range}.txt")
            g.write(path, "basic", (0, 0, 0, 1, use_range), (1, 0, 0))  # what graph_maker writes
            back = ck.graph_read(path); d = back.get(); back.close()
            rng = np.random.default_rng(3)
            near = d["pos"].reshape(-1, 3)[rng.integers(0, n, 600)] + rng.normal(0, radius * 0.6, (600, 3)).astype(np.float32)
            far = rng.uniform(-1, 3, (600, 3)).astype(np.float32)
            pts = np.concatenate([near, far]).astype(np.float32)
            if ck.kind == "ref":
                gi = RefGraphIntegrator(ck, tmp, f"probe{use_range}")
                out[f"info{use_range}"] = gi.info()[:3]
                out[f"connect{use_range}"] = gi.connect(pts)
                gi.close()
            else:
                out[f"info{use_range}"] = ck.graph_connect_info(d["vrange"] if use_range else None, n, radius)
                out[f"connect{use_range}"] = ck.graph_connect(d["pos"], d["vlight"], d["vrange"] if use_range else None, radius, pts)
    g.close()
    return out


# ---- ZSobolSampler (samplers.h:225-364) + SobolSample randomizers (util/lowdiscrepancy.h:168-258) -------
def zsobol_configs():
    """(spp, film resolution, RandomizeStrategy, seed): powers of 4, of 2 only, non-powers of two, odd films."""
    return [(1, (1280, 720), 2, 0), (16, (1280, 720), 2, 0), (256, (1280, 720), 2, 7), (8, (64, 48), 2, 3),
            (2, (33, 17), 3, 5), (64, (40, 30), 3, -2), (4, (48, 36), 1, 1), (32, (4096, 4096), 0, 0),
            (3, (20, 20), 2, 9), (100, (1, 1), 3, 4)]


def zsobol_samples_list(cfg, n=96):
    spp, res, _, seed = cfg
    rng = np.random.default_rng(1000 + spp + res[0])
    s = np.stack([rng.integers(0, res[0], n), rng.integers(0, res[1], n), rng.integers(0, spp, n)], axis=1)
    s[0] = (res[0] - 1, res[1] - 1, spp - 1)
    s[1] = (0, 0, 0)
    return s.astype(np.int32)


def case_zsobol(ck):
    out = {"sobol_matrices_dim01": ck.sobol_matrices()}
    for i, cfg in enumerate(zsobol_configs()):
        spp, res, rz, seed = cfg
        # 24 rounds of {Get1D, Get2D} = 72 dimensions: past the camera sample and several path vertices
        out[f"stream{i}"] = ck.zsobol_samples(spp, res, rz, seed, zsobol_samples_list(cfg), 24)
    return out


CASES = {"film": case_film, "zsobol": case_zsobol, "graphconnect": case_graphconnect, "scalar": case_scalar, "dda": case_dda, "transform": case_transform, "uniformgrid": case_uniformgrid,
         "rgbgrid": case_rgbgrid, "nanovdb": case_nanovdb, "graph": case_graph, "graphfile": case_graphfile}


# ---- VolPathIntegrator::Li as a whole (cpu/integrators.cpp:962-1106, :1282-1399) -----------------------
# Not part of CASES: the comparison is by tolerance, not bit-exact.  The reference spawns rays at the
# boundary sphere with its error-bound offsets (interaction.cpp), the restatement does not; positions then
# differ in the last bits, and because the reference seeds every shadow ray's RNG from Hash(ray.o),
# Hash(ray.d) (cpu/integrators.cpp:1338), next-event estimates de-correlate completely.  Hence two kinds:
#   "lockstep": emissive medium + black light: every random number comes from the sampler streams, the
#               two implementations walk the same path and per-sample radiance agrees to rounding;
#   "image":    the full estimator (NEE on), compared as converged images against the reference's own
#               seed-to-seed noise.
def li_scenes():
    from acceleratedvolrenderer_b200._abi import GridDesc
    out = []
    dens = scenes.fbm_density(48, 40, 52, seed=4)
    gd, rd = scenes.cloud_scene(dens, width=48, height=36, spp=4, sigma_t=14.0, albedo=0.95, g=0.4, max_depth=8)
    em = GridDesc(dens, gd.sigma_a, gd.sigma_s, render_from_medium=gd.m, g=0.4, temperature=dens * 2500 + 200,
                  lescale=dens * 0.5, le_norm=1.0)
    black = scenes.cloud_scene(dens, width=48, height=36, spp=4, sigma_t=14.0, albedo=0.95, g=0.4, max_depth=8,
                               light_scale=0.0)[1]
    out.append(("lockstep_grid", "uniformgrid", em, black))
    # box boundary: the reference side is a 12-triangle "trianglemesh" cube under the default BVH aggregate
    black_box = scenes.cloud_scene(dens, width=48, height=36, spp=4, sigma_t=14.0, albedo=0.95, g=0.4, max_depth=8,
                                   light_scale=0.0, boundary="box")[1]
    out.append(("lockstep_grid_box", "uniformgrid", em, black_box))
    d2 = dens.copy(); d2[d2 < 0.02] = 0
    nd, rdn = scenes.cloud_scene_nanovdb(d2, temperature=(d2 * 2500 + 200).astype(np.float32), le_scale=0.4, width=40,
                                         height=30, spp=4, sigma_t=18.0, albedo=0.95, g=0.4, max_depth=6, light_scale=0.0)
    out.append(("lockstep_nanovdb", "nanovdb", nd, rdn))
    # the same scene through pbrt's default sampler ("zsobol", fastowen) and its full Owen variant
    out.append(("lockstep_nanovdb_zsobol", "nanovdb", nd, rdn.with_(sampler=1, randomize=2)))
    out.append(("lockstep_grid_zsobol_owen", "uniformgrid", em, black.with_(sampler=1, randomize=3, sampler_seed=5)))
    gdc, rdc = scenes.cloud_scene(dens, width=32, height=24, spp=128, sigma_t=14.0, albedo=0.95, g=0.4, max_depth=6,
                                  chromatic=True)
    out.append(("image_chromatic", "uniformgrid", gdc, rdc))
    ndi, rdi = scenes.cloud_scene_nanovdb(d2, width=32, height=24, spp=128, sigma_t=18.0, albedo=0.95, g=0.4, max_depth=6)
    out.append(("image_nanovdb", "nanovdb", ndi, rdi))
    return out


def li_medium(ck, kind, desc):
    return ck.medium_uniformgrid(desc) if kind == "uniformgrid" else ck.medium_nanovdb(desc)


def li_mono_image(samples12, rdesc):
    """Per-pixel mean over samples of Average(L / pdf): the film estimate before the sensor response."""
    x0, y0, x1, y1, s0, s1 = rdesc.ranges()
    L = samples12[:, :4].astype(np.float64); pdf = samples12[:, 8:12].astype(np.float64)
    v = np.where(pdf > 0, L / np.where(pdf > 0, pdf, 1), 0).mean(axis=1)
    return v.reshape(y1 - y0, x1 - x0, s1 - s0).mean(axis=2)


def case_volpath_li(ck):
    """Reference outputs kept in the fixture (ck = Checker("ref")), or the restatement's (Checker("orc"))."""
    out = {}
    for name, kind, desc, rd in li_scenes():
        m = li_medium(ck, kind, desc)
        if name.startswith("lockstep"):
            smp, info = m.volpath_li(rd)
            out[f"{name}_samples"] = smp
            out[f"{name}_info"] = info
        else:
            for seed in (0, 1):
                smp, info = m.volpath_li(rd.with_(sampler_seed=seed))
                out[f"{name}_mono{seed}"] = li_mono_image(smp, rd).astype(np.float32)
            out[f"{name}_info"] = info
        m.close()
    return out
