"""GPU: BASELINE config 4 as a PIPELINE, from a graph the REFERENCE builds.

The reference's own FreeGraphBuilder::BuildGraph -> LightingCalculator::GetLightVector -> GetTransportMatrix ->
ComputeFinalLight (oracle/refprobe_graphbuild.cpp, driven as cmd/graph_maker.cpp:86-197 drives them) produces
vertices, edges, the light vector L0 and the per-vertex light after {0, 1, 3, 7} bounces.  The device then does
the same steps through the C ABI on that graph:
  (a) avr_graph_light_vector           vs GetLightVector      : >= 97 % of vertices within 1e-4 relative (the
      reference intersects its Sphere shape and respawns the ray, the device hits the sphere analytically: ray
      origins differ in the last bits, which now and then moves one tentative collision), means within 0.5 %;
  (b) avr_graph_create_from_edges + avr_graph_propagate from the reference's L0 vs ComputeFinalLight: bit-exact,
      identical iteration counts;
  (c) avr_render_graph of the propagated graph vs avr_render_volpath at max_depth = bounces + 1: image means
      within the band BASELINE.md section 3 records for the fork itself (graph_d1 / graph_d8 vs volpath < 0.5 %),
      widened by the Monte Carlo standard error of the two renders and the clustering bias of a 2 300-vertex
      graph: 4 %.
"""
import numpy as np
import pytest

from acceleratedvolrenderer_b200 import scenes

pytestmark = pytest.mark.gpu

CFG = dict(dimension_steps=80, iterations_per_step=8, max_depth=40, light_iterations=8, points_on_radius=2,
           radius_modifier=12.0, search_range_neighbours=4)
BOUNCES = [0, 1, 3, 7]


def test_config4_pipeline_from_a_reference_built_graph(ctx, ref):
    dens = scenes.fbm_density(48, 40, 52, seed=4)
    gd, rd = scenes.cloud_scene(dens, width=48, height=36, spp=64, sigma_t=14.0, albedo=0.95, g=0.0, max_depth=8)
    mo = ref.medium_uniformgrid(gd)
    pipe = mo.graph_pipeline(rd, bounces=BOUNCES, **CFG)
    mo.close()
    nv = pipe["pos"].shape[0]
    assert nv > 1500 and pipe["efrom"].size > nv

    md = ctx.create_medium("uniformgrid", gd, layout="octet32")
    # (a) the light vector
    L0 = md.graph_light_vector(rd, pipe["pos"], vertex_radius=pipe["vertex_radius"],
                               max_dist_to_center=pipe["max_dist_to_center"], light_iterations=CFG["light_iterations"],
                               points_on_radius=CFG["points_on_radius"], sampling_resolution=pipe["sampling_resolution"],
                               lam=pipe["lam"], list_index=pipe["list_index"])
    want = pipe["L0"]
    ok = np.abs(L0 - want) <= 1e-4 * np.maximum(np.abs(want), 1e-6)
    print(f"light vector: {ok.mean() * 100:.2f} % of {nv} vertices within 1e-4, mean {L0.mean():.6f} vs {want.mean():.6f}")
    assert ok.mean() >= 0.97, ok.mean()
    assert abs(L0.mean() - want.mean()) <= 5e-3 * want.mean(), (L0.mean(), want.mean())
    assert np.all(L0 >= 0) and L0.max() <= 1 / (4 * np.pi) + 1e-6

    # (b) transport matrix from the edge list + propagation
    g = ctx.create_graph_from_edges(pipe["vsamples"], pipe["efrom"], pipe["eto"], pipe["esamples"])
    totals = {}
    for k, b in enumerate(BOUNCES):
        got, it = g.compute_final_light(want, int(b))
        assert it == int(pipe["iters"][k])
        assert np.array_equal(got.view(np.uint32), pipe["totals"][k].view(np.uint32)), b
        # the device's own L0 through the same propagation stays as close as L0 itself is
        mine, _ = g.compute_final_light(L0, int(b))
        assert abs(mine.mean() - got.mean()) <= 5e-3 * got.mean()
        totals[int(b)] = mine

    # (c) graph render vs path tracing of the same depth (both on the device; independent sampler)
    means = {}
    for b in BOUNCES:
        g.set_vertices(pipe["pos"], totals[b], pipe["vrange"], pipe["vertex_radius"])
        gf, _ = ctx.render_graph(md, g, rd)
        vf, _ = ctx.render_volpath(md, rd.with_(max_depth=b + 1, spp=256))
        mg = gf[..., :3].sum() / gf[..., 3].sum(); mv = vf[..., :3].sum() / vf[..., 3].sum()
        means[b] = (float(mg), float(mv))
        print(f"bounces {b}: graph render mean {mg:.6f} vs volpath max_depth {b + 1} mean {mv:.6f} ({(mg / mv - 1) * 100:+.2f} %)")
        assert abs(mg - mv) <= 0.04 * mv, (b, mg, mv)
    # more bounces, more light, in both renderers
    assert means[0][0] < means[1][0] < means[3][0] < means[7][0]
    assert means[0][1] < means[1][1] < means[3][1] < means[7][1]
    g.close(); md.close()


def _merged_edges(efrom, eto, n):
    key = np.unique(efrom.astype(np.int64) * n + eto.astype(np.int64))
    return key.size, np.bincount((key // n).astype(np.int64), minlength=n)


def test_device_graph_builder_is_statistically_the_reference_builder(ctx, ref):
    """avr_graph_build_free (light paths traced per thread, the first-come vertex merge as a lexicographically-first
    independent set) against FreeGraphBuilder::BuildGraph on the same scene, sampler streams and configuration.
    The paths are the reference's up to last-bit differences of the scatter positions, so the counts nearly
    coincide; the bars are what "statistically the same graph" needs, each stated in place."""
    dens = scenes.fbm_density(48, 40, 52, seed=4)
    gd, rd = scenes.cloud_scene(dens, width=48, height=36, spp=64, sigma_t=14.0, albedo=0.95, g=0.0, max_depth=8)
    mo = ref.medium_uniformgrid(gd)
    pipe = mo.graph_pipeline(rd, bounces=BOUNCES, **CFG)
    mo.close()
    md = ctx.create_medium("uniformgrid", gd, layout="octet32")
    built = md.build_free_graph(rd, vertex_radius=pipe["vertex_radius"], max_dist_to_center=pipe["max_dist_to_center"],
                                bounds_center=pipe["bounds_center"], dimension_steps=CFG["dimension_steps"],
                                iterations_per_step=CFG["iterations_per_step"], max_depth=CFG["max_depth"],
                                sampling_resolution=pipe["sampling_resolution"], lam=pipe["lam"],
                                search_range_neighbours=CFG["search_range_neighbours"])
    nv_ref, nv = pipe["pos"].shape[0], built["pos"].shape[0]
    ne_ref, deg_ref = _merged_edges(pipe["efrom"], pipe["eto"], nv_ref)
    ne, deg = _merged_edges(built["efrom"], built["eto"], nv)
    print(f"vertices {nv} vs {nv_ref}; merged edges {ne} vs {ne_ref}; departures {built['vsamples'].sum()} vs "
          f"{pipe['vsamples'].sum()}; scatter points {built['n_points']}; rounds {built['cluster_rounds']}; "
          f"range mean {built['vrange'].mean():.5f} vs {pipe['vrange'].mean():.5f}")
    if nv == nv_ref:
        same = np.all(np.abs(built["pos"] - pipe["pos"]) <= 1e-5, axis=1).mean()
        ka = set((built["efrom"].astype(np.int64) * nv + built["eto"]).tolist())
        kb = set((pipe["efrom"].astype(np.int64) * nv + pipe["eto"]).tolist())
        print(f"vertices at the reference's position (1e-5): {same * 100:.2f} %; edge sets: Jaccard {len(ka & kb) / len(ka | kb):.4f}")
    assert abs(nv - nv_ref) <= 0.02 * nv_ref                     # vertex count within 2 %
    assert abs(ne - ne_ref) <= 0.03 * ne_ref                     # distinct edges within 3 %
    assert abs(int(built["vsamples"].sum()) - int(pipe["vsamples"].sum())) <= 0.005 * pipe["vsamples"].sum()
    assert abs(deg.mean() - deg_ref.mean()) <= 0.03 * deg_ref.mean()
    # out-degree histogram: total variation distance below 6 % (two independent 2 300-vertex samples of ONE
    # distribution over ~14 degree bins already differ by about 3 %)
    hmax = int(max(deg.max(), deg_ref.max())) + 1
    ha = np.bincount(deg, minlength=hmax) / nv; hb = np.bincount(deg_ref, minlength=hmax) / nv_ref
    print(f"out-degree histogram TVD {0.5 * np.abs(ha - hb).sum():.4f}")
    assert 0.5 * np.abs(ha - hb).sum() <= 0.06
    assert abs(built["vrange"].mean() - pipe["vrange"].mean()) <= 0.03 * pipe["vrange"].mean()
    # no two vertices closer than the vertex radius (what the first-come merge guarantees), every point merged
    from scipy.spatial import cKDTree
    pairs = cKDTree(built["pos"].astype(np.float64)).query_pairs(pipe["vertex_radius"] * 0.999)
    assert len(pairs) == 0
    assert built["efrom"].min() >= 0 and built["eto"].max() < nv
    # the device-built graph through the rest of the pipeline: image means of the graph render, device-built vs
    # reference-built graph, within 2 % at every bounce count
    L0 = md.graph_light_vector(rd, built["pos"], vertex_radius=pipe["vertex_radius"], max_dist_to_center=pipe["max_dist_to_center"],
                               light_iterations=CFG["light_iterations"], points_on_radius=CFG["points_on_radius"],
                               sampling_resolution=pipe["sampling_resolution"], lam=pipe["lam"])
    g = ctx.create_graph_from_edges(built["vsamples"], built["efrom"], built["eto"], built["esamples"])
    gr = ctx.create_graph_from_edges(pipe["vsamples"], pipe["efrom"], pipe["eto"], pipe["esamples"])
    for k, b in enumerate(BOUNCES):
        tot, _ = g.compute_final_light(L0, int(b))
        g.set_vertices(built["pos"], tot, built["vrange"], pipe["vertex_radius"])
        gr.set_vertices(pipe["pos"], pipe["totals"][k], pipe["vrange"], pipe["vertex_radius"])
        fa, _ = ctx.render_graph(md, g, rd); fb, _ = ctx.render_graph(md, gr, rd)
        ma = fa[..., :3].sum() / fa[..., 3].sum(); mb = fb[..., :3].sum() / fb[..., 3].sum()
        print(f"bounces {b}: graph render mean, device-built {ma:.6f} vs reference-built {mb:.6f} ({(ma / mb - 1) * 100:+.2f} %)")
        assert abs(ma - mb) <= 0.02 * mb, (b, ma, mb)
    g.close(); gr.close(); md.close()


def test_light_vector_edge_cases(ctx):
    from acceleratedvolrenderer_b200.api import AvrError
    dens = scenes.fbm_density(16, 16, 16, seed=2)
    gd, rd = scenes.cloud_scene(dens, width=8, height=8, spp=1, sigma_t=6.0)
    md = ctx.create_medium("uniformgrid", gd, layout="dense")
    kw = dict(vertex_radius=0.05, max_dist_to_center=1.3, light_iterations=2, points_on_radius=1, sampling_resolution=64,
              lam=[400, 500, 600, 700])
    # a vertex far outside the medium's bounding sphere: its disk rays never overlap the medium -> 0
    pos = np.array([[0.0, 0.0, 0.0], [5.0, 5.0, 5.0]], dtype=np.float32)
    L0 = md.graph_light_vector(rd, pos, **kw)
    assert 0 < L0[0] <= 1 / (4 * np.pi) + 1e-6 and L0[1] == 0
    assert md.graph_light_vector(rd, np.zeros((0, 3), np.float32), **kw).size == 0
    # points_on_radius = 0: the single ray through the centre (util.h:180-181)
    one = md.graph_light_vector(rd, pos[:1], **dict(kw, points_on_radius=0))
    assert 0 < one[0] <= 1 / (4 * np.pi) + 1e-6
    with pytest.raises(AvrError) as e:
        md.graph_light_vector(rd, pos, **dict(kw, light_iterations=0))
    assert "at least one light ray iteration" in str(e.value)   # the reference's ErrorExit text
    md.close()


def test_graph_side_work_shards_by_construction(ctx):
    """the light vector shards by vertex range (a rank computes a slice of the
    vertices, numbered for the sampler by `list_index`, and the slices ARE the whole vector, bit for bit), the graph
    render by sample index or by rank partition like the volpath render (partial host films add up to the whole)."""
    import math
    dens = scenes.fbm_density(32, 32, 32, seed=4)
    gd, rd = scenes.cloud_scene(dens, width=40, height=30, spp=6, sigma_t=14.0, albedo=0.95, g=0.0, max_depth=8)
    md = ctx.create_medium("uniformgrid", gd, layout="octet32")
    radius = float(rd.c.boundary[3])
    max_dist = radius * math.sqrt(3.0)
    vr = max_dist * 2 / 1000 * 12.0
    kw = dict(vertex_radius=vr, max_dist_to_center=max_dist, sampling_resolution=1024, lam=[450, 520, 610, 700])
    built = md.build_free_graph(rd, bounds_center=[float(rd.c.boundary[i]) for i in range(3)], dimension_steps=60,
                                iterations_per_step=4, max_depth=30, search_range_neighbours=4, **kw)
    pos = built["pos"]; nv = pos.shape[0]
    assert nv > 500
    lkw = dict(light_iterations=4, points_on_radius=2, **kw)
    idx = np.arange(nv, dtype=np.int32)
    whole = md.graph_light_vector(rd, pos, list_index=idx, **lkw)
    parts = [md.graph_light_vector(rd, pos[a:b], list_index=idx[a:b], **lkw) for a, b in ((0, nv // 3), (nv // 3, nv))]
    assert np.array_equal(np.concatenate(parts).view(np.uint32), whole.view(np.uint32))
    g = ctx.create_graph_from_edges(built["vsamples"], built["efrom"], built["eto"], built["esamples"])
    tot, _ = g.compute_final_light(whole, 3)
    g.set_vertices(pos, tot, built["vrange"], vr)
    full, _ = ctx.render_graph(md, g, rd)
    assert full[..., :3].sum() > 0 and np.all(full[..., 3] == rd.spp)
    acc = np.zeros_like(full)
    for r in range(3):
        f, _ = ctx.render_graph(md, g, rd.with_(rank=r, world=3))
        acc += f
    assert np.allclose(acc, full, rtol=1e-12, atol=1e-300)
    acc = np.zeros_like(full)
    for a, b in ((0, 2), (2, 6)):
        f, _ = ctx.render_graph(md, g, rd.with_(sample_begin=a, sample_end=b))
        acc += f
    assert np.allclose(acc, full, rtol=1e-12, atol=1e-300)
    g.close(); md.close()
