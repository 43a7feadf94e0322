"""Pixel filters that pbrt samples through its FilterSampler ("gaussian" = pbrt's default, "mitchell", "sinc"):
src/pbrt/filters.h:26-45, filters.cpp:133-147, util/sampling.h:603-790, GetCameraSample samplers.h:797-814; and the
"triangle" filter (filters.h:208-231), sampled analytically.

CPU: the restatement (oracle/restate_filter.hpp) against the reference's OWN filter objects through the probe, bit for
bit: the tabulated filter, Filter::Sample over random and edge-case u, and the camera rays + CameraSample::filterWeight
the real PerspectiveCamera / GetCameraSample produce.  GPU: the device code the render kernels run against the
restatement, bit for bit, and renders under a Gaussian filter against the oracle's."""
import numpy as np
import pytest

from acceleratedvolrenderer_b200 import scenes
from acceleratedvolrenderer_b200._abi import gaussian_filter_table

import golden_cases

FILTERS = {  # (kind, rx, ry, p0, p1) as oracle/refprobe_volpath.cpp make_filter
    "gaussian": (1, 1.5, 1.5, 0.5, 0.0),          # pbrt's defaults (filters.cpp:108-116)
    "gaussian-wide": (1, 2.25, 1.0, 0.8, 0.0),
    "mitchell": (2, 2.0, 2.0, 1.0 / 3.0, 1.0 / 3.0),
    "sinc": (3, 4.0, 4.0, 3.0, 0.0),
}


def sample_points(n=60000, seed=3):
    rng = np.random.default_rng(seed)
    u = rng.random((n, 2), dtype=np.float32)
    eps1 = np.float32(1) - np.float32(2 ** -24)
    edge = np.array([[0, 0], [eps1, eps1], [0, eps1], [eps1, 0], [0.5, 0.5], [1e-30, 1e-30], [0.25, 0.75]], dtype=np.float32)
    u[:edge.shape[0]] = edge
    return u


@pytest.mark.parametrize("name", list(FILTERS))
def test_filter_sampling_restated_matches_the_reference_filters(orc, ref, name):
    cfg = FILTERS[name]
    table = ref.filter_table(cfg)
    assert table.shape == (int(np.float32(32) * np.float32(cfg[2])), int(np.float32(32) * np.float32(cfg[1])))
    u = sample_points()
    want = ref.filter_sample(u, filter_cfg=cfg)
    got = orc.filter_sample(u, table=table, radius=cfg[1:3])
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.all(np.abs(want[:, 0]) <= cfg[1]) and np.all(np.abs(want[:, 1]) <= cfg[2])
    if name.startswith("gaussian"):
        assert want[:, 2].min() > 0 and np.ptp(want[:, 2]) < 1e-3 * want[:, 2].mean()   # importance sampled: ~constant weight
    else:
        assert (want[:, 2] < 0).any()                                                  # negative lobes keep their sign


@pytest.mark.parametrize("radius", [(2.0, 2.0), (0.7, 3.25)])
def test_triangle_filter_restated_matches_the_reference_filter(orc, ref, radius):
    """TriangleFilter::Sample (filters.h:224-227): SampleTent per axis, weight 1 — the one pbrt filter that is neither
    the box nor tabulated.  Bit for bit against the reference's object, incl. u = 0, 0.5 and 1 - ulp."""
    u = sample_points()
    want = ref.filter_sample(u, filter_cfg=(4, radius[0], radius[1], 0, 0))
    got = orc.filter_sample(u, table=None, radius=radius)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.all(want[:, 2] == 1) and np.all(np.abs(want[:, 0]) <= radius[0]) and np.all(np.abs(want[:, 1]) <= radius[1])
    assert abs(np.abs(want[:, 0]).mean() / radius[0] - 1 / 3) < 0.01      # a tent: E|x| = r / 3


def test_gaussian_table_helper_is_the_reference_table(ref):
    """_abi.gaussian_filter_table (float64 exp, rounded to fp32) against GaussianFilter::Evaluate's fp32 table (pbrt's
    polynomial FastExp, then the difference G(p) - G(radius), which cancels near the rim): equal to 0.5 % of a value
    plus 1e-7 absolute — the adapter and the parity tests pass the reference's OWN table; the helper is for standalone
    use, where a table that differs in the last digits is simply a slightly different filter."""
    t = ref.filter_table(FILTERS["gaussian"])
    mine = gaussian_filter_table((1.5, 1.5), 0.5)
    assert mine.shape == t.shape
    assert np.allclose(mine, t, rtol=5e-3, atol=1e-7)


@pytest.mark.parametrize("name", ["gaussian", "mitchell", "triangle"])
def test_camera_rays_and_weights_under_a_filter(orc, ref, name):
    """EvaluatePixelSample's prologue with the REAL PerspectiveCamera and the reference's own filter object: rays of the
    restatement bit-equal (fed with the reference's matrices), weights bit-equal to Filter::Sample's."""
    cfg = (4, 2.0, 2.0, 0.0, 0.0) if name == "triangle" else FILTERS[name]
    W, H, fov, eye, seed = golden_cases.camera_configs()[0]
    aspect = W / H
    sw = (-aspect, -1, aspect, 1) if aspect > 1 else (-1, -1 / aspect, 1, 1 / aspect)
    samples = golden_cases.camera_samples((W, H))
    rays, mats, weights = ref.camera_rays_reference(scenes.look_at(eye, (0, 0, 0), (0, 1, 0)), fov, sw, W, H, seed, 4, samples,
                                                    filter_cfg=cfg, want_weights=True)
    dens = scenes.fbm_density(8, 8, 8, seed=1)
    gd, rd = scenes.cloud_scene(dens, width=W, height=H, spp=4, fov=fov, eye=eye, sampler_seed=seed)
    for k in range(16):
        rd.c.camera_from_raster[k] = float(mats[0].ravel()[k])
        rd.c.render_from_camera[k] = float(mats[1].ravel()[k])
    rg = rd.with_filter("triangle" if name == "triangle" else ref.filter_table(cfg), cfg[1:3])
    m = orc.medium_uniformgrid(gd)
    _, mine = m.camera_rays(rg, samples)
    _, boxed = m.camera_rays(rd, samples)
    m.close()
    assert np.array_equal(mine.view(np.uint32), rays.view(np.uint32))
    assert not np.array_equal(mine, boxed)
    if name == "gaussian":
        assert weights.min() > 0
    elif name == "triangle":
        assert np.all(weights == 1)
    else:
        assert (weights < 0).any()
    # the restatement's CameraSample::filterWeight of the same samples (full-frame order: pick the tested ones)
    x0, y0, x1, y1, s0, s1 = rg.ranges()
    w_all = orc.camera_weights(rg).reshape(y1 - y0, x1 - x0, s1 - s0)
    mine_w = w_all[samples[:, 1] - y0, samples[:, 0] - x0, samples[:, 2] - s0]
    assert np.array_equal(mine_w.view(np.uint32), weights.view(np.uint32))


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(FILTERS))
def test_device_filter_sampling_is_bit_exact(ctx, orc, name):
    """avr_filter_sample = the device function start_path / write_sample call, against the restatement."""
    cfg = FILTERS[name]
    rx, ry = np.float32(cfg[1]), np.float32(cfg[2])
    rng = np.random.default_rng(5)
    nx, ny = int(np.float32(32) * rx), int(np.float32(32) * ry)
    if name.startswith("gaussian"):
        table = gaussian_filter_table((cfg[1], cfg[2]), cfg[3])
    else:   # any table will do for the sampling code: smooth with sign changes
        yy, xx = np.meshgrid(np.linspace(-1, 1, ny), np.linspace(-1, 1, nx), indexing="ij")
        table = (np.cos(3 * xx) * np.cos(2 * yy) * np.exp(-xx * xx - yy * yy)).astype(np.float32)
    u = sample_points(seed=int(rng.integers(1 << 30)))
    want = orc.filter_sample(u, table=table, radius=(cfg[1], cfg[2]))
    got = ctx.filter_sample(table, (cfg[1], cfg[2]), u)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.gpu
def test_device_triangle_filter_is_bit_exact_and_renders(ctx, orc):
    """kind AVR_FILTER_TRIANGLE: the device tent sampling against the restatement bit for bit; a render under it has
    unit weights (weightSum = spp) and the oracle's per-sample radiance."""
    u = sample_points(seed=11)
    for radius in ((2.0, 2.0), (0.7, 3.25)):
        want = orc.filter_sample(u, table=None, radius=radius)
        got = ctx.filter_sample("triangle", radius, u)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    dens = scenes.fbm_density(48, 40, 52, seed=4)
    gd, rd = scenes.cloud_scene(dens, width=40, height=30, spp=4, sigma_t=14.0, albedo=0.95, g=0.4, max_depth=8)
    rt = rd.with_filter("triangle", (2.0, 2.0))
    mo = orc.medium_uniformgrid(gd)
    md = ctx.create_medium("uniformgrid", gd, layout="octet32")
    want, _ = mo.render_samples(rt)
    got, _ = ctx.render_volpath_samples(md, rt)
    L, Lw = got[:, :4].astype(np.float64), want[:, :4].astype(np.float64)
    ok = np.abs(L - Lw).max(axis=1) <= 1e-3 * np.maximum(np.abs(Lw).max(axis=1), 1e-6)
    assert ok.mean() >= 0.99, ok.mean()
    boxed, _ = ctx.render_volpath_samples(md, rd)
    assert not np.allclose(boxed[:, :4], got[:, :4])
    film, _ = ctx.render_volpath(md, rt)
    assert np.all(film[..., 3] == 4)
    ofilm, _ = mo.render(rt, threads=8)
    a = film[..., :3].sum(axis=2); b = ofilm[..., :3].sum(axis=2)
    good = np.abs(a - b) <= 2e-3 * np.maximum(b, 1e-3 * b.max())
    assert good.mean() >= 0.97, good.mean()
    md.close(); mo.close()


@pytest.mark.gpu
@pytest.mark.parametrize("sampler", ["independent", "zsobol"])
def test_render_under_a_gaussian_filter_matches_oracle(ctx, orc, sampler):
    """The default pixel filter of an unmodified scene file: per-sample radiance under the exact arithmetic against the
    oracle (same camera offsets, so the same paths), and the device film = sum of weight * rgb over its samples with
    weightSum = sum of weights, against the oracle's film of the same samples."""
    dens = scenes.fbm_density(48, 40, 52, seed=4)
    gd, rd = scenes.cloud_scene(dens, width=40, height=30, spp=4, sigma_t=14.0, albedo=0.95, g=0.4, max_depth=8)
    table = gaussian_filter_table((1.5, 1.5), 0.5)
    rg = rd.with_filter(table, (1.5, 1.5)).with_(sampler={"independent": 0, "zsobol": 1}[sampler], randomize=2)
    mo = orc.medium_uniformgrid(gd)
    md = ctx.create_medium("uniformgrid", gd, layout="octet32")
    want, wstats = mo.render_samples(rg)
    got, gstats = ctx.render_volpath_samples(md, rg)
    L, Lw = got[:, :4].astype(np.float64), want[:, :4].astype(np.float64)
    ok = np.abs(L - Lw).max(axis=1) <= 1e-3 * np.maximum(np.abs(Lw).max(axis=1), 1e-6)
    assert ok.mean() >= 0.99, ok.mean()
    assert np.allclose(got[:, 4:8], want[:, 4:8], rtol=1e-6)
    boxed, _ = ctx.render_volpath_samples(md, rd.with_(sampler=rg.c.sampler, randomize=2))
    assert not np.allclose(boxed[:, :4], got[:, :4])                   # the filter moved the camera rays
    film, _ = ctx.render_volpath(md, rg)
    ofilm, _ = mo.render(rg, threads=8)
    wsum = film[..., 3]
    assert np.allclose(wsum, ofilm[..., 3], rtol=1e-6) and wsum.min() > 0 and abs(wsum.mean() / 4 - 1) < 0.1
    a = film[..., :3].sum(axis=2); b = ofilm[..., :3].sum(axis=2)
    good = np.abs(a - b) <= 2e-3 * np.maximum(b, 1e-3 * b.max())
    assert good.mean() >= 0.97, good.mean()
    # FAST arithmetic under the filter: same estimator, image mean within the Monte Carlo error of 16 renders
    means = []
    for k in range(16):
        f, _ = ctx.render_volpath(md, rg.with_(exact_math=0, sampler_seed=10 + k))
        means.append(f[..., :3].sum() / f[..., 3].sum())
    ref_means = []
    for k in range(16):
        f, _ = ctx.render_volpath(md, rg.with_(exact_math=1, sampler_seed=100 + k))
        ref_means.append(f[..., :3].sum() / f[..., 3].sum())
    means, ref_means = np.array(means), np.array(ref_means)
    se = np.sqrt(means.var(ddof=1) / 16 + ref_means.var(ddof=1) / 16)
    assert abs(means.mean() - ref_means.mean()) < 4 * se, (means.mean(), ref_means.mean(), se)
    md.close(); mo.close()


@pytest.mark.gpu
def test_filter_table_argument_errors(ctx):
    """A malformed avr_filter_table is refused with a message, never read out of bounds."""
    from acceleratedvolrenderer_b200.api import AvrError
    u = np.zeros((4, 2), dtype=np.float32)
    good = gaussian_filter_table()
    assert ctx.filter_sample(good, (1.5, 1.5), u).shape == (4, 3)
    with pytest.raises(AvrError) as e:
        ctx.filter_sample(good, (0.0, 1.5), u)
    assert "radius" in str(e.value)
    with pytest.raises(AvrError) as e:
        ctx.filter_sample(np.zeros((0, 48), dtype=np.float32), (1.5, 1.5), u)
    assert "nx, ny" in str(e.value)
    # an all-zero table: the reference's distribution falls back to uniform CDFs and the weight is 0 / 0
    z = ctx.filter_sample(np.zeros((8, 8), dtype=np.float32), (1.0, 1.0), np.full((4, 2), 0.5, dtype=np.float32))
    assert np.all(np.abs(z[:, :2]) <= 1.0)
