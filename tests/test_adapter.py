"""The adapter a maintainer adds to the reference tree (integration/avr_adapter.h), COMPILED against the reference's
headers (oracle/refprobe_adapter.cpp) and run on the reference's real objects: the descriptors it fills are checked
field by field against the ones scenes.py builds for the same scene (CPU), and the device renders through them (GPU).
"""
import ctypes as C
import math

import numpy as np
import pytest

from acceleratedvolrenderer_b200 import _abi, scenes

NSPEC = _abi.NSPEC


def _arr(ptr, n, dt=np.float32):
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dt, copy=True)


def _scene(**kw):
    dens = scenes.fbm_density(24, 20, 28, seed=3)
    args = dict(width=40, height=30, spp=4, sigma_t=9.0, albedo=0.9, g=0.35, max_depth=5)
    args.update(kw)
    return dens, scenes.cloud_scene(dens, **args)


def describe_medium(ref, mo):
    lib = ref.lib
    lib.ref_adapter_describe_medium.restype = C.c_void_p
    kind = C.c_int32(-1); desc = C.c_void_p()
    h = C.c_void_p(lib.ref_adapter_describe_medium(mo.h, C.byref(kind), C.byref(desc)))
    return h, kind.value, desc


def test_adapter_describes_the_reference_gridmedium(ref):
    dens, (gd, rd) = _scene()
    mo = ref.medium_uniformgrid(gd)
    h, kind, desc = describe_medium(ref, mo)
    assert kind == 0
    d = C.cast(desc, C.POINTER(_abi.avr_grid_desc)).contents
    want = gd.c
    assert (d.nx, d.ny, d.nz) == (want.nx, want.ny, want.nz)
    n = d.nx * d.ny * d.nz
    assert np.array_equal(_arr(d.density, n), gd.density.ravel())
    assert not d.temperature and not d.lescale and d.le_norm == want.le_norm
    assert list(d.p0) == list(want.p0) and list(d.p1) == list(want.p1)
    assert list(d.render_from_medium) == list(want.render_from_medium)
    assert list(d.medium_from_render) == list(want.medium_from_render)
    # the reference's constructor has applied "scale" to the spectra (media.cpp:229-230): same float products
    assert d.sigma_scale == 1.0
    assert np.array_equal(_arr(d.sigma_a, NSPEC), gd.sigma_a * np.float32(want.sigma_scale))
    assert np.array_equal(_arr(d.sigma_s, NSPEC), gd.sigma_s * np.float32(want.sigma_scale))
    assert not d.le                                             # Le = 0: not emissive
    assert d.g == want.g and d.temperature_offset == want.temperature_offset and d.temperature_scale == want.temperature_scale
    assert list(d.majorant_res) == [16, 16, 16]                 # media.cpp:229
    ref.lib.ref_adapter_free_medium(h, C.c_int32(kind))
    mo.close()


def test_adapter_describes_an_emissive_gridmedium(ref):
    dens, (gd, rd) = _scene()
    gd2 = _abi.GridDesc(dens, gd.sigma_a, gd.sigma_s, render_from_medium=gd.m, g=0.1, temperature=dens * 2000 + 300,
                        lescale=dens * 0.5, le_norm=0.25, temperature_offset=10.0, temperature_scale=1.5, scale=2.0)
    mo = ref.medium_uniformgrid(gd2)
    h, kind, desc = describe_medium(ref, mo)
    d = C.cast(desc, C.POINTER(_abi.avr_grid_desc)).contents
    n = d.nx * d.ny * d.nz
    assert np.array_equal(_arr(d.temperature, n), gd2.temperature.ravel())
    # LeScale arrives multiplied by LeNorm (media.cpp:295-306), le_norm = 1
    assert d.le_norm == 1.0 and np.array_equal(_arr(d.lescale, n), gd2.lescale.ravel() * np.float32(0.25))
    assert np.array_equal(_arr(d.sigma_a, NSPEC), gd2.sigma_a * np.float32(2.0))
    assert d.temperature_offset == 10.0 and d.temperature_scale == 1.5
    ref.lib.ref_adapter_free_medium(h, C.c_int32(kind))
    mo.close()


def test_adapter_describes_the_reference_nanovdbmedium(ref):
    dens = scenes.fbm_density(20, 24, 18, seed=5)
    dens[dens < 0.05] = 0
    nd, rd = scenes.cloud_scene_nanovdb(dens, voxel_size=1.0 / 24, width=16, height=12, spp=1, sigma_t=7.0, albedo=0.9, g=0.2)
    mo = ref.medium_nanovdb(nd)
    h, kind, desc = describe_medium(ref, mo)
    assert kind == 2
    d = C.cast(desc, C.POINTER(_abi.avr_nanovdb_desc)).contents
    want = nd.c
    g, wg = d.density, want.density
    assert list(g.index_min) == list(wg.index_min) and list(g.dims) == list(wg.dims)
    n = g.dims[0] * g.dims[1] * g.dims[2]
    assert np.array_equal(_arr(g.values, n), nd.density.values.ravel())     # the flattening of cmd/nanovdb2pbrt.cpp:100-110
    assert list(g.inv_mat) == list(wg.inv_mat) and list(g.vec) == list(wg.vec)
    assert list(g.world_bbox) == list(wg.world_bbox)
    assert not d.temperature.values
    assert list(d.render_from_medium) == list(want.render_from_medium)
    assert np.array_equal(_arr(d.sigma_s, NSPEC), nd.sigma_s * np.float32(want.sigma_scale))
    assert d.g == want.g and d.le_scale == want.le_scale
    assert list(d.majorant_res) == [64, 64, 64]                 # media.cpp:521
    ref.lib.ref_adapter_free_medium(h, C.c_int32(kind))
    mo.close()


def describe_render(ref, rd, width, height, fov, eye, sampler=0, randomize=2, seed=0, light_scale=1.0, filt=0):
    lib = ref.lib
    lib.ref_adapter_describe_render.restype = C.c_void_p
    wfc = scenes.look_at(eye, (0, 0, 0), (0, 1, 0)).astype(np.float32).ravel()
    aspect = width / height
    sw = np.array([-aspect, -1, aspect, 1] if aspect > 1 else [-1, -1 / aspect, 1, 1 / aspect], dtype=np.float32)
    cfg = np.array([width, height, rd.c.spp, sampler, randomize, seed, filt], dtype=np.int32)
    wi = np.array(list(rd.c.light_dir), dtype=np.float32)
    sphere = np.array(list(rd.c.boundary)[:4], dtype=np.float32)
    out_rgb = np.array(list(rd.c.output_rgb_from_sensor), dtype=np.float32)
    f32p, i32p = C.POINTER(C.c_float), C.POINTER(C.c_int32)
    desc = C.c_void_p()
    keep = (wfc, sw, cfg, wi, sphere, out_rgb)
    h = C.c_void_p(lib.ref_adapter_describe_render(wfc.ctypes.data_as(f32p), C.c_float(fov), sw.ctypes.data_as(f32p),
                                                   cfg.ctypes.data_as(i32p), wi.ctypes.data_as(f32p),
                                                   (rd.light_spectrum / np.float32(light_scale)).ctypes.data_as(f32p),
                                                   C.c_float(light_scale), sphere.ctypes.data_as(f32p),
                                                   rd.sensor_rgb.ctypes.data_as(f32p), C.c_float(rd.c.imaging_ratio),
                                                   out_rgb.ctypes.data_as(f32p), C.byref(desc)))
    if not desc.value:
        return h, None, keep
    return h, C.cast(desc, C.POINTER(_abi.avr_render_desc)).contents, keep


@pytest.mark.parametrize("sampler", [0, 1])
def test_adapter_describes_camera_light_sensor_sampler(ref, sampler):
    W, H, fov, eye = 40, 30, 20.0, (0.0, 0.35, -3.2)
    dens, (gd, rd) = _scene(width=W, height=H, fov=fov, eye=eye)
    h, d, keep = describe_render(ref, rd, W, H, fov, eye, sampler=sampler, randomize=3, seed=7, light_scale=3.0)
    want = rd.c
    assert (d.width, d.height, d.spp) == (W, H, want.spp)
    assert list(d.pixel_bounds) == [0, 0, W, H]
    # scenes.py composes the camera matrices in float64 and rounds once; pbrt composes them in fp32
    assert np.allclose(list(d.camera_from_raster), list(want.camera_from_raster), rtol=2e-6, atol=2e-7)
    assert np.allclose(list(d.render_from_camera), list(want.render_from_camera), rtol=2e-6, atol=2e-7)
    assert np.allclose(list(d.light_dir), list(want.light_dir), rtol=0, atol=1e-7)
    assert d.scene_radius == want.scene_radius                 # DistantLight::Preprocess, bit-equal
    assert d.boundary_type == 0 and list(d.boundary)[:4] == list(want.boundary)[:4]
    assert np.array_equal(_arr(d.light_spectrum, NSPEC), rd.light_spectrum)   # scale * Lemit
    assert np.array_equal(_arr(d.sensor_rgb, 3 * NSPEC), rd.sensor_rgb.ravel())
    assert d.imaging_ratio == want.imaging_ratio
    assert list(d.output_rgb_from_sensor) == list(want.output_rgb_from_sensor)
    assert math.isinf(d.max_component_value)
    assert list(d.filter_radius) == [0.5, 0.5]
    assert d.sampler == sampler and d.sampler_seed == 7
    if sampler == 1:
        assert d.randomize == 3                                # RandomizeStrategy::Owen
    ref.lib.ref_adapter_free_render(h)


def test_adapter_hands_over_a_nanovdb_grid_as_blocks(ref, tmp_path):
    """Describe(NanoVDBMedium, blocks = true): the in-memory FloatGrid as its leaf nodes (+ value tiles) instead of the
    flattened box — the same blocks the library's own .nvdb walker reads from the file of that grid, and, put back
    into the index box, the flattened grid bit for bit."""
    from acceleratedvolrenderer_b200.api import read_nvdb_blocks
    dens = scenes.fbm_density(44, 36, 40, seed=9)
    dens[dens < 0.15] = 0
    dens[8:24, 8:24, 16:32] = 0.75                  # whole leaves of one value: tiles if the builder prunes
    vs = 1.0 / 44
    nd, rd = scenes.cloud_scene_nanovdb(dens, voxel_size=vs, width=16, height=12, spp=1, sigma_t=7.0, albedo=0.9, g=0.2)
    mo = ref.medium_nanovdb(nd)
    lib = ref.lib
    lib.ref_adapter_describe_medium_blocks.restype = C.c_void_p
    desc = C.c_void_p()
    h = C.c_void_p(lib.ref_adapter_describe_medium_blocks(mo.h, C.byref(desc)))
    d = C.cast(desc, C.POINTER(_abi.avr_nanovdb_desc)).contents
    g = d.density
    assert not g.values and g.block_values and g.n_blocks > 20
    n = int(g.n_blocks)
    org = np.ctypeslib.as_array(g.block_origin, shape=(3 * n,)).reshape(n, 3).copy()
    vals = _arr(g.block_values, 512 * n).reshape(n, 8, 8, 8).copy()
    assert list(g.index_min) == list(nd.c.density.index_min) and list(g.dims) == list(nd.c.density.dims)
    path = tmp_path / "same.nvdb"
    ref.write_nvdb(path, [("density", dens, (0, 0, 0), vs, (0.0, 0.0, 0.0))])
    blk, _ = read_nvdb_blocks(path, "density")
    ka = np.lexsort(org.T[::-1]); kb = np.lexsort(blk["origins"].T[::-1])
    assert np.array_equal(org[ka], blk["origins"][kb])
    assert np.array_equal(vals[ka].view(np.uint32), blk["values"][kb].view(np.uint32))
    lo = np.asarray(list(g.index_min)); dims = np.asarray(list(g.dims))
    box = np.zeros(nd.density.values.shape, dtype=np.float32)
    for o, v in zip(org, vals):
        a = np.maximum(o, lo); b = np.minimum(o + 8, lo + dims)
        box[a[2] - lo[2]:b[2] - lo[2], a[1] - lo[1]:b[1] - lo[1], a[0] - lo[0]:b[0] - lo[0]] = \
            v[a[2] - o[2]:b[2] - o[2], a[1] - o[1]:b[1] - o[1], a[0] - o[0]:b[0] - o[0]]
    assert np.array_equal(box.view(np.uint32), nd.density.values.view(np.uint32))
    lib.ref_adapter_free_medium(h, C.c_int32(2))
    mo.close()


def test_adapter_describes_every_pixel_filter(ref):
    """pbrt's default pixel filter is "gaussian" (scene.cpp:94, filters.cpp:114): DescribeFilter hands the device the
    FilterSampler's own table (bit-equal to the one the probe tabulates from GaussianFilter::Evaluate) and the radius;
    "mitchell" the same; the box filter has no table; "triangle" (analytic tent sampling) is a kind of the descriptor
    without a table."""
    W, H, fov, eye = 40, 30, 20.0, (0.0, 0.35, -3.2)
    dens, (gd, rd) = _scene(width=W, height=H, fov=fov, eye=eye)
    h, d, keep = describe_render(ref, rd, W, H, fov, eye, filt=0)
    assert d is not None and not d.filter and list(d.filter_radius) == [0.5, 0.5]
    ref.lib.ref_adapter_free_render(h)
    for filt, cfg in ((1, (1, 1.5, 1.5, 0.5, 0.0)), (3, (2, 2.0, 2.0, 1.0 / 3.0, 1.0 / 3.0))):
        h, d, keep = describe_render(ref, rd, W, H, fov, eye, filt=filt)
        assert d is not None and d.filter
        t = d.filter.contents
        want = ref.filter_table(cfg)
        assert (t.ny, t.nx) == want.shape and list(d.filter_radius) == [cfg[1], cfg[2]]
        got = np.ctypeslib.as_array(t.f, shape=(t.ny * t.nx,)).reshape(t.ny, t.nx)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
        ref.lib.ref_adapter_free_render(h)
    h, d, keep = describe_render(ref, rd, W, H, fov, eye, filt=2)
    assert d is not None and d.filter and d.filter.contents.kind == _abi.FILTER_TRIANGLE
    assert list(d.filter_radius) == [2.0, 2.0]
    ref.lib.ref_adapter_free_render(h)


@pytest.mark.gpu
def test_device_builds_the_brick_pool_from_the_adapters_blocks(ctx, ref):
    """The block descriptor of the reference's in-memory NanoVDBMedium, handed to avr_medium_create_nanovdb as it is:
    the same brick pool as the one built from the flattened arrays (size, majorants, every lookup)."""
    from acceleratedvolrenderer_b200.api import AvrError, Medium
    dens = scenes.fbm_density(44, 36, 40, seed=9)
    dens[dens < 0.15] = 0
    nd, rd = scenes.cloud_scene_nanovdb(dens, voxel_size=1.0 / 44, width=16, height=12, spp=1, sigma_t=7.0, albedo=0.9, g=0.2)
    mo = ref.medium_nanovdb(nd)
    lib = ref.lib
    lib.ref_adapter_describe_medium_blocks.restype = C.c_void_p
    desc = C.c_void_p()
    h = C.c_void_p(lib.ref_adapter_describe_medium_blocks(mo.h, C.byref(desc)))
    mh = C.c_void_p()
    ctx.check(ctx.lib.avr_medium_create_nanovdb(ctx.h, desc, C.c_int(3), C.byref(mh)))
    via = Medium(ctx, mh, nd)
    direct = ctx.create_medium("nanovdb", nd, layout="brick16")
    assert via.bytes() == direct.bytes() and np.array_equal(via.majorants(), direct.majorants())
    rng = np.random.default_rng(4)
    xyz = rng.uniform(-2, np.array(dens.shape[::-1]) + 2, (30000, 3)).astype(np.float32)
    assert np.array_equal(via.lookup_density(xyz).view(np.uint32), direct.lookup_density(xyz).view(np.uint32))
    # blocks feed the brick pool only: any other layout is refused with a message
    with pytest.raises(AvrError) as e:
        ctx.check(ctx.lib.avr_medium_create_nanovdb(ctx.h, desc, C.c_int(1), C.byref(mh)))
    assert "brick16" in str(e.value)
    via.close(); direct.close()
    lib.ref_adapter_free_medium(h, C.c_int32(2))
    mo.close()


@pytest.mark.gpu
def test_device_renders_through_the_adapter(ctx, ref):
    """Medium and render descriptors as the adapter fills them from the reference's objects, handed to the C ABI as they
    are: the medium is the one scenes.py describes (majorants bit-equal), the film is bit-equal to the
    scenes.py render once both use the adapter's camera rays (the matrices differ by 2e-6 otherwise)."""
    from acceleratedvolrenderer_b200.api import Medium
    W, H, fov, eye = 40, 30, 20.0, (0.0, 0.35, -3.2)
    dens, (gd, rd) = _scene(width=W, height=H, fov=fov, eye=eye, spp=8)
    mo = ref.medium_uniformgrid(gd)
    hm, kind, desc = describe_medium(ref, mo)
    mh = C.c_void_p()
    ctx.check(ctx.lib.avr_medium_create_uniformgrid(ctx.h, desc, C.c_int(1), C.byref(mh)))
    via = Medium(ctx, mh, gd)
    direct = ctx.create_medium("uniformgrid", gd, layout="octet32")
    assert np.array_equal(via.majorants(), direct.majorants())
    hr, d, keep = describe_render(ref, rd, W, H, fov, eye, light_scale=3.0)
    d.max_depth = rd.c.max_depth
    film = np.zeros((H, W, 4)); stats = _abi.avr_render_stats()
    ctx.check(ctx.lib.avr_render_volpath(ctx.h, via.h, C.byref(d), film.ctypes.data_as(C.POINTER(C.c_double)), C.byref(stats)))
    want, _ = ctx.render_volpath(direct, rd)
    assert np.array_equal(film[..., 3], want[..., 3]) and film[..., 3].min() == 8
    # The camera matrices and the light direction differ in the last bits (scenes.py composes them in float64, pbrt in
    # fp32), and delta tracking amplifies a last-bit change of a ray into a different path wherever a comparison
    # flips, so per-pixel agreement of 8 samples is not a meaningful bar.  What IS exact: give the scenes.py
    # descriptor the adapter's rays and every other field must already be equivalent -- the films are bit-equal.
    for name in ("camera_from_raster", "render_from_camera", "light_dir"):
        src, dst = getattr(d, name), getattr(rd.c, name)
        for i in range(len(src)):
            dst[i] = src[i]
    same_rays, _ = ctx.render_volpath(direct, rd)
    assert np.array_equal(film, same_rays)
    # and the two renders of the unmodified descriptors are the same image statistically
    a, b = film[..., :3].sum(axis=2), want[..., :3].sum(axis=2)
    assert abs(a.mean() / b.mean() - 1) < 0.02, (a.mean(), b.mean())
    # pbrt's default pixel filter through the adapter: the FilterSampler table it hands over gives the film the
    # scenes.py descriptor gives with the same table and radius
    hg, dg, keepg = describe_render(ref, rd, W, H, fov, eye, light_scale=3.0, filt=1)
    dg.max_depth = rd.c.max_depth
    for name in ("camera_from_raster", "render_from_camera", "light_dir"):
        src, dst = getattr(d, name), getattr(dg, name)
        for i in range(len(src)):
            dst[i] = src[i]
    gfilm = np.zeros((H, W, 4)); gstats = _abi.avr_render_stats()
    ctx.check(ctx.lib.avr_render_volpath(ctx.h, via.h, C.byref(dg), gfilm.ctypes.data_as(C.POINTER(C.c_double)), C.byref(gstats)))
    want_g, _ = ctx.render_volpath(direct, rd.with_filter(ref.filter_table((1, 1.5, 1.5, 0.5, 0.0)), (1.5, 1.5)))
    assert np.array_equal(gfilm, want_g) and not np.array_equal(gfilm, film)
    assert abs(gfilm[..., 3].mean() / 8 - 1) < 0.1 and gfilm[..., 3].min() > 0
    ref.lib.ref_adapter_free_render(hg)
    ref.lib.ref_adapter_free_render(hr)
    ref.lib.ref_adapter_free_medium(hm, C.c_int32(kind))
    via.close(); direct.close(); mo.close()
