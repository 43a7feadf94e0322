"""Synthetic scene authoring for tests and bench (host side, numpy only; no compute path here).

The reference ships no scene files, so every parity / bench scene is authored here:
procedural density grids, the pbrt-style perspective camera matrices (cameras.cpp:404-460,
cameras.h CameraBase / ProjectiveCamera ctor), a distant light and sensor tables.
"""
import math

import numpy as np

from ._abi import NSPEC, GridDesc, RenderDesc, RGBGridDesc

LAMBDA = np.arange(360, 831, dtype=np.float64)


def const_spectrum(v):
    return np.full(NSPEC, v, dtype=np.float32)


def ramp_spectrum(v360, v830):
    return np.linspace(v360, v830, NSPEC).astype(np.float32)


def cie_xyz_tables():
    """CIE 1931 x̄,ȳ,z̄ via the multi-lobe Gaussian fit of Wyman, Sloan & Shirley (JCGT 2013).  Stands in
    for the reference's tabulated CIE curves (the in-pbrt adapter passes the real PixelSensor tables
    through avr_render_desc.sensor_rgb)."""
    def g(x, mu, s1, s2):
        s = np.where(x < mu, s1, s2)
        return np.exp(-0.5 * ((x - mu) / s) ** 2)
    x = LAMBDA
    X = 1.056 * g(x, 599.8, 37.9, 31.0) + 0.362 * g(x, 442.0, 16.0, 26.7) - 0.065 * g(x, 501.1, 20.4, 26.2)
    Y = 0.821 * g(x, 568.8, 46.9, 40.5) + 0.286 * g(x, 530.9, 16.3, 31.1)
    Z = 1.217 * g(x, 437.0, 11.8, 36.0) + 0.681 * g(x, 459.0, 26.0, 13.8)
    return np.stack([X, Y, Z]).astype(np.float32)


# XYZ -> linear sRGB (D65)
SRGB_FROM_XYZ = np.array([[3.2404542, -1.5371385, -0.4985314],
                          [-0.9692660, 1.8760108, 0.0415560],
                          [0.0556434, -0.2040259, 1.0572252]], dtype=np.float32)
CIE_Y_INTEGRAL = 106.856895


def fbm_density(nx, ny=None, nz=None, seed=7, octaves=5, frequency=3.0, threshold=0.3, sharp=3.0):
    """Cloud-like procedural density in [0,1]: trilinear value-noise fBm, thresholded so that a good
    part of the volume is empty (what makes the majorant grid pay off).  Returns [nz, ny, nx] fp32."""
    ny = ny or nx
    nz = nz or nx
    rng = np.random.default_rng(seed)
    zs = (np.arange(nz, dtype=np.float32) + 0.5) / nz
    ys = (np.arange(ny, dtype=np.float32) + 0.5) / ny
    xs = (np.arange(nx, dtype=np.float32) + 0.5) / nx
    out = np.zeros((nz, ny, nx), dtype=np.float32)
    amp, freq, norm = 1.0, frequency, 0.0
    for _ in range(octaves):
        n = int(math.ceil(freq)) + 2
        lat = rng.random((n, n, n), dtype=np.float32)

        def axis(c):
            f = c * freq
            i = np.floor(f).astype(np.int64)
            t = (f - i).astype(np.float32)
            t = t * t * (3 - 2 * t)
            return i, t
        iz, tz = axis(zs); iy, ty = axis(ys); ix, tx = axis(xs)
        a = lat[iz][:, iy][:, :, ix]
        b = lat[iz][:, iy][:, :, ix + 1]
        c = lat[iz][:, iy + 1][:, :, ix]
        d = lat[iz][:, iy + 1][:, :, ix + 1]
        e = lat[iz + 1][:, iy][:, :, ix]
        f = lat[iz + 1][:, iy][:, :, ix + 1]
        g = lat[iz + 1][:, iy + 1][:, :, ix]
        h = lat[iz + 1][:, iy + 1][:, :, ix + 1]
        TX = tx[None, None, :]; TY = ty[None, :, None]; TZ = tz[:, None, None]
        lo = (a * (1 - TX) + b * TX) * (1 - TY) + (c * (1 - TX) + d * TX) * TY
        hi = (e * (1 - TX) + f * TX) * (1 - TY) + (g * (1 - TX) + h * TX) * TY
        out += amp * (lo * (1 - TZ) + hi * TZ)
        norm += amp
        amp *= 0.5
        freq *= 2.0
    out /= norm
    # radial falloff so the cloud sits inside the box
    Z, Y, X = np.meshgrid(zs - 0.5, ys - 0.5, xs - 0.5, indexing="ij")
    r = np.sqrt(X * X + Y * Y + Z * Z) * 2.0
    out = out * np.clip(1.25 - r, 0, 1)
    out = np.clip((out - threshold * 0.6) * sharp, 0, 1).astype(np.float32)
    return out


def look_at(eye, look, up):
    """pbrt LookAt (util/transform.cpp): returns worldFromCamera 4x4 (float64)."""
    eye = np.asarray(eye, dtype=np.float64); look = np.asarray(look, dtype=np.float64)
    up = np.asarray(up, dtype=np.float64)
    d = look - eye
    d /= np.linalg.norm(d)
    right = np.cross(up / np.linalg.norm(up), d)
    right /= np.linalg.norm(right)
    newup = np.cross(d, right)
    m = np.eye(4)
    m[:3, 0] = right; m[:3, 1] = newup; m[:3, 2] = d; m[:3, 3] = eye
    return m


def perspective_camera(width, height, fov_deg, eye, look, up=(0, 1, 0)):
    """(camera_from_raster, render_from_camera) as float32 row-major 4x4, built the way
    ProjectiveCamera's ctor composes them (cameras.h: screenFromCamera = Perspective(fov, 1e-2, 1000),
    NDCFromScreen, rasterFromNDC); fov spans the shorter image axis.  Render space = world space."""
    n, f = 1e-2, 1000.0
    persp = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, f / (f - n), -f * n / (f - n)], [0, 0, 1, 0]], dtype=np.float64)
    inv_tan = 1.0 / math.tan(math.radians(fov_deg) / 2)
    screen_from_camera = np.diag([inv_tan, inv_tan, 1, 1]) @ persp
    aspect = width / height
    if aspect > 1:
        sw = (-aspect, -1.0, aspect, 1.0)
    else:
        sw = (-1.0, -1 / aspect, 1.0, 1 / aspect)
    trans = np.eye(4); trans[0, 3] = -sw[0]; trans[1, 3] = -sw[3]
    ndc_from_screen = np.diag([1 / (sw[2] - sw[0]), 1 / (sw[3] - sw[1]), 1, 1]) @ trans
    raster_from_ndc = np.diag([width, -height, 1, 1])
    raster_from_screen = raster_from_ndc @ ndc_from_screen
    camera_from_raster = np.linalg.inv(screen_from_camera) @ np.linalg.inv(raster_from_screen)
    render_from_camera = look_at(eye, look, up)
    return camera_from_raster.astype(np.float32), render_from_camera.astype(np.float32)


def scale_translate(scale, translate):
    m = np.eye(4, dtype=np.float64)
    for i in range(3):
        m[i, i] = scale[i]
        m[i, 3] = translate[i]
    return m.astype(np.float32)


def distant_light_scene_radius(boundary_type, boundary):
    """DistantLight::Preprocess (lights.h:269-271): sceneBounds.BoundingSphere(&c, &sceneRadius) over the
    aggregate's bounds, in fp32 as the reference computes it.  The bounds of a Shape "sphere" are the box
    c +- r (Sphere::Bounds through Transform::operator()(Bounds3f)), so sceneRadius = |(r, r, r)|, not r."""
    f = np.float32
    b = [f(v) for v in boundary]
    if boundary_type == 0:
        lo = [f(b[i] - b[3]) for i in range(3)]
        hi = [f(b[i] + b[3]) for i in range(3)]
    else:
        lo, hi = b[:3], b[3:6]
    ctr = [f(f(lo[i] + hi[i]) / f(2)) for i in range(3)]
    d = [f(hi[i] - ctr[i]) for i in range(3)]
    return float(np.sqrt(f(f(f(d[0] * d[0]) + f(d[1] * d[1])) + f(d[2] * d[2]))))


def cloud_scene(density, *, width, height, spp, sigma_t=8.0, albedo=0.99, g=0.0, max_depth=5, fov=20.0,
                eye=(0.0, 0.35, -3.2), light_dir=(0.4, 0.8, -0.45), light_scale=3.0, boundary="sphere",
                majorant_res=(16, 16, 16), exact_math=1, sampler_seed=0, chromatic=False):
    """A volume-only scene of the class the fork's graph pipeline accepts (SURVEY.md App. A gotchas):
    `uniformgrid` in medium bounds [0,1]^3 mapped to render-space [-.5,.5]^3 (x extent scaled by the
    grid aspect), an `interface` boundary around it, one distant light, perspective camera."""
    nz, ny, nx = density.shape
    m = max(nx, ny, nz)
    ext = (nx / m, ny / m, nz / m)
    rfm = scale_translate(ext, (-ext[0] / 2, -ext[1] / 2, -ext[2] / 2))
    if chromatic:
        sig_s = ramp_spectrum(0.8, 1.2) * (sigma_t * albedo)
        sig_a = ramp_spectrum(1.5, 0.5) * (sigma_t * (1 - albedo))
    else:
        sig_s = const_spectrum(sigma_t * albedo)
        sig_a = const_spectrum(sigma_t * (1 - albedo))
    gd = GridDesc(density, sig_a, sig_s, render_from_medium=rfm, g=g, majorant_res=majorant_res)
    cfr, rfc = perspective_camera(width, height, fov, eye, (0, 0, 0))
    half_diag = 0.5 * math.sqrt(sum(e * e for e in ext))
    if boundary == "sphere":
        btype, bvals = 0, (0.0, 0.0, 0.0, half_diag * 1.001)
    else:
        btype, bvals = 1, (-ext[0] / 2, -ext[1] / 2, -ext[2] / 2, ext[0] / 2, ext[1] / 2, ext[2] / 2)
    scene_radius = distant_light_scene_radius(btype, bvals)
    xyz = cie_xyz_tables()
    rd = RenderDesc(width, height, spp, camera_from_raster=cfr, render_from_camera=rfc, light_dir=light_dir,
                    light_spectrum=const_spectrum(light_scale), scene_radius=scene_radius,
                    boundary_type=btype, boundary=bvals, sensor_rgb=xyz, max_depth=max_depth,
                    sampler_seed=sampler_seed, exact_math=exact_math, imaging_ratio=1.0 / CIE_Y_INTEGRAL * 100,
                    output_rgb_from_sensor=SRGB_FROM_XYZ.ravel())
    return gd, rd


def cloud_scene_nanovdb(density, *, temperature=None, voxel_size=1.0 / 64, le_scale=1.0, sigma_t=8.0, albedo=0.99,
                        g=0.0, **kw):
    """The same scene class with the volume as a "nanovdb" medium (what pbrt-v4-scenes' Disney cloud is):
    a FloatGrid with uniform voxel size whose world box is centred on the render-space origin."""
    from ._abi import NanoVDBDesc, NanoVDBGrid
    dg = NanoVDBGrid.from_dense(density, (0, 0, 0), voxel_size, (0.0, 0.0, 0.0))
    tg = None if temperature is None else NanoVDBGrid.from_dense(temperature, (0, 0, 0), voxel_size, (0.0, 0.0, 0.0))
    wb = dg.world_bbox
    ctr = (wb[:3] + wb[3:]) / 2
    ext = wb[3:] - wb[:3]
    rfm = scale_translate((1, 1, 1), tuple(-ctr))
    nd = NanoVDBDesc(dg, const_spectrum(sigma_t * (1 - albedo)), const_spectrum(sigma_t * albedo), temperature=tg,
                     render_from_medium=rfm, g=g, le_scale=le_scale)
    # camera / light / film exactly as cloud_scene builds them for a grid of the same extent
    m = float(ext.max())
    fake = np.zeros((max(1, int(round(ext[2] / m * 8))), max(1, int(round(ext[1] / m * 8))), max(1, int(round(ext[0] / m * 8)))),
                    dtype=np.float32)
    _, rd = cloud_scene(fake, sigma_t=sigma_t, albedo=albedo, g=g, **kw)
    half_diag = 0.5 * float(np.linalg.norm(ext))
    if rd.c.boundary_type == 0:
        for i, v in enumerate((0.0, 0.0, 0.0, half_diag * 1.001)):
            rd.c.boundary[i] = v
    else:
        for i, v in enumerate(tuple(-ext / 2) + tuple(ext / 2)):
            rd.c.boundary[i] = float(v)
    rd.c.scene_radius = distant_light_scene_radius(rd.c.boundary_type, list(rd.c.boundary))
    return nd, rd


def rgb_voxels(rgb_like, rng):
    """Synthetic {scale,c0,c1,c2} sigmoid-polynomial voxels (what RGBUnboundedSpectrum(cs, rgb) yields,
    util/spectrum.cpp:236-247) for an [nz,ny,nx] scale field: smooth, bounded coefficient ranges."""
    scale = np.asarray(rgb_like, dtype=np.float32)
    shp = scale.shape
    c0 = rng.uniform(-2e-5, 2e-5, shp).astype(np.float32)
    c1 = rng.uniform(-2e-2, 2e-2, shp).astype(np.float32)
    c2 = rng.uniform(-4, 4, shp).astype(np.float32)
    return np.stack([scale, c0, c1, c2], axis=-1).astype(np.float32)


def random_graph(n, avg_deg=4, seed=3, albedo=0.95):
    """Random geometric-ish transport graph in CSR (rows sub-stochastic: sum <= albedo), the shape
    LightingCalculator::GetTransportMatrix produces (lighting_calculator.cpp:61-82)."""
    rng = np.random.default_rng(seed)
    deg = rng.poisson(avg_deg, n).astype(np.int64)
    deg = np.clip(deg, 0, max(1, n - 1))
    rowptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(deg, out=rowptr[1:])
    nnz = int(rowptr[-1])
    rows = np.repeat(np.arange(n, dtype=np.int64), deg)
    # neighbours close in index (vertices created along light paths are spatially coherent)
    off = np.rint(rng.normal(0, max(2.0, n ** (1 / 3) * 4), nnz)).astype(np.int64)
    col = np.clip(rows + off, 0, n - 1)
    key = rows * n + col
    key = np.unique(key)
    rows = (key // n); col = (key % n).astype(np.int32)
    counts = np.bincount(rows, minlength=n)
    rowptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(counts, out=rowptr[1:])
    w = rng.random(col.size).astype(np.float32) + 0.05
    rowsum = np.bincount(rows, weights=w, minlength=n).astype(np.float32)
    val = (w / np.maximum(rowsum[rows], 1e-20) * albedo).astype(np.float32)
    return rowptr, col, val


def random_edge_graph(n, n_edges, seed=3):
    """A FreeGraph-like record set (what FreeGraphBuilder leaves behind, src/graph/free/free_graph_builder.cpp):
    vertices with positions / sample counts, an edge list with repeated (from, to) pairs, self loops and a few
    edges to vertices that do not exist (Graph::AddEdge merges the first and drops the last), paths."""
    rng = np.random.default_rng(seed)
    pos = rng.normal(size=(n, 3)).astype(np.float32)
    efrom = rng.integers(0, n, n_edges).astype(np.int32)
    eto = np.clip(efrom + np.rint(rng.normal(0, 3, n_edges)).astype(np.int64), -2, n + 1).astype(np.int32)
    dup = rng.integers(0, n_edges, n_edges // 5)          # repeat a fifth of the pairs
    efrom = np.r_[efrom, efrom[dup]].astype(np.int32); eto = np.r_[eto, eto[dup]].astype(np.int32)
    esamples = rng.integers(1, 6, efrom.size).astype(np.int32)
    vsamples = (np.bincount(efrom, weights=esamples, minlength=n)[:n] + rng.integers(1, 4, n)).astype(np.int32)
    plen = rng.integers(1, 6, max(1, n // 10))
    poff = np.r_[0, np.cumsum(plen)].astype(np.int64)
    pverts = rng.integers(0, n, int(poff[-1])).astype(np.int32)
    return dict(pos=pos, vsamples=vsamples, efrom=efrom, eto=eto, esamples=esamples, poff=poff, pverts=pverts,
                vlight=rng.random(n).astype(np.float32) * 3, vrange=(rng.random(n).astype(np.float32) + 0.01) * 0.2,
                vtype=rng.integers(0, 5, n).astype(np.int32), vcoors=rng.integers(-50, 50, (n, 3)).astype(np.int32))
