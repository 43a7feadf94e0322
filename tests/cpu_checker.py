"""ctypes bindings for the CPU checkers (TEST INFRASTRUCTURE ONLY).

`Checker("orc")` loads oracle/liboracle.so (the CPU restatement, always available after
`make -C oracle`); `Checker("ref")` loads oracle/_ref/libavr_refprobe.so (the reference's own
templates compiled from /root/reference; present in the build container and shipped to the GPU box
as a built artefact).  Both export the same flat signatures (orc_* / ref_*).
"""
import ctypes as C
import os

import numpy as np

from acceleratedvolrenderer_b200._abi import GridDesc, NanoVDBDesc, NanoVDBGrid, RGBGridDesc, RenderDesc, TMAJ_DTYPE

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORC_PATH = os.path.join(ROOT, "oracle", "liboracle.so")
REF_PATH = os.path.join(ROOT, "oracle", "_ref", "libavr_refprobe.so")

f32p = C.POINTER(C.c_float)
i32p = C.POINTER(C.c_int32)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)
i64p = C.POINTER(C.c_int64)
f64p = C.POINTER(C.c_double)


def _p(a, t):
    return a.ctypes.data_as(t)


def f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def have_ref():
    return os.path.exists(REF_PATH)


class Checker:
    def __init__(self, kind="orc"):
        self.kind = kind
        path = ORC_PATH if kind == "orc" else REF_PATH
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `make -C oracle`")
        self.lib = C.CDLL(path)
        self.pfx = kind + "_"
        for name in ("medium_create_uniformgrid", "medium_create_rgbgrid", "medium_create_nanovdb"):
            getattr(self.lib, self.pfx + name).restype = C.c_void_p
        if kind == "ref":
            self.lib.ref_nanovdb_grid_build.restype = C.c_void_p
        self.fn("kind").restype = C.c_char_p

    def fn(self, name):
        return getattr(self.lib, self.pfx + name)

    def has(self, name):
        return hasattr(self.lib, self.pfx + name)

    # ---- primitives ------------------------------------------------------------------------------
    def fast_exp(self, x):
        x = f32(x); out = np.empty_like(x)
        self.fn("fast_exp")(_p(x, f32p), C.c_int64(x.size), _p(out, f32p))
        return out

    def libm(self, kind, x):
        x = f32(x); out = np.empty_like(x)
        self.fn("libm")(C.c_int(kind), _p(x, f32p), C.c_int64(x.size), _p(out, f32p))
        return out

    def mix_bits(self, x):
        x = np.ascontiguousarray(x, dtype=np.uint64); out = np.empty_like(x)
        self.fn("mix_bits")(_p(x, u64p), C.c_int64(x.size), _p(out, u64p))
        return out

    def zsobol_samples(self, spp, res, randomize, seed, samples, rounds):
        """ZSobolSampler streams: per {px, py, sampleIndex}, `rounds` x {Get1D, Get2D} -> (n, rounds, 3) floats."""
        samples = np.ascontiguousarray(samples, dtype=np.int32)
        n = samples.shape[0]
        out = np.zeros((n, rounds, 3), dtype=np.float32)
        self.fn("zsobol_samples")(C.c_int32(spp), C.c_int32(res[0]), C.c_int32(res[1]), C.c_int32(randomize),
                                  C.c_int32(seed), _p(samples, i32p), C.c_int64(n), C.c_int32(rounds), _p(out, f32p))
        return out

    def sobol_matrices(self):
        out = np.zeros(104, dtype=np.uint32)
        self.fn("sobol_matrices")(out.ctypes.data_as(C.c_void_p))
        return out

    def hash_float(self, x):
        x = f32(x); out = np.empty(x.size, dtype=np.uint64)
        self.fn("hash_float")(_p(x, f32p), C.c_int64(x.size), _p(out, u64p))
        return out

    def hash_point3(self, x):
        x = f32(x); n = x.size // 3; out = np.empty(n, dtype=np.uint64)
        self.fn("hash_point3")(_p(x, f32p), C.c_int64(n), _p(out, u64p))
        return out

    def rng_stream(self, seq, seed, use_seed, adv, n):
        u = np.empty(n, dtype=np.uint32); f = np.empty(n, dtype=np.float32)
        self.fn("rng_stream")(C.c_uint64(seq), C.c_uint64(seed), C.c_int(use_seed), C.c_int64(adv), C.c_int64(n),
                              _p(u, u32p), _p(f, f32p))
        return u, f

    def independent_sampler(self, px, py, seed, sample_index, dim, n):
        out = np.empty(n, dtype=np.float32)
        self.fn("independent_sampler")(C.c_int(px), C.c_int(py), C.c_int(seed), C.c_int(sample_index), C.c_int(dim),
                                       C.c_int64(n), _p(out, f32p))
        return out

    def sample_discrete3(self, w, u):
        w = f32(w); u = f32(u); out = np.empty(u.size, dtype=np.int32)
        self.fn("sample_discrete3")(_p(w, f32p), _p(u, f32p), C.c_int64(u.size), _p(out, i32p))
        return out

    def sample_exponential(self, u, a):
        u = f32(u); a = f32(a); out = np.empty_like(u)
        self.fn("sample_exponential")(_p(u, f32p), _p(a, f32p), C.c_int64(u.size), _p(out, f32p))
        return out

    def hg_sample(self, wo, g, u):
        wo = f32(wo); u = f32(u); n = wo.size // 3
        wi = np.empty((n, 3), dtype=np.float32); pdf = np.empty(n, dtype=np.float32)
        self.fn("hg_sample")(_p(wo, f32p), C.c_float(g), _p(u, f32p), C.c_int64(n), _p(wi, f32p), _p(pdf, f32p))
        return wi, pdf

    def hg_p(self, cos_theta, g):
        c = f32(cos_theta); out = np.empty_like(c)
        self.fn("hg_p")(_p(c, f32p), C.c_float(g), C.c_int64(c.size), _p(out, f32p))
        return out

    def sample_wavelengths(self, u):
        u = f32(u); out = np.empty((u.size, 8), dtype=np.float32)
        self.fn("sample_wavelengths")(_p(u, f32p), C.c_int64(u.size), _p(out, f32p))
        return out

    def blackbody(self, T, lam):
        T = f32(T); lam = f32(lam); out = np.empty((T.size, 4), dtype=np.float32)
        self.fn("blackbody")(_p(T, f32p), C.c_int64(T.size), _p(lam, f32p), _p(out, f32p))
        return out

    def grid_lookup(self, grid, p):
        grid = f32(grid); p = f32(p); nz, ny, nx = grid.shape; n = p.size // 3
        out = np.empty(n, dtype=np.float32)
        self.fn("grid_lookup")(_p(grid, f32p), C.c_int(nx), C.c_int(ny), C.c_int(nz), _p(p, f32p), C.c_int64(n),
                               _p(out, f32p))
        return out

    def dda(self, bounds, res, vox, rays, sigma_t=1.0, max_per_ray=200):
        bounds = f32(bounds); res = np.ascontiguousarray(res, dtype=np.int32); vox = f32(vox); rays = f32(rays)
        n = rays.size // 8
        counts = np.zeros(n, dtype=np.int32)
        seg_t = np.zeros((n, max_per_ray, 2), dtype=np.float32)
        seg_s = np.zeros((n, max_per_ray), dtype=np.float32)
        seg_v = np.full((n, max_per_ray), -1, dtype=np.int32)
        self.fn("dda")(_p(bounds, f32p), _p(res, i32p), _p(vox, f32p), _p(rays, f32p), C.c_int64(n),
                       C.c_float(sigma_t), C.c_int32(max_per_ray), _p(counts, i32p), _p(seg_t, f32p),
                       _p(seg_s, f32p), _p(seg_v, i32p))
        return counts, seg_t, seg_s, seg_v

    def apply_inverse_ray(self, m, minv, rays7):
        return self._xf_ray("apply_inverse_ray", m, minv, rays7)

    def apply_ray(self, m, minv, rays7):
        return self._xf_ray("apply_ray", m, minv, rays7)

    def _xf_ray(self, name, m, minv, rays7):
        m = f32(m); minv = f32(minv); rays7 = f32(rays7); n = rays7.size // 7
        out = np.empty((n, 7), dtype=np.float32)
        self.fn(name)(_p(m, f32p), _p(minv, f32p), _p(rays7, f32p), C.c_int64(n), _p(out, f32p))
        return out

    # ---- media -----------------------------------------------------------------------------------
    def medium_uniformgrid(self, desc: GridDesc):
        h = self.fn("medium_create_uniformgrid")(C.byref(desc.c))
        return CheckerMedium(self, h, desc)

    def medium_rgbgrid(self, desc: RGBGridDesc):
        h = self.fn("medium_create_rgbgrid")(C.byref(desc.c))
        return CheckerMedium(self, h, desc)

    # NanoVDB: under "ref" the grid is a real nanovdb::FloatGrid built with the vendored GridBuilder and
    # read back (index box, float inverse map, world box, flattened values); under "orc" the same
    # numbers come from the host-side helper NanoVDBGrid.from_dense.  The golden comparison of the
    # returned arrays pins the helper to NanoVDB itself.
    def _ref_grid_handle(self, source):
        dense, oijk, vs, ow = source
        dense = f32(dense)
        dims = np.asarray(dense.shape[::-1], dtype=np.int32); oijk = np.asarray(oijk, dtype=np.int32)
        ow = np.asarray(ow, dtype=np.float64)
        return C.c_void_p(self.lib.ref_nanovdb_grid_build(_p(dense, f32p), _p(dims, i32p), _p(oijk, i32p), C.c_double(vs),
                                                          _p(ow, f64p)))

    def nanovdb_grid(self, dense, origin_ijk, voxel_size, origin_world) -> NanoVDBGrid:
        py = NanoVDBGrid.from_dense(dense, origin_ijk, voxel_size, origin_world)
        if self.kind != "ref":
            return py
        h = self._ref_grid_handle(py.source)
        imin = np.zeros(3, dtype=np.int32); dims = np.zeros(3, dtype=np.int32)
        inv = np.zeros(9, dtype=np.float32); vec = np.zeros(3, dtype=np.float32); wb = np.zeros(6, dtype=np.float64)
        self.lib.ref_nanovdb_grid_info(h, _p(imin, i32p), _p(dims, i32p), _p(inv, f32p), _p(vec, f32p), _p(wb, f64p))
        vals = np.empty((dims[2], dims[1], dims[0]), dtype=np.float32)
        self.lib.ref_nanovdb_grid_flatten(h, _p(vals, f32p))
        self.lib.ref_nanovdb_grid_free(h)
        return NanoVDBGrid(vals, imin, inv, vec, wb, source=py.source)

    def write_nvdb(self, path, grids, codec=0):
        """"ref" only: writes a .nvdb file with the reference's vendored nanovdb::io::writeGrid (one segment per grid).
        grids: list of (name, dense [nz, ny, nx], origin_ijk, voxel_size, origin_world)."""
        assert self.kind == "ref"
        self.lib.ref_nanovdb_grid_build_named.restype = C.c_void_p
        hs = []
        for name, dense, origin_ijk, voxel_size, origin_world in grids:
            d = f32(dense)
            dims = np.array(d.shape[::-1], dtype=np.int32); org = np.array(origin_ijk, dtype=np.int32)
            ow = np.array(origin_world, dtype=np.float64)
            hs.append(C.c_void_p(self.lib.ref_nanovdb_grid_build_named(_p(d, f32p), _p(dims, i32p), _p(org, i32p),
                                                                       C.c_double(voxel_size), _p(ow, f64p), name.encode())))
        arr = (C.c_void_p * len(hs))(*hs)
        rc = self.lib.ref_nanovdb_write_file(str(path).encode(), arr, C.c_int32(len(hs)), C.c_int32(codec))
        for h in hs:
            self.lib.ref_nanovdb_grid_free(h)
        assert rc == 0, f"ref_nanovdb_write_file failed ({rc})"

    def medium_nanovdb(self, desc: NanoVDBDesc):
        if self.kind == "ref":
            dh = self._ref_grid_handle(desc.density.source)
            th = self._ref_grid_handle(desc.temperature.source) if desc.temperature is not None else None
            h = self.lib.ref_medium_create_nanovdb(C.byref(desc.c), dh, th)
            self.lib.ref_nanovdb_grid_free(dh)
            if th is not None:
                self.lib.ref_nanovdb_grid_free(th)
        else:
            h = self.fn("medium_create_nanovdb")(C.byref(desc.c))
        return CheckerMedium(self, h, desc)

    # ---- graph -----------------------------------------------------------------------------------
    def graph_propagate(self, rowptr, col, val, L0, bounces):
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int64); col = np.ascontiguousarray(col, dtype=np.int32)
        val = f32(val); L0 = f32(L0); n = L0.size
        total = np.empty(n, dtype=np.float32)
        self.fn("graph_propagate").restype = C.c_int
        it = self.fn("graph_propagate")(C.c_int32(n), C.c_int64(col.size), _p(rowptr, i64p), _p(col, i32p),
                                        _p(val, f32p), _p(L0, f32p), C.c_int32(bounces), _p(total, f32p))
        return total, it

    def graph_build(self, vertex_radius, vid, pos, *, vtype=None, vlight=None, vsamples=None, vrange=None, vcoors=None,
                    eid=(), efrom=(), eto=(), esamples=None, pid=(), poff=(0,), pverts=()):
        i32 = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.int32)  # noqa: E731
        opt = lambda a, t: None if a is None else _p(a, t)  # noqa: E731
        vid = i32(vid); pos = f32(pos); vtype = i32(vtype); vsamples = i32(vsamples); vcoors = i32(vcoors)
        vlight = None if vlight is None else f32(vlight); vrange = None if vrange is None else f32(vrange)
        eid = i32(eid); efrom = i32(efrom); eto = i32(eto); esamples = i32(esamples)
        pid = i32(pid); poff = np.ascontiguousarray(poff, dtype=np.int64); pverts = i32(pverts)
        fn = self.fn("graph_build"); fn.restype = C.c_void_p
        h = fn(C.c_float(vertex_radius), C.c_int32(vid.size), _p(vid, i32p), _p(pos, f32p), opt(vtype, i32p),
               opt(vlight, f32p), opt(vsamples, i32p), opt(vrange, f32p), opt(vcoors, i32p), C.c_int32(eid.size),
               _p(eid, i32p), _p(efrom, i32p), _p(eto, i32p), opt(esamples, i32p), C.c_int32(pid.size), _p(pid, i32p),
               _p(poff, i64p), _p(pverts, i32p))
        return CheckerGraph(self, h)

    def graph_read(self, path):
        fn = self.fn("graph_read"); fn.restype = C.c_void_p
        return CheckerGraph(self, fn(path.encode()))

    def graph_connect(self, pos, light, search_range, vertex_radius, points):
        pos = f32(pos); light = f32(light); points = f32(points); n = light.size; npts = points.size // 3
        rng = None if search_range is None else f32(search_range)
        out = np.empty(npts, dtype=np.float32)
        self.fn("graph_connect")(C.c_int32(n), _p(pos, f32p), _p(light, f32p),
                                 _p(rng, f32p) if rng is not None else None, C.c_float(vertex_radius),
                                 _p(points, f32p), C.c_int64(npts), _p(out, f32p))
        return out


    def camera_weights(self, rdesc):
        """CameraSample::filterWeight of every pixel sample of the descriptor's ranges, pixel-major (oracle only)."""
        x0, y0, x1, y1, s0, s1 = rdesc.ranges()
        out = np.zeros((x1 - x0) * (y1 - y0) * (s1 - s0), dtype=np.float32)
        self.lib.orc_camera_weights(C.byref(rdesc.c), _p(out, f32p))
        return out

    def film_add_samples(self, rdesc, pixels, vals, weights=None):
        """RGBFilm::AddSample over explicit samples (pixels {px, py}; vals {L[4], lambda[4], pdf[4]}), in order;
        `weights` (oracle only): the filter weight of each sample, 1 when None.
        Returns the film sums (H, W, 4) doubles and RGBFilm::GetPixelRGB (H, W, 3)."""
        pixels = np.ascontiguousarray(pixels, dtype=np.int32); vals = f32(vals); n = pixels.shape[0]
        W, H = rdesc.c.width, rdesc.c.height
        film = np.zeros((H, W, 4), dtype=np.float64)
        if weights is not None:
            assert self.kind == "orc"
            weights = f32(weights)
            self.lib.orc_film_add_samples_w(C.byref(rdesc.c), _p(pixels, i32p), _p(vals, f32p), _p(weights, f32p),
                                            C.c_int64(n), _p(film, f64p))
            rgb = np.zeros((H, W, 3), dtype=np.float32)
            self.lib.orc_film_resolve(C.byref(rdesc.c), _p(film, f64p), _p(rgb, f32p))
            return film, rgb
        if self.kind == "ref":
            rgb = np.zeros((H, W, 3), dtype=np.float32)
            M = np.array(list(rdesc.c.output_rgb_from_sensor), dtype=np.float32)
            self.lib.ref_film_add_samples(C.c_int32(W), C.c_int32(H), _p(rdesc.sensor_rgb, f32p), C.c_float(rdesc.c.imaging_ratio),
                                          C.c_float(rdesc.c.max_component_value), _p(M, f32p), _p(pixels, i32p),
                                          _p(vals, f32p), C.c_int64(n), _p(film, f64p), _p(rgb, f32p))
            return film, rgb
        self.lib.orc_film_add_samples(C.byref(rdesc.c), _p(pixels, i32p), _p(vals, f32p), C.c_int64(n), _p(film, f64p))
        rgb = np.zeros((H, W, 3), dtype=np.float32)
        self.lib.orc_film_resolve(C.byref(rdesc.c), _p(film, f64p), _p(rgb, f32p))
        return film, rgb

    def filter_sample(self, u, filter_cfg=None, table=None, radius=None):
        """Filter::Sample per point u: (n, 3) = p.x, p.y, weight.  "ref": the reference's own filter object described by
        filter_cfg = (kind, rx, ry, p0, p1) (1 gaussian sigma, 2 mitchell B C, 3 sinc tau, 4 triangle, 0 box); "orc":
        the restatement over the tabulated filter `table` [ny, nx] and `radius` (table None: the triangle filter)."""
        u = f32(u); n = u.size // 2
        out = np.zeros((n, 3), dtype=np.float32)
        if self.kind == "ref":
            cfg = f32(filter_cfg)
            self.lib.ref_filter_sample(_p(cfg, f32p), _p(u, f32p), C.c_int64(n), _p(out, f32p))
        elif table is None:
            self.lib.orc_filter_sample(None, C.c_int32(0), C.c_int32(0), C.c_float(radius[0]),
                                       C.c_float(radius[1]), _p(u, f32p), C.c_int64(n), _p(out, f32p))
        else:
            t = f32(table)
            self.lib.orc_filter_sample(_p(t, f32p), C.c_int32(t.shape[1]), C.c_int32(t.shape[0]), C.c_float(radius[0]),
                                       C.c_float(radius[1]), _p(u, f32p), C.c_int64(n), _p(out, f32p))
        return out

    def filter_table(self, filter_cfg):
        """"ref" only: the table FilterSampler's constructor builds for the filter, [ny, nx] float32."""
        assert self.kind == "ref"
        cfg = f32(filter_cfg); nxny = np.zeros(2, dtype=np.int32)
        self.lib.ref_filter_table(_p(cfg, f32p), _p(nxny, i32p), None)
        t = np.zeros((int(nxny[1]), int(nxny[0])), dtype=np.float32)
        self.lib.ref_filter_table(_p(cfg, f32p), _p(nxny, i32p), _p(t, f32p))
        return t

    def camera_rays_reference(self, world_from_camera, fov, screen_window, width, height, seed, spp, samples,
                              filter_cfg=None, want_weights=False):
        """"ref" only: the reference's real PerspectiveCamera (constructor + GenerateRayDifferential) driven like
        EvaluatePixelSample.  Returns rays (n, 6) and {cameraFromRaster, renderFromCamera} (2, 4, 4)
        (+ CameraSample::filterWeight per sample with want_weights)."""
        assert self.kind == "ref"
        wfc = f32(np.asarray(world_from_camera).ravel()); sw = f32(screen_window)
        samples = np.ascontiguousarray(samples, dtype=np.int32); n = samples.shape[0]
        rays = np.zeros((n, 6), dtype=np.float32); mats = np.zeros((2, 4, 4), dtype=np.float32)
        cfg = None if filter_cfg is None else f32(filter_cfg)
        weights = np.zeros(n, dtype=np.float32)
        self.lib.ref_camera_rays(_p(wfc, f32p), C.c_float(fov), _p(sw, f32p), C.c_int32(width), C.c_int32(height),
                                 C.c_int32(seed), C.c_int32(spp), _p(samples, i32p), C.c_int64(n), _p(rays, f32p),
                                 _p(mats, f32p), _p(cfg, f32p) if cfg is not None else None, _p(weights, f32p))
        if want_weights:
            return rays, mats, weights
        return rays, mats

    def graph_connect_info(self, search_range, n, vertex_radius):
        rng = None if search_range is None else f32(search_range)
        out = np.zeros(3, dtype=np.float32)
        self.fn("graph_connect_info")(C.c_int32(n), _p(rng, f32p) if rng is not None else None, C.c_float(vertex_radius),
                                      _p(out, f32p))
        return out


class RefGraphIntegrator:
    """The reference's own GraphIntegrator::Initialize + ConnectToGraph (oracle/refprobe_graphint.cpp) over a graph
    file `<directory>/files/graphs/<name>.txt` written by the reference's writer."""

    def __init__(self, ck, directory, name):
        assert ck.kind == "ref"
        self.ck = ck
        fn = ck.lib.ref_graph_integrator_open; fn.restype = C.c_void_p
        h = fn(str(directory).encode(), str(name).encode())
        assert h, "ref_graph_integrator_open failed"
        self.h = C.c_void_p(h)

    def info(self):
        out = np.zeros(4, dtype=np.float32)
        self.ck.lib.ref_graph_integrator_info(self.h, _p(out, f32p))
        return out

    def connect(self, points):
        points = f32(points); n = points.size // 3
        out = np.empty(n, dtype=np.float32)
        self.ck.lib.ref_graph_connect(self.h, _p(points, f32p), C.c_int64(n), _p(out, f32p))
        return out

    def close(self):
        self.ck.lib.ref_graph_integrator_close(self.h)


class CheckerGraph:
    """A FreeGraph held by the checker (reference: the fork's own graph::FreeGraph; oracle: FreeGraphR)."""

    def __init__(self, ck, handle):
        self.ck, self.h = ck, C.c_void_p(handle)

    def close(self):
        if self.h:
            self.ck.fn("graph_free")(self.h)
            self.h = None

    def write(self, path, desc, flags, opts=(1, 1, 1)):
        fl = np.asarray(flags, dtype=np.int32); op = np.asarray(opts, dtype=np.int32)
        self.ck.fn("graph_write")(self.h, path.encode(), desc.encode(), _p(fl, i32p), _p(op, i32p))

    def counts(self):
        out = np.zeros(11, dtype=np.int32); r = C.c_float()
        self.ck.fn("graph_counts")(self.h, _p(out, i32p), C.byref(r))
        return out, np.float32(r.value)

    def get(self):
        c, _ = self.counts()
        nv, ne = int(c[8]), int(c[9])
        d = dict(vid=np.zeros(nv, np.int32), pos=np.zeros((nv, 3), np.float32), vtype=np.zeros(nv, np.int32),
                 vlight=np.zeros(nv, np.float32), vsamples=np.zeros(nv, np.int32), vrange=np.zeros(nv, np.float32),
                 vcoors=np.zeros((nv, 3), np.int32), eid=np.zeros(ne, np.int32), efrom=np.zeros(ne, np.int32),
                 eto=np.zeros(ne, np.int32), esamples=np.zeros(ne, np.int32))
        self.ck.fn("graph_get")(self.h, _p(d["vid"], i32p), _p(d["pos"], f32p), _p(d["vtype"], i32p), _p(d["vlight"], f32p),
                                _p(d["vsamples"], i32p), _p(d["vrange"], f32p), _p(d["vcoors"], i32p), _p(d["eid"], i32p),
                                _p(d["efrom"], i32p), _p(d["eto"], i32p), _p(d["esamples"], i32p))
        return d

    def transport(self, L0=None, bounces=0):
        """LightingCalculator::GetTransportMatrix as CSR (+ ComputeFinalLight(L0, T) when L0 is given)."""
        c, _ = self.counts()
        n, cap = int(c[8]), int(c[9])
        rowptr = np.zeros(n + 1, np.int64); col = np.zeros(max(cap, 1), np.int32); val = np.zeros(max(cap, 1), np.float32)
        total = np.zeros(n, np.float32); iters = C.c_int32(-1)
        fn = self.ck.fn("graph_transport"); fn.restype = C.c_int64
        l0 = None if L0 is None else f32(L0)
        nnz = fn(self.h, C.c_int64(cap), _p(rowptr, i64p), _p(col, i32p), _p(val, f32p),
                 _p(l0, f32p) if l0 is not None else None, C.c_int32(bounces), _p(total, f32p), C.byref(iters))
        return rowptr, col[:nnz].copy(), val[:nnz].copy(), total, iters.value


class CheckerMedium:
    def __init__(self, ck, handle, desc):
        self.ck, self.h, self.desc = ck, C.c_void_p(handle), desc

    def close(self):
        if self.h:
            self.ck.fn("medium_destroy")(self.h)
            self.h = None

    def is_emissive(self):
        return bool(self.ck.fn("medium_is_emissive")(self.h))

    def majorants(self):
        res = np.zeros(3, dtype=np.int32)
        mr = self.desc.majorant_res
        out = np.empty(max(int(np.prod(mr)), 64 ** 3), dtype=np.float32)
        self.ck.fn("medium_majorants")(self.h, _p(out, f32p), _p(res, i32p))
        return out[: int(np.prod(res))].reshape(res[2], res[1], res[0])

    def bounds(self):
        out = np.zeros(6, dtype=np.float32)
        self.ck.fn("medium_bounds")(self.h, _p(out, f32p))
        return out

    def sample_point(self, p, lam):
        p = f32(p); lam = f32(lam); n = p.size // 3
        out = np.empty((n, 12), dtype=np.float32)
        self.ck.fn("medium_sample_point")(self.h, _p(p, f32p), C.c_int64(n), _p(lam, f32p), _p(out, f32p))
        return out

    def ray_to_medium(self, rays7):
        rays7 = f32(rays7); n = rays7.size // 7
        out = np.empty((n, 9), dtype=np.float32)
        self.ck.fn("medium_ray_to_medium")(self.h, _p(rays7, f32p), C.c_int64(n), _p(out, f32p))
        return out

    def sample_tmaj(self, mode, rays7, rng, u, lam):
        rays7 = f32(rays7); rng = np.ascontiguousarray(rng, dtype=np.uint64); u = f32(u); lam = f32(lam)
        n = rays7.size // 7
        out = np.zeros(n, dtype=TMAJ_DTYPE)
        self.ck.fn("sample_tmaj")(self.h, C.c_int(mode), _p(rays7, f32p), C.c_int64(n), _p(rng, u64p), _p(u, f32p),
                                  _p(lam, f32p), out.ctypes.data_as(C.c_void_p))
        return out

    # oracle-only (the reference's integrators.cpp cannot be built here)
    def render_samples(self, rdesc: RenderDesc):
        x0, y0, x1, y1, s0, s1 = rdesc.ranges()
        n = (x1 - x0) * (y1 - y0) * (s1 - s0)
        out = np.empty((n, 12), dtype=np.float32)
        stats = np.zeros(5, dtype=np.int64)
        self.ck.fn("render_volpath_samples")(self.h, C.byref(rdesc.c), _p(out, f32p), _p(stats, i64p))
        return out, stats

    def render(self, rdesc: RenderDesc, threads=1, film=None):
        if film is None:
            film = np.zeros((rdesc.c.height, rdesc.c.width, 4), dtype=np.float64)
        stats = np.zeros(5, dtype=np.int64)
        self.ck.fn("render_volpath")(self.h, C.byref(rdesc.c), _p(film, f64p), C.c_int(threads), _p(stats, i64p))
        return film, stats

    def camera_rays(self, rdesc: RenderDesc, samples=None):
        """(pixel, sample) list {px, py, sampleIndex} (default: the render's own enumeration) and its camera
        rays, as the oracle's camera forms them."""
        if samples is None:
            x0, y0, x1, y1, s0, s1 = rdesc.ranges()
            samples = np.array([(x, y, s) for y in range(y0, y1) for x in range(x0, x1) for s in range(s0, s1)], dtype=np.int32)
        samples = np.ascontiguousarray(samples, dtype=np.int32)
        rays = np.zeros((samples.shape[0], 6), dtype=np.float32)
        orc = Checker("orc") if self.ck.kind != "orc" else self.ck
        orc.lib.orc_camera_rays_list(C.byref(rdesc.c), _p(samples, i32p), C.c_int64(samples.shape[0]), _p(rays, f32p))
        return samples, rays

    def render_graph_samples(self, rdesc: RenderDesc, pos, light, search_range, vertex_radius, world_from_render=None):
        """GraphIntegrator::Li (free-graph branch) per pixel sample ("orc" only): {L[4], lambda[4], pdf[4]}."""
        pos = f32(pos); light = f32(light); n = light.size
        rng = None if search_range is None else f32(search_range)
        wfr = f32(np.eye(4) if world_from_render is None else world_from_render).ravel()
        x0, y0, x1, y1, s0, s1 = rdesc.ranges()
        out = np.zeros(((x1 - x0) * (y1 - y0) * (s1 - s0), 12), dtype=np.float32)
        self.ck.lib.orc_render_graph_samples(self.h, C.byref(rdesc.c), C.c_int32(n), _p(pos, f32p), _p(light, f32p),
                                             _p(rng, f32p) if rng is not None else None, C.c_float(vertex_radius),
                                             _p(wfr, f32p), _p(out, f32p))
        return out

    def volpath_li(self, rdesc: RenderDesc, samples=None, rays=None):
        """Per-sample {L[4], lambda[4], pdf[4]} of the render.  "ref": the reference's own
        VolPathIntegrator::Li (oracle/refprobe_volpath.cpp) over the oracle's camera rays, sphere boundary
        only; also returns {light wi, sceneRadius} as the reference's DistantLight computes them.
        "orc": the restatement (render_samples).  The call releases the GIL: host threads may run
        disjoint sample lists concurrently (bench.py --impl reference)."""
        if self.ck.kind != "ref":
            out, _ = self.render_samples(rdesc)
            c = rdesc.c
            return out, np.array([c.light_dir[0], c.light_dir[1], c.light_dir[2], c.scene_radius], dtype=np.float32)
        btype = int(rdesc.c.boundary_type)  # 0: Shape "sphere"; 1: 12-triangle box under a BVH
        if rays is None:
            samples, rays = self.camera_rays(rdesc, samples)
        samples = np.ascontiguousarray(samples, dtype=np.int32); rays = f32(rays)
        n = samples.shape[0]
        sphere = np.array(list(rdesc.c.boundary)[:6], dtype=np.float32)
        ld = np.array(list(rdesc.c.light_dir), dtype=np.float32)
        out = np.zeros((n, 12), dtype=np.float32); info = np.zeros(4, dtype=np.float32)
        self.ck.lib.ref_volpath_li(self.h, _p(sphere, f32p), _p(ld, f32p), _p(rdesc.light_spectrum, f32p),
                                   C.c_int32(rdesc.c.max_depth), C.c_int32(rdesc.c.sampler_seed), C.c_int32(rdesc.c.spp),
                                   _p(samples, i32p), _p(rays, f32p), C.c_int64(n), _p(out, f32p), _p(info, f32p),
                                   _p(np.array([rdesc.c.sampler, rdesc.c.randomize, rdesc.c.width, rdesc.c.height, btype],
                                               dtype=np.int32), i32p))
        return out, info

    def graph_pipeline(self, rdesc: RenderDesc, *, dimension_steps, iterations_per_step, max_depth, light_iterations,
                       points_on_radius, bounces, radius_modifier=1.0, search_range_neighbours=0, sampler_seed=0):
        """"ref" only: the reference's own graph pipeline (FreeGraphBuilder::BuildGraph -> GetLightVector ->
        GetTransportMatrix -> ComputeFinalLight per bounce count; oracle/refprobe_graphbuild.cpp) on this medium
        inside the render desc's sphere boundary, lit by its distant light.  Returns a dict of arrays."""
        assert self.ck.kind == "ref" and int(rdesc.c.boundary_type) == 0
        lib = self.ck.lib
        lib.ref_graphpipe_run.restype = C.c_void_p
        sphere = np.array(list(rdesc.c.boundary)[:4], dtype=np.float32)
        wi = np.array(list(rdesc.c.light_dir), dtype=np.float32)
        icfg = np.array([dimension_steps, iterations_per_step, max_depth, light_iterations, points_on_radius,
                         1 if search_range_neighbours > 0 else 0, search_range_neighbours, sampler_seed], dtype=np.int32)
        fcfg = np.array([radius_modifier], dtype=np.float32)
        b = np.ascontiguousarray(bounces, dtype=np.int32)
        h = C.c_void_p(lib.ref_graphpipe_run(self.h, _p(sphere, f32p), _p(wi, f32p), _p(icfg, i32p), _p(fcfg, f32p),
                                             _p(b, i32p), C.c_int32(b.size)))
        counts = np.zeros(4, dtype=np.int32); info = np.zeros(9, dtype=np.float32)
        lib.ref_graphpipe_info(h, _p(counts, i32p), _p(info, f32p))
        nv, ne = int(counts[0]), int(counts[1])
        out = dict(pos=np.zeros((nv, 3), np.float32), vsamples=np.zeros(nv, np.int32), vrange=np.zeros(nv, np.float32),
                   efrom=np.zeros(ne, np.int32), eto=np.zeros(ne, np.int32), esamples=np.zeros(ne, np.int32),
                   L0=np.zeros(nv, np.float32), totals=np.zeros((b.size, nv), np.float32), iters=np.zeros(b.size, np.int32),
                   list_index=np.zeros(nv, np.int32))
        lib.ref_graphpipe_get(h, _p(out["pos"], f32p), _p(out["vsamples"], i32p), _p(out["vrange"], f32p),
                              _p(out["efrom"], i32p), _p(out["eto"], i32p), _p(out["esamples"], i32p), _p(out["L0"], f32p),
                              _p(out["totals"], f32p), _p(out["iters"], i32p), _p(out["list_index"], i32p))
        lib.ref_graphpipe_free(h)
        out.update(sampling_resolution=int(counts[2]), pixel_samples=int(counts[3]), vertex_radius=float(info[0]),
                   max_dist_to_center=float(info[1]), lam=info[2:6].copy(), bounds_center=info[6:9].copy(), bounces=b)
        return out

    def graph_light_rays(self, rays6, pixel, ts, iterations, seed, lam):
        rays6 = f32(rays6); pixel = np.ascontiguousarray(pixel, dtype=np.int32); ts = f32(ts); lam = f32(lam)
        n = rays6.size // 6
        out = np.empty(n, dtype=np.float32)
        self.ck.fn("graph_light_rays")(self.h, _p(rays6, f32p), _p(pixel, i32p), _p(ts, f32p), C.c_int64(n),
                                       C.c_int32(iterations), C.c_int32(seed), _p(lam, f32p), _p(out, f32p))
        return out
