"""ctypes mirror of include/avr.h (POD structs only).  No compute here."""
import ctypes as C

import numpy as np

NSPEC = 471

f32p = C.POINTER(C.c_float)
i32p_ = C.POINTER(C.c_int32)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return a.ctypes.data_as(f32p) if a is not None else None


class avr_grid_desc(C.Structure):
    _fields_ = [
        ("nx", C.c_int32), ("ny", C.c_int32), ("nz", C.c_int32),
        ("density", f32p), ("temperature", f32p), ("lescale", f32p),
        ("le_norm", C.c_float),
        ("p0", C.c_float * 3), ("p1", C.c_float * 3),
        ("render_from_medium", C.c_float * 16), ("medium_from_render", C.c_float * 16),
        ("sigma_a", f32p), ("sigma_s", f32p), ("le", f32p),
        ("sigma_scale", C.c_float), ("g", C.c_float),
        ("temperature_offset", C.c_float), ("temperature_scale", C.c_float),
        ("majorant_res", C.c_int32 * 3),
    ]


class avr_rgbgrid_desc(C.Structure):
    _fields_ = [
        ("nx", C.c_int32), ("ny", C.c_int32), ("nz", C.c_int32),
        ("sigma_a", f32p), ("sigma_s", f32p), ("le", f32p), ("le_illuminant", f32p),
        ("le_scale", C.c_float),
        ("p0", C.c_float * 3), ("p1", C.c_float * 3),
        ("render_from_medium", C.c_float * 16), ("medium_from_render", C.c_float * 16),
        ("sigma_scale", C.c_float), ("g", C.c_float),
        ("majorant_res", C.c_int32 * 3),
    ]


class avr_nanovdb_grid(C.Structure):
    _fields_ = [
        ("values", f32p),
        ("index_min", C.c_int32 * 3), ("dims", C.c_int32 * 3),
        ("inv_mat", C.c_float * 9), ("vec", C.c_float * 3),
        ("world_bbox", C.c_double * 6),
        ("n_blocks", C.c_int64), ("block_origin", C.POINTER(C.c_int32)), ("block_values", f32p),
    ]


class avr_nvdb_info(C.Structure):
    _fields_ = [
        ("grids_in_file", C.c_int32), ("codec", C.c_int32),
        ("leaf_nodes", C.c_int32), ("lower_nodes", C.c_int32), ("upper_nodes", C.c_int32),
        ("active_voxels", C.c_int64), ("grid_bytes", C.c_int64), ("grid_name", C.c_char * 64),
    ]


class avr_nanovdb_desc(C.Structure):
    _fields_ = [
        ("density", avr_nanovdb_grid), ("temperature", avr_nanovdb_grid),
        ("render_from_medium", C.c_float * 16), ("medium_from_render", C.c_float * 16),
        ("sigma_a", f32p), ("sigma_s", f32p),
        ("sigma_scale", C.c_float), ("g", C.c_float), ("le_scale", C.c_float),
        ("temperature_offset", C.c_float), ("temperature_scale", C.c_float),
        ("majorant_res", C.c_int32 * 3),
    ]


class avr_graph_file(C.Structure):
    _fields_ = [
        ("description", C.c_char * 64),
        ("vertex_radius", C.c_float),
        ("use_coors", C.c_int32), ("use_samples", C.c_int32), ("use_ray_vertex_types", C.c_int32),
        ("use_lighting", C.c_int32), ("use_render_search_range", C.c_int32),
        ("cur_vertex_id", C.c_int32), ("cur_edge_id", C.c_int32), ("cur_path_id", C.c_int32),
        ("n_vertices", C.c_int32), ("n_edges", C.c_int32), ("n_paths", C.c_int32),
        ("vertex_id", i32p_), ("vertex_pos", f32p), ("vertex_type", i32p_), ("vertex_light", f32p),
        ("vertex_samples", i32p_), ("vertex_range", f32p), ("vertex_coors", i32p_),
        ("edge_id", i32p_), ("edge_from", i32p_), ("edge_to", i32p_), ("edge_samples", i32p_),
        ("path_id", i32p_), ("path_offset", C.POINTER(C.c_int64)), ("path_vertices", i32p_),
    ]


class avr_graph_sidecar(C.Structure):
    _fields_ = [
        ("n_vertices", C.c_int32), ("nnz", C.c_int64),
        ("rowptr", C.POINTER(C.c_int64)), ("col", i32p_), ("val", f32p),
        ("vertex_pos", f32p), ("light", f32p), ("search_range", f32p),
    ]


class avr_light_vector_desc(C.Structure):
    _fields_ = [
        ("light_dir", C.c_float * 3), ("boundary_type", C.c_int32), ("boundary", C.c_float * 6),
        ("max_dist_to_center", C.c_float), ("vertex_radius", C.c_float),
        ("light_iterations", C.c_int32), ("points_on_radius", C.c_int32), ("sampler_seed", C.c_int32),
        ("sampling_resolution", C.c_int32), ("lambda_", C.c_float * 4),
    ]


class avr_graph_builder_desc(C.Structure):
    _fields_ = [
        ("light_dir", C.c_float * 3), ("boundary_type", C.c_int32), ("boundary", C.c_float * 6),
        ("bounds_center", C.c_float * 3), ("max_dist_to_center", C.c_float), ("vertex_radius", C.c_float),
        ("dimension_steps", C.c_int32), ("iterations_per_step", C.c_int32), ("max_depth", C.c_int32),
        ("sampler_seed", C.c_int32), ("sampling_resolution", C.c_int32), ("search_range_neighbours", C.c_int32),
        ("lambda_", C.c_float * 4),
    ]


class avr_built_graph(C.Structure):
    _fields_ = [
        ("n_vertices", C.c_int32), ("n_paths", C.c_int32), ("n_points", C.c_int32), ("cluster_rounds", C.c_int32),
        ("n_edges", C.c_int64),
        ("vertex_pos", f32p), ("vertex_samples", i32p_), ("search_range", f32p),
        ("edge_from", i32p_), ("edge_to", i32p_), ("edge_samples", i32p_),
    ]


class avr_filter_table(C.Structure):
    _fields_ = [("nx", C.c_int32), ("ny", C.c_int32), ("f", f32p), ("kind", C.c_int32)]


FILTER_TABULATED, FILTER_TRIANGLE = 0, 1


class avr_render_desc(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32),
        ("pixel_bounds", C.c_int32 * 4),
        ("spp", C.c_int32), ("sample_begin", C.c_int32), ("sample_end", C.c_int32),
        ("max_depth", C.c_int32), ("sampler_seed", C.c_int32), ("exact_math", C.c_int32),
        ("filter_radius", C.c_float * 2),
        ("camera_from_raster", C.c_float * 16), ("render_from_camera", C.c_float * 16),
        ("light_dir", C.c_float * 3),
        ("light_spectrum", f32p),
        ("scene_radius", C.c_float),
        ("boundary_type", C.c_int32),
        ("boundary", C.c_float * 6),
        ("sensor_rgb", f32p),
        ("imaging_ratio", C.c_float),
        ("output_rgb_from_sensor", C.c_float * 9),
        ("max_component_value", C.c_float),
        ("rank", C.c_int32), ("world", C.c_int32),
        ("sampler", C.c_int32), ("randomize", C.c_int32),
        ("filter", C.POINTER(avr_filter_table)),
    ]


class avr_render_stats(C.Structure):
    _fields_ = [
        ("camera_rays", C.c_int64), ("camera_fetches", C.c_int64), ("shadow_fetches", C.c_int64),
        ("segments", C.c_int64), ("scatter_events", C.c_int64),
        ("kernel_ms", C.c_float), ("launches", C.c_int32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


TMAJ_DTYPE = np.dtype([
    ("T_maj", np.float32, 4), ("Tr", np.float32, 4), ("r_u", np.float32, 4), ("r_l", np.float32, 4),
    ("t_event", np.float32), ("event", np.int32), ("n_fetch", np.int32), ("n_seg", np.int32),
])
assert TMAJ_DTYPE.itemsize == 80


def _set_arr(dst, src):
    src = np.asarray(src, dtype=np.float64).ravel()
    for i in range(len(dst)):
        dst[i] = float(src[i])


IDENTITY16 = np.eye(4, dtype=np.float32).ravel()


class GridDesc:
    """Parameters of MakeNamedMedium "uniformgrid" (reference media.cpp:249-330), keeping the numpy
    buffers alive for the lifetime of the ctypes struct."""

    def __init__(self, density, sigma_a, sigma_s, *, temperature=None, lescale=None, le=None, le_norm=1.0,
                 p0=(0, 0, 0), p1=(1, 1, 1), render_from_medium=None, medium_from_render=None, scale=1.0, g=0.0,
                 temperature_offset=0.0, temperature_scale=1.0, majorant_res=(16, 16, 16)):
        self.density = _f32(density)
        assert self.density.ndim == 3, "density must be [nz, ny, nx]"
        nz, ny, nx = self.density.shape
        self.temperature = None if temperature is None else _f32(temperature)
        self.lescale = None if lescale is None else _f32(lescale)
        self.sigma_a = _f32(np.broadcast_to(sigma_a, (NSPEC,)))
        self.sigma_s = _f32(np.broadcast_to(sigma_s, (NSPEC,)))
        self.le = None if le is None else _f32(np.broadcast_to(le, (NSPEC,)))
        self.majorant_res = tuple(int(r) for r in majorant_res)
        m = IDENTITY16 if render_from_medium is None else _f32(render_from_medium).ravel()
        if medium_from_render is None:
            mi = np.linalg.inv(np.asarray(m, dtype=np.float64).reshape(4, 4)).astype(np.float32).ravel()
        else:
            mi = _f32(medium_from_render).ravel()
        self.m, self.minv = m, mi
        c = avr_grid_desc()
        c.nx, c.ny, c.nz = nx, ny, nz
        c.density = _ptr(self.density)
        c.temperature = _ptr(self.temperature)
        c.lescale = _ptr(self.lescale)
        c.le_norm = le_norm
        _set_arr(c.p0, p0); _set_arr(c.p1, p1)
        _set_arr(c.render_from_medium, m); _set_arr(c.medium_from_render, mi)
        c.sigma_a = _ptr(self.sigma_a); c.sigma_s = _ptr(self.sigma_s); c.le = _ptr(self.le)
        c.sigma_scale = scale; c.g = g
        c.temperature_offset = temperature_offset; c.temperature_scale = temperature_scale
        for i in range(3):
            c.majorant_res[i] = self.majorant_res[i]
        self.c = c
        self.p0 = np.asarray(p0, dtype=np.float32); self.p1 = np.asarray(p1, dtype=np.float32)


class RGBGridDesc:
    """Parameters of MakeNamedMedium "rgbgrid" (reference media.cpp:380-453); voxels are
    {scale,c0,c1,c2} as produced by RGBUnboundedSpectrum(cs, rgb)."""

    def __init__(self, sigma_a=None, sigma_s=None, *, le=None, le_illuminant=None, le_scale=1.0, p0=(0, 0, 0),
                 p1=(1, 1, 1), render_from_medium=None, medium_from_render=None, scale=1.0, g=0.0,
                 majorant_res=(16, 16, 16)):
        self.sigma_a = None if sigma_a is None else _f32(sigma_a)
        self.sigma_s = None if sigma_s is None else _f32(sigma_s)
        self.le = None if le is None else _f32(le)
        self.le_illuminant = None if le_illuminant is None else _f32(le_illuminant)
        ref = self.sigma_a if self.sigma_a is not None else self.sigma_s
        assert ref is not None and ref.ndim == 4 and ref.shape[3] == 4, "voxels must be [nz, ny, nx, 4]"
        nz, ny, nx, _ = ref.shape
        self.majorant_res = tuple(int(r) for r in majorant_res)
        m = IDENTITY16 if render_from_medium is None else _f32(render_from_medium).ravel()
        if medium_from_render is None:
            mi = np.linalg.inv(np.asarray(m, dtype=np.float64).reshape(4, 4)).astype(np.float32).ravel()
        else:
            mi = _f32(medium_from_render).ravel()
        self.m, self.minv = m, mi
        c = avr_rgbgrid_desc()
        c.nx, c.ny, c.nz = nx, ny, nz
        c.sigma_a = _ptr(self.sigma_a); c.sigma_s = _ptr(self.sigma_s)
        c.le = _ptr(self.le); c.le_illuminant = _ptr(self.le_illuminant)
        c.le_scale = le_scale
        _set_arr(c.p0, p0); _set_arr(c.p1, p1)
        _set_arr(c.render_from_medium, m); _set_arr(c.medium_from_render, mi)
        c.sigma_scale = scale; c.g = g
        for i in range(3):
            c.majorant_res[i] = self.majorant_res[i]
        self.c = c
        self.p0 = np.asarray(p0, dtype=np.float32); self.p1 = np.asarray(p1, dtype=np.float32)


class NanoVDBGrid:
    """One flattened nanovdb::FloatGrid (see avr_nanovdb_grid): values [dz, dy, dx] over the inclusive
    index bounding box, the float inverse map (mInvMatF, mVecF) and the double world bounding box."""

    def __init__(self, values, index_min, inv_mat, vec, world_bbox, source=None):
        self.values = _f32(values)
        assert self.values.ndim == 3
        self.index_min = np.asarray(index_min, dtype=np.int32)
        self.dims = np.asarray(self.values.shape[::-1], dtype=np.int32)
        self.inv_mat = _f32(inv_mat).ravel()
        self.vec = _f32(vec).ravel()
        self.world_bbox = np.asarray(world_bbox, dtype=np.float64).ravel()
        self.source = source  # (dense, origin_ijk, voxel_size, origin_world): lets the reference probe rebuild it

    def fill(self, c):
        c.values = _ptr(self.values)
        for i in range(3):
            c.index_min[i] = int(self.index_min[i]); c.dims[i] = int(self.dims[i]); c.vec[i] = float(self.vec[i])
        for i in range(9):
            c.inv_mat[i] = float(self.inv_mat[i])
        for i in range(6):
            c.world_bbox[i] = float(self.world_bbox[i])

    @staticmethod
    def from_dense(dense, origin_ijk=(0, 0, 0), voxel_size=1.0, origin_world=(0.0, 0.0, 0.0)):
        """Host-side equivalent of nanovdb::GridBuilder<float>(0)(dense over a box).getHandle(voxel_size,
        origin): the index box is that of the non-background voxels (GridBuilder.h:259-271), the map a
        uniform scale + translation (GridBuilder.h:535-560, Map::set NanoVDB.h:2045-2063), the world box
        map(min) .. map(max+1) (GridStats.h:551-563)."""
        from fractions import Fraction
        dense = _f32(dense)
        nzv = np.argwhere(dense != 0)
        assert len(nzv), "empty grid"
        lo = nzv.min(axis=0); hi = nzv.max(axis=0)           # (z, y, x)
        vals = dense[lo[0]:hi[0] + 1, lo[1]:hi[1] + 1, lo[2]:hi[2] + 1]
        imin = np.asarray(origin_ijk, dtype=np.int64) + lo[::-1]
        imax1 = np.asarray(origin_ijk, dtype=np.int64) + hi[::-1] + 1
        ow = np.asarray(origin_world, dtype=np.float64)
        # Map::applyMap = fma(x, dx, T): one rounding of the exact product-sum (exact rationals here)
        fma = lambda i, t: float(Fraction(int(i)) * Fraction(float(voxel_size)) + Fraction(float(t)))  # noqa: E731
        wb = np.array([fma(imin[k], ow[k]) for k in range(3)] + [fma(imax1[k], ow[k]) for k in range(3)])
        inv = np.zeros(9, dtype=np.float32)
        inv[[0, 4, 8]] = np.float32(1.0 / voxel_size)
        return NanoVDBGrid(vals, imin, inv, ow.astype(np.float32), wb,
                           source=(dense, tuple(int(v) for v in origin_ijk), float(voxel_size), tuple(float(v) for v in ow)))


class NanoVDBDesc:
    """Parameters of MakeNamedMedium "nanovdb" (reference media.cpp:628-665) with the grids flattened."""

    def __init__(self, density: NanoVDBGrid, sigma_a=1.0, sigma_s=1.0, *, temperature: NanoVDBGrid = None,
                 render_from_medium=None, medium_from_render=None, scale=1.0, g=0.0, le_scale=1.0,
                 temperature_offset=0.0, temperature_scale=1.0, majorant_res=(64, 64, 64)):
        self.density, self.temperature = density, temperature
        self.sigma_a = _f32(np.broadcast_to(sigma_a, (NSPEC,)))
        self.sigma_s = _f32(np.broadcast_to(sigma_s, (NSPEC,)))
        self.majorant_res = tuple(int(r) for r in majorant_res)
        m = IDENTITY16 if render_from_medium is None else _f32(render_from_medium).ravel()
        if medium_from_render is None:
            mi = np.linalg.inv(np.asarray(m, dtype=np.float64).reshape(4, 4)).astype(np.float32).ravel()
        else:
            mi = _f32(medium_from_render).ravel()
        self.m, self.minv = m, mi
        c = avr_nanovdb_desc()
        density.fill(c.density)
        if temperature is not None:
            temperature.fill(c.temperature)
        _set_arr(c.render_from_medium, m); _set_arr(c.medium_from_render, mi)
        c.sigma_a = _ptr(self.sigma_a); c.sigma_s = _ptr(self.sigma_s)
        c.sigma_scale = scale; c.g = g; c.le_scale = le_scale
        c.temperature_offset = temperature_offset; c.temperature_scale = temperature_scale
        for i in range(3):
            c.majorant_res[i] = self.majorant_res[i]
        self.c = c
        wb = density.world_bbox.astype(np.float32)
        if temperature is not None:
            wt = temperature.world_bbox.astype(np.float32)
            wb = np.r_[np.minimum(wb[:3], wt[:3]), np.maximum(wb[3:], wt[3:])]
        self.p0, self.p1 = wb[:3].copy(), wb[3:].copy()   # NanoVDBMedium::bounds, media.cpp:536-551


SAMPLERS = {"independent": 0, "zsobol": 1}
RANDOMIZE = {"none": 0, "permutedigits": 1, "fastowen": 2, "owen": 3}


def gaussian_filter_table(radius=(1.5, 1.5), sigma=0.5):
    """GaussianFilter (filters.h:84-120: Evaluate(p) = max(0, G(p.x) - G(r.x)) * max(0, G(p.y) - G(r.y)), G = the
    normal density of std sigma) tabulated as FilterSampler's constructor does (filters.cpp:133-143), in fp32."""
    f32 = np.float32
    rx, ry = f32(radius[0]), f32(radius[1])
    nx, ny = int(f32(32) * rx), int(f32(32) * ry)

    def gauss(x, s):
        # util/math.h Gaussian(x, mu = 0, sigma): 1 / sqrt(2 pi sigma^2) * FastExp(-x^2 / (2 sigma^2)) -- evaluated by
        # the caller's checker where bit-exactness matters; here a float64 evaluation rounded to fp32
        return 1.0 / np.sqrt(2 * np.pi * s * s) * np.exp(-(x * x) / (2 * s * s))
    xs = (-rx + (np.arange(nx, dtype=f32) + f32(0.5)) / f32(nx) * (rx + rx)).astype(np.float64)
    ys = (-ry + (np.arange(ny, dtype=f32) + f32(0.5)) / f32(ny) * (ry + ry)).astype(np.float64)
    gx = np.maximum(0.0, gauss(xs, sigma) - gauss(float(rx), sigma))
    gy = np.maximum(0.0, gauss(ys, sigma) - gauss(float(ry), sigma))
    return np.outer(gy, gx).astype(f32)


class RenderDesc:
    """Camera / light / film / sampler parameters of one volpath render (see avr_render_desc)."""

    def __init__(self, width, height, spp, *, camera_from_raster, render_from_camera, light_dir, light_spectrum,
                 scene_radius, boundary_type, boundary, sensor_rgb, max_depth=5, sampler_seed=0, exact_math=1,
                 filter_radius=(0.5, 0.5), imaging_ratio=1.0, output_rgb_from_sensor=None,
                 max_component_value=float("inf"), pixel_bounds=None, sample_begin=0, sample_end=0, rank=0, world=1,
                 sampler="independent", randomize="fastowen"):
        self.light_spectrum = _f32(np.broadcast_to(light_spectrum, (NSPEC,)))
        self.sensor_rgb = _f32(sensor_rgb).reshape(3, NSPEC)
        c = avr_render_desc()
        c.width, c.height, c.spp = width, height, spp
        pb = pixel_bounds if pixel_bounds is not None else (0, 0, width, height)
        for i in range(4):
            c.pixel_bounds[i] = int(pb[i])
        c.sample_begin, c.sample_end = sample_begin, sample_end
        c.max_depth, c.sampler_seed, c.exact_math = max_depth, sampler_seed, exact_math
        _set_arr(c.filter_radius, filter_radius)
        _set_arr(c.camera_from_raster, camera_from_raster)
        _set_arr(c.render_from_camera, render_from_camera)
        ld = np.asarray(light_dir, dtype=np.float64)
        ld = (ld / np.linalg.norm(ld)).astype(np.float32)
        _set_arr(c.light_dir, ld)
        c.light_spectrum = _ptr(self.light_spectrum)
        c.scene_radius = scene_radius
        c.boundary_type = boundary_type
        b = list(boundary) + [0.0] * (6 - len(boundary))
        _set_arr(c.boundary, b)
        c.sensor_rgb = _ptr(self.sensor_rgb)
        c.imaging_ratio = imaging_ratio
        _set_arr(c.output_rgb_from_sensor, np.eye(3).ravel() if output_rgb_from_sensor is None
                 else output_rgb_from_sensor)
        c.max_component_value = max_component_value
        c.rank, c.world = rank, world
        # "Sampler" names and "randomization" values as the scene file spells them (samplers.cpp:157-165, 417-431)
        c.sampler = SAMPLERS[sampler] if isinstance(sampler, str) else int(sampler)
        c.randomize = RANDOMIZE[randomize] if isinstance(randomize, str) else int(randomize)
        self.c = c

    def __getattr__(self, name):
        # scalar fields of the C struct read through (width, height, spp, sample_begin, ...)
        c = self.__dict__.get("c")
        if c is not None and name in {f[0] for f in avr_render_desc._fields_}:
            return getattr(c, name)
        raise AttributeError(name)

    def ranges(self):
        c = self.c
        x0, y0, x1, y1 = list(c.pixel_bounds)
        if (x0, y0, x1, y1) == (0, 0, 0, 0):
            x1, y1 = c.width, c.height
        s0, s1 = c.sample_begin, c.sample_end
        if s0 == 0 and s1 == 0:
            s1 = c.spp
        return x0, y0, x1, y1, s0, s1

    def with_filter(self, table, radius):
        """Variant with a tabulated pixel filter (avr_filter_table): table = [ny, nx] float32 values of the filter as
        FilterSampler tabulates them (see filter_table()), radius = the filter's radius; table None = box filter,
        table "triangle" = pbrt's triangle filter (sampled analytically, no table)."""
        r = self.with_()
        _set_arr(r.c.filter_radius, radius)
        if table is None:
            r.c.filter = None
            r.filter_values = r.filter_c = None
        elif isinstance(table, str):
            assert table == "triangle"
            r.filter_values = None
            r.filter_c = avr_filter_table(0, 0, None, FILTER_TRIANGLE)
            r.c.filter = C.pointer(r.filter_c)
        else:
            r.filter_values = _f32(table)
            r.filter_c = avr_filter_table(int(r.filter_values.shape[1]), int(r.filter_values.shape[0]), _ptr(r.filter_values))
            r.c.filter = C.pointer(r.filter_c)
        return r

    def with_(self, **kw):
        """Shallow variant with some scalar fields of the C struct replaced."""
        import copy
        r = copy.copy(self)
        r.c = avr_render_desc()
        C.memmove(C.byref(r.c), C.byref(self.c), C.sizeof(avr_render_desc))
        for k, v in kw.items():
            if k == "pixel_bounds":
                for i in range(4):
                    r.c.pixel_bounds[i] = int(v[i])
            else:
                setattr(r.c, k, v)
        return r
