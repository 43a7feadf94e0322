"""Host-side mirror of the reference's medium / integrator interface over the C ABI (include/avr.h).

Names follow the reference: `Context.create_medium("uniformgrid" | "rgbgrid" | "nanovdb", ...)` is
Medium::Create (media.cpp:667-689); `Medium.sample_point / sample_ray_segments / sample_t_maj`
are Medium::SamplePoint / SampleRay+DDAMajorantIterator::Next / SampleT_maj (media.h);
`Context.render("volpath" | "graph", ...)` is Integrator::Create(name)->Render()
(cpu/integrators.cpp:3658-3700); `Graph.compute_final_light` is LightingCalculator::ComputeFinalLight.
Every call goes through libavr_b200.so; a missing library or a non-zero avr_status raises
(the reference ErrorExit()s) — nothing here computes on the CPU.
"""
import ctypes as C
import os

import numpy as np

from ._abi import (avr_graph_sidecar, avr_light_vector_desc, avr_graph_builder_desc, avr_built_graph, NSPEC, TMAJ_DTYPE, GridDesc, NanoVDBDesc, NanoVDBGrid, RenderDesc, RGBGridDesc, avr_graph_file,
                   avr_grid_desc, avr_render_desc, avr_render_stats, avr_rgbgrid_desc)

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libavr_b200.so")

LAYOUTS = {"dense": 0, "octet32": 1, "octet16": 2, "brick16": 3}
TMAJ_RATIO, TMAJ_DELTA, TMAJ_SHADOW = 0, 1, 2

f32p = C.POINTER(C.c_float)
f64p = C.POINTER(C.c_double)
i32p = C.POINTER(C.c_int32)
i64p = C.POINTER(C.c_int64)
u64p = C.POINTER(C.c_uint64)

_lib = None

# every symbol include/avr.h declares (tests check that the .so exports all of them)
ABI_SYMBOLS = [
    "avr_ctx_create", "avr_ctx_destroy", "avr_last_error", "avr_version", "avr_measure_read_bandwidth",
    "avr_medium_create_uniformgrid", "avr_medium_create_rgbgrid", "avr_medium_create_nanovdb", "avr_medium_destroy",
    "avr_nvdb_read_grid", "avr_nvdb_read_blocks", "avr_nvdb_free_grid", "avr_medium_create_nanovdb_file",
    "avr_medium_is_emissive", "avr_medium_bounds", "avr_medium_bytes", "avr_medium_majorants", "avr_medium_lookup_density", "avr_medium_sample_point",
    "avr_dda_segments", "avr_medium_ray_to_medium", "avr_sample_tmaj", "avr_sample_tmaj_device",
    "avr_render_volpath", "avr_render_volpath_device", "avr_render_volpath_samples", "avr_film_resolve",
    "avr_graph_create", "avr_graph_create_from_edges", "avr_graph_get_csr", "avr_graph_file_read",
    "avr_graph_file_write", "avr_graph_file_free", "avr_graph_sidecar_read", "avr_graph_sidecar_write", "avr_graph_sidecar_free", "avr_graph_destroy", "avr_graph_propagate", "avr_graph_propagate_device", "avr_graph_propagate_noalloc",
    "avr_graph_spmv_rows_device", "avr_graph_axpy_rows_device", "avr_graph_set_vertices", "avr_graph_connect",
    "avr_render_graph", "avr_graph_light_rays", "avr_graph_light_vector", "avr_graph_build_free", "avr_built_graph_free", "avr_debug_math",
    "avr_comm_unique_id", "avr_comm_init_rank", "avr_comm_init_all", "avr_comm_destroy", "avr_comm_rank", "avr_comm_world", "avr_comm_graph_exchange", "avr_filter_sample",
    "avr_render_volpath_multi", "avr_graph_propagate_multi",
]


class AvrError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"avr_status {status}: {msg}")
        self.status = status


def load_library():
    """dlopen libavr_b200.so.  Raises if it has not been built — there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(make -C acceleratedvolrenderer_b200/csrc). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    lib.avr_last_error.restype = C.c_char_p
    lib.avr_last_error.argtypes = [C.c_void_p]
    lib.avr_version.restype = C.c_char_p
    lib.avr_medium_bytes.restype = C.c_int64
    for name in ABI_SYMBOLS:
        fn = getattr(lib, name)
        if name not in ("avr_last_error", "avr_version", "avr_medium_bytes", "avr_ctx_destroy", "avr_medium_destroy",
                        "avr_graph_destroy", "avr_graph_file_free", "avr_graph_sidecar_free", "avr_comm_destroy",
                        "avr_built_graph_free", "avr_nvdb_free_grid"):
            fn.restype = C.c_int
    lib.avr_ctx_destroy.restype = None
    lib.avr_medium_destroy.restype = None
    lib.avr_graph_destroy.restype = None
    lib.avr_graph_file_free.restype = None
    lib.avr_graph_file_free.argtypes = [C.c_void_p]
    lib.avr_graph_sidecar_free.restype = None
    lib.avr_graph_sidecar_free.argtypes = [C.c_void_p]
    lib.avr_comm_destroy.restype = None
    lib.avr_built_graph_free.restype = None
    lib.avr_nvdb_free_grid.restype = None
    lib.avr_nvdb_free_grid.argtypes = [C.c_void_p]
    lib.avr_built_graph_free.argtypes = [C.c_void_p]
    lib.avr_comm_destroy.argtypes = [C.c_void_p]
    _lib = lib
    return lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a, t=f32p):
    return a.ctypes.data_as(t)


def read_nvdb_grid(path, grid_name=None):
    """avr_nvdb_read_grid: (NanoVDBGrid or None when the file has no such grid, info dict).  Host only."""
    from ._abi import avr_nanovdb_grid, avr_nvdb_info
    lib = load_library()
    g = avr_nanovdb_grid(); info = avr_nvdb_info()
    st = lib.avr_nvdb_read_grid(str(path).encode(), grid_name.encode() if grid_name else None, C.byref(g), C.byref(info))
    if st != 0:
        raise AvrError(st, lib.avr_last_error(None).decode())
    meta = {k: getattr(info, k) for k, _ in avr_nvdb_info._fields_}
    meta["grid_name"] = info.grid_name.decode()
    if not g.values:
        return None, meta
    dims = list(g.dims)
    n = dims[0] * dims[1] * dims[2]
    vals = np.ctypeslib.as_array(g.values, shape=(n,)).astype(np.float32, copy=True).reshape(dims[2], dims[1], dims[0])
    grid = NanoVDBGrid(vals, list(g.index_min), list(g.inv_mat), list(g.vec), list(g.world_bbox))
    lib.avr_nvdb_free_grid(C.byref(g))
    return grid, meta


def read_nvdb_blocks(path, grid_name=None):
    """avr_nvdb_read_blocks: the grid without flattening.  Returns (dict or None when the file has no such grid, info):
    origins (n, 3) int32 — multiples of 8 in index space —, values (n, 8, 8, 8) float32 indexed [z, y, x], index_min,
    dims, inv_mat, vec, world_bbox.  Host only."""
    from ._abi import avr_nanovdb_grid, avr_nvdb_info
    lib = load_library()
    g = avr_nanovdb_grid(); info = avr_nvdb_info()
    st = lib.avr_nvdb_read_blocks(str(path).encode(), grid_name.encode() if grid_name else None, C.byref(g), C.byref(info))
    if st != 0:
        raise AvrError(st, lib.avr_last_error(None).decode())
    meta = {k: getattr(info, k) for k, _ in avr_nvdb_info._fields_}
    meta["grid_name"] = info.grid_name.decode()
    if not g.block_values:
        return None, meta
    n = int(g.n_blocks)
    out = {
        "origins": np.ctypeslib.as_array(g.block_origin, shape=(max(n, 1) * 3,))[:3 * n].astype(np.int32, copy=True).reshape(n, 3),
        "values": np.ctypeslib.as_array(g.block_values, shape=(max(n, 1) * 512,))[:512 * n].astype(np.float32, copy=True).reshape(n, 8, 8, 8),
        "index_min": np.array(list(g.index_min), dtype=np.int32), "dims": np.array(list(g.dims), dtype=np.int32),
        "inv_mat": np.array(list(g.inv_mat), dtype=np.float32), "vec": np.array(list(g.vec), dtype=np.float32),
        "world_bbox": np.array(list(g.world_bbox), dtype=np.float64),
    }
    lib.avr_nvdb_free_grid(C.byref(g))
    return out, meta


class Context:
    """One per GPU/process (avr_ctx)."""

    def __init__(self, device=0):
        self.lib = load_library()
        h = C.c_void_p()
        st = self.lib.avr_ctx_create(C.c_int(device), C.byref(h))
        if st != 0:
            raise AvrError(st, self.lib.avr_last_error(None).decode())
        self.h = h
        self.device = device

    def check(self, st):
        if st != 0:
            raise AvrError(st, self.lib.avr_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.avr_ctx_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def measure_read_bandwidth(self, nbytes, iters=20):
        """GB/s of a streaming read kernel over `nbytes` (64 MiB: the L2 peak; several GB: the HBM read peak)."""
        out = C.c_float(0)
        self.check(self.lib.avr_measure_read_bandwidth(self.h, C.c_int64(nbytes), C.c_int32(iters), C.byref(out)))
        return float(out.value)

    # Medium::Create(name, params...) — media.cpp:667-689
    def create_medium(self, name, desc, layout="dense"):
        h = C.c_void_p()
        if name == "uniformgrid":
            assert isinstance(desc, GridDesc)
            self.check(self.lib.avr_medium_create_uniformgrid(self.h, C.byref(desc.c), C.c_int(LAYOUTS[layout]),
                                                              C.byref(h)))
        elif name == "rgbgrid":
            assert isinstance(desc, RGBGridDesc)
            self.check(self.lib.avr_medium_create_rgbgrid(self.h, C.byref(desc.c), C.byref(h)))
        elif name == "nanovdb":
            assert isinstance(desc, NanoVDBDesc)
            self.check(self.lib.avr_medium_create_nanovdb(self.h, C.byref(desc.c), C.c_int(LAYOUTS[layout]),
                                                          C.byref(h)))
        else:
            # the reference: ErrorExit(loc, "%s: medium unknown.", name)
            raise AvrError(1, f"{name}: medium unknown.")
        return Medium(self, h, desc)

    # MakeNamedMedium "nanovdb" "string filename" ...: NanoVDBMedium::Create from the file (media.cpp:628-665)
    def create_medium_nanovdb_file(self, filename, params: NanoVDBDesc, gridname=None, temperaturename=None, layout="dense"):
        h = C.c_void_p()
        self.check(self.lib.avr_medium_create_nanovdb_file(self.h, str(filename).encode(),
                                                           gridname.encode() if gridname else None,
                                                           temperaturename.encode() if temperaturename else None,
                                                           C.byref(params.c), C.c_int(LAYOUTS[layout]), C.byref(h)))
        return Medium(self, h, params)

    # Integrator::Create("volpath") -> Render(): returns the film [H, W, 4] float64 (rgbSum, weightSum)
    def render_volpath(self, medium, rdesc: RenderDesc, film=None):
        if film is None:
            film = np.zeros((rdesc.c.height, rdesc.c.width, 4), dtype=np.float64)
        stats = avr_render_stats()
        self.check(self.lib.avr_render_volpath(self.h, medium.h, C.byref(rdesc.c), _p(film, f64p), C.byref(stats)))
        return film, stats.as_dict()

    def render_volpath_device(self, medium, rdesc: RenderDesc, film_dev_ptr, stream=None, want_stats=True):
        stats = avr_render_stats()
        self.check(self.lib.avr_render_volpath_device(self.h, medium.h, C.byref(rdesc.c), C.c_void_p(film_dev_ptr),
                                                      C.c_void_p(stream or 0),
                                                      C.byref(stats) if want_stats else None))
        return stats.as_dict()

    def render_volpath_samples(self, medium, rdesc: RenderDesc):
        x0, y0, x1, y1, s0, s1 = rdesc.ranges()
        n = (x1 - x0) * (y1 - y0) * (s1 - s0)
        out = np.zeros((n, 12), dtype=np.float32)
        stats = avr_render_stats()
        self.check(self.lib.avr_render_volpath_samples(self.h, medium.h, C.byref(rdesc.c), _p(out), C.byref(stats)))
        return out, stats.as_dict()

    def render_graph(self, medium, graph, rdesc: RenderDesc, world_from_render=None):
        film = np.zeros((rdesc.c.height, rdesc.c.width, 4), dtype=np.float64)
        wfr = _f32(np.eye(4) if world_from_render is None else world_from_render).ravel()
        stats = avr_render_stats()
        self.check(self.lib.avr_render_graph(self.h, medium.h, graph.h, C.byref(rdesc.c), _p(wfr), _p(film, f64p),
                                             C.byref(stats)))
        return film, stats.as_dict()

    def filter_sample(self, table, radius, u):
        """FilterSampler::Sample (table = "triangle": TriangleFilter::Sample) through the render kernels' device code:
        (n, 3) = p.x, p.y, weight."""
        from ._abi import FILTER_TRIANGLE, avr_filter_table
        u = _f32(u); n = u.size // 2
        if isinstance(table, str):
            assert table == "triangle"
            t = avr_filter_table(0, 0, None, FILTER_TRIANGLE)
        else:
            table = _f32(table)
            t = avr_filter_table(int(table.shape[1]), int(table.shape[0]), _p(table))
        rad = (C.c_float * 2)(float(radius[0]), float(radius[1]))
        out = np.empty((n, 3), dtype=np.float32)
        self.check(self.lib.avr_filter_sample(self.h, C.byref(t), rad, _p(u), C.c_int64(n), _p(out)))
        return out

    def film_resolve(self, rdesc: RenderDesc, film):
        rgb = np.zeros((rdesc.c.height, rdesc.c.width, 3), dtype=np.float32)
        film = np.ascontiguousarray(film, dtype=np.float64)
        st = self.lib.avr_film_resolve(C.byref(rdesc.c), _p(film, f64p), _p(rgb))
        if st != 0:
            raise AvrError(st, "avr_film_resolve: bad argument")
        return rgb

    def debug_math(self, kind, x):
        x = _f32(x); out = np.empty_like(x)
        self.check(self.lib.avr_debug_math(self.h, C.c_int(kind), _p(x), C.c_int64(x.size), _p(out)))
        return out

    def create_graph(self, rowptr, col, val):
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
        col = np.ascontiguousarray(col, dtype=np.int32)
        val = _f32(val)
        h = C.c_void_p()
        self.check(self.lib.avr_graph_create(self.h, C.c_int32(rowptr.size - 1), C.c_int64(col.size), _p(rowptr, i64p),
                                             _p(col, i32p), _p(val), C.byref(h)))
        return Graph(self, h, rowptr.size - 1, col.size)

    # LightingCalculator::GetTransportMatrix on the device, from vertex sample counts and an edge list
    def create_graph_from_edges(self, vertex_samples, edge_from, edge_to, edge_samples):
        vs = np.ascontiguousarray(vertex_samples, dtype=np.int32)
        ef = np.ascontiguousarray(edge_from, dtype=np.int32); et = np.ascontiguousarray(edge_to, dtype=np.int32)
        es = np.ascontiguousarray(edge_samples, dtype=np.int32)
        h = C.c_void_p()
        self.check(self.lib.avr_graph_create_from_edges(self.h, C.c_int32(vs.size), _p(vs, i32p), C.c_int64(ef.size),
                                                        _p(ef, i32p), _p(et, i32p), _p(es, i32p), C.byref(h)))
        g = Graph(self, h, vs.size, 0)
        g.nnz = g.csr_sizes()[1]
        return g


COMM_ID_BYTES = 128


def comm_unique_id():
    """avr_comm_unique_id: 128 bytes that rank 0 creates and hands to every other rank."""
    lib = load_library()
    buf = (C.c_uint8 * COMM_ID_BYTES)()
    st = lib.avr_comm_unique_id(buf)
    if st != 0:
        raise AvrError(st, lib.avr_last_error(None).decode())
    return bytes(buf)


class Comm:
    """avr_comm: this rank's end of the NCCL communicator behind the C ABI (one rank per GPU)."""

    def __init__(self, ctx, unique_id: bytes, rank: int, world: int):
        self.ctx, self.lib, self.rank, self.world = ctx, ctx.lib, rank, world
        buf = (C.c_uint8 * COMM_ID_BYTES).from_buffer_copy(unique_id)
        h = C.c_void_p()
        ctx.check(self.lib.avr_comm_init_rank(ctx.h, buf, C.c_int32(rank), C.c_int32(world), C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.avr_comm_destroy(self.h)
            self.h = None

    # Integrator::Render over all ranks: the complete film lands (is ADDED) in rank 0's host array
    def render_volpath(self, medium, rdesc: RenderDesc, film=None):
        if film is None and self.rank == 0:
            film = np.zeros((rdesc.c.height, rdesc.c.width, 4), dtype=np.float64)
        stats = avr_render_stats()
        self.ctx.check(self.lib.avr_render_volpath_multi(self.h, medium.h, C.byref(rdesc.c),
                                                         _p(film, f64p) if film is not None else None, C.byref(stats)))
        return film, stats.as_dict()

    # LightingCalculator::ComputeFinalLight, rows split over the ranks; total_ptr: n floats, this rank's rows written
    def propagate(self, graph, L0_ptr, bounces, total_ptr, stream=None, want_iters=True):
        it = C.c_int32(-1)
        self.ctx.check(self.lib.avr_graph_propagate_multi(self.h, graph.h, C.c_void_p(L0_ptr), C.c_int32(bounces),
                                                          C.c_void_p(total_ptr), C.byref(it) if want_iters else None,
                                                          C.c_void_p(stream or 0)))
        return int(it.value)

    def graph_exchange(self):
        """1: SpMV writes into the peers' memory; -1: ncclAllGather per bounce; 0: not decided yet."""
        return int(self.lib.avr_comm_graph_exchange(self.h))

    def row_range(self, n):
        chunk = (n + self.world - 1) // self.world
        return min(self.rank * chunk, n), min((self.rank + 1) * chunk, n)


class GraphSidecar:
    """Binary ".avrg" sidecar of a graph (include/avr.h: avr_graph_sidecar): the CSR transport matrix plus the
    per-vertex position, light vector L0 and render search range that `graph_maker` never writes to disk
    (cmd/graph_maker.cpp:167-169).  Host-only."""

    def __init__(self, rowptr, col, val, vertex_pos=None, light=None, search_range=None):
        f32 = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float32)  # noqa: E731
        self.rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
        self.col = np.ascontiguousarray(col, dtype=np.int32)
        self.val = _f32(val)
        self.vertex_pos, self.light, self.search_range = f32(vertex_pos), f32(light), f32(search_range)

    @property
    def n_vertices(self):
        return self.rowptr.size - 1

    def write(self, path):
        lib = load_library()
        c = avr_graph_sidecar()
        c.n_vertices, c.nnz = self.n_vertices, self.col.size
        c.rowptr = _p(self.rowptr, i64p)
        if self.col.size:
            c.col, c.val = _p(self.col, i32p), _p(self.val)
        keep = []
        for name in ("vertex_pos", "light", "search_range"):
            a = getattr(self, name)
            if a is not None:
                want = self.n_vertices * (3 if name == "vertex_pos" else 1)
                if a.size != want:
                    raise ValueError(f"{name}: {a.size} values for {self.n_vertices} vertices")
                flat = np.ascontiguousarray(a.reshape(-1)); keep.append(flat)
                setattr(c, name, _p(flat))
        st = lib.avr_graph_sidecar_write(str(path).encode(), C.byref(c))
        if st != 0:
            raise AvrError(st, lib.avr_last_error(None).decode())

    @staticmethod
    def read(path):
        lib = load_library()
        h = C.POINTER(avr_graph_sidecar)()
        st = lib.avr_graph_sidecar_read(str(path).encode(), C.byref(h))
        if st != 0:
            raise AvrError(st, lib.avr_last_error(None).decode())
        c = h.contents
        n, nnz = c.n_vertices, c.nnz

        def arr(ptr, cnt, dt):
            if not ptr:
                return None
            return np.ctypeslib.as_array(ptr, shape=(cnt,)).astype(dt, copy=True) if cnt else np.zeros(0, dt)

        g = GraphSidecar(arr(c.rowptr, n + 1, np.int64), arr(c.col, nnz, np.int32), arr(c.val, nnz, np.float32),
                         arr(c.vertex_pos, 3 * n, np.float32), arr(c.light, n, np.float32), arr(c.search_range, n, np.float32))
        if g.vertex_pos is not None:
            g.vertex_pos = g.vertex_pos.reshape(-1, 3)
        lib.avr_graph_sidecar_free(h)
        return g


class GraphFile:
    """The fork's graph text file (src/graph/graph.cpp:284-416, :644-663) as numpy arrays.  Host-only:
    needs the library but no GPU.  Fields that a flag switches off are None."""

    FIELDS = ("vertex_id", "vertex_pos", "vertex_type", "vertex_light", "vertex_samples", "vertex_range",
              "vertex_coors", "edge_id", "edge_from", "edge_to", "edge_samples", "path_id", "path_offset",
              "path_vertices")

    def __init__(self, vertex_radius, vertex_id, vertex_pos, *, description="basic", vertex_type=None, vertex_light=None,
                 vertex_samples=None, vertex_range=None, vertex_coors=None, edge_id=None, edge_from=None, edge_to=None,
                 edge_samples=None, path_id=None, path_offset=None, path_vertices=None, cur_ids=(0, 0, 0)):
        i32 = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.int32)  # noqa: E731
        f32 = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float32)  # noqa: E731
        self.description, self.vertex_radius, self.cur_ids = description, float(vertex_radius), tuple(cur_ids)
        self.vertex_id, self.vertex_pos = i32(vertex_id), f32(vertex_pos)
        self.vertex_type, self.vertex_light, self.vertex_samples = i32(vertex_type), f32(vertex_light), i32(vertex_samples)
        self.vertex_range, self.vertex_coors = f32(vertex_range), i32(vertex_coors)
        self.edge_id, self.edge_from, self.edge_to, self.edge_samples = i32(edge_id), i32(edge_from), i32(edge_to), i32(edge_samples)
        self.path_id, self.path_vertices = i32(path_id), i32(path_vertices)
        self.path_offset = None if path_offset is None else np.ascontiguousarray(path_offset, dtype=np.int64)

    # FreeGraph::ReadFromDisk
    @staticmethod
    def read(path):
        lib = load_library()
        h = C.POINTER(avr_graph_file)()
        st = lib.avr_graph_file_read(str(path).encode(), C.byref(h))
        if st != 0:
            raise AvrError(st, lib.avr_last_error(None).decode())
        c = h.contents
        nv, ne, npth = c.n_vertices, c.n_edges, c.n_paths

        def arr(ptr, n, dt):
            return None if not ptr else np.ctypeslib.as_array(ptr, shape=(n,)).astype(dt, copy=True)

        npv = int(c.path_offset[npth]) if c.path_offset else 0
        g = GraphFile(c.vertex_radius, arr(c.vertex_id, nv, np.int32), arr(c.vertex_pos, 3 * nv, np.float32),
                      description=c.description.decode(), vertex_type=arr(c.vertex_type, nv, np.int32),
                      vertex_light=arr(c.vertex_light, nv, np.float32), vertex_samples=arr(c.vertex_samples, nv, np.int32),
                      vertex_range=arr(c.vertex_range, nv, np.float32), vertex_coors=arr(c.vertex_coors, 3 * nv, np.int32),
                      edge_id=arr(c.edge_id, ne, np.int32), edge_from=arr(c.edge_from, ne, np.int32),
                      edge_to=arr(c.edge_to, ne, np.int32), edge_samples=arr(c.edge_samples, ne, np.int32),
                      path_id=arr(c.path_id, npth, np.int32), path_offset=arr(c.path_offset, npth + 1, np.int64),
                      path_vertices=arr(c.path_vertices, npv, np.int32),
                      cur_ids=(c.cur_vertex_id, c.cur_edge_id, c.cur_path_id))
        g.flags = (c.use_coors, c.use_samples, c.use_ray_vertex_types, c.use_lighting, c.use_render_search_range)
        g.counts = (nv, ne, npth)
        if g.vertex_pos is not None:
            g.vertex_pos = g.vertex_pos.reshape(-1, 3)
        if g.vertex_coors is not None:
            g.vertex_coors = g.vertex_coors.reshape(-1, 3)
        lib.avr_graph_file_free(h)
        return g

    # Graph::WriteToDisk; flags = (useCoors, useSamples, useRayVertexTypes, useLighting, useRenderSearchRange),
    # options = (writeVertices, writeEdges, writePaths) as graph::StreamFlags / StreamOptions
    def write(self, path, flags, options=(1, 1, 1)):
        lib = load_library()
        c = avr_graph_file()
        c.description = self.description.encode()
        c.vertex_radius = self.vertex_radius
        (c.use_coors, c.use_samples, c.use_ray_vertex_types, c.use_lighting, c.use_render_search_range) = [int(bool(f)) for f in flags]
        c.cur_vertex_id, c.cur_edge_id, c.cur_path_id = self.cur_ids
        c.n_vertices = self.vertex_id.size if options[0] and self.vertex_id is not None else 0
        c.n_edges = self.edge_id.size if options[1] and self.edge_id is not None else 0
        c.n_paths = self.path_id.size if options[2] and self.path_id is not None else 0
        for name in self.FIELDS:
            a = getattr(self, name)
            if a is not None and a.size:
                setattr(c, name, a.ctypes.data_as(dict(avr_graph_file._fields_)[name]))
        st = lib.avr_graph_file_write(str(path).encode(), C.byref(c))
        if st != 0:
            raise AvrError(st, lib.avr_last_error(None).decode())


class Medium:
    def __init__(self, ctx, h, desc):
        self.ctx, self.h, self.desc = ctx, h, desc
        self.lib = ctx.lib

    def close(self):
        if self.h:
            self.lib.avr_medium_destroy(self.h)
            self.h = None

    def is_emissive(self):  # Medium::IsEmissive
        return bool(self.lib.avr_medium_is_emissive(self.h))

    def bytes(self):
        return int(self.lib.avr_medium_bytes(self.h))

    def bounds(self):  # Medium bounds in medium space (NanoVDBMedium::GetBounds)
        out = np.zeros(6, dtype=np.float32)
        self.ctx.check(self.lib.avr_medium_bounds(self.h, _p(out)))
        return out

    def majorants(self):
        mr = self.desc.majorant_res
        out = np.empty(int(np.prod(mr)), dtype=np.float32)
        res = np.zeros(3, dtype=np.int32)
        self.ctx.check(self.lib.avr_medium_majorants(self.h, _p(out), _p(res, i32p)))
        return out.reshape(res[2], res[1], res[0])

    def lookup_density(self, p):
        p = _f32(p); n = p.size // 3
        out = np.empty(n, dtype=np.float32)
        self.ctx.check(self.lib.avr_medium_lookup_density(self.h, _p(p), C.c_int64(n), _p(out)))
        return out

    def sample_point(self, p, lam):  # Medium::SamplePoint
        p = _f32(p); lam = _f32(lam); n = p.size // 3
        out = np.empty((n, 12), dtype=np.float32)
        self.ctx.check(self.lib.avr_medium_sample_point(self.h, _p(p), C.c_int64(n), _p(lam), _p(out)))
        return out

    def ray_to_medium(self, rays7):
        rays7 = _f32(rays7); n = rays7.size // 7
        out = np.empty((n, 9), dtype=np.float32)
        self.ctx.check(self.lib.avr_medium_ray_to_medium(self.h, _p(rays7), C.c_int64(n), _p(out)))
        return out

    def dda_segments(self, rays8, sigma_t=1.0, max_per_ray=200):  # DDAMajorantIterator
        rays8 = _f32(rays8); n = rays8.size // 8
        counts = np.zeros(n, dtype=np.int32)
        seg_t = np.zeros((n, max_per_ray, 2), dtype=np.float32)
        seg_s = np.zeros((n, max_per_ray), dtype=np.float32)
        seg_v = np.full((n, max_per_ray), -1, dtype=np.int32)
        self.ctx.check(self.lib.avr_dda_segments(self.h, _p(rays8), C.c_int64(n), C.c_float(sigma_t),
                                                 C.c_int32(max_per_ray), _p(counts, i32p), _p(seg_t), _p(seg_s),
                                                 _p(seg_v, i32p)))
        return counts, seg_t, seg_s, seg_v

    def sample_t_maj(self, mode, rays7, rng, u, lam):  # SampleT_maj
        rays7 = _f32(rays7); rng = np.ascontiguousarray(rng, dtype=np.uint64); u = _f32(u); lam = _f32(lam)
        n = rays7.size // 7
        out = np.zeros(n, dtype=TMAJ_DTYPE)
        self.ctx.check(self.lib.avr_sample_tmaj(self.h, C.c_int(mode), _p(rays7), C.c_int64(n), _p(rng, u64p), _p(u),
                                                _p(lam), out.ctypes.data_as(C.c_void_p)))
        return out

    def sample_t_maj_device(self, mode, exact, rays_ptr, n, rng_ptr, u_ptr, lam, out_ptr):
        lam = _f32(lam)
        self.ctx.check(self.lib.avr_sample_tmaj_device(self.h, C.c_int(mode), C.c_int(1 if exact else 0),
                                                       C.c_void_p(rays_ptr), C.c_int64(n), C.c_void_p(rng_ptr),
                                                       C.c_void_p(u_ptr), _p(lam), C.c_void_p(out_ptr)))

    def graph_light_vector(self, rdesc, pos, *, vertex_radius, max_dist_to_center, light_iterations, points_on_radius,
                           sampling_resolution, lam, sampler_seed=0, list_index=None):
        """LightingCalculator::GetLightVector: L0 per vertex (positions in render = world space); the light and the
        medium's bounding shape come from the render desc."""
        pos = _f32(pos); n = pos.size // 3
        d = avr_light_vector_desc()
        for i in range(3):
            d.light_dir[i] = rdesc.c.light_dir[i]
        d.boundary_type = rdesc.c.boundary_type
        for i in range(6):
            d.boundary[i] = rdesc.c.boundary[i]
        d.max_dist_to_center, d.vertex_radius = max_dist_to_center, vertex_radius
        d.light_iterations, d.points_on_radius = light_iterations, points_on_radius
        d.sampler_seed, d.sampling_resolution = sampler_seed, sampling_resolution
        for i in range(4):
            d.lambda_[i] = float(lam[i])
        li = None if list_index is None else np.ascontiguousarray(list_index, dtype=np.int32)
        out = np.empty(n, dtype=np.float32)
        self.ctx.check(self.lib.avr_graph_light_vector(self.h, C.byref(d), _p(pos), _p(li, i32p) if li is not None else None,
                                                       C.c_int32(n), _p(out)))
        return out

    def build_free_graph(self, rdesc, *, vertex_radius, max_dist_to_center, bounds_center, dimension_steps,
                         iterations_per_step, max_depth, sampling_resolution, lam, search_range_neighbours=0, sampler_seed=0):
        """FreeGraphBuilder::BuildGraph on the device (avr_graph_build_free): dict of host arrays."""
        d = avr_graph_builder_desc()
        for i in range(3):
            d.light_dir[i] = rdesc.c.light_dir[i]
            d.bounds_center[i] = float(bounds_center[i])
        d.boundary_type = rdesc.c.boundary_type
        for i in range(6):
            d.boundary[i] = rdesc.c.boundary[i]
        d.max_dist_to_center, d.vertex_radius = max_dist_to_center, vertex_radius
        d.dimension_steps, d.iterations_per_step, d.max_depth = dimension_steps, iterations_per_step, max_depth
        d.sampler_seed, d.sampling_resolution, d.search_range_neighbours = sampler_seed, sampling_resolution, search_range_neighbours
        for i in range(4):
            d.lambda_[i] = float(lam[i])
        h = C.POINTER(avr_built_graph)()
        self.ctx.check(self.lib.avr_graph_build_free(self.h, C.byref(d), C.byref(h)))
        c = h.contents
        nv, ne = c.n_vertices, c.n_edges

        def arr(ptr, n, dt):
            if not ptr:
                return None
            return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dt, copy=True) if n else np.zeros(0, dt)
        out = dict(pos=arr(c.vertex_pos, 3 * nv, np.float32).reshape(-1, 3), vsamples=arr(c.vertex_samples, nv, np.int32),
                   vrange=arr(c.search_range, nv, np.float32), efrom=arr(c.edge_from, ne, np.int32),
                   eto=arr(c.edge_to, ne, np.int32), esamples=arr(c.edge_samples, ne, np.int32),
                   n_paths=c.n_paths, n_points=c.n_points, cluster_rounds=c.cluster_rounds)
        self.lib.avr_built_graph_free(h)
        return out

    def graph_light_rays(self, rays6, pixel, ts, iterations, seed, lam):
        rays6 = _f32(rays6); pixel = np.ascontiguousarray(pixel, dtype=np.int32); ts = _f32(ts); lam = _f32(lam)
        n = rays6.size // 6
        out = np.empty(n, dtype=np.float32)
        self.ctx.check(self.lib.avr_graph_light_rays(self.h, _p(rays6), _p(pixel, i32p), _p(ts), C.c_int64(n),
                                                     C.c_int32(iterations), C.c_int32(seed), _p(lam), _p(out)))
        return out


class Graph:
    def __init__(self, ctx, h, n, nnz):
        self.ctx, self.h, self.n, self.nnz = ctx, h, n, nnz
        self.lib = ctx.lib

    def close(self):
        if self.h:
            self.lib.avr_graph_destroy(self.h)
            self.h = None

    def csr_sizes(self):
        n = C.c_int32(); nnz = C.c_int64()
        self.ctx.check(self.lib.avr_graph_get_csr(self.h, C.byref(n), C.byref(nnz), None, None, None))
        return int(n.value), int(nnz.value)

    def csr(self):
        n, nnz = self.csr_sizes()
        rowptr = np.zeros(n + 1, np.int64); col = np.zeros(nnz, np.int32); val = np.zeros(nnz, np.float32)
        self.ctx.check(self.lib.avr_graph_get_csr(self.h, None, None, _p(rowptr, i64p), _p(col, i32p) if nnz else None,
                                                  _p(val) if nnz else None))
        return rowptr, col, val

    def compute_final_light(self, L0, bounces):  # LightingCalculator::ComputeFinalLight
        L0 = _f32(L0)
        total = np.empty(self.n, dtype=np.float32)
        it = C.c_int32(0)
        self.ctx.check(self.lib.avr_graph_propagate(self.h, _p(L0), C.c_int32(bounces), _p(total), C.byref(it)))
        return total, int(it.value)

    def propagate_device(self, L0_ptr, bounces, total_ptr, stream=None, want_iters=False):
        it = C.c_int32(-1)
        self.ctx.check(self.lib.avr_graph_propagate_device(self.h, C.c_void_p(L0_ptr), C.c_int32(bounces),
                                                           C.c_void_p(total_ptr), C.byref(it) if want_iters else None,
                                                           C.c_void_p(stream or 0)))
        return int(it.value)

    NOALLOC_MAX_VERTICES = 1 << 20  # AVR_NOALLOC_MAX_VERTICES

    def propagate_noalloc(self, L0_ptr, bounces, total_ptr, stream=None, want_iters=False):
        """avr_graph_propagate_noalloc: as propagate_device, scratch in the library's static device array
        (no allocation inside the call; at most NOALLOC_MAX_VERTICES vertices; calls must not overlap)."""
        it = C.c_int32(-1)
        self.ctx.check(self.lib.avr_graph_propagate_noalloc(self.h, C.c_void_p(L0_ptr), C.c_int32(bounces),
                                                            C.c_void_p(total_ptr), C.byref(it) if want_iters else None,
                                                            C.c_void_p(stream or 0)))
        return int(it.value)

    def spmv_rows_device(self, cur_in_ptr, cur_out_ptr, r0, r1, flag_ptr, stream=None):
        self.ctx.check(self.lib.avr_graph_spmv_rows_device(self.h, C.c_void_p(cur_in_ptr), C.c_void_p(cur_out_ptr),
                                                           C.c_int32(r0), C.c_int32(r1), C.c_void_p(flag_ptr),
                                                           C.c_void_p(stream or 0)))

    def axpy_rows_device(self, cur_ptr, total_ptr, r0, r1, flag_ptr, stream=None):
        self.ctx.check(self.lib.avr_graph_axpy_rows_device(self.h, C.c_void_p(cur_ptr), C.c_void_p(total_ptr),
                                                           C.c_int32(r0), C.c_int32(r1), C.c_void_p(flag_ptr),
                                                           C.c_void_p(stream or 0)))

    def compute_final_light_sharded(self, L0_dev, bounces, rank, world, group=None):
        """Row-sharded ComputeFinalLight over `world` ranks (one GPU each): every rank holds the whole CSR and
        multiplies its vertex range; slices travel by ncclAllGather (sharding.propagate_sharded).
        L0_dev: float32 CUDA tensor of n elements.  Returns (total tensor, iterations)."""
        import torch
        from . import sharding
        full = torch.empty(self.n, dtype=torch.float32, device=L0_dev.device)
        flag = torch.zeros(2, dtype=torch.int32, device=L0_dev.device)
        stream = torch.cuda.current_stream(L0_dev.device).cuda_stream

        def spmv_rows(cur, r0, r1):
            self.spmv_rows_device(cur.data_ptr(), full.data_ptr(), r0, r1, flag.data_ptr(), stream=stream)
            return full[r0:r1]

        return sharding.propagate_sharded(spmv_rows, L0_dev, bounces, rank, world, group)

    def set_vertices(self, pos, light, search_range, vertex_radius):
        pos = _f32(pos); light = _f32(light)
        rng = None if search_range is None else _f32(search_range)
        self.ctx.check(self.lib.avr_graph_set_vertices(self.h, _p(pos), _p(light), _p(rng) if rng is not None else None,
                                                       C.c_float(vertex_radius)))

    def connect(self, points):  # GraphIntegrator::ConnectToGraph
        points = _f32(points); n = points.size // 3
        out = np.empty(n, dtype=np.float32)
        self.ctx.check(self.lib.avr_graph_connect(self.h, _p(points), C.c_int64(n), _p(out)))
        return out
