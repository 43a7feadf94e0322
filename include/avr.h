/*
 * avr.h — C ABI of libavr_b200.so: the B200-native (sm_100a) volumetric hot path of
 * tsvdh/AcceleratedVolRenderer (a pbrt-v4 fork).
 *
 * Every entry point names the reference interface it replaces (file:line relative to the
 * reference tree).  Plain C: pointers, sizes, POD structs.  No C++/torch types cross this boundary.
 * Every call returns avr_status; nothing in the library exits or aborts (the reference's
 * ErrorExit/CHECK behaviour, util/error.h + util/check.h:36, is re-created by the host adapter,
 * which maps a non-zero status to ErrorExit).  There is NO CPU fallback: with no CUDA device every
 * compute entry point returns AVR_ERR_CUDA.
 *
 * Spectra cross the boundary the way the reference stores them after construction:
 * DenselySampledSpectrum tables (util/spectrum.h:374), AVR_NSPEC = 471 fp32 values for
 * lambda = 360..830 nm.
 */
#ifndef AVR_H
#define AVR_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AVR_NSPEC 471      /* Lambda_min=360 .. Lambda_max=830, util/spectrum.h:30-33 */
#define AVR_NLAMBDA 4      /* NSpectrumSamples, util/spectrum.h:36 */

typedef int avr_status;
enum {
    AVR_OK = 0,
    AVR_ERR_BAD_ARG = 1,
    AVR_ERR_OOM = 2,
    AVR_ERR_CUDA = 3,
    AVR_ERR_UNSUPPORTED = 4,
    AVR_ERR_NCCL = 5
};

typedef struct avr_ctx avr_ctx;
typedef struct avr_medium avr_medium;
typedef struct avr_graph avr_graph;
typedef struct avr_comm avr_comm;

/* ---- context -------------------------------------------------------------------------------- */
/* One context per GPU/process.  Owns streams, scratch and the last error string. */
avr_status avr_ctx_create(int device, avr_ctx **out);
void avr_ctx_destroy(avr_ctx *ctx);
/* Last error message of the context (or of the failed avr_ctx_create when ctx == NULL). */
const char *avr_last_error(const avr_ctx *ctx);
/* Library version / build string ("avr_b200 <n> sm_100a ..."). */
const char *avr_version(void);

/* Measurement aid (no counterpart in the reference): read bandwidth of a streaming 128-bit-load kernel over a
 * buffer of `bytes`, `iters` passes after a warm-up pass — the L2 peak of the density-fetch roofline with a
 * 64 MiB buffer (resident in the 126 MB L2), the HBM read peak with a multi-GB one. */
avr_status avr_measure_read_bandwidth(avr_ctx *ctx, int64_t bytes, int32_t iters, float *gb_per_s);

/* ---- media ---------------------------------------------------------------------------------- */
/* HBM layouts for scalar grids.  All give the same SampledGrid::Lookup tap values (fp16 layouts: rounded)
 * (util/containers.h:804-835); the octet layouts store, per interpolation cell, its 8 corner taps
 * contiguously (one 32-byte sector for fp32, one 16-byte vector for fp16) in 4x4x4-cell bricks. */
enum {
    AVR_LAYOUT_DENSE_F32 = 0,  /* reference layout: values[(z*ny+y)*nx+x], 8 bounds-checked taps */
    AVR_LAYOUT_OCTET_F32 = 1,  /* bit-identical taps, 1 sector per fetch                       */
    AVR_LAYOUT_OCTET_F16 = 2,  /* taps rounded to fp16 (perf option; NOT per-ray bit-parity)   */
    AVR_LAYOUT_BRICK_F16 = 3   /* sparse pool of 8x8x8-cell fp16 bricks behind an indirection grid: empty bricks cost
                                * nothing, ~1.1x the fp32 grid where occupied, no 2^32-cell limit, the dense grid is
                                * released after the build; taps rounded to fp16 like layout 2                    */
};

/* Replaces GridMedium::Create + ctor, media.cpp:212-336 ("uniformgrid").  Host pointers; the
 * library copies at create time (reference copies into pstd::vector, util/containers.h:771-774). */
typedef struct avr_grid_desc {
    int32_t nx, ny, nz;
    const float *density;      /* nx*ny*nz, x fastest                                         */
    const float *temperature;  /* NULL or nx*ny*nz ("temperature", media.cpp:250)             */
    const float *lescale;      /* NULL or nx*ny*nz ("Lescale", media.cpp:295), NOT yet *LeNorm */
    float le_norm;             /* LeNorm = 1/SpectrumToPhotometric(Le) or 1 (media.cpp:288-292) */
    float p0[3], p1[3];        /* medium-space bounds ("p0","p1")                             */
    float render_from_medium[16]; /* Transform::m,    row-major                               */
    float medium_from_render[16]; /* Transform::mInv, row-major                               */
    const float *sigma_a;      /* AVR_NSPEC, unscaled (ctor applies sigma_scale, media.cpp:229) */
    const float *sigma_s;      /* AVR_NSPEC                                                   */
    const float *le;           /* NULL or AVR_NSPEC ("Le")                                    */
    float sigma_scale;         /* "scale"                                                     */
    float g;                   /* HG asymmetry "g"                                            */
    float temperature_offset;  /* "temperatureoffset"                                         */
    float temperature_scale;   /* "temperaturescale"                                          */
    int32_t majorant_res[3];   /* {16,16,16} for uniformgrid (media.cpp:229); {64,64,64} is what
                                  NanoVDBMedium uses (media.cpp:521); 0 -> 16                  */
} avr_grid_desc;

/* Replaces RGBGridMedium::Create + ctor, media.cpp:339-459 ("rgbgrid").  Voxels arrive already
 * converted by the host with RGBUnboundedSpectrum(cs, rgb) (util/spectrum.cpp:236-240) as
 * {scale, c0, c1, c2} fp32x4 per voxel. */
typedef struct avr_rgbgrid_desc {
    int32_t nx, ny, nz;
    const float *sigma_a;      /* NULL or nx*ny*nz*4                                          */
    const float *sigma_s;      /* NULL or nx*ny*nz*4                                          */
    const float *le;           /* NULL or nx*ny*nz*4 (scale,c0,c1,c2 of RGBIlluminantSpectrum) */
    const float *le_illuminant;/* AVR_NSPEC: the colour space's illuminant, needed iff le      */
    float le_scale;            /* "Lescale"                                                   */
    float p0[3], p1[3];
    float render_from_medium[16];
    float medium_from_render[16];
    float sigma_scale, g;
    int32_t majorant_res[3];
} avr_rgbgrid_desc;

/* One NanoVDB FloatGrid flattened by the host the way the reference's own cmd/nanovdb2pbrt.cpp:100-110
 * does it (tree().getValue over the index bounding box), plus the pieces of nanovdb::GridData the
 * reference reads: the float inverse map of worldToIndexF (NanoVDB.h Map::applyInverseMapF:
 * mInvMatF, mVecF), indexBBox() (inclusive on both ends) and worldBBox() (doubles).  Voxels outside
 * the index box read as the background value, which must be 0 (fog volumes). */
typedef struct avr_nanovdb_grid {
    const float *values;       /* dims[0]*dims[1]*dims[2] floats, x fastest; NULL = grid absent   */
    int32_t index_min[3];      /* indexBBox().min()                                               */
    int32_t dims[3];           /* indexBBox().max() - min() + 1                                   */
    float inv_mat[9];          /* Map::mInvMatF, row-major                                        */
    float vec[3];              /* Map::mVecF                                                      */
    double world_bbox[6];      /* worldBBox(): min xyz, max xyz                                   */
    /* Sparse alternative to `values` (values == NULL, block_values != NULL): the grid as n_blocks blocks of 8^3
     * voxels aligned to multiples of 8 in index space — NanoVDB's leaf nodes as they are, plus its non-zero value
     * tiles cut into such blocks.  Block b covers [block_origin[3 b + i], + 8) on axis i and holds
     * block_values[512 b + x + 8 y + 64 z]; voxels no block covers are 0; what lies outside the index box is not
     * read.  Accepted by avr_medium_create_nanovdb for the density grid with layout AVR_LAYOUT_BRICK_F16 and no
     * temperature grid: the brick pool and the majorant grid are then built straight from the blocks and the
     * index box is never held as a dense array, on the host or on the device. */
    int64_t n_blocks;
    const int32_t *block_origin;
    const float *block_values;
} avr_nanovdb_grid;

/* Replaces NanoVDBMedium::Create + ctor, media.cpp:511-665 ("nanovdb"; what pbrt-v4-scenes' Disney
 * cloud uses).  Sampling follows nanovdb::SampleFromVoxels<Tree,1,false> (node-centred, lerp
 * a + w*(b-a), order z,y,x; SampleFromVoxels.h:226-238), NOT SampledGrid::Lookup; the majorant grid
 * is 64^3 built as in media.cpp:560-616. */
typedef struct avr_nanovdb_desc {
    avr_nanovdb_grid density;      /* "gridname" (default "density")                              */
    avr_nanovdb_grid temperature;  /* "temperaturename"; values == NULL when the file has none     */
    float render_from_medium[16];
    float medium_from_render[16];
    const float *sigma_a;      /* AVR_NSPEC, unscaled (ConstantSpectrum(1) when the scene gives none) */
    const float *sigma_s;      /* AVR_NSPEC                                                   */
    float sigma_scale;         /* "scale"                                                     */
    float g;
    float le_scale;            /* "Lescale"                                                   */
    float temperature_offset;  /* "temperatureoffset" / "temperaturecutoff"                   */
    float temperature_scale;   /* "temperaturescale"                                          */
    int32_t majorant_res[3];   /* 0 -> {64,64,64} (media.cpp:521)                             */
} avr_nanovdb_desc;

avr_status avr_medium_create_uniformgrid(avr_ctx *ctx, const avr_grid_desc *desc, int layout,
                                         avr_medium **out);
/* layout: as for uniformgrid (the octet layouts hold the 8 stencil taps of one index-space cell). */
avr_status avr_medium_create_nanovdb(avr_ctx *ctx, const avr_nanovdb_desc *desc, int layout,
                                     avr_medium **out);
avr_status avr_medium_create_rgbgrid(avr_ctx *ctx, const avr_rgbgrid_desc *desc, avr_medium **out);

/* ".nvdb" ingest: the file part of NanoVDBMedium::Create (media.cpp:628-665) and its readGrid (:487-509).
 * avr_nvdb_read_grid finds the FloatGrid called `grid_name` (NULL or "" = the first grid) in a NanoVDB file
 * (file-format major version 32, the one the reference vendors; codecs NONE and ZIP) and flattens it over its index
 * bounding box — the result the fork's cmd/nanovdb2pbrt.cpp:100-110 gets from tree().getValue() per voxel, produced
 * by one walk over the tree (leaves copied, value tiles filled).  grid->values is allocated by the library
 * (avr_nvdb_free_grid); a file without that grid leaves grid->values == NULL and returns AVR_OK, as the reference's
 * empty GridHandle.  Grids that are not fog volumes are refused with the reference's message.  Host code only. */
typedef struct avr_nvdb_info {
    int32_t grids_in_file, codec;
    int32_t leaf_nodes, lower_nodes, upper_nodes;
    int64_t active_voxels, grid_bytes;
    char grid_name[64];
} avr_nvdb_info;
avr_status avr_nvdb_read_grid(const char *path, const char *grid_name, avr_nanovdb_grid *grid, avr_nvdb_info *info /* or NULL */);
/* The same grid WITHOUT flattening: grid->values stays NULL and the sparse fields of avr_nanovdb_grid are filled —
 * one block per leaf node that holds a non-zero value (transposed from NanoVDB's z-fastest order), value tiles with a
 * non-zero value cut into 8^3 blocks (at most 2^22 of them: AVR_ERR_UNSUPPORTED beyond, use avr_nvdb_read_grid).  A
 * file without that grid leaves block_values == NULL too.  Host memory is proportional to the leaves, not to the
 * index box. */
avr_status avr_nvdb_read_blocks(const char *path, const char *grid_name, avr_nanovdb_grid *grid, avr_nvdb_info *info /* or NULL */);
void avr_nvdb_free_grid(avr_nanovdb_grid *grid);
/* NanoVDBMedium::Create from the file: "filename", "gridname" (NULL = "density"), "temperaturename" (NULL =
 * "temperature"; the grid may be absent); every other parameter comes from `params`, whose two grids are ignored.
 * With layout AVR_LAYOUT_BRICK_F16 and no temperature grid in the file the density grid goes from the file's leaf
 * nodes into the brick pool (avr_nvdb_read_blocks); otherwise through the flattened box (avr_nvdb_read_grid). */
avr_status avr_medium_create_nanovdb_file(avr_ctx *ctx, const char *filename, const char *gridname,
                                          const char *temperaturename, const avr_nanovdb_desc *params, int layout,
                                          avr_medium **out);
void avr_medium_destroy(avr_medium *m);

/* Medium::IsEmissive, media.h:284/374. */
int avr_medium_is_emissive(const avr_medium *m);
/* Bytes of HBM held by the medium's grids (the STAT_MEMORY_COUNTER "Memory/Volume grids"). */
int64_t avr_medium_bytes(const avr_medium *m);

/* Copies the device-built MajorantGrid (media.h:105-133; built as in media.cpp:240-246 /
 * :364-377) to host: out has res[0]*res[1]*res[2] floats, x fastest.  Bit-exact contract. */
avr_status avr_medium_majorants(avr_medium *m, float *out, int32_t res[3]);

/* Medium-space bounds: GridMedium/RGBGridMedium "p0","p1"; NanoVDBMedium::GetBounds() = the union of
 * the grids' worldBBox() narrowed to float (media.h:616, media.cpp:536-551).  out: min xyz, max xyz. */
avr_status avr_medium_bounds(const avr_medium *m, float out[6]);

/* SampledGrid<Float>::Lookup(Point3f), util/containers.h:804-819, on the medium's density grid.
 * p in grid-normalised [0,1]^3 coordinates (after bounds.Offset), 3 floats per point.  For a nanovdb
 * medium: nanovdb::SampleFromVoxels<Tree,1,false>::operator()(xyz), p in INDEX space. */
avr_status avr_medium_lookup_density(avr_medium *m, const float *p, int64_t n, float *out);

/* MediumProperties Medium::SamplePoint(p, lambda), media.h:287-319 / :377-401.  p in render space,
 * lambda: 4 wavelengths shared by all points.  out: 12 floats per point {sigma_a[4], sigma_s[4],
 * Le[4]}. */
avr_status avr_medium_sample_point(avr_medium *m, const float *p, int64_t n, const float lambda[4],
                                   float *out);

/* ---- kernel 2: DDA traversal ------------------------------------------------------------------ */
/* DDAMajorantIterator ctor + Next(), media.h:141-203.  rays: 8 floats each {o.xyz, d.xyz, tMin,
 * tMax} in MEDIUM space (what GridMedium::SampleRay hands the iterator after ApplyInverse + slab
 * test, media.h:322-337).  Emits up to max_per_ray segments per ray:
 *   seg_t[(i*max_per_ray+k)*2 + {0,1}] = tMin,tMax ; seg_sigma[...] = sigma_t * grid(voxel) ;
 *   seg_voxel[...] = x + res.x*(y + res.y*z) ;  counts[i] = number of segments the reference
 * iterator yields (may exceed max_per_ray; extra segments are counted, not stored).
 * Bit-exact with the reference (integer voxels, fp32 t values). */
avr_status avr_dda_segments(avr_medium *m, const float *rays, int64_t n, float sigma_t,
                            int32_t max_per_ray, int32_t *counts, float *seg_t, float *seg_sigma,
                            int32_t *seg_voxel);

/* {Grid,RGBGrid}Medium::SampleRay prologue, media.h:322-337: Transform::ApplyInverse(Ray,&tMax)
 * (util/transform.h:420-433, interval-arithmetic origin offset) + Bounds3::IntersectP slab test.
 * rays: 7 floats {o.xyz, d.xyz, tMax} render space.  out: 9 floats {o.xyz, d.xyz, tMin, tMax,
 * hit(0/1)} medium space. */
avr_status avr_medium_ray_to_medium(avr_medium *m, const float *rays, int64_t n, float *out);

/* ---- kernel 1a: majorant free-flight / transmittance sampling -------------------------------- */
enum {
    /* ratio tracking, channel 0: util::SampleTransmittance, src/graph/util.h:344-366 */
    AVR_TMAJ_RATIO = 0,
    /* delta tracking to the first real event: util::SampleScatter, src/graph/util.h:368-403, and
       GraphIntegrator::Li's callback, src/graph/graph_integrator.cpp:201-230 */
    AVR_TMAJ_DELTA = 1,
    /* chromatic ratio tracking with Russian roulette: the shadow-ray loop of
       VolPathIntegrator::SampleLd, cpu/integrators.cpp:1352-1380 */
    AVR_TMAJ_SHADOW = 2
};

typedef struct avr_tmaj_result {
    float T_maj[4];     /* return value of SampleT_maj (media.h:741-806)                        */
    float Tr[4];        /* RATIO: Tr in [0]; SHADOW: T_ray[4]; DELTA: unused (1)                 */
    float r_u[4];       /* SHADOW: r_u ; else 1                                                  */
    float r_l[4];       /* SHADOW: r_l ; else 1                                                  */
    float t_event;      /* DELTA: ray parameter t of the terminating event (normalised ray), else -1 */
    int32_t event;      /* DELTA: 0 absorbed, 1 scattered, 2 escaped ; else 2                    */
    int32_t n_fetch;    /* number of SamplePoint calls (density fetches) made                    */
    int32_t n_seg;      /* number of majorant segments consumed                                  */
} avr_tmaj_result;

/* SampleT_maj<GridMedium>(ray, tMax, u, rng, lambda, callback), media.h:741-806.
 * rays: 7 floats {o.xyz, d.xyz (not necessarily normalised), tMax}, render space.
 * rng: 2 uint64 per ray {sequenceIndex, seed} -> RNG(seq, seed) (util/rng.h:124-130);
 * u: first exponential sample per ray; lambda: 4 wavelengths, shared by the batch. */
avr_status avr_sample_tmaj(avr_medium *m, int mode, const float *rays, int64_t n,
                           const uint64_t *rng, const float *u, const float lambda[4],
                           avr_tmaj_result *out);

/* Same on DEVICE pointers (no copies; asynchronous on the context's stream).  exact_math: 1 = the
 * restated CPU arithmetic of the parity contract, 0 = fast intrinsics + folded transforms (the
 * reference's own GPU branch uses __expf, util/math.h:451-452). */
avr_status avr_sample_tmaj_device(avr_medium *m, int mode, int exact_math, const float *rays_dev, int64_t n,
                                  const uint64_t *rng_dev, const float *u_dev, const float lambda[4],
                                  avr_tmaj_result *out_dev);

/* ---- kernel 1b: volumetric path tracing -------------------------------------------------------- */
/* Replaces Integrator::Create("volpath"/"volpathcustom") + ImageTileIntegrator::Render
 * (cpu/integrators.cpp:72-232), RayIntegrator::EvaluatePixelSample (:235-298),
 * VolPathIntegrator::Li volume branch (:962-1106) and SampleLd (:1282-1399) for the scene class the
 * fork's graph pipeline accepts (src/graph/util.h:61-132): ONE grid medium inside ONE convex
 * boundary shape with Material "interface", ONE distant light, perspective camera, box pixel filter or a
 * FilterSampler-tabulated one (gaussian / mitchell / sinc),
 * Sampler "independent" or "zsobol". */
enum { AVR_BOUNDARY_SPHERE = 0, AVR_BOUNDARY_BOX = 1 };
enum { AVR_SAMPLER_INDEPENDENT = 0, AVR_SAMPLER_ZSOBOL = 1 };
/* RandomizeStrategy, util/lowdiscrepancy.h:261 (same order) */
enum { AVR_RANDOMIZE_NONE = 0, AVR_RANDOMIZE_PERMUTE_DIGITS = 1, AVR_RANDOMIZE_FAST_OWEN = 2, AVR_RANDOMIZE_OWEN = 3 };

/* A pixel filter that pbrt samples through its FilterSampler (filters.h:26-45, filters.cpp:133-147): "gaussian"
 * (pbrt's default, scene.cpp:94), "mitchell", "sinc".  f = the filter tabulated as the FilterSampler constructor
 * does it: nx = int(32 * radius.x), ny = int(32 * radius.y) values f(x, y) = filter.Evaluate(p) at
 * p = domain.Lerp(((x + .5) / nx, (y + .5) / ny)), domain = [-radius, radius]^2, row-major (x fastest); the
 * radius is avr_render_desc.filter_radius.  The library builds the PiecewiseConstant2D distribution from it
 * (util/sampling.h:625-649,737-754) and samples it per camera sample exactly as FilterSampler::Sample /
 * GetCameraSample do (samplers.h:797-814): pFilm = pPixel + p + 0.5, the sample is added to pixel pPixel alone with
 * weight f[pi] / pdf (RGBFilm::AddSample, film.h:239-255: rgbSum += weight * rgb, weightSum += weight). */
/* kind AVR_FILTER_TRIANGLE: pbrt's "triangle" filter, sampled analytically (TriangleFilter::Sample, filters.h:224-227:
 * SampleTent per axis, util/sampling.h:196-201, weight 1); nx, ny and f are not read. */
enum { AVR_FILTER_TABULATED = 0, AVR_FILTER_TRIANGLE = 1 };
typedef struct avr_filter_table {
    int32_t nx, ny;
    const float *f;                  /* ny * nx floats, host memory */
    int32_t kind;                    /* AVR_FILTER_TABULATED (0) or AVR_FILTER_TRIANGLE */
} avr_filter_table;

typedef struct avr_render_desc {
    int32_t width, height;           /* full film resolution                                   */
    int32_t pixel_bounds[4];         /* x0,y0,x1,y1 (exclusive); all 0 -> full frame           */
    int32_t spp;                     /* samplesPerPixel                                        */
    int32_t sample_begin, sample_end;/* sample-index range rendered by this call (multi-GPU
                                        sharding by sample index); both 0 -> [0, spp)          */
    int32_t max_depth;               /* "maxdepth" (default 5, cpu/integrators.cpp:1410)       */
    int32_t sampler_seed;            /* the sampler's "seed" (IndependentSampler / ZSobolSampler) */
    int32_t exact_math;              /* 1: restated CPU FastExp/logf/IEEE div (parity mode);
                                        0: ex2.approx / lg2.approx (reference's own GPU branch,
                                        util/math.h:451-452)                                   */
    float filter_radius[2];          /* pixel filter radius (box default 0.5; gaussian default 1.5) */
    float camera_from_raster[16];    /* PerspectiveCamera, cameras.cpp:404-460                 */
    float render_from_camera[16];
    float light_dir[3];              /* direction TOWARDS the distant light (wi), render space */
    const float *light_spectrum;     /* AVR_NSPEC, already multiplied by the light's scale     */
    float scene_radius;              /* DistantLight::Preprocess sceneRadius, lights.h:284-291 */
    int32_t boundary_type;
    float boundary[6];               /* sphere: cx,cy,cz,r ; box: min xyz, max xyz (render sp.) */
    const float *sensor_rgb;         /* 3*AVR_NSPEC: r_bar,g_bar,b_bar of PixelSensor, film.h:95 */
    float imaging_ratio;
    float output_rgb_from_sensor[9]; /* row-major 3x3 (film.h:257-275)                         */
    float max_component_value;       /* RGBFilm clamp, film.h:246                              */
    int32_t rank, world;             /* tile/sample partition: this call renders work items
                                        with (item % world) == rank; world 0 -> 1              */
    int32_t sampler;                 /* AVR_SAMPLER_INDEPENDENT ("independent", samplers.h:442-476)
                                        or AVR_SAMPLER_ZSOBOL ("zsobol", pbrt's default sampler,
                                        samplers.h:225-364: constructed from spp, the full film
                                        resolution, `randomize` and sampler_seed)               */
    int32_t randomize;               /* zsobol "randomization": AVR_RANDOMIZE_* (pbrt's default
                                        is fastowen, samplers.cpp:157); ignored for independent */
    const avr_filter_table *filter;  /* NULL: BoxFilter of filter_radius (weight 1); else a tabulated or triangle filter */
} avr_render_desc;

/* FilterSampler::Sample for n points u (2 floats each) under `filter` / radius: out = {p.x, p.y, weight} per
 * point, computed by the device code the render kernels run (parity probe). */
avr_status avr_filter_sample(avr_ctx *ctx, const avr_filter_table *filter, const float radius[2], const float *u,
                             int64_t n, float *out);

typedef struct avr_render_stats {
    int64_t camera_rays;       /* pixel samples evaluated                                       */
    int64_t camera_fetches;    /* "Integrator/Volume interactions" (cpu/integrators.cpp:958,997) */
    int64_t shadow_fetches;    /* SamplePoint calls on NEE shadow rays                           */
    int64_t segments;          /* DDA segments consumed (camera + shadow)                        */
    int64_t scatter_events;    /* real scattering events                                         */
    float kernel_ms;           /* device time of the render kernel(s), CUDA events              */
    int32_t launches;          /* kernels launched by this call                                 */
} avr_render_stats;

/* film: width*height*4 doubles {rgbSum[3], weightSum} per pixel (RGBFilm::Pixel, film.h:302-307),
 * host memory.  The call ADDS into film (so sample shards/ranks can be accumulated). */
avr_status avr_render_volpath(avr_ctx *ctx, avr_medium *m, const avr_render_desc *desc,
                              double *film, avr_render_stats *stats);
/* Same, film is a DEVICE pointer (for NCCL reduction by the caller); asynchronous on `stream`
 * (a cudaStream_t passed as void*, NULL = the context's stream).  stats may be NULL. */
avr_status avr_render_volpath_device(avr_ctx *ctx, avr_medium *m, const avr_render_desc *desc,
                                     double *film_dev, void *stream, avr_render_stats *stats);
/* Per-sample radiance dump for parity tests: L[4] + lambda[4] + pdf[4] per (pixel, sample) in
 * pixel_bounds x [sample_begin, sample_end), pixel-major.  Host memory, 12 floats per sample. */
avr_status avr_render_volpath_samples(avr_ctx *ctx, avr_medium *m, const avr_render_desc *desc,
                                      float *out, avr_render_stats *stats);
/* RGBFilm::GetPixelRGB, film.h:257-275: rgb = outputRGBFromSensorRGB * rgbSum/weightSum. */
avr_status avr_film_resolve(const avr_render_desc *desc, const double *film, float *rgb);

/* ---- kernel 3: radiance-graph propagation ----------------------------------------------------- */
/* Replaces LightingCalculator::GetTransportMatrix (src/graph/lighting_calculator.cpp:61-82): the
 * host hands over CSR (row = vertex, col = out-edge target, val = edge.samples/vertex.samples;
 * duplicates already summed as Eigen::setFromTriplets does). */
avr_status avr_graph_create(avr_ctx *ctx, int32_t n_vertices, int64_t nnz, const int64_t *rowptr,
                            const int32_t *col, const float *val, avr_graph **out);
void avr_graph_destroy(avr_graph *g);

/* LightingCalculator::GetTransportMatrix itself (src/graph/lighting_calculator.cpp:61-82) on the
 * device, from what a graph file / FreeGraph holds: per-vertex `samples` and the edge list
 * {from, to, samples}.  Follows Graph::AddEdge (src/graph/graph.cpp:202-235): edges whose endpoints are
 * not vertices (ids outside [0, n_vertices)) are dropped, edges with the same (from, to) are merged by
 * adding their samples; val = float(edge samples) / float(vertex samples of `from`).  Rows come out
 * with ascending column indices (std::map outEdges / Eigen column order). */
avr_status avr_graph_create_from_edges(avr_ctx *ctx, int32_t n_vertices, const int32_t *vertex_samples,
                                       int64_t n_edges, const int32_t *edge_from, const int32_t *edge_to,
                                       const int32_t *edge_samples, avr_graph **out);
/* Size and contents of the device CSR (parity surface for the call above; buffers may be NULL). */
avr_status avr_graph_get_csr(avr_graph *g, int32_t *n_vertices, int64_t *nnz, int64_t *rowptr, int32_t *col,
                             float *val);

/* LightingCalculator::ComputeFinalLight (src/graph/lighting_calculator.cpp:23-59):
 * total = L0; cur = L0; repeat `bounces`: cur = T*cur; stop BEFORE adding if any NaN/Inf;
 * total += cur.  iters_done = the function's return value. */
avr_status avr_graph_propagate(avr_graph *g, const float *L0, int32_t bounces, float *total,
                               int32_t *iters_done);
/* Same with L0/total resident on the device; iters_done may be NULL (then the call is asynchronous
 * on `stream`, NULL = the context's stream). */
avr_status avr_graph_propagate_device(avr_graph *g, const float *L0_dev, int32_t bounces, float *total_dev,
                                      int32_t *iters_done, void *stream);
/* ComputeFinalLight as avr_graph_propagate_device, for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  NO counterpart in the reference.
 * The call allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine): the ping-pong vectors and the NaN/Inf flags live in one file-scope
 * `static __device__` array of 2 * AVR_NOALLOC_MAX_VERTICES floats + 256 bytes (8 MiB + 256 B), which
 * every GPU that loads the library holds until exit.  Graphs with more vertices are refused with
 * AVR_ERR_UNSUPPORTED and a message naming the limit.  All calls on one GPU share that array and it keeps
 * the last call's contents (the call rewrites what must start clean): calls that overlap on the GPU —
 * two streams, two contexts, two host threads — RACE; serialise them.  L0_dev / total_dev: n floats on
 * the device, caller-owned; iters_done: host int or NULL (NULL = fully asynchronous on `stream`). */
#define AVR_NOALLOC_MAX_VERTICES 1048576
avr_status avr_graph_propagate_noalloc(avr_graph *g, const float *L0_dev, int32_t bounces, float *total_dev,
                                       int32_t *iters_done, void *stream);
/* Row-range shard of one iteration on device pointers, for the multi-GPU path (caller all-gathers
 * `cur_out` rows with NCCL): cur_out[r] = sum_j T[r,j]*cur_in[j] over rows [row_begin,row_end); nothing is
 * added to any total here (that is avr_graph_axpy_rows_device).  bad_flag_dev points to TWO device ints
 * {raised, first bad iteration}, which the caller initialises to {0, INT_MAX}: a NaN/Inf result sets
 * raised = 1 and lowers the second word to 0 (this entry always runs as iteration 0).  axpy:
 * total[r] += cur[r] unless bad_flag_dev[0] is raised.  Asynchronous on `stream`. */
avr_status avr_graph_spmv_rows_device(avr_graph *g, const float *cur_in_dev, float *cur_out_dev,
                                      int32_t row_begin, int32_t row_end, int32_t *bad_flag_dev,
                                      void *stream);
avr_status avr_graph_axpy_rows_device(avr_graph *g, const float *cur_dev, float *total_dev,
                                      int32_t row_begin, int32_t row_end, const int32_t *bad_flag_dev,
                                      void *stream);

/* Vertex data for rendering (what FreeGraph::ReadFromDisk yields, src/graph/graph.cpp:657 +
 * GraphIntegrator::Initialize, src/graph/graph_integrator.cpp:24-81): positions (world space),
 * lightScalar, optional renderSearchRange; builds the device spatial index. */
avr_status avr_graph_set_vertices(avr_graph *g, const float *pos, const float *light_scalar,
                                  const float *search_range /* or NULL */, float vertex_radius);
/* GraphIntegrator::ConnectToGraph, src/graph/graph_integrator.cpp:249-280: radius search with the
 * perc99 / max fallbacks and 1/dist^2 weighted mean.  points in WORLD space, 3 floats each. */
avr_status avr_graph_connect(avr_graph *g, const float *points, int64_t n, float *out);
/* GraphIntegrator::Li free-graph branch (src/graph/graph_integrator.cpp:180-244) rendered over the
 * same scene class as avr_render_volpath (Sampler "independent" or "zsobol"); world_from_render: 16 floats row-major.
 * desc->filter (a tabulated pixel filter) is honoured as in avr_render_volpath. */
avr_status avr_render_graph(avr_ctx *ctx, avr_medium *m, avr_graph *g, const avr_render_desc *desc,
                            const float *world_from_render, double *film, avr_render_stats *stats);

/* LightingCalculator::GetLightVector inner estimator (src/graph/lighting_calculator.cpp:84-155 via
 * ComputeRaysToSphere, src/graph/util.h:814-839): for each ray, `iterations` ratio-tracking
 * estimates with IndependentSampler streams StartPixelSample({x,y}, i) and
 * tMax_i = start_scatter_t + dist_in_sphere * Get1D(); out[i] = mean transmittance.
 * rays: 6 floats {o, d}; pixel: 2 ints per ray; ts: 2 floats {start_scatter_t, dist_in_sphere}. */
avr_status avr_graph_light_rays(avr_medium *m, const float *rays, const int32_t *pixel,
                                const float *ts, int64_t n, int32_t iterations, int32_t sampler_seed,
                                const float lambda[4], float *out);

/* LightingCalculator::GetLightVector (src/graph/lighting_calculator.cpp:84-155) whole: for every vertex the disk of
 * rays along the light direction through the vertex sphere (util::GetDiskPoints, src/graph/util.h:179-204), each
 * clipped against the medium's bounding shape and the vertex sphere (GetHits / GetStartEndT, :419-503), estimated
 * by ComputeRaysToSphere (:814-839) with `light_iterations` ratio-tracking samples, averaged over the rays that
 * qualify, times Inv4Pi.  pos: n vertex positions (render = world space); list_index: position of each vertex in the
 * reference's iteration order over its vertex map (it picks the sampler pixel, lighting_calculator.cpp:103,115),
 * NULL = the vertex id itself; L0: n floats.  The bounding shape is hit analytically (the reference intersects
 * its Shape and respawns the ray): hit distances agree to rounding, parity of L0 is by tolerance. */
typedef struct avr_light_vector_desc {
    float light_dir[3];          /* direction TOWARDS the distant light (as avr_render_desc.light_dir)           */
    int32_t boundary_type;       /* AVR_BOUNDARY_*: the medium's bounding primitive                             */
    float boundary[6];
    float max_dist_to_center;    /* util::PrimitiveData::maxDistToCenter = |bounds.Diagonal() / 2| (util.h:45-60) */
    float vertex_radius;         /* FreeGraph::GetVertexRadius()                                                 */
    int32_t light_iterations;    /* LightingCalculatorConfig::lightIterations                                    */
    int32_t points_on_radius;    /* ...::pointsOnRadiusLight                                                     */
    int32_t sampler_seed;        /* the IndependentSampler's seed                                                */
    int32_t sampling_resolution; /* Options->graph.samplingResolution.x (cmd/graph_maker.cpp:86-107)             */
    float lambda[4];             /* MediumData::defaultLambda                                                    */
} avr_light_vector_desc;
avr_status avr_graph_light_vector(avr_medium *m, const avr_light_vector_desc *desc, const float *pos,
                                  const int32_t *list_index, int32_t n, float *L0);

/* FreeGraphBuilder::BuildGraph (src/graph/free/free_graph_builder.cpp:143-214) on the device: the light paths of
 * TracePath (:19-141) traced one per thread under the reference's sampler streams, then the reference's first-come
 * merge of scatter points into vertices (a point is a new vertex iff no EXISTING vertex lies within the vertex
 * radius, else it joins the closest) computed exactly as a lexicographically-first maximal independent set in
 * parallel rounds over a hash grid, edges = consecutive points of a path (samples 1 each; duplicates are merged by
 * avr_graph_create_from_edges as Graph::AddEdge does), vertex samples = departures, and ComputeSearchRanges
 * (:498-548).  ReinforceSparseVertices (:281-471) is not run (the reference's configs switch it off by default).
 * Scatter positions differ from the reference's in the last bits (analytic boundary hits), so the graph is
 * statistically, not vertex-for-vertex, the reference's.  *out is allocated by the library. */
typedef struct avr_graph_builder_desc {
    float light_dir[3];          /* direction TOWARDS the distant light                                          */
    int32_t boundary_type;
    float boundary[6];
    float bounds_center[3];      /* util::PrimitiveData::boundsCenter                                            */
    float max_dist_to_center;    /* util::PrimitiveData::maxDistToCenter                                         */
    float vertex_radius;         /* GetSameSpotRadius(mediumData) * radiusModifier (or --node-radius)            */
    int32_t dimension_steps, iterations_per_step, max_depth;   /* GraphBuilderConfig                             */
    int32_t sampler_seed, sampling_resolution;
    int32_t search_range_neighbours;   /* renderSearchRange.neighboursToUse; 0 = inactive                        */
    float lambda[4];
} avr_graph_builder_desc;
typedef struct avr_built_graph {
    int32_t n_vertices, n_paths, n_points, cluster_rounds;
    int64_t n_edges;             /* unmerged: one per pair of consecutive scatter points                         */
    float *vertex_pos;           /* 3 * n_vertices                                                               */
    int32_t *vertex_samples;
    float *search_range;         /* n_vertices, or NULL                                                          */
    int32_t *edge_from, *edge_to, *edge_samples;
} avr_built_graph;
avr_status avr_graph_build_free(avr_medium *m, const avr_graph_builder_desc *desc, avr_built_graph **out);
void avr_built_graph_free(avr_built_graph *g);

/* ---- graph text files (src/graph/graph.cpp:284-416 WriteToStream/ReadFromStream, :644-663 FreeGraph;
 * layout in files/format.txt) ----------------------------------------------------------------------
 *   line 1: description (one token, e.g. "basic")      line 2: "free" vertexRadius
 *   line 3: five True/False flags: useCoors useSamples useRayVertexTypes useLighting useRenderSearchRange
 *   line 4: curVertexId curEdgeId curPathId numVertices numEdges numPaths
 *   vertices: id x y z [type] [lightScalar] [samples] [renderSearchRange] [cx cy cz]
 *   edges:    id from to [samples]            paths: id length v0 v1 ...
 * Floats are written as the reference's `ostream << float` does (6 significant digits, %g).  Arrays
 * that a flag switches off are NULL.  Pure host code (no CUDA device needed). */
typedef struct avr_graph_file {
    char description[64];
    float vertex_radius;
    int32_t use_coors, use_samples, use_ray_vertex_types, use_lighting, use_render_search_range;
    int32_t cur_vertex_id, cur_edge_id, cur_path_id;
    int32_t n_vertices, n_edges, n_paths;
    int32_t *vertex_id;        /* n_vertices                                                   */
    float *vertex_pos;         /* 3 * n_vertices                                               */
    int32_t *vertex_type;      /* RayVertexType, src/graph/graph.h:38-44                       */
    float *vertex_light;       /* lightScalar                                                  */
    int32_t *vertex_samples;
    float *vertex_range;       /* renderSearchRange                                            */
    int32_t *vertex_coors;     /* 3 * n_vertices                                               */
    int32_t *edge_id, *edge_from, *edge_to, *edge_samples;
    int32_t *path_id;
    int64_t *path_offset;      /* n_paths + 1                                                  */
    int32_t *path_vertices;    /* path_offset[n_paths]                                         */
} avr_graph_file;

/* FreeGraph::ReadFromDisk (graph.cpp:657-663).  *out is allocated by the library. */
avr_status avr_graph_file_read(const char *path, avr_graph_file **out);
/* Graph::WriteToDisk + FreeGraph::WriteToStream (graph.cpp:419-433, :644-648). */
avr_status avr_graph_file_write(const char *path, const avr_graph_file *g);
void avr_graph_file_free(avr_graph_file *g);

/* ---- binary sidecar of a graph -------------------------------------------------
 * `graph_maker` keeps the edge weights only in memory (cmd/graph_maker.cpp:167-169: lightVec and
 * transportMat are built, used and dropped; the files it writes hold vertices only), so the
 * propagation can only be fed from inside that tool.  The sidecar persists exactly what kernel 3 and
 * the graph render consume: the transport matrix in CSR (as avr_graph_get_csr returns it), and per
 * vertex the position, the light vector entry L0 (LightingCalculator::GetLightVector) and the
 * render search range.  Little-endian file "<name>.avrg":
 *   char magic[4] = "AVRG"; u32 version = 1; i32 n_vertices; u32 flags (1 pos, 2 light, 4 range);
 *   i64 nnz; i64 rowptr[n+1]; i32 col[nnz]; f32 val[nnz]; f32 pos[3n]; f32 light[n]; f32 range[n]
 *   (absent arrays skipped); u64 FNV-1a of every preceding byte.
 * Pure host code.  Arrays not present are NULL. */
typedef struct avr_graph_sidecar {
    int32_t n_vertices;
    int64_t nnz;
    int64_t *rowptr;      /* n_vertices + 1 */
    int32_t *col;         /* nnz            */
    float *val;           /* nnz: edge.samples / vertex.samples (lighting_calculator.cpp:74-77) */
    float *vertex_pos;    /* 3 * n_vertices, or NULL */
    float *light;         /* n_vertices: L0, or NULL  */
    float *search_range;  /* n_vertices: renderSearchRange, or NULL */
} avr_graph_sidecar;

avr_status avr_graph_sidecar_write(const char *path, const avr_graph_sidecar *s);
/* *out is allocated by the library; a truncated file, a bad checksum or an inconsistent CSR
 * (rowptr not monotone, col outside [0, n)) is AVR_ERR_BAD_ARG. */
avr_status avr_graph_sidecar_read(const char *path, avr_graph_sidecar **out);
void avr_graph_sidecar_free(avr_graph_sidecar *s);

/* ---- multi-GPU (SURVEY.md 8(b) avr_comm_init, 8(e)) ------------------------------------------------------
 * One rank per GPU: a process (MPI / torchrun) or a host thread, each with its own avr_ctx.  The reference
 * fans its tile loop out over host threads that all add into one film (ParallelFor2D,
 * cpu/integrators.cpp:168-226) and runs ComputeFinalLight on one core (src/graph/lighting_calculator.cpp:23-59);
 * here the ranks talk through NCCL over NVLink / NVSwitch.  NCCL is bound at run time (libnccl.so.2, the copy
 * already loaded in the process if there is one); a host without it gets AVR_ERR_NCCL from these entries and
 * loses nothing else.
 *   processes: rank 0 calls avr_comm_unique_id and hands the bytes to the others over any channel it has
 *              (MPI_Bcast, a file, torch.distributed), then every rank calls avr_comm_init_rank;
 *   threads:   avr_comm_init_all(ctx[], n, comm[]) (ncclCommInitAll), then thread i uses comm[i]. */
#define AVR_COMM_ID_BYTES 128
avr_status avr_comm_unique_id(uint8_t id[AVR_COMM_ID_BYTES]);
avr_status avr_comm_init_rank(avr_ctx *ctx, const uint8_t id[AVR_COMM_ID_BYTES], int32_t rank, int32_t world,
                              avr_comm **out);
avr_status avr_comm_init_all(avr_ctx **ctx, int32_t n, avr_comm **out);
void avr_comm_destroy(avr_comm *comm);
int32_t avr_comm_rank(const avr_comm *comm);
int32_t avr_comm_world(const avr_comm *comm);
/* How avr_graph_propagate_multi exchanges the vector on this communicator: 1 = inside the SpMV kernel, over peer
 * memory (every rank's vector block is mapped into every peer: cudaIpc* between processes, direct peer access
 * between threads of one process); -1 = ncclAllGather per bounce (peer mapping unavailable, more than 16 ranks, or
 * AVR_GRAPH_P2P=0 when the context was created); 0 = not decided yet (decided by the first propagation). */
int32_t avr_comm_graph_exchange(const avr_comm *comm);

/* avr_render_volpath over all ranks of `comm` (every rank calls it with the SAME desc; desc->rank / world are
 * overwritten): rank r renders the work items with item % world == r into its device film, one ncclReduce
 * (fp64 sum) to rank 0, which ADDS the complete film into its host `film` (other ranks may pass NULL).
 * stats: this rank's own counters. */
avr_status avr_render_volpath_multi(avr_comm *comm, avr_medium *m, const avr_render_desc *desc, double *film,
                                    avr_render_stats *stats);

/* avr_graph_propagate_device with the rows of T split over the ranks (equal vertex ranges; every rank holds
 * the whole CSR).  Per bounce one SpMV over the rank's rows whose results are stored straight into EVERY rank's
 * copy of the new vector over peer memory (NVLink / NVSwitch; the blocks are mapped once with cudaIpc*), then a
 * release / acquire handshake between the GPUs; a rank that produces a NaN/Inf posts the iteration to all peers,
 * so every rank reaches the reference's stop verdict without a collective.  Where peers cannot map each other
 * (or with AVR_GRAPH_P2P=0): SpMV + one in-place ncclAllGather + a finiteness pass per bounce
 * (avr_comm_graph_exchange tells which).  The whole bounce sequence is captured in a CUDA graph on first use: no
 * host synchronisation between bounces.  Collective: every rank calls it with the same graph and bounce count.  L0_dev: n floats (the same on every rank); total_dev: n floats of which THIS RANK'S rows
 * [rank * ceil(n / world), ...) are written; iters_done as avr_graph_propagate_device.  Bit-identical to the
 * single-GPU result. */
avr_status avr_graph_propagate_multi(avr_comm *comm, avr_graph *g, const float *L0_dev, int32_t bounces,
                                     float *total_dev, int32_t *iters_done, void *stream);

/* Parity probe for the EXACT path's restated scalar functions: kind 0 = logf (glibc algorithm),
 * 1 = sinf, 2 = cosf (glibc), 3 = FastExp CPU branch (util/math.h:453-474).  Host buffers. */
avr_status avr_debug_math(avr_ctx *ctx, int kind, const float *x, int64_t n, float *out);

#ifdef __cplusplus
}
#endif
#endif /* AVR_H */
