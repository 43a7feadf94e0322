""".nvdb ingest: avr_nvdb_read_grid / avr_medium_create_nanovdb_file against files written by the
REFERENCE's vendored nanovdb::io::writeGrid from grids built with its GridBuilder (oracle/refprobe.cpp), and against
the flattening the reference's own tree().getValue() gives (Checker.nanovdb_grid).  Bit-exact: the reader moves
floats, it computes nothing."""
import numpy as np
import pytest

from acceleratedvolrenderer_b200 import scenes
from acceleratedvolrenderer_b200.api import AvrError, read_nvdb_blocks, read_nvdb_grid


def sparse_cloud(shape=(70, 50, 90), seed=5):
    """A grid with empty 8^3 leaves, constant regions (tiles after pruning) and a negative index origin."""
    d = scenes.fbm_density(shape[2], shape[1], shape[0], seed=seed)
    d[d < 0.15] = 0
    d[10:26, 8:24, 16:32] = 0.75        # a constant block spanning whole leaves
    return d


@pytest.mark.parametrize("codec", [0, 1])
def test_reader_matches_the_reference_flattening(ref, tmp_path, codec):
    dens = sparse_cloud()
    temp = (dens * 900 + 100).astype(np.float32) * (dens > 0.3)
    origin_ijk, vs, ow = (-13, 5, -40), 0.0125, (0.25, -0.5, 1.0)
    path = tmp_path / "cloud.nvdb"
    ref.write_nvdb(path, [("density", dens, origin_ijk, vs, ow), ("temperature", temp, origin_ijk, vs, ow)], codec=codec)
    for name, arr in (("density", dens), ("temperature", temp)):
        got, info = read_nvdb_grid(path, name)
        want = ref.nanovdb_grid(arr, origin_ijk, vs, ow)       # the reference's grid read back voxel by voxel
        assert info["grids_in_file"] == 2 and info["grid_name"] == name and info["codec"] == codec
        assert info["leaf_nodes"] > 0 and info["active_voxels"] == int((arr != 0).sum())
        assert list(got.index_min) == list(want.index_min) and list(got.dims) == list(want.dims)
        assert np.array_equal(got.values.view(np.uint32), want.values.view(np.uint32))
        assert np.array_equal(got.inv_mat, want.inv_mat) and np.array_equal(got.vec, want.vec)
        assert np.array_equal(got.world_bbox, want.world_bbox)
    # the first grid when no name is given; a name the file does not hold is "absent", not an error (media.cpp:641-645)
    first, info = read_nvdb_grid(path, None)
    assert info["grid_name"] == "density" and first is not None
    none, info = read_nvdb_grid(path, "velocity")
    assert none is None and info["grids_in_file"] == 2


@pytest.mark.parametrize("codec", [0, 1])
def test_block_reader_is_the_flattening_cut_into_leaves(ref, tmp_path, codec):
    """avr_nvdb_read_blocks (no flattening): the 8^3 blocks, put back into the index box, ARE the flattened grid bit for
    bit; origins are multiples of 8, no block is empty, every leaf that holds a value and every value tile shows up."""
    dens = sparse_cloud()
    dens[40:48, 16:24, 8:16] = 0                    # one whole leaf of zeros in the middle of the cloud
    origin_ijk, vs, ow = (-13, 5, -40), 0.0125, (0.25, -0.5, 1.0)
    path = tmp_path / "cloud.nvdb"
    ref.write_nvdb(path, [("density", dens, origin_ijk, vs, ow)], codec=codec)
    flat, info = read_nvdb_grid(path, "density")
    blk, info2 = read_nvdb_blocks(path, "density")
    assert info == info2
    for k in ("index_min", "dims", "inv_mat", "vec", "world_bbox"):
        assert np.array_equal(blk[k], getattr(flat, k))
    org, vals = blk["origins"], blk["values"]
    assert org.shape[0] > 50 and np.all(org % 8 == 0) and np.all((vals != 0).any(axis=(1, 2, 3)))
    assert len({tuple(o) for o in org}) == org.shape[0]
    lo = np.asarray(flat.index_min); dims = np.asarray(flat.dims)
    box = np.zeros(flat.values.shape, dtype=np.float32)
    for o, v in zip(org, vals):
        a = np.maximum(o, lo); b = np.minimum(o + 8, lo + dims)
        assert np.all(a < b)                                   # every block meets the index box
        box[a[2] - lo[2]:b[2] - lo[2], a[1] - lo[1]:b[1] - lo[1], a[0] - lo[0]:b[0] - lo[0]] = \
            v[a[2] - o[2]:b[2] - o[2], a[1] - o[1]:b[1] - o[1], a[0] - o[0]:b[0] - o[0]]
    assert np.array_equal(box.view(np.uint32), flat.values.view(np.uint32))
    assert org.shape[0] * 512 < 0.8 * box.size                 # sparse: fewer voxels held than the box has
    none, _ = read_nvdb_blocks(path, "velocity")
    assert none is None


def test_reader_error_behaviour(tmp_path):
    with pytest.raises(AvrError) as e:
        read_nvdb_grid(tmp_path / "missing.nvdb", "density")
    assert "Unable to open file" in str(e.value)
    junk = tmp_path / "junk.nvdb"
    junk.write_bytes(b"not a nanovdb file at all" * 10)
    with pytest.raises(AvrError) as e:
        read_nvdb_grid(junk, "density")
    assert "not a valid nvdb file" in str(e.value)            # the reference's message (IO.h Segment::read)
    # a truncated file
    good = tmp_path / "t.nvdb"
    good.write_bytes(b"NanoVDB0" + (32 << 21 | 3 << 10 | 3).to_bytes(4, "little") + (1).to_bytes(2, "little") + b"\0\0" + b"\1" * 40)
    with pytest.raises(AvrError):
        read_nvdb_grid(good, "density")


@pytest.mark.gpu
def test_medium_from_file_is_the_medium_from_arrays(ctx, ref, tmp_path):
    """MakeNamedMedium "nanovdb" "string filename": the medium built from the file equals the one built from the
    flattened arrays — majorant grid, bounds and SamplePoint bit for bit, density and temperature grids both."""
    dens = sparse_cloud((40, 36, 44), seed=9)
    temp = (dens * 2500 + 200).astype(np.float32) * (dens > 0)
    vs = 1.0 / 44
    nd, rd = scenes.cloud_scene_nanovdb(dens, temperature=temp, voxel_size=vs, le_scale=0.4, width=16, height=12, spp=1,
                                        sigma_t=11.0, albedo=0.9, g=0.3)
    path = tmp_path / "m.nvdb"
    ref.write_nvdb(path, [("density", dens, (0, 0, 0), vs, (0.0, 0.0, 0.0)), ("temperature", temp, (0, 0, 0), vs, (0.0, 0.0, 0.0))])
    a = ctx.create_medium("nanovdb", nd, layout="octet32")
    b = ctx.create_medium_nanovdb_file(path, nd, layout="octet32")
    assert np.array_equal(a.majorants(), b.majorants())
    assert np.array_equal(a.bounds(), b.bounds()) and a.is_emissive() == b.is_emissive()
    rng = np.random.default_rng(2)
    p = rng.uniform(-0.5, 0.5, (5000, 3)).astype(np.float32)
    lam = np.array([420, 520, 610, 700], dtype=np.float32)
    assert np.array_equal(a.sample_point(p, lam).view(np.uint32), b.sample_point(p, lam).view(np.uint32))
    with pytest.raises(AvrError) as e:
        ctx.create_medium_nanovdb_file(path, nd, gridname="smoke")
    assert 'didn\'t find "density" grid' in str(e.value)       # media.cpp:637-638
    with pytest.raises(AvrError) as e:
        ctx.create_medium_nanovdb_file("", nd)
    assert 'Must supply "filename"' in str(e.value)            # media.cpp:632-633
    a.close(); b.close()


@pytest.mark.gpu
def test_sparse_nvdb_file_into_the_brick_pool(ctx, ref, tmp_path):
    """SURVEY 8(f) n4 as a whole: a SPARSE grid built with the vendored GridBuilder, written as a .nvdb file (ZIP
    codec), read by avr_medium_create_nanovdb_file into the sparse fp16 brick pool: lookups equal the fp16-rounded
    flattened grid (the octet16 medium from the same arrays, bit for bit), the pool is a fraction of the dense box,
    and a FAST render through it agrees with the octet32 render of the same medium to the fp16 rounding of the taps."""
    dens = sparse_cloud((72, 64, 80), seed=4)
    dens[:, :, 24:] = 0          # 70 % of the index box is empty on top of the cloud's own holes
    dens[0, 0, 79] = 0.01        # ... but the box still spans it
    vs = 1.0 / 80
    nd, rd = scenes.cloud_scene_nanovdb(dens, voxel_size=vs, width=32, height=24, spp=64, sigma_t=30.0, albedo=0.9, g=0.3)
    path = tmp_path / "sparse.nvdb"
    ref.write_nvdb(path, [("density", dens, (0, 0, 0), vs, (0.0, 0.0, 0.0))], codec=1)
    brick = ctx.create_medium_nanovdb_file(path, nd, layout="brick16")
    half = ctx.create_medium("nanovdb", nd, layout="octet16")
    full = ctx.create_medium("nanovdb", nd, layout="octet32")
    assert np.array_equal(brick.majorants(), full.majorants())
    rng = np.random.default_rng(3)
    xyz = rng.uniform(-2, np.array(dens.shape[::-1]) + 2, (40000, 3)).astype(np.float32)
    assert np.array_equal(brick.lookup_density(xyz).view(np.uint32), half.lookup_density(xyz).view(np.uint32))
    dense = ctx.create_medium("nanovdb", nd, layout="dense")
    pool_bytes = brick.bytes() - (dense.bytes() - dens.nbytes)   # both carry the same majorant grids and tables
    dense.close()
    assert pool_bytes < 0.6 * dens.nbytes and brick.bytes() < 0.25 * full.bytes(), (pool_bytes, dens.nbytes, full.bytes())
    a = np.zeros((24, 32, 4)); b = np.zeros_like(a)
    for seed in range(4):
        ctx.render_volpath(brick, rd.with_(exact_math=0, sampler_seed=seed), film=a)
        ctx.render_volpath(full, rd.with_(exact_math=0, sampler_seed=100 + seed), film=b)
    ma, mb = a[..., :3].sum() / a[..., 3].sum(), b[..., :3].sum() / b[..., 3].sum()
    assert abs(ma / mb - 1) < 0.02, (ma, mb)
    # the pool above came straight from the file's leaves (avr_nvdb_read_blocks); through the flattened box it is the
    # same medium: size, majorants, every lookup
    import os
    os.environ["AVR_NVDB_FLATTEN"] = "1"
    try:
        flat = ctx.create_medium_nanovdb_file(path, nd, layout="brick16")
    finally:
        del os.environ["AVR_NVDB_FLATTEN"]
    assert flat.bytes() == brick.bytes() and np.array_equal(flat.majorants(), brick.majorants())
    assert np.array_equal(flat.lookup_density(xyz).view(np.uint32), brick.lookup_density(xyz).view(np.uint32))
    flat.close()
    brick.close(); half.close(); full.close()


@pytest.mark.gpu
def test_blocks_with_a_negative_origin_and_value_tiles(ctx, ref, tmp_path):
    """Leaf-direct ingest where the index box does not start on a block boundary (negative, odd origin) and the grid
    holds constant regions that NanoVDB may store as value tiles: equal to the flattened route, lookup by lookup."""
    import os
    dens = sparse_cloud((70, 50, 90), seed=5)
    origin_ijk, vs = (-13, 5, -40), 1.0 / 90
    nd, rd = scenes.cloud_scene_nanovdb(dens, voxel_size=vs, width=16, height=12, spp=1, sigma_t=11.0, albedo=0.9, g=0.3)
    path = tmp_path / "neg.nvdb"
    ref.write_nvdb(path, [("density", dens, origin_ijk, vs, (0.0, 0.0, 0.0))], codec=1)
    a = ctx.create_medium_nanovdb_file(path, nd, layout="brick16")
    os.environ["AVR_NVDB_FLATTEN"] = "1"
    try:
        b = ctx.create_medium_nanovdb_file(path, nd, layout="brick16")
    finally:
        del os.environ["AVR_NVDB_FLATTEN"]
    assert a.bytes() == b.bytes() and np.array_equal(a.majorants(), b.majorants()) and a.majorants().max() > 0
    rng = np.random.default_rng(8)
    lo = np.array(origin_ijk, dtype=np.float32)
    xyz = (lo + rng.uniform(-2, np.array(dens.shape[::-1]) + 2, (40000, 3))).astype(np.float32)
    la, lb = a.lookup_density(xyz), b.lookup_density(xyz)
    assert np.array_equal(la.view(np.uint32), lb.view(np.uint32)) and (la > 0).mean() > 0.1
    a.close(); b.close()
