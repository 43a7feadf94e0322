"""CPU, world_size 2 over gloo: the multi-GPU host logic of acceleratedvolrenderer_b200/sharding.py.

The compute step is the CPU checker here (the only place a test may stand it in for the CUDA library);
what is under test is the partitioning, the collectives and the early-out agreement.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from acceleratedvolrenderer_b200 import scenes, sharding
from cpu_checker import Checker


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(fn, world, *args):
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_entry, args=(fn, r, world, port, q) + args) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    return q.get()


def _entry(fn, rank, world, port, q, *args):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        out = fn(rank, world, *args)
        if rank == 0:
            q.put(out)
    finally:
        dist.destroy_process_group()


def _scene():
    dens = scenes.fbm_density(24, 20, 28, seed=4)
    return scenes.cloud_scene(dens, width=16, height=12, spp=2, sigma_t=14.0, albedo=0.95, g=0.4, max_depth=6)


def _render_shard(rank, world, strong):
    gd, rd = _scene()
    ck = Checker("orc")
    m = ck.medium_uniformgrid(gd)
    r = sharding.split_samples(rd.with_(spp=5), rank, world) if strong else sharding.shard_samples(rd, rank, world, 2)
    film = np.zeros((rd.height, rd.width, 4), dtype=np.float64)
    if r is not None:
        m.render(r, threads=2, film=film)
    m.close()
    t = torch.from_numpy(film)
    sharding.reduce_film(t, dst=0)
    return t.numpy().copy()


@pytest.mark.parametrize("strong", [False, True])
def test_sample_sharded_film_equals_single_process(strong):
    got = _run(_render_shard, 2, strong)
    gd, rd = _scene()
    ck = Checker("orc")
    m = ck.medium_uniformgrid(gd)
    want, _ = m.render(rd.with_(spp=5 if strong else 4), threads=2)
    m.close()
    assert np.all(got[..., 3] == (5 if strong else 4))
    assert np.allclose(got, want, rtol=1e-12, atol=1e-300)
    assert got[..., :3].sum() > 0


def test_split_samples_covers_every_sample_once():
    _, rd = _scene()
    for spp in (1, 2, 5, 8):
        for world in (1, 2, 3, 8):
            seen = []
            for r in range(world):
                s = sharding.split_samples(rd.with_(spp=spp), r, world)
                if s is not None:
                    seen += list(range(s.sample_begin, s.sample_end))
            assert seen == list(range(spp))


def _csr_rows(rowptr, col, val):
    def spmv_rows(cur, r0, r1):
        x = cur.numpy()
        out = np.zeros(r1 - r0, dtype=np.float32)
        for i in range(r0, r1):  # per row, column order, fp32 accumulate: Eigen's row-major product
            acc = np.float32(0)
            for k in range(rowptr[i], rowptr[i + 1]):
                acc = np.float32(acc + np.float32(val[k] * x[col[k]]))
            out[i - r0] = acc
        return torch.from_numpy(out)
    return spmv_rows


def _propagate(rank, world, n, bounces, scale):
    rowptr, col, val = scenes.random_graph(n, avg_deg=4, seed=n)
    L0 = np.random.default_rng(3).random(n, dtype=np.float32)
    total, it = sharding.propagate_sharded(_csr_rows(rowptr, col, (val * scale).astype(np.float32)),
                                           torch.from_numpy(L0), bounces, rank, world)
    return total.numpy().copy(), it


@pytest.mark.parametrize("n,bounces,scale", [(301, 5, 1.0), (64, 0, 1.0), (301, 20, 4e7)])
def test_row_sharded_propagation_matches_oracle(orc, n, bounces, scale):
    got, it = _run(_propagate, 2, n, bounces, scale)
    rowptr, col, val = scenes.random_graph(n, avg_deg=4, seed=n)
    L0 = np.random.default_rng(3).random(n, dtype=np.float32)
    want, want_it = orc.graph_propagate(rowptr, col, (val * scale).astype(np.float32), L0, bounces)
    assert it == want_it
    assert (got.view(np.uint32) == want.view(np.uint32)).all()


def test_row_partition_pads_and_covers():
    for n in (1, 7, 64, 301):
        for world in (1, 2, 3, 8):
            chunk, parts = sharding.row_partition(n, world)
            assert chunk * world >= n
            assert [a for a, _ in parts][0] == 0 and parts[-1][1] == n
            assert all(b - a <= chunk for a, b in parts)
            assert sum(b - a for a, b in parts) == n
