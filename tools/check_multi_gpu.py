#!/usr/bin/env python
"""Multi-GPU parity check, run under torchrun (one rank per GPU, NCCL):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      tools/check_multi_gpu.py

1. sample-sharded volpath render + ncclReduce of the film == the same samples rendered on one GPU
   (fp64 film, rtol 1e-12: only the summation order differs);
2. row-sharded graph propagation with ncclAllGather == single-GPU propagation, bit for bit.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from acceleratedvolrenderer_b200 import scenes, sharding  # noqa: E402
from acceleratedvolrenderer_b200.api import Context  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    ctx = Context(local)
    ok = True

    dens = scenes.fbm_density(48, 40, 52, seed=4)
    gd, rd = scenes.cloud_scene(dens, width=64, height=48, spp=3, sigma_t=14.0, albedo=0.95, g=0.4, max_depth=8)
    med = ctx.create_medium("uniformgrid", gd, layout="octet32")
    H, W = rd.height, rd.width
    for exact in (1, 0):
        film = torch.zeros((H, W, 4), dtype=torch.float64, device=dev)
        torch.cuda.synchronize()   # the fill runs on torch's stream, the render on the library's: order them
        ctx.render_volpath_device(med, sharding.shard_samples(rd.with_(exact_math=exact), rank, world, 3), film.data_ptr())
        torch.cuda.synchronize()
        sharding.reduce_film(film, dst=0)
        if rank == 0:
            ref = torch.zeros_like(film)
            torch.cuda.synchronize()
            ctx.render_volpath_device(med, rd.with_(exact_math=exact, spp=3 * world), ref.data_ptr())
            torch.cuda.synchronize()
            same = torch.allclose(film, ref, rtol=1e-12, atol=1e-300) and bool((film[..., 3] == 3 * world).all())
            print(f"[film] exact={exact} world={world} sharded == single-GPU: {same}")
            ok = ok and same
    med.close()

    n = 20011
    rowptr, col, val = scenes.random_graph(n, avg_deg=6, seed=5)
    L0 = np.random.default_rng(2).random(n, dtype=np.float32)
    g = ctx.create_graph(rowptr, col, val)
    for bounces, scale in ((8, 1.0), (25, 3e7)):
        g2 = g if scale == 1.0 else ctx.create_graph(rowptr, col, (val * scale).astype(np.float32))
        total, it = g2.compute_final_light_sharded(torch.from_numpy(L0).to(dev), bounces, rank, world)
        torch.cuda.synchronize()
        want, want_it = g2.compute_final_light(L0, bounces)
        same = it == want_it and bool((total.cpu().numpy().view(np.uint32) == want.view(np.uint32)).all())
        flag = torch.tensor([0 if same else 1], device=dev)
        dist.all_reduce(flag)
        if rank == 0:
            print(f"[graph] bounces={bounces} scale={scale} iterations={it}/{want_it} all ranks bit-identical: {int(flag[0]) == 0}")
            ok = ok and int(flag[0]) == 0
        if g2 is not g:
            g2.close()
    # ---- the same two splits behind the C ABI (avr_comm_*, NCCL bound by the library itself) ------------
    comm = sharding.init_comm(ctx)
    med = ctx.create_medium("uniformgrid", gd, layout="octet32")
    for exact in (1, 0):
        full = rd.with_(exact_math=exact, spp=3 * world)
        film, _ = comm.render_volpath(med, full)
        if rank == 0:
            ref = torch.zeros((H, W, 4), dtype=torch.float64, device=dev)
            torch.cuda.synchronize()
            ctx.render_volpath_device(med, full, ref.data_ptr())
            torch.cuda.synchronize()
            same = np.allclose(film, ref.cpu().numpy(), rtol=1e-12, atol=1e-300) and bool((film[..., 3] == 3 * world).all())
            print(f"[C-ABI film] exact={exact} avr_render_volpath_multi == single-GPU: {same}")
            ok = ok and same
    med.close()
    L0d = torch.from_numpy(L0).to(dev)
    for bounces, scale in ((8, 1.0), (0, 1.0), (25, 3e7)):
        g2 = g if scale == 1.0 else ctx.create_graph(rowptr, col, (val * scale).astype(np.float32))
        total = torch.full((n,), float("nan"), device=dev)
        torch.cuda.synchronize()
        for rep in range(2):   # second call: the cached CUDA graph
            it = comm.propagate(g2, L0d.data_ptr(), bounces, total.data_ptr())
        torch.cuda.synchronize()
        want, want_it = g2.compute_final_light(L0, bounces)
        r0, r1 = comm.row_range(n)
        same = it == want_it and bool((total[r0:r1].cpu().numpy().view(np.uint32) == want[r0:r1].view(np.uint32)).all())
        flag = torch.tensor([0 if same else 1], device=dev)
        dist.all_reduce(flag)
        if rank == 0:
            print(f"[C-ABI graph] bounces={bounces} scale={scale} iterations={it}/{want_it} rows of every rank bit-identical: {int(flag[0]) == 0}")
            ok = ok and int(flag[0]) == 0
        if g2 is not g:
            g2.close()
    comm.close()
    g.close()
    ctx.close()
    dist.destroy_process_group()
    if rank == 0:
        print("MULTI-GPU CHECK", "OK" if ok else "FAILED")
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
