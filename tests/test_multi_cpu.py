"""Multi-GPU host logic on CPU: world_size-2 gloo run of the spp-split + single reduce + rank-0 tonemap driver
(path_tracer_b200/multi.py), with the CPU oracle standing in for the per-rank renderer."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import GOLDEN  # first: registers the package under its importable name (also in spawned workers)

import isf_py
import oracle_lib

import path_tracer_b200 as pt
from path_tracer_b200 import multi

W, H, SPP, BOUNCES = 48, 32, 7, 3
SCENE = os.path.join(GOLDEN, "scenes", "alpha_transparency", "scene.isf")


def test_sample_partition_covers_every_sample_once():
    for total in (0, 1, 7, 8, 128, 1001):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                b, s, c = multi.sample_partition(total, world, r)
                seen += [b + k * s for k in range(c)]
            assert sorted(seen) == list(range(total)), (total, world)
    with pytest.raises(ValueError):
        multi.sample_partition(8, 2, 2)


def test_single_process_driver_is_identity():
    acc = torch.zeros(4, 4, 3)
    img = multi.render_distributed(lambda b, s, c, a: a.add_(float(c)), lambda a, n: a / n, 5, acc)
    assert torch.equal(img, torch.ones(4, 4, 3))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        osc = oracle_lib.OracleScene(isf_py.load_isf(SCENE))
        prof = oracle_lib.make_profile(W, H, SPP, BOUNCES, 2)

        def render_fn(begin, stride, count, accum):
            part = osc.render_accum(prof, rng_mode=1, seed=5, sample_begin=begin, sample_stride=stride, n_samples=count,
                                    threads=2)
            accum.add_(torch.from_numpy(part))

        def tonemap_fn(accum, total):
            return oracle_lib.tonemap(accum.numpy(), total, 2)

        accum = torch.zeros(H, W, 3)
        img = multi.render_distributed(render_fn, tonemap_fn, SPP, accum)
        assert (img is None) == (rank != 0)
        if rank == 0:
            np.savez(out_path, img=img, accum=accum.numpy())
        # the scene-distribution leg (DistributedRenderer.distribute_scene): ragged byte arrays of one rank reach the
        # same-sized arrays of the others, on a process group of their own
        grp = dist.new_group(backend="gloo")
        sizes = [1, 48, 64 * 1000 + 17, 4096]
        want = [torch.arange(n, dtype=torch.int64).mul(7).add(k).to(torch.uint8) for k, n in enumerate(sizes)]
        have = [w.clone() if rank == 1 else torch.zeros_like(w) for w in want]
        multi.broadcast_arrays(have, src=1, group=grp)
        assert all(torch.equal(h, w) for h, w in zip(have, want))
        # ... and the all-gather form: part r of every array valid on rank r, the remainder on both
        have = []
        for w in want:
            n = multi.part_size(w.numel(), world)
            h = torch.zeros_like(w)
            h[rank * n:(rank + 1) * n] = w[rank * n:(rank + 1) * n]
            h[world * n:] = w[world * n:]
            have.append(h)
        multi.allgather_parts(have, group=grp)
        assert all(torch.equal(h, w) for h, w in zip(have, want))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_render_equals_single_process(tmp_path):
    out = str(tmp_path / "r.npz")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = np.load(out)
    osc = oracle_lib.OracleScene(isf_py.load_isf(SCENE))
    prof = oracle_lib.make_profile(W, H, SPP, BOUNCES, 2)
    full = osc.render_accum(prof, rng_mode=1, seed=5, threads=2)
    # same samples, different f32 summation order
    assert np.allclose(got["accum"], full, rtol=1e-5, atol=1e-6)
    ref = oracle_lib.tonemap(full, SPP, 2)
    assert (np.abs(got["img"].astype(int) - ref.astype(int)) <= 1).all()
    assert (got["img"] == ref).mean() > 0.999
