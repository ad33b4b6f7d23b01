"""world_size-2 test of the multi-GPU host logic on CPU (gloo): sample-range sharding + the single film
reduce. Each rank accumulates a deterministic stand-in film for ITS sample range (a hash of
(pixel, global sample index), the same keying the CUDA sampler uses); after the reduce rank 0 must hold
exactly the film a single process accumulates over all samples."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mitsuba_renderer_b200 import multigpu

H, W, SPP = 12, 17, 37


def fake_film(lo, hi):
    pix = np.arange(H * W, dtype=np.uint64)[:, None]
    smp = np.arange(lo, hi, dtype=np.uint64)[None, :]
    z = (pix << np.uint64(32) | smp) * np.uint64(0x9E3779B97F4A7C15)
    z ^= z >> np.uint64(31)
    val = ((z >> np.uint64(40)).astype(np.float64) / float(1 << 24))        # [0,1), exactly representable
    film = np.zeros((H * W, 5), np.float64)
    for c in range(4):
        film[:, c] = (val * (c + 1)).sum(1)
    film[:, 4] = hi - lo
    return film.astype(np.float32)


def worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = multigpu.sample_range(SPP, rank, world)
    film = torch.from_numpy(fake_film(lo, hi)).reshape(-1)
    multigpu.reduce_film(film, dst=0)
    if rank == 0:
        np.save(out, film.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_film_reduce_equals_single_rank(tmp_path):
    out = str(tmp_path / "film.npy")
    port = 29500 + os.getpid() % 2000
    mp.spawn(worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out).reshape(H * W, 5)
    want = fake_film(0, SPP)
    assert np.array_equal(got[:, 4], want[:, 4])                 # every sample counted exactly once
    assert np.allclose(got, want, rtol=1e-6, atol=1e-5)


def test_sample_ranges():
    for spp in (1, 5, 64, 1000):
        for world in (1, 2, 4, 8):
            r = [multigpu.sample_range(spp, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == spp and all(a[1] == b[0] for a, b in zip(r, r[1:]))
