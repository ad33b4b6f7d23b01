"""Host-side multi-GPU logic: one process per GPU, scene replicated, the sample range of
every pixel split across ranks, and ONE reduction of the {sum spec, sum alpha, sum weight} film. This
replaces BlockedRenderProcess::processResult's mutex-guarded merge (src/librender/renderproc.cpp:123-130)
and the TCP work-unit protocol of RemoteWorker (src/libcore/sched_remote.cpp) for this integrator."""
from __future__ import annotations


def sample_range(spp: int, rank: int, world: int):
    """Samples [lo, hi) of every pixel rendered by `rank`. Contiguous, disjoint, sizes differ by <= 1.
    Because RNG streams are keyed by (pixel, global sample index), the union over ranks is exactly the
    sample set a single GPU would render."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return spp * rank // world, spp * (rank + 1) // world


def reduce_film(film, dst: int = 0):
    """Sum the per-rank film accumulators onto rank `dst` (ncclReduce over NVLink on GPUs; gloo on CPU).
    `film` is a torch tensor of cropH*cropW*5 floats on the rank's device; reduced in place on dst."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM)
    return film


def render_sharded(ctx, scene, film, rank: int, world: int, seed: int = 0):
    """Render this rank's sample range into the device tensor `film` and reduce it to rank 0."""
    from .api import Context
    lo, hi = sample_range(scene.spp, rank, world)
    stats = ctx.render_device(Context.params_from_scene(scene, lo, hi, seed=seed), film.data_ptr())
    import torch
    torch.cuda.synchronize()
    reduce_film(film)
    return stats
