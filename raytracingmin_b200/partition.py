"""Sample-budget partition across GPUs and the one collective of the path.

Every (pixel, sub-pixel, sample) of the reference's loop nest (src/Renderer.cpp:222-243)
is independent, so rank g of G renders samples [g*S/G, (g+1)*S/G) of the `samples` loop
for ALL sub-pixels into its own float4[H][W] buffer, dividing by the GLOBAL ss*ss*S
(src/Renderer.cpp:240) so that the partial buffers add up to the reference's formula; the
RNG counter carries the global sample index, so the image does not depend on G (up to
FP32 summation order).  One sum-reduce to rank 0 is the only communication.
"""
from __future__ import annotations


def sample_slice(samples: int, rank: int, world: int) -> tuple[int, int]:
    """[begin, end) of the `samples` loop rank `rank` of `world` renders.  Slices are
    contiguous, disjoint, cover [0, samples) and differ in length by at most one."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    if samples < world:
        raise ValueError(f"cannot slice {samples} samples over {world} ranks (slice the image instead)")
    return samples * rank // world, samples * (rank + 1) // world


def plan_partition(samples: int, superSamples: int, rank: int, world: int) -> dict:
    """Keyword arguments for Renderer(...) that give rank `rank` its share of the
    ss*ss*samples sample budget of every pixel.

    Preferred: slice the ss*ss sub-pixel grid (src/Renderer.cpp:224-225) when it divides
    evenly — every rank then keeps full-length `samples` loops per sub-pixel (no extra
    per-unit overhead in the megakernel) and runs only 1/world of the FP64 primary pass.
    Otherwise slice the `samples` loop (src/Renderer.cpp:226).  Either way the partial
    framebuffers add up to the reference's image and the RNG sees global indices."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    grid = superSamples * superSamples
    if world > 1 and grid % world == 0:
        per = grid // world
        return dict(sub_begin=rank * per, sub_end=(rank + 1) * per, sample_begin=0, sample_end=samples)
    b, e = sample_slice(samples, rank, world)
    return dict(sub_begin=0, sub_end=grid, sample_begin=b, sample_end=e)


def reduce_partials(accum, dst: int = 0):
    """Sum the per-rank partial framebuffers into rank `dst` (in place, on the tensor's
    device, enqueued on the current stream: NCCL on GPUs, gloo on CPU tensors in tests)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum
