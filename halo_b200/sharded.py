"""Multi-GPU sharding of the path: one process per GPU, torch.distributed for the plumbing.

* MSM / commit: point-range sharding — rank g owns [g*k/G, (g+1)*k/G); the G partial sums (one Jacobian point,
  96 B) are all-gathered and added.  EC addition is not an NCCL reduction op, hence gather + local add.
* batches (commit columns, accumulation instances, NTT polynomials): round-robin by item, no data-path collective;
  results are all-gathered only when every rank needs them.
The collective calls are backend-agnostic (NCCL on the GPUs; gloo in the CPU tests of the host logic, where the
local compute is injected)."""
import numpy as np

from . import group


def point_range(k, rank, world):
    """Contiguous slice of k points owned by `rank` (sizes differ by at most one)."""
    base, rem = divmod(k, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def items_of(m, rank, world):
    """Round-robin item ownership for batched work."""
    return list(range(rank, m, world))


def _device_for(dist):
    import torch
    return torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")


def all_gather_words(dist, words):
    """All-gather of equally sized uint64 arrays -> (world, len) array."""
    import torch
    dev = _device_for(dist)
    t = torch.from_numpy(np.ascontiguousarray(words, dtype=np.uint64).view(np.int64).reshape(-1)).to(dev)
    out = torch.empty(dist.get_world_size() * t.numel(), dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(out, t)
    return out.cpu().numpy().view(np.uint64).reshape(dist.get_world_size(), -1)


def sum_points(curve, parts):
    """Sum of the per-rank Jacobian partial sums (host arithmetic inside the library: G - 1 additions)."""
    import ctypes
    from ._lib import lib, check, ptr
    parts = np.ascontiguousarray(parts, dtype=np.uint64).reshape(-1, 12)
    out = np.empty(12, dtype=np.uint64)
    check(lib.halo_point_sum(curve, ptr(parts.reshape(-1)), ctypes.c_size_t(parts.shape[0]), ptr(out)))
    return out


def msm_sharded(dist, curve, scalars, bases, local_msm=None, combine=None):
    """sum_i scalars[i]*bases[i] with every rank holding (or generating) the full inputs but computing only its
    point range.  bases=None uses the resident generators.  Returns the full result on every rank."""
    rank, world = dist.get_rank(), dist.get_world_size()
    scalars = np.ascontiguousarray(scalars, dtype=np.uint64).reshape(-1, 4)
    k = scalars.shape[0] if bases is None else min(scalars.shape[0], np.asarray(bases).reshape(-1, 8).shape[0])
    lo, hi = point_range(k, rank, world)
    if local_msm is None:
        local_msm = (lambda c, s, b, off: group.msm_resident(c, s, off)) if bases is None else \
                    (lambda c, s, b, off: group.point_dot_affine(c, s, b))
    b_loc = None if bases is None else np.asarray(bases, dtype=np.uint64).reshape(-1, 8)[lo:hi]
    part = local_msm(curve, scalars[lo:hi], b_loc, lo)
    parts = all_gather_words(dist, part)
    return (combine or sum_points)(curve, parts)


def msm_batch_sharded(dist, curve, columns, base_off=0, local_batch=None, gather=True):
    """b commits over the same resident generators, columns split round-robin over ranks (no data-path collective);
    gather=True returns all b results on every rank."""
    rank, world = dist.get_rank(), dist.get_world_size()
    columns = np.ascontiguousarray(columns, dtype=np.uint64)
    b = columns.shape[0]
    mine = items_of(b, rank, world)
    local_batch = local_batch or (lambda c, cols, off: group.msm_batch_resident(c, cols, off))
    res = local_batch(curve, columns[mine], base_off) if mine else np.zeros((0, 12), dtype=np.uint64)
    if not gather:
        return mine, res
    per = (b + world - 1) // world
    pad = np.zeros((per, 12), dtype=np.uint64)
    pad[: len(mine)] = res
    allr = all_gather_words(dist, pad).reshape(world, per, 12)
    out = np.zeros((b, 12), dtype=np.uint64)
    for r in range(world):
        for j, i in enumerate(items_of(b, r, world)):
            out[i] = allr[r, j]
    return out


def ntt_batch_sharded(dist, field, polys, inverse=False, local_ntt=None, gather=True):
    """Batched NTT sharded by polynomial (a single transform stays on one GPU)."""
    from . import poly
    rank, world = dist.get_rank(), dist.get_world_size()
    polys = np.ascontiguousarray(polys, dtype=np.uint64)
    b, n = polys.shape[0], polys.shape[1]
    mine = items_of(b, rank, world)
    local_ntt = local_ntt or (lambda f, x, inv: poly.ntt(f, x, inverse=inv))
    res = local_ntt(field, polys[mine], inverse) if mine else np.zeros((0, n, 4), dtype=np.uint64)
    if not gather:
        return mine, res
    per = (b + world - 1) // world
    pad = np.zeros((per, n, 4), dtype=np.uint64)
    pad[: len(mine)] = res
    allr = all_gather_words(dist, pad).reshape(world, per, n, 4)
    out = np.zeros_like(polys)
    for r in range(world):
        for j, i in enumerate(items_of(b, r, world)):
            out[i] = allr[r, j]
    return out
