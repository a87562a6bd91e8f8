"""Host-side multi-GPU logic of the raster path: one process per GPU, torch.distributed for the
plumbing.  Backend-agnostic (nccl on the GPU box, gloo in the CPU tests).

  * view-parallel batches (config C5): views are independent; the only exchange is the shadow map, rendered once on
    rank 0 and broadcast (NCCL over NVLink on the box).
  * sort-first row bands (config C4): rank r owns pixel rows y with (y // band_h) % world == r; geometry is replicated;
    the optional final exchange all-gathers every rank's frame and keeps the owned rows.
    Its shadow pass is split either into row slabs (all-gather) or by triangles into whole maps (all-reduce MIN).
No data-path collective exists besides those.
"""
import numpy as np


def views_for_rank(n_views, rank, world):
    """Indices of the camera views rank `rank` renders (contiguous blocks, sizes differ by at most one)."""
    base, rem = divmod(n_views, world)
    lo = rank * base + min(rank, rem)
    return list(range(lo, lo + base + (1 if rank < rem else 0)))


def owned_rows(height, rank, world, band_h=64):
    """Boolean mask over pixel rows owned by `rank` (matches row_owned() in csrc/reb_device.cuh and the oracle)."""
    y = np.arange(height)
    return ((y // band_h) % world) == rank if world > 1 else np.ones(height, bool)


def broadcast_shadow(shadow_tensor, src=0):
    """Shadow map rendered once on `src`, then shared: 4*S*S bytes per frame (16.8 MB at 2048^2)."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(shadow_tensor, src=src)
    return shadow_tensor


def slab_rows(size, world):
    """Rows per rank of a contiguous, equal split of a `size`-row map into `world` slabs whose height is a multiple of 32
    (reb_set_shadow_partition's granularity); None when the map does not divide that way."""
    if world <= 0 or size % (32 * world):
        return None
    return size // world


def allgather_shadow_slabs(shadow_tensor):
    """Shadow pass split by rows: every rank rendered slab `rank` of the map (reb_set_shadow_partition with band_h =
    slab_rows(S, world)); one in-place all-gather (4*S*S/N bytes sent per rank) completes the map on every rank."""
    import torch.distributed as dist
    if not (dist.is_initialized() and dist.get_world_size() > 1):
        return shadow_tensor
    world, rank = dist.get_world_size(), dist.get_rank()
    rows = slab_rows(shadow_tensor.shape[0], world)
    if rows is None:
        raise ValueError("shadow map rows must divide into equal slabs of a multiple of 32 rows")
    dist.all_gather_into_tensor(shadow_tensor, shadow_tensor[rank * rows:(rank + 1) * rows])
    return shadow_tensor


def allreduce_shadow_min(shadow_tensor):
    """Shadow pass split by TRIANGLES (reb_set_shadow_split): every rank rasterised its share of the triangle stream into a whole
    map; the element-wise minimum over the ranks (one in-place all-reduce of 4*S*S bytes; depths are non-negative floats) is the
    map a single GPU renders, bit for bit."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(shadow_tensor, op=dist.ReduceOp.MIN)
    return shadow_tensor


def contiguous_band(height, world):
    """band_h that gives every rank ONE contiguous band of the frame (multiple of 8): with per-draw culling this is what
    lets a rank skip most of the geometry."""
    return ((height + world - 1) // world + 7) // 8 * 8


def gather_bands(local_frame, band_h=64):
    """Sort-first final exchange: all-gather every rank's frame (clear values outside its bands) and compose the
    full image from the owned rows.  local_frame: torch tensor [H, W, ...]; returns the composed tensor on every rank."""
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(), dist.get_rank()
    parts = [torch.empty_like(local_frame) for _ in range(world)]
    dist.all_gather(parts, local_frame)
    out = torch.empty_like(local_frame)
    H = local_frame.shape[0]
    for r in range(world):
        m = torch.from_numpy(owned_rows(H, r, world, band_h)).to(local_frame.device)
        out[m] = parts[r][m]
    return out
