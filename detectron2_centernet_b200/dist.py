"""Multi-GPU plumbing of the ctdet hot path: one process per GPU, work sharded BY IMAGE, no
data-path collective.  The reference normalises the focal loss by the rank-LOCAL num_pos
(detectron2/modeling/meta_arch/centernet.py:360-368) and lets DDP average gradients
(engine/defaults.py:280-284); that is the default here too.  `sync_normalisers` is the one optional
collective: an all-reduce(SUM) of the 2-float vector {num_pos, sum(reg_mask)} (NCCL over
NVLink/NVSwitch on GPUs, gloo in the CPU tests), consumed by the fused loss kernel as a device
scalar -- no host sync."""
import torch
import torch.distributed as dist


def shard_range(n_items, rank, world):
    """Contiguous block of items owned by `rank` (like TrainingSampler's per-rank split,
    detectron2/data/samplers/distributed_sampler.py:43-55, but over one global batch)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def local_normalisers(targets):
    """f32[2] = {num_pos, sum(reg_mask)} of this rank's targets (on the targets' device)."""
    if "num_pos" in targets:
        npos = targets["num_pos"].sum().to(torch.float32)
    else:
        npos = (targets["hm"] == 1).sum().to(torch.float32)
    nmask = targets["reg_mask"].sum().to(torch.float32)
    return torch.stack([npos, nmask])


def sync_normalisers(targets, group=None):
    """All-reduced normalisers for `ctdet_losses(..., norm=...)`.  With one process it is the
    local vector.  Note: the reference does NOT do this; use it only when global normalisation
    is wanted."""
    v = local_normalisers(targets)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(v, op=dist.ReduceOp.SUM, group=group)
    return v
