"""Clip sharding across the GPUs of one box and the path's only exchange.

Every clip is independent, so rank r owns the contiguous clip range [r*N/G, (r+1)*N/G) and there is no traffic
during compute; after the last batch the int8 logits are gathered and the per-class counts summed.  The helpers
take a torch.distributed process group (NCCL on GPUs, gloo in the CPU tests).
"""


def shard_range(n_total, rank, world):
    """contiguous [lo, hi) of rank's clips; sizes differ by at most one and cover [0, n_total) exactly"""
    base, rem = divmod(n_total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_logits(dist, logits_local, n_total, world):
    """all ranks -> [n_total, 10] int8 in clip order (ragged shards are padded to the largest shard for the collective)"""
    import torch
    sizes = [shard_range(n_total, r, world)[1] - shard_range(n_total, r, world)[0] for r in range(world)]
    m = max(sizes)
    pad = torch.zeros((m, logits_local.shape[1]), dtype=logits_local.dtype, device=logits_local.device)
    pad[: logits_local.shape[0]] = logits_local
    out = torch.empty((world * m, logits_local.shape[1]), dtype=logits_local.dtype, device=logits_local.device)
    dist.all_gather_into_tensor(out, pad)
    return torch.cat([out[r * m: r * m + sizes[r]] for r in range(world)], dim=0)


def class_counts(dist, logits_local, n_classes=10):
    """argmax histogram over all ranks (int64 [n_classes])"""
    import torch
    c = torch.bincount(logits_local.to(torch.int64).argmax(dim=1), minlength=n_classes) if logits_local.shape[0] else \
        torch.zeros(n_classes, dtype=torch.int64, device=logits_local.device)
    dist.all_reduce(c)
    return c


def reduce_counts(dist, counts_local):
    """sum of the per-rank class histograms that K2 accumulates on the device (edison_kws_counts_dev): int64 [10] on every rank"""
    import torch
    c = counts_local.to(torch.int64)
    dist.all_reduce(c)
    return c
