"""Entity-wise sharding across the GPUs of one box.

Rank r owns the contiguous instance range [r*n, (r+1)*n) of the flattened order and runs the whole
single-GPU pipeline on it: the sharded phase needs no collective.  Only where ONE merged draw list is
wanted do the ranks exchange data: the per-view draw keys + counts (KBs) and the compacted
(key, instance) lists, all-gathered over NCCL/NVLink through torch.distributed.  Because shards are
contiguous and in rank order, "rank 0's run of key k, then rank 1's run of key k, ..." is exactly the
single-GPU stable-sort order, so the merge is run placement, not a sort.

Everything here is host-side plumbing around torch tensors; on CPU (gloo) it is exercised by
tests/test_sharding.py with world_size 2.
"""
from typing import List, Tuple

import torch
import torch.distributed as dist


def shard_range(n_total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, near-equal ranges; rank order = input order."""
    lo = n_total * rank // world
    hi = n_total * (rank + 1) // world
    return lo, hi


def _all_gather_padded(t: torch.Tensor, counts: List[int]) -> List[torch.Tensor]:
    """all_gather of variable-length 1-D tensors: pad to the max, gather once, trim."""
    world = dist.get_world_size()
    cap = max(max(counts), 1)
    buf = torch.zeros(cap, dtype=t.dtype, device=t.device)
    buf[: t.numel()] = t
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf)
    return [o[:c] for o, c in zip(out, counts)]


def _gather_scalar(value: int, dev) -> torch.Tensor:
    world = dist.get_world_size()
    mine = torch.tensor([value], dtype=torch.int64, device=dev)
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine)
    return torch.cat(out)


def merge_view_list(keys: torch.Tensor, idx: torch.Tensor, draw_keys: torch.Tensor, draw_counts: torch.Tensor,
                    idx_offset: int):
    """Merges this rank's compacted list of ONE view with every other rank's.

    keys, idx            this rank's visible (key, local instance) list, sorted by key (stable)
    draw_keys, draw_counts  this rank's draws: key and instance_count per run
    idx_offset           first global instance number of this rank's shard

    Returns (merged_keys, merged_global_idx, merged_draw_keys, merged_draw_counts, merged_first_instance);
    identical on every rank and identical to what one GPU holding all shards would emit.
    Keys are int64 tensors (the uint64 bit pattern; keys use at most 62 bits so order is preserved).
    """
    world = dist.get_world_size()
    dev = keys.device
    # 1) counts first (tiny): how many list entries / draws every rank holds
    mine = torch.tensor([keys.numel(), draw_keys.numel()], dtype=torch.int64, device=dev)
    sizes = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(sizes, mine)
    n_list = [int(s[0]) for s in sizes]
    n_draw = [int(s[1]) for s in sizes]
    # 2) draws of every rank (KBs), then 3) the instance lists (the only large transfer: 4 B per element;
    #    the keys are NOT sent, an element's key is the key of the run it lies in)
    all_dk = _all_gather_padded(draw_keys, n_draw)
    all_dc = _all_gather_padded(draw_counts.to(torch.int64), n_draw)
    all_i = _all_gather_padded(idx.to(torch.int32), n_list)

    # merged draws: union of keys, counts summed over ranks, in key order
    cat_dk = torch.cat(all_dk)
    merged_dk, inverse = torch.unique(cat_dk, sorted=True, return_inverse=True)
    cat_dc = torch.cat(all_dc)
    merged_dc = torch.zeros(merged_dk.numel(), dtype=torch.int64, device=dev).index_add_(0, inverse, cat_dc)
    first = torch.cumsum(merged_dc, 0) - merged_dc

    # run placement: rank r's run of key k starts at first[k] + sum of the counts of ranks < r for k
    total = int(merged_dc.sum())
    out_i = torch.empty(total, dtype=torch.int64, device=dev)
    before = torch.zeros(merged_dk.numel(), dtype=torch.int64, device=dev)
    shard = [int(x) for x in _gather_scalar(idx_offset, dev)]
    off = 0
    for r in range(world):
        nd = n_draw[r]
        if nd:
            slot = inverse[off: off + nd]              # merged draw index of each of rank r's runs
            run_len = all_dc[r]
            run_src = torch.cumsum(run_len, 0) - run_len
            # destination of every element = where its run starts in the merged list + its position in the run
            shift = first[slot] + before[slot] - run_src
            dst = torch.repeat_interleave(shift, run_len) + torch.arange(n_list[r], device=dev)
            out_i[dst] = all_i[r].to(torch.int64) + shard[r]
            before.index_add_(0, slot, run_len)
        off += nd
    out_k = torch.repeat_interleave(merged_dk, merged_dc)
    return out_k, out_i, merged_dk, merged_dc, first
