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


def _gather_rows(t: torch.Tensor) -> torch.Tensor:
    """all_gather of equally sized 1-D tensors into one [world, n] tensor (one collective, no list of copies)."""
    world = dist.get_world_size()
    out = torch.empty((world, t.numel()), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out.view(-1), t.contiguous())
    return out


def _merge_place(lists, seg_out, seg_src, seg_add, seg_key, n_segs, out_i, out_k, peer_row=None):
    """ss_merge_place / ss_merge_place_peers on the current CUDA stream (the library must be there: no torch fallback
    on a GPU).  peer_row: device tensor of one buffer address per rank; segments are numbered run-major, rank-minor."""
    import ctypes as C

    from . import scenesubmit
    lib = scenesubmit.load_library()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if peer_row is not None:
        world = peer_row.numel()
        seg_rank = torch.arange(world, dtype=torch.int32, device=out_i.device).repeat(n_segs // world).contiguous()
        st = lib.ss_merge_place_peers(stream, C.c_void_p(peer_row.data_ptr()), C.c_void_p(seg_rank.data_ptr()),
                                      C.c_void_p(seg_out.data_ptr()), C.c_void_p(seg_src.data_ptr()),
                                      C.c_void_p(seg_add.data_ptr()), C.c_void_p(seg_key.data_ptr()), C.c_uint32(n_segs),
                                      C.c_void_p(out_i.data_ptr()), C.c_void_p(out_k.data_ptr()))
        if st != 0:
            raise RuntimeError(f"ss_merge_place_peers failed with status {st}")
        return
    st = lib.ss_merge_place(stream, C.c_void_p(lists.data_ptr()),
                            C.c_void_p(seg_out.data_ptr()), C.c_void_p(seg_src.data_ptr()), C.c_void_p(seg_add.data_ptr()),
                            C.c_void_p(seg_key.data_ptr()), C.c_uint32(n_segs), C.c_void_p(out_i.data_ptr()),
                            C.c_void_p(out_k.data_ptr()))
    if st != 0:
        raise RuntimeError(f"ss_merge_place failed with status {st}")


class PeerLists:
    """Every rank's instance-index output buffers, mapped into this process once (CUDA IPC; reads go over NVLink).
    With it merge_view_lists does not all-gather the lists: the placement kernel reads each segment from the buffer of
    the rank that produced it (ss_merge_place_peers).  The caller keeps the usual contract of shared buffers: a slot is
    read by peers between its owner's ss_frame_sync (which precedes the merge call on every rank) and the owner's next
    ss_frame into that slot (which must come after a barrier that follows the merge)."""

    def __init__(self, ctx, device: int):
        from . import scenesubmit
        self._scenesubmit, self.device = scenesubmit, device
        world, rank = dist.get_world_size(), dist.get_rank()
        handles, own = ctx.export_idx_lists()
        everyone = [None] * world
        dist.all_gather_object(everyone, handles)
        self.n_views = ctx.max_views
        self._opened = []
        table = torch.zeros((len(handles), world), dtype=torch.int64)
        for r in range(world):
            for k, h in enumerate(everyone[r]):
                if r == rank:
                    table[k, r] = own[k]
                else:
                    p = scenesubmit.ipc_open(device, h)
                    self._opened.append(p)
                    table[k, r] = p
        self.table = table.to(torch.device("cuda", device))  # [slot * max_views + view][rank] -> device address

    def row(self, slot: int, view: int) -> torch.Tensor:
        return self.table[slot * self.n_views + view]

    def close(self):
        for p in self._opened:
            self._scenesubmit.ipc_close(self.device, p)
        self._opened = []


def merge_view_lists(views, idx_offset: int, peers: "PeerLists" = None, frame_slot: int = 0):
    """Merges this rank's compacted lists of SEVERAL views with every other rank's.

    views        list of (idx, draw_keys, draw_counts) per view: this rank's visible instance list (sorted by key,
                 stable), and its draws: key and instance_count per run.  `idx` may be longer than the list (a
                 device buffer of the library's capacity): only its first sum(draw_counts) elements are read.
    idx_offset   first global instance number of this rank's shard

    Returns, per view, (merged_keys, merged_global_idx, merged_draw_keys, merged_draw_counts, merged_first_instance);
    identical on every rank and identical to what one GPU holding all shards would emit.  Keys are int64 tensors
    (the uint64 bit pattern; keys use at most 62 bits so order is preserved).

    Collectives: one for all the sizes and shard offsets, one for all the draws, one per view for the instance lists
    (4 B per element; the keys are NOT sent, an element's key is the key of the run it lies in).  One host
    synchronisation (the sizes).  The merged list is written in output order: for merged run k, rank 0's elements,
    then rank 1's, ... - a segmented gather, not a sort.
    """
    world = dist.get_world_size()
    V = len(views)
    dev = views[0][0].device
    i64 = dict(dtype=torch.int64, device=dev)
    # 1) sizes of every view on every rank + shard offsets (tiny): the only host sync
    mine = torch.empty(2 * V + 1, **i64)
    for v, (idx, dk, dc) in enumerate(views):
        mine[2 * v] = dc.sum() if dc.numel() else 0
        mine[2 * v + 1] = dk.numel()
    mine[2 * V] = idx_offset
    sizes = _gather_rows(mine).cpu()
    n_list = sizes[:, 0:2 * V:2]          # [world, V]
    n_draw = sizes[:, 1:2 * V:2]
    shard = sizes[:, 2 * V].to(dev)
    cap_draw = int(n_draw.max().clamp(min=1))
    # 2) the draws of all views (KBs): keys and counts side by side, padded to the widest
    packed = torch.zeros((V, 2, cap_draw), **i64)
    for v, (idx, dk, dc) in enumerate(views):
        packed[v, 0, : dk.numel()] = dk
        packed[v, 1, : dc.numel()] = dc.to(torch.int64)
    all_draws = _gather_rows(packed.view(-1)).view(world, V, 2, cap_draw)

    out = []
    for v, (idx, dk, dc) in enumerate(views):
        nd = [int(x) for x in n_draw[:, v]]
        nl = [int(x) for x in n_list[:, v]]
        cap = max(max(nl), 1)
        lists = None
        if peers is None:
            # 3) the instance lists: equal-length slices (the tail past a rank's own count is never read)
            src = idx.to(torch.int32) if idx.dtype != torch.int32 else idx
            if src.numel() < cap:
                src = torch.cat([src, torch.zeros(cap - src.numel(), dtype=torch.int32, device=dev)])
            lists = _gather_rows(src[:cap])                                 # [world, cap] int32

        # merged draws: union of the ranks' keys in key order, counts summed
        keys_r = [all_draws[r, v, 0, : nd[r]] for r in range(world)]
        cnts_r = [all_draws[r, v, 1, : nd[r]] for r in range(world)]
        merged_dk, inverse = torch.unique(torch.cat(keys_r), sorted=True, return_inverse=True)
        K = merged_dk.numel()
        # per (merged run, rank): how many elements and where they start in that rank's list
        seg_len = torch.zeros((K, world), **i64)
        seg_src = torch.zeros((K, world), **i64)
        off = 0
        for r in range(world):
            if nd[r]:
                slot = inverse[off: off + nd[r]]
                seg_len[slot, r] = cnts_r[r]
                seg_src[slot, r] = (torch.cumsum(cnts_r[r], 0) - cnts_r[r]) + (r * cap if peers is None else 0)
            off += nd[r]
        merged_dc = seg_len.sum(1)
        first = torch.cumsum(merged_dc, 0) - merged_dc
        total = int(sum(nl))
        # segments in output order (run-major, rank-minor): element e of the output comes from
        # lists.flat[seg_src + (e - seg_out_start)], shifted by its rank's shard offset
        flat_len = seg_len.view(-1)
        if dev.type == "cuda":
            # on the GPU the placement is one kernel of the library (ss_merge_place); instance numbers come back as int32
            seg_out = torch.zeros(K * world + 1, **i64)
            torch.cumsum(flat_len, 0, out=seg_out[1:])
            out_i = torch.empty(total, dtype=torch.int32, device=dev)
            out_k = torch.empty(total, **i64)
            _merge_place(lists, seg_out, seg_src.view(-1).contiguous(), shard.repeat(K).contiguous(),
                         merged_dk.repeat_interleave(world).contiguous(), K * world, out_i, out_k,
                         None if peers is None else peers.row(frame_slot, v))
        else:
            seg_out = torch.cumsum(flat_len, 0) - flat_len
            gather_from = torch.repeat_interleave(seg_src.view(-1) - seg_out, flat_len, output_size=total)
            gather_from += torch.arange(total, **i64)
            add = torch.repeat_interleave(shard.repeat(K), flat_len, output_size=total)
            out_i = lists.view(-1)[gather_from].to(torch.int64) + add
            out_k = torch.repeat_interleave(merged_dk, merged_dc, output_size=total)
        out.append((out_k, out_i, merged_dk, merged_dc, first))
    return out


def merge_view_list(keys: torch.Tensor, idx: torch.Tensor, draw_keys: torch.Tensor, draw_counts: torch.Tensor,
                    idx_offset: int):
    """One view (see merge_view_lists). `keys` is accepted for symmetry with the library's outputs and not used:
    an element's key is the key of its run."""
    return merge_view_lists([(idx, draw_keys, draw_counts)], idx_offset)[0]
