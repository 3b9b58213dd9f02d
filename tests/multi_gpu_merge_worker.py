"""One rank of the multi-GPU merge check (run under torch.distributed.run by tests/test_sharding.py and by hand:
python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tests/multi_gpu_merge_worker.py).
Every rank owns a contiguous shard of one scene on its own GPU, runs a frame, opens the other ranks' output buffers
over CUDA IPC and calls ss_merge_views; the merged lists of every rank must equal the lists of ONE context that
holds the whole scene (built on rank 0's GPU and broadcast as arrays)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vulkanengine_b200 import scenesubmit as S  # noqa: E402
from vulkanengine_b200 import synth  # noqa: E402


def dev_to_numpy(ptr, n, dtype):
    nbytes = n * dtype.itemsize
    if nbytes == 0:
        return np.empty(0, dtype)

    class _Dev:
        __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (int(ptr), False), "version": 2}
    return torch.as_tensor(_Dev(), device="cuda").cpu().numpy().view(dtype).copy()


def open_peers(ctx, local_rank, rank, world, max_views, fif=3):
    """peer[r][slot][view] = (idx, draws, draw_keys, counts) device pointers valid in this process."""
    mine = ctx.export_slot_buffers()
    everyone = [None] * world
    dist.all_gather_object(everyone, mine)
    table = []
    for r in range(world):
        if r == rank:
            table.append(None)
            continue
        ptrs = [S.ipc_open(local_rank, h) for h in everyone[r]]
        per_slot = []
        for si in range(fif):
            counts = ptrs[fif * max_views * 3 + si]
            per_slot.append([(ptrs[(si * max_views + v) * 3], ptrs[(si * max_views + v) * 3 + 1], ptrs[(si * max_views + v) * 3 + 2],
                              counts + 8 * v) for v in range(max_views)])
        table.append(per_slot)
    return table


def main():
    rank, world, local_rank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n_total, n_views = int(os.environ.get("SS_MERGE_N", "300000")), 5
    pipe = int(os.environ.get("SS_MERGE_PIPE", "0"))
    views = synth.make_views(S.camera_view_proj, S.cascade_matrices, n_views)
    lo, hi = n_total * rank // world, n_total * (rank + 1) // world
    sc = synth.flat_scene(hi - lo, first=lo, n_keys=300)
    ctx = S.SceneSubmit(sc.n_nodes, hi - lo, len(sc.slots), max_views=n_views, device=local_rank, flags=pipe)
    ctx.load_scene(sc)
    peers_tab = open_peers(ctx, local_rank, rank, world, n_views)
    slot = ctx.frame(views)
    ctx.frame_sync(slot)
    dist.barrier()  # every rank's frame of this slot is complete
    peers = []
    for v in range(n_views):
        row = []
        for r in range(world):
            base = n_total * r // world
            if r == rank:
                row.append(S.peer_of_outputs(ctx.map_outputs(slot, v), base))
            else:
                row.append((*peers_tab[r][slot][v], base))
        peers.append(row)
    mg = S.Merger(world, n_views, max_draws_per_rank=1 << 12, max_visible_per_view=n_total, device=local_rank)
    mg.merge(peers)
    got = []
    for v in range(n_views):
        o = mg.map(v)
        got.append({"idx": dev_to_numpy(o.instance_idx, o.visible, np.dtype(np.uint32)), "keys": dev_to_numpy(o.keys, o.visible, np.dtype(np.uint64)),
                    "draws": dev_to_numpy(o.draws, o.n_draws, S.DRAW_DTYPE), "draw_keys": dev_to_numpy(o.draw_keys, o.n_draws, np.dtype(np.uint64))})
    dist.barrier()  # nobody reuses a slot while a peer still reads it
    # the single-context answer, on rank 0
    ok = True
    ref = [None]
    if rank == 0:
        whole = synth.flat_scene(n_total, n_keys=300)
        one = S.SceneSubmit(whole.n_nodes, n_total, len(whole.slots), max_views=n_views, device=local_rank)
        one.load_scene(whole)
        s1 = one.frame(views)
        st1 = one.frame_sync(s1)
        ref = [[one.download_view(s1, v, st1, want_rows=False) for v in range(n_views)]]
        one.close()
    dist.broadcast_object_list(ref, src=0)
    for v in range(n_views):
        for k in ("idx", "keys", "draws", "draw_keys"):
            if not np.array_equal(got[v][k], ref[0][v][k]):
                ok = False
                print(f"rank {rank} view {v}: merged {k} differ from the single-context lists", flush=True)
    flag = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(flag)
    if rank == 0:
        print(f"multi-GPU merge over {world} ranks, {n_total} instances, pipeline flags {pipe}: "
              f"{'identical to one context' if flag.item() == 0 else 'MISMATCH'}; visible per view {[len(g['idx']) for g in got]}", flush=True)
    mg.close()
    ctx.close()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 0 else 1)


if __name__ == "__main__":
    main()
