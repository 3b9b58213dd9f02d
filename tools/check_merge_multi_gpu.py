"""One-off check on N GPUs (torchrun): every rank runs its entity-wise shard through libscenesubmit, the ranks merge their
lists over NCCL (sharding.merge_view_lists, GPU placement kernel), and rank 0 compares with ONE context that holds the
whole scene.   torchrun --nproc-per-node N tools/check_merge_multi_gpu.py [instances_per_rank]"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vulkanengine_b200 import scenesubmit as S  # noqa: E402
from vulkanengine_b200 import sharding, synth  # noqa: E402


class _Dev:
    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def dev_tensor(ptr, n, typestr):
    if n == 0:
        return torch.empty(0, dtype=torch.int64 if typestr == "<i8" else torch.int32, device="cuda")
    return torch.as_tensor(_Dev(ptr, n, typestr), device="cuda")


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 300_000
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    views = synth.make_views(S.camera_view_proj, S.cascade_matrices, 5)
    sc = synth.flat_scene(n, first=rank * n, n_keys=500)
    ctx = S.SceneSubmit(sc.n_nodes, n, len(sc.slots), max_views=5, device=local)
    ctx.load_scene(sc)
    slot = ctx.frame(views)
    ctx.frame_sync(slot)
    lists = []
    for v in range(5):
        o = ctx.map_outputs(slot, v)
        draws = dev_tensor(o.draws, o.n_draws * 5, "<i4").view(-1, 5)
        lists.append((dev_tensor(o.instance_idx, n, "<i4"), dev_tensor(o.draw_keys, o.n_draws, "<i8"), draws[:, 1].contiguous()))
    peers = sharding.PeerLists(ctx, local)
    results = {"all-gather": sharding.merge_view_lists(lists, rank * n),
               "peer reads": sharding.merge_view_lists(lists, rank * n, peers=peers, frame_slot=slot)}
    torch.cuda.synchronize()
    dist.barrier()   # nobody tears its buffers down while a peer may still be reading them
    if rank == 0:
        whole = synth.flat_scene(n * world, first=0, n_keys=500)
        c2 = S.SceneSubmit(whole.n_nodes, n * world, len(whole.slots), max_views=5, device=local)
        c2.load_scene(whole)
        s2 = c2.frame(views)
        st2 = c2.frame_sync(s2)
        for how, v in [(h, v) for h in results for v in range(5)]:
            merged = results[how]
            ref = c2.download_view(s2, v, st2, want_rows=False)
            k, i, dk, dc, first = merged[v]
            assert np.array_equal(i.cpu().numpy().astype(np.int64), ref["idx"].astype(np.int64)), v
            assert np.array_equal(k.cpu().numpy().view(np.uint64), ref["keys"]), v
            assert np.array_equal(dk.cpu().numpy().view(np.uint64), ref["draw_keys"]), v
            assert np.array_equal(dc.cpu().numpy(), ref["draws"]["instance_count"].astype(np.int64)), v
            assert np.array_equal(first.cpu().numpy(), ref["draws"]["first_instance"].astype(np.int64)), v
        print(f"merge over {world} GPUs == one GPU holding all {n * world} instances: 5 views identical, both exchange paths")
    dist.barrier()
    peers.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
