"""Multi-GPU host logic on CPU: two gloo ranks shard a scene entity-wise, each builds its per-view lists
(from the oracle here: there is no GPU), the merge step must reproduce the single-process lists."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_lib as O
from vulkanengine_b200 import sharding, synth

N_TOTAL = 6000
N_VIEWS = 3


def _views():
    return synth.make_views(lambda *a: O.camera_view_proj(*a), lambda *a: O.cascade_matrices(*a), N_VIEWS)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = sharding.shard_range(N_TOTAL, rank, world)
    sc = synth.flat_scene(hi - lo, first=lo, n_keys=200)
    res = O.OracleScene(sc, 0).frame(_views())
    for v in range(N_VIEWS):
        r = res["views"][v]
        keys = torch.from_numpy(r["keys"].view(np.int64).copy())
        idx = torch.from_numpy(r["idx"].astype(np.int64))
        dk = torch.from_numpy(r["draw_keys"].view(np.int64).copy())
        dc = torch.from_numpy(r["draws"]["instance_count"].astype(np.int64))
        mk, mi, mdk, mdc, mfirst = sharding.merge_view_list(keys, idx, dk, dc, lo)
        np.savez(os.path.join(out_dir, f"r{rank}_v{v}.npz"), k=mk.numpy(), i=mi.numpy(), dk=mdk.numpy(),
                 dc=mdc.numpy(), first=mfirst.numpy())
    dist.destroy_process_group()


def test_shard_ranges_cover_everything():
    for world in (1, 2, 3, 8):
        r = [sharding.shard_range(1000003, k, world) for k in range(world)]
        assert r[0][0] == 0 and r[-1][1] == 1000003
        assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))


def test_counter_rng_shards_are_slices_of_one_scene():
    whole = synth.flat_scene(N_TOTAL, n_keys=200)
    lo, hi = sharding.shard_range(N_TOTAL, 1, 2)
    part = synth.flat_scene(hi - lo, first=lo, n_keys=200)
    assert np.array_equal(part.trs, whole.trs[lo:hi]) and np.array_equal(part.slot, whole.slot[lo:hi])
    assert np.array_equal(part.matword, whole.matword[lo:hi]) and np.array_equal(part.inst_aabb, whole.inst_aabb[lo:hi])


def test_two_rank_merge_equals_single_process(tmp_path):
    world = 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    whole = O.OracleScene(synth.flat_scene(N_TOTAL, n_keys=200), 0).frame(_views())
    for v in range(N_VIEWS):
        ref = whole["views"][v]
        for rank in range(world):
            g = np.load(os.path.join(str(tmp_path), f"r{rank}_v{v}.npz"))
            assert np.array_equal(g["k"].view(np.uint64), ref["keys"])
            assert np.array_equal(g["i"], ref["idx"].astype(np.int64))
            assert np.array_equal(g["dk"].view(np.uint64), ref["draw_keys"])
            assert np.array_equal(g["dc"], ref["draws"]["instance_count"].astype(np.int64))
            assert np.array_equal(g["first"], ref["draws"]["first_instance"].astype(np.int64))


@pytest.mark.gpu
def test_merge_place_kernel_against_numpy():
    """ss_merge_place (the GPU placement step of the merge): random segments, empty ones among them, against the
    definition out[e] = lists[seg_src[s] + (e - seg_out[s])] + seg_add[s], keys[e] = seg_key[s]."""
    rng = np.random.default_rng(4)
    n_segs = 5000
    lens = rng.integers(0, 400, n_segs)
    lens[rng.random(n_segs) < 0.3] = 0
    lens[17] = 100_000
    seg_out = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    total = int(seg_out[-1])
    lists = rng.integers(0, 1 << 20, total + 1000).astype(np.int32)
    seg_src = np.array([rng.integers(0, len(lists) - l + 1) for l in lens], np.int64)
    seg_add = rng.integers(0, 1 << 24, n_segs).astype(np.int64)
    seg_key = rng.integers(0, 1 << 60, n_segs).astype(np.int64)
    want_i = np.concatenate([lists[s:s + l].astype(np.int64) + a for s, l, a in zip(seg_src, lens, seg_add)])
    want_k = np.repeat(seg_key, lens)
    dev = torch.device("cuda", 0)
    t = lambda x: torch.from_numpy(x).to(dev)
    out_i = torch.empty(total, dtype=torch.int32, device=dev)
    out_k = torch.empty(total, dtype=torch.int64, device=dev)
    sharding._merge_place(t(lists), t(seg_out), t(seg_src), t(seg_add), t(seg_key), n_segs, out_i, out_k)
    torch.cuda.synchronize()
    assert np.array_equal(out_i.cpu().numpy().astype(np.int64), want_i)
    assert np.array_equal(out_k.cpu().numpy(), want_k)


def dev_to_numpy(ptr, n, dtype):
    """n elements of `dtype` at device address ptr, through torch (bytes, then reinterpreted)."""
    nbytes = n * dtype.itemsize
    if nbytes == 0:
        return np.empty(0, dtype)

    class _Dev:
        __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (int(ptr), False), "version": 2}
    return torch.as_tensor(_Dev(), device="cuda").cpu().numpy().view(dtype).copy()


@pytest.mark.gpu
@pytest.mark.parametrize("shards,pipe", [(2, 0), (3, 0), (4, 2)])
def test_merge_views_equals_single_context(shards, pipe):
    """ss_merge_views: the merged lists of G entity-wise shards (G contexts on one GPU standing in for G ranks; the
    pointers are the ranks' own output buffers) equal what ONE context holding the whole scene produces: instance
    order, keys, indirect draws, draw keys, every view, in one library call."""
    import ctypes as C

    import numpy as np

    from vulkanengine_b200 import scenesubmit as S
    from vulkanengine_b200 import synth

    n_total, n_views = 90_000, 5
    views = synth.make_views(S.camera_view_proj, S.cascade_matrices, n_views)
    whole = synth.flat_scene(n_total, n_keys=300)
    one = S.SceneSubmit(whole.n_nodes, n_total, len(whole.slots), max_views=n_views, flags=pipe)
    one.load_scene(whole)
    slot1 = one.frame(views)
    st1 = one.frame_sync(slot1)
    ref = [one.download_view(slot1, v, st1) for v in range(n_views)]

    bounds = [n_total * r // shards for r in range(shards + 1)]
    ctxs, slots = [], []
    for r in range(shards):
        n = bounds[r + 1] - bounds[r]
        sc = synth.flat_scene(n, first=bounds[r], n_keys=300)
        c = S.SceneSubmit(sc.n_nodes, n, len(sc.slots), max_views=n_views, flags=pipe)
        c.load_scene(sc)
        slots.append(c.frame(views))
        c.frame_sync(slots[-1])
        ctxs.append(c)
    peers = [[S.peer_of_outputs(ctxs[r].map_outputs(slots[r], v), bounds[r]) for r in range(shards)] for v in range(n_views)]
    mg = S.Merger(shards, n_views, max_draws_per_rank=1 << 12, max_visible_per_view=n_total)
    for _ in range(2):  # twice: the tables are reused
        mg.merge(peers)
    stage = mg.stage_ms()  # ss_merge_stage_ms: CUDA-event times of the four kernels
    assert set(stage) == {"fetch", "rank", "segments", "place"} and all(0.0 <= x < 1e3 for x in stage.values()), stage
    for v in range(n_views):
        o = mg.map(v)
        assert o.visible == len(ref[v]["idx"]) and o.n_draws == len(ref[v]["draws"]), v

        def dl(ptr, n, dt):
            return dev_to_numpy(ptr, n, np.dtype(dt))
        assert np.array_equal(dl(o.instance_idx, o.visible, np.uint32), ref[v]["idx"]), f"view {v}: merged instance order"
        assert np.array_equal(dl(o.keys, o.visible, np.uint64), ref[v]["keys"]), f"view {v}: merged keys"
        assert np.array_equal(dl(o.draws, o.n_draws, S.DRAW_DTYPE), ref[v]["draws"]), f"view {v}: merged draws"
        assert np.array_equal(dl(o.draw_keys, o.n_draws, np.uint64), ref[v]["draw_keys"]), f"view {v}: merged draw keys"
    mg.close()
    for c in ctxs:
        c.close()
    one.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipe", [0, 2])
def test_merge_views_across_gpus(pipe):
    """The same over real peers: one process per GPU (torch.distributed.run), shards on different GPUs, the other ranks'
    buffers opened over CUDA IPC and read across NVLink by ss_merge_views.  Needs two GPUs; skipped on one."""
    import subprocess
    n_gpus = torch.cuda.device_count()
    if n_gpus < 2:
        pytest.skip("one GPU: the multi-GPU merge check needs two")
    world = min(n_gpus, 4)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ, SS_MERGE_PIPE=str(pipe))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(os.path.dirname(os.path.abspath(__file__)), "multi_gpu_merge_worker.py")]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "identical to one context" in r.stdout
