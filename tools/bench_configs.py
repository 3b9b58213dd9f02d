"""Timings of the BASELINE.json configs that are NOT the bench line (bench.py times configs[2]):
    cfg2  synthetic 1M instances, 3-level hierarchy, 1 camera          (configs[1])
    cfg3h configs[2] with the 3-level hierarchy variant (2^24 leaves)  (SURVEY.md 8d "the same 3-level variant")
    cfg3s configs[2], steady state: static streams unchanged, sorted order kept (SS_FRAME_REUSE_SORT)
    raster the consumer pass over configs[2]'s lists: entity-id pass of the camera view at 1920x1080 and the depth pass of
          each shadow cascade at 4096x4096 (CSMSettings::MapSize, SceneRenderer.h:179), one box (12 triangles) per instance
    cfg5  one GPU's share of configs[4]: 512 cameras x 1M instances in batches of 8 views, world rows and sorted
          order kept after the first batch (SS_FRAME_REUSE_WORLD | SS_FRAME_REUSE_SORT)
One JSON line per config on stdout.  CUDA events on the library's stream, after warm-up.  Not a bench.py
replacement: nothing here feeds BENCH/SCALE json.

    python tools/bench_configs.py [cfg2 cfg3h cfg3s cfg5] [--views-per-gpu 512]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vulkanengine_b200 import synth  # noqa: E402


def timed(torch, stream, fn, reps):
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        ev0.record(stream)
        out = [fn(i) for i in range(reps)]
        ev1.record(stream)
    ev1.synchronize()
    return ev0.elapsed_time(ev1), out


def frames(S, torch, sc, views, reps, flags=0, max_views=None, warm=5):
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=max_views or len(views),
                        max_draws_per_view=1 << 16)
    ctx.load_scene(sc)
    for _ in range(warm):
        st = ctx.frame_sync(ctx.frame(views))
    stream = torch.cuda.ExternalStream(ctx.stream, device=0)
    ms, slots = timed(torch, stream, lambda i: ctx.frame(views, flags), reps)
    st = ctx.frame_sync(slots[-1])
    stage = ctx.frame_stage_ms(slots[-1])
    res = {"ms_per_frame": ms / reps, "instances_per_s": sc.n_instances / (ms / reps * 1e-3),
           "visible": [int(st.visible[v]) for v in range(len(views))],
           "draws": [int(st.draws[v]) for v in range(len(views))], "launches": int(st.kernel_launches),
           "stage_ms": {k: round(float(v), 4) for k, v in stage.items()}}
    ctx.close()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("which", nargs="*", default=["cfg2", "cfg3h", "cfg3s", "cfg5"])
    ap.add_argument("--views-per-gpu", type=int, default=512)
    ap.add_argument("--reps", type=int, default=50)
    a = ap.parse_args()
    import torch
    from vulkanengine_b200 import scenesubmit as S
    if not torch.cuda.is_available():
        raise SystemExit("no CUDA device; libscenesubmit has no CPU path")
    torch.cuda.set_device(0)

    if "cfg2" in a.which:
        sc = synth.hierarchy_scene()
        views = synth.make_views(S.camera_view_proj, S.cascade_matrices, 1)
        r = frames(S, torch, sc, views, a.reps)
        print(json.dumps({"config": "cfg2: 2^20 leaves, 3-level hierarchy, 1 camera", "nodes": sc.n_nodes, **r}), flush=True)
    if "cfg3h" in a.which:
        sc = synth.hierarchy_scene(n_roots=1 << 14, seed=synth.SEED_BASE + 3)
        views = synth.make_views(S.camera_view_proj, S.cascade_matrices, 5)
        r = frames(S, torch, sc, views, a.reps)
        print(json.dumps({"config": "cfg3 hierarchy variant: 2^24 leaves under 2^14 roots x 32 x 32, 5 views",
                          "nodes": sc.n_nodes, **r}), flush=True)
        del sc
    if "cfg3s" in a.which:
        sc = synth.flat_scene(1 << 24)
        views = synth.make_views(S.camera_view_proj, S.cascade_matrices, 5)
        r = frames(S, torch, sc, views, a.reps, flags=S.FRAME_REUSE_SORT)
        print(json.dumps({"config": "cfg3 steady state: 2^24 flat, 5 views, sorted order kept (static keys)", **r}), flush=True)
        del sc
    if "raster" in a.which:
        sc = synth.flat_scene(1 << 24)
        pos, ind, slots = synth.box_geometry(sc.slots)
        sc = synth.SceneArrays(**{**sc.__dict__, "slots": slots})
        views = synth.make_views(S.camera_view_proj, S.cascade_matrices, 5)
        ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=5, max_draws_per_view=1 << 16)
        ctx.load_scene(sc)
        ctx.set_geometry(pos, ind)
        slot = ctx.frame(views)
        st = ctx.frame_sync(slot)
        stream = torch.cuda.ExternalStream(ctx.stream, device=0)
        res = {}
        for v, (w, h, flags) in enumerate([(1920, 1080, 0)] + [(4096, 4096, S.RASTER_DEPTH_CLAMP)] * 4):
            ctx.raster_view(slot, v, w, h, flags)   # warm-up (allocates the target on a size change)
            ms, _ = timed(torch, stream, lambda i: ctx.raster_view(slot, v, w, h, flags), 5)
            img = ctx.download_raster(v)
            n_vis = int(st.visible[v])
            res[f"view{v}"] = {"target": [w, h], "instances": n_vis, "triangles": n_vis * 12, "ms": round(ms / 5, 4),
                               "mtri_per_s": round(n_vis * 12 / (ms / 5 * 1e-3) / 1e6, 1),
                               "covered_samples": int(np.count_nonzero(img != np.uint64(S.RASTER_EMPTY)))}
        print(json.dumps({"config": "consumer pass over cfg3's lists: id pass 1920x1080 + 4 cascade depth passes 4096^2", **res}), flush=True)
        ctx.close()
        del sc
    if "cfg5" in a.which:
        sc = synth.hierarchy_scene()
        n_cams = a.views_per_gpu
        cams = synth.fibonacci_cameras(S.camera_view_proj, 4096)[:n_cams]
        ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=8, max_draws_per_view=1 << 16)
        ctx.load_scene(sc)
        batches = [np.ascontiguousarray(cams[b:b + 8]) for b in range(0, n_cams, 8)]
        reuse = S.FRAME_REUSE_WORLD | S.FRAME_REUSE_SORT
        for b in batches[:4]:
            ctx.frame_sync(ctx.frame(b))
        stream = torch.cuda.ExternalStream(ctx.stream, device=0)
        ms, slots = timed(torch, stream, lambda i: ctx.frame(batches[i], 0 if i == 0 else reuse), len(batches))
        st = ctx.frame_sync(slots[-1])
        stage = ctx.frame_stage_ms(slots[-1])
        vis = []
        for i in range(max(0, len(batches) - 3), len(batches)):  # the slots still holding their frames
            s_ = ctx.frame_sync(slots[i])
            vis += [int(s_.visible[v]) for v in range(len(batches[i]))]
        iv = n_cams * sc.n_instances
        print(json.dumps({"config": f"cfg5 share of one GPU: {n_cams} cameras x 2^20 instances, 8 views per frame",
                          "ms_total": ms, "ms_per_batch": ms / len(batches), "views_per_s": n_cams / (ms * 1e-3),
                          "instance_views_per_s": iv / (ms * 1e-3), "mean_visible_last_batches": float(np.mean(vis)),
                          "launches_last_batch": int(st.kernel_launches),
                          "stage_ms_last_batch": {k: round(float(v), 4) for k, v in stage.items()}}), flush=True)
        ctx.close()


if __name__ == "__main__":
    main()
