#!/usr/bin/env python
"""Frame time of configs[2] (2^24 instances x 5 views) on each pipeline of the library, with the frame's stage
times (CUDA events inside ss_frame).  Usage: python tools/r2_probe.py [log2_n] [frames]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vulkanengine_b200 import scenesubmit as S  # noqa: E402
from vulkanengine_b200 import synth  # noqa: E402


def main():
    log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    frames = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    only = sys.argv[3].split(",") if len(sys.argv) > 3 else None
    n = 1 << log2n
    sc = synth.flat_scene(n)
    views = synth.make_views(S.camera_view_proj, S.cascade_matrices, 5)
    cases = [("general", 0, 0), ("general, reuse sort", 0, S.FRAME_REUSE_SORT), ("resident", S.CONFIG_DRAW_ORDER_RESIDENT, 0)]
    ref_vis = None
    for name, cflags, fflags in cases:
        if only and not any(o in name for o in only):
            continue
        ctx = S.SceneSubmit(sc.n_nodes, n, len(sc.slots), max_views=5, max_draws_per_view=1 << 16, flags=cflags)
        ctx.load_scene(sc)
        for _ in range(4):
            st = ctx.frame_sync(ctx.frame(views, fflags))
        slots = [ctx.frame(views, fflags) for _ in range(3)]
        ms, stages = [], []
        for _ in range(frames):
            slot = ctx.frame(views, fflags)
            st = ctx.frame_sync(slot)
            ms.append(float(st.gpu_ms))
            stages.append(ctx.frame_stage_ms(slot))
        vis = [int(st.visible[v]) for v in range(5)]
        if ref_vis is None:
            ref_vis = vis
        stage = {k: round(float(np.median([x[k] for x in stages])), 4) for k in stages[0]}
        print(f"{name:26s} pipeline={S.PIPELINE_NAMES[int(st.pipeline)]:14s} median {np.median(ms):.4f} ms  min {min(ms):.4f}  "
              f"stages {stage}  launches {int(st.kernel_launches)}  visible_ok {vis == ref_vis}", flush=True)
        ctx.close()


if __name__ == "__main__":
    main()
