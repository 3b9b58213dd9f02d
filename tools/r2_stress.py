#!/usr/bin/env python
"""Repeat a frame many times on one pipeline and compare every frame's masks, counts and camera-view list with the
three-kernel pipeline's: catches rare races that a single parity run misses.
Usage: python tools/r2_stress.py <config flags> <n instances> <frames> [hier]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vulkanengine_b200 import scenesubmit as S  # noqa: E402
from vulkanengine_b200 import synth  # noqa: E402


def main():
    flags, n, frames = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    hier = len(sys.argv) > 4 and sys.argv[4] == "hier"
    if hier:
        r = max(1, n // (16 * 128))
        sc = synth.hierarchy_scene(n_roots=r, fan1=16, fan2=128, n_keys=200)
    else:
        sc = synth.flat_scene(n, n_keys=200)
    n = sc.n_instances
    views = synth.make_views(S.camera_view_proj, S.cascade_matrices, 5)

    def make(fl):
        c = S.SceneSubmit(sc.n_nodes, n, len(sc.slots), max_views=5, max_draws_per_view=1 << 16, flags=fl)
        c.load_scene(sc)
        return c
    ref = make(0)
    slot = ref.frame(views)
    st = ref.frame_sync(slot)
    rmask = ref.download_masks()
    rvis = [int(st.visible[v]) for v in range(5)]
    ridx = ref.download_view(slot, 0, st)["idx"]
    ref.close()
    ctx = make(flags)
    bad = 0
    for f in range(frames):
        slot = ctx.frame(views)
        st = ctx.frame_sync(slot)
        vis = [int(st.visible[v]) for v in range(5)]
        ok = vis == rvis and np.array_equal(ctx.download_masks(), rmask)
        if ok and f % 8 == 0:
            ok = np.array_equal(ctx.download_view(slot, 0, st)["idx"], ridx)
        bad += 0 if ok else 1
    print(f"flags {flags} n {n} hier {hier}: {bad} bad frames of {frames} "
          f"(pipeline {S.PIPELINE_NAMES[int(st.pipeline)]})", flush=True)
    ctx.close()
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
