#!/usr/bin/env python
"""The consumer (raster) pass over one view of configs[2]'s lists, for profiling: python tools/raster_probe.py VIEW [REPS].
View 0 is the id pass at 1920x1080, views 1-4 are the cascades' depth passes at 4096^2 with DepthClamp."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from vulkanengine_b200 import scenesubmit as S  # noqa: E402
from vulkanengine_b200 import synth  # noqa: E402


def main():
    v = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    sc = synth.flat_scene(1 << 24)
    pos, ind, slots = synth.box_geometry(sc.slots)
    sc = synth.SceneArrays(**{**sc.__dict__, "slots": slots})
    views = synth.make_views(S.camera_view_proj, S.cascade_matrices, 5)
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=5, max_draws_per_view=1 << 16)
    ctx.load_scene(sc)
    ctx.set_geometry(pos, ind)
    slot = ctx.frame(views)
    ctx.frame_sync(slot)
    w, h, flags = (1920, 1080, 0) if v == 0 else (4096, 4096, S.RASTER_DEPTH_CLAMP)
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ctx.raster_view(slot, v, w, h, flags)
        torch.cuda.synchronize()
        print(f"view {v}: {(time.perf_counter() - t0) * 1e3:.3f} ms", flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
