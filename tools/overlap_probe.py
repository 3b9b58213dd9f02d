"""Probe: do frames of two independent contexts on one GPU overlap (throughput per frame vs one context)?"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from vulkanengine_b200 import scenesubmit as S, synth
n = 1 << 24
sc = synth.flat_scene(n)
views = synth.make_views(S.camera_view_proj, S.cascade_matrices, 5)
for L in (1, 2, 3):
    ctxs = []
    for _ in range(L):
        c = S.SceneSubmit(sc.n_nodes, n, len(sc.slots), max_views=5, max_draws_per_view=1 << 16, frames_in_flight=2)
        c.load_scene(sc); ctxs.append(c)
    for _ in range(6):
        for c in ctxs: c.frame_sync(c.frame(views))
    torch.cuda.synchronize()
    K = 30
    t0 = time.perf_counter()
    for k in range(K):
        ctxs[k % L].frame(views)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / K * 1e3
    print(f"lanes={L}: {dt:.3f} ms/frame")
    for c in ctxs: c.close()
