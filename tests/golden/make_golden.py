"""Generates tests/golden/*.npz from the reference-arithmetic build (oracle/_ref/libss_ref.so: the
reference's vendored GLM 0.9.9.9 + entt 3.15.0 compiled from /root/reference).  Run in the build
container (where /root/reference exists):  python tests/golden/make_golden.py
The vectors let the GPU box (no /root/reference) pin the oracle and the product's host functions."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as O  # noqa: E402
from golden_cases import entt_scene_cases, run_ref_pass  # noqa: E402


def main():
    assert O.ref_available(), "build oracle/_ref first: make -C oracle ref"
    rng = np.random.default_rng(20261106)

    # A1: TRS -> mat4 through glm (both trig modes); includes signed zeros, negative scale, big angles
    n = 4096
    trs = np.concatenate([rng.uniform(-500, 500, (n, 3)), rng.uniform(-np.pi, np.pi, (n, 3)),
                          rng.uniform(-2, 2, (n, 3))], axis=1).astype(np.float32)
    trs[0] = [0, 0, 0, 0, 0, 0, 1, 1, 1]
    trs[1] = [1, 2, 3, 0, 0, 0, -1, 1, 1]
    trs[2] = [-0.0, 0.0, -0.0, -0.0, 0.0, 0.0, 1, -1, 1]
    trs[3] = [10, 20, 30, 100.0, -250.0, 1000.0, 0.5, 2, 3]
    mats = {m: np.stack([O.trs_to_mat4(t, m, "ref") for t in trs]) for m in (0, 1)}
    np.savez_compressed(os.path.join(HERE, "trs_to_mat4.npz"), trs=trs, mat_det=mats[0], mat_libm=mats[1])

    # A9: EditorCamera view-projection + UpdateCascadeMap through glm
    cams, outs, snapped = [], [], []
    for _ in range(64):
        fov = rng.uniform(0.5, 1.6)
        aspect = rng.uniform(0.8, 2.2)
        zn, zf = 0.1, float(rng.choice([100.0, 1000.0, 2000.0]))
        pos = rng.uniform(-700, 700, 3).astype(np.float32)
        pitch, yaw = rng.uniform(-1.5, 1.5), rng.uniform(-3.1, 3.1)
        lam = float(rng.choice([0.5, 0.92, 0.95]))
        light = rng.uniform(-1, 1, 3).astype(np.float32)
        light[1] = -abs(light[1]) - 0.1
        v, p, vp = O.camera_view_proj(fov, aspect, zn, zf, pos, pitch, yaw, "ref")
        splits, lm = O.cascade_matrices(vp, zn, zf, lam, -75.0, 75.0, light, "ref")
        cams.append(np.concatenate([[fov, aspect, zn, zf], pos, [pitch, yaw, lam], light]).astype(np.float32))
        outs.append(np.concatenate([v, p, vp, splits, lm.reshape(-1)]).astype(np.float32))
        # the CascadeOffset branch (SceneRenderer.cpp:662-679): CascadeOrigin at the world origin and at the camera, 4096^2 map
        for org in (np.zeros(3, np.float32), pos):
            s2, lm2 = O.cascade_matrices(vp, zn, zf, lam, -75.0, 75.0, light, "ref", origin=org, map_size=4096.0)
            snapped.append(np.concatenate([org, s2, lm2.reshape(-1)]).astype(np.float32))
    # the reference's own defaults (EditorLayer.cpp:75, SceneRenderer.h:173-181, Components.h:144)
    np.savez_compressed(os.path.join(HERE, "views.npz"), cams=np.stack(cams), outs=np.stack(outs), snapped=np.stack(snapped))

    # A2/A4/A6: a whole submission pass over an entt registry with std::map<MeshKey> grouping
    packs = {}
    for name, case in entt_scene_cases().items():
        res = run_ref_pass(case, trig_mode=0)
        for k, v in res.items():
            packs[f"{name}__{k}"] = v
    np.savez_compressed(os.path.join(HERE, "entt_pass.npz"), **packs)
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
