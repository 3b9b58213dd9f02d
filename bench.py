#!/usr/bin/env python
"""bench.py — the scene-submission hot path on B200 (BASELINE.json: culled+compacted instances/sec and
ms/frame @16M instances, 5 views).

A step = one frame of the hot path (transform -> cull -> key -> sort -> compact -> emit) over one
synthetic scene of configs[2] ("synthetic 16M instances, main camera + 4 cascaded shadow-map
frusta"), 2^24 instances per GPU.  N > 1 GPUs: entity-wise shards, 2^24 instances per rank (weak
scaling; N = 4 is configs[3]'s 64M), no data-path collective in the sharded phase.

  value      instances/s with every input stream resident in HBM, general pipeline: the keys are sorted
             every frame (CUDA events on the frame stream, max over ranks)
  e2e        the same metric through the C ABI from HOST buffers: pinned TRS upload (36 B/instance) +
             view matrices in, per-view indirect draws + draw keys + counts out, every step; beside it
             e2e.dirty_tenth (a tenth of the transforms sent as sparse records) and e2e.h2d_probe (the
             host's copy bandwidth with all ranks copying at once: the ceiling of the full upload)
  roofline   the frame's longest kernel (k_emit: compaction + gather of the world rows) against the
             measured HBM peak; roofline_other_kernels holds the transform+cull+key kernel the same way
  steady_state   the draw-order-resident layout (SS_CONFIG_DRAW_ORDER_RESIDENT): sorted when the static
             streams change, not per frame.  Reported beside `value`, never instead of it
  merge      N > 1: one globally sorted list per view (ss_merge_views over CUDA IPC peer mappings)
  strong_64m, cfg5, other_configs   configs[3] over the N GPUs, configs[4] (64 views), configs[0/1]
  raster     one GPU: the consumer pass over the frame's lists (id pass + cascade depth passes, csrc/raster.cu)
  cpu_baseline / --impl reference
             the reference's CPU path (entt-order walk, glm T*R*S, std::map<MeshKey> grouping, plus
             the same cull) from oracle/_ref on the box's host cores (--impl reference: the whole scene)
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_PER_GPU = 1 << 24
N_VIEWS = 5
WORKLOAD = "synthetic 16M instances, main camera + 4 cascaded shadow-map frusta (configs[2])"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--instances", type=int, default=N_PER_GPU, help="instances per GPU (default 2^24)")
    ap.add_argument("--views", type=int, default=N_VIEWS)
    ap.add_argument("--trig", default="deterministic", choices=["deterministic", "libm"])
    ap.add_argument("--cpu-sample", type=int, default=0,
                    help="instances in the CPU sample (0: the whole scene for --impl reference, 2^20 for the cpu_baseline leg)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip steady state, strong scaling, cfg5 and the other configs")
    return ap.parse_args()


TRIG_NOTE = {"native": "deterministic polynomial sin/cos (include/scenesubmit_dettrig.h), bit-identical to the oracle",
             "reference": "libm sinf/cosf through glm (GLM/detail/type_quat.inl:208-217)"}


def bench_config(n, n_views, seed, world):
    """The same `config` object in both arms (the driver compares them)."""
    return {"workload": WORKLOAD, "instances_per_gpu": n, "views": n_views, "keys": 4096, "seed": seed,
            "l2": "inputs_exceed_l2 (>1 GB of streams per frame)",
            "sharding": "entity-wise, no collective in the sharded phase" if world > 1 else "none"}


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) >= 9:
                for k, nm in enumerate(names):
                    if r[5 + k].lower().startswith("active"):
                        reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def algorithmic_bytes(n, n_views, visible, draws, hierarchical):
    """SURVEY.md §8d B_alg: compulsory I/O of one frame."""
    per_inst = 36 + 8 + 24 + (4 if hierarchical else 0)
    b = n * per_inst + n * 48 + n * ((n_views + 7) // 8)
    b += sum(int(v) * (48 + 8 + 4) for v in visible) + sum(int(d) * 20 for d in draws)
    return b


def instances_kernel_bytes(n, n_views, hierarchical):
    """Algorithmic bytes of the transform+cull+key launch: the streams it must read and the world rows and
    masks it must write (the sort-key scratch it also writes is an intermediate and is not counted)."""
    return n * (36 + 8 + 24 + (4 if hierarchical else 0)) + n * 48 + n * ((n_views + 7) // 8)


def emit_kernel_bytes(n, visible, unique_visible):
    """Algorithmic bytes of the emit launch: the sorted (key|instance) words it must read (8 B each), one 48 B
    world row per instance visible in at least one view, and per view 48 B row + 8 B key + 4 B index out."""
    return n * 8 + unique_visible * 48 + sum(int(v) * (48 + 8 + 4) for v in visible)


def ncu_traffic(n, n_views):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture of
    this workload (profiles/r01_ncu_full_summary.csv); only valid for the workload it was taken on."""
    out = {}
    if n != N_PER_GPU or n_views != N_VIEWS:
        return out
    try:
        import csv
        name = "r02_ncu_full_summary.csv" if os.path.exists(os.path.join(ROOT, "profiles", "r02_ncu_full_summary.csv")) else \
            "r01_ncu_full_summary.csv"
        out["source"] = f"profiles/{name} (one ncu --set full capture of this workload; not re-measured in this run)"
        rows = list(csv.reader(open(os.path.join(ROOT, "profiles", name))))
        hdr = rows[0]
        ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        iname = hdr.index("Kernel Name") if "Kernel Name" in hdr else 0
        irep = hdr.index("report") if "report" in hdr else None
        for r in rows[2:]:
            if irep is not None and "general" not in r[irep]:
                continue  # the capture of the general pipeline's frame (the resident frame is in the same file)
            for key in ("k_instances", "k_emit"):
                if key in r[iname] and key not in out:
                    out[key] = int((float(r[ir]) + float(r[iw])) * 1e9)
    except Exception:
        pass
    return out


def bind_near_gpu(torch, device):
    """Run this rank on the CPUs next to its GPU (NVML's affinity for the device), so that the pinned upload buffer it
    allocates afterwards lies in that socket's memory: on a two-socket box eight ranks uploading at once otherwise pull
    half their bytes across the socket link.  Returns a short description for the bench line; a no-op where NVML or
    the affinity call is not available."""
    try:
        import pynvml
        pynvml.nvmlInit()
        props = torch.cuda.get_device_properties(device)
        bus = "%08x:%02x:%02x.0" % (getattr(props, "pci_domain_id", 0), props.pci_bus_id, props.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        pynvml.nvmlDeviceSetCpuAffinity(h)
        cpus = sorted(os.sched_getaffinity(0))
        return {"cpus": f"{cpus[0]}-{cpus[-1]}", "n": len(cpus)}
    except Exception as e:  # noqa: BLE001
        return {"unbound": type(e).__name__}


def cpu_reference(sc, views, sample, threads, min_seconds=10.0):
    """Times oracle/_ref (kind 'reference') or the plain-C oracle (kind 'port') on the first `sample`
    instances of the same scene with `threads` host threads. Returns (instances/s, kind, seconds, n)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    n = min(sample, sc.n_instances)
    if O.ref_available():
        class Flat(C.Structure):
            _fields_ = [("n", C.c_uint32), ("trs", C.c_void_p), ("parent", C.c_void_p), ("n_nodes", C.c_uint32),
                        ("node_base", C.c_uint32), ("slot", C.c_void_p), ("matword", C.c_void_p),
                        ("inst_aabb", C.c_void_p), ("slot_local", C.c_void_p), ("slot_mesh_handle", C.c_void_p),
                        ("slot_submesh", C.c_void_p), ("material_handle", C.c_void_p), ("trig_mode", C.c_int)]
        trs = np.ascontiguousarray(sc.trs[:sc.node_base + n])
        slot = np.ascontiguousarray(sc.slot[:n])
        mw = np.ascontiguousarray(sc.matword[:n])
        box = np.ascontiguousarray(sc.inst_aabb[:n])
        local = np.ascontiguousarray(sc.slots["local"])
        mh = np.ascontiguousarray(sc.mesh_handles[sc.slots["mesh_rank"]])
        sub = np.ascontiguousarray(sc.slots["submesh_index"])
        mat = np.ascontiguousarray(sc.material_handles)
        parent = None if sc.parent is None else np.ascontiguousarray(sc.parent[:sc.node_base + n])
        p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
        f = Flat(n, p(trs), p(parent), len(trs), sc.node_base, p(slot), p(mw), p(box), p(local), p(mh), p(sub), p(mat), 1)
        vps = np.ascontiguousarray(views["view_proj"])
        vfl = np.ascontiguousarray(views["flags"])
        vis = np.zeros(len(views), np.uint32)
        drw = np.zeros(len(views), np.uint32)
        fn = O.ref().ssr_flat_frame
        best, total, reps = None, 0.0, 0
        while (total < min_seconds or reps == 0) and reps < 50:
            s = fn(C.byref(f), p(vps), p(vfl), C.c_uint32(len(views)), C.c_int(threads), p(vis), p(drw), None, None)
            best = s if best is None else min(best, s)
            total += s
            reps += 1
        return n / best, "reference", best, n
    sub_sc = type(sc)(**{**sc.__dict__, "trs": sc.trs[:sc.node_base + n], "slot": sc.slot[:n], "matword": sc.matword[:n],
                         "inst_aabb": sc.inst_aabb[:n],
                         "parent": None if sc.parent is None else sc.parent[:sc.node_base + n]})
    osc = O.OracleScene(sub_sc, 1)
    best, total, reps = None, 0.0, 0
    while (total < min_seconds or reps == 0) and reps < 50:
        s = osc.frame(views, threads=threads, want_rows=True)["seconds"]
        best = s if best is None else min(best, s)
        total += s
        reps += 1
    return n / best, "port", best, n


def timed_frames(torch, ctx, stream, views, steps, flags=0):
    """`steps` frames back to back through the slot ring, CUDA events on the library's stream; (ms total, slots)."""
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    slots = []
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for _ in range(steps):
            slots.append(ctx.frame(views, flags))
        ev1.record(stream)
    ev1.synchronize()
    return ev0.elapsed_time(ev1), slots


def max_over_ranks(torch, dist, world, x):
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def open_merge_peers(S, dist, ctx, local_rank, rank, world, max_views, fif=3):
    """peer[r][slot][view] = (idx, draws, draw_keys, counts) pointers of rank r's outputs as this process sees them
    (CUDA IPC, opened once); None for this rank itself."""
    everyone = [None] * world
    dist.all_gather_object(everyone, ctx.export_slot_buffers())
    table = []
    for r in range(world):
        if r == rank:
            table.append(None)
            continue
        ptrs = [S.ipc_open(local_rank, h) for h in everyone[r]]
        table.append([[(ptrs[(si * max_views + v) * 3], ptrs[(si * max_views + v) * 3 + 1], ptrs[(si * max_views + v) * 3 + 2],
                        ptrs[fif * max_views * 3 + si] + 8 * v) for v in range(max_views)] for si in range(fif)])
    return table


def close_merge_peers(S, local_rank, table, max_views, fif=3):
    for per_rank in table:
        if per_rank is None:
            continue
        for si in range(fif):
            for v in range(max_views):
                for ptr in per_rank[si][v][:3]:
                    S.ipc_close(local_rank, ptr)
            S.ipc_close(local_rank, per_rank[si][0][3])


def bench_merge(torch, dist, S, ctx, views, n, rank, local_rank, world, n_views):
    """The one communicating step: ss_merge_views builds the merged list of every view on every rank, reading the other
    ranks' instance lists in place over NVLink.  Timed on the device, max over ranks; not part of `value`."""
    table = open_merge_peers(S, dist, ctx, local_rank, rank, world, n_views)
    mg = S.Merger(world, n_views, max_draws_per_rank=1 << 16, max_visible_per_view=min(world * n, (1 << 32) - 1), device=local_rank)
    slot = ctx.frame(views)
    st = ctx.frame_sync(slot)
    dist.barrier()
    peers = [[S.peer_of_outputs(ctx.map_outputs(slot, v), r * n) if r == rank else (*table[r][slot][v], r * n)
              for r in range(world)] for v in range(n_views)]
    mg.merge(peers)  # warm-up
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        mg.merge(peers, stream=torch.cuda.current_stream().cuda_stream)
    e1.record()
    e1.synchronize()
    ms = max_over_ranks(torch, dist, world, e0.elapsed_time(e1) / reps)
    stage = mg.stage_ms()
    merged = [mg.map(v) for v in range(n_views)]
    total = sum(int(m.visible) for m in merged)
    mine = sum(int(st.visible[v]) for v in range(n_views))
    t = torch.tensor([mine], dtype=torch.int64, device="cuda")
    dist.all_reduce(t)
    remote_bytes = (total - mine) * 4
    dist.barrier()
    mg.close()
    close_merge_peers(S, local_rank, table, n_views)
    return {"ms": ms, "stage_ms_rank0": stage, "merged_elements": total, "equals_sum_of_ranks": bool(int(t.item()) == total),
            "list_bytes_over_nvlink_per_rank": int(remote_bytes),
            "nvlink_gbs_per_rank": remote_bytes / (ms * 1e-3) / 1e9 if ms > 0 else None,
            "note": "ss_merge_views (C ABI, four kernels for all views, no host sync): every rank ends with the merged list of "
                    "every view; instance lists are read in place over NVLink (CUDA IPC peers); not part of `value`"}


def bench_cfg5(torch, S, synth, local_rank, rank, world, total_views=4096):
    """configs[4]: 4096 camera views x 2^20 instances (3-level hierarchy), sharded by VIEW: this rank takes its share
    of the cameras, 8 per frame; after the first frame the transforms and the order are kept (SS_FRAME_REUSE_*)."""
    sc = synth.hierarchy_scene()
    cams = synth.fibonacci_cameras(S.camera_view_proj, total_views)
    lo, hi = total_views * rank // world, total_views * (rank + 1) // world
    mine = cams[lo:hi]
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=8, device=local_rank, max_draws_per_view=1 << 16)
    ctx.load_scene(sc)
    batches = [np.ascontiguousarray(mine[b:b + 8]) for b in range(0, len(mine), 8)]
    reuse = S.FRAME_REUSE_WORLD | S.FRAME_REUSE_SORT
    for b in batches[:3]:
        ctx.frame_sync(ctx.frame(b))
    stream = torch.cuda.ExternalStream(ctx.stream, device=local_rank)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    slots = []
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for i, b in enumerate(batches):
            slots.append(ctx.frame(b, 0 if i == 0 else reuse))
        ev1.record(stream)
    ev1.synchronize()
    ms = ev0.elapsed_time(ev1)
    vis, out_bytes = [], 0
    for i in range(max(0, len(batches) - 3), len(batches)):
        st = ctx.frame_sync(slots[i])
        vis += [int(st.visible[v]) for v in range(len(batches[i]))]
    stage = ctx.frame_stage_ms(slots[-1])
    mean_vis = float(np.mean(vis)) if vis else 0.0
    # per batch of 8 views: masks pass reads 64 B rows + writes 1 B; emit reads 8 B words + rows of visible items, writes 60 B each
    n = sc.n_instances
    batch_bytes = n * (64 + 1 + 8) + 8 * mean_vis * 60 + min(n, 8 * mean_vis) * 48
    ctx.close()
    return {"views": len(mine), "ms_total": ms, "ms_per_batch": ms / max(len(batches), 1), "mean_visible_per_view": mean_vis,
            "stage_ms_last_batch": {k: round(float(v), 4) for k, v in stage.items()}, "batch_algorithmic_bytes": batch_bytes, "n": n}


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    n = a.instances

    from vulkanengine_b200 import synth

    if a.impl == "reference":
        # the reference's own CPU implementation of the path on this box's host cores; rank 0 only.  Nothing of the
        # product is imported here: the view matrices come from oracle/_ref too (the reference's GLM).
        if rank != 0:
            return
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as O
        which = "ref" if O.ref_available() else "oracle"
        cores = os.cpu_count() or 1
        sample = min(a.cpu_sample if a.cpu_sample else n, n)
        sc = synth.flat_scene(sample, first=0)
        views = synth.make_views(lambda *q: O.camera_view_proj(*q, which=which), lambda *q: O.cascade_matrices(*q, which=which), a.views)
        for _ in range(max(a.warmup, 1)):
            cpu_reference(sc, views, min(sample, 1 << 16), cores, min_seconds=0.0)
        t0 = time.perf_counter()
        vals = []
        for _ in range(a.steps):
            ips, kind, secs, ns = cpu_reference(sc, views, sample, cores, min_seconds=0.0)  # one step = one pass over the sample
            vals.append((ips, secs))
            if time.perf_counter() - t0 > 200:
                break
        ips = float(np.median([v[0] for v in vals]))
        secs = float(np.median([v[1] for v in vals]))
        # a second, smaller sample shows how the rate moves with the size (the std::map and the vectors grow with it)
        ips_small, _, _, ns_small = cpu_reference(sc, views, min(sample, 1 << 20), cores, min_seconds=0.0)
        line = {"impl": "reference", "metric": "culled+compacted instances/sec", "value": ips, "unit": "instances/s",
                "n_gpus": a.gpus, "steps": len(vals), "warmup": a.warmup, "ms_per_step": secs * 1e3,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": bench_config(n, a.views, int(sc.seed), 1),
                "trig": TRIG_NOTE,
                "cpu_baseline": {"value": ips, "unit": "instances/s", "cores": cores, "kind": kind,
                                 "sample": f"first {ns} instances x {a.views} views of the same scene, one pass per step, median; "
                                           f"the first {ns_small} instances alone run at {ips_small:.3g} instances/s",
                                 "note": "glm T*R*S + product + cull + one std::map<MeshKey> per view, range-partitioned over the "
                                         "cores; leaves out the per-entity AssetManager::GetAsset probe and shared_ptr copies of "
                                         "Scene.cpp:144 (conservative for the ratio)"},
                "e2e": {"value": ips, "unit": "instances/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line), flush=True)
        return

    import torch
    import torch.distributed as dist
    from vulkanengine_b200 import scenesubmit as S

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; libscenesubmit has no CPU path")
    torch.cuda.set_device(local_rank)
    all_cpus = os.sched_getaffinity(0)
    numa = bind_near_gpu(torch, local_rank)  # before any pinned allocation: first touch places the pages
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    sc = synth.flat_scene(n, first=rank * n)
    views = synth.make_views(S.camera_view_proj, S.cascade_matrices, a.views)
    trig = S.TRIG_DETERMINISTIC if a.trig == "deterministic" else S.TRIG_LIBM

    def make_ctx(scene, flags=0, n_views=a.views):
        c = S.SceneSubmit(scene.n_nodes, scene.n_instances, len(scene.slots), max_views=n_views, device=local_rank,
                          trig_mode=trig, max_draws_per_view=1 << 16, flags=flags)
        c.load_scene(scene)
        return c

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ctx = make_ctx(sc)

    # ---- device-resident timing: the general pipeline (transform -> sort -> compact, every frame) ----------------
    for _ in range(max(a.warmup, 3)):
        st = ctx.frame_sync(ctx.frame(views))
    launches_per_frame = int(st.kernel_launches)
    stream = torch.cuda.ExternalStream(ctx.stream, device=local_rank)
    # clocks are sampled DURING load: the sampler runs through an untimed stretch of identical frames
    # (a frame is ~2 ms, nvidia-smi samples every 100 ms) and through the timed region that follows directly
    sampler = ClockSampler(local_rank)
    sampler.start()
    t_load = time.perf_counter()
    while time.perf_counter() - t_load < 1.2:
        ctx.frame_sync(ctx.frame(views))
    barrier()
    ms_total, slots = timed_frames(torch, ctx, stream, views, a.steps)
    barrier()
    clocks = sampler.stop()
    st = ctx.frame_sync(slots[-1])
    stages = [ctx.frame_stage_ms(s_) for s_ in sorted(set(slots[-3:]))]
    stage = {k: float(np.mean([x[k] for x in stages])) for k in stages[0]}
    visible = [int(st.visible[v]) for v in range(a.views)]
    draws = [int(st.draws[v]) for v in range(a.views)]
    ms_step = max_over_ranks(torch, dist, world, ms_total) / a.steps
    value = world * n / (ms_step * 1e-3)
    masks = ctx.download_masks() if rank == 0 else None

    merge = bench_merge(torch, dist, S, ctx, views, n, rank, local_rank, world, a.views) if world > 1 else None

    # ---- end to end through the C ABI from host buffers ----------------------------------------------------------
    e2e = None
    if not a.no_e2e:
        host_trs = torch.from_numpy(sc.trs).pin_memory()
        h2d = host_trs.numel() * 4 + views.nbytes
        for _ in range(2):
            ctx.upload_trs_ptr(host_trs.data_ptr(), n)
            s_ = ctx.frame(views)
            st_ = ctx.frame_sync(s_)
            for v in range(a.views):
                ctx.download_draws(s_, v, st_)
        e_steps = max(3, min(a.steps, 30))

        def consume(slot_):
            st_k = ctx.frame_sync(slot_)
            nbytes = C.sizeof(S.FrameStats)
            for v in range(a.views):
                dr, dk = ctx.download_draws(slot_, v, st_k)
                nbytes += dr.nbytes + dk.nbytes
            return nbytes

        def run_e2e(upload):
            # Pipelined the way a host with frames in flight runs it: step k's upload and ss_frame are issued, then
            # step k-1's result is fenced and read.  Every step still moves all of its inputs.
            upload()
            consume(ctx.frame(views))  # untimed: staging buffers of this upload path exist before the clock starts
            barrier()
            t0 = time.perf_counter()
            prev, d2h_ = None, 0
            for _ in range(e_steps):
                upload()
                s_k = ctx.frame(views)
                if prev is not None:
                    d2h_ = consume(prev)
                prev = s_k
            d2h_ = consume(prev)
            torch.cuda.synchronize()
            return max_over_ranks(torch, dist, world, time.perf_counter() - t0), d2h_

        dt, d2h = run_e2e(lambda: ctx.upload_trs_ptr(host_trs.data_ptr(), n))
        e2e = {"value": world * n * e_steps / dt, "unit": "instances/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": dt / e_steps * 1e3, "host_binding": numa,
               "h2d_gbs_per_rank": h2d / (dt / e_steps) / 1e9,
               "note": "pinned TRS (36 B/instance) + views in; counts, indirect draws and draw keys of every view out; "
                       "compacted instance rows stay in HBM for the GPU consumer (the renderer); two frames in flight "
                       "(step k's result is read after step k+1 is issued); bound by the host-to-device copy"}
        # the shape of a physics tick (JoltPhysicsWorld.cpp:112-128): a tenth of the entities moved, sent as sparse records
        n_dirty = n // 10
        dirty = np.sort(np.random.default_rng(7).choice(n, n_dirty, replace=False)).astype(np.uint32)
        dirty_t = torch.from_numpy(dirty.view(np.int32)).pin_memory()
        recs_t = torch.from_numpy(np.ascontiguousarray(sc.trs[dirty])).pin_memory()
        dt_s, _ = run_e2e(lambda: ctx.upload_trs_sparse_ptr(dirty_t.data_ptr(), recs_t.data_ptr(), n_dirty))
        e2e["dirty_tenth"] = {"value": world * n * e_steps / dt_s, "ms_per_step": dt_s / e_steps * 1e3,
                              "h2d_bytes_per_step": int(n_dirty * 40 + views.nbytes),
                              "note": "the same loop when 10 % of the transforms change per step (ss_upload_trs_sparse); "
                                      "beside the full-upload number, not instead of it"}
        # the ceiling of the full-upload number: every rank copies the same pinned stream at once, nothing else running
        dev_probe = torch.empty_like(host_trs, device=f"cuda:{local_rank}")
        dev_probe.copy_(host_trs, non_blocking=True)
        torch.cuda.synchronize()
        barrier()
        t0 = time.perf_counter()
        for _ in range(3):
            dev_probe.copy_(host_trs, non_blocking=True)
        torch.cuda.synchronize()
        dt_p = max_over_ranks(torch, dist, world, time.perf_counter() - t0) / 3
        e2e["h2d_probe"] = {"gbs_per_rank": host_trs.numel() * 4 / dt_p / 1e9, "gbs_all_ranks": world * host_trs.numel() * 4 / dt_p / 1e9,
                            "ceiling_instances_per_s": world * n / dt_p,
                            "note": "plain pinned host-to-device copies of the TRS stream on all ranks at once (max over ranks): "
                                    "what the full-upload e2e number cannot exceed on this host"}
        del host_trs, dirty_t, recs_t, dev_probe
    ctx.close()

    # ---- steady state: the draw-order-resident layout (sorted when the static streams change, not per frame) -----
    steady = None
    if not a.no_extras:
        rctx = make_ctx(sc, flags=S.CONFIG_DRAW_ORDER_RESIDENT)
        st0 = rctx.frame_sync(rctx.frame(views))  # builds the layout: key words + sort + permutation of every stream
        relayout_ms = float(st0.gpu_ms)
        for _ in range(4):
            rctx.frame_sync(rctx.frame(views))
        rstream = torch.cuda.ExternalStream(rctx.stream, device=local_rank)
        barrier()
        rms, rslots = timed_frames(torch, rctx, rstream, views, a.steps)
        barrier()
        rst = rctx.frame_sync(rslots[-1])
        rstage = rctx.frame_stage_ms(rslots[-1])
        rctx.upload_trs(sc.trs)  # every transform changed: the permuted stream is rebuilt inside the next frame
        rctx.flush_uploads()
        rst2 = rctx.frame_sync(rctx.frame(views))
        rms_step = max_over_ranks(torch, dist, world, rms) / a.steps
        steady = {"ms_per_frame": rms_step, "value": world * n / (rms_step * 1e-3), "unit": "instances/s",
                  "pipeline": S.PIPELINE_NAMES[int(rst.pipeline)], "stage_ms": {k: round(float(v), 4) for k, v in rstage.items()},
                  "same_lists": [int(rst.visible[v]) for v in range(a.views)] == visible,
                  "relayout_ms": relayout_ms, "frame_after_full_trs_upload_ms": float(rst2.gpu_ms),
                  "note": "SS_CONFIG_DRAW_ORDER_RESIDENT: streams kept in (key, instance) order, a frame is two streaming "
                          "passes (no sort, no gather); relayout_ms = first frame (key words + sort + permutation), paid when "
                          "ss_set_* changes a static stream; reported beside `value`, which sorts every frame"}
        rctx.close()

    # ---- strong scaling on configs[3]: 2^26 instances over the ranks ---------------------------------------------
    strong = None
    if world > 1 and not a.no_extras:
        n64 = (1 << 26) // world
        sc64 = synth.flat_scene(n64, first=rank * n64)
        c64 = make_ctx(sc64)
        for _ in range(3):
            c64.frame_sync(c64.frame(views))
        s64 = torch.cuda.ExternalStream(c64.stream, device=local_rank)
        barrier()
        ms64, sl64 = timed_frames(torch, c64, s64, views, a.steps)
        barrier()
        c64.frame_sync(sl64[-1])
        ms64 = max_over_ranks(torch, dist, world, ms64) / a.steps
        strong = {"instances_total": n64 * world, "instances_per_rank": n64, "ms_per_frame": ms64,
                  "value": n64 * world / (ms64 * 1e-3), "unit": "instances/s",
                  "one_gpu_estimate_ms": ms_step * (1 << 26) / n,
                  "speedup_vs_estimate": ms_step * (1 << 26) / n / ms64,
                  "note": "configs[3] (64M instances, entity-wise shards, no collective in the sharded phase); the one-GPU time is "
                          "this rank count's 2^24 frame time scaled by 4 (2^26 instances do not fit the per-view capacity "
                          "budget of one context at full size), so the speed-up includes nothing but that linear estimate"}
        c64.close()
        del sc64

    # ---- configs[4]: 4096 cameras x 2^20 instances, sharded by view ----------------------------------------------
    cfg5 = None
    if not a.no_extras:
        barrier()
        r5 = bench_cfg5(torch, S, synth, local_rank, rank, world)
        ms5 = max_over_ranks(torch, dist, world, r5["ms_total"])
        peak5, _ = peaks()
        cfg5 = {"views_total": 4096, "views_per_rank": r5["views"], "ms_total": ms5, "ms_per_batch_of_8": r5["ms_per_batch"],
                "views_per_s": 4096 / (ms5 * 1e-3), "instance_views_per_s": 4096 * r5["n"] / (ms5 * 1e-3),
                "mean_visible_per_view": r5["mean_visible_per_view"], "stage_ms_last_batch": r5["stage_ms_last_batch"],
                "roofline": {"bound": "hbm", "achieved": r5["batch_algorithmic_bytes"] / (r5["ms_per_batch"] * 1e-3) / 1e9,
                             "peak": peak5, "unit": "GB/s",
                             "frac": r5["batch_algorithmic_bytes"] / (r5["ms_per_batch"] * 1e-3) / 1e9 / peak5,
                             "note": "per batch of 8 views: rows + words in, masks out, 60 B per visible (instance, view) out; "
                                     "the batch is emit/HBM-bound, not FP32-issue-bound"},
                "note": "sharded by view, no exchange; SS_FRAME_REUSE_WORLD | SS_FRAME_REUSE_SORT after the first batch"}

    # ---- the other single-GPU configs (rank 0, one GPU only): configs[1] and the hierarchy variant of configs[2] ---
    other = None
    if world == 1 and not a.no_extras:
        other = {}
        for name, scene, nv in (("cfg2_1m_hierarchy_1_camera", synth.hierarchy_scene(), 1),
                                ("cfg3_hierarchy_variant_16m", synth.hierarchy_scene(n_roots=1 << 14, seed=synth.SEED_BASE + 3), a.views)):
            vws = synth.make_views(S.camera_view_proj, S.cascade_matrices, nv)
            for label, flags in (("general", 0), ("resident", S.CONFIG_DRAW_ORDER_RESIDENT)):
                c2 = make_ctx(scene, flags=flags, n_views=nv)
                for _ in range(4):
                    c2.frame_sync(c2.frame(vws))
                s2 = torch.cuda.ExternalStream(c2.stream, device=local_rank)
                ms2, sl2 = timed_frames(torch, c2, s2, vws, a.steps)
                st2 = c2.frame_sync(sl2[-1])
                other[f"{name}_{label}"] = {"ms_per_frame": ms2 / a.steps, "instances_per_s": scene.n_instances / (ms2 / a.steps * 1e-3),
                                            "nodes": scene.n_nodes, "visible": [int(st2.visible[v]) for v in range(nv)],
                                            "stage_ms": {k: round(float(v), 4) for k, v in c2.frame_stage_ms(sl2[-1]).items()}}
                c2.close()
            del scene

    # ---- consumer of the lists: id pass of the camera view + depth passes of the cascades ----------
    raster = None
    if world == 1 and not a.no_extras and a.views == N_VIEWS:
        pos, ind, box_slots = synth.box_geometry(sc.slots)
        rsc = synth.SceneArrays(**{**sc.__dict__, "slots": box_slots})
        rctx = make_ctx(rsc)
        rctx.set_geometry(pos, ind)
        rslot = rctx.frame(views)
        rst = rctx.frame_sync(rslot)
        rstream = torch.cuda.ExternalStream(rctx.stream, device=local_rank)
        raster = {"note": "csrc/raster.cu over the frame's lists, one box (12 triangles) per instance: entity-id pass of the camera "
                          "view at 1920x1080 (SelectionPass) and DepthClamp depth passes of the four cascades at 4096^2 (ShadowPass); "
                          "bit-equal to the CPU restatement (tests/test_raster.py); the reference's rasteriser is the Vulkan device: "
                          "unavailable on this box"}
        for v, (w_, h_, fl_) in enumerate([(1920, 1080, 0)] + [(4096, 4096, S.RASTER_DEPTH_CLAMP)] * 4):
            rctx.raster_view(rslot, v, w_, h_, fl_)  # allocates the target
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 3
            with torch.cuda.stream(rstream):
                ev0.record(rstream)
                for _ in range(reps):
                    rctx.raster_view(rslot, v, w_, h_, fl_)
                ev1.record(rstream)
            ev1.synchronize()
            ms_r = ev0.elapsed_time(ev1) / reps
            n_vis = int(rst.visible[v])
            raster[f"view{v}"] = {"target": [w_, h_], "instances": n_vis, "triangles": n_vis * 12, "ms": ms_r,
                                  "mtri_per_s": n_vis * 12 / (ms_r * 1e-3) / 1e6}
        rctx.close()
        del rsc, pos, ind

    if rank == 0:
        peak, peak_src = peaks()
        # per-kernel rooflines: algorithmic bytes of the launch / its CUDA-event time (ss_frame_stage_ms)
        unique_visible = int(np.count_nonzero(masks))
        traffic = ncu_traffic(n, a.views)
        kb = instances_kernel_bytes(n, a.views, False)
        eb = emit_kernel_bytes(n, visible, unique_visible)
        roofs = []
        for name, key, nbytes, ms_k in (("k_emit (5-view compaction + row gather)", "k_emit", eb, stage["emit"]),
                                        ("k_instances_x2 (transform+cull+key)", "k_instances", kb, stage["instances"])):
            ach = nbytes / (ms_k * 1e-3) / 1e9
            roofs.append({"kernel": name, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                          "traffic": traffic.get(key), "traffic_source": traffic.get("source"), "peak_source": peak_src,
                          "algorithmic_bytes_per_launch": nbytes, "kernel_ms": ms_k})
        roofs.sort(key=lambda r: -r["kernel_ms"])  # the dominant kernel first
        balg = algorithmic_bytes(n, a.views, visible, draws, False)
        line = {"metric": "culled+compacted instances/sec", "value": value, "unit": "instances/s", "n_gpus": world,
                "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": bench_config(n, a.views, int(sc.seed), world), "trig": TRIG_NOTE,
                "pipeline": "general: transform+cull+key -> onesweep sort -> compact+gather, all per frame",
                "ms_per_frame": ms_step, "visible_per_view": visible, "draws_per_view": draws,
                "stage_ms": stage, "sort_passes_run": int(st.sort_passes_run),
                "frame_algorithmic_gb": balg / 1e9, "frame_frac_of_hbm_peak": balg / (ms_step * 1e-3) / 1e9 / peak,
                "roofline": roofs[0], "roofline_other_kernels": roofs[1:],
                "gpu_launches": launches_per_frame * a.steps, "clocks": clocks}
        if steady:
            steady["frame_frac_of_hbm_peak"] = balg / (steady["ms_per_frame"] * 1e-3) / 1e9 / peak
            line["steady_state"] = steady
        if e2e:
            line["e2e"] = e2e
        if merge:
            line["merge"] = merge
        if strong:
            line["strong_64m"] = strong
        if cfg5:
            line["cfg5"] = cfg5
        if other:
            line["other_configs"] = other
        if raster:
            line["raster"] = raster
        if not a.no_cpu_baseline and world == 1:
            os.sched_setaffinity(0, all_cpus)  # the CPU baseline gets every core again
            cores = os.cpu_count() or 1
            cs = a.cpu_sample if a.cpu_sample else 1 << 20
            ips, kind, secs, ns = cpu_reference(sc, views, cs, cores)
            ips1, _, secs1, _ = cpu_reference(sc, views, min(cs, 1 << 18), 1)
            line["cpu_baseline"] = {"value": ips, "unit": "instances/s", "cores": cores, "kind": kind,
                                    "sample": f"first {ns} instances x {a.views} views of the same scene "
                                              f"({secs * 1e3:.1f} ms best of repeats)",
                                    "single_thread_value": ips1}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
