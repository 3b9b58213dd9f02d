#!/usr/bin/env python
"""bench.py — the scene-submission hot path on B200 (BASELINE.json: culled+compacted instances/sec and
ms/frame @16M instances, 5 views).

A step = one frame of the hot path (transform -> cull -> key -> sort -> compact -> emit) over one
synthetic scene of configs[2] ("synthetic 16M instances, main camera + 4 cascaded shadow-map
frusta"), 2^24 instances per GPU.  N > 1 GPUs: entity-wise shards, 2^24 instances per rank (weak
scaling; N = 4 is configs[3]'s 64M), no data-path collective in the sharded phase.

  value      instances/s with every input stream resident in HBM (CUDA events, max over ranks)
  e2e        the same metric through the C ABI from HOST buffers: pinned TRS upload (36 B/instance) +
             view matrices in, per-view indirect draws + draw keys + counts out, every step
  roofline   the dominant kernel (transform+cull+key) against the measured HBM peak
  cpu_baseline / --impl reference
             the reference's CPU path (entt-order walk, glm T*R*S, std::map<MeshKey> grouping, plus
             the same cull) from oracle/_ref on the box's host cores, on a bounded sample
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
        rows = list(csv.reader(open(os.path.join(ROOT, "profiles", "r01_ncu_full_summary.csv"))))
        hdr = rows[0]
        ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        for r in rows[2:]:
            for key in ("k_instances", "k_emit"):
                if key in r[0] and key not in out:
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
    ctx = S.SceneSubmit(sc.n_nodes, n, len(sc.slots), max_views=a.views, device=local_rank, trig_mode=trig,
                        max_draws_per_view=1 << 16)
    ctx.load_scene(sc)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing -----------------------------------------------------------------
    for _ in range(max(a.warmup, 3)):
        st = ctx.frame_sync(ctx.frame(views))
    launches_per_frame = int(st.kernel_launches)
    stream = torch.cuda.ExternalStream(ctx.stream, device=local_rank)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # clocks are sampled DURING load: the sampler runs through an untimed stretch of identical frames
    # (a frame is ~2 ms, nvidia-smi samples every 100 ms) and through the timed region that follows directly
    sampler = ClockSampler(local_rank)
    sampler.start()
    t_load = time.perf_counter()
    while time.perf_counter() - t_load < 1.2:
        ctx.frame_sync(ctx.frame(views))
    barrier()
    slots = []
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for _ in range(a.steps):
            slots.append(ctx.frame(views))
        ev1.record(stream)
    ev1.synchronize()
    barrier()
    clocks = sampler.stop()
    ms_total = ev0.elapsed_time(ev1)
    st = ctx.frame_sync(slots[-1])
    stage = ctx.frame_stage_ms(slots[-1])
    # per-stage times averaged over the last frames_in_flight frames still in their slots
    stages = [ctx.frame_stage_ms(s) for s in sorted(set(slots[-3:]))]
    stage = {k: float(np.mean([x[k] for x in stages])) for k in stage}
    visible = [int(st.visible[v]) for v in range(a.views)]
    draws = [int(st.draws[v]) for v in range(a.views)]

    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / a.steps
    value = world * n / (ms_step * 1e-3)

    # ---- merged draw list (only where one list is needed): all-gather over NCCL, timed separately ------
    merge = None
    if world > 1:
        from vulkanengine_b200 import sharding

        class _Dev:
            def __init__(self, ptr, n, typestr):
                self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}

        def dev_tensor(ptr, n, typestr):
            if n == 0:
                return torch.empty(0, dtype=torch.int64 if typestr == "<i8" else torch.int32, device="cuda")
            return torch.as_tensor(_Dev(ptr, n, typestr), device="cuda")

        def merge_all_views(slot_, st_):
            lists = []
            for v in range(a.views):
                o = ctx.map_outputs(slot_, v)
                # the whole device buffer (capacity n): ranks gather equal-length slices, nothing is copied to pad
                idx = dev_tensor(o.instance_idx, n, "<i4")
                dk = dev_tensor(o.draw_keys, o.n_draws, "<i8")
                draws = dev_tensor(o.draws, o.n_draws * 5, "<i4").view(-1, 5)
                lists.append((idx, dk, draws[:, 1].contiguous()))
            merged = sharding.merge_view_lists(lists, rank * n, peers=peers, frame_slot=slot_)
            return sum(m[1].numel() * 4 for m in merged)  # instance indices crossing NVLink (keys are rebuilt from the runs)

        s_ = ctx.frame(views)
        st_ = ctx.frame_sync(s_)
        peers = None
        merge_all_views(s_, st_)  # warm-up (NCCL channels, allocator)
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        merge_all_views(s_, st_)
        g1.record()
        g1.synchronize()
        gt = torch.tensor([g0.elapsed_time(g1)], dtype=torch.float64, device="cuda")
        dist.all_reduce(gt, op=dist.ReduceOp.MAX)
        # the same with every rank reading its peers' lists in place over NVLink (CUDA IPC) instead of an all-gathered copy
        peers = sharding.PeerLists(ctx, local_rank)
        merge_all_views(s_, st_)
        barrier()
        m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        m0.record()
        merged_bytes = merge_all_views(s_, st_)
        m1.record()
        m1.synchronize()
        mt = torch.tensor([m0.elapsed_time(m1)], dtype=torch.float64, device="cuda")
        dist.all_reduce(mt, op=dist.ReduceOp.MAX)
        barrier()
        peers.close()
        merge = {"ms": float(mt.item()), "ms_with_all_gather": float(gt.item()), "list_bytes_over_nvlink": int(merged_bytes),
                 "note": "sizes and draws all-gathered (KBs); each rank's placement kernel reads the peers' instance lists in "
                         "place over NVLink (ms) or from an NCCL all-gathered copy (ms_with_all_gather); not part of the "
                         "sharded phase timed in value"}

    # ---- end to end through the C ABI from host buffers --------------------------------------------
    e2e = None
    if not a.no_e2e:
        host_trs = torch.from_numpy(sc.trs).pin_memory()
        h2d = host_trs.numel() * 4 + views.nbytes
        d2h = 0
        for _ in range(2):
            ctx.upload_trs_ptr(host_trs.data_ptr(), n)
            s_ = ctx.frame(views)
            st_ = ctx.frame_sync(s_)
            for v in range(a.views):
                ctx.download_draws(s_, v, st_)
        barrier()
        t0 = time.perf_counter()
        e_steps = max(3, min(a.steps, 30))
        # Pipelined the way a host with frames in flight runs it: step k's upload and ss_frame are issued, then step
        # k-1's result is fenced and read.  A whole-stream TRS upload goes to the library's second TRS buffer on its own
        # stream, so the copy for step k+1 overlaps step k's kernels; every step still moves all of its inputs.
        def consume(slot_):
            st_k = ctx.frame_sync(slot_)
            nbytes = C.sizeof(S.FrameStats)
            for v in range(a.views):
                dr, dk = ctx.download_draws(slot_, v, st_k)
                nbytes += dr.nbytes + dk.nbytes
            return nbytes
        prev = None
        for _ in range(e_steps):
            ctx.upload_trs_ptr(host_trs.data_ptr(), n)
            s_ = ctx.frame(views)
            if prev is not None:
                d2h = consume(prev)
            prev = s_
        d2h = consume(prev)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": world * n * e_steps / float(dt.item()), "unit": "instances/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": float(dt.item()) / e_steps * 1e3, "host_binding": numa,
               "note": "pinned TRS (36 B/instance) + views in; counts, indirect draws and draw keys of every view out; "
                       "compacted instance rows stay in HBM for the GPU consumer; two frames in flight (step k's result "
                       "is read after step k+1 is issued)"}

    if rank == 0:
        peak, peak_src = peaks()
        # per-kernel rooflines: algorithmic bytes of the launch / its CUDA-event time (ss_frame_stage_ms)
        masks = ctx.download_masks()
        unique_visible = int(np.count_nonzero(masks))
        traffic = ncu_traffic(n, a.views)
        kb = instances_kernel_bytes(n, a.views, False)
        eb = emit_kernel_bytes(n, visible, unique_visible)
        roofs = []
        for name, key, nbytes, ms_k in (("k_emit (5-view compaction + row gather)", "k_emit", eb, stage["emit"]),
                                        ("k_instances_x2 (transform+cull+key)", "k_instances", kb, stage["instances"])):
            ach = nbytes / (ms_k * 1e-3) / 1e9
            roofs.append({"kernel": name, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                          "traffic": traffic.get(key), "peak_source": peak_src, "algorithmic_bytes_per_launch": nbytes,
                          "kernel_ms": ms_k})
        roofs.sort(key=lambda r: -r["kernel_ms"])  # the dominant kernel first
        balg = algorithmic_bytes(n, a.views, visible, draws, False)
        line = {"metric": "culled+compacted instances/sec", "value": value, "unit": "instances/s", "n_gpus": world,
                "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "instances_per_gpu": n, "views": a.views, "keys": 4096,
                           "trig": a.trig, "seed": int(sc.seed), "l2": "inputs_exceed_l2 (2 GB of streams per frame)",
                           "sharding": "entity-wise, no collective in the sharded phase" if world > 1 else "none"},
                "ms_per_frame": ms_step, "visible_per_view": visible, "draws_per_view": draws,
                "stage_ms": stage, "sort_passes_run": int(st.sort_passes_run),
                "frame_algorithmic_gb": balg / 1e9, "frame_frac_of_hbm_peak": balg / (ms_step * 1e-3) / 1e9 / peak,
                "roofline": roofs[0], "roofline_other_kernels": roofs[1:],
                "gpu_launches": launches_per_frame * a.steps, "clocks": clocks}
        if e2e:
            line["e2e"] = e2e
        if merge:
            line["merge"] = merge
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
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
