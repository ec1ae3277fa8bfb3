#!/usr/bin/env python
"""bench.py — Mrays/s (primary + shadow + mirror scene queries) of the per-pixel hot path on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg3] [--impl ours|reference]

A step = one frame of the workload (ray-gen -> traversal -> shading -> final pixel write).  N > 1 is launched by
torchrun, one rank per GPU: the frame is split into interleaved 32-row bands (scene replicated), every rank writes
its bands straight into rank 0's framebuffer over NVLink (CUDA IPC peer mapping), i.e. strong scaling of one frame.
Prints ONE JSON line on rank 0.  See DESIGN.md §6 for how each field is measured.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

_OUT = sys.stdout
METRIC = "Mrays/s (primary+shadow+mirror)"
UNIT = "Mrays/s"


def workload(name, scenes):
    """BASELINE.json configs[1..4]; returns (scene arrays, render params, description)."""
    if name == "cfg1":       # the reference's own scene at its own resolution: launch-bound on a GPU, kept for the per-config CPU baselines
        return scenes.rt_unity(), scenes.rt_unity_params(), "RayTracing.unity as shipped, 50x50, 1 light, bounces 5 (no mirror material)"
    if name == "cfg2":
        return scenes.rt_unity(True), scenes.rt_unity_params(3840, 2160, 4), "RayTracing.unity, sphere mirror restored, 3840x2160, depth 4, 1 light"
    if name == "cfg3":
        return scenes.cfg3_scene(), scenes.cfg3_params(), "1 003 520-triangle displaced geodesic sphere + ground quad, 3840x2160, 2 point lights, depth 0, 1 spp"
    if name == "cfg4":
        return scenes.cfg4_scene(), scenes.cfg4_params(), "125 baked instances x 81 920 = 10 240 000 triangles, 7680x4320, 2 point lights"
    if name == "cfg4i":      # the same picture class as cfg4 from ONE base mesh + 125 instance matrices (rt_scene_mesh_instance)
        return scenes.cfg4_instanced(), scenes.cfg4_params(), "ONE 81 920-triangle base mesh x 125 instances (rt_scene_mesh_instance; = 10 240 000 baked triangles), 7680x4320, 2 point lights"
    if name == "cfg5":
        return scenes.cfg5_scene(), scenes.cfg5_params(), "100 000 spheres (10% mirrors) + 5 000 000-triangle sphere + ground, 3840x2160 x 16 spp, depth 4, 2 lights"
    raise SystemExit(f"unknown workload {name}")


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={gpu_index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush(); self.f.seek(0)
        sm, mx, pw, reasons = [], [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2])); pw.append(float(c[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # "under load": samples in the upper half of the power range seen
        thr = (min(pw) + max(pw)) / 2 if max(pw) > min(pw) + 20 else min(pw)
        load = [s for s, w in zip(sm, pw) if w >= thr] or sm
        return {"sm_mhz": float(np.median(load)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)), "samples": len(sm),
                "reasons": sorted(reasons)}


def csrc_hash():
    """content hash of the kernel sources: ncu-derived figures under profiles/ carry the hash they were captured on and are
    dropped from the bench line when the sources have changed since"""
    import hashlib
    h = hashlib.sha1()
    d = os.path.join(ROOT, "unity-raytracer_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh", ".h")) and f != "rt_microbench.cu":
            h.update(f.encode()); h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:12]


def bind_to_gpu_numa_node(local_rank):
    """N > 1: run this rank's host threads on the CPUs of the NUMA node its GPU hangs off, so that the pages of the shared host
    frame it touches first (its own bands, SharedFrame) are allocated there and its GPU's copy-out does not cross the socket
    interconnect.  Best effort: returns the node or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        if len(bus.split(":")[0]) == 8:
            bus = bus[4:]                                         # 00000000:1b:00.0 -> 0000:1b:00.0
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return node
    except Exception:
        pass
    return None


class SharedFrame:
    """One host frame for all ranks of the node: a POSIX shared-memory file that every rank maps and page-locks
    (rt_host_register).  Each rank's rt_render copies its own bands into it over its own GPU's PCIe link."""

    def __init__(self, rt, tag, spec, rank, world, dist, res=None, band_rows=32):
        self.rt, self.arrays, self.paths, self.rank = rt, {}, [], rank
        names = [None]
        if rank == 0:
            names[0] = f"/dev/shm/rtb200_{os.getpid()}_{tag}"
        if world > 1:
            dist.broadcast_object_list(names, src=0)
        for k, (shape, dt) in spec.items():
            path = f"{names[0]}_{k}"
            n = int(np.prod(shape)) * np.dtype(dt).itemsize
            if rank == 0:
                with open(path, "wb") as f:
                    f.truncate(n)
            self.paths.append(path)
        if world > 1:
            dist.barrier()
        for (k, (shape, dt)), path in zip(spec.items(), self.paths):
            a = np.memmap(path, dtype=dt, mode="r+", shape=shape)
            if res is not None and world > 1:
                # first touch by the rank that will fill them: the pages of a band land on the NUMA node of the GPU that copies
                # into it (bind_to_gpu_numa_node), not all on rank 0's
                X, Y = res
                rows = a.reshape((Y, X) + tuple(shape[1:]))
                for y0 in range(rank * band_rows, Y, world * band_rows):
                    rows[y0:y0 + band_rows] = 0
            elif rank == 0:
                a[...] = 0                                   # touch the pages once
            self.arrays[k] = a
        if world > 1:
            dist.barrier()
        for a in self.arrays.values():
            if rt is not None:                               # (None: the CPU test of this class, tests/test_dist_gloo.py)
                rt.host_register(a)

    def close(self, world, dist):
        for a in self.arrays.values():
            if self.rt is not None:
                self.rt.host_unregister(a)
        self.arrays = {}
        if world > 1:
            dist.barrier()
        if self.rank == 0:
            for pth in self.paths:
                try:
                    os.unlink(pth)
                except OSError:
                    pass


def block_sample(rng, res_x, res_y, n_pixels, block=16):
    """a pixel sample for the CPU arms: whole block x block pixel tiles drawn at random, each tile in scan order — spread over
    the frame like a random sample, but with the ray-to-ray coherence (cache reuse) a scan-order loop over the frame has"""
    bx, by = (res_x + block - 1) // block, (res_y + block - 1) // block
    nb = max(1, min(bx * by, int(np.ceil(n_pixels / (block * block)))))
    tiles = np.sort(rng.choice(bx * by, nb, replace=False))
    px = []
    for t in tiles:
        x0, y0 = (t % bx) * block, (t // bx) * block
        xs = np.arange(x0, min(x0 + block, res_x)); ys = np.arange(y0, min(y0 + block, res_y))
        px.append((ys[:, None] * res_x + xs[None, :]).ravel())
    return np.concatenate(px).astype(np.int64)


def measured_peak_hbm():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def microbench_peaks(device):
    """L2-resident read bandwidth (GB/s) and FP32 FMA rate (TFLOP/s) measured on this box by librtb200_microbench.so
    (unity-raytracer_b200/csrc/rt_microbench.cu): the denominators of SURVEY.md §8(d)'s T_l2 and T_issue."""
    import ctypes as C
    try:
        lib = C.CDLL(os.path.join(ROOT, "unity-raytracer_b200", "librtb200_microbench.so"))
        lib.rtmb_l2_read_gbs.restype = C.c_double; lib.rtmb_l2_read_gbs.argtypes = [C.c_int, C.c_int, C.c_int]
        lib.rtmb_fp32_tflops.restype = C.c_double; lib.rtmb_fp32_tflops.argtypes = [C.c_int, C.c_int]
        l2 = lib.rtmb_l2_read_gbs(device, 32, 200)
        fp = lib.rtmb_fp32_tflops(device, 1 << 16)
        if l2 > 0 and fp > 0:
            return l2, fp, "measured live (rt_microbench.cu: 32 MiB L2-resident ld.cg sweep; 8-chain FFMA loop)"
    except OSError:
        pass
    return 18000.0, 64.7, "fallback (profiles/r1_microbench.json, measured on this pool's B200)"


def visit_counts(workload):
    """visit counters of the instrumented build (-DRT_COUNT_VISITS, tools/visit_counts.py; deterministic for a given BVH):
    the newest profiles/r*_visit_counts_<workload>.json"""
    best = None
    for f in sorted(os.listdir(os.path.join(ROOT, "profiles"))):
        if f.endswith(f"_visit_counts_{workload}.json"):
            best = f
    if best is None:
        return None, None
    with open(os.path.join(ROOT, "profiles", best)) as fh:
        return json.load(fh), "profiles/" + best


def three_bounds(workload, builder, dom, ms_launch, launches_per_frame, t_hbm_ms, device):
    """T_hbm, T_l2 and T_issue lower bounds of the dominant kernel against its measured time.  T_issue on two
    counts: `per_ray` = the flops of the ALGORITHM, i.e. what each ray's own traversal of this BVH needs (packet = 0 run of the
    instrumented build); `packet_lanes` = the flops the packet kernel actually executes (every lane of every packet visit),
    which books the packet's redundant tests as work and is therefore the more flattering figure."""
    recs, src = visit_counts(workload)
    try:
        rec, solo = recs[f"builder={builder},packet=1"], recs[f"builder={builder},packet=0"]
    except Exception:
        return None
    cls = "any" if dom.startswith("shadow") else "closest"
    c, c0 = rec[cls], solo[cls]
    l2_gbs, fp_tf, psrc = microbench_peaks(device)
    lane_nodes = c["lane_node_visits"] + c["packet_lane_node_visits"]; lane_prims = c["lane_prim_tests"] + c["packet_lane_prim_tests"]
    flops_packet = 48 * lane_nodes + 53 * lane_prims                    # 2 boxes x 24 flops per node visit, 53 per triangle test (§8(d) model)
    flops_ray = 48 * (c0["lane_node_visits"] + c0["packet_lane_node_visits"]) + 53 * (c0["lane_prim_tests"] + c0["packet_lane_prim_tests"])
    l2_bytes = 64 * (c["lane_node_visits"] + c["packet_node_fetches"]) + 48 * (c["lane_prim_tests"] + c["packet_prim_fetches"])
    t_frame = ms_launch * launches_per_frame
    t_l2 = l2_bytes / (l2_gbs * 1e9) * 1e3
    t_issue_ray, t_issue_packet = flops_ray / (fp_tf * 1e12) * 1e3, flops_packet / (fp_tf * 1e12) * 1e3
    bounds = {"hbm_ms": t_hbm_ms, "l2_ms": t_l2, "issue_ms": t_issue_ray}
    binding = max(bounds, key=bounds.get)
    rays = c["lane_rays"] + c["packet_rays"]
    return {"per_frame_ms": bounds, "issue_ms_packet_lanes": t_issue_packet, "binding": binding.replace("_ms", ""), "measured_ms": t_frame,
            "frac_of_binding": bounds[binding] / t_frame,
            "frac_of_binding_by_count": {"per_ray": t_issue_ray / t_frame, "packet_lanes": t_issue_packet / t_frame},
            "flops": {"per_ray": flops_ray, "packet_lanes": flops_packet}, "l2_request_bytes": l2_bytes, "l2_peak_gbs": l2_gbs, "fp32_peak_tflops": fp_tf,
            "peaks": psrc,
            "per_ray": {"node_visits_private": (c0["lane_node_visits"] + c0["packet_lane_node_visits"]) / max(1, c0["lane_rays"] + c0["packet_rays"]),
                        "prim_tests_private": (c0["lane_prim_tests"] + c0["packet_lane_prim_tests"]) / max(1, c0["lane_rays"] + c0["packet_rays"]),
                        "node_visits_packet": lane_nodes / max(1, rays), "prim_tests_packet": lane_prims / max(1, rays),
                        "node_fetches": (c["lane_node_visits"] + c["packet_node_fetches"]) / max(1, rays),
                        "prim_fetches": (c["lane_prim_tests"] + c["packet_prim_fetches"]) / max(1, rays)},
            "counts": f"{src} (instrumented build, tools/visit_counts.py)",
            "note": "flops: 24 per box test, 53 per triangle test (FMA peak counts 2 flops per instruction; the slab/triangle arithmetic is "
                    "-fmad=false mul/add/min/max, 1 flop per instruction, so 50% is the ceiling)"}


def cpu_baseline(scene, params, seconds_target, threads, kind_label="port"):
    """The oracle's accelerated mode (CPU BVH over the same exact primitive tests) on a bounded random pixel sample of
    the same workload, all host threads.  Returns the cpu_baseline object."""
    from oracle import oracle as O
    ora = O.OracleScene(scene, build_bvh=True)
    npx = params["res_x"] * params["res_y"]
    rng = np.random.default_rng(12345)
    X, Y = params["res_x"], params["res_y"]
    cal = block_sample(rng, X, Y, min(npx, 20000))
    r = ora.render(params, O.ACCELERATED, threads, pixels=cal)
    rate = sum(r["stats"]) / max(r["seconds"], 1e-6)
    rays_per_px = sum(r["stats"]) / len(cal)
    n = int(min(npx, max(len(cal), seconds_target * rate / max(rays_per_px, 1e-9))))
    px = None if n >= npx else block_sample(rng, X, Y, n)             # the whole frame goes in scan order
    r = ora.render(params, O.ACCELERATED, threads, pixels=px)
    n = npx if px is None else len(px)
    rays = sum(r["stats"])
    # the reference's own algorithm (linear scan over every primitive, Scene.cs:43-122) on a small fixed pixel subset
    sub = rng.choice(npx, 256 if scene.n_prims > 100000 else min(npx, 65536), replace=False)
    rf = O.OracleScene(scene, build_bvh=False).render(params, O.FAITHFUL, threads, pixels=sub)
    faithful = {"value": sum(rf["stats"]) / max(rf["seconds"], 1e-9) / 1e6, "unit": UNIT, "cores": threads,
                "sample": f"{len(sub)} random pixels, {sum(rf['stats'])} rays, {rf['seconds']:.2f} s, linear scan as in the reference (no BVH)"}
    return {"value": rays / r["seconds"] / 1e6, "unit": UNIT, "cores": threads, "kind": kind_label, "faithful_scan": faithful,
            "sample": f"{n} pixels of the {params['res_x']}x{params['res_y']} frame ({'the whole frame in scan order' if px is None else 'random 16x16 tiles, scan order inside a tile'}; {rays} rays, {r['seconds']:.2f} s), "
                      "C++ restatement of the reference's C# (oracle/, CPU-BVH mode, std::thread); the reference itself is a "
                      "single-threaded Mono loop with a linear scan and cannot run here (no Unity/.NET)",
            "seconds": r["seconds"]}, ora


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path for this workload on the host cores.  The reference is C#/Unity and
    cannot execute here, so this is the oracle port (kind "port"), multithreaded, each step a bounded pixel sample."""
    if rank != 0:
        return
    rtb = importlib.import_module("unity-raytracer_b200")
    scene, params, desc = workload(args.workload, rtb.scenes)
    from oracle import oracle as O
    threads = os.cpu_count() or 1
    ora = O.OracleScene(scene, build_bvh=True)
    npx = params["res_x"] * params["res_y"]
    rng = np.random.default_rng(12345)
    cal = block_sample(rng, params["res_x"], params["res_y"], min(npx, 20000))
    r = ora.render(params, O.ACCELERATED, threads, pixels=cal)
    budget = min(3.0, 150.0 / max(1, args.steps + args.warmup))                # seconds of CPU work per step
    n = int(min(npx, max(2000, budget * len(cal) / max(r["seconds"], 1e-6))))
    # whole 16x16 pixel tiles in scan order (the whole frame in scan order when the budget allows it): a CPU loop over the
    # frame keeps its rays coherent, a random permutation of single pixels would throw that cache reuse away
    px = None if n >= npx else block_sample(rng, params["res_x"], params["res_y"], n)
    n = npx if px is None else len(px)
    for _ in range(args.warmup):
        ora.render(params, O.ACCELERATED, threads, pixels=px)
    rays, secs = 0, 0.0
    for _ in range(args.steps):
        r = ora.render(params, O.ACCELERATED, threads, pixels=px)
        rays += sum(r["stats"]); secs += r["seconds"]
    v = rays / secs / 1e6
    sample = (f"{n} pixels of the {params['res_x']}x{params['res_y']} frame per step ({'whole frame, scan order' if px is None else 'random 16x16 tiles, scan order inside a tile'}; "
              f"{rays // max(1, args.steps)} rays/step)")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * secs / max(1, args.steps), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": f"{args.workload}: {desc}", "rays_per_step": rays // max(1, args.steps), "sample": sample},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample + "; oracle/ C++ restatement of the reference's C# "
                             "hot path with a binned-SAH CPU BVH (leaves of <= 4), scalar code (the reference itself: single-threaded Mono linear scan, cannot run here)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), file=_OUT, flush=True)


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    rtb = importlib.import_module("unity-raytracer_b200")
    H = rtb.host
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the hot path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    run_ours.numa_node = None
    if world > 1:
        run_ours.numa_node = bind_to_gpu_numa_node(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        # waits that must not occupy a GPU (an NCCL barrier is a spinning kernel): host-side group
        run_ours.host_group = dist.new_group(backend="gloo")
    scene, params, desc = workload(args.workload, rtb.scenes)
    rt = rtb.RayTracer((local_rank,))
    rt.set_option("builder", 1 if args.builder == "sah" else 0)
    for kv in args.opt:
        k, v = kv.split("="); rt.set_option(k, float(v))
    stream = torch.cuda.Stream()
    rt.set_stream(stream.cuda_stream)
    rt.set_partition(rank, world, args.band_rows)
    t0 = time.perf_counter(); rt.set_scene(scene); commit_wall = time.perf_counter() - t0
    cstats = rt.stats()
    X, Y = params["res_x"], params["res_y"]
    want = H.RT_OUT_RGBA32F
    handles = [None]
    gather = "single GPU"
    if world > 1 and args.local_out:
        gather = "NONE (--local-out, development run)"
    elif world > 1:                      # peer-store gather: every rank writes its bands into rank 0's framebuffer
        if rank == 0:
            rt.alloc_outputs(X, Y, want); handles[0] = rt.export_outputs()
        dist.broadcast_object_list(handles, src=0)
        if rank != 0:
            rt.import_outputs(handles[0], X, Y, want)
        gather = "fused peer stores into rank 0's framebuffer over NVLink (CUDA IPC), no collective"

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(max(args.warmup, 3)):
        rt.render_device(params, want)
    barrier()
    st = rt.stats()
    rays_rank = st["rays_primary"] + st["rays_shadow"] + st["rays_mirror"]
    tot = torch.tensor([rays_rank, st["kernel_launches"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tot)
    rays_step, launches_step = int(tot[0].item()), int(tot[1].item())

    # ---- value: device-resident, per-step CUDA events on the launching stream, L2 flushed between steps ----
    sampler = ClockSampler(local_rank) if rank == 0 else None
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    for a, b in ev:
        with torch.cuda.stream(stream):
            flush.zero_()
        barrier()
        a.record(stream); rt.render_device(params, want); b.record(stream)
    barrier()
    ms = torch.tensor([a.elapsed_time(b) for a, b in ev], dtype=torch.float64, device="cuda")
    ms_by_rank = [float(ms.mean().item())]
    if world > 1:
        mine = ms.mean().reshape(1)
        allm = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allm, mine)
        ms_by_rank = [float(x.item()) for x in allm]
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.sum().item())
    clocks = sampler.stop() if sampler else None
    value = rays_step * args.steps / (ms_total * 1e-3) / 1e6

    # ---- e2e: through the C-ABI with HOST buffers, D2H of the frame inside the timed region ----
    # N = 1: rt_render into a pinned Color[].  N > 1: ONE host frame in POSIX shared memory, mapped and page-locked by every
    # rank; each rank's rt_render renders its bands and copies them into that frame over its OWN GPU's PCIe link (a frame
    # gathered on GPU 0 first would leave through one link: 2.3 ms however many GPUs rendered it).
    shared = None
    if world == 1:
        host_out = rt.alloc_host_outputs(params, want, pinned=True)
    else:
        shared = SharedFrame(rt, "e2e", {"rgba32f": ((X * Y, 4), np.float32)}, rank, world, dist, res=(X, Y), band_rows=args.band_rows)
        host_out = dict(shared.arrays)
    walls = []
    for i in range(args.steps + 2):
        barrier()
        t0 = time.perf_counter()
        rt.render(params, want, out=host_out)                           # rt_render: H2D params, kernels, D2H (own bands)
        walls.append(time.perf_counter() - t0)
    wt = torch.tensor(walls[2:], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(wt, op=dist.ReduceOp.MAX)                       # a frame is complete when the last rank has returned
    e2e_s = float(wt.sum().item())
    e2e = {"value": rays_step * args.steps / e2e_s / 1e6, "unit": UNIT, "h2d_bytes_per_step": 88 * world,
           "d2h_bytes_per_step": X * Y * 16, "ms_per_step": 1e3 * e2e_s / args.steps,
           "what": ("rt_render() into a pinned host Color[] (RGBA32F, the reference's PixelColors format); scene committed once" if world == 1 else
                    f"{world} ranks, each rt_render() into ONE shared host Color[] (POSIX shared memory, page-locked by every rank): a rank renders its "
                    "bands into its own framebuffer and copies them out over its own GPU's PCIe link while it renders; per-step wall = max over ranks; every rank runs on the CPUs of its GPU's NUMA node and first-touches its own bands of the frame; "
                    "scene committed once")}

    # ---- N > 1: both assembled frames must be the 1-GPU frame, bit for bit, in all four outputs ----
    gather_check = None
    if world > 1 and not args.local_out:
        barrier()
        allw = H.RT_OUT_ALL
        sh4 = SharedFrame(rt, "chk", {"rgba32f": ((X * Y, 4), np.float32), "rgba8": ((X * Y, 4), np.uint8),
                                      "hit_id": ((X * Y, 3), np.int32), "hit_t": ((X * Y,), np.float32)}, rank, world, dist)
        rt.render(params, allw, out=dict(sh4.arrays))                  # per-rank copy-out into the shared frame
        barrier()
        if rank == 0:
            rt.alloc_outputs(X, Y, allw); handles[0] = rt.export_outputs()
        dist.broadcast_object_list(handles, src=0)
        if rank != 0:
            rt.import_outputs(handles[0], X, Y, allw)
        rt.render_device(params, allw); barrier()                      # peer-store gather into rank 0's framebuffers
        if rank == 0:
            dev = rt.read_outputs(params, allw)
            solo = rtb.RayTracer((local_rank,))
            solo.set_option("builder", 1 if args.builder == "sah" else 0)
            solo.set_scene(scene)
            ref = solo.render(params, allw)
            solo.close()
            same = {}
            for k in ("rgba32f", "rgba8", "hit_id", "hit_t"):
                same["shared_host_frame." + k] = bool(np.array_equal(ref[k], sh4.arrays[k], equal_nan=True))
                same["gpu0_gather." + k] = bool(np.array_equal(ref[k], dev[k], equal_nan=True))
            gather_check = {"bit_identical_to_1gpu_frame": all(same.values()), "by_output": same,
                            "bytes_compared": int(2 * sum(ref[k].nbytes for k in ("rgba32f", "rgba8", "hit_id", "hit_t")))}
            if not all(same.values()):
                raise SystemExit(f"bench.py: an assembled N-GPU frame differs from the 1-GPU frame: {same}")
        barrier()
        sh4.close(world, dist)
    if shared is not None:
        shared.close(world, dist)

    # ---- roofline of the dominant kernel: per-kernel CUDA events on the launching stream over K more steps ----
    rt.set_option("kernel_timing", 1)
    ksum, kcnt = {}, {}
    for _ in range(args.steps):
        with torch.cuda.stream(stream):
            flush.zero_()
        s2 = rt.render_device(params, want, sync_stats=True)
        for k, v in s2["ms_kernel"].items():
            ksum[k] = ksum.get(k, 0.0) + v; kcnt[k] = kcnt.get(k, 0) + s2["launches_kernel"][k]
    rt.set_option("kernel_timing", 0)

    # ---- moving camera (N = 1): the tile-order hints are one frame stale, the camera orbits the scene 2 degrees per step ----
    moving = None
    if world == 1:
        moving = moving_camera_leg(torch, rt, stream, flush, scene, params, want, args.steps, rays_step, value)

    extra = {}
    if world > 1 and not args.no_extra:
        extra["cfg4"] = extra_cfg4(torch, dist, rtb, rank, world, local_rank, args)
        extra["in_process"] = extra_in_process(torch, dist, rtb, rank, world, local_rank, args, scene, params)

    if rank == 0:
        dom = max(ksum, key=ksum.get)
        n_launch = max(1, kcnt[dom])
        ms_launch = ksum[dom] / n_launch
        items = (st["rays_primary"])                                      # work items of this rank per frame
        # ALGORITHMIC bytes per launch: the scene records the kernel may touch, each once, + its per-item
        # wave-buffer / framebuffer traffic.  k_shadow_primary is the fused shadow+shade+pixel-write kernel.
        scene_bytes = 48 * cstats["n_prims"] + 64 * cstats["n_nodes"]
        fused = ksum.get("shade_primary", 0.0) == 0.0
        per_item = {"trace_primary": 8, "shade_primary": 8 + 4 + 16, "trace_queue": 32 + 8, "shade_queue": 32 + 8 + 4 + 16, "fold": 48, "finalize": 16 + 8 + 16,
                    "shadow_primary": (8 + 16) if fused else (8 + 4), "shadow_queue": 32 + 8 + 4}[dom]
        extra_b = 16 * cstats["n_prims"] if dom.startswith("shad") else 0      # pinfo (normals) read by the shading-side kernels
        launches_per_frame = max(1, kcnt[dom] // args.steps)
        alg_bytes = scene_bytes + extra_b + per_item * items / launches_per_frame
        peak, peak_src = measured_peak_hbm()
        achieved = alg_bytes / (ms_launch * 1e-3) / 1e9
        # ncu-derived facts of the same kernel on this workload (profiles/, captured once per change; not measured live): used
        # only while the kernel sources are the ones the capture ran on
        traffic, issue, traffic_note = None, None, "no ncu capture under profiles/ for this workload"
        try:
            prof_file = sorted(f for f in os.listdir(os.path.join(ROOT, "profiles")) if f.endswith("_ncu_kernels.json"))[-1]
            with open(os.path.join(ROOT, "profiles", prof_file)) as f:
                allp = json.load(f)
            prof = allp.get(args.workload, {}).get("k_" + dom)
            if prof:
                if allp.get("csrc_hash") == csrc_hash():
                    traffic = prof.get("dram_bytes_per_launch")
                    issue = {"bound": "sm issue slots", "achieved_pct_of_peak": prof.get("issue_slot_util_pct"),
                             "warp_instructions_per_launch": prof.get("warp_instructions"), "active_lanes_per_instruction": prof.get("active_lanes"),
                             "source": prof.get("source")}
                    traffic_note = f"profiles/{prof_file}, captured on csrc {allp.get('csrc_hash')} (= this build)"
                else:
                    traffic_note = f"profiles/{prof_file} was captured on csrc {allp.get('csrc_hash')}, this build is {csrc_hash()}: dropped as stale"
        except Exception:
            pass
        tb = None
        t_hbm_frame = alg_bytes * launches_per_frame / (peak * 1e9) * 1e3
        if world == 1:
            tb = three_bounds(args.workload, args.builder, dom, ms_launch, launches_per_frame, t_hbm_frame, local_rank)
        hbm = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_note,
               "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes}
        roofline = dict(hbm)
        if tb is not None and tb["binding"] == "issue":
            # the binding bound of this kernel is FP32 issue, not HBM (the scene is L2-resident): report THAT at the top level
            fl = tb["flops"]["per_ray"] / launches_per_frame
            roofline = {"bound": "issue", "achieved": fl / (ms_launch * 1e-3) / 1e12, "peak": tb["fp32_peak_tflops"], "unit": "TFLOP/s",
                        "frac": tb["frac_of_binding_by_count"]["per_ray"], "frac_packet_lanes": tb["frac_of_binding_by_count"]["packet_lanes"],
                        "traffic": traffic, "traffic_source": traffic_note, "peak_source": tb["peaks"],
                        "algorithmic_flops_per_launch": fl, "hbm": hbm}
        roofline.update({"kernel": "k_" + dom, "ms_per_launch": ms_launch,
                         "kernel_share_of_step": ksum[dom] / max(1e-9, sum(ksum.values())),
                         "ms_per_step_by_kernel": {k: v / args.steps for k, v in ksum.items() if v > 0},
                         "issue": issue, "three_bounds": tb,
                         "note": "BVH traversal is bound by SM issue slots / dependent-load latency, not by HBM (the scene is L2-resident). `bound` names the "
                                 "largest of the three lower bounds of SURVEY §8(d) (three_bounds); `achieved`/`frac` use the flops of each ray's own traversal "
                                 "(per_ray), `frac_packet_lanes` counts every lane of every packet visit; the HBM figures are under `hbm`"})
        # ---- commit path as a stage of its own: the reference rebuilds its Scene every frame ----
        steady = None
        inst_scene = any(isinstance(m, rtb.scenes.MeshInstance) for m in scene.meshes)
        if world == 1 and not args.no_commit and inst_scene:
            steady = instancing_leg(torch, rtb, rt, scene, params, want, local_rank, args)
        elif world == 1 and not args.no_commit:
            rt2 = rtb.RayTracer((local_rank,))                     # device LBVH builder = the per-frame (dynamic scene) path
            rt2.set_scene(scene)                                   # first commit allocates
            full = []
            for _ in range(3):
                t0 = time.perf_counter(); rt2.set_scene(scene); full.append(time.perf_counter() - t0)
            c2 = rt2.stats()
            re = []
            for _ in range(3):
                t0 = time.perf_counter(); rt2.update_lights(scene.lights, scene.ambient); rt2.recommit(); re.append(time.perf_counter() - t0)
            s3 = rt2.render_device(params, want, sync_stats=True); s3 = rt2.render_device(params, want, sync_stats=True)
            rf = []
            for _ in range(4):                                     # tree kept, boxes recomputed (first call allocates)
                t0 = time.perf_counter(); rt2.update_lights(scene.lights, scene.ambient); rt2.refit(); rf.append(time.perf_counter() - t0)
            rf_build = rt2.stats()["ms_commit_build"]
            steady = {"builder": "lbvh (device)", "full_commit_wall_ms": 1e3 * sorted(full)[1], "mprims_per_s": scene.n_prims / sorted(full)[1] / 1e6,
                      "ms_upload": c2["ms_commit_upload"], "ms_bvh_build": c2["ms_commit_build"], "h2d_bytes": c2["h2d_bytes"],
                      "recommit_after_update_wall_ms": 1e3 * sorted(re)[1], "refit_after_update_wall_ms": 1e3 * sorted(rf)[1],
                      "refit_device_ms": rf_build, "render_ms_with_this_bvh": s3["ms_render"],
                      "what": "steady-state re-commit of the whole scene from pageable host arrays through the C-ABI (rt_scene_begin .. "
                              "rt_scene_commit, buffers already allocated), rt_scene_update_lights + rt_scene_recommit (BVH rebuilt from "
                              "device-resident arrays) and rt_scene_update_lights + rt_scene_refit (tree kept, boxes recomputed)"}
            rt2.close()
            steady["frame_loop"] = frame_loop(rtb, scene, params, local_rank)
        cb = None
        if world == 1 and not args.no_cpu:
            cb, _ = cpu_baseline(scene.baked() if inst_scene else scene, params, 12.0, os.cpu_count() or 1)
            cb.pop("seconds", None)
        skipped = st.get("rays_shadow_skipped", 0)
        frac_skipped = skipped / max(1, rays_rank)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic",
                "value_traversed": value * (1.0 - frac_skipped),
                "config": {"workload": f"{args.workload}: {desc}", "rays_per_step": rays_step,
                           "rays_note": f"rank 0: {st['rays_primary']} primary + {st['rays_shadow']} shadow + {st['rays_mirror']} mirror queries, counted as the reference "
                                        f"issues them; {skipped} of the shadow queries (light behind the surface, outcome cannot change the pixel) "
                                        "are answered without traversal: `value_traversed` counts only the queries that walk the BVH",
                           "l2": "flushed between timed steps (256 MiB memset), per-step CUDA events",
                           "partition": f"{world} rank(s), interleaved {args.band_rows}-row bands, scene replicated", "gather": gather, "gather_check": gather_check, "output": "RGBA32F full frame on GPU 0",
                           "options": args.opt, "bvh": f"builder={args.builder} (commit is outside the timed region and reported under `commit`)",
                           "tile_order": "heavy-first from the previous frame's measured tile costs (warm-up frames establish it); static camera — see `moving_camera`"},
                "ms_per_step_by_rank": ms_by_rank, "rays_rank0": rays_rank, "clocks": clocks, "e2e": e2e, "gpu_launches": launches_step * args.steps, "roofline": roofline, "cpu_baseline": cb,
                "moving_camera": moving, "extra": extra or None, "csrc_hash": csrc_hash(),
                "commit": {"ms_upload": cstats["ms_commit_upload"], "ms_bvh_build": cstats["ms_commit_build"], "h2d_bytes": cstats["h2d_bytes"],
                           "wall_ms": 1e3 * commit_wall, "n_prims": cstats["n_prims"], "n_nodes": cstats["n_nodes"], "builder": args.builder,
                           "steady_state": steady}}
        print(json.dumps(line), file=_OUT, flush=True)
    barrier()
    rt.close()
    if world > 1:
        dist.destroy_process_group()


def moving_camera_leg(torch, rt, stream, flush, scene, params, want, steps, rays_static, value_static):
    """Same workload, but the camera orbits the scene's centre by 2 degrees per frame (always looking at it): the tile-order hints
    come from the previous camera position, the L2 is flushed as for `value`.  Ray counts change a little with the view."""
    bb = rt.scene_aabb().astype(np.float64)
    centre = 0.5 * (bb[:3] + bb[3:])
    pos0 = np.asarray(params["pos"], np.float64)
    ms, rays = [], 0

    def view(k):
        a = np.radians(2.0 * k)
        c, s_ = np.cos(a), np.sin(a)
        R = np.array(((c, 0, s_), (0, 1, 0), (-s_, 0, c)))                 # about the world y axis through the scene centre
        q = dict(params)
        q["pos"] = tuple(np.float32(centre + R @ (pos0 - centre)))
        for key in ("fwd", "right", "up"):
            v = R @ np.asarray(params[key], np.float64)
            q[key] = tuple(np.float32(v / np.linalg.norm(v)))              # RTS:178-180 normalises the basis on the managed side
        return q
    for k in range(3):
        rt.render_device(view(k - 3), want)
    for k in range(steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            flush.zero_()
        torch.cuda.synchronize()
        a.record(stream); st = rt.render_device(view(k), want, sync_stats=True); b.record(stream)
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b)); rays += st["rays_primary"] + st["rays_shadow"] + st["rays_mirror"]
    rt.render_device(params, want); rt.render_device(params, want)         # back to the static view (hints re-established)
    v = rays / (sum(ms) * 1e-3) / 1e6
    return {"value": v, "unit": UNIT, "ms_per_step": sum(ms) / len(ms), "rays_per_step": rays // max(1, steps), "vs_static_camera": v / value_static,
            "what": "camera orbits the scene centre 2 degrees per frame: tile-order hints are one frame stale; L2 flushed between frames; device-resident like `value`"}


def instancing_leg(torch, rtb, rt, scene, params, want, dev, args):
    """Workloads made of instanced meshes (rt_scene_mesh_instance): what the instancing buys and costs against the baked twin —
    the scene the reference would build by baking every instance every frame (SceneMesh.cs:21-31).  Device memory of the
    scene, wall time to re-pose EVERY instance (one rt_scene_update_instances call + rt_scene_recommit) against re-sending the baked
    meshes (rt_scene_update_mesh x N + rt_scene_recommit), frame time instanced vs baked, frames compared bit for bit."""
    S = rtb.scenes
    free0 = torch.cuda.mem_get_info()[0]
    ri = rtb.RayTracer((dev,)); ri.set_option("builder", 1 if args.builder == "sah" else 0)
    t0 = time.perf_counter(); ri.set_scene(scene); commit_i = time.perf_counter() - t0
    mem_i = free0 - torch.cuda.mem_get_info()[0]
    baked = scene.baked()
    free1 = torch.cuda.mem_get_info()[0]
    rb = rtb.RayTracer((dev,)); rb.set_option("builder", 1 if args.builder == "sah" else 0)
    t0 = time.perf_counter(); rb.set_scene(baked); commit_b = time.perf_counter() - t0
    mem_b = free1 - torch.cuda.mem_get_info()[0]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def frame_ms(r, n=5):
        out = []
        for _ in range(n + 2):
            flush.zero_(); torch.cuda.synchronize()
            out.append(r.render_device(params, want, sync_stats=True)["ms_render"])
        return sorted(out[2:])[len(out[2:]) // 2]
    ms_i, ms_b = frame_ms(ri), frame_ms(rb)
    allw = rtb.host.RT_OUT_ALL
    fi, fb = ri.render(params, allw), rb.render(params, allw)
    same = all(bool(np.array_equal(fi[k], fb[k], equal_nan=True)) for k in ("rgba32f", "rgba8", "hit_id", "hit_t"))
    del fi, fb
    inst = [(i, m) for i, m in enumerate(scene.meshes) if isinstance(m, S.MeshInstance)]
    boxes = [S.bake_tris(scene.base_meshes[m.base], m.m3x4)[2] for _, m in inst]
    rep_i, rep_b = [], []
    idx_all, m_all, b_all = np.array([i for i, _ in inst], np.int32), np.stack([m.m3x4 for _, m in inst]), np.stack(boxes)
    for _ in range(4):
        t0 = time.perf_counter()
        ri.update_instances(idx_all, m_all, b_all)                              # one rt_scene_update_instances call for all of them
        ri.recommit(); rep_i.append(time.perf_counter() - t0)
    st_i = ri.stats()
    for _ in range(2):
        t0 = time.perf_counter()
        for i, m in enumerate(baked.meshes):
            rb.update_mesh(i, m[0], m[1], m[3])
        rb.recommit(); rep_b.append(time.perf_counter() - t0)
    ri.close(); rb.close()
    del flush
    torch.cuda.empty_cache()
    if not same:
        raise SystemExit("bench.py: the instanced frame differs from the frame of its baked twin")
    return {"kind": "instancing", "instances": len(inst), "base_meshes": len(scene.base_meshes), "triangles_if_baked": int(scene.n_mesh_tris),
            "device_scene_bytes": {"instanced": int(mem_i), "baked": int(mem_b)},
            "commit_wall_s": {"instanced": commit_i, "baked": commit_b},
            "frame_ms_device": {"instanced": ms_i, "baked": ms_b, "ratio": ms_i / ms_b},
            "repose_all_wall_ms": {"instanced": 1e3 * sorted(rep_i)[1], "baked": 1e3 * sorted(rep_b)[0],
                                   "instanced_device_ms": st_i["ms_commit_upload"] + st_i["ms_commit_build"]},
            "frames_bit_identical": same,
            "what": "instanced = rt_scene_base_mesh + rt_scene_mesh_instance (object-space mesh stored once, triangles baked on the fly with the "
                    "reference's MultiplyPoint3x4 arithmetic); baked = the same matrices applied on the host and sent through rt_scene_mesh. "
                    "repose = new matrix for EVERY instance + rt_scene_recommit (instanced: TLAS rebuilt on the host, 128 B per instance uploaded; "
                    "baked: every triangle re-sent and the BVH rebuilt); L2 flushed between the timed frames"}


def extra_cfg4(torch, dist, rtb, rank, world, local_rank, args, steps=5):
    """BASELINE.json configs[3] — the config the multi-GPU split is quoted on (10.24 M triangles in 125 baked meshes, 7680x4320,
    the only config whose scene does not fit the L2) — at this N: device-resident value (peer-store gather into GPU 0) and e2e
    (every rank copies its bands into one shared host frame)."""
    H = rtb.host
    scene, params, desc = workload("cfg4", rtb.scenes)
    X, Y = params["res_x"], params["res_y"]
    want = H.RT_OUT_RGBA32F
    rt = rtb.RayTracer((local_rank,))
    rt.set_option("builder", 0)                                       # device LBVH: a 10 M-triangle host SAH build takes seconds per rank
    stream = torch.cuda.Stream(); rt.set_stream(stream.cuda_stream)
    rt.set_partition(rank, world, args.band_rows)
    t0 = time.perf_counter(); rt.set_scene(scene); commit_s = time.perf_counter() - t0
    handles = [None]
    if rank == 0:
        rt.alloc_outputs(X, Y, want); handles[0] = rt.export_outputs()
    dist.broadcast_object_list(handles, src=0)
    if rank != 0:
        rt.import_outputs(handles[0], X, Y, want)

    def barrier():
        torch.cuda.synchronize(); dist.barrier()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        rt.render_device(params, want)
    barrier()
    st = rt.stats()
    tot = torch.tensor([st["rays_primary"] + st["rays_shadow"] + st["rays_mirror"]], dtype=torch.float64, device="cuda")
    dist.all_reduce(tot)
    rays = int(tot.item())
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        with torch.cuda.stream(stream):
            flush.zero_()
        barrier()
        a.record(stream); rt.render_device(params, want); b.record(stream)
    barrier()
    ms = torch.tensor([a.elapsed_time(b) for a, b in ev], dtype=torch.float64, device="cuda")
    mine = ms.mean().reshape(1); allm = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(allm, mine)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_step = float(ms.mean().item())
    shared = SharedFrame(rt, "cfg4", {"rgba32f": ((X * Y, 4), np.float32)}, rank, world, dist, res=(X, Y), band_rows=args.band_rows)
    walls = []
    for i in range(steps + 2):
        barrier()
        t0 = time.perf_counter(); rt.render(params, want, out=dict(shared.arrays)); walls.append(time.perf_counter() - t0)
    wt = torch.tensor(walls[2:], dtype=torch.float64, device="cuda")
    dist.all_reduce(wt, op=dist.ReduceOp.MAX)
    e2e_ms = 1e3 * float(wt.mean().item())
    cs = rt.stats()
    shared.close(world, dist)
    rt.close()
    del flush
    torch.cuda.empty_cache()
    return {"workload": f"cfg4: {desc}", "value": rays / (ms_step * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_step, "steps": steps, "rays_per_step": rays,
            "ms_per_step_by_rank": [float(x.item()) for x in allm],
            "e2e": {"value": rays / (e2e_ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": e2e_ms, "d2h_bytes_per_step": X * Y * 16},
            "commit_wall_s": commit_s, "n_prims": cs["n_prims"], "n_nodes": cs["n_nodes"], "builder": "lbvh (device)",
            "note": "scaling efficiency = value(N) / (N * value(1)) from the N = 1 line of `bench.py --workload cfg4 --builder lbvh` kept under profiles/"}


def extra_in_process(torch, dist, rtb, rank, world, local_rank, args, scene, params):
    """The same frame through ONE context that owns all N devices (rt_create(ids, N)) — the form a single Unity process would
    use.  Rank 0 runs it while the other ranks wait: rt_render into page-locked host arrays (every device
    copies its bands out over its own link) and rt_render_device + rt_read_outputs (peer-store gather into device 0), all four
    outputs compared bit for bit with the 1-GPU frame."""
    H = rtb.host
    hg = run_ours.host_group
    torch.cuda.synchronize(); dist.barrier(group=hg)                   # the other ranks wait on the HOST: their GPUs stay idle
    res = None
    if rank == 0:
        allw = H.RT_OUT_ALL
        solo = rtb.RayTracer((0,)); solo.set_option("builder", 1 if args.builder == "sah" else 0); solo.set_scene(scene)
        ref = solo.render(params, allw); solo.close()
        ctx = rtb.RayTracer(tuple(range(world)))
        ctx.set_option("builder", 1 if args.builder == "sah" else 0)
        ctx.set_scene(scene)
        out = rtb.RayTracer.alloc_host_outputs(params, allw, pinned=True)
        ctx.render(params, allw, out=out)
        same = {"rt_render." + k: bool(np.array_equal(ref[k], out[k], equal_nan=True)) for k in ("rgba32f", "rgba8", "hit_id", "hit_t")}
        ctx.render_device(params, allw); dev = ctx.read_outputs(params, allw)
        same.update({"rt_render_device." + k: bool(np.array_equal(ref[k], dev[k], equal_nan=True)) for k in ("rgba32f", "rgba8", "hit_id", "hit_t")})
        col = {"rgba32f": out["rgba32f"]}
        for _ in range(3):
            ctx.render(params, H.RT_OUT_RGBA32F, out=col)
        walls = []
        for _ in range(10):
            t0 = time.perf_counter(); ctx.render(params, H.RT_OUT_RGBA32F, out=col); walls.append(time.perf_counter() - t0)
        stx = ctx.stats()
        rays = stx["rays_primary"] + stx["rays_shadow"] + stx["rays_mirror"]
        ctx.close()
        ms = 1e3 * sum(walls) / len(walls)
        res = {"devices": world, "bit_identical": all(same.values()), "by_output": same,
               "e2e": {"value": rays / (ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms, "d2h_bytes_per_step": params["res_x"] * params["res_y"] * 16},
               "what": "one rt_ctx over all N devices (rt_create(ids, N)), rt_render into a pinned host Color[]: one host thread per device queues the "
                       "launches, every device copies its own bands out"}
        if not all(same.values()):
            raise SystemExit(f"bench.py: the in-process {world}-device frame differs from the 1-GPU frame: {same}")
    torch.cuda.synchronize(); dist.barrier(group=hg)
    return res


def frame_loop(rtb, scene, params, dev, frames=30):
    """The reference's Update() is UpdateScene() then CastPixelRays() (RayTracingSetup.cs:171-199): scene re-sent from pageable
    host arrays, frame returned in host memory (RGBA32F).  Two ways to re-send: "full" = rt_scene_begin .. rt_scene_commit
    (device LBVH rebuilt), "refit" = rt_scene_update_mesh of every mesh + rt_scene_refit (host-built SAH tree kept).  Each
    (a) on one context, back to back, and (b) on two contexts of the same GPU, scene k+1 sent by a second host thread while
    frame k renders (one frame of latency for the caller)."""
    import threading
    want = rtb.host.RT_OUT_RGBA32F
    res = {"frames": frames,
           "what": "wall ms per frame of scene re-send (pageable arrays) + rt_render (host RGBA32F out); pipelined = scene k+1 sent "
                   "into a second context by a second host thread while frame k renders"}
    for mode in ("full", "refit"):
        ctx = [rtb.RayTracer((dev,)), rtb.RayTracer((dev,))]
        outs = [rtb.RayTracer.alloc_host_outputs(params, want, pinned=True) for _ in range(2)]

        def send(c):
            if mode == "full":
                c.set_scene(scene)
            else:
                for i, (v, n, _m, ab) in enumerate(scene.meshes):
                    c.update_mesh(i, v, n, ab)
                c.refit()
        for c, o in zip(ctx, outs):
            if mode == "refit":
                c.set_option("builder", 1)
            c.set_scene(scene); c.render(params, want, out=o); c.render(params, want, out=o)
        ref = outs[0]["rgba32f"].copy()
        t0 = time.perf_counter()
        for _ in range(frames):
            send(ctx[0]); ctx[0].render(params, want, out=outs[0])
        seq = (time.perf_counter() - t0) / frames
        ready = [threading.Semaphore(0), threading.Semaphore(0)]
        free = [threading.Semaphore(1), threading.Semaphore(1)]

        def sender():
            for k in range(frames):
                free[k & 1].acquire(); send(ctx[k & 1]); ready[k & 1].release()
        th = threading.Thread(target=sender)
        t0 = time.perf_counter(); th.start()
        for k in range(frames):
            ready[k & 1].acquire(); ctx[k & 1].render(params, want, out=outs[k & 1]); free[k & 1].release()
        th.join()
        pipe = (time.perf_counter() - t0) / frames
        same = bool(np.array_equal(ref, outs[0]["rgba32f"], equal_nan=True) and np.array_equal(ref, outs[1]["rgba32f"], equal_nan=True))
        for c in ctx:
            c.close()
        res[mode] = {"one_context_ms": 1e3 * seq, "two_contexts_pipelined_ms": 1e3 * pipe, "frames_identical": same}
    return res


def main():
    # stdout carries exactly ONE JSON line: libraries that print to fd 1 (NCCL's "NCCL version ..." banner) go to stderr
    global _OUT
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="cfg3")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-commit", action="store_true", help="skip the commit-path / frame-loop legs (N = 1)")
    ap.add_argument("--no-extra", action="store_true", help="N > 1: skip the cfg4 and in-process legs")
    ap.add_argument("--band-rows", type=int, default=32, help="height of the interleaved row bands of the N-rank partition (multiple of 4)")
    ap.add_argument("--opt", action="append", default=[], metavar="KEY=VALUE", help="library option passed to rt_set_option (development)")
    ap.add_argument("--local-out", action="store_true", help="N>1: every rank keeps its bands in its own framebuffer (development: "
                    "isolates the cost of the peer-store gather; the frame is then NOT assembled on GPU 0)")
    ap.add_argument("--builder", default="sah", choices=["sah", "lbvh"],
                    help="BVH builder at commit: host binned-SAH (default; static scene, best frame time) or device LBVH (fast commit)")
    args = ap.parse_args()
    rank, world, local_rank = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
