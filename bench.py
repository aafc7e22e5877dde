#!/usr/bin/env python3
"""bench.py — the headline benchmark of BASELINE.json: Mrays/s (and Msamples/s),
device-timed, on the Cornell box at 3840x2160 x 1024 spp (config C4 of SURVEY.md
§8d), on N B200s of one node, with the reference's CPU path timed beside it.

  python bench.py --gpus 1 --steps K --warmup W            # our arm, one GPU
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...   # one rank per GPU
  python bench.py --impl reference ...                     # the reference's own CPU code

A "step" is one full render pass of the hot path (primary FP64 pass, radiance
megakernel, accumulate) over the rank's slice of the sample budget, followed at
N>1 by the single NCCL sum-reduce of the partial framebuffers to rank 0.
Total work is fixed as N grows (the `samples` loop is sliced): scaling = strong.
Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# Workloads.  c4 is the configuration the metric is quoted on.
WORKLOADS = {
    "c4": dict(scene="cornellBox", width=3840, height=2160, samples=64, superSamples=4,
               name="cornellBox 3840x2160 x 1024 spp (64 samples x 4x4 sub-pixels)"),
    "c2": dict(scene="cornellBox", width=960, height=504, samples=25, superSamples=4,
               name="cornellBox 960x504 x 400 spp (file defaults)"),
    "c3": dict(scene="simple2", width=1920, height=1080, samples=256, superSamples=1,
               name="simpleSetting2 1920x1080 x 256 spp"),
    "c1": dict(scene="simple1", width=960, height=504, samples=100, superSamples=1,
               name="simpleSetting1 960x504 x 100 spp (file defaults)"),
    "c5": dict(scene="synth10k", file="synth10k.json", width=1920, height=1080, samples=256, superSamples=1,
               name="synthetic 10k spheres 1920x1080 x 256 spp (closest hit through the hierarchy)"),
    # PlaneObject extension: the Cornell box with squares instead of r = 1e4 wall spheres
    "c2p": dict(scene="cornellBoxPlanes", file="cornellBoxPlanes.json", width=960, height=504, samples=25, superSamples=4,
                name="planar cornellBox (3 spheres + 6 planes, extension) 960x504 x 400 spp"),
    "c4p": dict(scene="cornellBoxPlanes", file="cornellBoxPlanes.json", width=3840, height=2160, samples=64, superSamples=4,
                name="planar cornellBox (3 spheres + 6 planes, extension) 3840x2160 x 1024 spp"),
    "tiny": dict(scene="cornellBox", width=192, height=108, samples=8, superSamples=2,
                 name="cornellBox 192x108 x 32 spp (plumbing check only)"),
}


def algorithmic_flops(n_spheres, tests, rays, samples, n_sub):
    """SURVEY.md §8(d) counting convention, taken from the reference source:
    17/test + 21/ray + 91/continued bounce + 18/sample + 37/sub-pixel."""
    return 17.0 * tests + 21.0 * rays + 91.0 * (rays - samples) + 18.0 * samples + 37.0 * n_sub


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([v.strip() for v in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except Exception:
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def load_workload(name: str):
    import raytracingmin_b200 as rtm
    w = WORKLOADS[name]
    sc = rtm.load_scene(os.path.join(ROOT, "scenes", w.get("file", w["scene"] + ".head.json")), "abc")
    return sc.with_dims(w["width"], w["height"], w["samples"], w["superSamples"]), w


# --------------------------------------------------------------------------------------
# CPU legs: the reference's own object code (oracle/_ref/libref_abc.so, built from
# /root/reference/src with the 8-line L-ABC patch) or, if that was not built, the C
# restatement (oracle/librtm_oracle.so).  Test infrastructure used ONLY as a timed baseline.
def cpu_render_sample(sc, wl, target_seconds: float):
    """Times the CPU path on a bounded sample of the workload: the same scene and ss,
    1/8 x 1/8 of the image area (every 8th pixel's worth of solid angle is NOT sampled:
    the image is simply rendered at lower resolution, same camera) and as many `samples`
    as fit in ~target_seconds.  Returns dict(value Mrays/s, msamples, cores, kind, sample)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    # the thread count is passed explicitly: torchrun exports OMP_NUM_THREADS=1 to its workers, which would
    # silently time a single-threaded run while the line says "all host cores"
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    w, h = max(16, sc.width // 8), max(9, sc.height // 8)
    small = sc.with_dims(w, h, 1, sc.superSamples)
    use_ref = O.ref_available("abc") and sc.n_planes == 0   # the reference's objects cannot hold a plane

    def run(S):
        s2 = small.with_dims(samples=S)
        t0 = time.perf_counter()
        if use_ref:
            rs = O.RefScene("abc", s2)
            _, rays, samples = rs.render(1234, cores)
            rs.close()
        else:
            _, rays, samples = O.oracle_render(s2, O.LEVEL_ABC, 1234, cores)
        return time.perf_counter() - t0, rays, samples

    S, t = 1, 0.0
    while True:                      # grow until the run is long enough to extrapolate from
        t, rays, samples = run(S)
        if t >= min(1.0, target_seconds) or S >= 4096:
            break
        S *= 4
    if t < 0.7 * target_seconds and S < 4096:
        S = int(max(1, min(4096, round(S * target_seconds / t))))
        t, rays, samples = run(S)
    return {"value": rays / t / 1e6, "unit": "Mrays/s", "msamples_per_s": samples / t / 1e6, "cores": cores,
            "kind": "reference" if use_ref else "port",
            "sample": f"{wl['scene']} {w}x{h} ss={sc.superSamples} samples={S} ({samples / 1e6:.1f} Msamples, "
                      f"{rays / 1e6:.1f} Mrays, {t:.1f} s; OpenMP over rows, private mt19937 per row)",
            "seconds": t}


def cpu_cli_as_is(sc, wl, budget_s: float = 25.0):
    """The reference's CLI AS IS (oracle/_ref/ref_abc_cli: its own main.cpp + Render(), src/Renderer.cpp:200-258 — one
    shared, racing mt19937, an OpenMP fork/join per image row, stb file writes), run on a copy of the workload's scene
    file scaled down until it fits `budget_s` (the stock loop is ~16x slower than the fair harness).  Wall time of the
    whole process, as a user of the reference sees it.  None when the binary was not built (no /root/reference)."""
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_abc_cli")
    if not os.path.exists(exe) or sc.n_planes or sc.n > 64:
        return None
    import tempfile
    src = json.load(open(os.path.join(ROOT, "scenes", wl.get("file", wl["scene"] + ".head.json"))))
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    # ~1.5 Msamples/s (Cornell) .. 11 Msamples/s (simpleSetting1) on 16 cores in round 1: size the run for the slow end
    target = 1.2e6 * budget_s
    w, h, S, ss = sc.width, sc.height, sc.samples, sc.superSamples
    while w * h * S * ss * ss > target and (w > 60 or S > 1):
        if w > 60:
            w, h = w // 2, h // 2
        else:
            S = max(1, S // 2)
    src["00 width"], src["00 height"], src["00 samples"], src["00 superSamples"] = w, h, S, ss
    with tempfile.TemporaryDirectory() as tmp:
        jf = os.path.join(tmp, "scene.json")
        json.dump(src, open(jf, "w"))
        env = dict(os.environ, OMP_NUM_THREADS=str(cores))
        t0 = time.perf_counter()
        rc = subprocess.run([exe, "-json", jf], cwd=tmp, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL,
                            timeout=max(120.0, 8 * budget_s)).returncode
        dt = time.perf_counter() - t0
        ok = rc == 0 and os.path.exists(os.path.join(tmp, "result.bmp"))
    samples = w * h * S * ss * ss
    return {"msamples_per_s": samples / dt / 1e6, "wall_s": dt, "cores": cores, "ok": ok,
            "sample": f"{wl['scene']} {w}x{h} ss={ss} samples={S}: ref_abc_cli -json <file> (stock Render(): shared racing mt19937, "
                      f"per-row fork/join, result.jpg + result.bmp written), whole-process wall time"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    sc, wl = load_workload(args.workload)
    per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_render_sample(sc, wl, per_step / 4)
    vals, last = [], None
    t_all = 0.0
    for _ in range(args.steps):
        last = cpu_render_sample(sc, wl, per_step)
        vals.append(last["value"])
        t_all += last["seconds"]
    v = float(np.mean(vals))
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_all / max(1, args.steps),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl["name"], "note": "reference CPU code on the host cores; each step is a bounded sample"},
        "msamples_per_s": last["msamples_per_s"],
        "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": last["cores"], "kind": last["kind"], "sample": last["sample"]},
        "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# --------------------------------------------------------------------------------------
def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    import raytracingmin_b200 as rtm
    from raytracingmin_b200 import api

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N (one rank per GPU)")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    sc, wl = load_workload(args.workload)
    from raytracingmin_b200.partition import plan_partition, reduce_partials
    plan = plan_partition(sc.samples, sc.superSamples, rank, world)
    n_pix = sc.width * sc.height
    n_sub = n_pix * (plan["sub_end"] - plan["sub_begin"])      # sub-pixels THIS rank renders

    # device-resident handle: the library writes its partial framebuffer straight into a
    # torch tensor (the NCCL send buffer); no host sync between the last kernel and the reduce.
    knobs = {}
    for kv in args.knob:       # rtm_options.debug_* A/B knobs, e.g. --knob debug_bvh_walk=2
        k, v = kv.split("=", 1)
        knobs[k] = float(v) if k == "debug_bvh_ctrav" else int(v)
    r = rtm.Renderer(sc, device=local, seed=0x5EED, variant=args.variant, estimator=args.estimator, **plan, **knobs)
    accum = torch.zeros((sc.height, sc.width, 4), dtype=torch.float32, device=dev)
    r.set_accum_buffer(accum.data_ptr())
    stream = torch.cuda.current_stream()
    frame_bytes = n_pix * 16
    peer = world > 1 and args.reduce == "peer"
    if peer:
        # peer-memory gather (include/rtm_b200.h): rank 0 exports 2 x world framebuffer slots (double-buffered per
        # step), every rank's accumulate kernel stores its partial framebuffer straight into its slot in rank 0's HBM
        # over NVLink, a one-element all_reduce orders the streams, rank 0 sums the slots with one kernel.
        objs = [r.peer_export(2 * world) if rank == 0 else None]
        dist.broadcast_object_list(objs, src=0)
        base = objs[0][1] if rank == 0 else r.peer_open(objs[0][0])
        flag = torch.zeros(1, dtype=torch.float32, device=dev)
        step_no = [0]

    def step_nccl():
        r.render_async(stream.cuda_stream)
        reduce_partials(accum, dst=0)

    def step_peer():
        half = (step_no[0] & 1) * world
        step_no[0] += 1
        r.set_accum_buffer(base + (half + rank) * frame_bytes)
        r.render_async(stream.cuda_stream)
        dist.all_reduce(flag)          # stream-ordered: returns on a rank after every rank's stores were issued
        if rank == 0:
            r.sum_partials(base + half * frame_bytes, world, accum.data_ptr(), stream.cuda_stream)

    step = step_peer if peer else step_nccl

    def fence():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    fence()

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kms = np.zeros(3)
    rays = samples = tests = traced = 0
    fence()
    e0.record(stream)
    # per-step kernel events / counters are read after the whole timed region (no sync inside it),
    # except that counters are per pass: read them once per step only when steps are long.
    for i in range(args.steps):
        step()
    e1.record(stream)
    fence()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    # counters + per-kernel events of the LAST pass (every pass does identical work: the
    # image is a pure function of (scene, seed)).
    st = r.read_stats()
    kms = np.array(r.read_kernel_ms())
    rays, samples, tests, traced = st.rays, st.samples, st.sphere_tests, st.rays_traced

    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    cnt = torch.tensor([rays, samples, tests, traced], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    ms_total = float(t.item())
    rays_all, samples_all, tests_all, traced_all = (float(v) for v in cnt.tolist())
    ms_step = ms_total / args.steps

    # ---- e2e: the drop-in call with HOST buffers.  Scene arrays H2D (constant upload) and
    # the framebuffer D2H are inside the timed region; at N>1 rank 0 reads the reduced image.
    e2e_steps = max(1, min(args.steps, 3))
    if world == 1:
        # the drop-in call itself: renderer_render(handle, float rgb_out[H][W][3] in HOST memory, ...)
        host_rgb = torch.zeros((sc.height, sc.width, 3), dtype=torch.float32, pin_memory=True).numpy()   # pinned host memory
        r.render(image_out=host_rgb)
        fence()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            r.render(image_out=host_rgb)
        e2e_s = (time.perf_counter() - t0) / e2e_steps
        d2h = n_pix * 12
    else:
        # N > 1: every rank renders its slice (renderer_render_async), one NCCL sum-reduce into rank 0's accumulation buffer,
        # then rank 0 reads the image back THROUGH THE C ABI (renderer_read_rgb: repack kernel + one 12 B/pixel copy into a
        # host buffer, on the stream the reduce ran on) — the same bytes the N = 1 leg's renderer_render delivers
        host_rgb = torch.zeros((sc.height, sc.width, 3), dtype=torch.float32, pin_memory=True).numpy()   # pinned host memory
        if rank == 0:
            r.read_rgb(host_rgb, stream.cuda_stream)
        fence()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            step()
            if rank == 0:
                r.read_rgb(host_rgb, stream.cuda_stream)
            fence()
        e2e_s = (time.perf_counter() - t0) / e2e_steps
        d2h = n_pix * 12
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te.item())
    h2d = rtm_scene_bytes(sc)

    # untimed: the N-rank image against the SAME frame rendered by one handle on rank 0 (full sample / sub-pixel range).
    # The RNG is keyed by global indices and the divisor is global, so the two may differ by FP32 summation order only.
    image_check = None
    if world > 1:
        step()
        fence()
        if rank == 0:
            got = accum[..., :3].clone()
            with rtm.Renderer(sc, device=local, seed=0x5EED, variant=args.variant, estimator=args.estimator, **knobs) as r1:
                one = torch.from_numpy(r1.render().image).to(dev)
            diff = (got - one).abs()
            image_check = {"max_abs_diff": float(diff.max().item()), "max_rel_diff": float((diff / one.abs().clamp_min(1e-3)).max().item()),
                           "image_max": float(one.max().item()), "what": f"{world}-rank reduced framebuffer vs one handle rendering the whole frame on rank 0"}
            del got, one, diff
        fence()

    reduce_check = None
    if peer:
        # untimed: the same pass through the NCCL sum-reduce must give the same image (up to summation order)
        got = accum.clone()
        r.set_accum_buffer(accum.data_ptr())
        step_nccl()
        fence()
        if rank == 0:
            reduce_check = float((got - accum).abs().max().item())
        accum.copy_(got)

    line = None
    if rank == 0:
        fp32_peak, implied_mhz = api.fp32_peak_tflops(local)
        # dominant kernel = the radiance megakernel; its algorithmic flops per launch
        bounce_rays = (rays - samples)                      # this rank's continued bounces
        # tests of the bounce rays only (17 each), 21/ray finalisation + 91/bounce for each traced
        # ray, 18/sample accumulate.  The primary pass's 37/sub-pixel + N tests + 21 are in its own kernel.
        # Hierarchy workloads (n > 64): tests ACTUALLY executed (kernel counters), never N*rays; the counter
        # then includes the primary pass's own tests (< 1 % of the total at >= 64 spp), left in.
        brute = sc.n <= 64
        f_kernel = (17.0 * (st.sphere_tests - (n_sub * sc.n if brute else 0)) + 21.0 * bounce_rays + 91.0 * bounce_rays
                    + 18.0 * samples)
        f_total = algorithmic_flops(sc.n, tests_all, rays_all, samples_all, sc.n_sub)
        achieved = f_kernel / (kms[1] * 1e-3) / 1e12
        fb_gbs = n_pix * 16 / (kms[2] * 1e-3) / 1e9 if kms[2] > 0 else None
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        line = {
            "metric": "Mrays/s", "value": rays_all / (ms_step * 1e-3) / 1e6, "unit": "Mrays/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": wl["name"], "scene": wl["scene"], "width": sc.width, "height": sc.height,
                       "samples": sc.samples, "superSamples": sc.superSamples, "spp": sc.spp, "n_spheres": sc.n,
                       **({"n_planes": sc.n_planes, "flops_note": "plane tests are not counted (the 17-flop figure is the sphere test's)"}
                          if sc.n_planes else {}),
                       "variant": args.variant,
                       **({"knobs": knobs} if knobs else {}),
                       **({"estimator": "nee (extension, DESIGN.md §11): `value` counts the PATH rays of the reference's estimator; "
                                        f"the kernel also traced {(st.rays_traced - n_sub - (st.rays - st.samples)) / max(1, st.samples):.2f} "
                                        "shadow rays per sample; the roofline's flops count their sphere tests, nothing else of them"}
                          if args.estimator == "nee" else {}),
                       "closest_hit": "brute force (packed pairs)" if sc.n <= 64 else
                                      f"hierarchy: {st.sphere_tests / max(1, st.rays_traced):.1f} sphere tests + "
                                      f"{st.bvh_nodes / max(1, st.rays_traced):.1f} nodes per traced ray",
                       "partition": (("sub-pixel grid" if plan["sub_end"] - plan["sub_begin"] < sc.superSamples ** 2 else "samples loop")
                                     + f" sliced over {world} rank(s)"
                                     + ((", partial framebuffers stored into rank 0's HBM over NVLink by the accumulate kernel "
                                         "+ one sum kernel (peer-memory gather)") if peer else
                                        (", one NCCL sum-reduce to rank 0" if world > 1 else ""))),
                       "l2": f"inputs larger than L2: the {n_sub * 16 / 1e9:.2f} GB sub-pixel record buffer is rewritten and re-read every step"
                             if n_sub * 16 > 126e6 else "working set below L2 (plumbing workload)",
                       "semantics": "L-ABC (SURVEY.md §0.1), FP64 bit-exact primary pass + FP32 bounces, counter-based pcg3d hash per (sub-pixel, sample, depth)"},
            "msamples_per_s": samples_all / (ms_step * 1e-3) / 1e6,
            "rays_per_sample": rays_all / samples_all,
            # `value` counts the reference's rays (sum of path lengths: one PathTracing call each, src/Renderer.cpp:57), primary
            # rays included; the device traces the primary ray once per sub-pixel, not once per sample:
            "rays_traced_per_s_M": traced_all / (ms_step * 1e-3) / 1e6,
            "rays_traced_note": "closest-hit queries actually executed per second (primary query shared by a sub-pixel's samples)",
            "e2e": {"value": rays_all / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": e2e_s * 1e3, "steps": e2e_steps},
            "gpu_launches": (3 + (1 if peer else 0)) * args.steps,   # primary_f64 + radiance + accumulate per step (+ the sum kernel of --reduce peer)
            "kernel_ms": {"primary_f64": kms[0], "radiance_mega": kms[1], "accumulate": kms[2]},
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": achieved / fp32_peak if fp32_peak else None,
                         "traffic": ncu_traffic_bytes(args.workload, args.variant) if (world == 1 and args.estimator == "reference") else None,
                         "traffic_note": "bytes per launch, dram read+write from the committed ncu --set full capture (profiles/); "
                                         "the design's record round trip (16 B read + 16 B written per sub-pixel) explains it; the "
                                         "ALGORITHMIC bytes of SURVEY.md 8(d) are one framebuffer write (16 B per pixel)",
                         "record_roundtrip_bytes_per_launch": 32.0 * n_sub,
                         "algorithmic_bytes_per_launch": 16.0 * n_pix,
                         "kernel": ("radiance_mega_kernel" if sc.n <= 64 else "radiance_bvh_kernel") if args.variant == "mega"
                                   else "wf_shade_kernel + wf_extend_*_kernel (all waves)",
                         "peak_source": f"FP32 FMA-chain microbenchmark run live on this GPU (implied {implied_mhz:.0f} MHz); "
                                        "MEASURED_PEAKS.json holds no FP32 figure",
                         "flops_per_launch": f_kernel, "whole_pass_tflops": f_total / (ms_step * 1e-3) / 1e12,
                         "framebuffer_store_GBs": fb_gbs, "hbm_peak_GBs": peaks.get("hbm_gbs")},
            "clocks": clocks,
        }
        if ms_step < 2.0:
            line["note"] = ("frame of %.2f ms: launch-bound (3 kernel launches + counters memset per pass), the device rate is not "
                            "sustained work and the end-to-end rate is dominated by the framebuffer copy" % ms_step)
        if image_check is not None:
            line["image_check"] = image_check
        if reduce_check is not None:
            line["reduce_check_max_abs_diff_vs_nccl"] = reduce_check
    if world > 1:
        dist.barrier()
    return line, (r, sc, wl)


def ncu_traffic_bytes(workload: str, variant: str):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, per launch, from the committed
    `ncu --set full` capture of this very command (profiles/r02, else r01); None when no capture exists for it."""
    names = {("c4", "mega"): ["r02/ncu_mega_r02_c4_summary.txt", "r02/ncu_mega_r02a_c4_summary.txt", "r01/ncu_megav6_c4_summary.txt"],
             ("c3", "mega"): ["r01/ncu_megav5_rt_c3_summary.txt"], ("c2p", "mega"): ["r01/ncu_planes_v2_c2p_summary.txt"],
             ("c5", "mega"): ["r02/ncu_bvh_r02_c5_summary.txt", "r02/ncu_bvh_r02a_c5_summary.txt", "r01/ncu_bvh_v4_c5_summary.txt"]}.get((workload, variant), [])
    unit = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
    for name in names:
        total, seen = 0.0, 0
        try:
            for ln in open(os.path.join(ROOT, "profiles", name)):
                if ln.startswith(("dram__bytes_read.sum", "dram__bytes_write.sum")):
                    u = ln[ln.index("[") + 1:ln.index("]")]
                    total += float(ln.split("=")[1]) * unit[u]
                    seen += 1
        except Exception:
            continue
        if seen == 2:
            return total
    return None


def rtm_scene_bytes(sc) -> int:
    """bytes of scene data renderer_render uploads per call: the handle's FP32 per-sphere block (RtmSceneF32Host) plus the
    kernel parameters (RtmHotScene + RtmPassParams) of the three launches."""
    return 19248 + 2 * (1312 + 248) + 248   # sizeof(RtmSceneF32Host) + params of primary (RtmPassParams), radiance (hot + params), accumulate


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--reduce", default="nccl", choices=["peer", "nccl"],
                    help="N > 1: how the partial framebuffers reach rank 0 (peer-memory gather kernel path, or ncclReduce)")
    ap.add_argument("--estimator", default="reference", choices=["reference", "nee"],
                    help="nee = next-event estimation (extension); the headline metric is quoted on `reference`")
    ap.add_argument("--knob", action="append", default=[], metavar="debug_x=v",
                    help="rtm_options.debug_* knob for A/B runs (repeatable); the default line uses none")
    ap.add_argument("--variant", default="mega", choices=["mega", "wavefront"],
                    help="kernel organisation of the FP32 bounce loop (DESIGN.md: megakernel vs wavefront)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    args.warmup = max(args.warmup, 3)
    line, (r, sc, wl) = run_gpu_arm(args)
    if line is not None:
        if int(os.environ.get("WORLD_SIZE", "1")) == 1 and not args.no_cpu_baseline:
            cb = cpu_render_sample(sc, wl, 12.0)
            cb.pop("seconds")
            cli = cpu_cli_as_is(sc, wl)
            if cli is not None:
                cb["cli_as_is"] = cli     # BOTH CPU figures — the stock CLI and the fair harness — in one record
            line["cpu_baseline"] = cb
        print(json.dumps(line), flush=True)
    r.close()
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
