#!/usr/bin/env python
"""bench.py — throughput of the render hot path (BASELINE.json metric: Msamples/s, with Mrays/s beside it).

  python bench.py --gpus N --steps K --warmup W            our CUDA path (one process per GPU under torchrun)
  python bench.py --impl reference --gpus N --steps K ...   the CPU restatement of the reference on the host cores

A "step" renders one full frame of the workload: every rank renders `spp` samples of every pixel of the full frame
(rank g takes the sample indices s = g mod N: weak scaling, N*spp samples per pixel in total), the accumulation
buffers are summed onto rank 0 with one NCCL reduce and rank 0 tonemaps.  value = W*H*spp*N / time, time measured
with CUDA events on the launching stream, max over ranks.

Workload (BASELINE.json configs[2]): synthetic:console seed 3 (PS5-class stand-in, ~1.0 M triangles,
one 1024^2 texture, one point light), 1920x1080, 128 spp per GPU, 5 bounces, COOK_TORRANCE, FILMIC.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (recipe, n_tris, scene seed, width, height, spp per GPU, bounces, tonemap)
    "console_1080p_128spp_5b": ("console", 1000000, 3, 1920, 1080, 128, 5, "FILMIC"),
    "foliage_1080p_256spp_4b": ("foliage", 5000000, 4, 1920, 1080, 256, 4, "FILMIC"),
    "city_4k_1024spp_8b": ("city", 20000000, 5, 3840, 2160, 1024, 8, "ACES"),
    "console_small": ("console", 100000, 3, 480, 270, 8, 5, "FILMIC"),  # quick functional check
}
DEFAULT_WORKLOAD = "console_1080p_128spp_5b"
CANON_FILE = os.path.join(ROOT, "profiles", "canonical_counts.json")
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "ncu_traffic.json")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=None, help="override samples per pixel per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop, self.th = index, [], threading.Event(), None

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([x.strip() for x in out.strip().split(",")])
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=10)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 6:
                continue
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except ValueError:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def canonical_counts(workload):
    """Frozen canonical per-ray work of the workload: profiles/canonical_counts.json, written by
    tools/canonical_counts.py from the oracle's canonical BVH2."""
    try:
        with open(CANON_FILE) as f:
            return json.load(f)[workload]
    except Exception:
        return None


def ray_bytes(n_node, n_tri):
    return 64.0 * n_node + 48.0 * n_tri + 48.0  # B_ray of SURVEY.md 8(d)


# ----------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle (C++ restatement of the reference renderer; the Rust binary cannot be built here) on all host
# cores.  This is the ONLY place bench.py touches oracle/: as the timed baseline, never as the product.
def cpu_render_sample(pkg, wl, spp_sample, seed=0, scene_arrays=None):
    """Times `spp_sample` samples per pixel of the workload's full frame with the oracle on every host core.
    Returns (Msamples/s, Mrays/s, cores, seconds, build seconds)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    recipe, n_tris, sseed, w, h, spp, bounces, tm = wl
    if scene_arrays is None:
        scene_arrays = pkg.Scene.synthetic(recipe, n_tris, sseed, device=-1).arrays()  # host-only handle: no GPU work
    t0 = time.time()
    osc = oracle_lib.OracleScene(scene_arrays)
    build_s = time.time() - t0
    cores = os.cpu_count() or 1
    prof = oracle_lib.make_profile(w, h, spp, bounces, pkg.abi.TONEMAPS[tm])
    _, ms, cnt = osc.render_accum(prof, rng_mode=1, seed=seed, n_samples=spp_sample, threads=cores, return_counters=True)
    samples = w * h * spp_sample
    return samples / ms / 1e3, float(cnt[0] + cnt[1]) / ms / 1e3, cores, ms / 1e3, build_s, osc


def run_reference(args, wl):
    """--impl reference: rank 0 times the CPU restatement on a bounded sample per step; other ranks exit."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import __graft_entry__
    pkg = __graft_entry__.load_package()
    recipe, n_tris, sseed, w, h, spp, bounces, tm = wl
    sa = pkg.Scene.synthetic(recipe, n_tris, sseed, device=-1).arrays()
    spp_step = 1  # one sample per pixel of the full frame per step (a bounded sample of the 128-spp frame)
    vals, rays, secs = [], [], []
    osc_keep = None
    for i in range(args.warmup + args.steps):
        if i == 0:
            ms_v, mr_v, cores, s, build_s, osc_keep = cpu_render_sample(pkg, wl, spp_step, scene_arrays=sa)
        else:
            import oracle_lib
            prof = oracle_lib.make_profile(w, h, spp, bounces, pkg.abi.TONEMAPS[tm])
            _, ms, cnt = osc_keep.render_accum(prof, rng_mode=1, sample_begin=i, n_samples=spp_step, threads=cores,
                                               return_counters=True)
            ms_v, mr_v, s = w * h * spp_step / ms / 1e3, float(cnt[0] + cnt[1]) / ms / 1e3, ms / 1e3
        if i >= args.warmup:
            vals.append(ms_v)
            rays.append(mr_v)
            secs.append(s)
    total_s = sum(secs)
    value = w * h * spp_step * len(secs) / total_s / 1e6
    sample = "%d spp of %d per step, full %dx%d frame, %d bounces, all %d host threads" % (spp_step, spp, w, h, bounces, cores)
    line = {
        "impl": "reference", "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_s / len(secs), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "mrays_per_s": sum(rays) / len(rays),
        "config": {"workload": args.workload, "scene": "synthetic:%s seed %d, %d triangles" % (recipe, sseed, len(sa.tri_pos)),
                   "resolution": [w, h], "spp_per_gpu": spp, "bounces": bounces, "tonemap": tm, "brdf": "COOK_TORRANCE",
                   "note": "CPU restatement of the reference (oracle/, C++, all-hits ray_cast semantics); the Rust binary "
                           "cannot be built in this image (no cargo/rustc, crates un-vendored)"},
        "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------------
def run_ours(args, wl):
    import torch
    import torch.distributed as dist
    import __graft_entry__
    pkg = __graft_entry__.load_package()
    from path_tracer_b200 import multi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    recipe, n_tris, sseed, w, h, spp, bounces, tm = wl
    t0 = time.time()
    scene = pkg.Scene.synthetic(recipe, n_tris, sseed, device=local)
    build_s = time.time() - t0
    info = scene.info()
    prof = pkg.Profile(width=w, height=h, samples=spp, bounces=bounces, tonemap=tm)
    total_spp = spp * world
    R = multi.DistributedRenderer(scene, prof, seed=0)
    stream = torch.cuda.current_stream()
    rgb_host = torch.empty(h, w, 3, dtype=torch.uint8).pin_memory()

    def step():  # device-resident inputs: render this rank's samples, reduce, tonemap on rank 0
        return R.render(total_spp)

    def step_e2e():  # host buffers: H2D of the whole scene from pinned memory, render, D2H of the image
        nbytes = scene.upload(stream.cuda_stream)
        img = R.render(total_spp)
        if img is not None:
            rgb_host.copy_(img, non_blocking=True)
        return nbytes

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(n):
            fn()
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(args.warmup):
        step()
    barrier()
    with ClockSampler(local) as clk:
        ms_total = timed(step, args.steps)
    clocks = clk.summary()
    samples_per_step = w * h * total_spp
    value = samples_per_step * args.steps / ms_total / 1e3  # Msamples/s, whole job

    # one instrumented step: ray counts, launches, per-kernel time (events after every launch)
    pkg.set_kernel_timing(True)
    barrier()
    R.render(total_spp, want_stats=True)
    barrier()
    R.render(total_spp, want_stats=True)
    st = R.last_stats
    pkg.set_kernel_timing(False)
    rays_rank = st["extend_rays"] + st["shadow_rays"]
    rays_all = torch.tensor([float(rays_rank)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(rays_all)
    mrays = float(rays_all.item()) * args.steps / ms_total / 1e3

    # end to end through the public API with host buffers
    for _ in range(2):
        h2d = step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    e2e_value = samples_per_step * args.steps / ms_e2e / 1e3

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = peaks()
    canon = canonical_counts(args.workload)
    roof = None
    props = torch.cuda.get_device_properties(local)
    sm_count = props.multi_processor_count
    sm_mhz = clocks.get("sm_max_mhz") or 1965.0
    fp32_peak = sm_count * 128 * sm_mhz * 1e6 / 1e12  # T f32 instructions / s (x2 flops only if every op were an FMA)
    if canon:
        b_ext = ray_bytes(canon["n_node_closest"], canon["n_tri_closest"])
        b_sh = ray_bytes(canon["n_node_any"], canon["n_tri_any"])
        ext_bytes = b_ext * st["extend_rays"]  # algorithmic bytes of all extend launches of one step
        achieved = ext_bytes / (st["ms_extend"] * 1e-3) / 1e9 if st["ms_extend"] > 0 else None
        traffic = None
        try:
            with open(TRAFFIC_FILE) as f:
                per_ray = json.load(f).get(args.workload, {}).get("k_extend_dram_bytes_per_ray")
                traffic = per_ray * st["extend_rays"] / max(1, st["extend_launches"]) if per_ray else None
        except Exception:
            pass
        roof = {"bound": "hbm", "kernel": "k_extend", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak if achieved else None, "traffic": traffic, "peak_source": peak_src,
                "bytes_per_ray": b_ext, "n_node": canon["n_node_closest"], "n_tri": canon["n_tri_closest"],
                "rays_per_launch": st["extend_rays"] / max(1, st["extend_launches"]),
                "ms_per_launch": st["ms_extend"] / max(1, st["extend_launches"]),
                "algorithmic_bytes_per_launch": ext_bytes / max(1, st["extend_launches"]),
                "kernel_ms_per_step": {"extend": st["ms_extend"], "shade": st["ms_shade"], "shadow": st["ms_shadow"],
                                       "other": st["ms_other"], "total": st["ms_total"]},
                # second roofline of SURVEY 8(d): canonical f32 operations per ray (30 per node visit, 52 per triangle test)
                # against the FP32 issue peak SMs x 128 lanes x max SM clock; the scene being L2-resident or not, this is
                # the bound the kernel is actually nearest to (see profiles/: issue slots and L1 wavefronts co-limit)
                "fp32": {"flops_per_ray": canon.get("flops_per_ray_closest"),
                         "achieved_tflops": (canon.get("flops_per_ray_closest", 0.0) * st["extend_rays"] / (st["ms_extend"] * 1e-3) / 1e12)
                         if st["ms_extend"] > 0 else None,
                         "peak_tinstr": fp32_peak, "sm_count": sm_count, "sm_max_mhz": sm_mhz,
                         "frac": (canon.get("flops_per_ray_closest", 0.0) * st["extend_rays"] / (st["ms_extend"] * 1e-3) / 1e12 / fp32_peak)
                         if st["ms_extend"] > 0 and fp32_peak else None},
                "shadow_kernel": {"bytes_per_ray": b_sh,
                                  "achieved": b_sh * st["shadow_rays"] / (st["ms_shadow"] * 1e-3) / 1e9 if st["ms_shadow"] > 0 else None}}
    launches = (st["kernel_launches"] + 1) * args.steps  # + tonemap
    line = {
        "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "mrays_per_s": mrays,
        "config": {"workload": args.workload, "scene": "synthetic:%s seed %d, %d triangles, %d BVH4 nodes, %.0f MB on device"
                   % (recipe, sseed, info["n_tris"], info["n_nodes"], info["device_bytes"] / 1e6),
                   "resolution": [w, h], "spp_per_gpu": spp, "spp_total": total_spp, "bounces": bounces, "tonemap": tm,
                   "brdf": "COOK_TORRANCE", "split": "sample index mod n_gpus, one NCCL reduce of W*H*3 f32 per step",
                   "l2": "inputs larger than L2: scene %.0f MB + %.0f MB of wave state per wave vs 126 MB L2; no flush"
                         % (info["device_bytes"] / 1e6, 8.4e6 * 228 / 1e6),
                   "scene_build_s": build_s, "accel": scene.accel_info()},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(w * h * 3), "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": int(launches),
        "rays_per_step": {"extend": st["extend_rays"], "shadow": st["shadow_rays"], "alpha_recasts": st["alpha_recasts"]},
    }
    if roof:
        line["roofline"] = roof
    if world == 1 and not args.no_cpu_baseline:
        # bounded CPU sample: k spp of the same frame, sized from a 1-spp probe to ~cpu_seconds
        sa = scene.arrays()
        v1, r1, cores, s1, ob, osc = cpu_render_sample(pkg, wl, 1, scene_arrays=sa)
        k = max(1, min(spp, int(args.cpu_seconds / max(s1, 1e-3))))
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib
        oprof = oracle_lib.make_profile(w, h, spp, bounces, pkg.abi.TONEMAPS[tm])
        _, ms, cnt = osc.render_accum(oprof, rng_mode=1, sample_begin=1, n_samples=k, threads=cores, return_counters=True)
        line["cpu_baseline"] = {"value": w * h * k / ms / 1e3, "unit": "Msamples/s", "cores": cores, "kind": "port",
                                "mrays_per_s": float(cnt[0] + cnt[1]) / ms / 1e3,
                                "sample": "%d of %d spp of the full %dx%d frame (%.1f s of CPU work on %d threads)"
                                          % (k, spp, w, h, ms / 1e3, cores)}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    wl = list(WORKLOADS[args.workload])
    if args.spp:
        wl[5] = args.spp
    if args.impl == "reference":
        run_reference(args, tuple(wl))
    else:
        run_ours(args, tuple(wl))


if __name__ == "__main__":
    main()
