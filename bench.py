#!/usr/bin/env python
"""bench.py — throughput of the render hot path (BASELINE.json metric: Msamples/s, with Mrays/s beside it).

  python bench.py --gpus N --steps K --warmup W            our CUDA path (one process per GPU under torchrun)
  python bench.py --impl reference --gpus N --steps K ...   the CPU restatement of the reference on the host cores

A "step" renders one full frame of the workload.  The frame's samples are split by sample index over the N ranks
(rank g renders the samples s = g mod N of the FULL frame: BASELINE.json's "spp-split", renderer/mod.rs:105-106 is the
loop being split), the accumulation buffers are summed onto rank 0 with one NCCL reduce and rank 0 tonemaps:
a fixed-total-spp split, i.e. STRONG scaling — value = W*H*spp / time for every N.  Time is measured with CUDA events
on the launching stream, max over ranks.  The weak-scaling figure (the workload's spp on EVERY rank) is reported as
the extra key `weak`.

Headline workload (BASELINE.json configs[2]): synthetic:console seed 3 (PS5-class stand-in, ~1.0 M
triangles, one 1024^2 texture, one point light), 1920x1080, 128 spp, 5 bounces, COOK_TORRANCE, FILMIC.
configs[3] (foliage, 5 M triangles with alpha, 256 spp) and configs[4] (city, 20 M triangles, 4K, 1024 spp, 8 bounces,
ACES) are measured in the same run as the sub-records `configs` of the line (same split, same timing rules).
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
    # name: (recipe, n_tris, scene seed, width, height, spp (whole frame), bounces, tonemap)
    "console_1080p_128spp_5b": ("console", 1000000, 3, 1920, 1080, 128, 5, "FILMIC"),
    "foliage_1080p_256spp_4b": ("foliage", 5000000, 4, 1920, 1080, 256, 4, "FILMIC"),
    "city_4k_1024spp_8b": ("city", 20000000, 5, 3840, 2160, 1024, 8, "ACES"),
    "console_small": ("console", 100000, 3, 480, 270, 8, 5, "FILMIC"),  # quick functional check
}
DEFAULT_WORKLOAD = "console_1080p_128spp_5b"
SUB_WORKLOADS = ["foliage_1080p_256spp_4b", "city_4k_1024spp_8b"]  # BASELINE.json configs[3], configs[4]
# acceleration-structure builder per workload: the 20 M-triangle city is built on the GPU (PLOC: 20 ms instead of 6 s of
# host SAH build per rank, and the tree traces 15 % faster than the SAH one on this scene, profiles/r07/builders_s7.jsonl)
ACCEL = {"city_4k_1024spp_8b": "ploc"}
CANON_FILE = os.path.join(ROOT, "profiles", "canonical_counts.json")
NCU_FILE = os.path.join(ROOT, "profiles", "ncu_traffic.json")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=None, help="override the frame's samples per pixel")
    ap.add_argument("--sub", default=None, help="comma-separated sub-record workloads ('' = none; default: configs[3], configs[4])")
    ap.add_argument("--sub-steps", type=int, default=3)
    ap.add_argument("--no-weak", action="store_true")
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
        self.nvml = self.handle = None
        try:  # the same counters through NVML, in process: an nvidia-smi process per sample (on every rank at once) takes
            # tens of ms of driver time, which is visible in a timed region of 50 ms (5 frames of 9 ms on 8 GPUs)
            import pynvml
            pynvml.nvmlInit()
            self.nvml, self.handle = pynvml, pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(index))
        except Exception:
            self.nvml = None

    @staticmethod
    def _physical_index(index):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if index < len(ids) and ids[index].isdigit():
                return int(ids[index])
        return index

    def _sample_nvml(self):
        n, h = self.nvml, self.handle
        sm = n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM)
        mx = n.nvmlDeviceGetMaxClockInfo(h, n.NVML_CLOCK_SM)
        r = n.nvmlDeviceGetCurrentClocksEventReasons(h) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
            else n.nvmlDeviceGetCurrentClocksThrottleReasons(h)
        act = lambda bit: "Active" if r & bit else "Not Active"
        return [str(sm), str(mx), act(0x8), act(0x40), act(0x20), act(0x4)]  # hw, hw thermal, sw thermal, sw power cap

    def _run(self):
        while not self.stop.is_set():
            try:
                if self.nvml is not None:
                    self.rows.append(self._sample_nvml())
                else:
                    out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                    self.rows.append([x.strip() for x in out.strip().split(",")])
            except Exception:
                pass
            self.stop.wait(0.02 if self.nvml is not None else 0.1)  # >= 2 samples in the shortest timed region (8 GPUs: 46 ms)

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
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
                "source": "NVML, in process" if self.nvml is not None else "nvidia-smi"}


def load_json(path, key):
    try:
        with open(path) as f:
            return json.load(f).get(key)
    except Exception:
        return None


def ray_bytes(n_node, n_tri):
    return 64.0 * n_node + 48.0 * n_tri + 48.0  # B_ray of SURVEY.md 8(d)


def config_of(name, wl, world):
    """The `config` object: identical for both arms (the driver compares them)."""
    recipe, n_tris, sseed, w, h, spp, bounces, tm = wl
    return {"workload": name, "scene": "synthetic:%s seed %d, %d triangles requested" % (recipe, sseed, n_tris),
            "resolution": [w, h], "spp_total": spp, "bounces": bounces, "tonemap": tm, "brdf": "COOK_TORRANCE",
            "split": "fixed total spp, sample index mod n_gpus, one reduce of W*H*3 f32 per step",
            "l2": "inputs larger than L2 (scene + wave state of a frame vs 126 MB); no flush"}


# ----------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle (C++ restatement of the reference renderer; the Rust binary cannot be built here) on all host
# cores.  This is the ONLY place bench.py touches oracle/: as the timed baseline, never as the product.  The scene
# comes from the generator source compiled into the oracle (oracle/synth_glue.cpp): libpt_b200.so is not loaded.
def oracle_scene(wl):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import __graft_entry__
    __graft_entry__.load_package()  # python structs only (abi.py); no shared library is opened
    import oracle_lib
    recipe, n_tris, sseed = wl[0], wl[1], wl[2]
    t0 = time.time()
    osc = oracle_lib.OracleScene.synthetic(recipe, n_tris, sseed)
    return oracle_lib, osc, time.time() - t0


def cpu_sample(oracle_lib, osc, wl, sample_begin, n_samples, cores):
    recipe, n_tris, sseed, w, h, spp, bounces, tm = wl
    from path_tracer_b200.abi import TONEMAPS
    prof = oracle_lib.make_profile(w, h, spp, bounces, TONEMAPS[tm])
    _, ms, cnt = osc.render_accum(prof, rng_mode=1, sample_begin=sample_begin, n_samples=n_samples, threads=cores,
                                  return_counters=True)
    return w * h * n_samples / ms / 1e3, float(cnt[0] + cnt[1]) / ms / 1e3, ms / 1e3


def run_reference(args, name, wl):
    """--impl reference: rank 0 times the CPU restatement on a bounded sample per step; other ranks exit."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    recipe, n_tris, sseed, w, h, spp, bounces, tm = wl
    oracle_lib, osc, build_s = oracle_scene(wl)
    cores = os.cpu_count() or 1
    spp_step = 1  # one sample per pixel of the full frame per step (a bounded sample of the frame's spp)
    vals, rays, secs = [], [], []
    for i in range(args.warmup + args.steps):
        v, r, s = cpu_sample(oracle_lib, osc, wl, i, spp_step, cores)
        if i >= args.warmup:
            vals.append(v)
            rays.append(r)
            secs.append(s)
    total_s = sum(secs)
    value = w * h * spp_step * len(secs) / total_s / 1e6
    sample = "%d spp of %d per step, full %dx%d frame, %d bounces, all %d host threads" % (spp_step, spp, w, h, bounces, cores)
    line = {
        "impl": "reference", "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_s / len(secs), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "mrays_per_s": sum(rays) / len(rays),
        "config": config_of(name, wl, args.gpus),
        "note": "CPU restatement of the reference (oracle/, C++, all-hits ray_cast semantics) on %d triangles; the Rust "
                "binary cannot be built in this image (no cargo/rustc, crates un-vendored)" % osc.n_tris,
        "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------------
class Env:
    """Process-wide state of our arm: ranks, stream, helpers."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        import __graft_entry__
        self.torch, self.dist = torch, dist
        self.pkg = __graft_entry__.load_package()
        from path_tracer_b200 import multi
        self.multi = multi
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback; use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        self.stream = torch.cuda.current_stream()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def timed(self, fn, n):
        """n calls of fn between two events on the launching stream, barrier + synchronize on both sides, max over ranks."""
        torch = self.torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        e0.record(self.stream)
        for _ in range(n):
            fn()
        e1.record(self.stream)
        self.barrier()
        return self.max_over_ranks(e0.elapsed_time(e1))

    def max_over_ranks(self, x):
        t = self.torch.tensor([float(x)], device="cuda", dtype=self.torch.float64)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, x):
        t = self.torch.tensor([float(x)], device="cuda", dtype=self.torch.float64)
        if self.world > 1:
            self.dist.all_reduce(t)
        return float(t.item())


def roofline_of(name, st, clocks, props):
    """The roofline object of the dominant kernel (k_extend).  `bound` is chosen from evidence, not assumed: when the
    ncu capture of the workload (profiles/ncu_traffic.json) shows the kernel's DRAM traffic far below its algorithmic
    bytes, the node / triangle fetches are served by the 126 MB L2 and the bound SURVEY 8(d) names for that case is the
    FP32 issue rate; the algorithmic-bytes figure is then reported as `hbm` beside it, never as the fraction."""
    canon = load_json(CANON_FILE, name)
    if not canon or st["ms_extend"] <= 0:
        return None
    peak_hbm, peak_src = peaks()
    ncu = load_json(NCU_FILE, name) or {}
    rays, sec = st["extend_rays"], st["ms_extend"] * 1e-3
    launches = max(1, st["extend_launches"])
    b_ray = ray_bytes(canon["n_node_closest"], canon["n_tri_closest"])
    f_ray = canon.get("flops_per_ray_closest", 30.0 * canon["n_node_closest"] + 52.0 * canon["n_tri_closest"])
    alg_gbs = b_ray * rays / sec / 1e9
    dram_per_ray = ncu.get("k_extend_dram_bytes_per_ray")
    dram_gbs = dram_per_ray * rays / sec / 1e9 if dram_per_ray else None
    sm_mhz = clocks.get("sm_max_mhz") or 1965.0
    fp32_peak = props.multi_processor_count * 128 * sm_mhz * 1e6 / 1e12  # T f32 instructions/s at the max SM clock
    fp32_ach = f_ray * rays / sec / 1e12
    l2_served = dram_per_ray is not None and dram_per_ray < 0.5 * b_ray
    hbm = {"algorithmic_gbs": alg_gbs, "algorithmic_frac_of_peak": alg_gbs / peak_hbm, "dram_gbs": dram_gbs,
           "dram_frac": dram_gbs / peak_hbm if dram_gbs else None, "peak": peak_hbm, "peak_source": peak_src,
           "bytes_per_ray": b_ray, "dram_bytes_per_ray_ncu": dram_per_ray}
    fp32 = {"flops_per_ray": f_ray, "achieved": fp32_ach, "peak": fp32_peak, "frac": fp32_ach / fp32_peak,
            "sm_count": props.multi_processor_count, "sm_max_mhz": sm_mhz}
    roof = {"kernel": "k_extend", "n_node": canon["n_node_closest"], "n_tri": canon["n_tri_closest"],
            "rays_per_launch": rays / launches, "ms_per_launch": st["ms_extend"] / launches,
            "traffic": dram_per_ray * rays / launches if dram_per_ray else None,
            "algorithmic_bytes_per_launch": b_ray * rays / launches,
            "kernel_ms_per_step": {"extend": st["ms_extend"], "shade": st["ms_shade"], "shadow": st["ms_shadow"],
                                   "other": st["ms_other"], "total": st["ms_total"]},
            "hbm": hbm, "fp32": fp32,
            # explanatory counters of the committed ncu --set full capture of this workload (profiles/): issue slots, L1
            # data pipe, lanes per instruction
            "ncu": {k: v for k, v in ncu.items() if k.endswith("_pct") or k.endswith("_lanes") or k == "source"}}
    if l2_served or (dram_per_ray is None and alg_gbs > 1.2 * peak_hbm):
        roof.update({"bound": "fp32-issue", "achieved": fp32_ach, "peak": fp32_peak, "unit": "Tinstr/s", "frac": fp32_ach / fp32_peak,
                     "why": "DRAM traffic (ncu) is %.0f B/ray against %.0f algorithmic: nodes and triangles come from L2; "
                            "canonical un-fused f32 operations per ray against SMs x 128 lanes x max clock"
                            % (dram_per_ray or 0.0, b_ray)})
    else:
        roof.update({"bound": "hbm", "achieved": alg_gbs, "peak": peak_hbm, "unit": "GB/s", "frac": alg_gbs / peak_hbm})
    b_sh = ray_bytes(canon["n_node_any"], canon["n_tri_any"])
    if st["ms_shadow"] > 0:
        roof["shadow_kernel"] = {"bytes_per_ray": b_sh, "algorithmic_gbs": b_sh * st["shadow_rays"] / (st["ms_shadow"] * 1e-3) / 1e9}
    return roof


def measure(env, name, wl, steps, warmup, want_weak):
    """One workload at the current world size -> dict (on rank 0) with value / e2e / roofline / breakdown."""
    torch, pkg = env.torch, env.pkg
    recipe, n_tris, sseed, w, h, spp, bounces, tm = wl
    pkg.set_accel_builder(ACCEL.get(name, "sah"))
    t0 = time.time()
    scene = pkg.Scene.synthetic(recipe, n_tris, sseed, device=env.local)
    build_s = time.time() - t0
    pkg.set_accel_builder("sah")
    info = scene.info()
    prof = pkg.Profile(width=w, height=h, samples=spp, bounces=bounces, tonemap=tm)
    R = env.multi.DistributedRenderer(scene, prof, seed=0)
    rgb_host = torch.empty(h, w, 3, dtype=torch.uint8).pin_memory()
    samples_per_step = w * h * spp

    def step():  # device-resident inputs: render this rank's share of the frame's samples, reduce, tonemap on rank 0
        return R.render(spp)

    for _ in range(warmup):
        step()
    env.barrier()
    with ClockSampler(env.local) as clk:
        ms_total = env.timed(step, steps)
    clocks = clk.summary()
    value = samples_per_step * steps / ms_total / 1e3  # Msamples/s, whole job

    # one instrumented step: ray counts, launches, per-kernel time (events after every launch), phase breakdown
    pkg.set_kernel_timing(True)
    env.barrier()
    R.render(spp, want_stats=True)
    env.barrier()
    phases = R.render_phases(spp)
    st = R.last_stats
    pkg.set_kernel_timing(False)
    rays_all = env.sum_over_ranks(st["extend_rays"] + st["shadow_rays"])
    ext_all, sh_all = env.sum_over_ranks(st["extend_rays"]), env.sum_over_ranks(st["shadow_rays"])
    mrays = rays_all * steps / ms_total / 1e3
    ms_render = env.max_over_ranks(phases["ms_render"])
    dropped = env.sum_over_ranks(st["shadow_dropped"])
    alpha_all = env.sum_over_ranks(st["alpha_tests"])

    # end to end through the public API with host buffers.  One GPU: the C-ABI calls a host makes per frame
    # (INTEGRATION.md): pt_scene_upload of the scene from pinned host memory + pt_render_async into a pinned host image
    # - both asynchronous, so the upload of frame k+1 overlaps the rendering of frame k; the timed region ends with the
    # last image on the host.  Several GPUs under torchrun: upload, this rank's pt_render_accum_device, the NCCL reduce,
    # pt_tonemap_device and the read-back of the image on rank 0.
    if env.world == 1:
        def step_e2e():
            nbytes = scene.upload(env.stream.cuda_stream)
            pkg.render_async(scene, prof, rgb_host.data_ptr(), stream=env.stream.cuda_stream)
            return nbytes
    else:
        def step_e2e():
            # up to 4 GPUs every rank re-sends the whole scene from its pinned host copy; from 8 up each rank sends 1/N of
            # every array and an in-place NCCL all-gather replicates it over NVSwitch - eight uploads of the same bytes
            # exceed what one host delivers (profiles/r07/s26_8gpu_bench.json: 11.1 ms per step against a 9.7 ms frame)
            nbytes = R.distribute_scene()
            img = R.render(spp)
            if img is not None:
                rgb_host.copy_(img, non_blocking=True)
            return nbytes
    h2d = 0
    for _ in range(2):
        h2d = step_e2e()
    ms_e2e = env.timed(step_e2e, steps)
    e2e_value = samples_per_step * steps / ms_e2e / 1e3
    h2d_all = env.sum_over_ranks(h2d)  # host -> device bytes of one step over all ranks

    weak = None
    if want_weak and env.world > 1:
        def step_weak():  # the workload's full spp on EVERY rank: N*spp samples per pixel in total
            return R.render(spp * env.world)
        step_weak()
        ms_w = env.timed(step_weak, steps)
        weak = {"value": samples_per_step * env.world * steps / ms_w / 1e3, "unit": "Msamples/s", "ms_per_step": ms_w / steps,
                "spp_per_gpu": spp, "spp_total": spp * env.world, "scaling": "weak"}

    out = None
    if env.rank == 0:
        props = torch.cuda.get_device_properties(env.local)
        out = {
            "workload": name, "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": env.world,
            "steps": steps, "warmup": warmup, "ms_per_step": ms_total / steps, "scaling": "strong", "mrays_per_s": mrays,
            "config": config_of(name, wl, env.world),
            "scene": {"n_tris": info["n_tris"], "n_nodes": info["n_nodes"], "device_mb": info["device_bytes"] / 1e6,
                      "build_s": build_s, "accel": scene.accel_info()},
            "spp_per_gpu": [env.multi.sample_partition(spp, env.world, g)[2] for g in range(env.world)],
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": int(h2d_all), "d2h_bytes_per_step": int(w * h * 3),
                    "ms_per_step": ms_e2e / steps,
                    "call": "pt_scene_upload + pt_render_async (pinned host image out), pipelined" if env.world == 1 else
                            "scene to every rank (DistributedRenderer.distribute_scene: pt_scene_upload per rank below 8 GPUs, "
                            "pt_scene_upload_part + NCCL all-gather from 8 up) + pt_render_accum_device per rank + NCCL reduce + "
                            "pt_tonemap_device + D2H on rank 0",
                    "h2d_bytes_per_step_all_ranks": int(h2d_all)},
            "gpu_launches": int((st["kernel_launches"] + 1) * steps),
            "rays_per_step": {"extend": ext_all, "shadow": sh_all, "alpha_tests": alpha_all,
                              "shadow_dropped": dropped},
            # where one frame goes (one instrumented frame, CUDA events on the stream): the slowest rank's render, then
            # the reduce and the tonemap as rank 0 sees them (the reduce figure includes waiting for that slowest rank)
            "breakdown_ms": {"render_max_over_ranks": ms_render, "reduce_rank0": phases["ms_reduce"],
                             "tonemap_rank0": phases["ms_tonemap"], "render_rank0": phases["ms_render"],
                             "render_device_rank0": phases["ms_render_device"],
                             "note": "per-kernel timing is on in this frame (serial schedule, an event after every launch); "
                                     "render_* include the host's enqueue of that frame, render_device is first launch -> last launch"},
            "roofline": roofline_of(name, st, clocks, props),
        }
        if weak:
            out["weak"] = weak
    scene.close()
    del R
    torch.cuda.empty_cache()
    return out


def run_ours(args, name, wl):
    env = Env()
    head = measure(env, name, wl, args.steps, args.warmup, not args.no_weak)
    subs = []
    sub_names = SUB_WORKLOADS if args.sub is None else [s for s in args.sub.split(",") if s]
    if name != DEFAULT_WORKLOAD and args.sub is None:
        sub_names = []
    for sn in sub_names:
        try:
            r = measure(env, sn, WORKLOADS[sn], max(1, min(args.steps, args.sub_steps)), 3, False)
        except Exception as e:  # a sub-record must not take the headline down with it
            r = {"workload": sn, "error": repr(e)} if env.rank == 0 else None
        if r is not None:
            subs.append(r)
    if env.rank != 0:
        if env.world > 1:
            env.dist.destroy_process_group()
        return
    line = {"metric": head["metric"], "value": head["value"], "unit": head["unit"], "n_gpus": env.world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "mrays_per_s": head["mrays_per_s"],
            "config": head["config"], "clocks": head["clocks"], "e2e": head["e2e"], "gpu_launches": head["gpu_launches"],
            "rays_per_step": head["rays_per_step"], "breakdown_ms": head["breakdown_ms"], "scene": head["scene"],
            "spp_per_gpu": head["spp_per_gpu"]}
    if head.get("roofline"):
        line["roofline"] = head["roofline"]
    if head.get("weak"):
        line["weak"] = head["weak"]
    if subs:
        line["configs"] = subs
    if env.world == 1 and not args.no_cpu_baseline:
        # bounded CPU sample: k spp of the same frame, sized from a 1-spp probe to ~cpu_seconds
        oracle_lib, osc, _ = oracle_scene(wl)
        cores = os.cpu_count() or 1
        v1, r1, s1 = cpu_sample(oracle_lib, osc, wl, 0, 1, cores)
        k = max(1, min(wl[5], int(args.cpu_seconds / max(s1, 1e-3))))
        v, r, s = cpu_sample(oracle_lib, osc, wl, 1, k, cores)
        line["cpu_baseline"] = {"value": v, "unit": "Msamples/s", "cores": cores, "kind": "port", "mrays_per_s": r,
                                "sample": "%d of %d spp of the full %dx%d frame (%.1f s of CPU work on %d threads)"
                                          % (k, wl[5], wl[3], wl[4], s, cores)}
    print(json.dumps(line), flush=True)
    if env.world > 1:
        env.dist.destroy_process_group()


def main():
    args = parse_args()
    wl = list(WORKLOADS[args.workload])
    if args.spp:
        wl[5] = args.spp
    if args.impl == "reference":
        run_reference(args, args.workload, tuple(wl))
    else:
        run_ours(args, args.workload, tuple(wl))


if __name__ == "__main__":
    main()
