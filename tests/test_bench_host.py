"""Host logic of bench.py and of the one-process-per-GPU driver that needs no GPU: the clock sampler's summary, the
roofline object (bound chosen from the committed ncu traffic, never a fraction above 1), the spp split and the parts of
the scene distribution."""
import importlib.util
import os
import types

import pytest

from conftest import ROOT

import path_tracer_b200.multi as M


@pytest.fixture(scope="module")
def bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_clock_sampler_summary(bench):
    s = bench.ClockSampler(0)
    assert s.summary() == {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
    s.rows = [["1965", "1965", "Not Active", "Not Active", "Not Active", "Not Active"],
              ["1800", "1965", "Not Active", "Not Active", "Not Active", "Active"],
              ["1950", "1965", "Not Active", "Not Active", "Not Active", "Not Active"],
              ["garbage"], ["[N/A]", "x", "", "", "", ""]]
    out = s.summary()
    assert out["sm_mhz"] == 1950.0 and out["sm_max_mhz"] == 1965.0  # median under load, max
    assert out["reasons"] == ["sw_power_cap"] and out["samples"] == 3
    s.rows.append(["1000", "1965", "Active", "Active", "Active", "Not Active"])
    assert set(s.summary()["reasons"]) == {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"}


def test_ray_bytes_is_surveys_formula(bench):
    assert bench.ray_bytes(10.0, 2.0) == 64 * 10 + 48 * 2 + 48  # B_ray = 64 n_node + 48 n_tri + 48


@pytest.mark.parametrize("name", ["console_1080p_128spp_5b", "foliage_1080p_256spp_4b", "city_4k_1024spp_8b"])
def test_roofline_object_is_bounded_and_says_why(bench, name):
    """With the committed counters the kernel's DRAM traffic is far below its algorithmic bytes, so the bound must be the
    FP32 issue rate and the fraction a fraction; the HBM view stays beside it as an explanation."""
    props = types.SimpleNamespace(multi_processor_count=148)
    clocks = {"sm_max_mhz": 1965.0}
    # a frame's worth of rays at roughly the measured rates of the three workloads
    st = {"ms_extend": 33.0, "ms_shade": 13.0, "ms_shadow": 20.0, "ms_other": 3.0, "ms_total": 69.0,
          "extend_rays": 4.5e8, "shadow_rays": 2.3e8, "extend_launches": 6}
    r = bench.roofline_of(name, st, clocks, props)
    assert r is not None and r["kernel"] == "k_extend"
    assert r["bound"] in ("fp32-issue", "hbm")
    assert 0.0 < r["frac"] <= 1.2 and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-12
    assert r["bound"] == "fp32-issue" and "why" in r  # ncu: tens of bytes per ray against hundreds algorithmic
    assert r["hbm"]["dram_bytes_per_ray_ncu"] < 0.5 * r["hbm"]["bytes_per_ray"]
    assert r["traffic"] == pytest.approx(r["hbm"]["dram_bytes_per_ray_ncu"] * st["extend_rays"] / 6)
    assert r["fp32"]["peak"] == pytest.approx(148 * 128 * 1965e6 / 1e12)
    # no ncu entry and an impossible algorithmic rate: still not reported as an HBM fraction above 1.2
    st_fast = dict(st, ms_extend=1.0)
    r2 = bench.roofline_of(name, st_fast, clocks, props)
    assert r2["bound"] == "fp32-issue" or r2["frac"] <= 1.2


def test_roofline_without_kernel_time_is_none(bench):
    st = {"ms_extend": 0.0, "ms_shade": 0.0, "ms_shadow": 0.0, "ms_other": 0.0, "ms_total": 1.0, "extend_rays": 1.0,
          "shadow_rays": 0.0, "extend_launches": 1}
    assert bench.roofline_of("console_1080p_128spp_5b", st, {}, types.SimpleNamespace(multi_processor_count=148)) is None


def test_workloads_match_baseline_configs(bench):
    import json
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert base["metric"].lower().startswith("msamples")
    w = bench.WORKLOADS[bench.DEFAULT_WORKLOAD]
    assert (w[3], w[4], w[5], w[6]) == (1920, 1080, 128, 5)  # configs[2]: 1080p, 128 spp, 5 bounces
    assert bench.WORKLOADS["foliage_1080p_256spp_4b"][5] == 256 and bench.WORKLOADS["city_4k_1024spp_8b"][3:7] == (3840, 2160, 1024, 8)
    cfg = bench.config_of(bench.DEFAULT_WORKLOAD, w, 8)
    assert cfg["workload"] == bench.DEFAULT_WORKLOAD and "model" not in cfg and cfg["spp_total"] == 128


@pytest.mark.parametrize("total,world", [(128, 1), (128, 8), (7, 3), (2, 4), (0, 2), (1024, 8)])
def test_sample_partition_covers_every_sample_once(total, world):
    seen = []
    for rank in range(world):
        begin, stride, count = M.sample_partition(total, world, rank)
        assert (begin, stride) == (rank, world) and count >= 0
        seen += [begin + k * stride for k in range(count)]
    assert sorted(seen) == list(range(total))
    with pytest.raises(ValueError):
        M.sample_partition(total, world, world)


@pytest.mark.parametrize("nbytes,parts", [(0, 2), (100, 2), (511, 2), (512, 2), (148298056 // 5, 8), (1 << 30, 3)])
def test_part_size_matches_the_c_abi_rule(nbytes, parts):
    """pt_scene_upload_part: shard = (size / parts) rounded down to 256; parts * shard + remainder == size."""
    n = M.part_size(nbytes, parts)
    assert n % 256 == 0 and n * parts <= nbytes and nbytes - n * parts < parts * 256 + parts
    assert n == ((nbytes // parts) // 256) * 256
