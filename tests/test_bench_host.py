"""Host-side logic of bench.py that needs no GPU: the weak-scaling sizes, the clock / throttle parser, the
`--impl reference` arm (the reference's CPU path, or the oracle port when oracle/_ref is not built) and the
loud failure of the product arm on a box without a CUDA device."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_weak_scaling_sizes():
    # flops per GPU constant: n = 32768 * N^(1/3) rounded to 512; N = 1 is the size the metric is quoted on
    assert [bench.workload_n(n) for n in (1, 2, 4, 8)] == [32768, 41472, 52224, 65536]
    assert bench.workload_n(8, 2000) == 2000
    for n_gpus in (2, 4, 8):
        per_gpu = bench.workload_n(n_gpus) ** 3 / n_gpus
        assert abs(per_gpu / 32768 ** 3 - 1.0) < 0.03


def test_both_arms_name_the_same_config():
    import argparse
    a = argparse.Namespace(dtype="f32", alg=6)
    assert bench.make_config(a, 4, 52224) == bench.make_config(a, 4, 52224)
    assert "n=52224" in bench.make_config(a, 4, 52224)["workload"]


def test_clock_sampler_parses_reasons():
    s = bench.ClockSampler(0)
    s.proc = type("P", (), {"terminate": lambda self: None, "wait": lambda self, timeout=None: 0, "kill": lambda self: None})()
    s.lines = ["0, 1965, 1965, 712.3, 0x0000000000000004, Not Active, Not Active, Not Active, Active",
               "0, 1905, 1965, 733.0, 0x0000000000000004, Not Active, Not Active, Not Active, Active",
               "0, 1875, 1965, 741.9, 0x0000000000000000, Not Active, Not Active, Not Active, Not Active",
               "garbage line", ""]
    got = s.stop()
    assert got["sm_mhz"] == 1905 and got["sm_max_mhz"] == 1965 and got["samples"] == 3
    assert got["reasons"] == ["sw_power_cap"] and abs(got["power_w_max"] - 741.9) < 1e-9
    s.lines = ["0, 1200, 1965, 500.0, 0x8, Active, Active, Not Active, Not Active"]
    assert s.stop()["reasons"] == ["hw_slowdown", "hw_thermal_slowdown"]
    s.lines = []
    assert s.stop()["reasons"] == ["no samples"]


def _run(args, env_extra=None, timeout=600):
    env = dict(os.environ)
    env.update(env_extra or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                          env=env, timeout=timeout, cwd=ROOT)


@pytest.mark.parametrize("dtype", ["f32", "f64"])
def test_reference_arm_prints_one_contract_line(dtype):
    r = _run(["--impl", "reference", "--steps", "1", "--warmup", "0", "--dtype", dtype, "--n", "768"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    line = json.loads(lines[0])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e", "gpu_launches"):
        assert key in line, key
    assert line["impl"] == "reference" and line["metric"] == bench.METRIC and line["unit"] == "TFLOP/s"
    assert line["dtype"] == dtype and line["vs_baseline"] is None and line["gpu_launches"] == 0
    assert line["value"] > 0 and line["e2e"]["value"] == line["value"] == line["cpu_baseline"]["value"]
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    assert "n=768" in line["cpu_baseline"]["sample"] and "workload" in line["config"]


def test_reference_arm_other_ranks_exit_silently():
    r = _run(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"], {"RANK": "1", "WORLD_SIZE": "2"}, timeout=60)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_product_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    r = _run(["--steps", "1", "--warmup", "0"], timeout=300)
    assert r.returncode != 0 and r.stdout.strip() == ""
    assert "no CPU fallback" in r.stderr
