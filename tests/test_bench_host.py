"""Host-side pieces of bench.py that need no GPU: the roofline records, the flop count of a PPO iteration and the
configuration table (BASELINE.json's five configs)."""
import json
import os

import bench

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tensor_roofline_record():
    peak, src = bench.tensor_peak()
    assert peak > 0 and isinstance(src, str)
    r = bench.gemm_roofline({"tflops": 90.0}, 3)
    assert r["bound"] == "tensor" and r["unit"] == "TFLOP/s" and r["peak"] == round(peak, 1)
    assert abs(r["frac"] - 90.0 / peak) < 1e-3 and abs(r["issued_frac"] - 270.0 / peak) < 1e-3
    assert r["issued"] == 270.0
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):   # tf32 issues at half the bf16 rate
        m = json.load(open(p))
        assert abs(peak - float(m.get("bf16_tflops_sustained") or m["bf16_tflops"]) / 2.0) < 1e-6


def test_ppo_flops_follow_survey_8d():
    """0.234 / 1.68 TFLOP per PPO iteration for configs 1 and 3 (fwd + dX + dW over actor and critic)."""
    c1 = dict(obs_dim=48, act_dim=12, hidden=[128, 128, 128], T=24, N=4096, epochs=5)
    c3 = dict(obs_dim=235, act_dim=12, hidden=[512, 256, 128], T=24, N=4096, epochs=5)
    assert abs(bench.ppo_flops_per_iter(c1) / 1e12 - 0.234) < 0.01
    assert abs(bench.ppo_flops_per_iter(c3) / 1e12 - 1.68) < 0.03


def test_config_table_matches_baseline_json():
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert set(bench.CONFIGS) == {"cfg1", "cfg2", "cfg3", "cfg4", "cfg5"} and bench.HEADLINE == "cfg3"
    assert len(base["configs"]) == len(bench.CONFIGS)
    assert bench.CONFIGS["cfg3"]["envs"] == 4096 and bench.CONFIGS["cfg4"]["envs"] == 16384 and bench.CONFIGS["cfg5"]["envs"] == 8192
    assert bench.METRIC and bench.UNIT == "env-steps/s"
