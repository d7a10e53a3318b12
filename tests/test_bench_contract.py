"""bench.py's output contract: exactly ONE JSON line on stdout with the agreed keys.
CPU: the reference arm (`--impl reference`, the oracle port on the host cores).
GPU: the product arm with a few steps."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
             "scaling", "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"}


def run_bench(*args, timeout=600):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True,
                         text=True, timeout=timeout, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, out.stdout[:500]                 # one line, nothing else on stdout
    return json.loads(lines[0])


def test_reference_arm_line(oracle):
    d = run_bench("--impl", "reference", "--steps", "3", "--warmup", "1")
    assert BASE_KEYS <= set(d) and d["impl"] == "reference"
    assert d["metric"] == "routed_requests_per_sec" and d["unit"] == "requests/s" and d["value"] > 0
    assert d["steps"] == 3 and d["warmup"] == 1 and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_print_nothing(oracle):
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                          "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


@pytest.mark.gpu
def test_product_arm_line(oracle, pk):
    """The default line at a driver-like K: headline C1 plus every sub-record the judge reads."""
    d = run_bench("--steps", "200", "--warmup", "5", timeout=900)
    assert BASE_KEYS <= set(d) and "impl" not in d
    assert d["n_gpus"] == 1 and d["steps"] == 200 and d["warmup"] >= 3 and d["scaling"] == "weak"
    r0 = d["roofline"]
    assert d["gpu_launches"] == -(-200 // 32) and r0["batches_per_launch"] == 200 / d["gpu_launches"]   # 32 staged batches per launch
    assert abs(r0["algorithmic_bytes_per_launch"] / r0["batches_per_launch"] - 8_584_333) < 200_000
    assert abs(r0["avg_launch_us"] * d["gpu_launches"] - d["ms_per_step"] * 200 * 1e3) < 1e-6 * d["ms_per_step"] * 200 * 1e3 + 1e-9
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and 0 < r["frac"] < 1.2 and r["peak"] > 1000
    assert abs(r["achieved"] / r["peak"] - r["frac"]) < 1e-9
    assert "traffic" in r and r["traffic_source"]                     # a number only with a capture of THESE kernel sources
    e = d["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] > 8_000_000 and e["d2h_bytes_per_step"] == 24 * 4096
    assert e["value"] < d["value"]                                   # copies cost something
    assert e["pcie_peak_gbs"] > 5 and 0 < e["frac_of_pcie"] < 1.5 and e["other_path"]["value"] > 0
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["value"] > 0 and cb["cores"] >= 1 and cb["one_core"] > 0
    c = d["clocks"]
    assert c["sm_max_mhz"] and c["sm_mhz"] and isinstance(c["reasons"], list)
    # config: the same keys as the reference arm prints (the driver compares them)
    assert d["config"] == C1_CONFIG
    assert "L2" in d["l2"] and "programmatic dependent launch" in d["issue"]
    modes = d["issue_modes_ms_per_step"]
    assert {"fused", "pdl4", "pdl", "serial", "plain", "plain_pdl", "four_chains"} <= set(modes) and modes["fused"] == d["ms_per_step"]
    assert d["serial"]["ms_per_step"] == modes["serial"] and 0 < d["serial"]["frac"] <= r["frac"] * 1.25
    ss = d["steady_state"]
    assert ss["steps"] >= 2000 and 0 < ss["frac"] < 1.2 and ss["ms_per_step"] > 0
    for w, nreq in (("c2", 65536), ("c3", 1024)):                     # BASELINE configs[2], [3]
        s = d["workloads"][w]
        assert 0 < s["frac"] < 1.2 and s["e2e"]["value"] > 0 and s["parity_gate"] and s["ms_per_step"] > 0
        assert abs(s["value"] - nreq / (s["ms_per_step"] / 1e3)) / s["value"] < 1e-9
    c4 = d["c4"]                                                      # BASELINE configs[4]
    assert c4["scaling"] == "strong" and c4["batches_per_shard"] == 245 and c4["e2e"]["value"] > 0
    assert abs(c4["value"] - 1e6 / (c4["pass_ms"] / 1e3)) / c4["value"] < 1e-9
    g = d["gcm"]                                                      # SURVEY §8(f) rank 1, cipher.go:110-141
    for k in ("65536x256B", "32768x4KiB"):
        assert g[k]["kernel_ms"] > 0 and g[k]["call_ms"] >= g[k]["kernel_ms"] and g[k]["openssl_key_once_ms"] > 0 and g[k]["parity_gate"]
    lat = d["latency"]                                                # the front-end at the service.Service seam
    assert lat["gpus"] == 1 and lat["p50_us"] > 0 and lat["p99_us"] >= lat["p50_us"] and len(lat["rates"]) >= 3
    assert e["with_pack"]["value"] == lat["sustained_per_s"] > 0


C1_CONFIG = {"workload": "C1: 4096 requests x 2048 B payloads, 8 providers (BASELINE configs[1])",
             "batch_requests": 4096, "payload_bytes": 2048, "providers": 8}


def test_reference_arm_config_is_the_product_arms(oracle):
    """`config` must compare equal between the two arms (the driver's same_config check): both
    print exactly this object for the default workload (the GPU test checks the product arm)."""
    ref = run_bench("--impl", "reference", "--steps", "2", "--warmup", "1")
    assert ref["config"] == C1_CONFIG


@pytest.mark.gpu
def test_c4_strong_scaling_line(oracle, pk):
    """BASELINE configs[4] as a bench workload: one rank's pass over the whole 1M-request stream."""
    d = run_bench("--workload", "c4", "--no-cpu", "--steps", "500", "--warmup", "5")
    assert BASE_KEYS <= set(d) and d["scaling"] == "strong" and d["n_gpus"] == 1
    c = d["config"]
    assert c["batches_per_shard"] == 245 and c["shard_requests_rank0"] == 1_000_000 and "configs[4]" in c["workload"]
    assert d["gpu_launches"] == d["steps"] and d["steps"] % 245 == 0
    assert abs(d["value"] - 1e6 / (d["pass_ms"] / 1e3)) / d["value"] < 1e-9
    assert 0 < d["e2e"]["value"] < d["value"] and 0 < d["roofline"]["frac"] < 1.2
