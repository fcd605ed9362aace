"""bench.py's output contract, checked on the CPU: the reference arm prints exactly ONE JSON line on stdout with the
keys the driver reads, and the per-launch DRAM traffic lookup matches the committed ncu capture."""
import importlib.util
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench_module():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_reference_arm_prints_one_json_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--ref-seconds", "2"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "realtime_factor_full_pipeline" and d["unit"] == "audio-sec/sec"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["steps"] == 1
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]


def test_reference_arm_other_ranks_print_nothing():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0", "--ref-seconds", "2"], capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_traffic_lookup_follows_the_committed_capture():
    b = _bench_module()
    # the newest capture that holds a kernel wins (r02 for the kernels of this round, r01 for the one-CTA GRU)
    with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:
        t = json.load(f)
    with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
        t1 = json.load(f)
    w = t["workload"]
    assert b.ncu_traffic("tc_gru_kernel", w["streams"], w["hours"], w["sample_rate"]) == t1["kernels"]["tc_gru_kernel"]["dram_bytes_per_launch"]
    for k in ("tc_gru2_kernel", "stft_kernel_d1", "istft_stream_kernel"):
        v = b.ncu_traffic(k, w["streams"], w["hours"], w["sample_rate"])
        assert v == t["kernels"][k]["dram_bytes_per_launch"] and v > 1e9
        # DRAM traffic of a launch stays close to its algorithmic bytes (no wasted re-reads)
    frames = w["streams"] * b.ncu_traffic_chunks("stft_kernel_d1") * b.FRAMES_PER_CHUNK
    assert 0.9 < b.ncu_traffic("stft_kernel_d1", w["streams"], w["hours"], w["sample_rate"]) / (frames * 2280) < 1.1
    assert 0.9 < b.ncu_traffic("istft_stream_kernel", w["streams"], w["hours"], w["sample_rate"]) / (frames * b.ISTFT_BYTES_PER_FRAME) < 1.1
    assert b.ncu_traffic("tc_gru2_kernel", w["streams"] + 1, w["hours"], w["sample_rate"]) is None  # another workload: no figure
