"""bench.py's reference arm and helpers on CPU: `--impl reference` times the oracle (the C++ port of the
reference algorithm) on the host cores and prints ONE JSON line with the contract's keys; under
torchrun only rank 0 works.  (The GPU arm needs a B200; its line is checked by the driver.)"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_line(oracle):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                          "--count", "64", "--cpu-sample", "16"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "sequences folded/sec" and d["unit"] == "seq/s"
    assert d["higher_is_better"] is True and d["vs_baseline"] is None and d["value"] > 0 and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "seq/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["nb_mode"] == 100 and d["config"]["max_stack"] == 50 and "workload" in d["config"] and "sample" in d["config"]
    # ranks other than 0 exit 0 without work and without output
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1"],
                         capture_output=True, text=True, timeout=120, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_bench_helpers():
    sys.path.insert(0, ROOT)
    import bench
    s, k = bench.cpu_sample(list(range(100)), 10)
    assert s == list(range(0, 100, 10)) and k == 10
    assert bench.cpu_sample([1, 2, 3], 50) == ([1, 2, 3], 1)
    st = bench.stats_ms([5.0, 1.0, 3.0, 2.0, 4.0])
    assert st["median_ms"] == 3.0 and st["min_ms"] == 1.0 and st["max_ms"] == 5.0 and st["p95_ms"] == 5.0
    assert set(bench.WORKLOADS) == {"cfg2", "cfg3", "cfg4"} and bench.WORKLOADS["cfg3"]["max_stack"] == 200
    # the committed instruction counters cover every workload the roofline is reported for
    c = json.load(open(bench.COUNTERS))
    for w in bench.WORKLOADS:
        assert c[w]["expand_fast"]["warp_inst_per_step"] > 0 and 0 < c[w]["expand_fast"]["threads_per_inst"] <= 32
