"""CPU: the parts of bench.py that run without a GPU — the reference arm (the CPU port with all host
threads) and the parsers that read the committed ncu summaries."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_the_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--cpu-n", "3000"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "SPH particle-steps/sec" and line["unit"] == "particle-steps/s"
    assert line["higher_is_better"] is True and line["value"] > 0 and line["steps"] == 1
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert line["config"]["workload"].startswith("uniform box, N=1000000")


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                          "--cpu-n", "3000"], capture_output=True, text=True, timeout=300, cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_ncu_summary_parsers():
    sys.path.insert(0, ROOT)
    import bench
    traffic, src = bench.ncu_traffic("knn_cell_kernel")
    inst, src2 = bench.ncu_value("knn_cell_kernel", "smsp__inst_executed.sum")
    assert src == src2 and src.endswith("_ncu_full_step.txt")
    assert 1e7 < traffic < 1e10 and 1e8 < inst < 1e10
    assert bench.ncu_traffic("no_such_kernel") == (None, None)
    assert set(bench.ALGO_BYTES) >= {"sort", "knn", "density", "pi", "force", "integrate"}
    assert sum(bench.ALGO_BYTES[k] for k in ("sort", "knn", "density", "force", "integrate")) == 728.0   # SURVEY.md §8d total
