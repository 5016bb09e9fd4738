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
    assert line["cpu_baseline"]["sample_particles"] == 3000 and line["warmup"] == 0
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    # the default workload is BASELINE configs[3] (uniform, N = 10^7, seed 3) with moving particles; the CPU arm
    # times a bounded sample of it and extrapolates with the N log N law
    assert line["config"]["workload"].startswith("BASELINE configs[3]: uniform box") and "N=10000000" in line["config"]["workload"]
    assert line["config"]["particles_total"] == 10000000 and line["config"]["move_h_per_step"] == 0.2
    ex = line["cpu_baseline"]["extrapolated_to_config"]
    assert ex["particles"] == 10000000 and 0 < ex["value"] < line["value"]


def test_workload_names_follow_baseline_configs():
    sys.path.insert(0, ROOT)
    import bench
    class A(object):
        n = 1e6; kind = "uniform"; scaling = "weak"; move = 0.2
    assert bench.workload_name(A).startswith("BASELINE configs[1]") and "seed 1" in bench.workload_name(A)
    A.n = 1e7
    assert "configs[3]" in bench.workload_name(A) and "seed 3" in bench.workload_name(A)
    assert "per GPU x 8 GPUs (weak)" in bench.workload_name(A, 8)
    A.scaling = "strong"
    assert "split over 4 GPUs (strong)" in bench.workload_name(A, 4)
    A.kind = "clustered"; A.n = 1.25e7; A.scaling = "weak"
    assert "configs[4]" in bench.workload_name(A, 8) and "seed 4" in bench.workload_name(A, 8)
    A.n = 4e6
    assert "configs[2]" in bench.workload_name(A) and "seed 2" in bench.workload_name(A)


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


def test_numa_helpers_are_harmless_without_a_gpu():
    """bench.py binds every rank to the CPUs next to its GPU; without NVML / sysfs entries nothing must change."""
    sys.path.insert(0, ROOT)
    from sphb200 import domain
    assert domain._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert domain._parse_cpulist("") == set()
    before = os.sched_getaffinity(0)
    info = domain.gpu_numa_info(0)
    assert set(info) >= {"bus", "node", "cpus"}
    changed = domain.bind_to_gpu_numa(0)
    assert changed in (False, True)
    if info["node"] < 0:
        assert changed is False and os.sched_getaffinity(0) == before
    os.sched_setaffinity(0, before)


def test_committed_ncu_summary_feeds_the_roofline():
    """bench.py takes the k-NN kernel's DRAM traffic and instruction count from the newest committed capture."""
    sys.path.insert(0, ROOT)
    import bench
    traffic, src = bench.ncu_traffic(bench.PHASE_KERNEL["knn"])
    assert src == "r02_ncu_full_step.txt" and traffic > 1e7
    assert bench.ncu_particles(src) == 1.0e6
    inst, src2 = bench.ncu_value(bench.PHASE_KERNEL["knn"], "smsp__inst_executed.sum")
    assert src2 == src and 5e8 < inst < 2e9
    for ph in ("density", "pi", "force", "crossing"):
        t, s = bench.ncu_traffic(bench.PHASE_KERNEL[ph])
        assert s == src and t > 0
