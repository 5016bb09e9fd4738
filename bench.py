#!/usr/bin/env python
"""bench.py — SPH particle-steps/s of the B200 path vs the reference's CPU step (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--n PARTICLES_PER_GPU] [--impl ours|reference]

One "step" is one pass of the reference driver's while-body (code_running.py:217-540: crossing time,
dt, clamp, neighbour search, density, viscosity Pi, force, kick-drift, heating, smoothing-length
update) over one batch of synthetic particles.  Workload at N=1: BASELINE.json configs[1], the
uniform box of 10^6 gas particles at the reference's default number density, k = 40.  With --gpus N
(launched under torchrun, one rank per GPU) the box is N times longer along x, cut into N slabs of
10^6 particles, ghost particles exchanged over NCCL every step (sphb200.domain): weak scaling.

Printed JSON line (rank 0): value = particle-steps/s with the state resident in HBM; e2e = the same
step with the state in pinned HOST buffers, uploaded and downloaded inside the timed region;
roofline = the dominant kernel's algorithmic bytes / its CUDA-event time against the measured HBM
peak; cpu_baseline = the oracle (CPU port of the reference) timed on this box on a bounded sample.
`--impl reference` times that CPU port alone with all host threads.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "SPH particle-steps/sec"
UNIT = "particle-steps/s"
K_NEIGH = 40
DT_SCALE = 1e-8      # nsc.dt_0 is caller-set; see DESIGN.md "stability of the reference step"

# algorithmic bytes per particle per launch (DESIGN.md "Kernels")
ALGO_BYTES = {"sort": 216.0, "knn": 32.0, "density": 76.0, "pi": 116.0, "force": 156.0, "integrate": 248.0,
              "pack": 0.0, "crossing": 0.0, "dt_clamp": 0.0}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--n", type=float, default=1e6, help="particles per GPU")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--kind", default="uniform", choices=["uniform", "clustered"])
    ap.add_argument("--cpu-n", type=float, default=1e5, help="particles of the bounded CPU sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def ncu_traffic(kernel_substr):
    """dram read + write bytes per launch of one kernel from the committed ncu `--set full` summary
    (profiles/*_ncu_full_step.txt); None when no capture is present."""
    import glob
    import re
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_full_step.txt")), reverse=True):
        try:
            txt = open(path).read()
        except OSError:
            continue
        for block in txt.split("== ")[1:]:
            if kernel_substr in block.split("\n", 1)[0]:
                vals = {}
                for m in re.finditer(r"(dram__bytes_(?:read|write)\.sum)\s+(\w+)\s+([0-9.]+)", block):
                    scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}.get(m.group(2), 1.0)
                    vals[m.group(1)] = float(m.group(3)) * scale
                if len(vals) == 2:
                    return sum(vals.values()), os.path.basename(path)
    return None, None


def ncu_value(kernel_substr, metric):
    """One metric of one kernel from the committed ncu `--set full` summary; None when absent."""
    import glob
    import re
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_full_step.txt")), reverse=True):
        try:
            txt = open(path).read()
        except OSError:
            continue
        for block in txt.split("== ")[1:]:
            if kernel_substr in block.split("\n", 1)[0]:
                m = re.search(re.escape(metric) + r"\s+(?:\S+\s+)?([0-9.]+)\s*$", block, re.M)
                if m:
                    return float(m.group(1)), os.path.basename(path)
    return None, None


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as fh:
            p = json.load(fh)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# CPU legs (the only place outside tests/ and smoke() that executes oracle/)
# ------------------------------------------------------------------------------------------------
def make_state(n, kind, seed):
    from sphb200 import ics
    st = ics.uniform_box(n, seed) if kind == "uniform" else ics.clustered_cloud(n, seed)
    from sphb200 import constants as C
    st["dt_0"] = C.dt_0 * DT_SCALE
    return st


def cpu_steps(n, kind, steps, warmup, threads):
    """Time `steps` passes of the CPU port of the reference step (oracle.sph_oracle.step) on `n`
    particles.  threads == 1 is the reference as it runs (numpy/scipy single-threaded on these
    paths); threads > 1 lets cKDTree.query and the row-chunked hydro_update use that many threads."""
    from oracle import sph_oracle as O
    st = make_state(n, kind, 1)
    nfn = lambda p, dist, k: O.neighbors(p, dist, k, workers=(-1 if threads > 1 else 1))
    hfn = lambda *a: O.hydro_update(*a, species=False, threads=threads)
    s = O.startup(st, neighbors_fn=nfn, hydro_fn=hfn)
    for _ in range(warmup):
        s = O.step(s, neighbors_fn=nfn, hydro_fn=hfn)
    t0 = time.perf_counter()
    for _ in range(steps):
        s = O.step(s, neighbors_fn=nfn, hydro_fn=hfn)
    dt = time.perf_counter() - t0
    return n * steps / dt, dt / steps


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n = int(args.cpu_n)
    value, per_step = cpu_steps(n, args.kind, args.steps, min(args.warmup, 1), threads)
    sample = "%d steps of the CPU port on N=%d particles of the same %s box recipe (k=%d), %d threads" % (
        args.steps, n, args.kind, K_NEIGH, threads)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": per_step * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args), "k": K_NEIGH, "sample_particles": n},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "pair_interactions_per_s": 2.0 * K_NEIGH * value, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_name(args):
    n = int(args.n)
    if args.kind == "uniform":
        return "uniform box, N=%d gas particles per GPU at the reference's default number density, k=%d, seed 1" % (n, K_NEIGH)
    return "clustered cloud (lane_emden_modified.py:317-321 profile), N=%d per GPU, k=%d, seed 2" % (n, K_NEIGH)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
UPLOAD_FIELDS = ("points", "velocities", "total_accel", "mass", "sizes", "T", "E_internal", "mu_array",
                 "gamma_array", "particle_type", "densities_0")
# what a step changes (densities_0 is the step's density, code_running.py:540): 104 B/particle back to the host
DOWNLOAD_FIELDS = ("points", "velocities", "total_accel", "sizes", "T", "E_internal", "densities_0")


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device visible; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # a stuck collective must fail loudly and soon (stack of every thread on stderr), not hang the run
        import datetime
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ.get("BENCH_WATCHDOG_S", "420")), exit=True)
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))
    if rank == 0:
        ge.build()
    if world > 1:
        dist.barrier()
    from sphb200 import _lib
    from sphb200.session import PHASES
    _lib.load()
    n = int(args.n)

    if world > 1:
        from sphb200.domain import DistributedSession
        sess = DistributedSession.uniform_slabs(n, seed=1, k=K_NEIGH, dt_scale=DT_SCALE, device=local_rank)
        parallelism = "slab%d + NCCL ghost exchange" % world
    else:
        from sphb200.session import Session
        st = make_state(n, args.kind, 1 if args.kind == "uniform" else 2)
        sess = Session(n, K_NEIGH, st["dist"], device=local_rank, dt_0=st["dt_0"])
        sess.upload(st)
        parallelism = "single GPU"
    sess.startup()
    sess.check()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        sess.step()
    sess.check()
    barrier()
    launches0 = sess.launches()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    phases = sess.phases() if hasattr(sess, "phases") else PHASES
    K = args.steps
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(len(phases) + 1)] for _ in range(K)]
    barrier()
    t_wall0 = time.perf_counter()
    for s in range(K):
        for i, ph in enumerate(phases):
            ev[s][i].record()
            sess.phase(ph)
        ev[s][-1].record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop() if rank == 0 else None
    sess.check()
    total_ms = ev[0][0].elapsed_time(ev[K - 1][-1])
    launches = sess.launches() - launches0
    phase_ms = {ph: sum(ev[s][i].elapsed_time(ev[s][i + 1]) for s in range(K)) / K for i, ph in enumerate(phases)}
    if os.environ.get("BENCH_VERBOSE") and rank == 0:
        for s in range(K):
            sys.stderr.write("step %d: %s\n" % (s, " ".join("%s=%.3f" % (ph, ev[s][i].elapsed_time(ev[s][i + 1]))
                                                              for i, ph in enumerate(phases))))
    n_local = float(sess.n_owned() if hasattr(sess, "n_owned") else n)
    stats = torch.tensor([total_ms, n_local], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = stats.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = stats.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        total_ms = float(tmax[0]); n_total = float(tsum[1])
    else:
        n_total = n_local
    value = n_total * K / (total_ms * 1e-3)

    # ---- end to end: state in pinned host buffers, uploaded and downloaded every step ----------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(sess, world, dev, K, max(1, min(args.warmup, 2)), n_total, barrier)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = peaks()
    kernels = {}
    for ph in phases:
        ms = phase_ms[ph]
        b = ALGO_BYTES.get(ph, 0.0) * n_local
        kernels[ph] = {"ms": round(ms, 4), "algo_gb_per_launch": round(b / 1e9, 6),
                       "gbs": round(b / 1e9 / (ms * 1e-3), 1) if ms > 0 else None,
                       "frac_hbm": round(b / 1e9 / (ms * 1e-3) / peak, 4) if ms > 0 else None}
    top = max(phase_ms, key=lambda p: phase_ms[p])
    achieved = kernels[top]["gbs"] or 0.0
    top_kernel = {"knn": "knn_cell_kernel", "density": "density_kernel", "pi": "pi_kernel", "force": "force_kernel",
                  "crossing": "crossing_kernel", "integrate": "integrate_kernel", "sort": "reorder_kernel"}.get(top, top)
    traffic, traffic_src = ncu_traffic(top_kernel)
    roofline = {"kernel": top, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "note": "the k-NN kernel is bound by instruction issue / the fp64 pipe, not HBM (ncu: issue slots 74 % "
                        "busy, fp64 pipe 26 %, DRAM 2 %; profiles/README.md); its algorithmic traffic is 32 B/particle",
                "algo_bytes_per_particle": ALGO_BYTES.get(top), "avg_launch_ms": phase_ms[top],
                "density_plus_force_frac": round((ALGO_BYTES["density"] + ALGO_BYTES["force"]) * n_local / 1e9 /
                                                 ((phase_ms.get("density", 0) + phase_ms.get("force", 0)) * 1e-3) / peak, 4)
                if (phase_ms.get("density", 0) + phase_ms.get("force", 0)) > 0 else None}
    # the bound that does bind this kernel ("report both fractions and say which binds"):
    # warp instructions per launch from the committed ncu capture / the live launch time, against the SM's
    # issue rate (4 warp instructions per clock and SM at the sampled SM clock)
    inst, inst_src = ncu_value(top_kernel, "smsp__inst_executed.sum")
    if inst is not None and clocks and clocks.get("sm_mhz"):
        # the capture is of one N = 1e6 launch: scale to this run's particles per GPU
        inst_launch = inst * n_local / 1.0e6
        issue_peak = 148 * 4 * float(clocks["sm_mhz"]) * 1e6
        roofline["issue"] = {"bound": "instruction issue", "achieved": inst_launch / (phase_ms[top] * 1e-3) / 1e9,
                             "peak": issue_peak / 1e9, "unit": "G warp-instr/s", "frac": inst_launch / (phase_ms[top] * 1e-3) / issue_peak,
                             "warp_instr_per_particle": inst / 1.0e6, "source": inst_src,
                             "what": "instructions of %s only; the phase time also holds the work-list kernel" % top_kernel}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": args.warmup,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args), "k": K_NEIGH, "particles_total": int(n_total),
                       "parallelism": parallelism, "dt_0_scale": DT_SCALE,
                       "l2": "inputs larger than L2 (state + records + neighbour lists ~%d MB per GPU vs 126 MB L2)"
                             % int(n_local * 560 / 1e6)},
            "pair_interactions_per_s": 2.0 * K_NEIGH * value, "gpu_launches": int(launches),
            "wall_ms_per_step": t_wall / K * 1e3, "clocks": clocks, "roofline": roofline, "kernels": kernels}
    if e2e is not None:
        line["e2e"] = e2e
    if world == 1 and not args.no_cpu:
        cn = int(args.cpu_n)
        v, per = cpu_steps(cn, args.kind, 5, 0, 1)                     # ~10 s of CPU work at the default N = 1e5
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": 1, "kind": "port", "host_cpus": os.cpu_count(),
                                "sample": "5 steps of the CPU port (oracle/sph_oracle.py, numpy + scipy cKDTree, "
                                          "single-threaded as the reference runs) on N=%d particles of the same box "
                                          "recipe, %.2f s per step" % (cn, per)}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_e2e(sess, world, dev, K, W, n_total, barrier):
    """The same step with the particle state living in pinned HOST memory: every step uploads the
    full state (sess.step_host), runs the step and downloads the advanced state + densities."""
    import torch
    import torch.distributed as dist
    hs = sess.make_host_state(UPLOAD_FIELDS, DOWNLOAD_FIELDS)
    for _ in range(W):
        sess.step_host(hs)
    barrier()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    for _ in range(K):
        sess.step_host(hs)
    b.record()
    barrier()
    wall = time.perf_counter() - t0
    sess.check()
    ms = torch.tensor([a.elapsed_time(b), wall * 1e3, float(hs["h2d_bytes"]), float(hs["d2h_bytes"])],
                      dtype=torch.float64, device=dev)
    if world > 1:
        mx = ms.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = ms.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        ms = torch.stack([mx[0], mx[1], sm[2], sm[3]])
    ms = [float(x) for x in ms]
    return {"value": n_total * K / (ms[0] * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(ms[2]),
            "d2h_bytes_per_step": int(ms[3]), "ms_per_step": ms[0] / K, "wall_ms_per_step": ms[1] / K,
            "api": "Session.step_host: pinned host state -> sphb200_session_upload -> sphb200_session_phase (the nine "
                   "phases of the step; the copy of the late fields runs under the k-NN) -> sphb200_session_download "
                   "-> pinned host"}


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
