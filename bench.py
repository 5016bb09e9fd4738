#!/usr/bin/env python
"""bench.py — SPH particle-steps/s of the B200 path vs the reference's CPU step (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--n PARTICLES] [--kind uniform|clustered]
                    [--scaling weak|strong] [--impl ours|reference] [--move FRACTION_OF_H]

One "step" is one pass of the reference driver's while-body (code_running.py:217-540: crossing time,
dt, clamp, neighbour search, density, viscosity Pi, force, kick-drift, heating, smoothing-length
update) over one batch of synthetic particles.

Workloads (BASELINE.json `configs`):
  N=1 default            configs[3], single-GPU half: uniform box, N = 10^7 gas particles at the reference's
                         default number density, k = 40, seed 3.  `--n 1e6` gives configs[1] (seed 1).
  --gpus N (default)     weak scaling of that box: N x 10^7 particles, slab decomposition along x, ghost
                         exchange over NCCL every step (sphb200.domain).  The JSON line also carries a short
                         `strong` leg: the SAME 10^7 particles split N ways (configs[3], strong-scaling half).
  --gpus N --kind clustered --n 1.25e7    configs[4]: clustered cloud, 1.25 x 10^7 particles per GPU.
  --scaling strong       make the strong split the headline of the line instead.

Particles MOVE: before every step each particle is displaced by `--move` (default 0.2) x the median smoothing
length per component (sphb200_session_perturb; a pure function of particle id and step number, so it is the
same for every decomposition).  The reference's own step has no stable regime that moves particles by a
sizeable fraction of h, so dt is kept tiny and the motion is imposed: the neighbour search
starts from stale radii, the cell order changes every step, and particles migrate between slabs.
The displacement kernel runs inside the timed region (phase "perturb").

Printed JSON line (rank 0): value = particle-steps/s with the state resident in HBM; e2e = the same step
with the state in pinned HOST buffers, uploaded and downloaded inside the timed region; roofline = the
dominant kernel's algorithmic bytes / its CUDA-event time against the measured HBM peak (plus fp64-pipe and
issue-rate fractions for the pair passes); cpu_baseline = the oracle (CPU port of the reference) timed on this
box on a bounded sample.  `--impl reference` times that CPU port alone with all host threads on
min(N, 10^6) particles of the same recipe.
"""
import argparse
import json
import math
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
PERTURB_SEED0 = 9000

# algorithmic bytes per particle per launch (DESIGN.md "Kernels")
ALGO_BYTES = {"sort": 216.0, "knn": 32.0, "density": 76.0, "pi": 116.0, "force": 156.0, "integrate": 248.0,
              "pack": 0.0, "crossing": 0.0, "dt_clamp": 0.0, "perturb": 0.0, "exchange": 0.0}
# fp64 instructions per evaluated pair of the pair passes (executed D* thread-instructions of the ncu source page
# divided by the pairs, profiles/r02_sass_pair_loops.txt; a DFMA counts as one instruction)
FP64_INSTR_PER_PAIR = {"density": 17.6, "pi": 36.4, "force": 31.6, "crossing": 9.2}
PHASE_KERNEL = {"knn": "knn_group_kernel", "density": "tile_pass_kernel<DensityPass", "pi": "tile_pass_kernel<PiPass",
                "force": "tile_pass_kernel<ForcePass", "crossing": "tile_pass_kernel<CrossingPass",
                "integrate": "integrate_kernel", "sort": "reorder_kernel"}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--n", "--particles", dest="n", type=float, default=1e7, help="particles per GPU (weak) / in total (strong)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--kind", default="uniform", choices=["uniform", "clustered"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--move", type=float, default=0.2, help="displacement per step and component, in units of the median h (0 = static)")
    ap.add_argument("--cpu-n", type=float, default=0, help="particles of the CPU sample (0 = min(N, 1e6) for --impl reference, 1e5 for the cpu_baseline leg)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-verify", action="store_true", help="skip the multi-GPU parity leg")
    ap.add_argument("--no-strong-leg", action="store_true", help="skip the secondary strong-scaling leg of a weak multi-GPU run")
    return ap.parse_args()


def seed_for(kind, n):
    """config 2 (uniform 10^6) seed 1, config 3 (clustered) seed 2, config 4 (uniform 10^7) seed 3,
    config 5 (clustered, weak) seed 4."""
    if kind == "uniform":
        return 1 if n <= 2000000 else 3
    return 2 if n <= 5000000 else 4


def workload_name(args, world=1, scaling=None):
    n = int(args.n)
    scaling = scaling or args.scaling
    seed = seed_for(args.kind, n)
    what = ("uniform box at the reference's default number density" if args.kind == "uniform"
            else "clustered cloud (imf/lane_emden_modified.py:317-321 profile)")
    if world == 1:
        size = "N=%d gas particles" % n
    elif scaling == "weak":
        size = "N=%d gas particles per GPU x %d GPUs (weak)" % (n, world)
    else:
        size = "N=%d gas particles split over %d GPUs (strong)" % (n, world)
    cfg = {("uniform", 1): "configs[1]", ("uniform", 3): "configs[3]", ("clustered", 2): "configs[2]",
           ("clustered", 4): "configs[4]"}[(args.kind, seed)]
    move = ", every particle displaced %.2g h per component and step" % args.move if args.move > 0 else ", static particles"
    return "BASELINE %s: %s, %s, k=%d, seed %d%s" % (cfg, what, size, K_NEIGH, seed, move)


def ncu_summaries():
    import glob
    return sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_full_step.txt")), reverse=True)


def ncu_traffic(kernel_substr):
    """dram read + write bytes per launch of one kernel from the newest committed ncu `--set full` summary
    (profiles/*_ncu_full_step.txt) that holds it; None when no capture is present."""
    import re
    for path in ncu_summaries():
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
    import re
    for path in ncu_summaries():
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


def ncu_particles(path):
    """Particles per launch of a committed capture (first line `# N=...`); 10^6 for round-1 captures."""
    import re
    try:
        m = re.search(r"^#.*\bN=(\d+)", open(os.path.join(ROOT, "profiles", path)).read(), re.M)
        return float(m.group(1)) if m else 1.0e6
    except OSError:
        return 1.0e6


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


def cpu_steps(n, kind, seed, steps, warmup, threads, move):
    """Time `steps` passes of the CPU port of the reference step (oracle.sph_oracle.step) on `n`
    particles, displaced between steps like the GPU arm.  threads == 1 is the reference as it runs
    (numpy/scipy single-threaded on these paths); threads > 1 lets cKDTree.query and the row-chunked
    hydro_update use that many threads.  The species densities of nsc:624-628 are left out on BOTH arms
    (the device step does not compute them either; SURVEY.md §8d lists them as optional)."""
    import numpy as np
    from oracle import sph_oracle as O
    from sphb200 import ics
    st = make_state(n, kind, seed)
    nfn = lambda p, dist, k: O.neighbors(p, dist, k, workers=(-1 if threads > 1 else 1))
    hfn = lambda *a: O.hydro_update(*a, species=False, threads=threads)
    s = O.startup(st, neighbors_fn=nfn, hydro_fn=hfn)
    amp = move * float(np.median(s["sizes"]))
    ids = np.arange(n)
    t0 = 0.0
    for it in range(warmup + steps):
        if it == warmup:
            t0 = time.perf_counter()
        if amp > 0:
            s["points"] = s["points"] + ics.perturbation(ids, amp, PERTURB_SEED0 + it)
        s = O.step(s, neighbors_fn=nfn, hydro_fn=hfn)
    dt = time.perf_counter() - t0
    return n * steps / dt, dt / steps


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    threads = os.cpu_count() or 1
    n_cfg = int(args.n)
    n = int(args.cpu_n) if args.cpu_n > 0 else min(n_cfg, 1000000)
    seed = seed_for(args.kind, n_cfg)
    value, per_step = cpu_steps(n, args.kind, seed, args.steps, args.warmup, threads, args.move)
    # the CPU step is a KD-tree build + query (N log N) plus linear passes: extrapolate the measured time per
    # step to the configured N with the N log N law (an upper bound on the CPU's particle-steps/s at that N)
    n_total = n_cfg * (world if args.scaling == "weak" else 1)
    extrap = None
    if n_total > n:
        t_cfg = per_step * (n_total * math.log(n_total)) / (n * math.log(n))
        extrap = {"particles": n_total, "s_per_step": t_cfg, "value": n_total / t_cfg, "law": "t ~ N log N from the measured sample"}
    sample = ("%d timed steps (+%d warm-up) of the CPU port on N=%d particles of the same %s recipe (seed %d, k=%d, same "
              "displacement per step), %d host threads (cKDTree workers=-1, row blocks of hydro_update on a thread pool), "
              "%.2f s per step; species densities excluded on both arms" % (args.steps, args.warmup, n, args.kind, seed,
                                                                          K_NEIGH, threads, per_step))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step * 1e3,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": run_config(args, world, n_total, "host CPU, %d threads" % threads),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                             "sample_particles": n, "extrapolated_to_config": extrap},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "pair_interactions_per_s": 2.0 * K_NEIGH * value, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_config(args, world, n_total, parallelism):
    """The `config` object: identical for both arms except `parallelism`."""
    return {"workload": workload_name(args, world), "k": K_NEIGH, "particles_total": int(n_total),
            "parallelism": parallelism, "dt_0_scale": DT_SCALE, "move_h_per_step": args.move,
            "l2": "inputs larger than L2 (state + records + neighbour lists ~%d MB per GPU vs 126 MB L2)"
                  % int(n_total / max(world, 1) * 480 / 1e6)}


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
UPLOAD_FIELDS = ("points", "velocities", "total_accel", "mass", "sizes", "T", "E_internal", "mu_array",
                 "gamma_array", "particle_type", "densities_0")
# what a step changes (densities_0 is the step's density, code_running.py:540): 104 B/particle back to the host
DOWNLOAD_FIELDS = ("points", "velocities", "total_accel", "sizes", "T", "E_internal", "densities_0")


def build_session(args, world, local_rank, scaling, n_arg):
    """The session of this rank for the given scaling mode; returns (session, parallelism, median h)."""
    import torch
    n = int(n_arg)
    seed = seed_for(args.kind, n)
    if world > 1:
        from sphb200.domain import DistributedSession
        sess = DistributedSession.synthetic(args.kind, n, seed, K_NEIGH, DT_SCALE, local_rank, scaling=scaling)
        parallelism = "slab%d + NCCL ghost exchange (%s)" % (world, scaling)
    else:
        from sphb200.session import Session
        st = make_state(n, args.kind, seed)
        sess = Session(n, K_NEIGH, st["dist"], device=local_rank, dt_0=st["dt_0"])
        sess.upload(st)
        del st
        parallelism = "single GPU"
    sess.startup()
    sess.check()
    return sess, parallelism, float(sess.median_h())


def timed_steps(sess, phases, K, W, amp, barrier, step0=0):
    """W warm-up + K timed steps, one CUDA event per phase boundary.  Returns (total ms, per-phase ms, wall s)."""
    import torch
    for it in range(W):
        if amp > 0:
            sess.perturb(amp, PERTURB_SEED0 + step0 + it)
        sess.step()
    sess.check()
    names = (("perturb",) if amp > 0 else ()) + tuple(phases)
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(len(names) + 1)] for _ in range(K)]
    barrier()
    t_wall0 = time.perf_counter()
    for s in range(K):
        for i, ph in enumerate(names):
            ev[s][i].record()
            if ph == "perturb":
                sess.perturb(amp, PERTURB_SEED0 + step0 + W + s)
            else:
                sess.phase(ph)
        ev[s][-1].record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    total_ms = ev[0][0].elapsed_time(ev[K - 1][-1])
    phase_ms = {ph: sum(ev[s][i].elapsed_time(ev[s][i + 1]) for s in range(K)) / K for i, ph in enumerate(names)}
    per_step = sorted(ev[s][0].elapsed_time(ev[s][-1]) for s in range(K))
    timed_steps.last_median_ms = per_step[K // 2] if K % 2 else 0.5 * (per_step[K // 2 - 1] + per_step[K // 2])
    if os.environ.get("BENCH_VERBOSE") == "2":                   # every rank: mean phase times (who waits for whom)
        sys.stderr.write("rank %s n=%s: %s\n" % (os.environ.get("RANK", "0"), getattr(getattr(sess, "sess", sess), "n", "?"),
                                                 " ".join("%s=%.3f" % (ph, v) for ph, v in phase_ms.items())))
    if os.environ.get("BENCH_VERBOSE") and int(os.environ.get("RANK", "0")) == 0:
        for s in range(K):
            sys.stderr.write("step %d: %s\n" % (s, " ".join("%s=%.3f" % (ph, ev[s][i].elapsed_time(ev[s][i + 1]))
                                                              for i, ph in enumerate(names))))
    return total_ms, phase_ms, t_wall


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
        faulthandler.dump_traceback_later(int(os.environ.get("BENCH_WATCHDOG_S", "900")), exit=True)
        from sphb200 import domain as _domain
        _domain.bind_to_gpu_numa(local_rank)      # pinned staging buffers next to the GPU's PCIe root (no-op on one NUMA node)
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=300))
    if rank == 0:
        ge.build()
    if world > 1:
        dist.barrier()
    from sphb200 import _lib
    _lib.load()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allred(vals, op):
        t = torch.tensor(vals, dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=op)
        return [float(x) for x in t]

    # ---- multi-GPU parity leg (not timed): a small slab run against the single-process oracle -----------
    verify = None
    if world > 1 and not args.no_verify:
        verify = run_verify(world, rank, local_rank)

    sess, parallelism, h_med = build_session(args, world, local_rank, args.scaling, args.n)
    amp = args.move * h_med
    K, W = args.steps, args.warmup
    phases = sess.phases()
    launches0 = sess.launches()
    sess.counters(reset=True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    total_ms, phase_ms, t_wall = timed_steps(sess, phases, K, W, amp, barrier)
    median_ms = timed_steps.last_median_ms
    clocks = sampler.stop() if rank == 0 else None
    sess.check()
    launches = sess.launches() - launches0
    counters = sess.counters(reset=True)
    n_local = float(sess.n_owned())
    total_ms = allred([total_ms], dist.ReduceOp.MAX)[0] if world > 1 else total_ms
    n_total = allred([n_local], dist.ReduceOp.SUM)[0] if world > 1 else n_local
    value = n_total * K / (total_ms * 1e-3)
    fp64_peak = sess.fp64_peak() if rank == 0 else None           # G DFMA warp-instructions/s... see below

    # ---- end to end: state in pinned host buffers, uploaded and downloaded every step ----------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(sess, world, dev, K, max(1, min(W, 2)), n_total, barrier, amp, W + K)

    # ---- secondary leg of a weak multi-GPU run: the strong split of configs[3] -------------------
    strong = None
    if world > 1 and args.scaling == "weak" and args.kind == "uniform" and not args.no_strong_leg:
        sess.close()
        del sess
        torch.cuda.empty_cache()
        s2, _, h2 = build_session(args, world, local_rank, "strong", 1e7)
        k2 = max(5, K // 2)
        s2.counters(reset=True)
        ms2, ph2, _ = timed_steps(s2, s2.phases(), k2, 3, args.move * h2, barrier)
        s2.check()
        c2 = s2.counters(reset=True)
        ms2 = allred([ms2], dist.ReduceOp.MAX)[0]
        nt2 = allred([float(s2.n_owned())], dist.ReduceOp.SUM)[0]
        strong = {"workload": "BASELINE configs[3]: uniform box, N=%d in total split over %d GPUs (strong), seed 3" % (int(nt2), world),
                  "particles_total": int(nt2), "steps": k2, "warmup": 3, "ms_per_step": ms2 / k2,
                  "value": nt2 * k2 / (ms2 * 1e-3), "unit": UNIT, "phase_ms": {p: round(v, 4) for p, v in ph2.items()},
                  "per_step_rank0": {k: (v / float(k2 + 3) if k != "knn_worklist_last" else v) for k, v in c2.items()},
                  "particles_rank0": int(s2.n_owned()), "particles_with_ghosts_rank0": int(s2.sess.n),
                  "ms_per_step_median_rank0": round(timed_steps.last_median_ms, 4),
                  "note": "mean over the timed steps; a step in which a rank's ghost count outgrows its buffers "
                          "re-allocates them (tens of ms, once): the median is the typical step"}
        sess = s2

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = peaks()
    pairs = n_local * K_NEIGH
    kernels = {}
    for ph, ms in phase_ms.items():
        b = ALGO_BYTES.get(ph, 0.0) * n_local
        kernels[ph] = {"ms": round(ms, 4), "algo_gb_per_launch": round(b / 1e9, 6),
                       "gbs": round(b / 1e9 / (ms * 1e-3), 1) if ms > 0 else None,
                       "frac_hbm": round(b / 1e9 / (ms * 1e-3) / peak, 4) if ms > 0 else None}
        if ph in FP64_INSTR_PER_PAIR and fp64_peak and ms > 0:
            # fp64 thread-instructions of the pair loop per second against the measured DFMA issue rate
            rate = FP64_INSTR_PER_PAIR[ph] * pairs / (ms * 1e-3)
            kernels[ph]["frac_fp64"] = round(rate / fp64_peak["dfma_per_s"], 4)
    top = max((p for p in phase_ms if p != "exchange"), key=lambda p: phase_ms[p])
    achieved = kernels[top]["gbs"] or 0.0
    top_kernel = PHASE_KERNEL.get(top, top)
    traffic, traffic_src = ncu_traffic(top_kernel)
    if traffic is not None:
        traffic *= n_local / ncu_particles(traffic_src)           # the capture's N -> this run's particles per GPU
    dens_force_ms = phase_ms.get("density", 0) + phase_ms.get("force", 0)
    roofline = {"kernel": top, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "note": "the k-NN and the pair passes are bound by instruction issue, the fp64 pipe and shared-memory "
                        "bank conflicts, not by HBM (profiles/README.md); per-phase frac_hbm / frac_fp64 are in `kernels`",
                "algo_bytes_per_particle": ALGO_BYTES.get(top), "avg_launch_ms": phase_ms[top],
                "density_plus_force_frac": round((ALGO_BYTES["density"] + ALGO_BYTES["force"]) * n_local / 1e9 /
                                                 (dens_force_ms * 1e-3) / peak, 4) if dens_force_ms > 0 else None,
                "fp64_peak": fp64_peak}
    # the bound that does bind the dominant kernel ("report both fractions and say which binds"):
    # warp instructions per launch from the committed ncu capture / the live launch time, against the SM's
    # issue rate (4 warp instructions per clock and SM at the sampled SM clock)
    inst, inst_src = ncu_value(top_kernel, "smsp__inst_executed.sum")
    if inst is not None and clocks and clocks.get("sm_mhz"):
        per_particle = inst / ncu_particles(inst_src)
        inst_launch = per_particle * n_local
        issue_peak = 148 * 4 * float(clocks["sm_mhz"]) * 1e6
        roofline["issue"] = {"bound": "instruction issue", "achieved": inst_launch / (phase_ms[top] * 1e-3) / 1e9,
                             "peak": issue_peak / 1e9, "unit": "G warp-instr/s", "frac": inst_launch / (phase_ms[top] * 1e-3) / issue_peak,
                             "warp_instr_per_particle": per_particle, "source": inst_src,
                             "what": "instructions of %s* only; the phase time also holds the escape-path kernels" % top_kernel}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": run_config(args, world, n_total, parallelism),
            "pair_interactions_per_s": 2.0 * K_NEIGH * value, "gpu_launches": int(launches),
            "wall_ms_per_step": t_wall / K * 1e3, "ms_per_step_median_rank0": round(median_ms, 4), "clocks": clocks, "roofline": roofline, "kernels": kernels,
            "per_step": {k: (v / float(K + W) if k != "knn_worklist_last" else v) for k, v in counters.items()}}
    if verify is not None:
        line["verify"] = verify
    if strong is not None:
        line["strong"] = strong
    if e2e is not None:
        line["e2e"] = e2e
    if world == 1 and not args.no_e2e:
        sess.close()
        torch.cuda.empty_cache()
        line["e2e_dropin"] = run_dropin(1000000, 3)
    if world == 1 and not args.no_cpu:
        cn = int(args.cpu_n) if args.cpu_n > 0 else 100000
        v, per = cpu_steps(cn, args.kind, seed_for(args.kind, int(args.n)), 5, 0, 1, args.move)   # ~10 s of CPU work at N = 1e5
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": 1, "kind": "port", "host_cpus": os.cpu_count(),
                                "sample": "5 steps of the CPU port (oracle/sph_oracle.py, numpy + scipy cKDTree, "
                                          "single-threaded as the reference runs) on N=%d particles of the same box "
                                          "recipe, %.2f s per step" % (cn, per)}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_verify(world, rank, local_rank):
    """Slab-decomposed run of N_total = 4x10^4 moving particles, 2 steps, gathered on every rank and compared
    on rank 0 with the single-process oracle: neighbour rows as sets of particle ids (bit-exact), h bit-exact,
    positions / densities to 1e-10.  The oracle is the CHECKER here (never timed, never on the product path)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from sphb200 import ics
    from sphb200 import constants as C
    from sphb200.domain import DistributedSession
    n = 40000
    st = ics.uniform_box(n, 5)
    st["dt_0"] = C.dt_0 * 1e-6
    ds = DistributedSession.from_global_state(st, local_rank, dt_0=st["dt_0"])
    ds.startup()
    msg = "ok"
    ref = None
    if rank == 0:
        from oracle import sph_oracle as O
        ref = O.startup(st)
    amp = None
    try:
        hs = ds.gather("sizes")
        if rank == 0:
            assert np.array_equal(hs, ref["sizes"]), "start-up h differs"
        amp = 0.2 * float(np.median(hs))
        ids = np.arange(n)
        for it in range(2):
            ds.perturb(amp, 300 + it)
            ds.step(); ds.check()
            got = {f: ds.gather(f) for f in ("points", "sizes", "densities", "velocities")}
            rows = ds.gather_neighbor_ids()
            if rank == 0:
                ref["points"] = ref["points"] + ics.perturbation(ids, amp, 300 + it)
                ref = O.step(ref)
                assert np.array_equal(np.sort(rows, axis=1), np.sort(ref["neighbor"], axis=1)), "neighbour sets differ (step %d)" % it
                for f, g in got.items():
                    err = float(np.max(np.abs(g - ref[f])) / np.max(np.abs(ref[f])))
                    assert err < 1e-10, "%s differs by %.2e (step %d)" % (f, err, it)
        stats = ds.counters(reset=True)
        if rank == 0:
            msg = "ok: N=%d over %d slabs, 2 moving steps, neighbour sets bit-exact, x/v/h/rho within 1e-10; %d particles migrated" % (
                n, world, int(stats.get("migrated", 0)))
    except AssertionError as e:
        msg = "FAILED: %s" % e
    ds.close()
    flag = torch.tensor([0.0 if msg.startswith("ok") else 1.0], dtype=torch.float64, device=torch.device("cuda", local_rank))
    dist.all_reduce(flag, op=dist.ReduceOp.MAX)
    torch.cuda.empty_cache()
    return msg


def run_dropin(n, steps):
    """The path a user of the reference gets when they only swap the module: `nsc.neighbors` plus the loop-generation
    calls of the driver's while-body (code_running.py:437,451-458,222) with numpy arrays in and out, every call
    crossing PCIe both ways (the uploaded neighbour array is reused between the calls of one step)."""
    import numpy as np
    from sphb200 import ics
    from sphb200 import navier_stokes_cleaned as nsc
    st = ics.uniform_box(n, 1)
    p, m, pt, v = st["points"], st["mass"], st["particle_type"], st["velocities"]
    E, g, mu, T = st["E_internal"], st["gamma_array"], st["mu_array"], st["T"]
    nsc.d, nsc.dt_0 = st["d"], st["dt_0"] if "dt_0" in st else nsc.dt_0
    h2d = d2h = 0
    t0 = None
    for it in range(steps + 1):
        if it == 1:
            t0 = time.perf_counter()
        nb, _, dd, nontriv, sizes = nsc.neighbors(p, st["dist"], K_NEIGH)                 # :437
        rho = nsc.density(p, m, pt, nb)                                                    # :451
        rho_d = nsc.dust_density(p, m, nb, pt, sizes)                                      # :452
        nd = nsc.num_dens(m, p, mu, nb)                                                    # :453
        gp = nsc.del_pressure(p, m, pt, nb, E, g)                                          # :456
        va, vh = nsc.artificial_viscosity(nb, p, pt, sizes, m, rho, v, T, g, mu)           # :458
        ct = nsc.crossing_time(nb, v, sizes, pt)                                           # :222
    per = (time.perf_counter() - t0) / steps
    h2d = n * (24 + 8 * K_NEIGH + 6 * 24 + 30 * 8)      # points for the search, the neighbour array once, per-call vectors
    d2h = n * (16 * K_NEIGH + 16 + 3 * 8 + 24 + 32)
    return {"value": n / per, "unit": UNIT, "ms_per_step": per * 1e3, "particles": n, "steps": steps,
            "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
            "api": "nsc.* drop-in: neighbors + density + dust_density + num_dens + del_pressure + artificial_viscosity + "
                   "crossing_time, numpy in / numpy out (loop generation, the calls of code_running.py:437,451-458,222)"}


def run_e2e(sess, world, dev, K, W, n_total, barrier, amp, step0):
    """The same step with the particle state living in pinned HOST memory: every step uploads the
    full state (sess.step_host), runs the step and downloads the advanced state + densities."""
    import torch
    import torch.distributed as dist
    hs = sess.make_host_state(UPLOAD_FIELDS, DOWNLOAD_FIELDS)
    for it in range(W):
        sess.step_host(hs, perturb=(amp, PERTURB_SEED0 + step0 + it) if amp > 0 else None)
    barrier()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    for it in range(K):
        sess.step_host(hs, perturb=(amp, PERTURB_SEED0 + step0 + W + it) if amp > 0 else None)
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
            "api": "Session.step_host: pinned host state -> sphb200_session_upload -> sphb200_session_phase (the phases "
                   "of the step; the copy of the late fields runs under the k-NN) -> sphb200_session_download "
                   "-> pinned host"}


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
