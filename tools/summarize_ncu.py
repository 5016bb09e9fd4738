"""Turns gpurun_out/launches.csv (ncu --metrics gpu__time_duration.sum) and a `--set full` .ncu-rep into
the small text summaries committed under profiles/ (development aid; run in the build container)."""
import collections
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "smsp__thread_inst_executed_per_inst_executed.ratio"]


EXTRA = ["l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
         "smsp__inst_executed_pipe_fp64.sum", "smsp__inst_executed_pipe_lsu.sum", "smsp__inst_executed_pipe_alu.sum",
         "smsp__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
         "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_dynamic"]


def launches(path, out):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    h = rows[0]
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        a = agg.setdefault(r[ki][:70], [0, 0.0])
        a[0] += 1
        a[1] += float(r[vi].replace(",", ""))
    tot = sum(a[1] for a in agg.values())
    with open(out, "w") as fh:
        fh.write("# kernel, launches, total us, share of GPU time (ncu per-launch times: cold-cache, serialised)\n")
        for n, a in sorted(agg.items(), key=lambda x: -x[1][1]):
            fh.write("%-72s %5d %10.1f %6.2f%%\n" % (n, a[0], a[1] / 1e3, 100 * a[1] / tot))


def full(rep, out, header=""):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--print-kernel-base", "demangled"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    h, u = rows[0], rows[1]
    with open(out, "w") as fh:
        if header:
            fh.write("# %s\n" % header)
        for v in rows[2:]:
            fh.write("== %s\n" % v[h.index("Kernel Name")])
            for i, n in enumerate(h):
                if n in KEYS or (n.startswith("smsp__average_warps_issue_stalled") and n.endswith("_per_issue_active.ratio")
                                 and float(v[i] or 0) > 0.25) or n in EXTRA:
                    fh.write("  %-75s %-14s %s\n" % (n, u[i], v[i]))


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3])
    else:
        full(sys.argv[2], sys.argv[3], " ".join(sys.argv[4:]))
