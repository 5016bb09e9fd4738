"""Host<->device copy bandwidth of every rank, alone and all ranks at once, before and after binding the
process to the CPUs of its GPU's NUMA node (development aid; launched with torch.distributed.run)."""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from sphb200 import domain  # noqa: E402


def bw(dev, host, devbuf, direction, reps=6):
    st = torch.cuda.current_stream(dev)
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record(st)
    for _ in range(reps):
        if direction in ("h2d", "both"):
            devbuf[0].copy_(host[0], non_blocking=True)
        if direction in ("d2h",):
            host[1].copy_(devbuf[1], non_blocking=True)
    if direction == "both":
        s2 = bw.s2
        s2.wait_stream(st)
        with torch.cuda.stream(s2):
            for _ in range(reps):
                host[1].copy_(devbuf[1], non_blocking=True)
        st.wait_stream(s2)
    b.record(st)
    st.synchronize()
    nbytes = host[0].numel() * 8 * reps * (2 if direction == "both" else 1)
    return nbytes / (a.elapsed_time(b) * 1e-3) / 1e9


def main():
    rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    bw.s2 = torch.cuda.Stream(device=dev)
    info = domain.gpu_numa_info(lr)
    aff0 = sorted(os.sched_getaffinity(0))
    lines = ["rank %d: %s  affinity before: %d cpus [%s..%s]" % (rank, info, len(aff0), aff0[0], aff0[-1])]
    nelem = 64 * 1024 * 1024            # 512 MB per buffer
    devbuf = [torch.empty(nelem, dtype=torch.float64, device=dev) for _ in range(2)]
    for bound in (False, True):
        if bound:
            ok = domain.bind_to_gpu_numa(lr)
            lines.append("rank %d: bound=%s affinity now %d cpus" % (rank, ok, len(os.sched_getaffinity(0))))
        host = [torch.empty(nelem, dtype=torch.float64).pin_memory() for _ in range(2)]
        host[0].fill_(1.0); host[1].fill_(2.0)         # first touch under the current affinity
        bw(dev, host, devbuf, "h2d", 2)
        for direction in ("h2d", "d2h", "both"):
            # alone: ranks take turns
            alone = 0.0
            for r in range(world):
                dist.barrier()
                if r == rank:
                    alone = bw(dev, host, devbuf, direction)
            dist.barrier()
            together = bw(dev, host, devbuf, direction)
            t = torch.tensor([alone, together], dtype=torch.float64, device=dev)
            out = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(out, t)
            if rank == 0:
                print("bound=%s %-4s alone GB/s: %s | all ranks at once: %s (sum %.0f)" % (
                    bound, direction, " ".join("%.0f" % float(o[0]) for o in out),
                    " ".join("%.0f" % float(o[1]) for o in out), sum(float(o[1]) for o in out)), flush=True)
        del host
    allines = [None] * world
    dist.all_gather_object(allines, lines)
    if rank == 0:
        for ls in allines:
            for l in ls:
                print(l)
        os.system("nvidia-smi topo -m 2>&1 | head -30; lscpu | grep -i -E 'numa|socket|model name|^CPU\\(s\\)'; "
                  "cat /sys/fs/cgroup/cpuset.cpus.effective 2>/dev/null")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
