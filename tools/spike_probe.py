"""Per-step wall time of the slab-decomposed step with allocator counters (development aid)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist

def main():
    rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    from sphb200 import domain
    ds = domain.DistributedSession.uniform_slabs(1000000, 1, 40, 1e-8, lr)
    ds.startup()
    out = []
    for it in range(40):
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        ds.step()
        torch.cuda.synchronize()
        st = torch.cuda.memory_stats()
        out.append("%d:%.2f/%d/%d" % (it, (time.perf_counter() - t0) * 1e3, st["num_device_alloc"], st["num_device_free"]))
    print("rank %d  step:ms/cudaMallocs/cudaFrees  " % rank + " ".join(out), flush=True)
    dist.destroy_process_group()
main()
