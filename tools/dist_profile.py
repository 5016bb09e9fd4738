"""Host-side profile of the slab-decomposed step (development aid).
torchrun --nproc-per-node 2 tools/dist_profile.py [n_per_rank]"""
import cProfile
import os
import pstats
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1000000
    rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    from sphb200.domain import DistributedSession
    ds = DistributedSession.uniform_slabs(n, 1, 40, 1e-8, lr)
    ds.startup()
    for _ in range(3):
        ds.step()
    torch.cuda.synchronize(); dist.barrier()
    if os.environ.get("HOST_PROFILE"):
        pr = cProfile.Profile()
        pr.enable()
        for _ in range(10):
            ds.step()
        torch.cuda.synchronize()
        pr.disable()
        if rank == 0:
            st = pstats.Stats(pr); st.sort_stats("cumulative").print_stats(45)
    else:
        from torch.profiler import profile, ProfilerActivity
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            for _ in range(5):
                ds.step()
            torch.cuda.synchronize()
        if rank == 0:
            print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=70))
            ka = prof.key_averages()
            busy = sum(getattr(e, "self_device_time_total", 0) for e in ka)
            print("GPU busy (sum of kernel + memcpy time) per step: %.3f ms" % (busy / 5 / 1e3))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
