"""all_to_all_single with slab-like uneven splits vs batched isend/irecv (development aid)."""
import os, sys, time
import torch, torch.distributed as dist

def main():
    rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    dev = torch.device("cuda", lr)
    rows, ncol = 47000, 19
    counts = [rows if abs(r - rank) == 1 else 0 for r in range(world)]
    send = torch.randn(sum(counts), ncol, dtype=torch.float64, device=dev)
    recv = torch.empty_like(send)
    def a2a():
        dist.all_to_all_single(recv, send, output_split_sizes=counts, input_split_sizes=counts)
    def p2p():
        ops = []; o = 0
        for r in range(world):
            if counts[r]:
                ops.append(dist.P2POp(dist.isend, send[o:o + counts[r]], r))
                ops.append(dist.P2POp(dist.irecv, recv[o:o + counts[r]], r))
                o += counts[r]
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    small_s = torch.randn(sum(counts), 3, dtype=torch.float64, device=dev); small_r = torch.empty_like(small_s)
    def a2a_small():
        dist.all_to_all_single(small_r, small_s, output_split_sizes=counts, input_split_sizes=counts)
    for name, fn in (("a2a 7MB/neigh", a2a), ("isend/irecv 7MB/neigh", p2p), ("a2a 1.1MB/neigh", a2a_small)):
        for _ in range(5):
            fn()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(20):
            fn()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / 20 * 1e3
        if rank == 0:
            print("%-24s %.3f ms" % (name, dt))
    dist.destroy_process_group()

main()
