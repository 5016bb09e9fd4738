"""Where the slab driver's "exchange" phase goes (development aid): every sub-step bracketed by a device
synchronize.  torchrun --nproc-per-node 2 tools/exchange_times.py [n_per_rank]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def main():
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1000000
    rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    from sphb200 import domain
    ds = domain.DistributedSession.uniform_slabs(n, 1, 40, 1e-8, lr)
    ds.startup()
    for _ in range(3):
        ds.step()
    acc = {}

    def tick(name, fn):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        out = fn()
        torch.cuda.synchronize(); acc[name] = acc.get(name, 0.0) + (time.perf_counter() - t0) * 1e3
        return out

    steps = 10
    for _ in range(steps):
        for ph in ("crossing", "dt_clamp"):
            ds.phase(ph)
        rows = tick("owned_rows", ds._owned_rows)
        rows = tick("migrate", lambda: ds.ex.migrate(rows))
        if os.environ.get("HALO_DETAIL"):
            ex = ds.ex; w = 1.25 * ds.hmax; x = rows[:, 0]; ex.w = float(w)
            def sel():
                ex.send_idx = []
                for r in range(ex.world):
                    if r == ex.rank or not (ex.bounds[r] - w < ex.hi and ex.bounds[r + 1] + w > ex.lo):
                        ex.send_idx.append(torch.zeros(0, dtype=torch.int64, device=rows.device)); continue
                    m = (x >= ex.bounds[r] - w) & (x < ex.bounds[r + 1] + w)
                    ex.send_idx.append(torch.nonzero(m, as_tuple=False).flatten())
            tick("halo.select", sel)
            parts = tick("halo.gather", lambda: [rows[i] for i in ex.send_idx])
            send = tick("halo.cat", lambda: torch.cat(parts, dim=0).contiguous())
            sc = tick("halo.sc", lambda: torch.tensor([int(p.shape[0]) for p in parts], dtype=torch.int64, device=send.device))
            rc = torch.empty_like(sc)
            tick("halo.a2a_counts", lambda: dist.all_to_all_single(rc, sc))
            both = tick("halo.tolist", lambda: torch.cat([sc, rc]).tolist())
            sl, rcl = both[:ex.world], both[ex.world:]
            recv = tick("halo.empty", lambda: torch.empty((sum(rcl), send.shape[1]), dtype=send.dtype, device=send.device))
            tick("halo.a2a_rows", lambda: dist.all_to_all_single(recv, send, output_split_sizes=rcl, input_split_sizes=sl))
            ex.recv_counts = rcl; ghosts = recv
        else:
            ghosts = tick("build_halo", lambda: ds.ex.build_halo(rows, 1.25 * ds.hmax))
        local = tick("cat", lambda: torch.cat([rows, ghosts], dim=0).contiguous())
        ds.rows = rows; ds.n_own = int(rows.shape[0])
        tick("resize", lambda: ds.sess.resize(int(local.shape[0]), ds.n_own))
        tick("upload_rows", lambda: ds.sess.upload_rows(local))

        def rest():
            ds.gid = local[:, 18].contiguous()
            ds._send_cat = torch.cat(ds.ex.send_idx)
            ds._send_counts = [int(i.numel()) for i in ds.ex.send_idx]
            ds._n_ghost = int(ghosts.shape[0])
            ds._ghost_idx = torch.arange(ds.n_own, ds.n_own + ds._n_ghost, dtype=torch.int64, device=ds.device)
        tick("lists", rest)
        tick("search", ds._search)
        for ph in ("pack", "density", "pi", "force", "integrate"):
            ds.phase(ph)
    if os.environ.get("HALO_DETAIL"):
        print("rank %d: " % rank + "  ".join("%s=%.3f" % (k, v / steps) for k, v in acc.items()), flush=True)
    if rank == 0:
        print("ms per step (each sub-step synchronised): " + "  ".join("%s=%.3f" % (k, v / steps) for k, v in acc.items()))
        print("n_own=%d n_ghost=%d" % (ds.n_own, ds._n_ghost))
        for c in ("knn_worklist", "knn_widen", "knn_rescan"):
            try:
                print(c, ds.sess.ctx.counter(c))
            except Exception as e:
                print(c, "n/a", e)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
