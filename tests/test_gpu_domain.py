"""GPU, 2 ranks over NCCL (skipped on a single-GPU box): the slab-decomposed session against the CPU
oracle's single-process run — neighbour sets bit-exact, state within 1e-10."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
RTOL = 1e-10


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, n, nsteps):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from oracle import sph_oracle as O
        from sphb200 import ics
        from sphb200.domain import DistributedSession
        st = ics.uniform_box(n, 0, n_dust=3)
        st["dt_0"] = O.dt_0 * 1e-6
        ds = DistributedSession.from_global_state(st, rank, dt_0=st["dt_0"])
        ds.startup()
        ref = O.startup(st)

        def rel(a, b):
            return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))

        assert np.array_equal(ds.gather("sizes"), ref["sizes"])            # k-th distances bit-exact
        assert rel(ds.gather("densities"), ref["densities"]) < RTOL
        for it in range(nsteps):
            ds.step(); ds.check()
            ref = O.step(ref)
            assert ds.scalar("ct") == pytest.approx(ref["ct"], rel=1e-13), it
            assert ds.scalar("dt") == pytest.approx(ref["dt"], rel=1e-13), it
            assert ds.scalar("d") == pytest.approx(ref["d"], rel=1e-12), it
            for name in ("points", "velocities", "total_accel", "E_internal", "T", "sizes", "densities", "densities_0",
                         "visc_heat", "pressure_accel", "visc_accel"):
                assert rel(ds.gather(name), ref[name]) < RTOL, (it, name, rel(ds.gather(name), ref[name]))
        # neighbour rows of the owned particles, as sets of global ids
        gid, _ = ds.download_owned("sizes")
        nb = ds.sess.download_tensor("neighbor")[:ds.n_own].cpu().numpy()
        g_all = ds.gid.cpu().numpy().astype(np.int64)
        got = np.sort(g_all[nb], axis=1)
        want = np.sort(ref["neighbor"][gid.cpu().numpy()], axis=1)
        assert np.array_equal(got, want)
        assert ds.n_own < n and ds.sess.n > ds.n_own                        # really decomposed, with ghosts
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_slabs_match_single_process_oracle(built_lib):
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), 20000, 3), nprocs=2, join=True)
