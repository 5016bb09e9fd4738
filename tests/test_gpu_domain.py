"""GPU, 2 and 4 ranks over NCCL (skipped where the box has fewer GPUs): the slab-decomposed session against the
CPU oracle's single-process run with particles that MOVE between steps (so that migration really takes the
all_to_all branch and the halo is rebuilt from scratch) — neighbour sets bit-exact, state within 1e-10
(SURVEY.md §8e)."""
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
        amp = 0.2 * float(np.median(ref["sizes"]))
        ids = np.arange(len(st["points"]))
        for it in range(nsteps):
            ds.perturb(amp, 40 + it)
            ref["points"] = ref["points"] + ics.perturbation(ids, amp, 40 + it)
            ds.step(); ds.check()
            ref = O.step(ref)
            assert ds.scalar("ct") == pytest.approx(ref["ct"], rel=1e-13), it
            assert ds.scalar("dt") == pytest.approx(ref["dt"], rel=1e-13), it
            assert ds.scalar("d") == pytest.approx(ref["d"], rel=1e-12), it
            for name in ("points", "velocities", "total_accel", "E_internal", "T", "sizes", "densities", "densities_0",
                         "visc_heat", "pressure_accel", "visc_accel"):
                assert rel(ds.gather(name), ref[name]) < RTOL, (it, name, rel(ds.gather(name), ref[name]))
            # neighbour rows of every particle, as sets of global ids
            got = np.sort(ds.gather_neighbor_ids(), axis=1)
            assert np.array_equal(got, np.sort(ref["neighbor"], axis=1)), it
        stats = ds.counters()
        tot = torch.tensor([float(stats["migrated"])], device=torch.device("cuda", rank), dtype=torch.float64)
        dist.all_reduce(tot)
        assert tot.item() > 0, "no particle migrated: the test does not exercise the exchange"
        assert ds.n_own < len(st["points"]) and ds.sess.n > ds.n_own        # really decomposed, with ghosts
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_slabs_match_single_process_oracle(built_lib, world):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(world, _free_port(), 20000, 3), nprocs=world, join=True)
