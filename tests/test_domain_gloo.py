"""CPU, world_size 2 (and 3, 8 for the exchange protocol) over gloo: the slab decomposition's host logic (sphb200.domain) — migration, ghost
selection and refresh, the distributed percentile — and the whole exchange protocol with the CPU oracle
standing in for the CUDA passes: owned particles must get exactly the single-process neighbour rows
and densities ("results must be identical to the 1-GPU run for owned particles")."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

WORLD = 2


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(fn, *args, world=WORLD):
    port = _free_port()
    mp.spawn(_entry, args=(fn, port, world, args), nprocs=world, join=True)


def _entry(rank, fn, port, world, args):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["OMP_NUM_THREADS"] = "1"
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fn(rank, *args)
    finally:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
def _w_exchange(rank, w=0.31):
    from sphb200.domain import SlabExchange, NCOL
    world = dist.get_world_size()
    rs = np.random.RandomState(0)
    n = 5000
    rows = np.zeros((n, NCOL))
    rows[:, :3] = rs.rand(n, 3) * [4.0, 1.0, 1.0]
    rows[:, 18] = np.arange(n)
    rows[:, 10] = 2.0 * np.arange(n)                       # a field to refresh later
    bounds = [-np.inf, 1.7, np.inf] if world == 2 else [-np.inf] + list(np.linspace(0, 4, world + 1)[1:-1]) + [np.inf]
    start = rows[rs.permutation(n)[rank::world]]           # ownership unrelated to position
    ex = SlabExchange(bounds)
    own = ex.migrate(torch.from_numpy(start))
    x = rows[:, 0]
    mine = (x >= bounds[rank]) & (x < bounds[rank + 1])
    assert sorted(own[:, 18].numpy().astype(int)) == sorted(np.nonzero(mine)[0])
    again = ex.migrate(own)                                # nobody moves: every rank must take the same shortcut
    assert again.shape == own.shape and torch.equal(again, own)
    if rank == 0 and own.shape[0]:                         # one particle of one rank moves: the others have no traffic
        own = own.clone(); own[0, 0] = bounds[1] + 1e-3
    own2 = ex.migrate(own)
    moved_t = torch.tensor([int(own[0, 18]) if rank == 0 else -1]); dist.broadcast(moved_t, 0)
    moved = int(moved_t)
    x = x.copy(); x[moved] = bounds[1] + 1e-3
    mine = (x >= bounds[rank]) & (x < bounds[rank + 1])
    assert sorted(own2[:, 18].numpy().astype(int)) == sorted(np.nonzero(mine)[0])
    own = own2
    ghosts = ex.build_halo(own, w)
    want = (~mine) & (x >= bounds[rank] - w) & (x < bounds[rank + 1] + w)
    assert sorted(ghosts[:, 18].numpy().astype(int)) == sorted(np.nonzero(want)[0])
    # ghost refresh: owners change a field, ghosts receive the new values in ghost order
    own[:, 10] = own[:, 18] * 3.0 + 1.0
    new = ex.update_ghosts(own[:, 10])
    assert torch.equal(new, ghosts[:, 18] * 3.0 + 1.0)
    new3 = ex.update_ghosts(own[:, 0:3])
    assert torch.equal(new3, ghosts[:, 0:3])
    # coverage test: a ball reaching past the halo is flagged
    h = torch.full((own.shape[0],), 0.3, dtype=torch.float64)
    ok = ex.covered(own[:, 0], h)
    want_ok = torch.ones_like(ok)
    if rank > 0:
        want_ok &= own[:, 0] - 0.3 >= bounds[rank] - w
    if rank < world - 1:
        want_ok &= own[:, 0] + 0.3 < bounds[rank + 1] + w
    assert torch.equal(ok, want_ok)


def test_migrate_halo_refresh():
    _run(_w_exchange)


@pytest.mark.parametrize("world,w", [(3, 0.31), (8, 0.31), (8, 0.7)])
def test_migrate_halo_refresh_more_ranks(world, w):
    """more slabs than the GPU tests can afford: middle ranks with two neighbours, ranks that exchange nothing
    with most others, and (w = 0.7 > slab width 0.5) halos that reach across two slabs."""
    _run(_w_exchange, w, world=world)


def _w_percentile(rank):
    from sphb200.domain import distributed_percentile, distributed_order_statistic
    rs = np.random.RandomState(3)
    for n0, n1, scale in ((1000, 37, 1e34), (0, 5, 1.0), (2501, 2500, 1e-20), (1, 1, 7.0)):
        allv = np.abs(rs.normal(size=n0 + n1)) * scale
        if n0 + n1 > 4:
            allv[1] = allv[2]
        local = np.sort(allv[:n0] if rank == 0 else allv[n0:])
        v, ns = distributed_percentile(torch.from_numpy(local), 90.0)
        assert ns == n0 + n1
        assert v == np.percentile(allv, 90), (v, np.percentile(allv, 90))
        k = (n0 + n1) // 3
        assert distributed_order_statistic(torch.from_numpy(local), k) == np.sort(allv)[k]


def test_distributed_percentile_matches_numpy():
    _run(_w_percentile)


def _w_protocol(rank):
    """owned + ghosts with the oracle as the compute engine == the single-process oracle."""
    from oracle import sph_oracle as O
    from sphb200 import ics
    from sphb200.domain import SlabExchange, rows_from_state, equal_count_bounds, COLS
    st = ics.uniform_box(4000, 5)
    k = 40
    g_idx, _, _, _, g_h = O.neighbors(st["points"], st["dist"], k)
    g_out = O.hydro_update(g_idx, st["points"], st["mass"], g_h, None, st["particle_type"], st["T"], st["mu_array"],
                           st["gamma_array"], st["velocities"], species=False)
    rows = rows_from_state(st)
    world = dist.get_world_size()
    bounds = equal_count_bounds(rows[:, 0], world)
    ex = SlabExchange(bounds)
    own = ex.migrate(torch.from_numpy(rows[rank::world].copy()))
    assert abs(own.shape[0] - 4000 // world) <= 1
    w = 1.25 * float(g_h.max())
    ghosts = ex.build_halo(own, w)
    local = torch.cat([own, ghosts]).numpy()
    n_own = own.shape[0]
    gid = local[:, 18].astype(int)
    col = lambda name: local[:, COLS[name][0]:COLS[name][1]].reshape(len(local), -1).squeeze()
    idx, _, _, _, h = O.neighbors(col("points"), st["dist"], k)
    assert ex.covered(own[:, 0], torch.from_numpy(h[:n_own])).all()
    assert np.array_equal(gid[idx[:n_own]], g_idx[gid[:n_own]])          # same rows, same order
    assert np.array_equal(h[:n_own], g_h[gid[:n_own]])
    h[n_own:] = ex.update_ghosts(torch.from_numpy(h[:n_own].copy())).numpy()   # owners' h for the ghosts
    assert np.array_equal(h, g_h[gid])
    out = O.hydro_update(idx, col("points"), col("mass"), h, None, col("particle_type"), col("T"), col("mu_array"),
                         col("gamma_array"), col("velocities"), species=False)
    # density and pressure acceleration need h_j only: already exact for owned particles
    assert np.array_equal(out[3][:n_own], g_out[3][gid[:n_own]])
    assert np.array_equal(out[0][:n_own], g_out[0][gid[:n_own]])
    # crossing time: global minimum = min over ranks of the owned-only minimum
    vmax = lambda nb, v: np.max(np.sum((v[nb] - v[:, None, :]) ** 2, axis=2), axis=1) ** 0.5
    ct_loc = np.min(h[:n_own] / vmax(idx, col("velocities"))[:n_own])
    t = torch.tensor([ct_loc]); dist.all_reduce(t, op=dist.ReduceOp.MIN)
    assert float(t) + 0.0001 == O.crossing_time(g_idx, st["velocities"], g_h, st["particle_type"])


def test_exchange_protocol_reproduces_single_process_results():
    _run(_w_protocol)


def test_exchange_protocol_three_ranks():
    _run(_w_protocol, world=3)
