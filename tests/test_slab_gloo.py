"""N > 1 path on CPU: the product's slab orchestration (slab.SlabStepper / slab.Ring) over torch.distributed with the
gloo backend, world_size 2 and 3, with the per-rank compute supplied by the CPU oracle stand-in (tests/cpu_slab.py).
Bar: after K full PIC steps the union of the ranks' particles and field rows is BIT-IDENTICAL to the single-rank oracle
run (partition independence: SURVEY.md F8 for the fields, fixed-point current for the deposit)."""
import importlib
import os
import socket
import sys
import tempfile

import numpy as np
import pytest
import torch.multiprocessing as mp

from util import PKG, ROOT, plasma, by_tag, copy_parts


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _problem(seed=5, nx=32, ny=48, tnx=8, tny=8, ppc=6, u_th=0.7):
    dx, dy = 0.1, 0.125
    dt = 0.5 / np.sqrt(1 / dx ** 2 + 1 / dy ** 2)
    parts = plasma().load_uniform(nx, ny, tnx, tny, dx, dy, ppc, u_th=u_th, seed=seed, two_stream=0.3)
    rng = np.random.default_rng(seed)
    planes = [0.02 * rng.standard_normal((ny, nx)) for _ in range(6)]
    return dict(nx=nx, ny=ny, tnx=tnx, tny=tny, dx=dx, dy=dy, dt=dt), parts, planes


def _worker(rank, world, port, steps, rows, outdir):
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist
    from cpu_slab import CpuSlabSim
    slab = importlib.import_module(PKG + ".slab")
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g, parts, planes = _problem()
        row0, nrows = rows[rank]
        sim = CpuSlabSim(g["nx"], g["ny"], g["tnx"], g["tny"], g["dx"], g["dy"], g["dt"], rank, world, row0, nrows)
        sim.set_interior(planes)
        sim.set_particles(parts)
        st = slab.SlabStepper(sim, slab.Ring(rank, world))
        st.refresh_guards()
        for _ in range(steps):
            st.step()
        np.savez(os.path.join(outdir, f"rank{rank}.npz"), rows=sim.rows_int, moved=st.moved,
                 **{f"f{k}": sim.planes[k][sim.rows_int] for k in range(6)}, **sim.parts)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,rows", [(2, [(0, 3), (3, 3)]), (3, [(0, 1), (1, 3), (4, 2)])])
def test_slab_ring_matches_single_rank_oracle(world, rows):
    from oracle import bindings as ob
    steps = 6
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(world, _free_port(), steps, rows, d), nprocs=world, join=True)
        outs = [np.load(os.path.join(d, f"rank{r}.npz")) for r in range(world)]
        g, parts, planes = _problem()
        lib = ob.load_oracle()
        P = ob.make_params(lib, g["nx"], g["ny"], g["tnx"], g["tny"], g["dx"], g["dy"], g["dt"])
        want, wf = copy_parts(parts), [a.copy() for a in planes]
        ob.pic_steps(lib, P, wf, want, steps)
        got = by_tag({k: np.concatenate([o[k] for o in outs]) for k in ("x", "y", "ux", "uy", "uz", "q", "tag")})
        want = by_tag(want)
        assert got["tag"].shape == want["tag"].shape and np.array_equal(got["tag"], want["tag"])   # nobody lost or duplicated
        for k in ("x", "y", "ux", "uy", "uz", "q"):
            assert np.array_equal(got[k], want[k]), k
        for o in outs:
            for k in range(6):
                assert np.array_equal(o[f"f{k}"], wf[k][o["rows"]]), k
        assert sum(int(o["moved"]) for o in outs) > 0          # the hand-off path really ran
        # every particle sits on the rank that owns its cell row
        for r, o in enumerate(outs):
            cy = np.floor(o["y"] * (1.0 / g["dy"])).astype(int)
            assert ((cy >= o["rows"][0]) & (cy <= o["rows"][-1])).all()


def _beam_problem():
    """thin dense band (6x the background) in the lowest tile rows: an even split is badly imbalanced"""
    g, parts, planes = _problem(seed=9, ny=64, ppc=3)
    extra = plasma().load_uniform(g["nx"], g["ny"], g["tnx"], g["tny"], g["dx"], g["dy"], 18, u_th=0.4, seed=10, drift=(0.3, 0.0, 0.0))
    band = extra["y"] < 12 * g["dy"]
    extra = {k: v[band] for k, v in extra.items()}
    extra["tag"] = (extra["tag"].astype(np.int64) + 10 ** 6).astype(np.int32)
    return g, {k: np.concatenate([parts[k], extra[k]]) for k in parts}, planes


def _rebalance_worker(rank, world, port, steps, outdir):
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist
    from cpu_slab import CpuSlabSim
    slab = importlib.import_module(PKG + ".slab")
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g, parts, planes = _beam_problem()
        gtr = g["ny"] // g["tny"]
        make = lambda r0, nr, n=0: CpuSlabSim(g["nx"], g["ny"], g["tnx"], g["tny"], g["dx"], g["dy"], g["dt"], rank, world, r0, nr)
        base, rem = divmod(gtr, world)
        sim = make(rank * base + min(rank, rem), base + (rank < rem))
        sim.set_interior(planes)
        sim.set_particles(parts)
        st = slab.SlabStepper(sim, slab.Ring(rank, world))
        st.refresh_guards()
        for _ in range(steps):
            st.step()
        before = int(sim.parts["x"].shape[0])
        sim, rows, changed = slab.rebalance(sim, rank, world, make, dist, w_cell=0.25)
        st = slab.SlabStepper(sim, slab.Ring(rank, world))
        st.refresh_guards()
        for _ in range(steps):
            st.step()
        np.savez(os.path.join(outdir, f"rank{rank}.npz"), rows=sim.rows_int, changed=changed, before=before,
                 after=int(sim.parts["x"].shape[0]), new_rows=np.array(rows),
                 **{f"f{k}": sim.planes[k][sim.rows_int] for k in range(6)}, **sim.parts)
    finally:
        dist.destroy_process_group()


def _rebalance_device_worker(rank, world, port, steps, outdir):
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist
    from cpu_slab import CpuSlabSim
    slab = importlib.import_module(PKG + ".slab")
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g, parts, planes = _beam_problem()
        gtr = g["ny"] // g["tny"]
        base, rem = divmod(gtr, world)
        sim = CpuSlabSim(g["nx"], g["ny"], g["tnx"], g["tny"], g["dx"], g["dy"], g["dt"], rank, world,
                         rank * base + min(rank, rem), base + (rank < rem))
        sim.set_interior(planes)
        sim.set_particles(parts)
        st = slab.SlabStepper(sim, slab.Ring(rank, world))
        st.refresh_guards()
        for _ in range(steps):
            st.step()
        before = int(sim.parts["x"].shape[0])
        ident = id(sim)
        timings = {}
        rows, changed = slab.rebalance_device(sim, rank, world, dist, w_cell=0.25, timings=timings)
        assert id(sim) == ident and (sim.row0, sim.nrows) == tuple(rows[rank])          # same handle, new rows
        rows2, again = slab.rebalance_device(sim, rank, world, dist, w_cell=0.25)        # the plan is a fixed point
        assert not again and rows2 == rows
        st = slab.SlabStepper(sim, slab.Ring(rank, world))
        st.refresh_guards()
        for _ in range(steps):
            st.step()
        np.savez(os.path.join(outdir, f"rank{rank}.npz"), rows=sim.rows_int, changed=changed, before=before,
                 after=int(sim.parts["x"].shape[0]), new_rows=np.array(rows), stages=np.array(sorted(timings)),
                 **{f"f{k}": sim.planes[k][sim.rows_int] for k in range(6)}, **sim.parts)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_device_rebalance_protocol_over_gloo(world):
    """The PRODUCT's collective re-partition (slab.rebalance_device: occupancy all-gather -> mpic_rebalance_plan in the C
    library -> one parcel per new owner -> blob all-gather -> apply in place -> barrier -> finish) on world_size 2 and 3 over
    gloo, with the CPU stand-in playing the device: bit-exact versus the single-rank oracle through the re-partition, the
    most loaded rank gets lighter, every rank keeps its handle."""
    from oracle import bindings as ob
    steps = 3
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_rebalance_device_worker, args=(world, _free_port(), steps, d), nprocs=world, join=True)
        outs = [np.load(os.path.join(d, f"rank{r}.npz")) for r in range(world)]
        g, parts, planes = _beam_problem()
        lib = ob.load_oracle()
        P = ob.make_params(lib, g["nx"], g["ny"], g["tnx"], g["tny"], g["dx"], g["dy"], g["dt"])
        want, wf = copy_parts(parts), [a.copy() for a in planes]
        ob.pic_steps(lib, P, wf, want, 2 * steps)
        got = by_tag({k: np.concatenate([o[k] for o in outs]) for k in ("x", "y", "ux", "uy", "uz", "q", "tag")})
        want = by_tag(want)
        assert np.array_equal(got["tag"], want["tag"])
        for k in ("x", "y", "ux", "uy", "uz", "q"):
            assert np.array_equal(got[k], want[k]), k
        for o in outs:
            for k in range(6):
                assert np.array_equal(o[f"f{k}"], wf[k][o["rows"]]), k
        assert all(bool(o["changed"]) for o in outs)
        assert max(int(o["after"]) for o in outs) < 0.8 * max(int(o["before"]) for o in outs)
        assert list(outs[0]["stages"]) == ["apply_in_place", "barrier_free_parcels", "disconnect_blob_allgather",
                                           "occupancy_allgather_plan", "pack_parcels"]


@pytest.mark.parametrize("world", [2, 3])
def test_runtime_rebalance_is_bit_exact_and_evens_the_load(world):
    """slab.rebalance (cost-driven re-partition of tile rows with their fields AND particles) in the
    middle of a run: the result still equals the single-rank oracle bit for bit, and the most loaded rank gets lighter."""
    from oracle import bindings as ob
    steps = 3
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_rebalance_worker, args=(world, _free_port(), steps, d), nprocs=world, join=True)
        outs = [np.load(os.path.join(d, f"rank{r}.npz")) for r in range(world)]
        g, parts, planes = _beam_problem()
        lib = ob.load_oracle()
        P = ob.make_params(lib, g["nx"], g["ny"], g["tnx"], g["tny"], g["dx"], g["dy"], g["dt"])
        want, wf = copy_parts(parts), [a.copy() for a in planes]
        ob.pic_steps(lib, P, wf, want, 2 * steps)
        got = by_tag({k: np.concatenate([o[k] for o in outs]) for k in ("x", "y", "ux", "uy", "uz", "q", "tag")})
        want = by_tag(want)
        assert np.array_equal(got["tag"], want["tag"])
        for k in ("x", "y", "ux", "uy", "uz", "q"):
            assert np.array_equal(got[k], want[k]), k
        for o in outs:
            for k in range(6):
                assert np.array_equal(o[f"f{k}"], wf[k][o["rows"]]), k
        assert all(bool(o["changed"]) for o in outs)
        assert max(int(o["after"]) for o in outs) < 0.8 * max(int(o["before"]) for o in outs)
        rows = outs[0]["new_rows"]
        assert rows[0][0] == 0 and all(rows[i][0] + rows[i][1] == rows[i + 1][0] for i in range(world - 1))


def test_balanced_slab_rows_policy():
    slab = importlib.import_module(PKG + ".slab")
    # uniform load -> even split
    assert slab.balanced_slab_rows(np.ones(16), 4) == [(0, 4), (4, 4), (8, 4), (12, 4)]
    # beam: 55 % of the load in 10 % of the rows -> the hot rows get narrow slabs
    c = np.full(40, 8.0)
    c[18:22] = 88.0
    parts = slab.balanced_slab_rows(c, 8)
    assert len(parts) == 8 and parts[0][0] == 0 and sum(nr for _, nr in parts) == 40
    assert all(parts[k][0] + parts[k][1] == parts[k + 1][0] for k in range(7))
    load = [c[r0:r0 + nr].sum() for r0, nr in parts]
    even = [c[r0:r0 + 5].sum() for r0 in range(0, 40, 5)]
    assert max(load) < 0.6 * max(even)
    assert max(load) <= 1.25 * c.sum() / 8 + c.max()
    # more ranks than heavy rows, every rank still owns a row
    assert all(nr >= 1 for _, nr in slab.balanced_slab_rows(np.array([100.0, 1, 1, 1]), 4))
    costs = slab.tile_row_costs(np.arange(12), 3, 64)
    assert costs.shape == (4,) and costs[1] > costs[0]


def test_policy_reproduces_the_measured_beam_partition():
    """configs[4] at 8 ranks (4096^2 cells, 8 ppc background + 80 ppc beam in rows [0.45, 0.55) of the box): with the cost
    model calibrated on B200 (slab.W_CELL = 8 particle-equivalents per cell) the exact bottleneck partition is the one the
    N=8 bench line reports (profiles/r01e_bench_n8_beam2.json: 48,48,24,8,8,24,48,48 tile rows; 3.44 -> 2.12 ms/step)."""
    slab = importlib.import_module(PKG + ".slab")
    ny = nx = 4096
    tn = 16
    lo, hi = int(round(0.45 * ny)), int(round(0.55 * ny))
    rows = np.arange(ny)
    per_tile_row = ((8 + 80 * ((rows >= lo) & (rows < hi))) * nx).reshape(-1, tn).sum(axis=1)
    costs = per_tile_row + 8.0 * (nx // tn) * tn * tn
    parts = slab.balanced_slab_rows(costs, 8)
    assert [nr for _, nr in parts] == [48, 48, 24, 8, 8, 24, 48, 48]
    load = [costs[r0:r0 + nr].sum() for r0, nr in parts]
    even = [costs[r * 32:(r + 1) * 32].sum() for r in range(8)]
    assert max(load) / np.mean(load) < 1.01 and max(even) / np.mean(even) > 1.9


def test_partition_is_optimal_against_brute_force():
    """mpic_partition_rows (host C behind slab.balanced_slab_rows): on random small problems the bottleneck equals the minimum
    over ALL contiguous partitions with non-empty slabs; mpic_rebalance_plan keeps the old partition unless the gain is real."""
    import itertools
    m = importlib.import_module(PKG)
    rng = np.random.default_rng(12)
    for _ in range(200):
        n = int(rng.integers(2, 11))
        k = int(rng.integers(1, min(n, 5) + 1))
        c = rng.integers(0, 50, n).astype(np.float64) + (rng.random(n) < 0.2) * 500.0
        parts = m.partition_rows(c, k)
        assert len(parts) == k and parts[0][0] == 0 and parts[-1][0] + parts[-1][1] == n
        assert all(nr >= 1 for _, nr in parts) and all(parts[i][0] + parts[i][1] == parts[i + 1][0] for i in range(k - 1))
        got = max(c[r0:r0 + nr].sum() for r0, nr in parts)
        best = min(max(c[a:b].sum() for a, b in zip((0,) + cut, cut + (n,))) for cut in itertools.combinations(range(1, n), k - 1))
        assert got == best, (c, k, parts)
    with pytest.raises(m.MpicError):
        m.partition_rows(np.ones(3), 4)
    # plan: occupancy -> cost = particles + w_cell * cells; a balanced slab set stays, an imbalanced one moves
    even = [(0, 4), (4, 4), (8, 4), (12, 4)]
    changed, rows, cost = m.rebalance_plan(np.full(16, 1000), 256, even, 8.0)
    assert not changed and rows == even and cost[0] == cost[1]
    occ = np.full(16, 100)
    occ[:4] = 50000
    changed, rows, cost = m.rebalance_plan(occ, 256, even, 8.0)
    assert changed and rows[0][1] < 4 and cost[1] < 0.95 * cost[0]
    changed2, rows2, _ = m.rebalance_plan(occ, 256, rows, 8.0)
    assert not changed2 and rows2 == rows                       # the plan is a fixed point
    changed3, rows3, _ = m.rebalance_plan(occ, 256, even, 8.0, min_gain=0.99)
    assert not changed3 and rows3 == even                       # gain threshold
