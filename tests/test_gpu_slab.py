"""GPU tests of the slab (N > 1) kernels and protocol.

1. N slab handles on ONE GPU, one thread per rank, product SlabStepper with an in-process ring (tests/util.ThreadRing):
   halo pack/unpack kernels, J guard-row add, leaver records written by the push kernel, arrival absorb + merged sort.
   Bit-exact versus the single-rank CPU oracle and versus the single-slab GPU path.
2. (needs >= 2 GPUs, skipped otherwise) the same through torchrun + NCCL, one process per GPU.
"""
import importlib
import os
import subprocess
import sys
import threading

import numpy as np
import pytest
from oracle import bindings as ob
from util import pkg, plasma, by_tag, copy_parts, PKG, ROOT, ThreadHub, ThreadRing

pytestmark = pytest.mark.gpu


def _problem(nx, ny, tn, ppc, u_th, seed=11):
    dx = dy = 0.1
    dt = pkg().courant_dt(dx, dy)
    parts = plasma().load_uniform(nx, ny, tn, tn, dx, dy, ppc, u_th=u_th, seed=seed, two_stream=0.2)
    rng = np.random.default_rng(seed)
    planes = [0.02 * rng.standard_normal((ny, nx)) for _ in range(6)]
    return dx, dy, dt, parts, planes


@pytest.mark.parametrize("world,shape", [(2, (64, 64, 16, 6, 0.6)), (3, (48, 96, 8, 5, 0.9)), (4, (32, 64, 16, 4, 0.5))])
def test_slabs_on_one_gpu_bit_exact(world, shape):
    m = pkg()
    slab = importlib.import_module(PKG + ".slab")
    nx, ny, tn, ppc, u_th = shape
    dx, dy, dt, parts, planes = _problem(nx, ny, tn, ppc, u_th)
    steps = 8
    hub = ThreadHub(world)
    sims, errs = [None] * world, []

    def run(rank):
        try:
            s = m.Sim(nx, ny, tn, tn, dx, dy, dt, max_particles=parts["x"].shape[0], rank=rank, nranks=world)
            sims[rank] = s
            r0, r1 = s.row0 * tn, (s.row0 + s.nrows) * tn
            s.upload_planes([a[r0:r1] for a in planes])
            s.upload_particles(parts)                      # the device keeps only the particles of its slab
            st = slab.SlabStepper(s, ThreadRing(rank, hub))
            st.refresh_guards()
            for _ in range(steps):
                st.step()
            s.sync()
            s.moved = st.moved
        except Exception as e:                              # noqa: BLE001
            errs.append((rank, repr(e)))

    th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join(300) for t in th]
    assert not errs, errs
    lib = ob.load_oracle()
    P = ob.make_params(lib, nx, ny, tn, tn, dx, dy, dt)
    want, wf = copy_parts(parts), [a.copy() for a in planes]
    counts, _ = ob.pic_steps(lib, P, wf, want, steps)
    got = by_tag({k: np.concatenate([s.download_particles()[k] for s in sims]) for k in ("x", "y", "ux", "uy", "uz", "q", "tag")})
    want = by_tag(want)
    assert np.array_equal(got["tag"], want["tag"])
    for k in ("x", "y", "ux", "uy", "uz", "q"):
        assert np.array_equal(got[k], want[k]), k
    gf = [np.concatenate([s.download_planes()[k] for s in sims]) for k in range(6)]
    for k in range(6):
        assert np.array_equal(gf[k], wf[k]), k
    assert np.array_equal(np.concatenate([s.tile_counts() for s in sims]), counts)
    assert sum(s.moved for s in sims) > 0
    assert sum(s.num_particles for s in sims) == parts["x"].shape[0]
    # tile patches WITH guards (what a dump holds) agree with the single-slab device path
    one = m.Sim(nx, ny, tn, tn, dx, dy, dt, max_particles=parts["x"].shape[0])
    one.upload_planes(planes)
    one.upload_particles(parts)
    one.pic_steps(steps)
    gtc = nx // tn
    for s in sims:
        gids = np.arange(s.row0 * gtc, (s.row0 + s.nrows) * gtc, dtype=np.int32)
        assert np.array_equal(s.download_tiles(gids), one.download_tiles(gids))


def test_field_only_slabs_match_reference_loop_fixture():
    """Yee loop on 2 slabs == the reference's own tile loop (fixture from oracle/_ref/libref.so)."""
    m = pkg()
    slab = importlib.import_module(PKG + ".slab")
    f = dict(np.load(os.path.join(ROOT, "tests", "golden", "ref_current.npz")))     # materialise: NpzFile is not thread-safe
    nx, ny, R, C, tr, tc, g, steps = [int(v) for v in f["meta"]]
    tnx, tny = nx // (C * tc), ny // (R * tr)
    world = 2
    hub = ThreadHub(world)
    outs, errs = [None] * world, []

    def run(rank):
        try:
            s = m.Sim(nx, ny, tnx, tny, float(f["dx"]), float(f["dy"]), float(f["dt"]), rank=rank, nranks=world)
            r0, r1 = s.row0 * tny, (s.row0 + s.nrows) * tny
            s.upload_planes([f["ic"][r0:r1, :, k] for k in range(6)])
            st = slab.SlabStepper(s, ThreadRing(rank, hub))
            st.refresh_guards()
            for _ in range(steps):
                st.field_step()
            gtc = nx // tnx
            gids = np.arange(s.row0 * gtc, (s.row0 + s.nrows) * gtc, dtype=np.int32)
            outs[rank] = (s.download_planes(), gids, s.download_tiles(gids))
        except Exception as e:                              # noqa: BLE001
            errs.append((rank, repr(e)))

    th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join(300) for t in th]
    assert not errs, errs
    for k in range(6):
        assert np.array_equal(np.concatenate([o[0][k] for o in outs]), f["out"][..., k]), k
    for _, gids, tiles in outs:
        assert np.array_equal(tiles, f["tiles"][gids])


@pytest.mark.parametrize("world,shape", [(2, (64, 64, 16, 6, 0.6)), (3, (48, 96, 8, 5, 0.9)), (5, (32, 80, 16, 3, 0.7))])
def test_single_process_peer_copy_ring_bit_exact(world, shape):
    """csrc/ring.cu (the C++ driver's multi-GPU path): N handles, one host thread, peer copies ordered by events.
    Uses every visible GPU round-robin (all handles share the device on a one-GPU box)."""
    import torch
    m = pkg()
    nx, ny, tn, ppc, u_th = shape
    dx, dy, dt, parts, planes = _problem(nx, ny, tn, ppc, u_th, seed=17)
    ndev = max(1, torch.cuda.device_count())
    sims = [m.Sim(nx, ny, tn, tn, dx, dy, dt, max_particles=parts["x"].shape[0], rank=r, nranks=world, device=r % ndev)
            for r in range(world)]
    for s in sims:
        r0, r1 = s.row0 * tn, (s.row0 + s.nrows) * tn
        s.upload_planes([a[r0:r1] for a in planes])
        s.upload_particles(parts)
    ring = m.SlabRing(sims)
    ring.refresh_guards()
    steps = 9
    ring.pic_steps(steps)
    ring.sync()
    lib = ob.load_oracle()
    P = ob.make_params(lib, nx, ny, tn, tn, dx, dy, dt)
    want, wf = copy_parts(parts), [a.copy() for a in planes]
    counts, _ = ob.pic_steps(lib, P, wf, want, steps)
    got = by_tag({k: np.concatenate([s.download_particles()[k] for s in sims]) for k in ("x", "y", "ux", "uy", "uz", "q", "tag")})
    want = by_tag(want)
    assert np.array_equal(got["tag"], want["tag"])
    for k in ("x", "y", "ux", "uy", "uz", "q"):
        assert np.array_equal(got[k], want[k]), k
    for k in range(6):
        assert np.array_equal(np.concatenate([s.download_planes()[k] for s in sims]), wf[k]), k
    assert np.array_equal(np.concatenate([s.tile_counts() for s in sims]), counts)
    # field-only ring == single-slab field loop
    one = m.Sim(nx, ny, tn, tn, dx, dy, dt)
    one.upload_planes(wf)
    one.field_steps(5)
    ring.field_steps(5)
    ring.sync()
    ref = one.download_planes()
    for k in range(6):
        assert np.array_equal(np.concatenate([s.download_planes()[k] for s in sims]), ref[k]), k


@pytest.mark.parametrize("world,shape", [(2, (64, 64, 16, 6, 0.6)), (3, (48, 96, 8, 5, 0.9)), (4, (32, 64, 16, 4, 0.5))])
def test_peer_memory_ring_bit_exact(world, shape):
    """csrc/peer.cu: halos stored by the Yee sweeps straight into the neighbours' guard rows, current rows added with remote
    atomics, leaver records shipped into the neighbours' arrival buffers, phases ordered by flags.  N handles driven by one
    host thread (begin on all, end on all), spread over every visible GPU (all on one device on a one-GPU box).
    Bit-exact versus the single-rank CPU oracle, fields, particles and tile occupancy."""
    import torch
    m = pkg()
    nx, ny, tn, ppc, u_th = shape
    dx, dy, dt, parts, planes = _problem(nx, ny, tn, ppc, u_th, seed=23)
    ndev = max(1, torch.cuda.device_count())
    sims = [m.Sim(nx, ny, tn, tn, dx, dy, dt, max_particles=parts["x"].shape[0], rank=r, nranks=world, device=r % ndev)
            for r in range(world)]
    for s in sims:
        r0, r1 = s.row0 * tn, (s.row0 + s.nrows) * tn
        s.upload_planes([a[r0:r1] for a in planes])
        s.upload_particles(parts)
    ring = m.PeerRing(sims)
    ring.refresh_guards()
    steps = 9
    ring.pic_steps(steps)
    ring.sync()
    lib = ob.load_oracle()
    P = ob.make_params(lib, nx, ny, tn, tn, dx, dy, dt)
    want, wf = copy_parts(parts), [a.copy() for a in planes]
    counts, _ = ob.pic_steps(lib, P, wf, want, steps)
    got = by_tag({k: np.concatenate([s.download_particles()[k] for s in sims]) for k in ("x", "y", "ux", "uy", "uz", "q", "tag")})
    want = by_tag(want)
    assert np.array_equal(got["tag"], want["tag"])
    for k in ("x", "y", "ux", "uy", "uz", "q"):
        assert np.array_equal(got[k], want[k]), k
    for k in range(6):
        assert np.array_equal(np.concatenate([s.download_planes()[k] for s in sims]), wf[k]), k
    assert np.array_equal(np.concatenate([s.tile_counts() for s in sims]), counts)
    # guard rows too (what a dump holds): tile patches WITH guards == single-slab device path
    one = m.Sim(nx, ny, tn, tn, dx, dy, dt)
    one.upload_planes(wf)
    gtc = nx // tn
    for s in sims:
        gids = np.arange(s.row0 * gtc, (s.row0 + s.nrows) * gtc, dtype=np.int32)
        assert np.array_equal(s.download_tiles(gids), one.download_tiles(gids))
    # field-only ring == single-slab field loop
    one.field_steps(5)
    ring.field_steps(5)
    ring.sync()
    ref = one.download_planes()
    for k in range(6):
        assert np.array_equal(np.concatenate([s.download_planes()[k] for s in sims]), ref[k]), k


def _torchrun_slabs(transport, port, marker):
    env = dict(os.environ, MPIC_TEST_TRANSPORT=transport)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "run_slab_nccl.py")],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert marker in r.stdout


def test_peer_memory_ring_two_processes_cuda_ipc():
    """The REAL multi-process path: two torchrun processes, CUDA-IPC mapping of the neighbour's planes / accumulators /
    arrival buffers, .sys-scope flags between processes.  Two GPUs when the box has them, else both ranks on cuda:0 (IPC
    works between processes on one device).  Bit-exact versus the single-rank oracle, migration included."""
    _torchrun_slabs("peer", 29613, "SLAB_PEER_PARITY_OK")


def test_nccl_slabs_two_processes():
    """SlabStepper between two processes: over NCCL with one GPU per rank; on a one-GPU box the same protocol over gloo with
    host staging (NCCL cannot put two ranks on one device) -- pack / unpack / J add / migration kernels are the same."""
    import torch
    if torch.cuda.device_count() >= 2:
        _torchrun_slabs("nccl", 29611, "SLAB_NCCL_PARITY_OK")
    else:
        _torchrun_slabs("staged", 29611, "SLAB_STAGED_PARITY_OK")


def _rebalance_gpu_worker(rank, world, port, steps, outdir):
    import torch
    import torch.distributed as dist
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    from test_slab_gloo import _beam_problem
    m = importlib.import_module(PKG)
    slab = importlib.import_module(PKG + ".slab")
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dev = rank % max(1, torch.cuda.device_count())
    try:
        g, parts, planes = _beam_problem()
        cap = parts["x"].shape[0]
        make = lambda r0, nr, n=0: m.Sim(g["nx"], g["ny"], g["tnx"], g["tny"], g["dx"], g["dy"], g["dt"], max_particles=cap,
                                          rank=rank, nranks=world, device=dev, rows=(r0, nr))
        sim = m.Sim(g["nx"], g["ny"], g["tnx"], g["tny"], g["dx"], g["dy"], g["dt"], max_particles=cap, rank=rank, nranks=world, device=dev)
        r0, r1 = sim.row0 * g["tny"], (sim.row0 + sim.nrows) * g["tny"]
        sim.upload_planes([a[r0:r1] for a in planes])
        sim.upload_particles(parts)
        st = slab.PeerStepper(sim, rank, world, dist)
        st.refresh_guards()
        for _ in range(steps):
            st.step()
        sim.sync()
        before = sim.num_particles
        handle = sim.h.value
        rows, changed = slab.rebalance_device(sim, rank, world, dist, w_cell=0.25)
        assert sim.h.value == handle and (sim.row0, sim.nrows) == tuple(rows[rank])     # same handle, new rows, in place
        own = sim.owner().reshape(-1, sim.gtc)
        for r, (r0, nr) in enumerate(rows):
            assert (own[r0:r0 + nr] == r).all()
        st = slab.PeerStepper(sim, rank, world, dist)
        st.refresh_guards()
        for _ in range(steps):
            st.step()
        sim.sync()
        dist.barrier()
        pl = sim.download_planes()
        np.savez(os.path.join(outdir, f"rank{rank}.npz"), row0=sim.row0 * g["tny"], changed=changed, before=before,
                 after=sim.num_particles, counts=sim.tile_counts(), **{f"f{k}": pl[k] for k in range(6)}, **sim.download_particles())
        dist.barrier()
        sim.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_runtime_rebalance_over_the_peer_ring(world):
    """One process per rank (CUDA IPC between them, all on one GPU when the box has one): peer-ring steps, cost-driven
    re-partition of tile rows with their fields and particles ON THE DEVICE through the C ABI (slab.rebalance_device:
    mpic_rebalance_plan, mpic_reslab_pack / _apply / _finish; parcels pulled through CUDA IPC, handles rebuilt in place),
    peer-ring steps on the new, uneven slabs.  Bit-exact versus the single-rank oracle; the most loaded rank gets lighter."""
    import tempfile
    import torch.multiprocessing as mp
    from test_slab_gloo import _beam_problem, _free_port
    steps = 3
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_rebalance_gpu_worker, args=(world, _free_port(), steps, d), nprocs=world, join=True)
        outs = [np.load(os.path.join(d, f"rank{r}.npz")) for r in range(world)]
    g, parts, planes = _beam_problem()
    lib = ob.load_oracle()
    P = ob.make_params(lib, g["nx"], g["ny"], g["tnx"], g["tny"], g["dx"], g["dy"], g["dt"])
    want, wf = copy_parts(parts), [a.copy() for a in planes]
    counts, _ = ob.pic_steps(lib, P, wf, want, 2 * steps)
    got = by_tag({k: np.concatenate([o[k] for o in outs]) for k in ("x", "y", "ux", "uy", "uz", "q", "tag")})
    want = by_tag(want)
    assert np.array_equal(got["tag"], want["tag"])
    for k in ("x", "y", "ux", "uy", "uz", "q"):
        assert np.array_equal(got[k], want[k]), k
    outs.sort(key=lambda o: int(o["row0"]))
    for k in range(6):
        assert np.array_equal(np.concatenate([o[f"f{k}"] for o in outs]), wf[k]), k
    assert np.array_equal(np.concatenate([o["counts"] for o in outs]), counts)
    assert all(bool(o["changed"]) for o in outs)
    assert max(int(o["after"]) for o in outs) < 0.8 * max(int(o["before"]) for o in outs)


@pytest.mark.parametrize("world,kind", [(2, "copy"), (3, "copy"), (4, "peer"), (3, "peer")])
def test_single_process_ring_rebalance_in_place(world, kind):
    """mpic_ring_rebalance (what host/pic2d --rebalance-every uses): steps, cost-driven re-partition with the handles rebuilt
    in place from device parcels, more steps on the uneven slabs -- for the peer-copy ring (csrc/ring.cu) and the peer-memory
    ring (csrc/peer.cu).  Bit-exact versus the single-rank oracle; a second call changes nothing (the plan is a fixed point)."""
    import torch
    from test_slab_gloo import _beam_problem
    m = pkg()
    g, parts, planes = _beam_problem()
    nx, ny, tn, dx, dy, dt = g["nx"], g["ny"], g["tnx"], g["dx"], g["dy"], g["dt"]
    ndev = max(1, torch.cuda.device_count())
    cap = parts["x"].shape[0]
    # capacities: own population + 10 % -- after the re-partition the heavy rank fits (its arrays are kept), the light ones grow
    cy = np.floor(parts["y"] * (1.0 / dy)).astype(np.int64)
    gtr = ny // g["tny"]
    own = []
    for r in range(world):
        r0, nr = m.slab_rows(gtr, world, r)
        own.append((cy >= r0 * g["tny"]) & (cy < (r0 + nr) * g["tny"]))
    sims = [m.Sim(nx, ny, tn, g["tny"], dx, dy, dt, max_particles=int(1.1 * own[r].sum()) + 64, rank=r, nranks=world, device=r % ndev)
            for r in range(world)]
    for r, s in enumerate(sims):
        r0, r1 = s.row0 * g["tny"], (s.row0 + s.nrows) * g["tny"]
        s.upload_planes([a[r0:r1] for a in planes])
        s.upload_particles({k: np.ascontiguousarray(v[own[r]]) for k, v in parts.items()})
    ring = m.SlabRing(sims) if kind == "copy" else m.PeerRing(sims)
    ring.refresh_guards()
    steps = 4
    ring.pic_steps(steps)
    ring.sync()
    before = [s.num_particles for s in sims]
    handles = [s.h.value for s in sims]
    changed, rows = ring.rebalance(w_cell=0.25)
    assert changed and [s.h.value for s in sims] == handles
    assert rows[0][0] == 0 and all(rows[k][0] + rows[k][1] == rows[k + 1][0] for k in range(world - 1)) and rows[-1][0] + rows[-1][1] == ny // g["tny"]
    assert max(s.num_particles for s in sims) < 0.85 * max(before) and sum(s.num_particles for s in sims) == sum(before)
    again, rows2 = ring.rebalance(w_cell=0.25)
    assert not again and rows2 == rows
    ring.pic_steps(steps)
    ring.sync()
    lib = ob.load_oracle()
    P = ob.make_params(lib, nx, ny, tn, g["tny"], dx, dy, dt)
    want, wf = copy_parts(parts), [a.copy() for a in planes]
    counts, _ = ob.pic_steps(lib, P, wf, want, 2 * steps)
    got = by_tag({k: np.concatenate([s.download_particles()[k] for s in sims]) for k in ("x", "y", "ux", "uy", "uz", "q", "tag")})
    want = by_tag(want)
    assert np.array_equal(got["tag"], want["tag"])
    for k in ("x", "y", "ux", "uy", "uz", "q"):
        assert np.array_equal(got[k], want[k]), k
    for k in range(6):
        assert np.array_equal(np.concatenate([s.download_planes()[k] for s in sims]), wf[k]), k
    assert np.array_equal(np.concatenate([s.tile_counts() for s in sims]), counts)
    own = sims[0].owner().reshape(-1, sims[0].gtc)
    for r, (r0, nr) in enumerate(rows):
        assert (own[r0:r0 + nr] == r).all()


def test_reslab_rejects_bad_parcels():
    """mpic_reslab_apply validates before it frees anything: parcels must tile the new rows exactly."""
    m = pkg()
    nx, ny, tn, ppc, u_th = 32, 64, 16, 2, 0.3
    dx, dy, dt, parts, planes = _problem(nx, ny, tn, ppc, u_th, seed=5)
    sims = [m.Sim(nx, ny, tn, tn, dx, dy, dt, max_particles=parts["x"].shape[0], rank=r, nranks=2, device=0) for r in range(2)]
    for s in sims:
        r0, r1 = s.row0 * tn, (s.row0 + s.nrows) * tn
        s.upload_planes([a[r0:r1] for a in planes])
        s.upload_particles(parts)
    a = sims[0].reslab_pack(0, 1)
    with pytest.raises(m.MpicError):
        sims[0].reslab_pack(1, 3)                        # row 2 is not on slab 0
    with pytest.raises(m.MpicError):
        sims[0].reslab_apply(0, 2, [a])                  # rows [0,2) are not covered by one parcel of row 0
    with pytest.raises(m.MpicError):
        sims[0].reslab_apply(0, 1, [a, a])               # overlap
    assert sims[0].num_particles > 0                     # the handle survived the rejected calls
    got = sims[0].download_planes()
    for k in range(6):
        assert np.array_equal(got[k], planes[k][:sims[0].nyl])
    sims[0].reslab_finish()


def test_blob_with_our_pid_but_a_foreign_nonce_is_not_trusted():
    """'same process' was decided from the pid alone, and two ranks in separate PID namespaces (one container per
    rank) can share a pid -- the raw device pointers of ANOTHER process would have been dereferenced.  Blobs now carry a random
    per-process value as well.  A blob of this process whose nonce is altered must go down the CUDA-IPC road (which refuses to
    open a handle inside the process that exported it) instead of being trusted; the untouched blob still connects."""
    m = pkg()
    nx, ny, tn = 32, 64, 16
    dx = dy = 0.1
    dt = m.courant_dt(dx, dy)
    sims = [m.Sim(nx, ny, tn, tn, dx, dy, dt, rank=r, nranks=2, device=0) for r in range(2)]
    blobs = [s.peer_export() for s in sims]
    NONCE = slice(32, 40)                                # PeerBlob: magic (8 bytes), six ints, then the nonce
    assert blobs[0][NONCE] == blobs[1][NONCE] and blobs[1][NONCE] != bytes(8)
    forged = bytearray(blobs[1])
    forged[32] ^= 0x5A
    with pytest.raises(m.MpicError):
        sims[0].peer_connect(bytes(forged), bytes(forged))
    sims[0].peer_connect(blobs[1], blobs[1])
    sims[1].peer_connect(blobs[0], blobs[0])
    for s in sims:
        s.peer_disconnect()
        s.close()


def test_peer_ring_without_the_fused_plane_block(monkeypatch):
    """mpic_peer_export used to allocate the fused step's second plane block (six more planes) for every ring and a
    failure there made the plain ring unusable.  With MPIC_PEER_NO_FUSED=1 -- the same state a failed set-up now leaves -- no
    second block is exported: the three-sweep ring still equals the single-slab loop bit for bit and the fused step says why
    it cannot run."""
    m = pkg()
    monkeypatch.setenv("MPIC_PEER_NO_FUSED", "1")
    nx, ny, tn = 64, 64, 16
    dx, dy, dt, parts, planes = _problem(nx, ny, tn, 2, 0.3, seed=31)
    sims = [m.Sim(nx, ny, tn, tn, dx, dy, dt, rank=r, nranks=2, device=0) for r in range(2)]
    for s in sims:
        r0, r1 = s.row0 * tn, (s.row0 + s.nrows) * tn
        s.upload_planes([a[r0:r1] for a in planes])
    ring = m.PeerRing(sims)
    ring.refresh_guards()
    ring.field_steps(7)
    ring.sync()
    one = m.Sim(nx, ny, tn, tn, dx, dy, dt)
    one.upload_planes(planes)
    one.field_steps(7)
    ref = one.download_planes()
    for k in range(6):
        assert np.array_equal(np.concatenate([s.download_planes()[k] for s in sims]), ref[k]), k
    with pytest.raises(m.MpicError, match="no second plane block"):
        sims[0].peer_field_step(fused=True)

