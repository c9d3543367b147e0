"""GPU parity tests for the particle path (push+gather, deposit, sort) against the authored CPU oracle
(oracle/pic_oracle.c; the reference has no particle code, SURVEY.md F1 -> parity vs the reference is 'unpinned').
Bar: BIT-EXACT for positions, momenta, fixed-point current, fields, tile counts and particle->tile indices
(stronger than north_star's 1e-10 rel-L2)."""
import numpy as np
import pytest
from oracle import bindings as ob
from util import pkg, plasma, by_tag, copy_parts

pytestmark = pytest.mark.gpu
lib = ob.load_oracle()


def _setup(nx, ny, tnx, tny, ppc, u_th, seed=3, two_stream=0.0, fields=0.05):
    dx, dy = 0.1, 0.1 * 1.25 if nx != ny else 0.1
    dt = pkg().courant_dt(dx, dy)
    P = ob.make_params(lib, nx, ny, tnx, tny, dx, dy, dt)
    parts = plasma().load_uniform(nx, ny, tnx, tny, dx, dy, ppc, u_th=u_th, seed=seed, two_stream=two_stream)
    rng = np.random.default_rng(seed)
    pl = [fields * rng.standard_normal((ny, nx)) for _ in range(6)]
    s = pkg().Sim(nx, ny, tnx, tny, dx, dy, dt, max_particles=parts["x"].shape[0] + 64)
    return s, P, parts, pl


@pytest.mark.parametrize("cfg", [(32, 32, 16, 16, 8, 0.3), (48, 30, 8, 6, 5, 0.8), (50, 25, 25, 25, 3, 0.5)])
def test_single_push_deposit_bit_exact(cfg):
    s, P, parts, pl = _setup(*cfg)
    s.upload_planes(pl)
    s.upload_particles(parts)
    s.push_deposit()
    s.fold_current()
    got = by_tag(s.download_particles())
    J = s.download_current()
    want = copy_parts(parts)
    Jw = ob.push_deposit(lib, P, pl, want)
    want = by_tag(want)
    for k in ("x", "y", "ux", "uy", "uz", "q"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(J, Jw)


@pytest.mark.parametrize("cfg", [(32, 32, 16, 16, 8, 0.3), (64, 48, 16, 8, 4, 0.6), (50, 50, 25, 25, 2, 0.4)])
def test_ten_pic_steps_bit_exact(cfg):
    s, P, parts, pl = _setup(*cfg, fields=0.01)
    s.upload_planes(pl)
    s.upload_particles(parts)
    s.pic_steps(10)
    got = by_tag(s.download_particles())
    gf = s.download_planes()
    want, wf = copy_parts(parts), [a.copy() for a in pl]
    counts, _ = ob.pic_steps(lib, P, wf, want, 10, do_sort=True)
    wkey = ob.tile_keys(lib, P, want["x"], want["y"])
    order = np.argsort(want["tag"], kind="stable")
    want = by_tag(want)
    for k in ("x", "y", "ux", "uy", "uz"):
        assert np.array_equal(got[k], want[k]), k
    for k in range(6):
        assert np.array_equal(gf[k], wf[k]), k
    # tile occupancy and particle->tile index: exact integers
    assert np.array_equal(s.tile_counts(), counts)
    raw = s.download_particles()
    gidx = s.particle_tile_index()
    assert (np.diff(gidx) >= 0).all()                        # sorted by tile
    assert np.array_equal(gidx[np.argsort(raw["tag"], kind="stable")], wkey[order])


@pytest.mark.parametrize("ppc", [12, 16, 20, 24])
def test_lane_group_walks_bit_exact(ppc):
    """The sparse-slab push variant splits its warps into lane groups of 32 / 16 / 8 per CTA (mean occupancy > 20 / > 10 /
    else): 12-20 ppc runs the 16-lane walk, 24 ppc the 32-lane walk of the ADAPT... and of the dense kernel; same bits."""
    s, P, parts, pl = _setup(48, 32, 16, 16, ppc, 0.5, fields=0.02)
    s.upload_planes(pl)
    s.upload_particles(parts)
    s.pic_steps(6)
    want, wf = copy_parts(parts), [a.copy() for a in pl]
    counts, _ = ob.pic_steps(lib, P, wf, want, 6)
    got, want = by_tag(s.download_particles()), by_tag(want)
    for k in ("x", "y", "ux", "uy", "uz"):
        assert np.array_equal(got[k], want[k]), k
    for a, b in zip(s.download_planes(), wf):
        assert np.array_equal(a, b)
    assert np.array_equal(s.tile_counts(), counts)


def test_mixed_density_tiles_bit_exact():
    """Sparse background + dense drifting band in one launch: CTAs of the same kernel take different lane-group walks."""
    nx, ny, tn, dx = 64, 64, 16, 0.1
    dt = pkg().courant_dt(dx, dx)
    bg = plasma().load_uniform(nx, ny, tn, tn, dx, dx, 4, u_th=0.3, seed=21)
    beam = plasma().load_uniform(nx, ny, tn, tn, dx, dx, 48, u_th=0.05, seed=22, drift=(0.5, 0.0, 0.0), band=(0.25, 0.5))
    beam["tag"] = (beam["tag"].astype(np.int64) + 10 ** 6).astype(np.int32)
    parts = {k: np.concatenate([bg[k], beam[k]]) for k in bg}
    rng = np.random.default_rng(5)
    pl = [0.02 * rng.standard_normal((ny, nx)) for _ in range(6)]
    P = ob.make_params(lib, nx, ny, tn, tn, dx, dx, dt)
    s = pkg().Sim(nx, ny, tn, tn, dx, dx, dt, max_particles=parts["x"].shape[0] + 64)
    s.upload_planes(pl)
    s.upload_particles(parts)
    s.pic_steps(8)
    want, wf = copy_parts(parts), [a.copy() for a in pl]
    counts, _ = ob.pic_steps(lib, P, wf, want, 8)
    got, want = by_tag(s.download_particles()), by_tag(want)
    for k in ("x", "y", "ux", "uy", "uz"):
        assert np.array_equal(got[k], want[k]), k
    for a, b in zip(s.download_planes(), wf):
        assert np.array_equal(a, b)
    assert np.array_equal(s.tile_counts(), counts)


def test_device_loader_matches_host_loader():
    nx, ny, tnx, tny, dx, dy = 64, 32, 16, 16, 0.1, 0.1
    m = pkg()
    s = m.Sim(nx, ny, tnx, tny, dx, dy, 0.03, max_particles=nx * ny * 12)
    s.load_uniform(6, u_th=0.05, drift=(0.1, 0.0, -0.2), two_stream=0.2, seed=99)
    s.load_uniform(4, u_th=0.01, drift=(0.5, 0.0, 0.0), density=10.0, band=(0.25, 0.5), seed=7, append=True)
    got = by_tag(s.download_particles())
    a = plasma().load_uniform(nx, ny, tnx, tny, dx, dy, 6, u_th=0.05, drift=(0.1, 0.0, -0.2), two_stream=0.2, seed=99)
    b = plasma().load_uniform(nx, ny, tnx, tny, dx, dy, 4, u_th=0.01, drift=(0.5, 0.0, 0.0), density=10.0, band=(0.25, 0.5),
                              seed=7, tag0=nx * ny * 6)
    want = by_tag({k: np.concatenate([a[k], b[k]]) for k in a})
    assert got["x"].shape == want["x"].shape
    for k in want:
        assert np.array_equal(got[k], want[k]), k
    counts = s.tile_counts()
    key = plasma().tile_of(want["x"], want["y"], nx, tnx, tny, dx, dy)
    assert np.array_equal(counts, np.bincount(key, minlength=8))


def test_empty_and_ragged_occupancy():
    """no particles at all; particles only in a band (most tiles empty); one particle."""
    m = pkg()
    nx, ny, tnx, tny, dx, dy = 32, 32, 8, 8, 0.1, 0.1
    dt = m.courant_dt(dx, dy)
    P = ob.make_params(lib, nx, ny, tnx, tny, dx, dy, dt)
    rng = np.random.default_rng(1)
    pl = [0.02 * rng.standard_normal((ny, nx)) for _ in range(6)]
    for parts in (plasma().load_uniform(nx, ny, tnx, tny, dx, dy, 3, u_th=0.9, band=(0.5, 0.625), seed=4),
                  {k: v[:1] for k, v in plasma().load_uniform(nx, ny, tnx, tny, dx, dy, 1, u_th=0.9, seed=5).items()},
                  {k: v[:0] for k, v in plasma().load_uniform(nx, ny, tnx, tny, dx, dy, 1).items()}):
        s = m.Sim(nx, ny, tnx, tny, dx, dy, dt, max_particles=4096)
        s.upload_planes(pl)
        s.upload_particles(parts)
        s.pic_steps(5)
        got = by_tag(s.download_particles())
        want, wf = copy_parts(parts), [a.copy() for a in pl]
        counts, _ = ob.pic_steps(lib, P, wf, want, 5)
        want = by_tag(want)
        for k in ("x", "y", "ux", "uy", "uz"):
            assert np.array_equal(got[k], want[k]), k
        assert np.array_equal(s.tile_counts(), counts)
        gf = s.download_planes()
        for k in range(6):
            assert np.array_equal(gf[k], wf[k])


def test_unsorted_upload_is_sorted_on_device():
    s, P, parts, pl = _setup(32, 32, 16, 16, 4, 0.5)
    perm = np.random.default_rng(8).permutation(parts["x"].shape[0])
    shuffled = {k: np.ascontiguousarray(v[perm]) for k, v in parts.items()}
    s.upload_planes(pl)
    s.upload_particles(shuffled)
    assert (np.diff(s.particle_tile_index()) >= 0).all()
    s.pic_steps(3)
    got = by_tag(s.download_particles())
    want, wf = copy_parts(parts), [a.copy() for a in pl]
    ob.pic_steps(lib, P, wf, want, 3)
    want = by_tag(want)
    for k in ("x", "y", "ux", "uy", "uz"):
        assert np.array_equal(got[k], want[k]), k


def test_diagnostics_and_conservation():
    """energy/charge diagnostics agree with the oracle's; Gauss residual stays at round-off (charge-conserving deposit)."""
    s, P, parts, pl = _setup(64, 64, 16, 16, 16, 0.05, two_stream=0.2, fields=0.0)
    s.upload_planes(pl)
    s.upload_particles(parts)
    d0 = s.diagnostics()
    assert abs(d0["kinetic_energy"] - ob.load_oracle().orc_kinetic_energy(
        *[ob._dp(parts[k]) for k in ("ux", "uy", "uz", "q")], parts["x"].shape[0], -1.0)) < 1e-12 * d0["kinetic_energy"]
    assert abs(d0["charge"] - parts["q"].sum()) < 1e-12 * abs(parts["q"].sum())
    s.pic_steps(200)
    d1 = s.diagnostics()
    assert d1["particles"] == parts["x"].shape[0]
    assert d1["gauss_residual"] < 1e-11
    tot0, tot1 = d0["field_energy"] + d0["kinetic_energy"], d1["field_energy"] + d1["kinetic_energy"]
    assert abs(tot1 - tot0) / tot0 < 2e-3


def test_host_buffer_entry_point_matches_resident_path():
    s, P, parts, pl = _setup(32, 32, 16, 16, 8, 0.3)
    hp, hf = copy_parts(parts), [a.copy() for a in pl]
    s.pic_steps_host(4, hf, hp)
    want, wf = copy_parts(parts), [a.copy() for a in pl]
    ob.pic_steps(lib, P, wf, want, 4)
    hp, want = by_tag(hp), by_tag(want)
    for k in ("x", "y", "ux", "uy", "uz"):
        assert np.array_equal(hp[k], want[k]), k
    for k in range(6):
        assert np.array_equal(hf[k], wf[k])


def test_pipelined_host_step_matches_oracle_and_resident_path(monkeypatch):
    """mpic_pic_steps_host with one step overlaps the PCIe transfers with the compute chunk by chunk (api.cu:
    pic_step_host_pipelined).  Same bits as the oracle on a small case (forced through the pipeline), and on a 2 M-particle
    case in shuffled host order the same bits as upload -> pic_steps(1) -> download; the resident state it leaves behind
    continues identically."""
    monkeypatch.setenv("MPIC_PIPELINE_MIN", "2")
    s, P, parts, pl = _setup(48, 32, 16, 16, 7, 0.4)
    hp, hf = copy_parts(parts), [a.copy() for a in pl]
    for _ in range(3):                                   # three calls: the resident leftovers of one call never leak into the next
        s.pic_steps_host(1, hf, hp)
    want, wf = copy_parts(parts), [a.copy() for a in pl]
    counts, _ = ob.pic_steps(lib, P, wf, want, 3)
    hp, want = by_tag(hp), by_tag(want)
    for k in ("x", "y", "ux", "uy", "uz", "q"):
        assert np.array_equal(hp[k], want[k]), k
    for k in range(6):
        assert np.array_equal(hf[k], wf[k])
    assert np.array_equal(s.tile_counts(), counts)
    s.close()
    monkeypatch.delenv("MPIC_PIPELINE_MIN")
    nx = ny = 256
    dx = 0.1
    dt = pkg().courant_dt(dx, dx)
    parts = plasma().load_uniform(nx, ny, 16, 16, dx, dx, 32, u_th=0.2, seed=9)
    rng = np.random.default_rng(4)
    order = rng.permutation(parts["x"].shape[0])
    parts = {k: np.ascontiguousarray(v[order]) for k, v in parts.items()}
    pl = [0.02 * rng.standard_normal((ny, nx)) for _ in range(6)]
    n = parts["x"].shape[0]
    a = pkg().Sim(nx, ny, 16, 16, dx, dx, dt, max_particles=n + 64)
    b = pkg().Sim(nx, ny, 16, 16, dx, dx, dt, max_particles=n + 64)
    hp, hf = copy_parts(parts), [x.copy() for x in pl]
    a.pic_steps_host(1, hf, hp)
    b.upload_planes(pl)
    b.upload_particles(parts)
    b.pic_steps(1)
    got, want = by_tag(hp), by_tag(b.download_particles())
    for k in ("x", "y", "ux", "uy", "uz", "q"):
        assert np.array_equal(got[k], want[k]), k
    for u, v in zip(hf, b.download_planes()):
        assert np.array_equal(u, v)
    a.pic_steps(2)
    b.pic_steps(2)
    got, want = by_tag(a.download_particles()), by_tag(b.download_particles())
    for k in ("x", "y", "ux", "uy", "uz"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(a.tile_counts(), b.tile_counts())


@pytest.mark.gpu
def test_branch_free_division_and_sqrt_are_ieee():
    """The push kernel's transcribed fast-path sequences (1/sqrt(x), 2/x for x >= 1) must be bit-identical to the
    compiler's IEEE sqrt.rn / div.rn on 2^31 pseudo-random arguments -- otherwise parity with the oracle's `/` and sqrt()
    would only be statistical."""
    for seed in (1, 0xC0FFEE):
        assert pkg().selftest_math(1 << 30, seed) == 0


def test_upload_rejects_positions_outside_the_box():
    """caller input must not reach device indexing unvalidated.  x == box_x, a negative y and a NaN are reported as
    MPIC_EINVAL by the upload (the particle is binned into cell 0, nothing is written out of bounds), and the handle stays
    usable for a valid upload afterwards."""
    m = pkg()
    nx = ny = 32
    dx = dy = 0.1
    s = m.Sim(nx, ny, 16, 16, dx, dy, m.courant_dt(dx, dy), max_particles=4096)
    good = plasma().load_uniform(nx, ny, 16, 16, dx, dy, 2, u_th=0.1, seed=3)
    for bad_x, bad_y in ((nx * dx, 0.5), (0.5, -1e-9), (float("nan"), 0.5)):
        p = copy_parts(good)
        p["x"][7], p["y"][7] = bad_x, bad_y
        with pytest.raises(m.MpicError, match="outside"):
            s.upload_particles(p)
    s.upload_particles(good)
    s.pic_steps(2)
    s.sync()
    assert s.num_particles == good["x"].shape[0]
    rc = m.lib().mpic_upload_particles(s.h, 4, None, None, None, None, None, None, None)
    assert rc == -1 and b"null" in m.lib().mpic_last_error()


@pytest.mark.parametrize("cfg", [(32, 32, 16, 16, 8, 0.3, 0.0), (48, 32, 8, 8, 40, 0.6, 0.2), (64, 32, 16, 16, 3, 0.9, 0.0),
                                 (16, 16, 8, 8, 300, 0.7, 0.0)])
def test_sort_is_stable_and_reproducible(cfg):
    """SURVEY.md B.8 / VERDICT r1 item 7: after 10 steps the slot order (the order a download returns) equals the oracle's
    STABLE sort by (tile, cell) element-wise -- tag[] and every component --, and two identical runs return identical arrays.
    Covers dense cells (two chunks per cell; packed 8-bit running mover counts), sparse ones (lane groups of 8), many cell
    crossers per step, and cells of 300 particles (running counts in shared memory)."""
    nx, ny, tnx, tny, ppc, u_th, ts = cfg
    outs = []
    for rep in range(2):
        s, P, parts, pl = _setup(nx, ny, tnx, tny, ppc, u_th, seed=5, two_stream=ts)
        shuf = np.random.default_rng(1).permutation(parts["x"].shape[0])       # arbitrary upload order
        up = {k: np.ascontiguousarray(v[shuf]) for k, v in parts.items()}
        s.upload_planes(pl)
        s.upload_particles(up)
        if rep == 0:                                   # the upload itself sorts stably (per-cell ascending storage index)
            got0 = s.download_particles()
            k0 = ob.cell_keys(lib, P, up["x"], up["y"])
            o0 = np.argsort(k0, kind="stable")
            assert np.array_equal(got0["tag"], up["tag"][o0])
        s.pic_steps(10)
        outs.append(s.download_particles())
        s.close()
    for k in outs[0]:
        assert np.array_equal(outs[0][k], outs[1][k]), k
    # oracle: same upload order, stable sort by cell key once (the upload), then 10 steps each followed by the stable cell sort
    want = {k: np.ascontiguousarray(v[shuf][o0]) for k, v in parts.items()}
    wf = [a.copy() for a in pl]
    ob.pic_steps(lib, P, wf, want, 10, do_sort=2)
    for k in ("tag", "x", "y", "ux", "uy", "uz", "q"):
        assert np.array_equal(outs[0][k], want[k]), k


def test_cell_of_70000_particles_keeps_exact_move_counts():
    """the packed push counts movers per (cell, direction) in 16-bit fields.  One cell with 70 000 particles
    that all cross its +x face in the first step (into the next TILE) overflows such a field; the kernel re-counts cells of
    65 536 or more from the direction codes, so slot order, phase space and tile occupancy still equal the oracle's bit for bit."""
    nx = ny = 32
    s, P, parts, pl = _setup(nx, ny, 16, 16, 2, 0.1, seed=21, fields=0.01)
    dx = dy = 0.1
    rng = np.random.default_rng(22)
    nf = 70000
    fat = {"x": (15 + 0.70 + 0.29 * rng.random(nf)) * dx, "y": (7 + 0.2 + 0.6 * rng.random(nf)) * dy,
           "ux": 3.0 + 0.1 * rng.standard_normal(nf), "uy": 0.2 * rng.standard_normal(nf), "uz": 0.1 * rng.standard_normal(nf),
           "q": np.full(nf, -1e-7), "tag": (np.arange(nf) + 10_000_000).astype(np.int32)}
    allp = {k: np.ascontiguousarray(np.concatenate([parts[k], fat[k]])) for k in parts}
    n = allp["x"].shape[0]
    assert n >= 6 * nx * ny                                                # dense enough for the packed walk
    s.close()
    s = pkg().Sim(nx, ny, 16, 16, dx, dy, pkg().courant_dt(dx, dy), max_particles=n + 64)
    s.upload_planes(pl)
    s.upload_particles(allp)
    k0 = ob.cell_keys(lib, P, allp["x"], allp["y"])
    o0 = np.argsort(k0, kind="stable")
    assert np.bincount(k0).max() >= 70000
    s.pic_steps(1)
    moved = s.download_particles()
    assert (np.floor(moved["x"][moved["tag"] >= 10_000_000] / dx) == 16).sum() > 65536      # they did cross, into the next tile
    s.pic_steps(3)
    s.sync()
    got = s.download_particles()
    want = {k: np.ascontiguousarray(v[o0]) for k, v in allp.items()}
    wf = [a.copy() for a in pl]
    counts, _ = ob.pic_steps(lib, P, wf, want, 4, do_sort=2)
    for k in ("tag", "x", "y", "ux", "uy", "uz", "q"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(s.tile_counts(), counts)
    gf = s.download_planes()
    for k in range(6):
        assert np.array_equal(gf[k], wf[k]), k


@pytest.mark.parametrize("cfg", [(32, 32, 16, 16, 40, 0.6, 0.0), (64, 32, 16, 16, 64, 0.05, 0.2), (48, 30, 8, 6, 30, 0.9, 0.0),
                                 (64, 64, 32, 32, 26, 0.7, 0.0)])
def test_packed_walk_bit_exact(cfg):
    """Dense slabs (>= 24 particles per cell) run push_packed_kernel: warps walk contiguous slot ranges across cell boundaries
    (carry cell + masked flushes).  Cases: half the particles crossing a cell face per step, a cold two-stream lattice (every
    chunk boundary is a cell boundary), a generic (non-templated) tile shape, 32x32 tiles."""
    nx, ny, tnx, tny, ppc, u_th, ts = cfg
    s, P, parts, pl = _setup(nx, ny, tnx, tny, ppc, u_th, seed=9, two_stream=ts, fields=0.02)
    s.upload_planes(pl)
    s.upload_particles(parts)
    s.pic_steps(5)
    want, wf = copy_parts(parts), [a.copy() for a in pl]
    counts, _ = ob.pic_steps(lib, P, wf, want, 5)
    got, want = by_tag(s.download_particles()), by_tag(want)
    for k in ("x", "y", "ux", "uy", "uz"):
        assert np.array_equal(got[k], want[k]), k
    for a, b in zip(s.download_planes(), wf):
        assert np.array_equal(a, b)
    assert np.array_equal(s.tile_counts(), counts)


_VARIANT_SCRIPT = r"""
import sys, numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[1] + "/tests")
from oracle import bindings as ob
from util import pkg, plasma, by_tag, copy_parts
lib = ob.load_oracle()
nx, ny, tn, dx = 48, 32, 16, 0.1
dt = pkg().courant_dt(dx, dx)
P = ob.make_params(lib, nx, ny, tn, tn, dx, dx, dt)
for ppc in (7, 33):
    parts = plasma().load_uniform(nx, ny, tn, tn, dx, dx, ppc, u_th=0.5, seed=2)
    rng = np.random.default_rng(3)
    pl = [0.02 * rng.standard_normal((ny, nx)) for _ in range(6)]
    s = pkg().Sim(nx, ny, tn, tn, dx, dx, dt, max_particles=parts["x"].shape[0] + 64)
    s.upload_planes(pl); s.upload_particles(parts); s.pic_steps(6)
    raw = s.download_particles()
    want, wf = copy_parts(parts), [a.copy() for a in pl]
    ob.pic_steps(lib, P, wf, want, 6, do_sort=2)
    for k in ("tag", "x", "y", "ux", "uy", "uz"):
        assert np.array_equal(raw[k], want[k]), (ppc, k)
    for a, b in zip(s.download_planes(), wf):
        assert np.array_equal(a, b)
print("variant ok")
"""


@pytest.mark.parametrize("gl", ["0", "1", "8", "16", "32"])
def test_every_push_variant_bit_exact_and_same_slot_order(gl):
    """MPIC_PUSH_GL forces one walk (0 adaptive lane groups, 1 packed, 8/16/32 fixed lane groups) for sparse AND dense cells:
    each must give the oracle's bits and the oracle's stable slot order (the variant is read once per process -> subprocess)."""
    import os, subprocess, sys
    from util import ROOT
    env = dict(os.environ, MPIC_PUSH_GL=gl)
    r = subprocess.run([sys.executable, "-c", _VARIANT_SCRIPT, ROOT], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "variant ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.parametrize("cfg", [(32, 32, 16, 16, 8, 0.4), (48, 30, 8, 6, 5, 0.8)])
def test_direct_cic_deposit_mode_bit_exact(cfg):
    """mpic_params.deposit = 1 (SURVEY.md B.5 option a, the literal "CIC charge and current deposition" of north_star): the
    current of one push and ten full PIC steps, bit-exact versus the oracle in the same mode; rho (node weights) through
    mpic_download_rho; and the mode really differs from the charge-conserving default."""
    nx, ny, tnx, tny, ppc, u_th = cfg
    dx, dy = 0.1, 0.1 * 1.25 if nx != ny else 0.1
    m = pkg()
    dt = m.courant_dt(dx, dy)
    P = ob.make_params(lib, nx, ny, tnx, tny, dx, dy, dt, deposit="cic")
    parts = plasma().load_uniform(nx, ny, tnx, tny, dx, dy, ppc, u_th=u_th, seed=8)
    rng = np.random.default_rng(8)
    pl = [0.02 * rng.standard_normal((ny, nx)) for _ in range(6)]
    s = m.Sim(nx, ny, tnx, tny, dx, dy, dt, max_particles=parts["x"].shape[0] + 64, deposit="cic")
    s.upload_planes(pl)
    s.upload_particles(parts)
    assert np.array_equal(s.download_rho(), ob.deposit_rho(lib, P, parts))
    s.push_deposit()
    s.fold_current()
    J = s.download_current()
    want = copy_parts(parts)
    Jw = ob.push_deposit(lib, P, pl, want)
    assert np.array_equal(J, Jw)
    z = copy_parts(parts)
    Jz = ob.push_deposit(lib, ob.make_params(lib, nx, ny, tnx, tny, dx, dy, dt), pl, z)
    assert not np.array_equal(Jw[0], Jz[0]) and np.array_equal(Jw[2], Jz[2])      # Jz is CIC in both modes
    g = np.sqrt(1 + want["ux"]**2 + want["uy"]**2 + want["uz"]**2)
    for c, k in enumerate(("ux", "uy", "uz")):                                      # total current = sum q v, each component
        assert abs(J[c].sum() / 2.0**52 * dx * dy - (want["q"] * want[k] / g).sum()) < 1e-12
    s.sort_tiles()
    s.step_b_half(); s.step_e_full(); s.step_b_half()
    s.pic_steps(9)
    got = by_tag(s.download_particles())
    gf = s.download_planes()
    want, wf = copy_parts(parts), [a.copy() for a in pl]
    counts, _ = ob.pic_steps(lib, P, wf, want, 10, do_sort=True)
    want = by_tag(want)
    for k in ("x", "y", "ux", "uy", "uz"):
        assert np.array_equal(got[k], want[k]), k
    for k in range(6):
        assert np.array_equal(gf[k], wf[k]), k
    assert np.array_equal(s.tile_counts(), counts)
    assert s.diagnostics()["gauss_residual"] > 1e-9                                # not charge conserving: Gauss's law drifts
