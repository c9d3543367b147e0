"""Full-size checks at BASELINE.json's named shapes, through size-independent properties (the oracle cannot run these
sizes in seconds): conservation laws, exact integer bookkeeping, and the physics the configs are named after.

  C2  two-stream 512x512 x 64 ppc, 2000 steps: electrostatic growth rate against cold-beam theory, energy budget,
      Gauss's law (the deposit is charge conserving), exact particle/charge bookkeeping; and a 2000-step BIT-EXACT run
      against the oracle on a small box ("conservation drift no worse than the reference's": identical bits => identical drift).
  C3  8192x8192 vacuum pulse: discrete energy conserved, y-invariance preserved exactly.
  C4  4096x512 slab (one GPU's share at 8 GPUs, 134 M particles) and the whole 4096x4096 x 64 ppc (1.07 B particles) on ONE
      GPU when 150 GB are free: tile occupancy == histogram of the particle->tile index == N, Gauss residual at round-off.
  C5  beam-plasma 4096x4096 (8 ppc background + 80 ppc beam in 10 % of the rows): loader occupancy exact, rebalancing
      policy evens the slabs, conservation through steps.
"""
import numpy as np
import pytest
from oracle import bindings as ob
from util import pkg, plasma, by_tag, copy_parts, PKG
import importlib

pytestmark = pytest.mark.gpu


def _free_gb():
    import torch
    free, _ = torch.cuda.mem_get_info(0)
    return free / 2**30


def test_c2_two_stream_growth_energy_and_gauss():
    m = pkg()
    nx = ny = 512
    ppc, dx = 64, 0.1
    dt = m.courant_dt(dx, dx)
    n = nx * ny * ppc
    s = m.Sim(nx, ny, 16, 16, dx, dx, dt, max_particles=n + 1024)
    s.load_uniform(ppc, u_th=0.001, two_stream=0.2, seed=12345)
    d0 = s.diagnostics()
    assert d0["particles"] == n
    t, wex, tot, gauss = [], [], [], []
    every, steps = 20, 2000
    for k in range(steps // every + 1):
        d = s.diagnostics()
        ex = s.download_planes()[0]
        t.append(k * every * dt)
        wex.append(0.5 * dx * dx * float(np.sum(ex * ex)))
        tot.append(d["field_energy"] + d["kinetic_energy"])
        gauss.append(d["gauss_residual"])
        assert d["particles"] == n
        assert abs(d["charge"] - d0["charge"]) < 1e-11 * abs(d0["charge"])      # q is carried, never recomputed
        if k < steps // every:
            s.pic_steps(every)
    t, wex, tot = np.array(t), np.array(wex), np.array(tot)
    # (1) Gauss's law: div E - rho stays where it started (E(0) = 0 with a neutralising background => residual is
    # measured relative to step 0 by the diagnostic); round-off only
    assert max(gauss) < 1e-9, max(gauss)
    # (2) linear phase: ln W_Ex grows at 2*gamma.  Cold symmetric beams (each omega_p^2/2): gamma_max = omega_p / (2 sqrt 2)
    # / gamma0^1.5 with u0 = 0.2 => 0.3433 omega_p; 2-D, dx = 0.1 and CIC smoothing lower it by a few per cent.
    wmax = wex.max()
    # (window: above the slowly growing noise floor of the quiet start, below saturation)
    sel = (wex > 3e-3 * wmax) & (wex < 0.3 * wmax) & (t < t[np.argmax(wex)])
    assert sel.sum() >= 5, (sel.sum(), wex[:10], wmax)
    slope = np.polyfit(t[sel], np.log(wex[sel]), 1)[0]
    gamma = 0.5 * slope
    gamma_th = 1.0 / (2.0 * np.sqrt(2.0)) / (1.0 + 0.2 ** 2) ** 0.75
    print(f"two-stream growth rate: measured {gamma:.4f} omega_p, cold-beam theory {gamma_th:.4f}; saturation W_Ex={wmax:.4e}, "
          f"energy drift {abs(tot[-1] - tot[0]) / tot[0]:.3e}")
    assert 0.70 * gamma_th < gamma < 1.10 * gamma_th, (gamma, gamma_th)
    # (3) the instability converts a sizeable part of the drift energy into field energy and the total is conserved
    assert wmax > 0.02 * tot[0]
    assert np.abs(tot - tot[0]).max() / tot[0] < 1e-2
    s.close()


def test_long_run_2000_steps_bit_exact_vs_oracle():
    """2000 steps of a small two-stream box: positions, momenta, fields, tile occupancy identical to the oracle's bits, so
    every conservation diagnostic drifts exactly as the CPU implementation's does."""
    m = pkg()
    lib = ob.load_oracle(omp=True)
    nx, ny, tn, ppc, dx = 64, 32, 16, 16, 0.1
    dt = m.courant_dt(dx, dx)
    parts = plasma().load_uniform(nx, ny, tn, tn, dx, dx, ppc, u_th=0.001, two_stream=0.2, seed=5)
    pl = [np.zeros((ny, nx)) for _ in range(6)]
    s = m.Sim(nx, ny, tn, tn, dx, dx, dt, max_particles=parts["x"].shape[0] + 64)
    s.upload_planes(pl)
    s.upload_particles(parts)
    s.pic_steps(2000)
    P = ob.make_params(lib, nx, ny, tn, tn, dx, dx, dt)
    want, wf = copy_parts(parts), [a.copy() for a in pl]
    counts, _ = ob.pic_steps(lib, P, wf, want, 2000)
    got, want = by_tag(s.download_particles()), by_tag(want)
    for k in ("x", "y", "ux", "uy", "uz"):
        assert np.array_equal(got[k], want[k]), k
    gf = s.download_planes()
    for k in range(6):
        assert np.array_equal(gf[k], wf[k]), k
    assert np.array_equal(s.tile_counts(), counts)
    assert float(np.abs(wf[0]).max()) > 1e-3          # the instability did develop: this is not a comparison of zeros


def test_c3_vacuum_pulse_8192():
    m = pkg()
    n = 8192
    dx = dy = 10.0 / n
    dt = m.courant_dt(dx, dy)
    x = (np.arange(n) + 0.5) * dx
    prof = 0.1 * np.sin(np.pi * x) * np.cos((x - 3.5) / 3.0 * np.pi / 2) ** 2 * (np.abs(x - 3.5) < 3.0)
    z = np.zeros((n, n))
    row = np.tile(prof, (n, 1))
    s = m.Sim(n, n, 32, 32, dx, dy, dt)
    s.upload_planes([z, row, z, z, z, row])
    del z
    e0 = s.diagnostics()["field_energy"]
    s.field_steps(100)
    e1 = s.diagnostics()["field_energy"]
    assert abs(e1 - e0) / e0 < 1e-6
    got = s.download_planes()
    # the pulse is uniform along y: every row must stay bit-identical to row 0 (periodic images included)
    for k in (1, 5):
        assert np.array_equal(got[k], np.broadcast_to(got[k][0], got[k].shape)), k
    for k in (0, 2, 3, 4):
        assert not got[k].any(), k
    # it moved: +x propagation by 100*dt
    xc = np.fft.irfft(np.fft.rfft(got[1][0]) * np.conj(np.fft.rfft(prof)), n)        # circular cross-correlation
    shift = int(np.argmax(xc))
    assert abs(shift * dx - 100 * dt) < 2 * dx, shift
    s.close()


def _occupancy_properties(s, n, steps):
    d0 = s.diagnostics()
    assert d0["particles"] == n
    s.pic_steps(steps)
    counts = s.tile_counts().astype(np.int64)
    assert counts.sum() == n
    d1 = s.diagnostics()
    assert d1["particles"] == n
    assert abs(d1["charge"] - d0["charge"]) < 1e-11 * abs(d0["charge"])
    # charge-conserving deposit: div E - rho keeps its step-0 value in every cell (0 for a uniform load; the beam's own
    # unneutralised excess for configs[4], where E starts at zero)
    assert abs(d1["gauss_residual"] - d0["gauss_residual"]) < 1e-9 * max(1.0, d0["gauss_residual"]), (d0, d1)
    return counts, d0, d1


def test_c4_slab_occupancy_and_tile_index():
    """One GPU's share of configs[3] at 8 GPUs: 4096 x 512 cells x 64 ppc = 134 M particles."""
    m = pkg()
    nx, ny, ppc, dx = 4096, 512, 64, 0.1
    n = nx * ny * ppc
    s = m.Sim(nx, ny, 16, 16, dx, dx, m.courant_dt(dx, dx), max_particles=n + 1024)
    s.load_uniform(ppc, u_th=0.05, seed=12345)
    assert np.array_equal(s.tile_counts(), np.full(s.ntiles_local, 256 * ppc, dtype=np.int32))     # step 0: exact
    counts, d0, d1 = _occupancy_properties(s, n, 5)
    gid = s.particle_tile_index()
    assert np.array_equal(np.bincount(gid, minlength=s.ntiles_local), counts)
    assert np.all(np.diff(gid) >= 0)                     # storage order == tile order (the sort's contract)
    assert counts.min() > 0.9 * 256 * ppc and counts.max() < 1.1 * 256 * ppc
    assert abs((d1["field_energy"] + d1["kinetic_energy"]) / (d0["field_energy"] + d0["kinetic_energy"]) - 1.0) < 1e-3
    s.close()


def test_c4_whole_domain_one_gpu_1e9_particles():
    """configs[3] complete on ONE B200: 4096^2 cells x 64 ppc = 1 073 741 824 particles (2 x 55.8 GB of SoA state)."""
    if _free_gb() < 150:
        pytest.skip("needs 150 GB of free HBM")
    m = pkg()
    nx = ny = 4096
    ppc, dx = 64, 0.1
    n = nx * ny * ppc
    s = m.Sim(nx, ny, 16, 16, dx, dx, m.courant_dt(dx, dx), max_particles=n + 1024)
    s.load_uniform(ppc, u_th=0.05, seed=12345)
    counts, d0, d1 = _occupancy_properties(s, n, 3)
    assert counts.min() > 0.9 * 256 * ppc and counts.max() < 1.1 * 256 * ppc
    s.close()


def test_c5_beam_plasma_occupancy_and_policy():
    """configs[4]: 4096^2, 8 ppc background + 80 ppc drifting beam in the band y in [0.45, 0.55): 0.27 B particles."""
    m = pkg()
    slab = importlib.import_module(PKG + ".slab")
    nx = ny = 4096
    dx, tn = 0.1, 16
    lo, hi = int(round(0.45 * ny)), int(round(0.55 * ny))
    n = nx * ny * 8 + nx * (hi - lo) * 80
    s = m.Sim(nx, ny, tn, tn, dx, dx, m.courant_dt(dx, dx), max_particles=n + 1024)
    s.load_uniform(8, u_th=0.05, seed=12345)
    s.load_uniform(80, u_th=0.01, drift=(0.5, 0.0, 0.0), band=(0.45, 0.55), density=10.0, seed=777, append=True)
    assert s.num_particles == n
    c0 = s.tile_counts().reshape(ny // tn, nx // tn).astype(np.int64)
    rows = np.arange(ny)
    per_row = 8 + 80 * ((rows >= lo) & (rows < hi))
    want = per_row.reshape(-1, tn).sum(axis=1) * tn            # particles per tile in each tile row
    assert np.array_equal(c0, np.broadcast_to(want[:, None], c0.shape))
    counts, d0, d1 = _occupancy_properties(s, n, 5)
    # an even 8-way slab split is badly imbalanced; the policy evens it
    rowc = slab.tile_row_costs(counts, nx // tn, tn * tn)
    even = [rowc[r * 32:(r + 1) * 32].sum() for r in range(8)]
    bal = [rowc[r0:r0 + nr].sum() for r0, nr in slab.balanced_slab_rows(rowc, 8)]
    # Imbalance of the even split under the cost model: with cost = particles only the two middle slabs carry 8 + 80 * 0.4 = 40 vs
    # 8 per cell -> max/mean = 40 / 16 = 2.5; the calibrated model adds W_CELL = 8 particle-equivalents per cell to every slab
    # ((40 + 8) / (16 + 8) = 2.0, measured 1.98 after five steps of beam spreading), hence the bound 1.7 (was 2.5 for the
    # particles-only model of early round 1)
    assert max(even) / np.mean(even) > 1.7
    assert max(bal) / np.mean(bal) < 1.25
    s.close()


def test_c3_dump_of_65536_tiles_off_the_step_path(tmp_path):
    """SURVEY.md f1 at scale: configs[2] (8192^2) in 32^2 tiles = 65 536 tile groups, 4.1 GB in the reference's HDF5 layout.
    The asynchronous dump costs the step path one patch kernel; the device-to-host copy and the file writer run next to the
    following steps.  The file is parsed back (mmap) and a sample of tiles is compared with the planes of the dump step."""
    import time
    from h5mini import H5Mini, GRID_DTYPE
    m = pkg()
    n, tn = 8192, 32
    dx = 10.0 / n
    s = m.Sim(n, n, tn, tn, dx, dx, m.courant_dt(dx, dx))
    x = (np.arange(n) + 0.5) * dx
    prof = 0.1 * np.sin(np.pi * x) * np.cos((x - 3.5) / 3.0 * np.pi / 2) ** 2 * (np.abs(x - 3.5) < 3.0)
    z = np.zeros((n, n))
    row = np.tile(prof, (n, 1))
    s.upload_planes([z, row, z, z, z, row])
    del z, row
    s.field_steps(5, fused=True)
    want = s.download_planes()
    path = str(tmp_path / "fields_rank_0_step_5.h5")
    s.dump_fields_async(path)                            # warm-up of the pinned staging (4 GB cudaMallocHost) is part of this call
    s.dump_wait()
    os_size = __import__("os").path.getsize(path)
    assert os_size > 65536 * 36 * 36 * 48
    s.sync()
    t0 = time.perf_counter()
    s.dump_fields_async(path)
    t_issue = time.perf_counter() - t0
    s.field_steps(20, fused=True)
    s.sync()
    t_steps = time.perf_counter() - t0
    s.dump_wait()
    t_all = time.perf_counter() - t0
    print(f"async dump of 65536 tiles: issue {1e3 * t_issue:.1f} ms, 20 steps done after {1e3 * t_steps:.1f} ms, file complete after {1e3 * t_all:.1f} ms")
    assert t_steps < t_all                               # the steps did not wait for the file
    h = H5Mini(path)
    links = h.links()
    assert len(links) == 65536 and h.eof == os_size
    gtc = n // tn
    pad = lambda a, r, c: np.take(np.take(a, np.arange(r * tn - 2, (r + 1) * tn + 2) % n, axis=0), np.arange(c * tn - 2, (c + 1) * tn + 2) % n, axis=1)
    for name, ohdr in links[::4099] + links[-1:]:
        gid = int(name[5:])
        r, c = divmod(gid, gtc)
        assert h.attrs(ohdr) == {"tileRow": r, "tileCol": c, "currentRank": 0}
        d = h.dataset(dict(h.links(ohdr))["fields"])
        for k, nm in enumerate(GRID_DTYPE.names):
            assert np.array_equal(d[nm], pad(want[k], r, c)), (gid, nm)


def test_c1_reference_default_200_steps_bit_exact_with_dumps(tmp_path):
    """configs[0] at its stated shape -- 128 x 128 cells, 16 x 16 tiles, 16 ppc thermal plasma (u_th = 0.05), periodic, 200 steps,
    a field dump every 25 steps (loop edge PIC_2D.cpp:74,414-419; dump layout HDF5_output.cpp:10-79) -- is the one named config
    the oracle can run IN FULL, so it is compared in full: at every dump step the particles in slot order (== the oracle's stable
    cell sort), the six planes, the tile occupancy and the per-particle tile index are array_equal to the oracle's, and the
    .h5 file written by mpic_dump_fields_async parses back to the oracle's tiles WITH their periodic guard cells.  The same
    configuration through the C++ driver host/pic2d writes the same dumps."""
    import glob
    import os
    import subprocess
    from h5mini import H5Mini, GRID_DTYPE, assemble
    from util import ROOT
    m = pkg()
    lib = ob.load_oracle()
    nx = ny = 128
    tn, ppc, u_th, steps, every = 16, 16, 0.05, 200, 25
    dx = dy = 0.1
    dt = m.courant_dt(dx, dy)
    P = ob.make_params(lib, nx, ny, tn, tn, dx, dy, dt)
    parts = plasma().load_uniform(nx, ny, tn, tn, dx, dy, ppc, u_th=u_th, seed=12345)
    n = parts["x"].shape[0]
    assert n == nx * ny * ppc
    planes = [np.zeros((ny, nx)) for _ in range(6)]
    s = m.Sim(nx, ny, tn, tn, dx, dy, dt, max_particles=n + 64)
    s.upload_planes(planes)
    s.upload_particles(parts)
    o0 = np.argsort(ob.cell_keys(lib, P, parts["x"], parts["y"]), kind="stable")
    want = {k: np.ascontiguousarray(v[o0]) for k, v in parts.items()}
    wf = [a.copy() for a in planes]
    gtc = nx // tn
    pad = lambda a, r, c: np.take(np.take(a, np.arange(r * tn - 2, (r + 1) * tn + 2) % ny, axis=0), np.arange(c * tn - 2, (c + 1) * tn + 2) % nx, axis=1)
    oracle_planes = {}
    for st in range(every, steps + 1, every):
        s.pic_steps(every)
        counts, _ = ob.pic_steps(lib, P, wf, want, every, do_sort=2)
        path = str(tmp_path / f"fields_rank_0_step_{st}.h5")
        s.dump_fields_async(path)
        got = s.download_particles()
        for k in ("tag", "x", "y", "ux", "uy", "uz", "q"):
            assert np.array_equal(got[k], want[k]), (st, k)
        gf = s.download_planes()
        for k in range(6):
            assert np.array_equal(gf[k], wf[k]), (st, k)
        assert np.array_equal(s.tile_counts(), counts) and counts.sum() == n
        assert np.array_equal(s.particle_tile_index(), ob.tile_keys(lib, P, want["x"], want["y"]))
        s.dump_wait()
        tiles = H5Mini(path).tiles()
        assert sorted(tiles) == list(range(gtc * gtc))
        for gid, (attrs, d) in tiles.items():
            r, c = divmod(gid, gtc)
            assert attrs == {"tileRow": r, "tileCol": c, "currentRank": 0}
            for k, nm in enumerate(GRID_DTYPE.names):
                assert np.array_equal(d[nm], pad(wf[k], r, c)), (st, gid, nm)
        oracle_planes[st] = [a.copy() for a in wf]
    assert max(np.abs(a).max() for a in wf) > 0                      # thermal noise did excite fields
    s.close()
    # the same run through the C++ driver (device loader, seed 12345 == plasma.load_uniform's default): same dumps
    pic2d = os.path.join(ROOT, "host", "pic2d")
    if not os.path.exists(pic2d):
        pytest.skip("host/pic2d not built (the library half of this test passed)")
    r = subprocess.run([pic2d, "--nx", "128", "--ny", "128", "--box-x", "12.8", "--box-y", "12.8", "--tile-nx", "16", "--tile-ny", "16",
                        "--ranks", "1", "--ic", "zero", "--ppc", "16", "--uth", "0.05", "--steps", "200", "--save-frequency", "25",
                        "--no-migrate-demo", "--quiet"], cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    for st in range(every, steps + 1, every):
        pa = sorted(glob.glob(str(tmp_path / "Simulation" / "Fields" / f"fields_rank_*_step_{st}.h5")))
        assert len(pa) == 1
        A, _, _ = assemble(pa, nx, ny, tn, tn, 2)
        for k, nm in enumerate(GRID_DTYPE.names):
            assert np.array_equal(A[nm], oracle_planes[st][k]), (st, nm)
