"""The C++ host layer on a GPU: the parameterised driver host/pic2d (L1 resident API), the reference's OWN unmodified
PIC_2D.cpp linked against host/*.h (host/_ref_dropin/pic2d_reference, L0 compat API), and the HDF5 dumps they write."""
import glob
import os
import subprocess
import numpy as np
import pytest
from h5mini import assemble, H5Mini, GRID_DTYPE
from util import ROOT

pytestmark = pytest.mark.gpu
PIC2D = os.path.join(ROOT, "host", "pic2d")
DROPIN = os.path.join(ROOT, "host", "_ref_dropin", "pic2d_reference")
G = os.path.join(ROOT, "tests", "golden")


def _run(cmd, cwd, timeout=300, ok=(0,)):
    r = subprocess.run(cmd, cwd=cwd, capture_output=True, text=True, timeout=timeout)
    assert r.returncode in ok, r.stdout[-2000:] + r.stderr[-2000:]
    return r.stdout


def test_driver_replays_the_shipped_dumps(tmp_path):
    """Defaults = the reference's literals (450x450, 9 ranks x 36 tiles, Test-3 pulse); with the old scheme + contraction the
    driver's own HDF5 files reproduce the reference's shipped dumps bit for bit, including the step-5 migration."""
    if not os.path.exists(PIC2D):
        pytest.skip("host/pic2d not built")
    g = np.load(os.path.join(G, "pulse_rows.npz"))
    out = _run([PIC2D, "--scheme", "1", "--fma", "1", "--steps", "100", "--out", "Simulation/Fields"], str(tmp_path))
    assert "Simulation complete." in out
    assert out.count("sent Tile GID=") == 3 and "Rank 0 sent Tile GID=3 to Rank 3" in out
    d = str(tmp_path / "Simulation" / "Fields")
    assert open(os.path.join(d, "params.txt")).read().startswith("box_x=10\nbox_y=10\nnx_global=450\nny_global=450\nguard=2\ninterior_nx=25\n")
    for st in (0, 25, 50, 100):
        paths = sorted(glob.glob(os.path.join(d, f"fields_rank_*_step_{st}.h5")))
        assert len(paths) == 9
        planes, owner, tiles = assemble(paths, 450, 450, 25, 25, 2)
        if st > 0:      # the step-0 IC is recomputed with this box's libm: 1-2 ulp from the shipped file
            assert np.abs(planes["Ey"] - np.tile(g[f"Ey_{st}"], (450, 1))).max() < 1e-15
        for n in ("Ex", "Ez", "Bx", "By"):
            assert not planes[n].any()
        if st in (0, 25):
            assert np.array_equal(np.array([owner[i] for i in range(324)]), g[f"owner_{st}"])
    # number of tiles per rank file after the migration, as in the reference's step-25 files: 35/35/35/39/36...
    sizes = [len(H5Mini(p).tiles()) for p in sorted(glob.glob(os.path.join(d, "fields_rank_*_step_25.h5")))]
    assert sizes == [35, 35, 35, 39, 36, 36, 36, 36, 36]


def test_unmodified_reference_main_runs_on_the_gpu_and_matches_the_resident_driver(tmp_path):
    """PIC_2D.cpp as shipped (1 rank -> 36 tiles of 75x75, per-tile updateBhalf/updateEfull/pack/unpack through the L0 ABI)
    is cut off after a while; its step-25 dump must equal the resident driver's, bit for bit, guards included."""
    if not (os.path.exists(DROPIN) and os.path.exists(PIC2D)):
        pytest.skip("drop-in binaries not built")
    a, b = tmp_path / "ref", tmp_path / "fast"
    a.mkdir(); b.mkdir()
    try:
        subprocess.run([DROPIN], cwd=str(a), capture_output=True, timeout=120)
    except subprocess.TimeoutExpired:
        pass
    fa = str(a / "Simulation" / "Fields" / "fields_rank_0_step_25.h5")
    assert os.path.exists(fa), "the reference main did not reach step 25 in 120 s"
    _run([PIC2D, "--ranks", "1", "--steps", "25", "--quiet"], str(b))
    fb = str(b / "Simulation" / "Fields" / "fields_rank_0_step_25.h5")
    ta, tb = H5Mini(fa).tiles(), H5Mini(fb).tiles()
    assert sorted(ta) == sorted(tb) == list(range(36))
    for gid in ta:
        assert ta[gid][0] == tb[gid][0]
        for n in GRID_DTYPE.names:
            assert np.array_equal(ta[gid][1][n], tb[gid][1][n]), (gid, n)


def test_driver_with_particles_one_vs_two_slabs(tmp_path):
    """Two-stream plasma through the C++ driver: --gpus 2 (peer-copy ring; both handles share the GPU on a one-GPU box)
    writes the same dumps as --gpus 1, and the conservation diagnostics behave."""
    if not os.path.exists(PIC2D):
        pytest.skip("host/pic2d not built")
    outs = []
    for gpus in (1, 2):
        d = tmp_path / f"g{gpus}"
        d.mkdir()
        outs.append(_run([PIC2D, "--nx", "64", "--ny", "64", "--box-x", "6.4", "--box-y", "6.4", "--tile-nx", "16", "--tile-ny", "16",
                          "--ranks", "2", "--ic", "zero", "--ppc", "8", "--uth", "0.05", "--two-stream", "0.2", "--steps", "40",
                          "--save-frequency", "20", "--diag-every", "20", "--gpus", str(gpus), "--no-migrate-demo"], str(d)))
    for st in (20, 40):
        pa = sorted(glob.glob(str(tmp_path / "g1" / "Simulation" / "Fields" / f"fields_rank_*_step_{st}.h5")))
        pb = sorted(glob.glob(str(tmp_path / "g2" / "Simulation" / "Fields" / f"fields_rank_*_step_{st}.h5")))
        assert len(pa) == 2 and len(pb) == 4        # ranks split columns, slabs split rows: 2 files per rank with --gpus 2
        A, _, _ = assemble(pa, 64, 64, 16, 16, 2)
        B, _, _ = assemble(pb, 64, 64, 16, 16, 2)
        for n in GRID_DTYPE.names:
            assert np.array_equal(A[n], B[n]), (st, n)
        assert np.abs(A["Ex"]).max() > 0
    diag = [l for l in outs[0].splitlines() if l.startswith("step ")]
    e0 = float(diag[0].split("total=")[1].split()[0]); e1 = float(diag[-1].split("total=")[1].split()[0])
    assert abs(e1 - e0) / e0 < 1e-3
    assert float(diag[-1].split("gauss_residual=")[1].split()[0]) < 1e-11


def test_driver_rebalances_in_place_and_dumps_asynchronously(tmp_path):
    """host/pic2d --gpus 3 --rebalance-every 10 on a beam-plasma box (the load sits in a band of rows): the load balancer
    (mpic_ring_rebalance: plan, device parcels, handles rebuilt in place) moves tile rows during the run, and the dumps --
    written asynchronously -- equal those of the single-GPU run bit for bit."""
    if not os.path.exists(PIC2D):
        pytest.skip("host/pic2d not built")
    common = ["--nx", "64", "--ny", "96", "--box-x", "6.4", "--box-y", "9.6", "--tile-nx", "16", "--tile-ny", "8", "--ranks", "1",
              "--ic", "zero", "--ppc", "4", "--uth", "0.05", "--beam-density", "8", "--beam-lo", "0.05", "--beam-hi", "0.25",
              "--steps", "30", "--save-frequency", "15", "--no-migrate-demo", "--async-dump"]
    outs = []
    for gpus, extra in ((1, []), (3, ["--rebalance-every", "10", "--rebalance-w-cell", "0.25"])):
        d = tmp_path / f"g{gpus}"
        d.mkdir()
        outs.append(_run([PIC2D] + common + ["--gpus", str(gpus)] + extra, str(d)))
    assert "rebalanced in" in outs[1], outs[1][-2000:]
    rows = [int(v) for v in outs[1].split("tile rows per GPU:")[1].splitlines()[0].split()]
    assert sum(rows) == 12 and rows != [4, 4, 4]
    for st in (15, 30):
        pa = sorted(glob.glob(str(tmp_path / "g1" / "Simulation" / "Fields" / f"fields_rank_*_step_{st}.h5")))
        pb = sorted(glob.glob(str(tmp_path / "g3" / "Simulation" / "Fields" / f"fields_rank_*_step_{st}.h5")))
        assert len(pa) == 1 and len(pb) == 3
        A, _, _ = assemble(pa, 64, 96, 16, 8, 2)
        B, _, _ = assemble(pb, 64, 96, 16, 8, 2)
        for n in GRID_DTYPE.names:
            assert np.array_equal(A[n], B[n]), (st, n)
        assert np.abs(A["Ex"]).max() > 0
