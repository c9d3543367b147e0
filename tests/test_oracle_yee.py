"""CPU tests: the Yee/tiling ORACLE (oracle/yee_oracle.c) pinned against
  (1) the reference's shipped dumps  (tests/golden/pulse_rows.npz <- Simulation/Fields/*.h5, v1 scheme + FMA build),
  (2) the reference's own sources compiled here (oracle/_ref/*.so) and the fixtures generated from them."""
import os
import numpy as np
import pytest
from oracle import bindings as ob

lib = ob.load_oracle()


def _pulse(golden_dir):
    return np.load(os.path.join(golden_dir, "pulse_rows.npz"))


def _planes_from_rows(g, step, ny):
    nx = int(g["nx"])
    z = lambda: np.zeros((ny, nx))
    ey = np.tile(g[f"Ey_{step}"], (ny, 1)).copy()
    bz = np.tile(g[f"Bz_{step}"], (ny, 1)).copy()
    return [z(), ey, z(), z(), z(), bz]


def _dt(g):
    dx, dy = float(g["box_x"]) / int(g["nx"]), float(g["box_y"]) / int(g["ny"])
    return 0.5 * (1.0 / np.sqrt(1.0 / (dx * dx) + 1.0 / (dy * dy))), dx, dy   # PIC_2D.cpp:71-72


def test_replay_shipped_dumps_full_grid(golden_dir):
    """v1 scheme + FMA contraction replays the reference's own dumps BIT-EXACTLY on the full 450x450 grid."""
    g = _pulse(golden_dir)
    dt, dx, dy = _dt(g)
    assert f"{dt:.6g}" == "0.00785674"                     # Simulation/Fields/params.txt: dt=0.00785674
    p = _planes_from_rows(g, 0, 450)
    done = 0
    for s in (25, 50, 100):
        ob.yee_steps(lib, p, dt, dx, dy, s - done, scheme=1, mode=1)
        done = s
        assert np.array_equal(p[1], np.tile(g[f"Ey_{s}"], (450, 1))), s
        assert np.array_equal(p[5], np.tile(g[f"Bz_{s}"], (450, 1))), s
        for k in (0, 2, 3, 4):
            assert not p[k].any()


def test_replay_shipped_dumps_to_last_step(golden_dir):
    """Same, on a 3-row strip (the pulse is y-uniform, every row does identical arithmetic) through the LAST shipped dump."""
    g = _pulse(golden_dir)
    dt, dx, dy = _dt(g)
    p = _planes_from_rows(g, 0, 3)
    done = 0
    for s in (1000, 10000, 63625):
        ob.yee_steps(lib, p, dt, dx, dy, s - done, scheme=1, mode=1)
        done = s
        assert np.array_equal(p[1][1], g[f"Ey_{s}"]), s
        assert np.array_equal(p[5][1], g[f"Bz_{s}"]), s


def test_current_scheme_is_not_the_dump_scheme(golden_dir):
    """the current half-B/E/half-B scheme differs from the dumps by ~7.8e-3 rel-L2 at step 25."""
    g = _pulse(golden_dir)
    dt, dx, dy = _dt(g)
    p = _planes_from_rows(g, 0, 3)
    ob.yee_steps(lib, p, dt, dx, dy, 25, scheme=0, mode=0)
    rel = np.linalg.norm(p[5][1] - g["Bz_25"]) / np.linalg.norm(g["Bz_25"])
    assert 1e-3 < rel < 3e-2


def test_energy_known_answers(golden_dir):
    """discrete energy of the CURRENT scheme on the Test-3 pulse (FMA build)."""
    g = _pulse(golden_dir)
    dt, dx, dy = _dt(g)
    p = _planes_from_rows(g, 0, 450)
    want = {0: 0.11250000000000002, 25: 0.11250000066586865, 50: 0.11250000148140987, 75: 0.11250000111595541,
            100: 0.11250000037798263}
    done = 0
    for s in sorted(want):
        ob.yee_steps(lib, p, dt, dx, dy, s - done, scheme=0, mode=1)
        done = s
        e = 0.5 * dx * dy * sum(float((a * a).sum()) for a in p)       # numpy pairwise sum, as in the survey probe
        assert abs(e - want[s]) < 1e-15 and abs(ob.field_energy(lib, p, dx, dy) - want[s]) < 1e-12, s


def test_oracle_matches_reference_loop(golden_dir):
    """Global-array oracle == the reference functions driven tile by tile with guard exchange and tile migration
    (fixture made from oracle/_ref/libref.so): bit-exact, and every dumped tile patch == periodic window (F8)."""
    f = np.load(os.path.join(golden_dir, "ref_current.npz"))
    nx, ny, R, C, tr, tc, g, steps = [int(v) for v in f["meta"]]
    p = [np.ascontiguousarray(f["ic"][..., k]) for k in range(6)]
    ob.yee_steps(lib, p, float(f["dt"]), float(f["dx"]), float(f["dy"]), steps, scheme=0, mode=0)
    for k in range(6):
        assert np.array_equal(p[k], f["out"][..., k]), k
    tnx, tny = nx // (C * tc), ny // (R * tr)
    full = [np.pad(a, g, mode="wrap") for a in p]
    for gid in range(R * tr * C * tc):
        r0, c0 = (gid // (C * tc)) * tny, (gid % (C * tc)) * tnx
        for k in range(6):
            assert np.array_equal(f["tiles"][gid][..., k], full[k][r0:r0 + tny + 2 * g, c0:c0 + tnx + 2 * g])
    # ownership after the three migrations of the fixture
    own = np.repeat(np.repeat(np.arange(R * C).reshape(R, C), tr, 0), tc, 1).ravel()
    own[5] = 3; own[7] = 3; own[14] = 2
    assert np.array_equal(f["owner"], own)


@pytest.mark.skipif(not ob.have_ref(), reason="oracle/_ref not built")
def test_oracle_matches_reference_live():
    """Same comparison against the live _ref library on a different shape (when oracle/_ref is present)."""
    ref = ob.load_ref("libref.so")
    rng = np.random.default_rng(3)
    nx, ny, g = 40, 24, 2
    f0 = rng.standard_normal((ny, nx, 6))
    dx, dy, dt = 0.1, 0.13, 0.03
    res = ob.ref_run_loop(ref, f0, 2, 1, 2, 5, g, dt, dx, dy, 7, 3)
    p = [np.ascontiguousarray(f0[..., k]) for k in range(6)]
    ob.yee_steps(lib, p, dt, dx, dy, 7, scheme=0, mode=0)
    for k in range(6):
        assert np.array_equal(p[k], res["global"][..., k])
    # v1 two-phase loop == scheme 1 (E and Bz bit-exact in FMA mode; Bx/By differ by gcc's addsub vectorisation)
    v1 = ob.load_ref("libref_v1fma.so")
    res = ob.ref_run_loop(v1, f0, 1, 2, 3, 2, g, dt, dx, dy, 1, 2)
    p = [np.ascontiguousarray(f0[..., k]) for k in range(6)]
    ob.yee_steps(lib, p, dt, dx, dy, 1, scheme=1, mode=1)
    for k in (5,):
        assert np.array_equal(p[k], res["global"][..., k])
    for k in range(6):
        assert np.allclose(p[k], res["global"][..., k], rtol=0, atol=1e-14)


def test_tile_functions_vs_reference_fixture(golden_dir):
    f = np.load(os.path.join(golden_dir, "ref_tiles.npz"))
    tnx, tny, g = [int(v) for v in f["shape"]]
    grid = np.ascontiguousarray(f["grid"])
    for d in range(8):
        out = np.zeros_like(f[f"pack{d}"])
        n = lib.orc_pack_send_buffer(ob._dp(grid), tnx, tny, g, d, ob._dp(out))
        assert n == out.shape[0] == lib.orc_strip_cells(d, tnx, tny, g)
        assert np.array_equal(out, f[f"pack{d}"]), d
        gg = grid.copy()
        strip = np.ascontiguousarray(f[f"strip{d}"])
        lib.orc_update_guard_region(ob._dp(gg), tnx, tny, g, d, ob._dp(strip))
        assert np.array_equal(gg, f[f"unpack{d}"]), d
    for (gtr, gtc, key) in ((6, 6, "nbr_6x6"), (3, 5, "nbr_3x5")):
        got = np.array([[lib.orc_get_neighbor_gid(gid, d, gtr, gtc) for d in range(8)] for gid in range(gtr * gtc)])
        assert np.array_equal(got, f[key])
    # Report Fig. 2b: neighbours of gid 0 on a 6x6 tile grid
    assert sorted(f["nbr_6x6"][0]) == [1, 5, 6, 7, 11, 30, 31, 35]
    import ctypes
    for n in range(1, 65):
        r, c = ctypes.c_int(), ctypes.c_int()
        lib.orc_find_best_grid(n, ctypes.byref(r), ctypes.byref(c))
        assert (r.value, c.value) == tuple(f["best_grid"][n - 1]) == tuple(f["best_tile_grid"][n - 1])
    E, B = np.ascontiguousarray(f["E"]), np.ascontiguousarray(f["B"])
    Bn = np.zeros_like(B)
    lib.orc_update_b_half_tile(ob._dp(E), ob._dp(B), ob._dp(Bn), tnx, tny, g, 0.013, 0.11, 0.07, 0, 0)
    assert np.array_equal(Bn, f["Bnew"])
    En = E.copy()
    lib.orc_update_e_full_tile(ob._dp(En), ob._dp(B), tnx, tny, g, 0.013, 0.11, 0.07, 0)
    assert np.array_equal(En, f["Enew"])
    assert lib.orc_compute_tag(17, 5) == 17 * 8 + 5


@pytest.mark.skipif(not (ob.have_ref() and ob.have_ref("libref_omp.so")), reason="oracle/_ref not built")
def test_openmp_variant_of_the_reference_loop_is_bit_identical():
    """libref_omp.so (the all-cores CPU baseline: same reference functions, loops over tiles under OpenMP) == libref.so."""
    rng = np.random.default_rng(11)
    g6 = rng.standard_normal((60, 90, 6))
    a = ob.ref_run_loop(ob.load_ref("libref.so"), g6, 1, 1, 4, 6, 2, 0.004, 0.1, 0.07, 7)
    b = ob.ref_run_loop(ob.load_ref("libref_omp.so"), g6, 1, 1, 4, 6, 2, 0.004, 0.1, 0.07, 7)
    assert np.array_equal(a["global"], b["global"]) and np.array_equal(a["tiles"], b["tiles"])
