"""Generate the committed golden fixtures from the REFERENCE (run once in the build container; needs /root/reference).

  pulse_rows.npz   from the reference's shipped dumps Simulation/Fields/fields_rank_{0..8}_step_<s>.h5 (real HDF5 written
                   by the reference, old v1 scheme + FMA build).  The Test-3 pulse is uniform in y and
                   stays so bit-for-bit (checked here for every step used), so one row of Ey and Bz per step carries
                   the whole 450x450 dump; Ex,Ez,Bx,By are exactly zero (checked).  Also: tile ownership at steps 0/25
                   (the step-5 migration demo, PIC_2D.cpp:399-412) and two complete tile patches incl. guards.
  ref_current.npz  outputs of the reference's CURRENT Field_update.cpp + Auxiliar_functions.cpp (compiled here as
                   oracle/_ref/libref.so, strict IEEE) driven through ref_run_loop on a 2-D oblique wave (report Test 2)
                   with a 2x2 rank grid, 3x2 tiles per rank and a tile migration; 10 steps.
  ref_tiles.npz    pack/unpack/neighbour/tag known answers from the reference functions.

Usage:  python tests/golden/make_golden.py
"""
import glob
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from h5mini import assemble, GRID_DTYPE  # noqa: E402

REF = "/root/reference"
FIELDS = REF + "/Simulation/Fields"
STEPS = [0, 25, 50, 100, 1000, 10000, 63625]


def pulse_rows():
    out = {"steps": np.array(STEPS), "nx": 450, "ny": 450, "tnx": 25, "tny": 25, "guard": 2,
           "box_x": 10.0, "box_y": 10.0}
    for s in STEPS:
        paths = sorted(glob.glob(f"{FIELDS}/fields_rank_*_step_{s}.h5"))
        assert len(paths) == 9, (s, paths)
        planes, owner, tiles = assemble(paths, 450, 450, 25, 25, 2)
        for n in ("Ex", "Ez", "Bx", "By"):
            assert not planes[n].any(), (s, n)
        for n in ("Ey", "Bz"):
            assert (planes[n] == planes[n][0:1]).all(), (s, n)
            out[f"{n}_{s}"] = planes[n][0].copy()
        # guard ring == periodic neighbours' interior
        full = {n: np.pad(planes[n], 2, mode="wrap") for n in GRID_DTYPE.names}
        for gid, (a, d) in tiles.items():
            r0, c0 = a["tileRow"] * 25, a["tileCol"] * 25
            for n in GRID_DTYPE.names:
                assert (d[n] == full[n][r0:r0 + 29, c0:c0 + 29]).all(), (s, gid, n)
        if s in (0, 25):
            out[f"owner_{s}"] = np.array([owner[g] for g in range(324)], dtype=np.int32)
        if s == 25:
            for gid in (3, 200):
                out[f"tile{gid}_25"] = np.stack([tiles[gid][1][n] for n in GRID_DTYPE.names], axis=-1)
    np.savez_compressed(os.path.join(HERE, "pulse_rows.npz"), **out)
    print("pulse_rows.npz written")


def ref_current():
    from oracle import bindings as ob
    ref = ob.load_ref("libref.so")
    nx, ny, R, C, tr, tc, g = 48, 60, 2, 2, 3, 2, 2
    box_x, box_y = 4.8, 6.0
    dx, dy = box_x / nx, box_y / ny
    dt = 0.5 / np.sqrt(1 / dx**2 + 1 / dy**2)
    A, kx, ky = 0.1, 2 * np.pi * 2 / box_x, 2 * np.pi * 3 / box_y
    j, i = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    ph = lambda x, y: np.sin(kx * x + ky * y)
    f = np.zeros((ny, nx, 6))
    f[..., 0] = A / np.sqrt(2) * ph((i + .5) * dx, j * dy)
    f[..., 1] = -A / np.sqrt(2) * ph(i * dx, (j + .5) * dy)
    f[..., 2] = 0.03 * np.cos(kx * i * dx) * np.sin(ky * j * dy)          # exercise Ez/Bx/By too
    f[..., 5] = -A * ph((i + .5) * dx, (j + .5) * dy)
    mig = np.array([[5, 0, 3, 3], [5, 1, 3, 3], [7, 3, 0, 2]], dtype=np.int32)
    ic = f.copy()
    res = ob.ref_run_loop(ref, f, R, C, tr, tc, g, dt, dx, dy, 10, 3, mig)
    np.savez_compressed(os.path.join(HERE, "ref_current.npz"), ic=ic, out=res["global"], owner=res["owner"],
                        order=res["order"], tiles=res["tiles"], mig=mig,
                        meta=np.array([nx, ny, R, C, tr, tc, g, 10]), dt=dt, dx=dx, dy=dy)
    print("ref_current.npz written")


def ref_tiles():
    from oracle import bindings as ob
    ref = ob.load_ref("libref.so")
    rng = np.random.default_rng(7)
    tnx, tny, g = 7, 5, 2
    grid = rng.standard_normal((tny + 2 * g, tnx + 2 * g, 6))
    out = {"grid": grid, "shape": np.array([tnx, tny, g])}
    for d in range(8):
        out[f"pack{d}"] = ob.ref_pack(ref, grid, tnx, tny, g, d)
        strip = rng.standard_normal(out[f"pack{d}"].shape)
        out[f"strip{d}"] = strip
        out[f"unpack{d}"] = ob.ref_unpack(ref, grid, tnx, tny, g, d, strip)
    gtr, gtc = 6, 6
    out["nbr_6x6"] = np.array([[ref.ref_get_neighbor_gid(gid, d, gtr, gtc) for d in range(8)] for gid in range(36)], dtype=np.int32)
    out["nbr_3x5"] = np.array([[ref.ref_get_neighbor_gid(gid, d, 3, 5) for d in range(8)] for gid in range(15)], dtype=np.int32)
    out["best_grid"] = np.array([ob.ref_best_grid(ref, n) for n in range(1, 65)], dtype=np.int32)
    out["best_tile_grid"] = np.array([ob.ref_best_tile_grid(ref, n) for n in range(1, 65)], dtype=np.int32)
    # one B and one E tile update on random data (Field_update.cpp:14-59)
    E = rng.standard_normal((tny + 2 * g, tnx + 2 * g, 3)); B = rng.standard_normal((tny + 2 * g, tnx + 2 * g, 3))
    out["E"], out["B"] = E, B
    out["Bnew"] = ob.ref_update_b_half(ref, E, B, tnx, tny, g, 0.013, 0.11, 0.07)
    out["Enew"] = ob.ref_update_e_full(ref, E, B, tnx, tny, g, 0.013, 0.11, 0.07)
    np.savez_compressed(os.path.join(HERE, "ref_tiles.npz"), **out)
    print("ref_tiles.npz written")


if __name__ == "__main__":
    pulse_rows()
    ref_current()
    ref_tiles()
