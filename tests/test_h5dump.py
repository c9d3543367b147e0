"""Dump layout: the libhdf5-free writer behind saveRankData / mpic_save_rank_data, checked with the
same minimal parser (tests/h5mini.py) that reads the reference's own shipped HDF5 files.  Runs without a GPU."""
import os
import numpy as np
import pytest
from h5mini import H5Mini, GRID_DTYPE, assemble
from util import pkg, ROOT

G = os.path.join(ROOT, "tests", "golden")


def test_roundtrip_of_reference_tiles(tmp_path):
    """Two complete tile patches (with guards) taken from the reference's step-25 dumps, written back and re-read."""
    m = pkg()
    g = np.load(os.path.join(G, "pulse_rows.npz"))
    grid = np.stack([g["tile3_25"], g["tile200_25"]])
    path = str(tmp_path / "fields_rank_3_step_25.h5")
    m.save_rank_data(path, [3, 200], [0, 11], [3, 2], [3, 4], 25, 25, 2, grid)
    h = H5Mini(path)
    assert (h.leaf_k, h.internal_k) == (4, 16)                 # same group B-tree parameters as the reference's files
    assert h.eof == os.path.getsize(path)
    t = h.tiles()
    assert sorted(t) == [3, 200]
    assert t[3][0] == {"tileRow": 0, "tileCol": 3, "currentRank": 3}
    assert t[200][0] == {"tileRow": 11, "tileCol": 2, "currentRank": 4}
    for k, gid in enumerate((3, 200)):
        d = t[gid][1]
        assert d.shape == (29, 29) and d.dtype == GRID_DTYPE
        for c, n in enumerate(GRID_DTYPE.names):
            assert np.array_equal(d[n], grid[k, :, :, c])


@pytest.mark.parametrize("ntiles", [0, 1, 8, 9, 36, 257, 1500])
def test_many_tiles_multi_level_btree(tmp_path, ntiles):
    """Group B-tree growth: 8 symbols per SNOD, 32 children per TREE node -> 1500 tiles need a two-level tree."""
    m = pkg()
    rng = np.random.default_rng(ntiles)
    tnx, tny, g = 4, 3, 2
    gids = rng.permutation(4000)[:ntiles]
    grid = rng.standard_normal((ntiles, tny + 2 * g, tnx + 2 * g, 6))
    path = str(tmp_path / "f.h5")
    m.save_rank_data(path, gids, gids // 50, gids % 50, np.full(ntiles, 7), tnx, tny, g, grid)
    h = H5Mini(path)
    names = [n for n, _ in h.links()]
    assert names == sorted(f"Tile_{i}" for i in gids)          # B-tree order is strcmp order, as libhdf5 requires
    t = h.tiles()
    assert len(t) == ntiles
    for k, gid in enumerate(gids):
        a, d = t[int(gid)]
        assert a == {"tileRow": int(gid) // 50, "tileCol": int(gid) % 50, "currentRank": 7}
        assert np.array_equal(np.stack([d[n] for n in GRID_DTYPE.names], axis=-1), grid[k])


def test_65536_tiles_three_level_btree(tmp_path):
    """The tile count of BASELINE.json configs[2] (8192^2 cells in 32^2 tiles = 65 536 groups) with small patches: the group
    B-tree needs three levels; every group is found again, in strcmp order, with its attributes and data."""
    m = pkg()
    nt, tnx, tny, g = 65536, 4, 4, 2
    rng = np.random.default_rng(65536)
    gids = np.arange(nt)
    grid = rng.standard_normal((nt, tny + 2 * g, tnx + 2 * g, 6))
    path = str(tmp_path / "f.h5")
    m.save_rank_data(path, gids, gids // 256, gids % 256, gids % 9, tnx, tny, g, grid)
    h = H5Mini(path)
    assert h.eof == os.path.getsize(path)
    links = h.links()
    assert [n for n, _ in links] == sorted(f"Tile_{i}" for i in gids)
    for name, ohdr in links[::997] + links[-1:]:
        gid = int(name[5:])
        d = h.dataset(dict(h.links(ohdr))["fields"])
        assert h.attrs(ohdr) == {"tileRow": gid // 256, "tileCol": gid % 256, "currentRank": gid % 9}
        assert np.array_equal(np.stack([d[n] for n in GRID_DTYPE.names], axis=-1), grid[gid])


def test_assemble_like_file_reader(tmp_path):
    """The consumer contract (File_reader.py:82-117): interiors placed by tileRow/tileCol rebuild the global planes."""
    m = pkg()
    nx, ny, tnx, tny, g = 24, 16, 6, 4, 2
    rng = np.random.default_rng(3)
    planes = rng.standard_normal((6, ny, nx))
    padded = np.pad(planes, ((0, 0), (g, g), (g, g)), mode="wrap")
    gtc = nx // tnx
    files = []
    for rank, gl in enumerate(([0, 1, 2, 3, 4, 5, 6, 7], [8, 9, 10, 11, 12, 13, 14, 15])):
        grid = np.stack([np.moveaxis(padded[:, (i // gtc) * tny:(i // gtc) * tny + tny + 2 * g,
                                            (i % gtc) * tnx:(i % gtc) * tnx + tnx + 2 * g], 0, -1) for i in gl])
        p = str(tmp_path / f"fields_rank_{rank}_step_0.h5")
        m.save_rank_data(p, gl, [i // gtc for i in gl], [i % gtc for i in gl], [rank] * len(gl), tnx, tny, g, grid)
        files.append(p)
    got, owner, _ = assemble(files, nx, ny, tnx, tny, g)
    for c, n in enumerate(GRID_DTYPE.names):
        assert np.array_equal(got[n], planes[c])
    assert [owner[i] for i in range(16)] == [0] * 8 + [1] * 8


def test_params_txt_matches_reference_formatting(tmp_path):
    m = pkg()
    p = str(tmp_path / "params.txt")
    dx = 10.0 / 450
    dt = 0.5 / np.sqrt(2.0 / dx ** 2)
    m.write_params(p, 10.0, 10.0, 450, 450, 2, 25, 25, 500.0, dt, int(500.0 / dt))
    # byte-for-byte the file the reference shipped (Simulation/Fields/params.txt)
    assert open(p).read() == ("box_x=10\nbox_y=10\nnx_global=450\nny_global=450\nguard=2\ninterior_nx=25\ninterior_ny=25\n"
                              "sim_time=500\ndt=0.00785674\ntotal_steps=63639\n")
