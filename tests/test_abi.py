"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/minipic_b200.h declares, the pure
integer helpers match the reference fixture, and compute entry points FAIL LOUDLY without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re
import numpy as np
import pytest
from util import pkg, ROOT, have_gpu, PKG


def _declared():
    h = open(os.path.join(ROOT, "include", "minipic_b200.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    return sorted(set(re.findall(r"\b(mpic_[a-z0-9_]+)\s*\(", h)))


def test_library_exports_every_declared_symbol():
    L = pkg().lib()
    names = _declared()
    assert len(names) > 40
    for n in names:
        assert hasattr(L, n), n
    assert L.mpic_abi_version() == 3


def test_integer_helpers_on_cpu():
    L = pkg().lib()
    f = np.load(os.path.join(ROOT, "tests", "golden", "ref_tiles.npz"))
    got = np.array([[L.mpic_get_neighbor_gid(gid, d, 6, 6) for d in range(8)] for gid in range(36)])
    assert np.array_equal(got, f["nbr_6x6"])
    for n in range(1, 65):
        r, c = C.c_int(), C.c_int()
        L.mpic_find_best_grid(n, C.byref(r), C.byref(c))
        assert (r.value, c.value) == tuple(f["best_grid"][n - 1])
    drow, dcol, opp = (C.c_int * 8)(), (C.c_int * 8)(), (C.c_int * 8)()
    L.mpic_direction_tables(drow, dcol, opp)
    assert list(drow) == [0, 0, -1, 1, -1, -1, 1, 1] and list(dcol) == [-1, 1, 0, 0, -1, 1, -1, 1]
    assert list(opp) == [1, 0, 3, 2, 7, 6, 5, 4]
    # slab rows partition every tile row exactly once
    for gtr in (1, 7, 32, 256):
        for n in (1, 2, 4, 8):
            if n > gtr:
                continue
            rows = [pkg().slab_rows(gtr, n, r) for r in range(n)]
            assert rows[0][0] == 0 and sum(nr for _, nr in rows) == gtr
            assert all(rows[k][0] + rows[k][1] == rows[k + 1][0] for k in range(n - 1))


@pytest.mark.skipif(have_gpu(), reason="this box has a GPU")
def test_compute_fails_loudly_without_gpu():
    m = pkg()
    with pytest.raises(m.MpicError, match="no CUDA device"):
        m.Sim(32, 32, 16, 16, 0.1, 0.1, 0.03)
    E = np.zeros((8, 8, 3))
    with pytest.raises(m.MpicError):
        m.update_e_full(E, E.copy(), 4, 4, 2, 0.01, 0.1, 0.1)


def test_create_rejects_bad_parameters_before_touching_the_device():
    """Argument validation needs no GPU: dt beyond one cell per step (the push assumes |v| dt < min(dx,dy)), tiles that do
    not divide the grid, capacities beyond the 32-bit particle index."""
    m = pkg()
    with pytest.raises(m.MpicError, match="exceeds min"):
        m.Sim(32, 32, 16, 16, 0.1, 0.1, 0.15, max_particles=100)
    with pytest.raises(m.MpicError, match="multiples of the tile"):
        m.Sim(30, 32, 16, 16, 0.1, 0.1, 0.03)
    with pytest.raises(m.MpicError, match="32-bit"):
        m.Sim(32, 32, 16, 16, 0.1, 0.1, 0.03, max_particles=2**31)


def test_product_never_touches_the_oracle():
    """The product package and its CUDA sources must not import, link or dlopen anything under oracle/."""
    bad = []
    for base in (os.path.join(ROOT, PKG), os.path.join(ROOT, "include"), os.path.join(ROOT, "host"), os.path.join(ROOT, "drivers")):
        for dp, _, fs in os.walk(base):
            for fn in fs:
                if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp", "Makefile")):
                    txt = open(os.path.join(dp, fn), errors="ignore").read()
                    if re.search(r"(from|import)\s+oracle|liboracle|oracle/_ref|libref", txt):
                        bad.append(os.path.join(dp, fn))
    assert not bad, bad
