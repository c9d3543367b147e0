"""ctypes bindings for the CHECKER libraries (oracle/liboracle*.so, oracle/_ref/*.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module; the
product package never does (tests/test_no_oracle_in_product.py enforces it).
"""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)
c_lp = C.POINTER(C.c_int64)


def _dp(a):
    return a.ctypes.data_as(c_dp)


def _ip(a):
    return a.ctypes.data_as(c_ip)


def _lp(a):
    return a.ctypes.data_as(c_lp)


def build(verbose=False):
    """(Re)build the checker libraries with oracle/Makefile (oracle/_ref only when /root/reference is present)."""
    r = subprocess.run(["make", "-C", HERE], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout)


def _load(path):
    if not os.path.exists(path):
        build()
    return C.CDLL(path)


class PicParams(C.Structure):
    _fields_ = [("nx", C.c_int), ("ny", C.c_int), ("tnx", C.c_int), ("tny", C.c_int),
                ("dx", C.c_double), ("dy", C.c_double), ("dt", C.c_double),
                ("inv_dx", C.c_double), ("inv_dy", C.c_double), ("box_x", C.c_double), ("box_y", C.c_double),
                ("h", C.c_double), ("cjx", C.c_double), ("cjy", C.c_double), ("cjz", C.c_double),
                ("deposit", C.c_int), ("reserved", C.c_int)]


_cache = {}


def load_oracle(omp=False):
    name = "liboracle_omp.so" if omp else "liboracle.so"
    if name in _cache:
        return _cache[name]
    lib = _load(os.path.join(HERE, name))
    lib.orc_field_energy.restype = C.c_double
    lib.orc_kinetic_energy.restype = C.c_double
    lib.orc_kinetic_energy.argtypes = [c_dp, c_dp, c_dp, c_dp, C.c_long, C.c_double]
    lib.orc_yee_steps_global.argtypes = [c_dp] * 6 + [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int]
    lib.orc_b_sweep_global.argtypes = [c_dp] * 6 + [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int]
    lib.orc_e_sweep_global.argtypes = [c_dp] * 6 + [C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int]
    lib.orc_field_energy.argtypes = [c_dp] * 6 + [C.c_int, C.c_int, C.c_double, C.c_double]
    lib.orc_update_b_half_tile.argtypes = [c_dp, c_dp, c_dp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int]
    lib.orc_update_e_full_tile.argtypes = [c_dp, c_dp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int]
    lib.orc_pack_send_buffer.argtypes = [c_dp, C.c_int, C.c_int, C.c_int, C.c_int, c_dp]
    lib.orc_update_guard_region.argtypes = [c_dp, C.c_int, C.c_int, C.c_int, C.c_int, c_dp]
    lib.orc_pic_make_params.argtypes = [C.POINTER(PicParams), C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double]
    lib.orc_push_deposit.argtypes = [C.POINTER(PicParams)] + [c_dp] * 6 + [c_lp] * 3 + [c_dp] * 6 + [C.c_long]
    lib.orc_apply_current.argtypes = [C.POINTER(PicParams)] + [c_dp] * 3 + [c_lp] * 3
    lib.orc_tile_keys.argtypes = [C.POINTER(PicParams), c_dp, c_dp, C.c_long, c_ip]
    lib.orc_cell_keys.argtypes = [C.POINTER(PicParams), c_dp, c_dp, C.c_long, c_ip]
    lib.orc_sort_perm.argtypes = [c_ip, C.c_long, C.c_int, c_ip, c_lp]
    lib.orc_deposit_rho.argtypes = [C.POINTER(PicParams), c_dp, c_dp, c_dp, C.c_long, c_lp]
    lib.orc_pic_steps.argtypes = [C.POINTER(PicParams)] + [c_dp] * 6 + [c_dp] * 6 + [c_ip, C.c_long, C.c_int, C.c_int, c_ip, c_lp]
    _cache[name] = lib
    return lib


def load_ref(name="libref.so"):
    key = "_ref/" + name
    if key in _cache:
        return _cache[key]
    lib = _load(os.path.join(HERE, "_ref", name))
    lib.ref_run_loop.restype = C.c_double
    lib.ref_run_loop.argtypes = [c_dp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                 C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, c_ip, C.c_int, c_ip, c_dp, c_ip]
    lib.ref_update_b_half.argtypes = [c_dp, c_dp, c_dp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]
    lib.ref_update_e_full.argtypes = [c_dp, c_dp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]
    lib.ref_pack_send_buffer.argtypes = [c_dp, C.c_int, C.c_int, C.c_int, C.c_int, c_dp]
    lib.ref_update_guard_region.argtypes = [c_dp, C.c_int, C.c_int, C.c_int, C.c_int, c_dp, C.c_int]
    lib.ref_heaviside.restype = C.c_double
    lib.ref_heaviside.argtypes = [C.c_double]
    _cache[key] = lib
    return lib


def have_ref(name="libref.so"):
    return os.path.exists(os.path.join(HERE, "_ref", name)) or os.path.exists("/root/reference/Field_update.cpp")


# ------------------------------------------------------------------------------------------ reference wrappers
def strip_cells(d, tnx, tny, g):
    return tny * g if d < 2 else tnx * g if d < 4 else g * g


def ref_pack(ref, grid6, tnx, tny, g, d):
    grid6 = np.ascontiguousarray(grid6, dtype=np.float64)
    out = np.zeros((strip_cells(d, tnx, tny, g), 6))
    n = ref.ref_pack_send_buffer(_dp(grid6), tnx, tny, g, d, _dp(out))
    assert n == out.shape[0]
    return out


def ref_unpack(ref, grid6, tnx, tny, g, d, strip):
    grid = np.ascontiguousarray(grid6, dtype=np.float64).copy()
    strip = np.ascontiguousarray(strip, dtype=np.float64)
    ref.ref_update_guard_region(_dp(grid), tnx, tny, g, d, _dp(strip), strip.shape[0])
    return grid


def ref_best_grid(ref, n):
    r, c = C.c_int(), C.c_int()
    ref.ref_find_best_grid(n, C.byref(r), C.byref(c))
    return r.value, c.value


def ref_best_tile_grid(ref, n):
    r, c = C.c_int(), C.c_int()
    ref.ref_find_best_tile_grid(n, C.byref(r), C.byref(c))
    return r.value, c.value


def ref_update_b_half(ref, E3, B3, tnx, tny, g, dt, dx, dy):
    E3 = np.ascontiguousarray(E3); B3 = np.ascontiguousarray(B3)
    out = np.zeros_like(B3)                 # the reference driver zero-initialises BhalfNew (PIC_2D.cpp:189)
    ref.ref_update_b_half(_dp(E3), _dp(B3), _dp(out), tnx, tny, g, dt, dx, dy)
    return out


def ref_update_e_full(ref, E3, B3, tnx, tny, g, dt, dx, dy):
    E = np.ascontiguousarray(E3).copy(); B3 = np.ascontiguousarray(B3)
    ref.ref_update_e_full(_dp(E), _dp(B3), tnx, tny, g, dt, dx, dy)
    return E


def ref_run_loop(ref, global6, R, C_, tile_rows, tile_cols, g, dt, dx, dy, steps, phases=3, mig=None, want_tiles=True):
    """Run the reference loop structure on (ny,nx,6) interior fields; see oracle/ref_wrap.cpp:ref_run_loop."""
    f = np.ascontiguousarray(global6, dtype=np.float64).copy()
    ny, nx = f.shape[:2]
    gtr, gtc = R * tile_rows, C_ * tile_cols
    total = gtr * gtc
    tnx, tny = nx // gtc, ny // gtr
    owner = np.zeros(total, dtype=np.int32)
    order = np.zeros((R * C_, total), dtype=np.int32)
    tiles = np.zeros((total, tny + 2 * g, tnx + 2 * g, 6)) if want_tiles else None
    mig = np.zeros((0, 4), dtype=np.int32) if mig is None else np.ascontiguousarray(mig, dtype=np.int32)
    secs = ref.ref_run_loop(_dp(f), nx, ny, R, C_, tile_rows, tile_cols, g, dt, dx, dy, steps, phases,
                            _ip(mig), mig.shape[0], _ip(owner), _dp(tiles) if want_tiles else None, _ip(order))
    if secs < 0:
        raise ValueError("ref_run_loop rejected the configuration (%g)" % secs)
    return {"global": f, "owner": owner, "order": order, "tiles": tiles, "seconds": secs}


# ------------------------------------------------------------------------------------------ oracle wrappers
def make_params(lib, nx, ny, tnx, tny, dx, dy, dt, qm=-1.0, deposit="zigzag"):
    p = PicParams()
    lib.orc_pic_make_params(C.byref(p), nx, ny, tnx, tny, dx, dy, dt, qm)
    p.deposit = {"zigzag": 0, "cic": 1}[deposit]
    return p


def yee_steps(lib, planes, dt, dx, dy, steps, scheme=0, mode=0):
    """planes: list of six (ny,nx) float64 arrays [Ex,Ey,Ez,Bx,By,Bz], advanced in place."""
    ny, nx = planes[0].shape
    for a in planes:
        assert a.flags.c_contiguous and a.dtype == np.float64
    lib.orc_yee_steps_global(*[_dp(a) for a in planes], nx, ny, dt, dx, dy, steps, scheme, mode)


def field_energy(lib, planes, dx, dy):
    ny, nx = planes[0].shape
    return lib.orc_field_energy(*[_dp(a) for a in planes], nx, ny, dx, dy)


def pic_steps(lib, P, planes, parts, steps, do_sort=True):
    """parts: dict x,y,ux,uy,uz,q (float64) and tag (int32), all advanced in place.  Returns (tile counts, J int64 (3,ny,nx)).
    do_sort: True/1 = stable sort by tile key; 2 = stable sort by (tile, cell in tile), the GPU's slot order."""
    n = parts["x"].shape[0]
    ntiles = (P.nx // P.tnx) * (P.ny // P.tny)
    counts = np.zeros(ntiles, dtype=np.int32)
    J = np.zeros((3, P.ny, P.nx), dtype=np.int64)
    lib.orc_pic_steps(C.byref(P), *[_dp(a) for a in planes],
                      *[_dp(parts[k]) for k in ("x", "y", "ux", "uy", "uz", "q")], _ip(parts["tag"]), n,
                      steps, int(do_sort), _ip(counts), _lp(J))
    return counts, J


def push_deposit(lib, P, planes, parts):
    """One gather+push+move+deposit pass (no field update, no sort).  Returns J int64 (3,ny,nx)."""
    n = parts["x"].shape[0]
    J = np.zeros((3, P.ny, P.nx), dtype=np.int64)
    lib.orc_push_deposit(C.byref(P), *[_dp(a) for a in planes], _lp(J[0]), _lp(J[1]), _lp(J[2]),
                         *[_dp(parts[k]) for k in ("x", "y", "ux", "uy", "uz", "q")], n)
    return J


def tile_keys(lib, P, x, y):
    key = np.zeros(x.shape[0], dtype=np.int32)
    lib.orc_tile_keys(C.byref(P), _dp(x), _dp(y), x.shape[0], _ip(key))
    return key


def cell_keys(lib, P, x, y):
    """(tile, cell in tile) sort key of every particle: the GPU layout's slot order (a refinement of tile_keys)."""
    key = np.zeros(x.shape[0], dtype=np.int32)
    lib.orc_cell_keys(C.byref(P), _dp(x), _dp(y), x.shape[0], _ip(key))
    return key


def deposit_rho(lib, P, parts):
    rho = np.zeros((P.ny, P.nx), dtype=np.int64)
    lib.orc_deposit_rho(C.byref(P), _dp(parts["x"]), _dp(parts["y"]), _dp(parts["q"]), parts["x"].shape[0], _lp(rho))
    return rho


def b_sweep(lib, planes, dt, dx, dy, scheme=0, mode=0):
    ny, nx = planes[0].shape
    lib.orc_b_sweep_global(*[_dp(a) for a in planes], nx, ny, dt, dx, dy, scheme, mode)


def e_sweep(lib, planes, dt, dx, dy, mode=0):
    ny, nx = planes[0].shape
    lib.orc_e_sweep_global(*[_dp(a) for a in planes], nx, ny, dt, dx, dy, mode)


def apply_current(lib, P, planes, J):
    """E <- E - dt*J (J: int64 (3,ny,nx) fixed point), as in orc_pic_steps."""
    lib.orc_apply_current(C.byref(P), _dp(planes[0]), _dp(planes[1]), _dp(planes[2]), _lp(J[0]), _lp(J[1]), _lp(J[2]))
