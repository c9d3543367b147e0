"""mpic_kinetic_energy_static -- the entry point for callers that cannot allocate device memory -- against mpic_diagnostics.

Such callers exist around this library: a code whose particle arrays sit in its own device pool (all of HBM handed to it at
start-up) has nothing left for the planes and sort tables of an mpic_create handle, and a diagnostics hook running between
the steps of a peer ring must not call cudaMalloc, whose device-wide synchronisation would hold up every rank that is
spinning on this rank's flags.  The entry point therefore takes the caller's device arrays and keeps its reduction work area
in a static device array of the library image (9.5 KB per GPU).  That array is shared by all calls and remembers the last
one, so this test calls it several times in a row on different data and wants the handle-based answer every time.

Here the role of "the caller's pool" is played by torch: torch allocates the four arrays, the library allocates nothing.
The file name sorts last on purpose: with `pytest -x` a failure here cannot hide any other test.
"""
import numpy as np
import pytest
from util import pkg, plasma

# the tolerance tests/test_gpu_pic.py::test_diagnostics_and_conservation holds the energy diagnostics to
RTOL = 1e-12


def _state(steps_between=5):
    """A small two-stream plasma on a handle (the existing entry point's side of the comparison)."""
    m = pkg()
    nx = ny = 128
    dx = dy = 0.1
    dt = m.courant_dt(dx, dy)
    parts = plasma().load_uniform(nx, ny, 16, 16, dx, dy, 16, u_th=0.05, two_stream=0.2, seed=11)
    rng = np.random.default_rng(11)
    s = m.Sim(nx, ny, 16, 16, dx, dy, dt, max_particles=parts["x"].shape[0] + 64)
    s.upload_planes([0.01 * rng.standard_normal((ny, nx)) for _ in range(6)])
    s.upload_particles(parts)
    return s


def _on_device(slot):
    """The particle arrays in memory the CALLER owns: torch tensors stand in for a host code's own device pool."""
    import torch
    t = {k: torch.from_numpy(np.ascontiguousarray(slot[k])).to("cuda:0") for k in ("ux", "uy", "uz", "q")}
    torch.cuda.synchronize()
    return t


def _static(t, n, qm=-1.0):
    return pkg().kinetic_energy_static(t["ux"].data_ptr(), t["uy"].data_ptr(), t["uz"].data_ptr(), t["q"].data_ptr(), n, qm)


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_diagnostics_on_consecutive_calls():
    import torch
    torch.cuda.set_device(0)
    s = _state()
    n = s.num_particles
    assert n == 128 * 128 * 16

    # first call: the state as uploaded (slot order of the handle, so both sides add the same numbers in the same order)
    d0 = s.diagnostics()
    t0 = _on_device(s.download_particles())
    ke0, ch0 = _static(t0, n)
    assert abs(ke0 - d0["kinetic_energy"]) <= RTOL * d0["kinetic_energy"]
    assert abs(ch0 - d0["charge"]) <= RTOL * abs(d0["charge"])

    # second call, straight after, on DIFFERENT data: nothing of the first call may survive in the shared work area
    s.pic_steps(5)
    d1 = s.diagnostics()
    assert d1["kinetic_energy"] != d0["kinetic_energy"]
    t1 = _on_device(s.download_particles())
    ke1, ch1 = _static(t1, n)
    assert abs(ke1 - d1["kinetic_energy"]) <= RTOL * d1["kinetic_energy"]
    assert abs(ch1 - d1["charge"]) <= RTOL * abs(d1["charge"])

    # third call on the first data again: same digits as the first time (the reduction is ordered, not atomic)
    assert _static(t0, n) == (ke0, ch0)

    # a shorter prefix leaves most blocks without work: their partials must read as zero, not as the last call's
    k = 1000
    ke_k, ch_k = _static(t1, k)
    u2 = sum(t1[c][:k].cpu().numpy() ** 2 for c in ("ux", "uy", "uz"))
    q = t1["q"][:k].cpu().numpy()
    want = float(np.sum(np.abs(q) * (u2 / (np.sqrt(1.0 + u2) + 1.0))))
    assert abs(ke_k - want) <= RTOL * want and abs(ch_k - q.sum()) <= RTOL * abs(q.sum())

    # another charge/mass ratio scales the energy only
    ke_m, ch_m = _static(t1, n, qm=-0.5)
    assert abs(ke_m - 2.0 * ke1) <= RTOL * ke_m and ch_m == ch1

    # empty population: zeros, no launch
    assert _static(t1, 0) == (0.0, 0.0)

    # host memory is refused, not dereferenced on the device
    host = np.ones(16)
    with pytest.raises(pkg().MpicError, match="not device memory"):
        pkg().kinetic_energy_static(host.ctypes.data, host.ctypes.data, host.ctypes.data, host.ctypes.data, 16)
    # ... and the refusal leaves the entry point usable
    assert _static(t0, n) == (ke0, ch0)


def test_static_scratch_entry_point_rejects_sizes_beyond_its_maximum():
    """Argument checks come before any device work, so they hold on a box without a GPU too."""
    m = pkg()
    assert m.KINETIC_STATIC_MAX_PARTICLES == 2**31 - 64
    with pytest.raises(m.MpicError, match="exceed the maximum"):
        m.kinetic_energy_static(256, 256, 256, 256, m.KINETIC_STATIC_MAX_PARTICLES + 1)
    with pytest.raises(m.MpicError, match="negative"):
        m.kinetic_energy_static(256, 256, 256, 256, -1)
    with pytest.raises(m.MpicError, match="non-zero"):
        m.kinetic_energy_static(256, 256, 256, 256, 10, qm=0.0)
    with pytest.raises(m.MpicError, match="null particle array"):
        m.kinetic_energy_static(0, 256, 256, 256, 10)
