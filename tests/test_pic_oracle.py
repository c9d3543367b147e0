"""Known-answer tests for the AUTHORED particle oracle (oracle/pic_oracle.c).  The reference has no particle code
(SURVEY.md F1), so these physics checks are what pins the oracle: parity for push/deposit/sort is "unpinned"
with respect to the reference and is GPU-vs-this-oracle."""
import importlib
import numpy as np
from oracle import bindings as ob

plasma = importlib.import_module("mini-pic--2d-pic-code-with-tiling-load-balance_b200.plasma")
lib = ob.load_oracle()


def _zero_planes(nx, ny):
    return [np.zeros((ny, nx)) for _ in range(6)]


def _one(x, y, ux, uy, uz, q=-1e-3):
    return {"x": np.array([x]), "y": np.array([y]), "ux": np.array([ux]), "uy": np.array([uy]),
            "uz": np.array([uz]), "q": np.array([q]), "tag": np.array([0], dtype=np.int32)}


def test_gyration_in_uniform_bz():
    nx = ny = 16; dx = dy = 0.1; dt = 0.03
    P = ob.make_params(lib, nx, ny, 8, 8, dx, dy, dt)
    pl = _zero_planes(nx, ny); pl[5][:] = 2.0
    p = _one(0.8, 0.8, 0.3, 0.0, 0.1)
    u0 = np.sqrt(0.3**2 + 0.1**2)
    gam = np.sqrt(1 + u0**2)
    ang = []
    for _ in range(50):
        ob.push_deposit(lib, P, pl, p)
        assert abs(np.sqrt(p["ux"][0]**2 + p["uy"][0]**2 + p["uz"][0]**2) - u0) < 1e-15   # |u| conserved
        assert abs(p["uz"][0] - 0.1) < 1e-16
        ang.append(np.arctan2(p["uy"][0], p["ux"][0]))
    dphi = np.diff(np.unwrap(ang))
    want = 2 * np.arctan(P.h * 2.0 / gam)        # rotation per step, SURVEY.md B.9 (q/m=-1: sign carried by h)
    assert np.allclose(dphi, -want, rtol=0, atol=1e-13)


def test_exb_drift():
    nx = ny = 16; dx = dy = 0.1; dt = 0.02
    P = ob.make_params(lib, nx, ny, 8, 8, dx, dy, dt)
    pl = _zero_planes(nx, ny); pl[5][:] = 1.0; pl[1][:] = 0.01          # Bz, Ey -> drift vx = Ey/Bz
    p = _one(0.8, 0.8, 0.0, 0.0, 0.0)
    xs = []
    x_unw, prev = 0.8, 0.8
    for _ in range(2000):
        ob.push_deposit(lib, P, pl, p)
        d = p["x"][0] - prev
        d -= P.box_x * np.round(d / P.box_x)
        x_unw += d; prev = p["x"][0]
        xs.append(x_unw)
    v = (xs[-1] - xs[0]) / (dt * (len(xs) - 1))
    assert abs(v - 0.01) < 2e-4


def test_wrap_and_tile_key_edges():
    nx, ny, tnx, tny = 32, 16, 8, 4; dx = dy = 0.1
    P = ob.make_params(lib, nx, ny, tnx, tny, dx, dy, 0.03)
    pl = _zero_planes(nx, ny)
    # leaves through the right/upper box edge and through the left/lower edge
    for (x, y, ux, uy) in ((3.199, 1.599, 0.5, 0.5), (0.001, 0.001, -0.5, -0.5)):
        p = _one(x, y, ux, uy, 0.0)
        ob.push_deposit(lib, P, pl, p)
        assert 0.0 <= p["x"][0] < P.box_x and 0.0 <= p["y"][0] < P.box_y
        k = ob.tile_keys(lib, P, p["x"], p["y"])[0]
        assert k == plasma.tile_of(p["x"], p["y"], nx, tnx, tny, dx, dy)[0]
        assert 0 <= k < (nx // tnx) * (ny // tny)
    x = np.nextafter(np.array([P.box_x]), 0.0); y = np.array([0.05])
    assert ob.tile_keys(lib, P, x, y)[0] in (nx // tnx - 1, 0)


def test_deposit_total_and_continuity():
    """Sum of Jx*dx*dy = q*vx per particle, and the discrete continuity equation holds to round-off:
    (rho1-rho0)/dt + div J = 0 at every node (zigzag is charge conserving)."""
    nx, ny, tnx, tny = 24, 16, 8, 8; dx, dy = 0.1, 0.125
    dt = 0.5 / np.sqrt(1 / dx**2 + 1 / dy**2)
    P = ob.make_params(lib, nx, ny, tnx, tny, dx, dy, dt)
    parts = plasma.load_uniform(nx, ny, tnx, tny, dx, dy, 9, u_th=0.6, seed=5)
    pl = [0.05 * np.random.default_rng(2).standard_normal((ny, nx)) for _ in range(6)]
    rho0 = ob.deposit_rho(lib, P, parts).astype(np.float64) / 2.0**52
    x0 = parts["x"].copy()
    J = ob.push_deposit(lib, P, pl, parts).astype(np.float64) / 2.0**52
    rho1 = ob.deposit_rho(lib, P, parts).astype(np.float64) / 2.0**52
    div = (J[0] - np.roll(J[0], 1, axis=1)) / dx + (J[1] - np.roll(J[1], 1, axis=0)) / dy
    res = (rho1 - rho0) / dt + div
    assert np.abs(res).max() < 1e-12 * np.abs(rho0).max() / dt
    g = np.sqrt(1 + parts["ux"]**2 + parts["uy"]**2 + parts["uz"]**2)
    assert abs(J[0].sum() * dx * dy - (parts["q"] * parts["ux"] / g).sum()) < 1e-13
    assert abs(J[2].sum() * dx * dy - (parts["q"] * parts["uz"] / g).sum()) < 1e-13
    assert (np.abs(parts["x"] - x0) > 0).any()


def test_sort_is_stable_and_counts_match():
    nx, ny, tnx, tny = 32, 32, 8, 8; dx = dy = 0.1; dt = 0.035
    P = ob.make_params(lib, nx, ny, tnx, tny, dx, dy, dt)
    parts = plasma.load_uniform(nx, ny, tnx, tny, dx, dy, 4, u_th=0.5, seed=9)
    pl = _zero_planes(nx, ny)
    before = {k: v.copy() for k, v in parts.items()}
    counts, _ = ob.pic_steps(lib, P, pl, parts, 3, do_sort=True)
    key = ob.tile_keys(lib, P, parts["x"], parts["y"])
    assert (np.diff(key) >= 0).all() and np.array_equal(np.bincount(key, minlength=16), counts)
    assert sorted(parts["tag"]) == sorted(before["tag"])
    # stability: replay unsorted, then one stable argsort per step must give the same slot order
    ref = {k: v.copy() for k, v in before.items()}
    pl2 = _zero_planes(nx, ny)
    for _ in range(3):
        ob.pic_steps(lib, P, pl2, ref, 1, do_sort=False)
        perm = np.argsort(ob.tile_keys(lib, P, ref["x"], ref["y"]), kind="stable")
        ref = {k: np.ascontiguousarray(v[perm]) for k, v in ref.items()}
    assert np.array_equal(ref["tag"], parts["tag"]) and np.array_equal(ref["x"], parts["x"])


def test_cold_plasma_oscillation_at_omega_p():
    """Uniform cold electrons with a sinusoidal velocity kick oscillate at omega_p = 1 (n0 = 1, q/m = -1)."""
    nx, ny, tnx, tny = 32, 8, 8, 8; dx = dy = 0.1
    dt = 0.5 / np.sqrt(2) * dx
    P = ob.make_params(lib, nx, ny, tnx, tny, dx, dy, dt)
    parts = plasma.load_uniform(nx, ny, tnx, tny, dx, dy, 16)
    parts["ux"] = 1e-3 * np.sin(2 * np.pi * parts["x"] / P.box_x)
    pl = _zero_planes(nx, ny)
    mode = np.exp(-2j * np.pi * (np.arange(nx) + 0.5) / nx)
    amp, t = [], []
    nsteps = int(2.2 * np.pi / dt)
    for s in range(nsteps):
        ob.pic_steps(lib, P, pl, parts, 1, do_sort=True)
        amp.append((pl[0].mean(axis=0) * mode).sum().imag); t.append((s + 1) * dt)
    amp = np.array(amp); t = np.array(t)
    zc = np.where(np.sign(amp[1:]) != np.sign(amp[:-1]))[0]
    t_cross = [t[i] - amp[i] * dt / (amp[i + 1] - amp[i]) for i in zc[:2]]
    assert len(t_cross) == 2
    assert abs((t_cross[1] - t_cross[0]) - np.pi) < 0.03 * np.pi      # half period = pi/omega_p


def test_two_stream_growth_rate_of_a_seeded_mode():
    """Two cold counter-streaming electron beams (u0 = +-0.2, each half the density): the seeded mode m = 3 of a 6.4 c/omega_p
    box grows at the rate of the cold-beam dispersion relation 1 = omega_b^2 [1/(w - kv)^2 + 1/(w + kv)^2] with
    omega_b^2 = omega_p^2 / (2 gamma0^3):  gamma^2 = omega_b^2 (sqrt(4x + 1) - (x + 1)),  x = (k v0)^2 / omega_b^2  ->  0.343."""
    nx, ny, tn, ppc, dx = 64, 8, 8, 64, 0.1
    dt = 0.5 / np.sqrt(2) * dx
    P = ob.make_params(lib, nx, ny, tn, tn, dx, dx, dt)
    parts = plasma.load_uniform(nx, ny, tn, tn, dx, dx, ppc, u_th=0.0, two_stream=0.2)
    m = 3
    k = 2 * np.pi * m / P.box_x
    parts["ux"] = parts["ux"] + 1e-5 * np.sin(k * parts["x"])
    pl = _zero_planes(nx, ny)
    mode = np.exp(-1j * k * (np.arange(nx) + 0.5) * dx)
    A, T = [], []
    for s in range(850):
        ob.pic_steps(lib, P, pl, parts, 1)
        A.append(abs((pl[0].mean(axis=0) * mode).sum()))
        T.append((s + 1) * dt)
    A, T = np.array(A), np.array(T)
    sel = (T > 12.0) & (A < 0.1 * A.max()) & (T < T[np.argmax(A)])
    assert sel.sum() > 100
    rate = np.polyfit(T[sel], np.log(A[sel]), 1)[0]
    g0 = np.sqrt(1.0 + 0.2 ** 2)
    wb2 = 0.5 / g0 ** 3
    x = (k * 0.2 / g0) ** 2 / wb2
    theory = np.sqrt(wb2 * (np.sqrt(4 * x + 1) - (x + 1)))
    assert abs(theory - 0.343) < 2e-3
    assert abs(rate - theory) < 0.05 * theory, (rate, theory)


def test_cell_key_sort_is_a_stable_refinement_of_the_tile_sort():
    """do_sort=2 (the GPU's slot order): keys non-decreasing, same tile counts and the same particle set per tile as the
    tile-key sort, and equal to one numpy stable argsort by cell key per step."""
    nx, ny, tnx, tny = 32, 24, 8, 4; dx = dy = 0.1; dt = 0.035
    P = ob.make_params(lib, nx, ny, tnx, tny, dx, dy, dt)
    parts = plasma.load_uniform(nx, ny, tnx, tny, dx, dy, 5, u_th=0.5, seed=4)
    a = {k: v.copy() for k, v in parts.items()}
    b = {k: v.copy() for k, v in parts.items()}
    c = {k: v.copy() for k, v in parts.items()}
    ca, _ = ob.pic_steps(lib, P, _zero_planes(nx, ny), a, 4, do_sort=1)
    cb, _ = ob.pic_steps(lib, P, _zero_planes(nx, ny), b, 4, do_sort=2)
    assert np.array_equal(ca, cb)
    kb = ob.cell_keys(lib, P, b["x"], b["y"])
    assert (np.diff(kb) >= 0).all()
    assert np.array_equal(ob.tile_keys(lib, P, b["x"], b["y"]), kb // (tnx * tny))
    ta, tb = ob.tile_keys(lib, P, a["x"], a["y"]), ob.tile_keys(lib, P, b["x"], b["y"])
    for t in range(len(ca)):
        assert sorted(a["tag"][ta == t]) == sorted(b["tag"][tb == t])
    pl = _zero_planes(nx, ny)
    for _ in range(4):
        ob.pic_steps(lib, P, pl, c, 1, do_sort=False)
        perm = np.argsort(ob.cell_keys(lib, P, c["x"], c["y"]), kind="stable")
        c = {k: np.ascontiguousarray(v[perm]) for k, v in c.items()}
    assert np.array_equal(c["tag"], b["tag"]) and np.array_equal(c["x"], b["x"])


def test_relativistic_gyration_gamma_10():
    """gamma = 10 (|u| = sqrt(99)): the Boris rotation conserves |u| to round-off, turns by 2 atan(h B / gamma) per step --
    the relativistic gyro-frequency omega_c / gamma to second order -- and the particle moves |v| dt = (|u|/gamma) dt per step."""
    nx = ny = 16; dx = dy = 0.1; dt = 0.03
    P = ob.make_params(lib, nx, ny, 8, 8, dx, dy, dt)
    pl = _zero_planes(nx, ny); pl[5][:] = 2.0
    u0 = np.sqrt(99.0)
    p = _one(0.8, 0.8, u0, 0.0, 0.0)
    gam = np.sqrt(1 + u0**2)
    assert abs(gam - 10.0) < 1e-14
    ang, step_len = [], []
    prev = (p["x"][0], p["y"][0])
    for _ in range(400):
        ob.push_deposit(lib, P, pl, p)
        assert abs(np.sqrt(p["ux"][0]**2 + p["uy"][0]**2 + p["uz"][0]**2) - u0) < 5e-14 * u0
        ang.append(np.arctan2(p["uy"][0], p["ux"][0]))
        d = np.array([p["x"][0] - prev[0], p["y"][0] - prev[1]])
        d -= np.array([P.box_x, P.box_y]) * np.round(d / np.array([P.box_x, P.box_y]))
        step_len.append(np.hypot(*d)); prev = (p["x"][0], p["y"][0])
    dphi = np.diff(np.unwrap(ang))
    want = 2 * np.arctan(P.h * 2.0 / gam)
    assert np.allclose(dphi, -want, rtol=0, atol=1e-12)
    assert abs(abs(want) / dt - 2.0 / gam) < 1e-5 * 2.0 / gam          # omega_c / gamma, up to the O(dt^2) phase error of Boris
    assert np.allclose(step_len, u0 / gam * dt, rtol=0, atol=1e-13)


def test_oblique_drift_continuity_with_corner_crossers():
    """A fast oblique beam at the CFL limit: most particles cross a cell face every step and many cross two (through a corner
    region) -- the zigzag split must still satisfy the discrete continuity equation at every node, and the total current must
    equal sum q v in all three components."""
    nx, ny, tnx, tny = 16, 24, 8, 8; dx, dy = 0.1, 0.1
    dt = 1.0 / np.sqrt(1 / dx**2 + 1 / dy**2)                           # cfl = 1: |v| dt up to 0.7 cells per axis
    P = ob.make_params(lib, nx, ny, tnx, tny, dx, dy, dt)
    parts = plasma.load_uniform(nx, ny, tnx, tny, dx, dy, 12, u_th=0.05, seed=21, drift=(3.0, -2.5, 1.0))
    pl = _zero_planes(nx, ny)
    crossed_both = 0
    for _ in range(6):
        rho0 = ob.deposit_rho(lib, P, parts).astype(np.float64) / 2.0**52
        c0 = (np.floor(parts["x"] / dx).astype(int), np.floor(parts["y"] / dy).astype(int))
        J = ob.push_deposit(lib, P, pl, parts).astype(np.float64) / 2.0**52
        rho1 = ob.deposit_rho(lib, P, parts).astype(np.float64) / 2.0**52
        c1 = (np.floor(parts["x"] / dx).astype(int), np.floor(parts["y"] / dy).astype(int))
        crossed_both += int(((c0[0] != c1[0]) & (c0[1] != c1[1])).sum())
        div = (J[0] - np.roll(J[0], 1, axis=1)) / dx + (J[1] - np.roll(J[1], 1, axis=0)) / dy
        res = (rho1 - rho0) / dt + div
        assert np.abs(res).max() < 1e-12 * np.abs(rho0).max() / dt
        g = np.sqrt(1 + parts["ux"]**2 + parts["uy"]**2 + parts["uz"]**2)
        for c, k in enumerate(("ux", "uy", "uz")):
            assert abs(J[c].sum() * dx * dy - (parts["q"] * parts[k] / g).sum()) < 1e-12
    print("corner crossers per step:", crossed_both / 6 / parts["x"].shape[0])
    assert crossed_both > 0.2 * 6 * parts["x"].shape[0]


def test_two_stream_total_momentum_stays_zero():
    """Counter-streaming beams exchange momentum only with each other (the fields stay electrostatic: no E x B momentum).  The
    scheme is charge-conserving, not momentum-conserving, so the net particle momentum is only nearly constant: through growth
    AND saturation it drifts by 3e-3 of the momentum of one beam (bound: 1e-2), while each beam loses most of its own."""
    nx, ny, tn, ppc, dx = 64, 8, 8, 64, 0.1
    dt = 0.5 / np.sqrt(2) * dx
    P = ob.make_params(lib, nx, ny, tn, tn, dx, dx, dt)
    parts = plasma.load_uniform(nx, ny, tn, tn, dx, dx, ppc, u_th=0.01, two_stream=0.2, seed=77)   # thermal spread: no mirror symmetry
    k = 2 * np.pi * 3 / P.box_x
    parts["ux"] = parts["ux"] + 1e-5 * np.sin(k * parts["x"]) * (parts["ux"] > 0)                    # only one beam is seeded
    beam = parts["ux"] > 0
    tags_beam = parts["tag"][beam].copy()
    mass = -parts["q"]                                                  # q/m = -1
    p_beam0 = (mass[beam] * parts["ux"][beam]).sum()
    pl = _zero_planes(nx, ny)
    net0 = np.array([(mass * parts["ux"]).sum(), (mass * parts["uy"]).sum()])      # sampling noise of the thermal spread
    worst = 0.0
    for s in range(1300):
        ob.pic_steps(lib, P, pl, parts, 1)
        if s % 10 == 0:
            net = np.array([(mass * parts["ux"]).sum(), (mass * parts["uy"]).sum()])
            worst = max(worst, np.abs(net - net0).max() / p_beam0)
    sel = np.isin(parts["tag"], tags_beam)                             # the same particles after 1300 sorts
    p_beam1 = (mass[sel] * parts["ux"][sel]).sum()
    print("net momentum / beam momentum, worst:", worst, " beam momentum kept:", p_beam1 / p_beam0)
    assert worst < 1e-2, worst
    assert p_beam1 < 0.9 * p_beam0                                      # saturation really happened: the beams slowed down


def test_direct_cic_mode_total_current_and_no_continuity():
    """deposit = 1 (SURVEY.md B.5 option a): the summed current still equals sum q v exactly (the weights of each lattice sum
    to one), one resting-cell particle lands on the four Ex points around it with the bilinear weights, and -- unlike the
    zigzag default -- the discrete continuity equation is violated (this is why the default is the charge-conserving one)."""
    nx, ny, tnx, tny = 24, 16, 8, 8; dx, dy = 0.1, 0.125
    dt = 0.5 / np.sqrt(1 / dx**2 + 1 / dy**2)
    P = ob.make_params(lib, nx, ny, tnx, tny, dx, dy, dt, deposit="cic")
    parts = plasma.load_uniform(nx, ny, tnx, tny, dx, dy, 9, u_th=0.6, seed=5)
    pl = _zero_planes(nx, ny)
    rho0 = ob.deposit_rho(lib, P, parts).astype(np.float64) / 2.0**52
    J = ob.push_deposit(lib, P, pl, parts).astype(np.float64) / 2.0**52
    rho1 = ob.deposit_rho(lib, P, parts).astype(np.float64) / 2.0**52
    g = np.sqrt(1 + parts["ux"]**2 + parts["uy"]**2 + parts["uz"]**2)
    for c, k in enumerate(("ux", "uy", "uz")):
        assert abs(J[c].sum() * dx * dy - (parts["q"] * parts[k] / g).sum()) < 1e-13
    div = (J[0] - np.roll(J[0], 1, axis=1)) / dx + (J[1] - np.roll(J[1], 1, axis=0)) / dy
    res = (rho1 - rho0) / dt + div
    assert np.abs(res).max() > 1e-6 * np.abs(rho0).max() / dt
    # one particle, no fields: x_mid = (3.3 + v dt / 2) dx ... weights on the Ex lattice (i + 1/2, j)
    p = _one(0.33 * 1.0, 0.40, 0.5, 0.0, 0.0, q=-2e-3)
    x0, y0 = p["x"][0], p["y"][0]
    J1 = ob.push_deposit(lib, P, pl, p).astype(np.float64) / 2.0**52
    v = 0.5 / np.sqrt(1.25)
    Xm, Ym = (x0 + 0.5 * v * dt) / dx, y0 / dy
    i, j = int(np.floor(Xm - 0.5)), int(np.floor(Ym))
    wx, wy = Xm - 0.5 - i, Ym - j
    want = np.zeros((ny, nx))
    amp = -2e-3 * v / (dx * dy)
    want[j, i], want[j, i + 1], want[j + 1, i], want[j + 1, i + 1] = amp * (1 - wx) * (1 - wy), amp * wx * (1 - wy), amp * (1 - wx) * wy, amp * wx * wy
    assert np.allclose(J1[0], want, rtol=0, atol=1e-15) and not J1[1].any()
