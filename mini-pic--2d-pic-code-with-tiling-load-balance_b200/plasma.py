"""Synthetic plasma loaders (no reference counterpart: the reference has no particles).

Counter-based and partition independent: particle `tag` (global ordinal at load time, the key of `Particle::tag`,
Auxiliar_functions.h:17) fully determines its initial state through a splitmix64 hash, so any rank can generate any
subset.  Only +,* on dyadic rationals are used (Irwin-Hall sum of 12 uniforms for the Maxwellian), so the host loader
here and the device loader kernel (csrc/loader.cu) produce identical bits.

Conventions: c = 1, omega_p = 1, electrons q/m = -1, reference density n0 = 1, macro-charge
q_p = -n0*dx*dy/ppc, immobile neutralising ions.  Particles are laid out tile-major (tile gid ascending, then cell
row-major inside the tile, then sub-lattice index), i.e. already sorted for the first push.
"""
import numpy as np

_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M
    z = x
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M
    return z ^ (z >> np.uint64(31))


def uniform01(seed, tag, stream):
    """53-bit uniform in [0,1) from (seed, tag, stream)."""
    with np.errstate(over="ignore"):
        k = splitmix64(np.uint64(seed) ^ splitmix64(tag.astype(np.uint64) * np.uint64(64) + np.uint64(stream)))
    return (k >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def gaussian_ih(seed, tag, stream0):
    """Irwin-Hall N(0,1) approximation: sum of 12 uniforms - 6 (exactly reproducible on any IEEE machine)."""
    s = np.zeros(tag.shape[0])
    for k in range(12):
        s = s + uniform01(seed, tag, stream0 + k)
    return s - 6.0


def tile_major_cells(nx, ny, tnx, tny, rows=None):
    """(cx, cy) of every cell in tile-major order; `rows` = (first, end) tile rows of a slab (default: all)."""
    gtc, gtr = nx // tnx, ny // tny
    r0, r1 = (0, gtr) if rows is None else rows
    trow, tcol, cj, ci = np.meshgrid(np.arange(r0, r1), np.arange(gtc), np.arange(tny), np.arange(tnx), indexing="ij")
    return (tcol * tnx + ci).ravel(), (trow * tny + cj).ravel()


def load_uniform(nx, ny, tnx, tny, dx, dy, ppc, u_th=0.0, drift=(0.0, 0.0, 0.0), seed=12345, two_stream=0.0,
                 density=1.0, rows=None, band=None, tag0=0):
    """Quiet-start uniform population: ppc particles per cell on a regular sub-lattice, Maxwellian momenta.

    two_stream = u0 > 0 gives two counter-streaming halves (even/odd sub-lattice index -> ux = +u0 / -u0).
    band = (y_lo, y_hi) as fractions of box_y restricts the population to cells with y_lo <= cy/ny < y_hi (beam).
    Returns a dict of x,y,ux,uy,uz,q (float64) and tag (int32), tile-major."""
    cx, cy = tile_major_cells(nx, ny, tnx, tny, rows)
    if band is not None:
        keep = (cy >= int(round(band[0] * ny))) & (cy < int(round(band[1] * ny)))
        cx, cy = cx[keep], cy[keep]
    nsx = int(np.ceil(np.sqrt(ppc)))
    nsy = (ppc + nsx - 1) // nsx
    k = np.arange(ppc)
    sx, sy = (k % nsx + 0.5) / nsx, (k // nsx + 0.5) / nsy
    x = ((cx[:, None] + sx[None, :]) * dx).ravel()
    y = ((cy[:, None] + sy[None, :]) * dy).ravel()
    # tag: global cell-major ordinal (independent of tiling and of the slab the cell lives on)
    tag = ((cy.astype(np.int64) * nx + cx)[:, None] * ppc + k[None, :]).ravel() + tag0
    n = x.shape[0]
    out = {"x": x, "y": y}
    for j, name in enumerate(("ux", "uy", "uz")):
        u = np.full(n, float(drift[j]))
        if u_th > 0.0:
            u = u + u_th * gaussian_ih(seed, tag, 16 * j)
        out[name] = u
    if two_stream:
        sign = np.where(np.tile(k, cx.shape[0]) % 2 == 0, 1.0, -1.0)
        out["ux"] = out["ux"] + sign * two_stream
    out["q"] = np.full(n, -density * dx * dy / ppc)
    out["tag"] = tag.astype(np.int32)
    return out


def concat_sorted(parts_list, nx, ny, tnx, tny, dx, dy):
    """Merge populations and restore tile-major order with a STABLE sort by tile gid."""
    cat = {k: np.concatenate([p[k] for p in parts_list]) for k in parts_list[0]}
    key = tile_of(cat["x"], cat["y"], nx, tnx, tny, dx, dy)
    perm = np.argsort(key, kind="stable")
    return {k: np.ascontiguousarray(v[perm]) for k, v in cat.items()}


def tile_of(x, y, nx, tnx, tny, dx, dy):
    """Tile gid of stored positions: floor(x * (1/dx)) with the reciprocal rounded once (DESIGN.md arithmetic contract);
    gid = (cy // tny) * globalTileCols + cx // tnx (Auxiliar_functions.cpp:44-46)."""
    cx = np.floor(x * (1.0 / dx)).astype(np.int64)
    cy = np.floor(y * (1.0 / dy)).astype(np.int64)
    return ((cy // tny) * (nx // tnx) + cx // tnx).astype(np.int32)
