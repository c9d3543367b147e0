"""TEST INFRASTRUCTURE: a CPU stand-in with the slab interface of `Sim`, driven by the CPU oracle.

Purpose: run the PRODUCT's multi-rank orchestration (slab.SlabStepper + slab.Ring over torch.distributed/gloo,
world_size 2+) on a box without GPUs.  Only the per-rank compute is replaced (by oracle/pic_oracle.c and yee_oracle.c);
ordering of phases, ring neighbours, message matching, the add-vs-copy halo semantics and the leaver-count protocol are
the product code under test.

Each rank holds GLOBAL-size arrays in which only its slab rows (+ 2 guard rows each side) are meaningful; everything
else is NaN, so any read outside what the protocol delivered poisons the result and the final bit-exact comparison
with the single-rank oracle fails.
"""
import numpy as np
import torch
from oracle import bindings as ob

G = 2          # guard rows
HDR = 16       # int64 header values per message, [0] = leaver count (same idea as the library's staging header)
KEYS = ("x", "y", "ux", "uy", "uz", "q")


class CpuSlabSim:
    def __init__(self, nx, ny, tnx, tny, dx, dy, dt, rank, nranks, row0, nrows):
        self.lib = ob.load_oracle()
        self.nx, self.ny, self.tny, self.dx, self.dy, self.dt = nx, ny, tny, dx, dy, dt
        self.tnx, self.gtc, self.row0, self.nrows = tnx, nx // tnx, row0, nrows
        self.P = ob.make_params(self.lib, nx, ny, tnx, tny, dx, dy, dt)
        self.cy0, self.nyl = row0 * tny, nrows * tny
        self.rows_int = np.arange(self.cy0, self.cy0 + self.nyl)
        self.rows_lo = np.arange(self.cy0 - G, self.cy0) % ny                       # guard rows below
        self.rows_hi = np.arange(self.cy0 + self.nyl, self.cy0 + self.nyl + G) % ny   # guard rows above
        self.planes = [np.full((ny, nx), np.nan) for _ in range(6)]
        self.J = np.zeros((3, ny, nx), dtype=np.int64)
        self.parts = {k: np.zeros(0) for k in KEYS}
        self.parts["tag"] = np.zeros(0, dtype=np.int32)
        nb = (HDR + 3 * G * nx) * 8
        self._send = [torch.zeros(nb, dtype=torch.uint8) for _ in range(2)]
        self._recv = [torch.zeros(nb, dtype=torch.uint8) for _ in range(2)]
        cap = 7 * 8 * max(64, 4 * nx * 64)
        self._msend = [torch.zeros(cap, dtype=torch.uint8) for _ in range(2)]
        self._mrecv = [torch.zeros(cap, dtype=torch.uint8) for _ in range(2)]
        self.leave = [0, 0]
        self.left_mask = None
        self.arrivals = []

    # ---- the Sim methods slab.rebalance() uses
    def tile_counts(self):
        cx = np.floor(self.parts["x"] * (1.0 / self.dx)).astype(np.int64) // self.tnx
        cy = np.floor(self.parts["y"] * (1.0 / self.dy)).astype(np.int64) // self.tny - self.row0
        return np.bincount(cy * self.gtc + cx, minlength=self.nrows * self.gtc).astype(np.int32)

    def download_planes(self):
        return [a[self.rows_int].copy() for a in self.planes]

    def download_particles(self):
        return {k: v.copy() for k, v in self.parts.items()}

    def upload_planes(self, local_planes):
        for a, b in zip(self.planes, local_planes):
            a[self.rows_int] = b

    def upload_particles(self, parts):
        self.set_particles(parts)

    # ---- the re-partition interface slab.rebalance_device() drives (product: csrc/reslab.cu).  Here a parcel is a pickled blob
    # of the packed rows instead of a CUDA-IPC handle; the protocol -- who packs which rows for whom, which blobs a rank applies,
    # rows changing in place -- is the product's.
    def tile_row_counts(self):
        return self.tile_counts().reshape(self.nrows, self.gtc).sum(axis=1).astype(np.int64)

    def reslab_pack(self, row_lo, row_hi):
        import pickle
        assert self.row0 <= row_lo < row_hi <= self.row0 + self.nrows
        cy = np.floor(self.parts["y"] * (1.0 / self.dy)).astype(np.int64) // self.tny
        sel = (cy >= row_lo) & (cy < row_hi)
        rows = np.arange(row_lo * self.tny, row_hi * self.tny)
        return pickle.dumps({"rows": (row_lo, row_hi), "f": [a[rows].copy() for a in self.planes],
                             "p": {k: v[sel].copy() for k, v in self.parts.items()}})

    def peer_disconnect(self):
        pass

    def reslab_apply(self, row0, nrows, parcels, all_rows=None, max_particles=0):
        import pickle
        ps = sorted((pickle.loads(b) for b in parcels), key=lambda d: d["rows"][0])
        assert ps[0]["rows"][0] == row0 and ps[-1]["rows"][1] == row0 + nrows
        assert all(a["rows"][1] == b["rows"][0] for a, b in zip(ps, ps[1:])), "parcels must tile the new rows"
        self.__init__(self.nx, self.ny, self.tnx, self.tny, self.dx, self.dy, self.dt, 0, 0, row0, nrows)   # in place: same object
        for d in ps:
            rows = np.arange(d["rows"][0] * self.tny, d["rows"][1] * self.tny)
            for a, b in zip(self.planes, d["f"]):
                a[rows] = b
        self.parts = {k: np.concatenate([d["p"][k] for d in ps]) for k in ps[0]["p"]}
        self.all_rows = all_rows

    def reslab_finish(self):
        pass

    # ---- state in / out
    def set_interior(self, global_planes):
        for a, b in zip(self.planes, global_planes):
            a[self.rows_int] = b[self.rows_int]

    def set_particles(self, parts):
        cy = np.floor(parts["y"] * (1.0 / self.dy)).astype(np.int64)
        keep = (cy >= self.cy0) & (cy < self.cy0 + self.nyl)
        self.parts = {k: np.ascontiguousarray(v[keep]) for k, v in parts.items()}

    def _poison(self, first):
        """after a sweep only the interior rows of the updated planes are valid; guards are stale until exchanged"""
        keep = np.zeros(self.ny, dtype=bool)
        keep[self.rows_int] = True
        for p in range(first, first + 3):
            self.planes[p][~keep] = np.nan

    # ---- Sim slab interface
    def halo_tensors(self):
        return {"send": self._send, "recv": self._recv}

    def _body(self, t, dtype):
        return t.numpy().view(np.int64)[HDR:].view(dtype).reshape(3, G, self.nx)

    def halo_pack(self, which):
        if which == 2:
            self._body(self._send[0], np.int64)[:] = self.J[:, self.rows_lo]
            self._body(self._send[1], np.int64)[:] = self.J[:, self.rows_hi]
            for d in range(2):
                self._send[d].numpy().view(np.int64)[0] = self.leave[d]
        else:
            for p in range(3):
                self._body(self._send[0], np.float64)[p] = self.planes[3 * which + p][self.rows_int[:G]]
                self._body(self._send[1], np.float64)[p] = self.planes[3 * which + p][self.rows_int[-G:]]

    def halo_unpack(self, which):
        if which == 2:
            self.J[:, self.rows_int[:G]] += self._body(self._recv[0], np.int64)
            self.J[:, self.rows_int[-G:]] += self._body(self._recv[1], np.int64)
            self.J[:, self.rows_lo] = 0
            self.J[:, self.rows_hi] = 0
        else:
            for p in range(3):
                self.planes[3 * which + p][self.rows_lo] = self._body(self._recv[0], np.float64)[p]
                self.planes[3 * which + p][self.rows_hi] = self._body(self._recv[1], np.float64)[p]

    def push_deposit(self):
        self.J += ob.push_deposit(self.lib, self.P, self.planes, self.parts)
        for k in KEYS:
            assert np.isfinite(self.parts[k]).all(), "push read a cell the exchange never delivered"
        outside = np.ones(self.ny, dtype=bool)
        outside[np.concatenate([self.rows_int, self.rows_lo, self.rows_hi])] = False
        assert not self.J[:, outside].any(), "deposit outside slab + guard rows"
        cy = np.floor(self.parts["y"] * (1.0 / self.dy)).astype(np.int64)
        gone = (cy < self.cy0) | (cy >= self.cy0 + self.nyl)
        down = gone & (cy == (self.cy0 - 1) % self.ny)
        up = gone & ~down
        self.left_mask = gone
        for d, mask in enumerate((down, up)):
            rec = np.stack([self.parts[k][mask] for k in KEYS] + [self.parts["tag"][mask].astype(np.int64).view(np.float64)], axis=1)
            self._msend[d].numpy().view(np.float64)[: rec.size] = rec.ravel()
            self.leave[d] = int(mask.sum())

    def step_b_half(self):
        with np.errstate(invalid="ignore"):
            ob.b_sweep(self.lib, self.planes, self.dt, self.dx, self.dy)
        self._poison(3)

    def step_e_full(self):
        with np.errstate(invalid="ignore"):
            ob.e_sweep(self.lib, self.planes, self.dt, self.dx, self.dy)
            ob.apply_current(self.lib, self.P, self.planes, self.J)
        self.J[:] = 0
        self._poison(0)

    def migration_tensors(self):
        return {"send": self._msend, "recv": self._mrecv, "record_bytes": 56}

    def migration_counts(self):
        arrive = [int(self._recv[d].numpy().view(np.int64)[0]) for d in range(2)]
        return list(self.leave), arrive

    def migration_absorb(self, n_lo, n_hi):
        for d, n in enumerate((n_lo, n_hi)):
            self.arrivals.append(self._mrecv[d].numpy().view(np.float64)[: n * 7].reshape(n, 7).copy())

    def sort_tiles(self):
        keep = ~self.left_mask if self.left_mask is not None else np.ones(self.parts["x"].shape[0], dtype=bool)
        arr = np.concatenate(self.arrivals) if self.arrivals else np.zeros((0, 7))
        new = {k: np.concatenate([self.parts[k][keep], arr[:, i]]) for i, k in enumerate(KEYS)}
        new["tag"] = np.concatenate([self.parts["tag"][keep], np.ascontiguousarray(arr[:, 6]).view(np.int64).astype(np.int32)])
        key = ob.tile_keys(self.lib, self.P, new["x"], new["y"])
        perm = np.argsort(key, kind="stable")
        self.parts = {k: np.ascontiguousarray(v[perm]) for k, v in new.items()}
        self.arrivals, self.left_mask, self.leave = [], None, [0, 0]
