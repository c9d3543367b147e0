"""mini-pic b200: host-side Python mirror of the C ABI in include/minipic_b200.h (ctypes; no torch types cross it).

The reference is a C++ program (PIC_2D.cpp) with no plugin layer; its real host-side mirror is host/*.h (C++, the
reference's exact signatures) and drivers/pic2d.cpp.  This Python layer exists for the test/benchmark harness and for
torch.distributed plumbing (slab halos over NCCL): it only forwards to libminipic_b200.so and fails loudly when the CUDA
library or a device is missing -- there is no CPU fallback anywhere in the product path.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MPIC_LIB") or os.path.join(_HERE, "libminipic_b200.so")   # MPIC_LIB: kernel-variant A/B runs

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)
c_lp = C.POINTER(C.c_int64)


PEER_INFO_BYTES = 512   # sizeof(mpic_peer_info)
PARCEL_INFO_BYTES = 192  # sizeof(mpic_parcel_info)


class MpicError(RuntimeError):
    pass


class Params(C.Structure):
    _fields_ = [("nx", C.c_int), ("ny", C.c_int), ("tnx", C.c_int), ("tny", C.c_int), ("guard", C.c_int),
                ("scheme", C.c_int), ("fma_mode", C.c_int), ("rank", C.c_int), ("nranks", C.c_int),
                ("row0", C.c_int), ("nrows", C.c_int), ("device", C.c_int),
                ("dx", C.c_double), ("dy", C.c_double), ("dt", C.c_double), ("qm", C.c_double),
                ("max_particles", C.c_int64), ("deposit", C.c_int), ("reserved", C.c_int)]


class HaloDesc(C.Structure):
    _fields_ = [("send", C.c_void_p * 2), ("recv", C.c_void_p * 2), ("bytes", C.c_int64)]


class MigrDesc(C.Structure):
    _fields_ = [("send", C.c_void_p * 2), ("recv", C.c_void_p * 2), ("cap", C.c_int64), ("record_bytes", C.c_int64)]


_lib = None


def lib():
    """Load libminipic_b200.so (built in-tree by __graft_entry__.build()).  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MpicError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    L.mpic_last_error.restype = C.c_char_p
    L.mpic_stream.restype = C.c_void_p
    L.mpic_stream.argtypes = [C.c_void_p]
    L.mpic_kernel_launches.restype = C.c_int64
    L.mpic_kernel_launches.argtypes = [C.c_void_p]
    L.mpic_num_particles.restype = C.c_int64
    L.mpic_num_particles.argtypes = [C.c_void_p]
    L.mpic_create.argtypes = [C.POINTER(Params), C.POINTER(C.c_void_p)]
    for name in ("mpic_destroy", "mpic_sync", "mpic_step_b_half", "mpic_step_e_full", "mpic_push_deposit",
                 "mpic_sort_tiles", "mpic_fold_current", "mpic_step_fields_fused", "mpic_peer_field_step_fused"):
        getattr(L, name).argtypes = [C.c_void_p]
    for name in ("mpic_field_steps", "mpic_field_steps_fused", "mpic_pic_steps", "mpic_halo_pack", "mpic_halo_unpack"):
        getattr(L, name).argtypes = [C.c_void_p, C.c_int]
    L.mpic_migration_describe.argtypes = [C.c_void_p, C.POINTER(MigrDesc)]
    L.mpic_migration_counts.argtypes = [C.c_void_p, c_ip, c_ip]
    L.mpic_migration_absorb.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.mpic_upload_planes.argtypes = [C.c_void_p, C.POINTER(c_dp)]
    L.mpic_download_planes.argtypes = [C.c_void_p, C.POINTER(c_dp)]
    L.mpic_upload_tiles.argtypes = [C.c_void_p, c_dp, c_ip, C.c_int]
    L.mpic_download_tiles.argtypes = [C.c_void_p, c_dp, c_ip, C.c_int]
    L.mpic_halo_describe.argtypes = [C.c_void_p, C.POINTER(HaloDesc)]
    L.mpic_upload_particles.argtypes = [C.c_void_p, C.c_int64] + [c_dp] * 6 + [c_ip]
    L.mpic_download_particles.argtypes = [C.c_void_p, C.c_int64] + [c_dp] * 6 + [c_ip, c_lp]
    L.mpic_load_uniform.argtypes = [C.c_void_p, C.c_int] + [C.c_double] * 8 + [C.c_uint64, C.c_int]
    L.mpic_get_tile_counts.argtypes = [C.c_void_p, c_ip]
    L.mpic_cell_crossers.argtypes = [C.c_void_p, c_lp]
    L.mpic_get_particle_tile_index.argtypes = [C.c_void_p, c_ip]
    L.mpic_download_current.argtypes = [C.c_void_p, C.POINTER(c_lp)]
    L.mpic_diagnostics.argtypes = [C.c_void_p, c_dp]
    L.mpic_download_rho.argtypes = [C.c_void_p, c_lp]
    L.mpic_gauss_residual.argtypes = [C.c_void_p, C.c_double, c_dp]
    L.mpic_get_owner.argtypes = [C.c_void_p, c_ip]
    L.mpic_migrate_tile.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.mpic_pic_steps_host.argtypes = [C.c_void_p, C.c_int, C.POINTER(c_dp), C.c_int64] + [c_dp] * 6 + [c_ip, c_lp]
    L.mpic_update_b_half.argtypes = [c_dp, c_dp, c_dp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]
    L.mpic_update_e_full.argtypes = [c_dp, c_dp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]
    L.mpic_pack_send_buffer.argtypes = [c_dp, C.c_int, C.c_int, C.c_int, C.c_int, c_dp]
    L.mpic_update_guard_region.argtypes = [c_dp, C.c_int, C.c_int, C.c_int, C.c_int, c_dp]
    L.mpic_save_rank_data.argtypes = [C.c_char_p, C.c_int, c_ip, c_ip, c_ip, c_ip, C.c_int, C.c_int, C.c_int, c_dp]
    L.mpic_dump_fields.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
    L.mpic_write_params.argtypes = [C.c_char_p, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_double, C.c_double, C.c_int]
    for name in ("mpic_ring_enable_peer_access", "mpic_ring_refresh_guards", "mpic_ring_migrate"):
        getattr(L, name).argtypes = [C.POINTER(C.c_void_p), C.c_int]
    for name in ("mpic_ring_exchange", "mpic_ring_field_steps", "mpic_ring_pic_steps"):
        getattr(L, name).argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int]
    L.mpic_set_owner.argtypes = [C.c_void_p, c_ip]
    L.mpic_slab_rows.argtypes = [C.c_int, C.c_int, C.c_int, c_ip, c_ip]
    L.mpic_selftest_math.argtypes = [C.c_int64, C.c_uint64, c_lp]
    L.mpic_peer_export.argtypes = [C.c_void_p, C.c_char_p]
    L.mpic_peer_connect.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
    for name in ("mpic_peer_refresh_begin", "mpic_peer_refresh_end", "mpic_peer_field_step", "mpic_peer_step_begin",
                 "mpic_peer_step_end"):
        getattr(L, name).argtypes = [C.c_void_p]
    L.mpic_peer_trace.argtypes = [C.c_void_p, C.c_int, c_dp, c_lp]
    L.mpic_migrated_total.restype = C.c_int64
    L.mpic_migrated_total.argtypes = [C.c_void_p]
    L.mpic_peer_disconnect.argtypes = [C.c_void_p]
    L.mpic_partition_rows.argtypes = [C.c_int, c_dp, C.c_int, c_ip]
    L.mpic_rebalance_plan.argtypes = [C.c_int, C.c_int, c_lp, C.c_int64, C.c_double, C.c_double, c_ip, c_ip, c_dp]
    L.mpic_get_tile_row_counts.argtypes = [C.c_void_p, c_lp]
    L.mpic_reslab_pack.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p]
    L.mpic_reslab_apply.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int64, C.c_char_p, C.c_int, c_ip]
    L.mpic_reslab_finish.argtypes = [C.c_void_p]
    L.mpic_get_rows.argtypes = [C.c_void_p, c_ip, c_ip]
    L.mpic_ring_rebalance.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_double, C.c_double, c_ip, c_ip]
    L.mpic_dump_fields_async.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
    L.mpic_dump_wait.argtypes = [C.c_void_p]
    L.mpic_kinetic_energy_static.argtypes = [C.c_void_p] * 4 + [C.c_int64, C.c_double, c_dp, C.c_void_p]
    _lib = L
    return L


def _chk(rc):
    if rc < 0:
        raise MpicError(f"libminipic_b200 error {rc}: {lib().mpic_last_error().decode()}")
    return rc


def _dp(a):
    return a.ctypes.data_as(c_dp)


def _ip(a):
    return a.ctypes.data_as(c_ip)


def slab_rows(global_tile_rows, nranks, rank):
    r0, nr = C.c_int(), C.c_int()
    lib().mpic_slab_rows(global_tile_rows, nranks, rank, C.byref(r0), C.byref(nr))
    return r0.value, nr.value


def partition_rows(row_costs, nranks):
    """mpic_partition_rows: contiguous partition of the rows into `nranks` non-empty slabs minimising the most expensive
    slab (exact).  Returns [(row0, nrows)]."""
    c = np.ascontiguousarray(row_costs, dtype=np.float64)
    out = np.zeros(2 * nranks, dtype=np.int32)
    _chk(lib().mpic_partition_rows(int(c.shape[0]), _dp(c), int(nranks), _ip(out)))
    return [(int(out[2 * k]), int(out[2 * k + 1])) for k in range(nranks)]


def rebalance_plan(row_particles, cells_per_tile_row, old_rows, w_cell, min_gain=0.05):
    """mpic_rebalance_plan: (changed, [(row0, nrows)] per rank, (bottleneck cost before, after))."""
    rp = np.ascontiguousarray(row_particles, dtype=np.int64)
    old = np.ascontiguousarray(np.asarray(old_rows, dtype=np.int32).ravel())
    new = np.zeros_like(old)
    cost = np.zeros(2)
    rc = _chk(lib().mpic_rebalance_plan(int(rp.shape[0]), old.shape[0] // 2, rp.ctypes.data_as(c_lp), int(cells_per_tile_row),
                                        float(w_cell), float(min_gain), _ip(old), _ip(new), _dp(cost)))
    return bool(rc), [(int(new[2 * k]), int(new[2 * k + 1])) for k in range(old.shape[0] // 2)], (float(cost[0]), float(cost[1]))


KINETIC_STATIC_SCRATCH_BYTES = 9728          # MPIC_KINETIC_STATIC_SCRATCH_BYTES
KINETIC_STATIC_MAX_PARTICLES = 2**31 - 64    # MPIC_KINETIC_STATIC_MAX_PARTICLES


def kinetic_energy_static(ux_ptr, uy_ptr, uz_ptr, q_ptr, n, qm=-1.0, stream=None):
    """mpic_kinetic_energy_static: (kinetic energy, total charge) of n particles whose ux, uy, uz, q arrays sit in device
    memory the CALLER owns (raw device addresses, e.g. tensor.data_ptr()).  For callers that cannot allocate on the GPU --
    all of it is in their own pool, or a cudaMalloc between ring steps would stall the neighbours: the library allocates
    nothing, the reduction's work area is a 9.5 KB static device array shared by every call on the device, so do not overlap
    two calls.  Blocks until `stream` (a cudaStream_t address; None = default stream) has drained."""
    out = np.zeros(2)
    _chk(lib().mpic_kinetic_energy_static(C.c_void_p(ux_ptr), C.c_void_p(uy_ptr), C.c_void_p(uz_ptr), C.c_void_p(q_ptr),
                                          int(n), float(qm), _dp(out), C.c_void_p(stream)))
    return float(out[0]), float(out[1])


def courant_dt(dx, dy, factor=0.5):
    """dt = factor * dt_CFL, PIC_2D.cpp:71-72."""
    return factor * (1.0 / np.sqrt(1.0 / (dx * dx) + 1.0 / (dy * dy)))


PART_KEYS = ("x", "y", "ux", "uy", "uz", "q")


class _DevMem:
    """__cuda_array_interface__ carrier so torch can wrap library-owned device memory without a copy."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def device_bytes(ptr, nbytes, device):
    import torch
    return torch.as_tensor(_DevMem(ptr, nbytes), device=f"cuda:{device}")


class Sim:
    """Device-resident simulation (L1 API).  Mirrors the phases of the reference loop PIC_2D.cpp:171-420."""

    def __init__(self, nx, ny, tnx, tny, dx, dy, dt, guard=2, scheme=0, fma_mode=0, qm=-1.0, max_particles=0,
                 rank=0, nranks=1, device=0, rows=None, deposit="zigzag"):
        """rows=(row0, nrows): this rank's tile rows (default: the even split of mpic_slab_rows; the rebalancer in slab.py
        passes uneven ones).  deposit: "zigzag" (charge conserving, default) or "cic" (direct CIC, SURVEY.md B.5 option a)."""
        L = lib()
        gtr = ny // tny
        row0, nrows = rows if rows is not None else slab_rows(gtr, nranks, rank)
        self.p = Params(nx, ny, tnx, tny, guard, scheme, fma_mode, rank, nranks, row0, nrows, device, dx, dy, dt, qm,
                        max_particles, {"zigzag": 0, "cic": 1}[deposit], 0)
        self.h = C.c_void_p()
        _chk(L.mpic_create(C.byref(self.p), C.byref(self.h)))
        self.nx, self.ny, self.tnx, self.tny, self.guard = nx, ny, tnx, tny, guard
        self.dx, self.dy, self.dt = dx, dy, dt
        self.nyl = nrows * tny
        self.row0, self.nrows = row0, nrows
        self.gtc, self.gtr = nx // tnx, gtr
        self.ntiles_local = nrows * self.gtc

    def close(self):
        if self.h:
            lib().mpic_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- fields
    def upload_planes(self, planes):
        arr = [np.ascontiguousarray(a, dtype=np.float64) for a in planes]
        assert all(a.shape == (self.nyl, self.nx) for a in arr)
        ptrs = (c_dp * 6)(*[_dp(a) for a in arr])
        _chk(lib().mpic_upload_planes(self.h, ptrs))

    def download_planes(self, out=None):
        """out: optional six preallocated (nyl, nx) float64 arrays (e.g. views of pinned host memory) filled in place."""
        if out is None:
            out = [np.empty((self.nyl, self.nx)) for _ in range(6)]
        assert all(a.shape == (self.nyl, self.nx) and a.dtype == np.float64 and a.flags.c_contiguous for a in out)
        ptrs = (c_dp * 6)(*[_dp(a) for a in out])
        _chk(lib().mpic_download_planes(self.h, ptrs))
        return out

    def upload_tiles(self, grid6, gids):
        gids = np.ascontiguousarray(gids, dtype=np.int32)
        grid6 = np.ascontiguousarray(grid6, dtype=np.float64)
        assert grid6.shape == (len(gids), self.tny + 2 * self.guard, self.tnx + 2 * self.guard, 6)
        _chk(lib().mpic_upload_tiles(self.h, _dp(grid6), _ip(gids), len(gids)))

    def download_tiles(self, gids):
        gids = np.ascontiguousarray(gids, dtype=np.int32)
        out = np.empty((len(gids), self.tny + 2 * self.guard, self.tnx + 2 * self.guard, 6))
        _chk(lib().mpic_download_tiles(self.h, _dp(out), _ip(gids), len(gids)))
        return out

    def step_b_half(self):
        _chk(lib().mpic_step_b_half(self.h))

    def step_e_full(self):
        _chk(lib().mpic_step_e_full(self.h))

    def field_steps(self, n, fused=False):
        """n iterations of half-B / E / half-B (PIC_2D.cpp:171-396).  fused=True: one kernel per step (bit-identical)."""
        _chk((lib().mpic_field_steps_fused if fused else lib().mpic_field_steps)(self.h, n))

    def cell_crossers(self):
        """Particles that changed cell in the last push (call between push_deposit and sort_tiles)."""
        n = C.c_int64()
        _chk(lib().mpic_cell_crossers(self.h, C.byref(n)))
        return int(n.value)

    def step_fields_fused(self):
        """Phases A..F of one step in ONE kernel, with the -dt*J term when particles are present (single slab)."""
        _chk(lib().mpic_step_fields_fused(self.h))

    # ---- slab plumbing (nranks > 1); the torch tensors are zero-copy views of the library's device staging buffers
    def halo_tensors(self):
        """{"send": [to_lower, to_upper], "recv": [from_lower, from_upper]} as uint8 CUDA tensors."""
        d = HaloDesc()
        _chk(lib().mpic_halo_describe(self.h, C.byref(d)))
        dev = self.p.device
        return {"send": [device_bytes(d.send[k], d.bytes, dev) for k in range(2)],
                "recv": [device_bytes(d.recv[k], d.bytes, dev) for k in range(2)]}

    def halo_pack(self, which):
        _chk(lib().mpic_halo_pack(self.h, which))

    def halo_unpack(self, which):
        _chk(lib().mpic_halo_unpack(self.h, which))

    def migration_tensors(self):
        d = MigrDesc()
        _chk(lib().mpic_migration_describe(self.h, C.byref(d)))
        dev, nb = self.p.device, d.cap * d.record_bytes
        return {"send": [device_bytes(d.send[k], nb, dev) for k in range(2)],
                "recv": [device_bytes(d.recv[k], nb, dev) for k in range(2)], "record_bytes": int(d.record_bytes)}

    def migration_counts(self):
        """(leave[2], arrive[2]) of the last push; one stream sync."""
        lv, ar = (C.c_int * 2)(), (C.c_int * 2)()
        _chk(lib().mpic_migration_counts(self.h, lv, ar))
        return [lv[0], lv[1]], [ar[0], ar[1]]

    @property
    def migrated_total(self):
        return int(lib().mpic_migrated_total(self.h))

    def migration_absorb(self, n_from_lo, n_from_hi):
        _chk(lib().mpic_migration_absorb(self.h, n_from_lo, n_from_hi))

    # ---- peer-memory ring (nranks > 1): neighbours' device memory mapped, halos / current / leavers move inside the kernels
    def peer_export(self):
        """512-byte blob describing this handle's peer-visible memory (CUDA IPC handles); give it to the two neighbours."""
        buf = C.create_string_buffer(PEER_INFO_BYTES)
        _chk(lib().mpic_peer_export(self.h, buf))
        return buf.raw

    def peer_connect(self, lo_blob, hi_blob):
        _chk(lib().mpic_peer_connect(self.h, lo_blob, hi_blob))

    def peer_disconnect(self):
        _chk(lib().mpic_peer_disconnect(self.h))

    # ---- run-time re-partition (csrc/reslab.cu)
    def tile_row_counts(self):
        out = np.zeros(self.nrows, dtype=np.int64)
        _chk(lib().mpic_get_tile_row_counts(self.h, out.ctypes.data_as(c_lp)))
        return out

    def reslab_pack(self, row_lo, row_hi):
        """Pack global tile rows [row_lo, row_hi) of this slab (fields + particles) into an exported device parcel; returns
        the 192-byte blob for the new owner."""
        buf = C.create_string_buffer(PARCEL_INFO_BYTES)
        _chk(lib().mpic_reslab_pack(self.h, int(row_lo), int(row_hi), buf))
        return buf.raw

    def reslab_apply(self, row0, nrows, parcels, all_rows=None, max_particles=0):
        """Rebuild this handle IN PLACE on tile rows [row0, row0 + nrows) from the parcels that tile them."""
        blob = b"".join(parcels)
        ar = None if all_rows is None else np.ascontiguousarray(np.asarray(all_rows, dtype=np.int32).ravel())
        _chk(lib().mpic_reslab_apply(self.h, int(row0), int(nrows), int(max_particles), blob, len(parcels),
                                     None if ar is None else _ip(ar)))
        self.row0, self.nrows = int(row0), int(nrows)
        self.nyl = self.nrows * self.tny
        self.ntiles_local = self.nrows * self.gtc
        self.p.row0, self.p.nrows = self.row0, self.nrows

    def reslab_finish(self):
        _chk(lib().mpic_reslab_finish(self.h))

    def peer_refresh_begin(self):
        _chk(lib().mpic_peer_refresh_begin(self.h))

    def peer_refresh_end(self):
        _chk(lib().mpic_peer_refresh_end(self.h))

    def peer_trace(self, enable=-1):
        """Per-phase mean ms of the peer PIC step since the last reset (enable=1 starts/resets, 0 stops, -1 only reads)."""
        out = np.zeros(11)
        n = C.c_int64()
        _chk(lib().mpic_peer_trace(self.h, enable, _dp(out), C.byref(n)))
        names = ("wait_prev", "push_deposit", "handoff", "_3", "_4", "b_half_1", "_6", "e_full", "_8", "b_half_2", "counts_absorb_sort")
        return {k: float(v) for k, v in zip(names, out) if not k.startswith("_")}, int(n.value)

    def peer_field_step(self, fused=False):
        _chk((lib().mpic_peer_field_step_fused if fused else lib().mpic_peer_field_step)(self.h))

    def peer_step_begin(self):
        _chk(lib().mpic_peer_step_begin(self.h))

    def peer_step_end(self):
        _chk(lib().mpic_peer_step_end(self.h))

    def torch_stream(self):
        import torch
        return torch.cuda.ExternalStream(self.stream, device=self.p.device)

    # ---- particles
    def upload_particles(self, parts):
        n = parts["x"].shape[0]
        arr = [np.ascontiguousarray(parts[k], dtype=np.float64) for k in PART_KEYS]
        tag = np.ascontiguousarray(parts["tag"], dtype=np.int32)
        _chk(lib().mpic_upload_particles(self.h, n, *[_dp(a) for a in arr], _ip(tag)))
        self.sync()

    def download_particles(self, out=None):
        """out: optional dict of preallocated arrays with capacity >= num_particles (e.g. pinned host memory); the returned
        dict holds views of their first num_particles entries."""
        n = self.num_particles
        if out is None:
            out = {k: np.empty(n) for k in PART_KEYS}
            out["tag"] = np.empty(n, dtype=np.int32)
        else:
            assert all(out[k].shape[0] >= n and out[k].flags.c_contiguous for k in out)
            out = {k: v[:n] for k, v in out.items()}
        nout = C.c_int64()
        _chk(lib().mpic_download_particles(self.h, n, *[_dp(out[k]) for k in PART_KEYS], _ip(out["tag"]), C.byref(nout)))
        assert nout.value == n
        return out

    @property
    def num_particles(self):
        return lib().mpic_num_particles(self.h)

    def load_uniform(self, ppc, u_th=0.0, drift=(0.0, 0.0, 0.0), two_stream=0.0, density=1.0, band=(0.0, 1.0),
                     seed=12345, append=False):
        _chk(lib().mpic_load_uniform(self.h, ppc, u_th, drift[0], drift[1], drift[2], two_stream, density, band[0], band[1],
                                     seed, 1 if append else 0))

    def push_deposit(self):
        _chk(lib().mpic_push_deposit(self.h))

    def fold_current(self):
        _chk(lib().mpic_fold_current(self.h))

    def sort_tiles(self):
        _chk(lib().mpic_sort_tiles(self.h))

    def pic_steps(self, n):
        _chk(lib().mpic_pic_steps(self.h, n))

    def pic_steps_host(self, n, planes, parts):
        """The host-buffer entry point (e2e path): planes and parts are updated in place (numpy or pinned torch memory)."""
        ptrs = (c_dp * 6)(*[C.cast(a.ctypes.data, c_dp) for a in planes])
        nout = C.c_int64()
        _chk(lib().mpic_pic_steps_host(self.h, n, ptrs, parts["x"].shape[0], *[_dp(parts[k]) for k in PART_KEYS],
                                       _ip(parts["tag"]), C.byref(nout)))
        return nout.value

    def tile_counts(self):
        out = np.empty(self.ntiles_local, dtype=np.int32)
        _chk(lib().mpic_get_tile_counts(self.h, _ip(out)))
        return out

    def particle_tile_index(self):
        out = np.empty(self.num_particles, dtype=np.int32)
        _chk(lib().mpic_get_particle_tile_index(self.h, _ip(out)))
        return out

    def download_current(self):
        out = np.empty((3, self.nyl, self.nx), dtype=np.int64)
        ptrs = (c_lp * 3)(*[out[k].ctypes.data_as(c_lp) for k in range(3)])
        _chk(lib().mpic_download_current(self.h, ptrs))
        return out

    def download_rho(self):
        """rho of the stored particles, int64 fixed point (value * 2^52), node lattice."""
        out = np.empty((self.nyl, self.nx), dtype=np.int64)
        _chk(lib().mpic_download_rho(self.h, out.ctypes.data_as(c_lp)))
        return out

    def diagnostics(self):
        out = np.zeros(5)
        _chk(lib().mpic_diagnostics(self.h, _dp(out)))
        return {"field_energy": out[0], "kinetic_energy": out[1], "gauss_residual": out[2], "charge": out[3],
                "particles": int(out[4])}

    def gauss_residual(self, mean_rho):
        """max |div E - rho + mean_rho| over this slab, mean_rho = GLOBAL mean charge density (multi-slab Gauss check)."""
        out = C.c_double()
        _chk(lib().mpic_gauss_residual(self.h, float(mean_rho), C.byref(out)))
        return float(out.value)

    def dump_fields(self, filename, rank=-1):
        """Phase H: fields_rank_<r>_step_<s>.h5 in the reference layout (HDF5_output.cpp:10-79)."""
        _chk(lib().mpic_dump_fields(self.h, os.fsencode(filename), rank))

    def dump_fields_async(self, filename, rank=-1):
        """Phase H off the step path: patches formed on the device, D2H on a side stream, file written by a worker thread."""
        _chk(lib().mpic_dump_fields_async(self.h, os.fsencode(filename), rank))

    def dump_wait(self):
        _chk(lib().mpic_dump_wait(self.h))

    # ---- ownership
    def owner(self):
        out = np.empty(self.gtr * self.gtc, dtype=np.int32)
        _chk(lib().mpic_get_owner(self.h, _ip(out)))
        return out

    def set_owner(self, owner):
        _chk(lib().mpic_set_owner(self.h, _ip(np.ascontiguousarray(owner, dtype=np.int32))))

    def migrate_tile(self, gid, dest):
        _chk(lib().mpic_migrate_tile(self.h, gid, dest))

    # ---- misc
    def sync(self):
        _chk(lib().mpic_sync(self.h))

    @property
    def stream(self):
        return lib().mpic_stream(self.h)

    @property
    def kernel_launches(self):
        return lib().mpic_kernel_launches(self.h)


# ---- L0 compat (one tile, host buffers): Field_update.h:36,54 and Auxiliar_functions.h:68-69
def selftest_math(n, seed=1):
    """Mismatches between the push kernel's branch-free 1/sqrt, a/b sequences and the compiler's IEEE ones (must be 0)."""
    bad = C.c_int64(-1)
    _chk(lib().mpic_selftest_math(int(n), int(seed), C.byref(bad)))
    return bad.value


def update_b_half(E3, Bold3, Bnew3, tnx, tny, guard, dt, dx, dy):
    _chk(lib().mpic_update_b_half(_dp(E3), _dp(Bold3), _dp(Bnew3), tnx, tny, guard, dt, dx, dy))


def update_e_full(E3, B3, tnx, tny, guard, dt, dx, dy):
    _chk(lib().mpic_update_e_full(_dp(E3), _dp(B3), tnx, tny, guard, dt, dx, dy))


def pack_send_buffer(grid6, tnx, tny, guard, d):
    out = np.empty((lib().mpic_strip_cells(d, tnx, tny, guard), 6))
    _chk(lib().mpic_pack_send_buffer(_dp(grid6), tnx, tny, guard, d, _dp(out)))
    return out


def update_guard_region(grid6, tnx, tny, guard, d, strip):
    _chk(lib().mpic_update_guard_region(_dp(grid6), tnx, tny, guard, d, _dp(np.ascontiguousarray(strip))))


def save_rank_data(filename, gids, tile_rows, tile_cols, current_rank, tnx, tny, guard, grid6):
    """saveRankData (HDF5_output.h:11) on host arrays; works without a GPU."""
    a = [np.ascontiguousarray(v, dtype=np.int32) for v in (gids, tile_rows, tile_cols, current_rank)]
    grid6 = np.ascontiguousarray(grid6, dtype=np.float64)
    assert grid6.shape == (len(a[0]), tny + 2 * guard, tnx + 2 * guard, 6)
    _chk(lib().mpic_save_rank_data(os.fsencode(filename), len(a[0]), *[_ip(v) for v in a], tnx, tny, guard, _dp(grid6)))


def write_params(path, box_x, box_y, nx, ny, guard, tnx, tny, sim_time, dt, total_steps):
    _chk(lib().mpic_write_params(os.fsencode(path), box_x, box_y, nx, ny, guard, tnx, tny, sim_time, dt, total_steps))


class SlabRing:
    """N slab handles driven by ONE host process with CUDA peer copies (csrc/ring.cu) -- what the C++ driver uses for
    --gpus N.  sims[r] must be rank r of len(sims)."""

    def __init__(self, sims):
        self.sims = list(sims)
        self.arr = (C.c_void_p * len(self.sims))(*[s.h for s in self.sims])
        self.n = len(self.sims)
        _chk(lib().mpic_ring_enable_peer_access(self.arr, self.n))

    def refresh_guards(self):
        _chk(lib().mpic_ring_refresh_guards(self.arr, self.n))

    def exchange(self, which):
        _chk(lib().mpic_ring_exchange(self.arr, self.n, which))

    def field_steps(self, nsteps):
        _chk(lib().mpic_ring_field_steps(self.arr, self.n, nsteps))

    def pic_steps(self, nsteps):
        _chk(lib().mpic_ring_pic_steps(self.arr, self.n, nsteps))

    def rebalance(self, w_cell=8.0, min_gain=0.05):
        """mpic_ring_rebalance: cost-driven re-partition of the tile rows, handles rebuilt in place.  Returns (changed, rows)."""
        changed = C.c_int(0)
        rows = np.zeros(2 * self.n, dtype=np.int32)
        _chk(lib().mpic_ring_rebalance(self.arr, self.n, float(w_cell), float(min_gain), C.byref(changed), _ip(rows)))
        for k, s in enumerate(self.sims):
            s.row0, s.nrows = int(rows[2 * k]), int(rows[2 * k + 1])
            s.nyl, s.ntiles_local = s.nrows * s.tny, s.nrows * s.gtc
            s.p.row0, s.p.nrows = s.row0, s.nrows
        return bool(changed.value), [(int(rows[2 * k]), int(rows[2 * k + 1])) for k in range(self.n)]

    def sync(self):
        for s in self.sims:
            s.sync()


class PeerRing:
    """Several slab handles of ONE process (one per GPU, or all on one GPU in the tests) connected as a peer-memory ring:
    the split begin/end calls let a single host thread drive all of them (mpic_peer_*; csrc/peer.cu)."""

    def __init__(self, sims):
        self.sims = list(sims)
        self.connect()

    def connect(self):
        n = len(self.sims)
        blobs = [s.peer_export() for s in self.sims]
        for r, s in enumerate(self.sims):
            s.peer_connect(blobs[(r - 1) % n], blobs[(r + 1) % n])

    def rebalance(self, w_cell=8.0, min_gain=0.05):
        """Same re-partition as SlabRing.rebalance for a peer-memory ring inside one process: unmap, re-partition, re-map."""
        ring = SlabRing(self.sims)
        changed, rows = ring.rebalance(w_cell, min_gain)
        if changed:
            self.connect()
            self.refresh_guards()
        return changed, rows

    def refresh_guards(self):
        for s in self.sims:
            s.peer_refresh_begin()
        for s in self.sims:
            s.peer_refresh_end()

    def field_steps(self, nsteps):
        for _ in range(nsteps):
            for s in self.sims:
                s.peer_field_step()

    def pic_steps(self, nsteps):
        for _ in range(nsteps):
            for s in self.sims:
                s.peer_step_begin()
            for s in self.sims:
                s.peer_step_end()

    def sync(self):
        for s in self.sims:
            s.sync()
