"""torchrun entry (one process per rank): K PIC steps on N slabs, rank 0 checks bit-exactness vs the CPU oracle.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tests/run_slab_nccl.py
MPIC_TEST_TRANSPORT = nccl (default; one GPU per rank) | peer (CUDA-IPC peer ring) | staged (SlabStepper over gloo with host
staging).  On a box with fewer GPUs than ranks every rank uses cuda:0 and the process group is gloo: CUDA IPC works between
processes on one device, so the REAL two-process peer path (IPC mapping, .sys-scope flags between processes) still runs."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
from util import PKG, by_tag, copy_parts  # noqa: E402


class HostStagedRing:
    """slab.Ring's interface over gloo: device staging buffers -> host -> isend/irecv (tagged) -> host -> device."""

    def __init__(self, rank, world):
        self.rank, self.world = rank, world
        self.lo, self.hi = (rank - 1) % world, (rank + 1) % world

    def sendrecv(self, send_lo, send_hi, recv_lo, recv_hi):
        reqs, back = [], []
        for t, dst, tag in ((send_lo, self.lo, 0), (send_hi, self.hi, 1)):
            if t is not None:
                reqs.append(dist.isend(t.cpu(), dst, tag=tag))
        # the lower neighbour's "to upper" (tag 1) lands in my recv_lo; the upper neighbour's "to lower" (tag 0) in recv_hi
        for t, src, tag in ((recv_hi, self.hi, 0), (recv_lo, self.lo, 1)):
            if t is not None:
                h = torch.empty(t.shape, dtype=t.dtype)
                reqs.append(dist.irecv(h, src, tag=tag))
                back.append((t, h))
        for r in reqs:
            r.wait()
        for t, h in back:
            t.copy_(h)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    transport = os.environ.get("MPIC_TEST_TRANSPORT", "nccl")
    one_device = torch.cuda.device_count() < world
    if one_device:
        local = 0
        assert transport != "nccl", "NCCL needs one GPU per rank"
    torch.cuda.set_device(local)
    if one_device or transport == "staged":
        dist.init_process_group("gloo")
    else:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    m = importlib.import_module(PKG)
    plasma = importlib.import_module(PKG + ".plasma")
    slab = importlib.import_module(PKG + ".slab")
    nx, ny, tn, ppc, steps = 64, 32 * world, 16, 8, 10
    dx = dy = 0.1
    dt = m.courant_dt(dx, dy)
    parts = plasma.load_uniform(nx, ny, tn, tn, dx, dy, ppc, u_th=0.7, seed=21, two_stream=0.2)
    rng = np.random.default_rng(21)
    planes = [0.02 * rng.standard_normal((ny, nx)) for _ in range(6)]
    s = m.Sim(nx, ny, tn, tn, dx, dy, dt, max_particles=parts["x"].shape[0], rank=rank, nranks=world, device=local)
    r0, r1 = s.row0 * tn, (s.row0 + s.nrows) * tn
    with torch.cuda.stream(s.torch_stream()):
        s.upload_planes([a[r0:r1] for a in planes])
        s.upload_particles(parts)
        if transport == "peer":
            st = slab.PeerStepper(s, rank, world)
        else:
            st = slab.SlabStepper(s, HostStagedRing(rank, world) if transport == "staged" else slab.Ring(rank, world))
        st.refresh_guards()
        for _ in range(steps):
            st.step()
        fsteps = 0
        if transport == "peer":                     # field-only steps on top: fused kernel (one flag exchange per step) mixed with the sweeps
            for _ in range(3):
                s.peer_field_step(fused=True)
            s.peer_field_step()
            for _ in range(2):
                s.peer_field_step(fused=True)
            fsteps = 6
        s.sync()
    dist.barrier()                                  # peer transport: nobody tears its memory down while a neighbour still steps
    mine = {"parts": s.download_particles(), "planes": s.download_planes(), "counts": s.tile_counts(),
            "moved": s.migrated_total}
    allr = [None] * world
    dist.all_gather_object(allr, mine)
    if rank == 0:
        from oracle import bindings as ob
        lib = ob.load_oracle()
        P = ob.make_params(lib, nx, ny, tn, tn, dx, dy, dt)
        want, wf = copy_parts(parts), [a.copy() for a in planes]
        counts, _ = ob.pic_steps(lib, P, wf, want, steps)
        if fsteps:
            ob.yee_steps(lib, wf, dt, dx, dy, fsteps)
        got = by_tag({k: np.concatenate([a["parts"][k] for a in allr]) for k in ("x", "y", "ux", "uy", "uz", "q", "tag")})
        want = by_tag(want)
        assert np.array_equal(got["tag"], want["tag"])
        for k in ("x", "y", "ux", "uy", "uz", "q"):
            assert np.array_equal(got[k], want[k]), k
        for k in range(6):
            assert np.array_equal(np.concatenate([a["planes"][k] for a in allr]), wf[k]), k
        assert np.array_equal(np.concatenate([a["counts"] for a in allr]), counts)
        assert sum(a["moved"] for a in allr) > 0
        print(f"SLAB_{transport.upper()}_PARITY_OK world={world} moved={sum(a['moved'] for a in allr)}")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
