import importlib
import os
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "mini-pic--2d-pic-code-with-tiling-load-balance_b200"


def pkg():
    return importlib.import_module(PKG)


def plasma():
    return importlib.import_module(PKG + ".plasma")


def have_gpu():
    try:
        return pkg().lib().mpic_device_count() > 0
    except Exception:
        return False


def by_tag(parts):
    o = np.argsort(parts["tag"], kind="stable")
    return {k: v[o] for k, v in parts.items()}


def copy_parts(parts):
    return {k: v.copy() for k, v in parts.items()}


def rel_l2(a, b):
    d = np.linalg.norm(np.asarray(a, dtype=np.float64) - np.asarray(b, dtype=np.float64))
    n = np.linalg.norm(np.asarray(b, dtype=np.float64))
    return d / n if n > 0 else d


class ThreadHub:
    """Mailboxes for ThreadRing: emulates N slab ranks as N threads of ONE process (all handles on one GPU)."""

    def __init__(self, world):
        import queue
        self.world = world
        self.q = {(r, side): queue.Queue() for r in range(world) for side in ("lo", "hi")}


class ThreadRing:
    """Drop-in for slab.Ring without torch.distributed: the same sendrecv contract through in-process queues.  Fully
    synchronising (device-wide) on both sides of every message: a correctness harness, not a fast path."""

    def __init__(self, rank, hub):
        self.rank, self.world, self.hub = rank, hub.world, hub
        self.lo, self.hi = (rank - 1) % hub.world, (rank + 1) % hub.world

    def sendrecv(self, send_lo, send_hi, recv_lo, recv_hi):
        import torch
        torch.cuda.synchronize()
        if send_lo is not None:
            self.hub.q[(self.lo, "hi")].put(send_lo.clone())
        if send_hi is not None:
            self.hub.q[(self.hi, "lo")].put(send_hi.clone())
        torch.cuda.synchronize()
        if recv_hi is not None:
            recv_hi.copy_(self.hub.q[(self.rank, "hi")].get(timeout=30))
        if recv_lo is not None:
            recv_lo.copy_(self.hub.q[(self.rank, "lo")].get(timeout=30))
        torch.cuda.synchronize()
