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
