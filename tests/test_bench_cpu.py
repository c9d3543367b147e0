"""Host-side logic of bench.py that needs no GPU: workload bookkeeping and the reference arm's JSON contract."""
import json
import os
import subprocess
import sys

import numpy as np

from util import ROOT, plasma

sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_expected_particles_matches_the_loader():
    """expected_particles (capacity planning of every rank) == what the loader produces, band populations included."""
    w = dict(nx=32, ny=40, tn=8, ppc=3, u_th=0.1, two_stream=0.0, dx=0.1, beam_ppc=5, beam_band=(0.45, 0.55), beam_drift=0.5,
             beam_u_th=0.01)
    bg = plasma().load_uniform(w["nx"], w["ny"], 8, 8, 0.1, 0.1, w["ppc"], u_th=0.1)
    beam = plasma().load_uniform(w["nx"], w["ny"], 8, 8, 0.1, 0.1, w["beam_ppc"], u_th=0.01, band=w["beam_band"])
    total = bg["x"].shape[0] + beam["x"].shape[0]
    assert bench.expected_particles(w, w["ny"], 0, w["ny"]) == total
    # slabs add up, and the band lands in the slabs that contain it
    parts = [bench.expected_particles(w, w["ny"], lo, lo + 8) for lo in range(0, 40, 8)]
    assert sum(parts) == total
    lo, hi = int(round(0.45 * 40)), int(round(0.55 * 40))
    for k, lo_row in enumerate(range(0, 40, 8)):
        rows_in_band = max(0, min(hi, lo_row + 8) - max(lo, lo_row))
        assert parts[k] == 32 * 8 * 3 + 32 * rows_in_band * 5


def test_workloads_name_the_baseline_configs():
    wl = bench.WORKLOADS
    assert wl["two_stream_512"]["nx"] == wl["two_stream_512"]["ny"] == 512 and wl["two_stream_512"]["ppc"] == 64
    assert wl["yee_8192"]["nx"] == 8192 and wl["yee_8192"]["fields_only"]
    assert wl["uniform_c4"]["nx"] * wl["uniform_c4"]["ny"] * wl["uniform_c4"]["ppc"] == 1073741824
    assert wl["uniform_c4_slab"]["ny"] * 8 == wl["uniform_c4"]["ny"]
    assert wl["beam_c5_slab"]["beam_band"] == (0.45, 0.55)


def test_reference_arm_line_for_the_field_workload():
    """`bench.py --impl reference` runs the reference's own field code (oracle/_ref) and prints the contract's keys."""
    from oracle import bindings as ob
    if not ob.have_ref():
        import pytest
        pytest.skip("oracle/_ref not built")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "yee_8192", "--steps", "1",
                        "--warmup", "0"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and d["metric"] == "Yee cell-updates/s" and d["unit"] == "cells/s"
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] == 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["value"] > 1e5


def test_rank_nonzero_of_the_reference_arm_does_no_work():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"], capture_output=True,
                       text=True, timeout=120, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_reference_arm_particle_line_is_honest():
    """The CPU arm runs the FULL named configuration for the steps it is given, prints those steps, the same `config`
    object the native arm prints, and uses every host core even when a launcher exported OMP_NUM_THREADS=1 (torchrun)."""
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "thermal_128", "--steps", "2",
                        "--warmup", "1"], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stderr
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and d["steps"] == 2 and d["warmup"] == 1
    assert d["config"] == json.loads(json.dumps(bench.workload_config("thermal_128", bench.WORKLOADS["thermal_128"], 1)))
    assert d["config"]["particles"] == 128 * 128 * 16 and "262144 particles" in d["cpu_baseline"]["sample"]
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] == bench.cpu_threads()
    assert d["e2e"] == {"value": d["value"], "unit": "particles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert abs(d["ms_per_step"] * 1e-3 * d["value"] - 262144) < 1.0          # ms_per_step is the measured time of one full step


def test_both_arms_build_the_same_config_object():
    for name, w in bench.WORKLOADS.items():
        for world in (1, 8):
            c = bench.workload_config(name, w, world)
            assert c["workload"] == name and json.loads(json.dumps(c)) == c or isinstance(w.get("beam_band"), tuple)
            if not w.get("strong"):
                assert c["ny_global"] == w["ny"] * world
    c8 = bench.workload_config("uniform_c4_slab", bench.WORKLOADS["uniform_c4_slab"], 8)
    assert c8["particles"] == 1073741824          # at 8 GPUs the slab workload IS configs[3]
