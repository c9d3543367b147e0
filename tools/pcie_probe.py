"""PCIe / host-memory floor of the host-buffer step: pinned H2D alone, D2H alone, and both directions at once (two streams),
885 MB each way -- the bytes bench.py's e2e leg moves per rank and step.

  python tools/pcie_probe.py                                        one GPU
  python -m torch.distributed.run --nproc-per-node N ... tools/pcie_probe.py    all N ranks AT THE SAME TIME (one process per GPU,
      barrier in front of every timed loop): what the N>1 e2e leg can reach at best -- the ranks share the host's memory system
      and PCIe root complexes, so the per-rank rate falls as N grows ("measure the real host floor").
Prints one JSON line (rank 0): per-rank times (max over ranks), aggregate GB/s, and the implied floor of an e2e step in ms."""
import json
import os
import torch

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist = None
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("gloo")

n = 884998144
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    if dist:
        dist.barrier()                      # every rank starts its timed loop together
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    for s in (s1, s2):
        torch.cuda.current_stream().wait_stream(s)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    if dist:
        allms = [None] * world
        dist.all_gather_object(allms, ms)
        ms = max(allms)
    return ms


def up():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)


def down():
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


def both():
    s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
    up(); down()
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)


def chunked(k=16):
    c = n // k
    for i in range(k):
        with torch.cuda.stream(s1):
            d_in[i * c:(i + 1) * c].copy_(h_in[i * c:(i + 1) * c], non_blocking=True)
        with torch.cuda.stream(s2):
            h_out[i * c:(i + 1) * c].copy_(d_out[i * c:(i + 1) * c], non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)


r = {"ranks_at_once": world, "bytes_each_way_per_rank": n, "h2d_ms": timed(up), "d2h_ms": timed(down), "both_ms": timed(both),
     "both_16_chunks_ms": timed(chunked)}
r["h2d_gbs_per_rank"] = n / r["h2d_ms"] / 1e6
r["d2h_gbs_per_rank"] = n / r["d2h_ms"] / 1e6
r["duplex_gbs_each_way_per_rank"] = n / r["both_ms"] / 1e6
r["aggregate_host_traffic_gbs_duplex"] = 2 * world * n / r["both_ms"] / 1e6
r["floor_ms_sequential_h2d_then_d2h"] = r["h2d_ms"] + r["d2h_ms"]
r["floor_ms_fully_overlapped"] = r["both_ms"]
if rank == 0:
    print(json.dumps(r))
if dist:
    dist.destroy_process_group()
