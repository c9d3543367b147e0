"""PCIe floor of the host-buffer step: pinned H2D alone, D2H alone, and both directions at once (two streams), 885 MB each
-- the bytes bench.py's e2e leg moves per step.  Prints GB/s and the implied floor in ms."""
import json
import torch
n = 884998144
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    for s in (s1, s2):
        torch.cuda.current_stream().wait_stream(s)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


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


r = {"bytes_each_way": n, "h2d_ms": timed(up), "d2h_ms": timed(down), "both_ms": timed(both), "both_16_chunks_ms": timed(chunked)}
r["h2d_gbs"] = n / r["h2d_ms"] / 1e6; r["d2h_gbs"] = n / r["d2h_ms"] / 1e6; r["duplex_gbs_each"] = n / r["both_ms"] / 1e6
print(json.dumps(r))
