"""Host <-> device copy bandwidth of this box, the way the end-to-end arm of bench.py uses it: 10 MB chunks from / to
page-locked memory, host-to-device alone, device-to-host alone, and both directions at the same time on two streams.
    python tools/pcie_probe.py [chunk MB] [chunks]"""
import json
import sys

import torch

mb = int(sys.argv[1]) if len(sys.argv) > 1 else 10
n = int(sys.argv[2]) if len(sys.argv) > 2 else 200
size = mb << 20
h_in = [torch.empty(size, dtype=torch.uint8).pin_memory() for _ in range(8)]
h_out = [torch.empty(size, dtype=torch.uint8).pin_memory() for _ in range(8)]
d_in = [torch.empty(size, dtype=torch.uint8, device="cuda") for _ in range(8)]
d_out = [torch.empty(size, dtype=torch.uint8, device="cuda") for _ in range(8)]
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(up: bool, down: bool) -> float:
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s1.wait_event(e0)
    s2.wait_event(e0)
    for j in range(n):
        if up:
            with torch.cuda.stream(s1):
                d_in[j % 8].copy_(h_in[j % 8], non_blocking=True)
        if down:
            with torch.cuda.stream(s2):
                h_out[j % 8].copy_(d_out[j % 8], non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1)
    torch.cuda.current_stream().wait_stream(s2)
    e1.record()
    torch.cuda.synchronize()
    return n * size / (e0.elapsed_time(e1) * 1e-3) / 1e9


for _ in range(2):
    run(True, True)
res = {"chunk_mb": mb, "chunks": n, "h2d_alone_gbs": round(run(True, False), 2), "d2h_alone_gbs": round(run(False, True), 2),
       "each_direction_when_both_gbs": round(run(True, True), 2), "gpu": torch.cuda.get_device_name(0)}
print(json.dumps(res))
