#!/usr/bin/env python
"""Host <-> device copy bandwidth of the box with 1, 2, 4, 8 GPUs copying AT THE SAME TIME (pinned memory, one stream per
direction per GPU): the ceiling of every end-to-end number that moves its inputs and outputs through the host.
usage: python tools/pcie_bw.py [out.json]   (prints one JSON object)"""
import json
import sys
import time

import torch

MB = 512


def measure(devs, mode):
    n = MB << 20
    bufs = []
    for d in devs:
        torch.cuda.set_device(d)
        bufs.append(dict(dev=d, h_in=torch.empty(n, dtype=torch.uint8).pin_memory(), h_out=torch.empty(n, dtype=torch.uint8).pin_memory(),
                         d_in=torch.empty(n, dtype=torch.uint8, device=f"cuda:{d}"), d_out=torch.empty(n, dtype=torch.uint8, device=f"cuda:{d}"),
                         s1=torch.cuda.Stream(d), s2=torch.cuda.Stream(d)))

    def once():
        for b in bufs:
            if mode in ("h2d", "both"):
                with torch.cuda.stream(b["s1"]):
                    b["d_in"].copy_(b["h_in"], non_blocking=True)
            if mode in ("d2h", "both"):
                with torch.cuda.stream(b["s2"]):
                    b["h_out"].copy_(b["d_out"], non_blocking=True)

    def sync():
        for b in bufs:
            torch.cuda.synchronize(b["dev"])

    for _ in range(2):
        once()
    sync()
    reps = 4
    t = time.perf_counter()
    for _ in range(reps):
        once()
    sync()
    dt = (time.perf_counter() - t) / reps
    per_dir = n * len(devs) / dt / 1e9
    return round(per_dir, 1)


def main():
    n_dev = torch.cuda.device_count()
    out = {"gpus_visible": n_dev, "buffer_mb": MB, "unit": "GB/s aggregate over the GPUs copying concurrently, per direction", "runs": []}
    for g in (1, 2, 4, 8):
        if g > n_dev:
            break
        devs = list(range(g))
        out["runs"].append({"gpus": g, "h2d": measure(devs, "h2d"), "d2h": measure(devs, "d2h"), "both_each_direction": measure(devs, "both")})
    print(json.dumps(out))
    if len(sys.argv) > 1:
        json.dump(out, open(sys.argv[1], "w"), indent=1)


if __name__ == "__main__":
    main()
