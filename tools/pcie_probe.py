#!/usr/bin/env python
"""What the host link can do on this box: H2D alone, D2H alone, both at once
(pinned memory, 1 GiB per direction, CUDA events).  The e2e number of bench.py
(host buffers in, host buffers out) cannot exceed the 'both' figure."""
import json

import torch


def main():
    n = 1 << 30
    h_a = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    h_b = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    d_a = torch.empty(n, dtype=torch.uint8, device="cuda")
    d_b = torch.empty(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(h2d, d2h, iters=5):
        best = 1e9
        for _ in range(iters):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            s1.wait_event(e0)
            s2.wait_event(e0)
            if h2d:
                with torch.cuda.stream(s1):
                    d_a.copy_(h_a, non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    h_b.copy_(d_b, non_blocking=True)
            torch.cuda.current_stream().wait_stream(s1)
            torch.cuda.current_stream().wait_stream(s2)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return n / best / 1e6

    out = {"h2d_gbs": round(run(True, False), 2), "d2h_gbs": round(run(False, True), 2),
           "both_gbs_each_direction": round(run(True, True), 2)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
