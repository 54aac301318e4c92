#!/usr/bin/env python
"""PCIe probe for the e2e legs: link generation / width and pinned copy bandwidth (one way and both ways at once)."""
import subprocess
import torch

print(subprocess.run(["nvidia-smi", "--query-gpu=pcie.link.gen.current,pcie.link.gen.max,pcie.link.width.current", "--format=csv"],
                     capture_output=True, text=True).stdout.strip())
for mb in (0.25, 4, 64):
    nb = int(mb * 2**20)
    h_a, h_b = torch.empty(nb, dtype=torch.uint8).pin_memory(), torch.empty(nb, dtype=torch.uint8).pin_memory()
    d_a, d_b = torch.empty(nb, dtype=torch.uint8, device="cuda"), torch.empty(nb, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    reps = 50

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        s1.synchronize(); s2.synchronize()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) * 1e3 / reps

    t_h2d = timed(lambda: d_a.copy_(h_a, non_blocking=True))
    t_d2h = timed(lambda: h_b.copy_(d_b, non_blocking=True))

    def both():
        with torch.cuda.stream(s1):
            d_a.copy_(h_a, non_blocking=True)
        with torch.cuda.stream(s2):
            h_b.copy_(d_b, non_blocking=True)
    import time
    for _ in range(3):
        both()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        both()
    torch.cuda.synchronize()
    t_both = (time.perf_counter() - t0) * 1e6 / reps
    print("%6.2f MiB: H2D %7.1f us (%5.1f GB/s)  D2H %7.1f us (%5.1f GB/s)  both at once %7.1f us (%5.1f GB/s each way)" %
          (mb, t_h2d, nb / t_h2d / 1e3, t_d2h, nb / t_d2h / 1e3, t_both, nb / t_both / 1e3))
