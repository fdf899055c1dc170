"""Host-side cost of one small-batch forward (C1 shape): wall time per call without synchronising, and a cProfile
listing.   python tools/host_profile.py"""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vqvae_b200 import VectorQuantizer  # noqa: E402

dev = torch.device("cuda:0")
mod = VectorQuantizer(64, 512, quantizer="ema").to(dev)
z = torch.randn(8, 5, 32, 32, 64, device=dev).to(torch.bfloat16)
with torch.no_grad():
    for _ in range(20):
        mod(z)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(2000):
        mod(z)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"host time per call {1e6 * (t1 - t0) / 2000:.1f} us; incl. drain {1e6 * (t2 - t0) / 2000:.1f} us")
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(2000):
        mod(z)
    pr.disable()
    torch.cuda.synchronize()
    pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
