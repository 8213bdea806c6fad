"""Differential fuzz on the GPU: CUDA path through the C ABI vs the CPU checker on random stage configurations (tests/_fuzz.py).
Usage: python tools/fuzz_gpu.py [seed] [configs]"""
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import _fuzz as F

seed, n = (int(sys.argv[1]) if len(sys.argv) > 1 else 0), (int(sys.argv[2]) if len(sys.argv) > 2 else 40)
t0 = time.time()
bad = F.run_gpu(seed, n)
for b in bad:
    print("MISMATCH", b)
print("fuzz gpu seed %d configs %d mismatching %d (%.0f s)" % (seed, n, len(bad), time.time() - t0))
