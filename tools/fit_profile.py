"""One c2-shaped run of the ancestor-fitting kernel for ncu (tools/fit_probe.py without the probes)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import fit_probe

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
print("c2 us/iter", fit_probe.one(64, 1000, 4, iters, 0))
