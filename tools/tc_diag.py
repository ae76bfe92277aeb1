"""Run each tcgen05 conv geometry of tests/test_gpu_ops.py in its own process (a trapping kernel poisons the
CUDA context) and print the rel-L2 errors.  Usage: python tools/tc_diag.py [index]"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

if len(sys.argv) > 1:
    import torch
    from test_gpu_ops import GEOMS, conv_case
    g = GEOMS[int(sys.argv[1])]
    res = conv_case(*g, dtype=torch.bfloat16, use_tc=1)
    print("TC", g, {k: "%.3e" % v for k, v in res.items()}, flush=True)
else:
    from test_gpu_ops import GEOMS
    for i in range(len(GEOMS)):
        try:
            r = subprocess.run([sys.executable, __file__, str(i)], capture_output=True, text=True, timeout=120)
            out = (r.stdout.strip().splitlines() or ["<no output>"])[-1]
            if r.returncode != 0:
                out += " | rc=%d %s" % (r.returncode, r.stderr.strip().splitlines()[-1:] )
        except subprocess.TimeoutExpired:
            out = "TIMEOUT"
        print(i, GEOMS[i], out, flush=True)
