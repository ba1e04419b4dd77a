#!/usr/bin/env python
"""A/B of the N = 8192 frequency-search layouts on the GPU box (library built with EXTRA=-DXC_EXPERIMENTS):
XC_LAYOUT = b2 | dual | staged | pp.  Each layout runs in its own process (the switch is read once per process) on the
C2 record and on a short record with windows that leave rx; results are compared bit for bit with the B = 2 layout,
times are CUDA-event medians."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CHILD = r'''
import sys, numpy as np, torch
sys.path.insert(0, %r)
from pydsproutines_b200 import signals
import pydsproutines_b200.xcorrRoutines as xr
out = sys.argv[1]
c = signals.config_c2()
cu = torch.from_numpy(c["cutout"]).cuda(); rx = torch.from_numpy(c["rx"]).cuda()
def run():
    return xr.cp_fastXcorr(cu, rx, freqsearch=True, copyToCpu=False)
q, f = run()
torch.cuda.synchronize()
ts = []
for _ in range(7):
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record(); run(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
# short record, odd count, shifts leaving rx on both sides via CyIppXcorrFFT-style range
cutout, rxs = signals.make_fastxcorr_case(16, 8192, 8192 + 3000, 777, 1234, amp=0.7)
q2, f2 = xr.fastXcorr(cutout, rxs, freqsearch=True)
q3, f3 = xr.fastXcorr(cutout, rxs, freqsearch=True, shifts=np.arange(5, 2500))
np.savez(out, q=np.asarray(q.cpu() if hasattr(q, "cpu") else q), f=np.asarray(f.cpu() if hasattr(f, "cpu") else f),
         q2=q2, f2=f2, q3=q3, f3=f3, ms=np.median(ts))
''' % ROOT


def main():
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    res = {}
    names = sys.argv[1:] or ["b2", "staged", "pp"]
    for name, env in [(n, {"XC_LAYOUT": n}) for n in names]:
        out = os.path.join(ROOT, "gpurun_out", "ab_%s.npz" % name)
        e = dict(os.environ, **env)
        r = subprocess.run([sys.executable, "-c", CHILD, out], env=e, cwd=ROOT, capture_output=True, text=True, timeout=600)
        if r.returncode != 0:
            print(name, "FAILED", r.stderr[-2000:])
            continue
        res[name] = np.load(out)
    ref = res[names[0]]
    for name, d in res.items():
        line = {"layout": name, "c2_ms": float(d["ms"])}
        for k in ("q", "f", "q2", "f2", "q3", "f3"):
            line["same_" + k] = bool(np.array_equal(d[k], ref[k]))
            if not line["same_" + k]:
                a, b = d[k].astype(np.float64), ref[k].astype(np.float64)
                line["ndiff_" + k] = int(np.sum(d[k] != ref[k]))
                if k.startswith("q"):
                    line["maxrel_" + k] = float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-30)))
                else:
                    line["maxdiff_" + k] = float(np.max(np.abs(a - b)))
        print(json.dumps(line))


if __name__ == "__main__":
    main()
