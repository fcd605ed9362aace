"""Print CTA 0's GRU timeline (FVAD_GRU_TRACE dump): per chunk of one step, cycles relative to the step start."""
import sys
import numpy as np
t = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(54, 25, 8).astype(np.int64)
step = int(sys.argv[2]) if len(sys.argv) > 2 else 20
base = t[step, 0, 6]
print(f"step {step}: h_ready at 0; previous step's h_ready {t[step-1,0,6]-base}; next {t[step+1,0,6]-base}")
print("chunk  accEmptyOK  fullOK  issued | epi: accFullOK  ldDone  mathDone")
for c in range(25):
    r = t[step, c] - base
    print(f"{c:3d}   {r[0]:8d} {r[1]:8d} {r[2]:8d} |  {r[3]:8d} {r[4]:8d} {r[5]:8d}")
d = np.diff(t[5:50, 0, 6])
print("cycles per step: mean", d.mean(), "min", d.min(), "max", d.max())
