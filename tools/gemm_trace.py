"""Print CTA 0's GEMM timeline (FVAD_GEMM_TRACE dump) for one M-tile: cycles relative to the tile's first event."""
import sys
import numpy as np
t = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(5, 4, 8, 32).astype(np.int64)
im = int(sys.argv[2]) if len(sys.argv) > 2 else 2
base = t[4, im, 0, 0]
nsl = int((t[1, im, 0] > 0).sum())
print(f"M-tile {im}: slabs per N-tile {nsl}; previous tile's last issue {t[2, im-1][t[2, im-1] > 0].max() - base if im else 0}")
for j in range(8):
    if t[4, im, j, 0] == 0:
        break
    r = lambda k, sl: t[k, im, j, sl] - base
    print(f"N-tile {j}: accEmptyOK {r(4,0):7d} | committed {r(4,1):7d} | epi start {r(3,0):7d} done {r(3,1):7d}")
    print("   slab: prodIssued / mmaFullOK / mmaIssued")
    for sl in range(nsl):
        print(f"   {sl:2d}  {r(0,sl):7d} {r(1,sl):7d} {r(2,sl):7d}")
nxt = t[4, im + 1, 0, 0] - base if im < 3 else 0
print("next M-tile starts at", nxt)
