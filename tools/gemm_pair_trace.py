"""Print the leader CTA's timeline of the CTA-pair GEMM (FVAD_GEMM_TRACE dump with FVAD_GEMM_PAIR=1) for one M-pair:
cycles relative to the pair's first accumulator hand-over."""
import sys
import numpy as np
t = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(5, 4, 8, 32).astype(np.int64)
im = int(sys.argv[2]) if len(sys.argv) > 2 else 2
base = t[4, im, 0, 0]
nsl = int((t[1, im, 0, :8] > 0).sum())
r = lambda k, j, sl: (t[k, im, j, sl] - base) if t[k, im, j, sl] else -1
print(f"M-pair {im}: {nsl} slabs per N-tile; previous pair's last commit {t[4, im-1, :, 1].max() - base if im else 0}")
print("A slabs:  stager a_empty ok / as_full ok / stored / staged (after the stagers' barrier) | MMA a_full ok")
for sl in range(nsl):
    print(f"   {sl}  {r(4,0,2+sl):7d} {r(4,1,2+sl):7d} {r(4,3,2+sl):7d} {r(4,2,2+sl):7d} | {r(3,0,2+sl):7d}")
for j in range(8):
    if t[4, im, j, 0] == 0:
        break
    print(f"N-tile {j}: accEmptyOK {r(4,j,0):7d} | committed {r(4,j,1):7d} | epi start {r(3,j,0):7d} done {r(3,j,1):7d}")
    print("   slab: prodIssued / mmaFullOK / mmaIssued")
    for sl in range(nsl):
        print(f"   {sl:2d}  {r(0,j,sl):7d} {r(1,j,sl):7d} {r(2,j,sl):7d}")
if im < 3:
    print("next M-pair starts at", t[4, im + 1, 0, 0] - base)
