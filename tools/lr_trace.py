"""Dev tool: print the hand-off timeline lr_mn_tc_kernel recorded (library built with -DLT_TRACE, B2F_LR_TRACE=<file>).

  tools/build_variant.sh lrtrace lr_tc.cu -DLT_TRACE
  B2F_LIB_PATH=build_variants/lrtrace/libb200fit.so B2F_LR_TRACE=gpurun_out/lrtrace.bin python tools/lr_bench.py --reps 1
  python tools/lr_trace.py gpurun_out/lrtrace.bin [first_tile [ntiles [FB]]]
"""
import sys

import numpy as np

TILES, SLOTS, ROLES = 48, 16, 6
t = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(ROLES, TILES, SLOTS).astype(np.int64)
first = int(sys.argv[2]) if len(sys.argv) > 2 else 20
ntiles = int(sys.argv[3]) if len(sys.argv) > 3 else 2
FB = int(sys.argv[4]) if len(sys.argv) > 4 else 4
t0 = t[0, first, 0]
ev = []
for it in range(first, first + ntiles):
    for s, lab in enumerate(("G1 top (wait d1_empty)", "G1 got margins buffer", "G1 all chunks issued", "G2 top (wait py_full)",
                             "G2 got (P-Y)")):
        ev.append((t[0, it, s], "mma     tile %d %s" % (it, lab)))
    for fb in range(FB):
        ev.append((t[0, it, 5 + fb], "mma     tile %d G2 block %d operand ready" % (it, fb)))
    ev.append((t[0, it, 9], "mma     tile %d G2 all blocks issued" % it))
    for s, lab in enumerate(("wait d1_full", "got margins", "margins read, buffer released", "softmax done (wait py_empty)",
                             "got py buffer", "(P-Y) stored")):
        ev.append((t[1, it, s], "softmax tile %d %s" % (it, lab)))
    for grp in range(2):
        r = 2 + grp
        ev.append((t[r, it, 0], "split%d  tile %d G1 chunks start" % (grp, it)))
        ev.append((t[r, it, 1], "split%d  tile %d G1 chunks done" % (grp, it)))
        ev.append((t[r, it, 2], "split%d  tile %d G2 blocks start" % (grp, it)))
        for fb in range(FB):
            ev.append((t[r, it, 3 + fb], "split%d  tile %d G2 block %d got a2 tile" % (grp, it, fb)))
            ev.append((t[r, it, 7 + fb], "split%d  tile %d G2 block %d stored" % (grp, it, fb)))
    for s, lab in enumerate(("G1 loads start", "G1 loads issued", "G2 loads start", "G2 loads issued")):
        ev.append((t[4, it, s], "xprod   tile %d %s" % (it, lab)))
ev = [(c - t0, s) for c, s in ev if c]
ev.sort()
for c, s in ev:
    print("%8d  %s" % (c, s))

a, b = 8, TILES - 2
print()
print("mean cycles per row tile (mma G2 issued, tiles %d..%d): %.0f" % (a, b, np.diff(t[0, a:b, 9]).mean()))
m = t[0, a:b]
print("mma: wait d1_empty %.0f | G1 issue loop %.0f | wait py_full %.0f | G2 loop %.0f (first operand wait %.0f)" % (
    (m[:, 1] - m[:, 0]).mean(), (m[:, 2] - m[:, 1]).mean(), (m[:, 4] - m[:, 3]).mean(), (m[:, 9] - m[:, 4]).mean(),
    (m[:, 5] - m[:, 4]).mean()))
s = t[1, a:b]
print("softmax: wait d1_full %.0f | tmem read %.0f | math %.0f | wait py_empty %.0f | store %.0f" % (
    (s[:, 1] - s[:, 0]).mean(), (s[:, 2] - s[:, 1]).mean(), (s[:, 3] - s[:, 2]).mean(), (s[:, 4] - s[:, 3]).mean(),
    (s[:, 5] - s[:, 4]).mean()))
for grp in range(2):
    g = t[2 + grp, a:b]
    waits = sum((g[:, 3 + fb] - (g[:, 2] if fb == 0 else g[:, 6 + fb])).mean() for fb in range(FB))
    work = sum((g[:, 7 + fb] - g[:, 3 + fb]).mean() for fb in range(FB))
    print("split%d: G1 chunks %.0f | G2: waiting for the a2 tile %.0f, converting %.0f | gap G2 end -> next G1 start %.0f" % (
        grp, (g[:, 1] - g[:, 0]).mean(), waits, work, (t[2 + grp, a + 1:b + 1, 0] - g[:, 1]).mean()))
p = t[4, a:b]
print("xprod: G1 loads %.0f | G2 loads %.0f" % ((p[:, 1] - p[:, 0]).mean(), (p[:, 3] - p[:, 2]).mean()))
c = t[5, a:b]
print("inside the last G1 chunk of a tile (warp 8): wait xs_full %.0f | load + convert %.0f | wait a1_empty %.0f | tcgen05.st + wait %.0f | "
      "fence + arrive %.0f" % tuple((c[:, i + 1] - c[:, i]).mean() for i in range(5)))
print("inside the last G2 chunk of a tile (warp 8): wait xs_full %.0f | load + convert + store %.0f" % (
    (c[:, 7] - c[:, 6]).mean(), (c[:, 8] - c[:, 7]).mean()))
print("warp 8, cycles waited per tile phase: xs_full in G1 chunks %.0f | a1_empty %.0f | a2_empty %.0f | xs_full in G2 chunks %.0f" %
      tuple(c[:, 9 + k].mean() for k in range(4)))
