"""Dev tool: print the hand-off timeline that km_tcb_kernel recorded (B2F_TC_TRACE=<file>), in SM cycles.

  B2F_TC_TRACE=gpurun_out/trace.bin python tools/km_bench.py --algo 5 --reps 1 ; python tools/tc_trace.py gpurun_out/trace.bin
"""
import sys

import numpy as np

TILES, SLOTS, ROLES = 64, 32, 12
t = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(ROLES, TILES, SLOTS).astype(np.int64)
first = int(sys.argv[2]) if len(sys.argv) > 2 else 20
ntiles = int(sys.argv[3]) if len(sys.argv) > 3 else 3
t0 = t[0, first, 0]
names = {0: "mma", 9: "split", 10: "fin"}
for w in range(8):
    names[1 + w] = "e%dq%d" % (w >> 2, w & 3)
ev = []
for it in range(first, first + ntiles):
    for nt in range(8):
        for s, lab in enumerate(("loop top", "operands ready", "issued")):
            ev.append((t[0, it, 3 * nt + s], "mma   tile %d nt %d %s" % (it, nt, lab)))
        ev.append((t[0, it, 24 + nt], "mma   tile %d nt %d accumulator free, issuing" % (it, nt)))
        ev.append((t[11, it, nt], "bprod tile %d nt %d stage free, TMA issued" % (it, nt)))
        for r in range(1, 9):
            if t[r, it, 3 * nt]:
                for s, lab in enumerate(("wait t_full", "got accumulator", "released accumulator")):
                    ev.append((t[r, it, 3 * nt + s], "%-5s tile %d nt %d %s" % (names[r], it, nt, lab)))
    for r in range(1, 9):
        ev.append((t[r, it, 31], "%-5s tile %d partial result handed to finish" % (names[r], it)))
    for s, lab in enumerate(("start", "atom0 converted, wait xa_empty", "atom0 got buffer", "atom0 stored",
                             "atom1 converted, wait xa_empty", "atom1 got buffer", "atom1 stored")):
        ev.append((t[9, it, s], "split tile %d %s" % (it, lab)))
    for s, lab in enumerate(("wait a_full", "got partial results", "labels + list done", "accumulated")):
        ev.append((t[10, it, s], "fin   tile %d %s" % (it, lab)))
    for g in range(4):
        for s, lab in enumerate(("batch start", "loads issued", "reductions issued")):
            ev.append((t[10, it, 4 + 3 * g + s], "fin   tile %d batch %d %s" % (it, g, lab)))
ev = [(c - t0, s) for c, s in ev if c]
ev.sort()
for c, s in ev:
    print("%8d  %s" % (c, s))
per_tile = np.diff(t[1, 8:TILES, 31]).mean()
print("mean cycles per row tile (epilogue warp 0 hand-offs, tiles 8..63): %.0f" % per_tile)
for r in range(1, 9):
    busy = 0
    n = 0
    for it in range(8, TILES):
        for nt in range(8):
            if t[r, it, 3 * nt]:
                busy += t[r, it, 3 * nt + 2] - t[r, it, 3 * nt + 1]
                n += 1
    print("%s: %.0f cycles per accumulator tile while holding it (%d tiles)" % (names[r], busy / max(n, 1), n))
    wait = sum(t[r, it, 3 * nt + 1] - t[r, it, 3 * nt] for it in range(8, TILES) for nt in range(8) if t[r, it, 3 * nt])
    print("%s: %.0f cycles waiting for t_full per accumulator tile" % (names[r], wait / max(n, 1)))
own = [(it, nt) for it in range(8, TILES) for nt in range(8) if t[0, it, 3 * nt]]
mw = np.mean([t[0, it, 3 * nt + 1] - t[0, it, 3 * nt] for it, nt in own])
mi = np.mean([t[0, it, 3 * nt + 2] - t[0, it, 3 * nt + 1] for it, nt in own])
print("mma: %.0f cycles until operands ready, %.0f cycles waiting for the accumulator + issuing, per centre tile" % (mw, mi))
