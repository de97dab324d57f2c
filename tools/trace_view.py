"""Condensed view of a tools/trace_x3.py dump: one steady-state tile pair, block issues folded into per-round spans."""
import sys
lines = open(sys.argv[1]).read().splitlines()[1:]
ev = []
for l in lines:
    p = l.split()
    if len(p) >= 4 and p[0].isdigit():
        ev.append((int(p[0]), p[1], p[2], p[3]))
starts = [t for t, n, r, w in ev if n == 'mma_start' and r == 'r0' and w == 'slot0']
print("tile-pair starts:", starts, "periods:", [b - a for a, b in zip(starts, starts[1:])])
k = int(sys.argv[2]) if len(sys.argv) > 2 else 1
t0 = starts[k]
t1 = starts[k + 1] if len(starts) > k + 1 else ev[-1][0]
mma = {}
epi = {}
for t, n, r, w in ev:
    if t < t0 - 3000 or t > t1 + 6000:
        continue
    if n == 'mma_start': mma.setdefault((r, w), [None, None])[0] = t - t0
    if n == 'mma_issued': mma.setdefault((r, w), [None, None])[1] = t - t0
    if n == 'epi_wake': epi.setdefault((r, w), [None, None])[0] = t - t0
    if n == 'epi_done': epi.setdefault((r, w), [None, None])[1] = t - t0
print("issuer (start..issued):")
for (r, w), (a, b) in sorted(mma.items(), key=lambda x: (x[1][0] is None, x[1][0])):
    print(f"  {r} {w}: {a} .. {b}  ({(b - a) if a is not None and b is not None else '?'})")
print("epilogue (wake..done):")
for (r, w), (a, b) in sorted(epi.items(), key=lambda x: (x[1][0] is None, x[1][0])):
    print(f"  {r} {w}: {a} .. {b}  ({(b - a) if a is not None and b is not None else '?'})")
