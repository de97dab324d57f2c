"""Histogram of pc samples and executed instructions over SASS index bins (first kernel instance matching pat)."""
import csv, subprocess, sys, collections
rep, pat = sys.argv[1], sys.argv[2]
binw = int(sys.argv[3]) if len(sys.argv) > 3 else 100
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{pat}", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
i_s, i_src, i_ie = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
body = []
for r in rows[hi + 1:]:
    if r and r[0] == "Address": break
    if len(r) == len(hdr): body.append(r)
tot_s = sum(int(r[i_s] or 0) for r in body); tot_i = sum(int(r[i_ie] or 0) for r in body)
print("instructions", len(body), "samples", tot_s, "warp-instr", tot_i)
for b in range(0, len(body), binw):
    seg = body[b:b + binw]
    s = sum(int(r[i_s] or 0) for r in seg); ie = sum(int(r[i_ie] or 0) for r in seg)
    ops = collections.Counter(r[i_src].split()[1 if r[i_src].startswith("@") else 0].split(".")[0] for r in seg if r[i_src].split())
    print(f"#{b:<5} samples {s:>7} {100*s/tot_s:5.1f}%  instr {ie:>11} {100*ie/tot_i:5.1f}%  {dict(ops.most_common(6))}")
