"""Summarise an ncu source page: stall totals and the hottest SASS lines with their source line."""
import csv, collections, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{pat}", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
i_s, i_src = hdr.index("# Samples"), hdr.index("Source")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = collections.Counter()
body = []
for r in rows[hi + 1:]:
    if r and r[0] == "Address": break          # next kernel instance
    if len(r) == len(hdr): body.append(r)
for r in body:
    for i in stall:
        try: tot[hdr[i]] += int(r[i] or 0)
        except ValueError: pass
print("total samples", sum(int(r[i_s] or 0) for r in body))
print(tot.most_common(10))
for k, r in enumerate(body): r.append(k)
for r in sorted(body, key=lambda r: -int(r[i_s] or 0))[:topn]:
    st = sorted(((int(r[i] or 0), hdr[i]) for i in stall), reverse=True)[:2]
    print(f"{r[i_s]:>7} #{r[-1]:<5} {r[i_src][:90]:<90} {st}")
