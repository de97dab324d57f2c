"""Per-CUDA-source-line sample and instruction counts from an ncu report (needs -lineinfo + --import-source on)."""
import csv, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{pat}", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
fname, res = None, []
hdr = None
for r in rows:
    if r and r[0] == "File Name": fname = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0] not in ("", "Line No"):
        i_s, i_ie = hdr.index("# Samples"), hdr.index("Instructions Executed")
        try: res.append((int(r[i_s] or 0), int(r[i_ie] or 0), fname, r[0], r[1]))
        except ValueError: pass
tot = sum(x[0] for x in res)
print("total samples", tot, " total warp-instr", sum(x[1] for x in res))
for s, ie, f, ln, src in sorted(res, reverse=True)[:topn]:
    print(f"{s:>7} {100*s/tot:5.1f}%  inst {ie:>10}  {f}:{ln:<4} {src.strip()[:100]}")
