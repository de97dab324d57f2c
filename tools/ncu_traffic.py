"""profiles/ncu_traffic.json: per-launch DRAM traffic (dram__bytes_read.sum + dram__bytes_write.sum) of our kernels
from `ncu --set full` captures.   usage: ncu_traffic.py <workload> <raw.csv> [<raw.csv> ...]
(raw.csv = `ncu -i X.ncu-rep --page raw --csv`).  bench.py reports these numbers as roofline.traffic."""
import csv, json, os, re, sys
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles", "ncu_traffic.json")
out = json.load(open(root)) if os.path.exists(root) else {}
wl = sys.argv[1]
acc = {}
for path in sys.argv[2:]:
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    ki, ri, wi = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    for r in rows[2:]:
        name = re.search(r"(\w+_kernel)", r[ki])
        if not name:
            continue
        b = float(r[ri]) * UNIT[units[ri]] + float(r[wi]) * UNIT[units[wi]]
        acc.setdefault(name.group(1), []).append(b)
out.setdefault(wl, {}).update({k: sum(v) / len(v) for k, v in acc.items()})
out["_how"] = "ncu --set full --clock-control none, dram__bytes_read.sum + dram__bytes_write.sum per launch (mean over captured launches)"
json.dump(out, open(root, "w"), indent=1, sort_keys=True)
print(json.dumps(out.get(wl), indent=1))
