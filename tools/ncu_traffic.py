"""Update profiles/ncu_traffic.json from an .ncu-rep: python tools/ncu_traffic.py <report.ncu-rep> <kernel-regex> <key> [source-note]
Averages dram__bytes_read.sum + dram__bytes_write.sum over the launches whose kernel name matches the regex."""
import csv, io, json, os, re, subprocess, sys
rep, pat, key = sys.argv[1:4]
note = sys.argv[4] if len(sys.argv) > 4 else os.path.basename(rep)
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h, units = rows[0], rows[1]
ik, ir, iw = h.index("Kernel Name"), h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
vals = [float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]] for r in rows[2:] if re.search(pat, r[ik])]
assert vals, f"no launch matches {pat}"
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_traffic.json")
rec = json.load(open(path))
rec[key] = {"dram_bytes_per_launch": sum(vals) / len(vals), "launches": len(vals), "source": note}
json.dump(rec, open(path, "w"), indent=1)
print(key, rec[key])
