"""Per-source-line share of executed instructions and stall samples in one kernel of an ncu report.
    python scripts/ncu_lines.py X.ncu-rep [min_pct]"""
import csv, subprocess, sys
rep = sys.argv[1]; min_pct = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
cur_file = None; lines = {}
hdr = None
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; ie = r.index("Instructions Executed"); isamp = r.index("# Samples"); continue
    if hdr is None or r[0] == "" or r[2] != "-": continue   # only the per-line summary rows
    try: n = float(r[ie]); s = float(r[isamp])
    except ValueError: continue
    k = (cur_file, int(r[0])); a = lines.setdefault(k, [0.0, 0.0, r[1]]); a[0] += n; a[1] += s
tot = sum(v[0] for v in lines.values()); tots = sum(v[1] for v in lines.values())
print(f"total instructions {tot:.0f}, samples {tots:.0f}")
for k, v in sorted(lines.items()):
    if v[0] >= tot * min_pct / 100 or v[1] >= tots * min_pct / 100:
        print(f"{k[0]:>16s}:{k[1]:<4d} inst {v[0]/tot*100:5.1f}%  samples {v[1]/tots*100:5.1f}%  {v[2][:100]}")
