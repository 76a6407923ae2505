"""Print the hottest SASS lines (warp-stall samples) of an `ncu --page source --csv` export."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) > col["stall_long_sb"] and r[col["# Samples"]].isdigit()]
tot = sum(int(r[col["# Samples"]]) for r in data)
print("total samples", tot, "SASS lines", len(data))
keys = ["stall_long_sb", "stall_short_sb", "stall_barrier", "stall_mio", "stall_wait", "stall_math", "stall_lg"]
agg = {k: sum(int(r[col[k]] or 0) for r in data) for k in keys}
print(agg)
cum = 0
for r in sorted(data, key=lambda r: -int(r[col["# Samples"]]))[:n]:
    cum += int(r[col["# Samples"]])
    print(f'{r[col["# Samples"]]:>5} cum {100*cum/tot:4.1f}% ' + " ".join(f'{k[6:9]}={r[col[k]]}' for k in keys) + "  " + r[col["Source"]][:90])
