"""Summarise an `ncu --page source --csv` dump: top SASS instructions by executed count and by stall samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
# find header row
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]; idx = {n: i for i, n in enumerate(hdr)}
data = [r for r in rows[h + 1:] if len(r) == len(hdr)]
def f(r, n):
    try: return float(r[idx[n]])
    except Exception: return 0.0
tot_i = sum(f(r, "Instructions Executed") for r in data); tot_s = sum(f(r, "# Samples") for r in data)
print("total warp-instr %.0f  samples %.0f  sass lines %d" % (tot_i, tot_s, len(data)))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 25
print("== by instructions executed")
for r in sorted(data, key=lambda r: -f(r, "Instructions Executed"))[:N]:
    print("%6.2f%% %6.2f%%smp  %s" % (100 * f(r, "Instructions Executed") / tot_i, 100 * f(r, "# Samples") / max(tot_s, 1), r[idx["Source"]][:110]))
print("== by stall samples")
stalls = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
for r in sorted(data, key=lambda r: -f(r, "# Samples"))[:N]:
    top = sorted(((f(r, s), s) for s in stalls), reverse=True)[:2]
    print("%6.2f%%smp %6.2f%%  %-90s %s" % (100 * f(r, "# Samples") / max(tot_s, 1), 100 * f(r, "Instructions Executed") / tot_i, r[idx["Source"]][:90], " ".join("%s=%.0f" % (s[6:], v) for v, s in top)))
