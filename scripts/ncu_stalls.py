"""Per-kernel stall breakdown from an .ncu-rep source page: totals per stall reason and the top SASS lines
for long_scoreboard / wait (development aid).  usage: ncu_stalls.py REP KERNEL_REGEX [reason ...]"""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
reasons = sys.argv[3:] or ["stall_long_sb", "stall_wait"]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]
body = [r for r in rows[2:] if len(r) == len(h) and r[0] != h[0]]
# several launches are concatenated: keep the first
addr0 = body[0][0]
for n in range(1, len(body)):
    if body[n][0] == addr0:
        body = body[:n]
        break
cols = [c for c in h if c.startswith("stall_") and "Not Issued" not in c]
tot = {c: sum(int(r[h.index(c)] or 0) for r in body) for c in cols}
alls = sum(tot.values())
print("stall totals:", ", ".join(f"{c[6:]} {v/alls*100:.1f}%" for c, v in sorted(tot.items(), key=lambda x: -x[1]) if v))
si = h.index("Source")
for reason in reasons:
    ci = h.index(reason)
    top = sorted(((int(r[ci] or 0), i, r[si].strip()) for i, r in enumerate(body)), reverse=True)[:14]
    print(f"-- top {reason} ({tot[reason]/alls*100:.1f}% of samples)")
    for v, i, s_ in top:
        print(f"   {v/alls*100:5.2f}%  line {i:5d}  {s_[:80]}")
