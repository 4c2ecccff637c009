"""Dynamic (executed) SASS opcode histogram of one kernel from an .ncu-rep captured with --import-source on
(development aid).  usage: ncu_ophist.py REP KERNEL_REGEX [top]"""
import csv, io, subprocess, sys, collections
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]
body = [r for r in rows[2:] if len(r) == len(h) and r[0] != h[0]]
addr0 = body[0][0]
for n in range(1, len(body)):
    if body[n][0] == addr0:
        body = body[:n]
        break
si = h.index("Source")
ci = h.index("# Instructions Executed") if "# Instructions Executed" in h else [i for i, c in enumerate(h) if "Instructions Executed" in c][0]
hist = collections.Counter()
for r in body:
    t = r[si].strip().split()
    if not t:
        continue
    op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
    hist[op.split(".")[0]] += int(r[ci] or 0)
tot = sum(hist.values())
fp64 = sum(v for k, v in hist.items() if k in ("DFMA", "DMUL", "DADD", "DSETP", "DMNMX"))
print(f"executed warp instructions {tot}; fp64-pipe {fp64} ({fp64/tot*100:.1f}%); model issue clocks 2*fp64+other = {tot+fp64}")
for k, v in hist.most_common(top):
    print(f"  {k:10s} {v:12d} {v/tot*100:5.1f}%")
