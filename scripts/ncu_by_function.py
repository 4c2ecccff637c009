"""Executed issue slots of one kernel attributed to the device functions of the source (development aid).
Joins the per-SASS-instruction execution counts of an .ncu-rep (captured with --import-source on) with the
inline line information of the SAME build's object file (nvdisasm -gi), instruction by instruction.
usage: ncu_by_function.py REP KERNEL_REGEX OBJECT MANGLED_SUBSTRING [depth]"""
import collections, csv, io, os, re, subprocess, sys, tempfile
rep, kern, obj, sym = sys.argv[1:5]
depth = int(sys.argv[5]) if len(sys.argv) > 5 else 1
FP64 = ("DFMA", "DMUL", "DADD", "DSETP", "DMNMX")

def exec_counts():
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    h = rows[1]
    body = [r for r in rows[2:] if len(r) == len(h) and r[0] != h[0]]
    a0 = body[0][0]
    for n in range(1, len(body)):
        if body[n][0] == a0:
            body = body[:n]
            break
    si = h.index("Source")
    ci = [i for i, c in enumerate(h) if "Instructions Executed" in c][0]
    return [(r[si].strip(), int(r[ci] or 0)) for r in body]

def func_table(path):
    """line -> enclosing function name, from the (simply formatted) source file"""
    tab, cur = {}, None
    pat = re.compile(r"^\s*(?:template\s*<[^>]*>\s*)?(?:__device__|__global__|__host__)[^;{]*?\b(\w+)\s*\(")
    lines = open(path).read().split("\n")
    depth_b = 0
    for n, l in enumerate(lines, 1):
        m = pat.match(l)
        if m and depth_b <= 1:
            cur = m.group(1)
        tab[n] = cur
        depth_b += l.count("{") - l.count("}")
    return tab

with tempfile.TemporaryDirectory() as td:
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=td, check=True, capture_output=True)
    cubin = [f for f in os.listdir(td) if f.endswith(".cubin")][0]
    dis = subprocess.run(["nvdisasm", "-gi", os.path.join(td, cubin)], capture_output=True, text=True).stdout.split("\n")
start = [i for i, l in enumerate(dis) if l.startswith(".text.") and sym in l][0]
insts, stack = [], []
fresh = True
for l in dis[start + 1:]:
    if l.startswith(".text.") or l.startswith("\t.section") and insts:
        break
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m:
        if fresh:
            stack = []
            fresh = False
        stack.append((os.path.basename(m.group(1)), int(m.group(2))))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        insts.append((m.group(2).strip(), list(stack)))
        fresh = True
counts = exec_counts()
assert len(counts) == len(insts), (len(counts), len(insts))
tabs = {}
by = collections.defaultdict(lambda: [0, 0])
for (text, st), (src, n) in zip(insts, counts):
    t = text.split()
    op = (t[1] if t[0].startswith("@") and len(t) > 1 else t[0]).split(".")[0]
    names = []
    for f, ln in reversed(st):  # outermost first
        path = os.path.join(os.path.dirname(os.path.abspath(obj)), "..", "..", "csrc", f)
        if f not in tabs:
            tabs[f] = func_table(path) if os.path.exists(path) else {}
        fn = tabs[f].get(ln)
        if fn and (not names or names[-1] != fn):
            names.append(fn)
    key = "/".join(names[1:1 + depth]) if len(names) > 1 else (names[0] if names else "?")
    by[key][0 if op in FP64 else 1] += n
tot = sum(2 * a + b for a, b in by.values())
print(f"{'function':60s} {'fp64':>12s} {'other':>12s}  slots%")
for k, (a, b) in sorted(by.items(), key=lambda x: -(2 * x[1][0] + x[1][1])):
    print(f"{k:60s} {a:12d} {b:12d}  {(2*a+b)/tot*100:5.1f}")
