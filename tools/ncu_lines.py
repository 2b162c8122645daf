"""Per-source-line instruction counts and stall samples of one kernel.
usage: ncu_lines.py <report.ncu-rep> <kernel-mangled-substring> [lib.so] [top]
Joins `ncu --page source --csv` (SASS level) with `nvdisasm -g` line info of the cubin."""
import csv, os, re, subprocess, sys, tempfile, collections
rep, kern = sys.argv[1], sys.argv[2]
lib = sys.argv[3] if len(sys.argv) > 3 else "x-edr-pointcloud-segmentation_b200/libpcseg_b200.so"
top = int(sys.argv[4]) if len(sys.argv) > 4 else 45
lib = os.path.abspath(lib); rep = os.path.abspath(rep)
d = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", lib], cwd=d, stdout=subprocess.DEVNULL)
cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cub], cwd=d, capture_output=True, text=True).stdout.splitlines()
line_of = {}
infn = False; cur = None
for l in dis:
    if l.startswith(".text."):
        infn = kern in l; cur = None; continue
    if not infn: continue
    m = re.match(r'\s*//## File "(.*)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r'\s*/\*([0-9a-f]{4,6})\*/', l)
    if m: line_of[int(m.group(1), 16)] = cur
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hi = [i for i, r in enumerate(rows) if "Address" in r][0]
hdr = rows[hi]
ia, ie, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
base = None
tot_e = tot_s = 0
for r in rows[hi + 1:]:
    if len(r) <= ie: continue
    try: addr = int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])
    except ValueError: continue
    if base is None: base = addr
    key = line_of.get(addr - base)
    e = int(float(r[ie] or 0)); s = int(float(r[isamp] or 0))
    a = agg[key]; a[0] += e; a[1] += s
    for i, h in stall_cols:
        v = int(float(r[i] or 0))
        if v: a[2][h] += v
    tot_e += e; tot_s += s
print("total warp-instructions %d, samples %d" % (tot_e, tot_s))
byfile = collections.Counter()
for k, a in agg.items(): byfile[k[0] if k else None] += a[0]
print("by file:", dict(byfile))
print("%-28s %12s %6s %8s %6s  top stalls" % ("line", "instr", "%", "samples", "%"))
sampf = collections.Counter()
for k, a in agg.items(): sampf[k[0] if k else None] += a[1]
print("samples by file:", dict(sampf))
order = (lambda kv: -kv[1][0]) if os.environ.get("BY") == "instr" else (lambda kv: -kv[1][1])
for k, a in sorted(agg.items(), key=order)[:top]:
    st = ", ".join("%s %d" % (h[6:], v) for h, v in a[2].most_common(3))
    print("%-28s %12d %6.2f %8d %6.2f  %s" % ("%s:%d" % k if k else "?", a[0], 100.0 * a[0] / max(tot_e, 1), a[1], 100.0 * a[1] / max(tot_s, 1), st))
