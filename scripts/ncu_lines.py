"""Per-source-line aggregation of an ncu SASS source page: python scripts/ncu_lines.py rep.ncu-rep obj.o kernel_substr [launch_skip]
Maps the report's SASS rows (in order) onto nvdisasm -g line info of the same function in obj.o."""
import csv, re, subprocess, sys, tempfile, os, glob, collections
rep, obj, kern = sys.argv[1], sys.argv[2], sys.argv[3]
skip = sys.argv[4] if len(sys.argv) > 4 else "0"
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", skip, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[1]
data = rows[2:]
ci = {n: i for i, n in enumerate(hdr)}
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
dis = subprocess.run(["nvdisasm", "-g", "-c", glob.glob(tmp + "/*.cubin")[0]], capture_output=True, text=True).stdout.splitlines()
# locate function
start = None
for i, l in enumerate(dis):
    if l.startswith(".text.") and kern in l and l.rstrip().endswith(":"):
        start = i
        break
assert start is not None, "kernel not found"
lines = []  # (file,line) per instruction
cur = ("?", 0)
inl = ""
for l in dis[start + 1:]:
    if l.startswith("//---------------------"):
        break
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        lines.append(cur)
print("sass rows in report: %d, instructions in disassembly: %d" % (len(data), len(lines)))
n = min(len(data), len(lines))
agg = collections.defaultdict(lambda: collections.Counter())
tot = collections.Counter()
for k in range(n):
    r = data[k]
    key = lines[k]
    for name in ("# Samples", "Instructions Executed", "stall_long_sb", "stall_short_sb", "stall_wait", "stall_not_selected",
                 "stall_math", "stall_mio", "stall_branch_resolving", "stall_dispatch", "stall_selected", "stall_no_inst", "stall_lg"):
        if name in ci:
            v = float(r[ci[name]] or 0)
            agg[key][name] += v
            tot[name] += v
print("totals:", dict(tot))
top = sorted(agg.items(), key=lambda kv: -kv[1]["# Samples"])[:45]
src_cache = {}
def src(f, ln):
    for d in ("hnsw-nsg_b200/csrc",):
        p = os.path.join(d, f)
        if os.path.exists(p):
            if p not in src_cache:
                src_cache[p] = open(p).read().splitlines()
            L = src_cache[p]
            return L[ln - 1].strip()[:90] if 0 < ln <= len(L) else ""
    return ""
print("%-18s %8s %6s %10s %7s %7s %7s %7s  source" % ("file:line", "samples", "%", "inst", "longsb", "shortsb", "wait", "notsel"))
for (f, ln), c in top:
    print("%-18s %8d %6.2f %10d %7d %7d %7d %7d  %s" % ("%s:%d" % (f, ln), c["# Samples"], 100 * c["# Samples"] / tot["# Samples"],
          c["Instructions Executed"], c["stall_long_sb"], c["stall_short_sb"], c["stall_wait"], c["stall_not_selected"], src(f, ln)))
