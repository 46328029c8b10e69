"""Per-source-line instruction counts and stall samples of one kernel of an ncu capture.
Joins `ncu --page source --print-source sass` (per SASS instruction, in address order) with the line table of
`nvdisasm -g` on the matching cubin function (same instruction order).

    python tools/ncu_lines.py gpurun_out/cost_r01.ncu-rep sgbm_cost.o 'k_cost_fast<5, 16, false>' [kernel_index]
"""
import csv
import io
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, obj, pat = sys.argv[1], sys.argv[2], sys.argv[3]
kidx = int(sys.argv[4]) if len(sys.argv) > 4 else 0
objp = os.path.join(ROOT, "stereo_match_b200", "_obj", obj)

raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
kernels, cur = [], None
for row in csv.reader(io.StringIO(raw)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "rows": [], "hdr": None}
        kernels.append(cur)
    elif cur is not None and row and row[0] == "Address":
        cur["hdr"] = row
    elif cur is not None and row and row[0].startswith("0x"):
        cur["rows"].append(row)
# the page lists every kernel twice (same addresses): keep one of each
ded = []
for kk in kernels:
    if ded and ded[-1]["name"] == kk["name"] and [r[0] for r in ded[-1]["rows"]] == [r[0] for r in kk["rows"]]:
        continue
    ded.append(kk)
k = ded[kidx]
H = k["hdr"]
iex, ist = H.index("Instructions Executed"), H.index("Warp Stall Sampling (All Samples)")
stall_cols = [i for i, c in enumerate(H) if c.startswith("stall_") and "Not Issued" not in c]

tmp = "/tmp/ncu_lines_cubin"
os.makedirs(tmp, exist_ok=True)
for f in os.listdir(tmp):
    os.remove(os.path.join(tmp, f))
subprocess.run(["cuobjdump", "-xelf", "all", objp], cwd=tmp, capture_output=True)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
names = re.findall(r"\.global\s+(\S+)", dis)
dem = {n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip() for n in set(names)}
target = [n for n in dem if pat in dem[n]]
assert target, "no function matching %r" % pat
fn = target[0]
# walk the listing of that function: '//## File "...", line N' comments precede instructions
lines, cur_line, inside = [], None, False
for ln in dis.splitlines():
    if re.match(r"\s*\.text\.%s:" % re.escape(fn), ln) or ln.strip() == fn + ":":
        inside = True
        continue
    if inside and re.match(r"\s*\.section|\s*\.text\.", ln) and fn not in ln:
        if lines:
            break
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur_line = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
        lines.append(cur_line)
rows = k["rows"]
print("# %s\n# sass rows %d, disasm instructions %d" % (k["name"], len(rows), len(lines)))
n = min(len(rows), len(lines))
agg = {}
tot_i = tot_s = 0
for i in range(n):
    key = lines[i]
    a = agg.setdefault(key, [0, 0, {}])
    ie, st = int(rows[i][iex]), int(rows[i][ist])
    a[0] += ie
    a[1] += st
    tot_i += ie
    tot_s += st
    for c in stall_cols:
        v = int(rows[i][c])
        if v:
            a[2][H[c]] = a[2].get(H[c], 0) + v
src_cache = {}
def src(key):
    if key is None:
        return ""
    f, l = key
    if f not in src_cache:
        p = os.path.join(ROOT, "stereo_match_b200", "csrc", f)
        src_cache[f] = open(p).read().splitlines() if os.path.exists(p) else []
    return src_cache[f][l - 1].strip()[:90] if 0 < l <= len(src_cache[f]) else ""
print("%-26s %7s %7s  %s" % ("file:line", "inst%", "stall%", "top stalls | source"))
for key, (ie, st, sd) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    if ie * 1000 < tot_i and st * 1000 < tot_s:
        continue
    top = ",".join("%s:%d" % (s.replace("stall_", ""), v * 100 // max(st, 1)) for s, v in sorted(sd.items(), key=lambda kv: -kv[1])[:3])
    print("%-26s %6.2f%% %6.2f%%  %-34s | %s" % ("%s:%d" % key if key else "?", 100.0 * ie / tot_i, 100.0 * st / max(tot_s, 1), top, src(key)))
