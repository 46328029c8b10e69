"""Per-opcode DYNAMIC instruction table of one kernel of an ncu capture (source page, `Instructions Executed` per SASS
line), normalised per warp-step of the wavefront kernels, next to its stall-reason shares.

    python tools/ncu_opcodes.py gpurun_out/wave_r02.ncu-rep <kernel index> <warp-steps per launch> > profiles/opcodes_...txt

warp-steps per launch of the cfg3 sweeps at batch 128: pairs * bands * warps * steps = 128 * 12 * 8 * 1176 = 14450688 (32-row bands; 16-row bands: 128 * 24 * 4 * 1144 = 14057472).
"""
import collections
import csv
import io
import re
import subprocess
import sys

rep, kidx, steps = sys.argv[1], int(sys.argv[2]), float(sys.argv[3])
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
kern, cur, H = [], None, None
for r in csv.reader(io.StringIO(raw)):
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}
        kern.append(cur)
    elif r and r[0] == "Address":
        H = r
    elif cur is not None and r and r[0].startswith("0x"):
        cur["rows"].append(r)
ded = []
for k in kern:                      # the page lists every kernel twice
    if ded and ded[-1]["name"] == k["name"] and len(ded[-1]["rows"]) == len(k["rows"]):
        continue
    ded.append(k)
k = ded[kidx]
iS, iE, iSt = H.index("Source"), H.index("Instructions Executed"), H.index("Warp Stall Sampling (All Samples)")
ops, pipes = collections.Counter(), collections.Counter()
ALU = ("VIMNMX", "VIADDMNMX", "IADD3", "LOP3", "PRMT", "SEL", "SHF", "LEA", "VIADD", "ISETP", "PLOP3", "MOV", "FLO", "BREV", "POPC", "VABSDIFF", "P2R", "R2P")
FMA = ("IMAD", "FFMA", "FMUL", "FADD", "HFMA2")
LSU = ("LDS", "STS", "LDG", "STG", "SHFL", "LDGSTS", "RED", "ATOM", "CCTL", "LDC", "SYNCS", "VOTE", "LDL", "STL")
tot = 0
for r in k["rows"]:
    op = re.sub(r"^@!?U?P\d\s+", "", r[iS].strip()).split()[0]
    n = int(r[iE])
    ops[op] += n
    tot += n
    base = op.split(".")[0]
    pipes["alu" if base.startswith(ALU) else "fma" if base.startswith(FMA) else "lsu/mio" if base.startswith(LSU) else "other"] += n
out = sys.stdout
out.write("# %s\n" % k["name"])
out.write("# dynamic warp-instructions: %d = %.1f per warp-step (%.0f warp-steps); per packed register and direction (32 per step): %.2f\n"
          % (tot, tot / steps, steps, tot / steps / 32))
out.write("# ideal recurrence per register and direction: VIMNMX3 + VIADDMNMX + IADD3 + 1/2 VIMNMX3 (min) + 1/2..3/4 add (S) = 4 .. 4.25\n")
out.write("# by pipe: " + ", ".join("%s %.1f" % (p, v / steps) for p, v in pipes.most_common()) + "\n")
out.write("%-34s %10s %8s\n" % ("opcode", "per step", "share"))
for op, n in ops.most_common(40):
    out.write("%-34s %10.2f %7.1f%%\n" % (op, n / steps, 100.0 * n / tot))
ts = sum(int(r[iSt]) for r in k["rows"])
out.write("\n# stall reasons (share of warp samples)\n")
for i, c in enumerate(H):
    if c.startswith("stall_") and "Not Issued" not in c:
        v = sum(int(r[i]) for r in k["rows"])
        if v > 0.01 * ts:
            out.write("%-28s %6.1f%%\n" % (c, 100.0 * v / ts))
