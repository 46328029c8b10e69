"""Regenerates profiles/: SASS listings of the kernels that run in the headline configuration and
text summaries of the ncu captures in gpurun_out/ (the .ncu-rep files themselves are scratch).

    python tools/make_profiles.py [round_tag]
"""
import csv
import io
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "stereo_match_b200", "_obj")
OUT = os.path.join(ROOT, "profiles")
TAG = sys.argv[1] if len(sys.argv) > 1 else "r02"

# (object file, regex on the demangled kernel name, output name)
KERNELS = [
    # the headline sweeps: the lean kernel, 128 threads (the instantiation bench.py and the ncu captures run)
    # the headline sweeps: the lean kernel with 32-row bands (256 threads), the instantiation bench.py and ncu run
    ("sgbm_wave8.o", r"k_wave8<8, 256, false>", "k_wave8_G8_256_pass1"),
    ("sgbm_wave8.o", r"k_wave8<8, 256, true>", "k_wave8_G8_256_pass2_wta"),
    ("sgbm_wave8.o", r"k_wave8<8, 128, false>", "k_wave8_G8_128_pass1"),
    ("sgbm_wave8.o", r"k_wave8<8, 128, true>", "k_wave8_G8_128_pass2_wta"),
    ("sgbm_wave8.o", r"k_wave8<16, 256, true>", "k_wave8_G16_256_pass2_wta"),
    # the generic template (every other mode / range), 128-thread instantiations where the host picks them
    ("sgbm_wave_g8s.o", r"k_wave<8, 8, 128, false, true, false, true, true>", "k_wave_R8_G8_128_pass1"),
    ("sgbm_wave_g8s.o", r"k_wave<8, 8, 128, false, true, true, true, true>", "k_wave_R8_G8_128_pass2_wta"),
    ("sgbm_wave_g8.o", r"k_wave<4, 8, 256, false, false, true, false, true>", "k_wave_R4_G8_sgbm_last"),
    ("sgbm_wave_g8.o", r"k_wave<8, 8, 256, false, true, true, false, false>", "k_wave_R8_G8_hh4_pass2_wta"),
    ("wls.o", r"k_wls_solve", "k_wls_solve"),
    ("sgbm_cost.o", r"k_cost_fast2<5, 16, 1>", "k_cost_fast2_BS5"),
    ("sgbm_cost.o", r"k_descriptor_planes8b", "k_descriptor_planes8b"),
    ("sgbm_cost.o", r"k_cost_fast<5, 16, false, true, 1>", "k_cost_fast_BS5"),
    ("sgbm_cost.o", r"k_cost_fast<5, 16, false, true, 3>", "k_cost_fast_BS5_color"),
    ("cloud.o", r"k_cloud_write<short>", "k_cloud_write_i16"),
    ("sgbm_cost.o", r"k_descriptors", "k_descriptors"),
    ("sgbm_cost.o", r"k_descriptor_planes8", "k_descriptor_planes8"),
    ("sgbm_post.o", r"k_finish_rec_x2", "k_finish_rec_x2"),
    ("sgbm_post.o", r"k_median3_x2", "k_median3_x2"),
    ("speckle.o", r"k_speckle_runs", "k_speckle_runs"),
    ("speckle.o", r"k_speckle_merge", "k_speckle_merge"),
    ("speckle.o", r"k_speckle_count", "k_speckle_count"),
    ("speckle.o", r"k_speckle_apply", "k_speckle_apply"),
    ("geom.o", r"k_remap<1>", "k_remap_c1"),
    ("geom.o", r"k_reproject<short>", "k_reproject_i16"),
    ("bm.o", r"k_bm_fast<15>", "k_bm_fast_BS15"),
]


def sass_listings():
    os.makedirs(os.path.join(OUT, "sass"), exist_ok=True)
    summary = []
    for obj, pat, name in KERNELS:
        path = os.path.join(OBJ, obj)
        syms = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
        names = re.findall(r"Function : (\S+)", syms)
        dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
        hits = [m for m, d in zip(names, dem) if re.search(pat, d)]
        for k, m in enumerate(hits[:2]):
            txt = subprocess.run(["cuobjdump", "-sass", "-fun", m, path], capture_output=True, text=True).stdout
            lines = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", ln) for ln in txt.splitlines()
                     if re.match(r"\s+/\*[0-9a-f]{4,5}\*/", ln) or "Function" in ln]
            ops = {}
            for ln in lines:
                mm = re.match(r"\s+/\*[0-9a-f]{4,5}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
                if mm:
                    op = mm.group(1).split(".")[0]
                    ops[op] = ops.get(op, 0) + 1
            fn = "%s%s.sass" % (name, "" if k == 0 else "_%d" % k)
            with open(os.path.join(OUT, "sass", fn), "w") as f:
                f.write("// %s\n// static opcode histogram: %s\n" % (
                    subprocess.run(["c++filt", m], capture_output=True, text=True).stdout.strip(),
                    ", ".join("%s %d" % kv for kv in sorted(ops.items(), key=lambda kv: -kv[1])[:16])))
                f.write("\n".join(lines) + "\n")
            summary.append((fn, sum(ops.values()), ops))
    return summary


RAW_METRICS = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]


def ncu_summary(rep, out_name):
    p = os.path.join(ROOT, "gpurun_out", rep)
    if not os.path.exists(p):
        return False
    raw = subprocess.run(["ncu", "-i", p, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    if len(rows) < 3:
        return False
    H = rows[0]
    with open(os.path.join(OUT, out_name), "w") as f:
        f.write("# ncu --set full --clock-control none, from gpurun_out/%s (one launch per kernel, replayed by ncu;\n"
                "# durations are cold-cache and serialised -- the bench numbers come from CUDA events, not from here)\n" % rep)
        ki = H.index("Kernel Name")
        for r in rows[2:]:
            f.write("\nkernel: %s\n" % r[ki])
            for m in RAW_METRICS:
                if m in H:
                    f.write("  %-72s %s %s\n" % (m, r[H.index(m)], rows[1][H.index(m)]))
            for c in H:
                if "issue_stalled" in c and c.endswith("per_issue_active.ratio"):
                    f.write("  stall %-66s %s\n" % (c.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), r[H.index(c)]))
    return True


def traffic_json(rep, out_name):
    """dram bytes per k_wave launch (read + write) from the full capture -> profiles/traffic_<tag>.json,
    which bench.py quotes as roofline.traffic."""
    import json
    p = os.path.join(ROOT, "gpurun_out", rep)
    if not os.path.exists(p):
        return False
    raw = subprocess.run(["ncu", "-i", p, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    if len(rows) < 3:
        return False
    H = rows[0]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    per = []
    for r in rows[2:]:
        tot = 0.0
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = H.index(m)
            tot += float(r[i].replace(",", "")) * scale[rows[1][i]]
        per.append({"kernel": re.sub(r"\(.*", "", r[H.index("Kernel Name")]), "dram_bytes": tot,
                    "grid": r[H.index("launch__grid_size")]})
    with open(os.path.join(OUT, out_name), "w") as f:
        json.dump({"source": "ncu --set full --clock-control none, gpurun_out/%s, tools/run_sgbm_once.py 128 (cfg3, batch 128)" % rep,
                   "batch": 128, "launches": per, "avg_dram_bytes_per_launch": sum(x["dram_bytes"] for x in per) / len(per)}, f, indent=1)
    return True


def launch_list(csvname, out_name):
    p = os.path.join(ROOT, "gpurun_out", csvname)
    if not os.path.exists(p):
        return False
    rows = list(csv.reader(open(p)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"]
    if not hdr:
        return False
    H = rows[hdr[0]]
    ki, vi = H.index("Kernel Name"), H.index("Metric Value")
    agg = {}
    order = []
    for r in rows[hdr[0] + 1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except (ValueError, IndexError):
            continue
        k = re.sub(r"\(.*", "", r[ki])
        if k not in agg:
            agg[k] = []
            order.append(k)
        agg[k].append(v)
    tot = sum(sum(v) for v in agg.values())
    with open(os.path.join(OUT, out_name), "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none (launch list of one bench.py run, from gpurun_out/%s)\n" % csvname)
        f.write("# per-launch times are cold-cache and serialised: compare SHARES with bench.py's stage shares, not absolutes\n")
        f.write("%-64s %6s %12s %12s %7s\n" % ("kernel", "n", "avg_us", "total_us", "share"))
        for k in order:
            v = agg[k]
            f.write("%-64s %6d %12.1f %12.1f %6.1f%%\n" % (k[:64], len(v), sum(v) / len(v) / 1e3, sum(v) / 1e3, 100 * sum(v) / tot))
    return True


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    for fn, n, ops in sass_listings():
        print("sass/%s: %d instructions" % (fn, n))
    for rep, out in (("wave_%s.ncu-rep" % TAG, "ncu_k_wave_%s.txt" % TAG), ("cost_%s.ncu-rep" % TAG, "ncu_k_cost_%s.txt" % TAG),
                     ("all_%s.ncu-rep" % TAG, "ncu_all_kernels_%s.txt" % TAG)):
        print(out, ncu_summary(rep, out))
    print("ncu_k_bm", ncu_summary("bm_%s.ncu-rep" % TAG, "ncu_k_bm_%s.txt" % TAG))
    print("traffic", traffic_json("wave_%s.ncu-rep" % TAG, "traffic_%s.json" % TAG))
    print("launches", launch_list("launches_%s.csv" % TAG, "launches_%s.txt" % TAG))
