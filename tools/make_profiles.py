"""Builds the committed profile summaries from one gpurun capture:
python tools/make_profiles.py TAG   (expects gpurun_out/bench_512_TAG.json or bench_TAG.json, launches_TAG.csv, prof_TAG_step.ncu-rep)"""
import collections, csv, json, re, subprocess, sys
tag = sys.argv[1]
rep = f"gpurun_out/prof_{tag}_step.ncu-rep"
import os
summ = ''; t = {}
if os.path.exists(rep):
    # 1. --set full summary
    summ = subprocess.run([sys.executable, "tools/ncu_summary.py", rep], capture_output=True, text=True).stdout
    open(f"profiles/{tag}_ncu_full_step512_summary.txt", "w").write(
        "# ncu --set full --clock-control none --import-source on, one RK substep of bench.py at 512^3 (12 launches)\n"
        "# GB_rd/GB_wr = dram__bytes_read.sum / dram__bytes_write.sum per launch; fp64% = sm__pipe_fp64_cycles_active; st_* = warp stall reasons per issue\n" + summ)
    # 2. traffic table
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(out.splitlines())); h, u = r[0], r[1]
    def gb(row, name):
        i = h.index(name); return float(row[i].replace(",", "")) / {"byte": 1e9, "Kbyte": 1e6, "Mbyte": 1e3, "Gbyte": 1.}[u[i]]
    t = {}
    for row in r[2:]:
        kn = row[h.index("Kernel Name")]
        v = round(gb(row, "dram__bytes_read.sum") + gb(row, "dram__bytes_write.sum"), 3)
        key = None
        if "pres2_out_rk_kernel" in kn: key = "pres2_out_rk_kernel"
        elif "rk_kernel" in kn: key = "rk_kernel"
        elif "evisc_kernel" in kn: key = "evisc_kernel"
        elif "smag2_diff_uvw_tile_kernel" in kn: key = "advec2_smag2_uvw_kernel"
        elif "smag2_scalar_tile_kernel" in kn: key = "advec2_smag2_c_kernel"
        elif "pres2_in_kernel" in kn: key = "pres2_in_kernel"
        elif "pres2_tdma_kernel" in kn: key = "pres2_tdma_kernel"
        elif "fft_pow2_col_kernel<(int)-1" in kn: key = "fft_fwd_y"
        elif "fft_pow2_col_kernel<1" in kn: key = "fft_bwd_y"
        elif "fft_pow2_kernel<(int)-1, 0" in kn: key = "fft_fwd_x"
        elif "fft_pow2_kernel<(int)-1, 1" in kn: key = "fft_fwd_y"
        elif "fft_pow2_kernel<1, 1" in kn: key = "fft_bwd_y"
        elif "fft_pow2_kernel<1, 0" in kn: key = "fft_bwd_x"
        if key: t[key] = v
    json.dump({"n": 512, "source": f"profiles/{tag}_ncu_full_step512_summary.txt (ncu --set full --clock-control none, one RK substep of bench.py at 512^3)",
               "gb_per_launch": t}, open(f"profiles/{tag[:3]}_traffic.json", "w"), indent=1)
# 3. launch list + summary
src = f"gpurun_out/launches_{tag}.csv"
open(f"profiles/{tag}_ncu_launches_bench512.csv", "w").write(open(src).read())
rows = [x for x in csv.reader(l for l in open(src) if l.startswith('"'))]
hh = rows[0]; ki, vi, ui = hh.index("Kernel Name"), hh.index("Metric Value"), hh.index("Metric Unit")
agg = collections.OrderedDict(); tot = 0.; n = 0
for x in rows[1:]:
    v = float(x[vi].replace(",", "")); ms = v / 1e6 if x[ui] == "ns" else (v / 1e3 if x[ui] == "us" else v)
    short = re.sub(r"\(.*", "", x[ki]).replace("void ", "").replace("mhh::<unnamed>::", "").replace("mhh::", "")[:70]
    a = agg.setdefault(short, [0, 0.]); a[0] += 1; a[1] += ms; tot += ms; n += 1
lines = ["# ncu --metrics gpu__time_duration.sum --clock-control none -c 400, python bench.py --steps 2 --warmup 1 --no-cpu-baseline (512^3)",
         f"# cold-cache, serialised launch times: compare SHARES with bench.py's live CUDA-event shares (profiles/{tag}_bench_512.json), not absolutes",
         f"# {n} launches captured, total {tot:.2f} ms", "%-72s %4s %10s %9s %7s" % ("kernel", "n", "ms_total", "ms_avg", "share")]
for k, (c, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    lines.append("%-72s %4d %10.3f %9.4f %7.4f" % (k, c, ms, ms / c, ms / tot))
open(f"profiles/{tag}_ncu_launches_bench512_summary.txt", "w").write("\n".join(lines) + "\n")
bsrc = f"gpurun_out/bench_512_{tag}.json" if os.path.exists(f"gpurun_out/bench_512_{tag}.json") else f"gpurun_out/bench_{tag}.json"
open(f"profiles/{tag}_bench_512.json", "w").write(open(bsrc).read())
print(summ); print("\n".join(lines[:16])); print(t)
