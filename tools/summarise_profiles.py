"""Turns the raw captures of tools/profile_round.sh (gpurun_out/<tag>_*) into the committed summaries under profiles/.
usage: python tools/summarise_profiles.py <tag>          (needs `ncu` on PATH to read the .ncu-rep files)"""
import csv
import io
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")
tag = sys.argv[1] if len(sys.argv) > 1 else "r1b"

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
           "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
           "sm__inst_executed_pipe_tensor.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "smsp__inst_executed.sum", "lts__t_sector_hit_rate.pct", "launch__grid_size", "launch__block_size", "launch__cluster_size",
           "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max",
           "smsp__cycles_active.avg", "sm__cycles_active.avg"]


def raw_page(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    return hdr, units, body


def full_summary(rep, title, f):
    hdr, units, body = raw_page(rep)
    f.write("# %s\n" % title)
    ik = hdr.index("Kernel Name")
    for r in body:
        f.write("\nKernel Name%s%s\n" % (" " * 62, r[ik]))
        for m in METRICS:
            if m in hdr:
                i = hdr.index(m)
                f.write("%-72s %s %s\n" % (m, r[i], units[i]))


def launch_shares(path, f, header):
    rows = list(csv.reader(l for l in open(path) if not l.startswith("==")))
    hdr = rows[0]
    ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = {}
    for r in rows[1:]:
        if len(r) != len(hdr):
            continue
        v = float(r[iv].replace(",", ""))
        v = v / 1000.0 if r[iu] in ("ns", "nsecond") else (v * 1000.0 if r[iu] in ("ms", "msecond") else v)
        a = agg.setdefault(r[ik], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    f.write(header)
    f.write("# kernel | launches | total us | avg us | share\n")
    for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
        f.write("%-95s %5d %10.1f %9.1f %6.1f%%\n" % (k[:95], n, us, us / n, 100.0 * us / tot))


if __name__ == "__main__":
    os.makedirs(PROF, exist_ok=True)
    with open(os.path.join(PROF, "%s_ncu_full_tc_kernels.txt" % tag), "w") as f:
        f.write("# ncu --set full --clock-control none --import-source on (B200), one launch per kernel; extracted with tools/summarise_profiles.py\n")
        f.write("# (ncu -i X.ncu-rep --page raw --csv).  Commands: tools/profile_round.sh.  dram__bytes_* are per launch.\n")
        rep = os.path.join(OUT, "%s_mlp.ncu-rep" % tag)
        if os.path.exists(rep):
            full_summary(rep, "MLP kernels: tools/time_train_mlp.py 4096 2 (4,096 rays x 128 samples, bf16 path), -k regex:tc_(fwd|bwd|wgrad)_kernel -s 6 -c 3", f)
        rep = os.path.join(OUT, "%s_chamfer.ncu-rep" % tag)
        if os.path.exists(rep):
            f.write("\n")
            full_summary(rep, "Chamfer sweeps: tools/time_chamfer.py 28830 --raster (P = 28,830 per cloud, raster-order clouds: sub-sampled screen, screen and exact sweep over the stage lists), -k regex:chamfer_tc_kernel -s 3 -c 3", f)
        for name in ("time_train_mlp.json", "time_chamfer.json"):
            p = os.path.join(OUT, "%s_%s" % (tag, name))
            if os.path.exists(p):
                f.write("\n# %s (the same command WITHOUT ncu, CUDA events):\n" % name)
                f.write("".join("#   " + l for l in open(p)))
    p = os.path.join(OUT, "%s_launches.csv" % tag)
    if os.path.exists(p):
        with open(os.path.join(PROF, "%s_launch_shares.txt" % tag), "w") as f:
            launch_shares(p, f, "# ncu --metrics gpu__time_duration.sum --clock-control none -c 900: python bench.py --steps 2 --warmup 3 --no-cpu-baseline\n"
                                "# (B200, cfg2, bf16 path, CUDA-graph step: graph replays are profiled kernel by kernel; the 900 launches = set-up, eager warm-up\n"
                                "#  passes of the captures and the replayed steps; cold-cache serialised times: the SHARE per kernel is what counts)\n")
    for name in ("timeline.txt", "bench_plain.json"):
        p = os.path.join(OUT, "%s_%s" % (tag, name))
        if os.path.exists(p):
            with open(os.path.join(PROF, "%s_%s" % (tag, name)), "w") as f:
                f.write(open(p).read())
    print("profiles written for", tag)
