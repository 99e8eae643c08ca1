"""Writes profiles/ncu_traffic.json from an `ncu --set full` capture of the three MLP kernels (tools/profile_round.sh:
gpurun_out/<tag>_mlp.ncu-rep, `tools/time_train_mlp.py 4096 2`): dram__bytes_read.sum + dram__bytes_write.sum per launch of
tc_fwd_kernel<save>, tc_bwd_kernel and tc_wgrad_kernel at 4,096 rays = the DRAM traffic of the MLP kernels per cfg2 step.
bench.py reads the file for `roofline.traffic` and drops it when the recorded hash of csrc/mlp_tc.cu is not the tree's.
usage: python tools/ncu_traffic.py <tag>        (needs `ncu` on PATH to read the .ncu-rep; no GPU needed)"""
import csv
import hashlib
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
rep = os.path.join(ROOT, "gpurun_out", "%s_mlp.ncu-rep" % tag)
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units, body = rows[0], rows[1], rows[2:]
ik, ir, iw, it = (hdr.index(k) for k in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))


def to_bytes(v, unit):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]


kern = {}
for r in body:
    name = r[ik].split("(")[0]
    kern[name] = {"dram_read_bytes": to_bytes(r[ir], units[ir]), "dram_write_bytes": to_bytes(r[iw], units[iw]),
                  "ncu_duration": "%s %s" % (r[it], units[it])}
total = sum(k["dram_read_bytes"] + k["dram_write_bytes"] for k in kern.values())
src = hashlib.sha256(open(os.path.join(ROOT, "my-nope-nerf_b200", "csrc", "mlp_tc.cu"), "rb").read()).hexdigest()[:16]
out = {"bytes_per_step": total, "rays": 4096, "kernels": kern, "mlp_tc_sha16": src,
       "source": "ncu --set full --clock-control none, gpurun_out/%s_mlp.ncu-rep (tools/profile_round.sh %s), one launch of each kernel at 4,096 rays" % (tag, tag)}
json.dump(out, open(os.path.join(ROOT, "profiles", "ncu_traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
