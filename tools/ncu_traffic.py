"""From the `ncu --set full` captures of the kernels bench.py dispatches (gpurun_out/r02_c2_<op>_<density>.ncu-rep,
made by tools/gpu_ncu_c2.sh) write
  profiles/r02_c2_dram_traffic.json   dram__bytes_read.sum + dram__bytes_write.sum per launch, per (op, density)
                                      -- what bench.py reports as roofline.traffic
  profiles/r02_<name>_ncu_summary.txt the metrics of tools/ncu_summary.py for every capture found
usage (on the GPU box, where ncu reads the reports): python tools/ncu_traffic.py gpurun_out"""
import csv
import glob
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out")
out_dir = os.path.join(ROOT, sys.argv[2] if len(sys.argv) > 2 else "gpurun_out")
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
kernels = []
for rep in sorted(glob.glob(os.path.join(src, "r02_*.ncu-rep"))):
    name = os.path.basename(rep)[:-len(".ncu-rep")]
    summary = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), rep],
                             capture_output=True, text=True).stdout
    with open(os.path.join(out_dir, name + "_ncu_summary.txt"), "w") as f:
        f.write(f"{name}: ncu --set full --clock-control none, one steady-state launch (tools/gpu_ncu_c2.sh)\n" + summary)
    m = re.match(r"r02_c2_(spmm|sddmm)_d(\d+)$", name)
    if not m:
        continue
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    get = lambda k: float(vals[hdr.index(k)].replace(",", "")) * UNIT.get(units[hdr.index(k)], 1.0)
    kernels.append({"op": m.group(1), "density": int(m.group(2)) / 1000.0,
                    "kernel": re.sub(r"\(.*", "", vals[hdr.index("Kernel Name")]),
                    "dram_bytes": get("dram__bytes_read.sum") + get("dram__bytes_write.sum"),
                    "duration_us_under_ncu": float(vals[hdr.index("gpu__time_duration.sum")].replace(",", "")) *
                    {"us": 1.0, "ms": 1e3, "ns": 1e-3}.get(units[hdr.index("gpu__time_duration.sum")], 1.0),
                    "source": f"profiles/{name}_ncu_summary.txt"})
with open(os.path.join(out_dir, "r02_c2_dram_traffic.json"), "w") as f:
    json.dump({"what": "dram__bytes_read.sum + dram__bytes_write.sum of ONE steady-state launch of the kernel bench.py "
                       "dispatches for each C2 case (ncu --set full, cold cache, serialised)", "kernels": kernels}, f, indent=1)
print(json.dumps(kernels, indent=1))
