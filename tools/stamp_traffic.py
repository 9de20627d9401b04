#!/usr/bin/env python
"""profiles/stamp_traffic.json from an `ncu --set full` capture of csim_stamp_spec taken on bench.py's own
command line:  dram__bytes_read.sum + dram__bytes_write.sum of one launch, and the launch size.
usage: tools/stamp_traffic.py <report.ncu-rep> <netlist> <instances>"""
import csv
import io
import json
import os
import subprocess
import sys

rep, netlist, inst = sys.argv[1], sys.argv[2], int(sys.argv[3])
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, v = rows[0], rows[1], rows[2]
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def get(name):
    i = hdr.index(name)
    return float(v[i].replace(",", "")) * SCALE[units[i]]


assert "csim_stamp_spec" in v[hdr.index("Kernel Name")]
grid, block = int(float(v[hdr.index("launch__grid_size")])), int(float(v[hdr.index("launch__block_size")]))
assert grid * block >= inst > (grid - 1) * block, (grid, block, inst)
out = {"netlist": netlist, "instances": inst, "dram_bytes_read": get("dram__bytes_read.sum"),
       "dram_bytes_write": get("dram__bytes_write.sum"),
       "dram_bytes_per_launch": get("dram__bytes_read.sum") + get("dram__bytes_write.sum"),
       "source": "ncu --set full --clock-control none, one launch of csim_stamp_spec<TRAN>: " + os.path.basename(rep)}
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
json.dump(out, open(os.path.join(root, "profiles", "stamp_traffic.json"), "w"), indent=1)
print(json.dumps(out))
