"""Record the DRAM traffic of one k_rejuvenate launch from an ncu raw page into profiles/traffic.json, stamped with the
hash of the kernel sources (run on the source state the capture was taken from).
usage: python tools/update_traffic.py <raw.csv> <shape> <particles> <launch_t> <source-name>"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

raw, shape, particles, launch_t, source = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
rows = list(csv.reader(open(raw)))
hdr, units = rows[0], rows[1]
d = dict(zip(hdr, rows[2]))
u = dict(zip(hdr, units))


def to_bytes(k):
    v, unit = float(d[k]), u[k].lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12}[unit]


path = os.path.join(ROOT, "profiles", "traffic.json")
try:
    t = json.load(open(path))
except (OSError, ValueError):
    t = {}
t["k_rejuvenate_" + shape] = {
    "dram_bytes": to_bytes("dram__bytes_read.sum") + to_bytes("dram__bytes_write.sum"),
    "dram_read": to_bytes("dram__bytes_read.sum"), "dram_write": to_bytes("dram__bytes_write.sum"),
    "launch_t": launch_t, "gpu_time_ms": float(d["gpu__time_duration.sum"]), "source": source,
    "src_sha": bench.kernel_source_hash(),
    "config": {"shape": shape, "particles": particles, "timesteps": 120, "moves": 5, "substeps": 5}}
json.dump(t, open(path, "w"), indent=1)
print(json.dumps(t["k_rejuvenate_" + shape]))
