"""Writes profiles/se2_traffic.json (the `roofline.traffic` figure of bench.py) from an ncu report of the volume
kernel:  python tools/make_traffic_json.py gpurun_out/k1_r2_final.ncu-rep profiles/r2_summary.md"""
import csv, json, os, subprocess, sys
rep, source_note = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
def get(name):
    i = hdr.index(name)
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
    return float(vals[i]) * scale
rd, wr = get("dram__bytes_read.sum"), get("dram__bytes_write.sum")
kname = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "se2_fast_kernel"
out = {"kernel": kname, "dram_bytes_per_launch": rd + wr, "dram_bytes_read": rd, "dram_bytes_write": wr,
       "algorithmic_bytes_per_launch": 4096 * 4096 * 73,
       "source": f"{source_note} ({os.path.basename(rep)}: ncu --set full, one launch of the 4096x4096x72 build; dram__bytes_read.sum + dram__bytes_write.sum)",
       "generated_by": "tools/make_traffic_json.py"}
json.dump(out, open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "se2_traffic.json"), "w"), indent=1)
print(out)
