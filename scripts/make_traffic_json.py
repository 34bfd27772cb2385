"""profiles/dominant_traffic.json from an `ncu --set full` raw page: DRAM bytes per launch of the kernels that can be the
dominant one.  usage: make_traffic_json.py RAW_CSV SOURCE_NOTE"""
import csv, json, re, sys
raw, note = sys.argv[1:3]
rows = list(csv.reader(open(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
c = {n: hdr.index(n) for n in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum")}
scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
tscale = {"ms": 1e3, "us": 1.0, "ns": 1e-3, "s": 1e6}
classes = {"shade": r"k_shade<", "resolve": r"k_resolve<|k_pixels<", "cull_secondary": r"k_cull<float, 0, SpawnSource", "cull_primary": r"k_cull<float, 1, (Gen)?CameraSource",
           "gen": r"k_gen<", "primary_points": r"k_primary_points<"}
out = {"source": note, "note": "per kernel class: mean of dram__bytes_read.sum + dram__bytes_write.sum over the captured launches of that kernel "
       "(one per recursion level of a frame); "
       "bench.py's kernel_bytes_per_launch / kernel_flops_per_launch are likewise means over that kernel's launches", "classes": {}}
for cls, rx in classes.items():
    launches = []
    for r in data:
        if not re.search(rx, r[c["Kernel Name"]]):
            continue
        f = lambda k: float(r[c[k]].replace(",", ""))
        launches.append({"us": f("gpu__time_duration.sum") * tscale[units[c["gpu__time_duration.sum"]]],
                         "dram_read_bytes": f("dram__bytes_read.sum") * scale[units[c["dram__bytes_read.sum"]]],
                         "dram_write_bytes": f("dram__bytes_write.sum") * scale[units[c["dram__bytes_write.sum"]]]})
    if launches:
        out["classes"][cls] = {"kernel": rx, "launches": launches,
                               "dram_bytes_per_launch": sum(l["dram_read_bytes"] + l["dram_write_bytes"] for l in launches) / len(launches)}
json.dump(out, open("profiles/dominant_traffic.json", "w"), indent=1)
print({k: round(v["dram_bytes_per_launch"] / 1e9, 3) for k, v in out["classes"].items()})
