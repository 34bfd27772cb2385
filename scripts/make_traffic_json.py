"""profiles/dominant_traffic.json from an `ncu --set full` raw page: DRAM bytes per launch of the dominant kernel.
usage: make_traffic_json.py RAW_CSV KERNEL_REGEX KERNEL_CLASS SOURCE_NOTE"""
import csv, json, re, sys
raw, rx, cls, note = sys.argv[1:5]
rows = list(csv.reader(open(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
c = {n: hdr.index(n) for n in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum")}
scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
tscale = {"ms": 1e3, "us": 1.0, "ns": 1e-3, "s": 1e6}
launches = []
for r in data:
    if not re.search(rx, r[c["Kernel Name"]]):
        continue
    f = lambda k: float(r[c[k]].replace(",", ""))
    launches.append({"us": f("gpu__time_duration.sum") * tscale[units[c["gpu__time_duration.sum"]]],
                     "dram_read_bytes": f("dram__bytes_read.sum") * scale[units[c["dram__bytes_read.sum"]]],
                     "dram_write_bytes": f("dram__bytes_write.sum") * scale[units[c["dram__bytes_write.sum"]]]})
out = {"kernel_class": cls, "kernel": rx, "source": note, "launches": launches,
       "dram_bytes_per_launch": sum(l["dram_read_bytes"] + l["dram_write_bytes"] for l in launches) / max(len(launches), 1),
       "note": "mean of dram__bytes_read.sum + dram__bytes_write.sum over the captured launches of the kernel (one per recursion level of a frame); "
               "bench.py's kernel_bytes_per_launch is likewise a mean over that kernel's launches"}
json.dump(out, open("profiles/dominant_traffic.json", "w"), indent=1)
print(json.dumps(out)[:400])
