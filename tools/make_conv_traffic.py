"""profiles/r02_conv_traffic.json from an ncu launch list with dram__bytes_read.sum / dram__bytes_write.sum
(usage: make_conv_traffic.py launches.csv batch out.json source-note)"""
import csv, io, json, sys
txt = [l for l in open(sys.argv[1]).read().splitlines() if not l.startswith("==")]
rows = list(csv.DictReader(io.StringIO("\n".join(txt))))
by = {}
for r in rows:
    by.setdefault(r["ID"], {})[r["Metric Name"]] = (float(r["Metric Value"]), r["Metric Unit"], r["Kernel Name"])
ids = sorted(by, key=int)
names = [by[i]["gpu__time_duration.sum"][2] for i in ids]
first = max(k for k, i in enumerate(ids) if "firstConv" in names[k] and k + 114 <= len(ids))
sel = ids[first:first + 114]
mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
val = lambda i, m: by[i][m][0] * mult[by[i][m][1]]
rd = sum(val(i, "dram__bytes_read.sum") for i in sel)
wr = sum(val(i, "dram__bytes_write.sum") for i in sel)
json.dump({"batch": int(sys.argv[2]), "dram_bytes_per_step": rd + wr, "dram_read_bytes": rd, "dram_write_bytes": wr, "launches": 114,
           "note": "sum of dram__bytes_read.sum + dram__bytes_write.sum over the 114 launches (first-layer kernel + 113 convIgemm) of "
                   "one BODY_25 forward, 656x368; ncu flushes caches between launches, so inter-layer L2 reuse is not credited",
           "source": sys.argv[4]}, open(sys.argv[3], "w"), indent=1)
print("batch %s: read %.3f GB, write %.3f GB" % (sys.argv[2], rd / 1e9, wr / 1e9))
