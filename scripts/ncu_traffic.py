"""`ncu --page raw --csv` dump -> profiles/*_dram_traffic.json: DRAM bytes per launch of every captured kernel
(dram__bytes_read.sum + dram__bytes_write.sum), which bench.py reports as roofline.traffic."""
import csv, json, sys

rows = list(csv.reader(open(sys.argv[1])))
h, units = rows[0], rows[1]
mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
ki = h.index("Kernel Name")
ri, wi, ti = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum"), h.index("gpu__time_duration.sum")
tmul = {"ns": 1e-3, "us": 1.0, "ms": 1e3}[units[ti]]
out = {}
for r in rows[2:]:
    if len(r) != len(h):
        continue
    name = r[ki].split("(")[0].replace("pdt::", "")
    if name in out:
        continue
    rd, wr = float(r[ri]) * mult[units[ri]], float(r[wi]) * mult[units[wi]]
    out[name] = {"dram_read_bytes": rd, "dram_write_bytes": wr, "traffic_bytes": rd + wr,
                 "duration_us_under_ncu": float(r[ti]) * tmul}
json.dump({"source": sys.argv[3] if len(sys.argv) > 3 else "ncu --set full --clock-control none", "kernels": out},
          open(sys.argv[2], "w"), indent=1)
print(json.dumps(out, indent=1))
