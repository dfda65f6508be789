#!/usr/bin/env python
"""benchmarks/ncu_digest.py RAW.csv OUT.json ALGORITHMIC_BYTES [note] -- per-launch DRAM traffic of one `ncu --set full`
capture (`ncu -i X.ncu-rep --page raw --csv`), as bench.py reports it under roofline.traffic."""
import csv
import json
import sys

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main():
    raw, out, alg = sys.argv[1], sys.argv[2], float(sys.argv[3])
    rows = list(csv.reader(open(raw)))
    hdr, units, launches = rows[0], rows[1], rows[2:]

    def val(r, name):
        i = hdr.index(name)
        return float(r[i].replace(",", "")) * UNIT.get(units[i], 1)
    per = []
    for r in launches:
        per.append({"kernel": r[hdr.index("Kernel Name")], "duration_us": val(r, "gpu__time_duration.sum") / (1e3 if units[hdr.index("gpu__time_duration.sum")] == "ns" else 1) if units[hdr.index("gpu__time_duration.sum")] in ("ns", "us") else None,
                    "dram_read": val(r, "dram__bytes_read.sum"), "dram_write": val(r, "dram__bytes_write.sum"),
                    "grid": int(float(r[hdr.index("launch__grid_size")])), "registers": int(float(r[hdr.index("launch__registers_per_thread")]))})
    tot = sum(p["dram_read"] + p["dram_write"] for p in per) / len(per)
    d = {"source": raw, "launches_captured": len(per), "kernel": per[0]["kernel"], "dram_bytes_read_per_launch": sum(p["dram_read"] for p in per) / len(per),
         "dram_bytes_write_per_launch": sum(p["dram_write"] for p in per) / len(per), "dram_bytes_per_launch": tot,
         "algorithmic_bytes_per_launch": alg, "traffic_over_algorithmic": tot / alg if alg else None,
         "grid": per[0]["grid"], "registers_per_thread": per[0]["registers"],
         "note": sys.argv[4] if len(sys.argv) > 4 else "ncu --set full --clock-control none; cold, serialised launch"}
    json.dump(d, open(out, "w"), indent=1)
    print(json.dumps(d))


if __name__ == "__main__":
    main()
