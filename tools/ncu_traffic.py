"""DRAM traffic per launch of the hot kernels, from `ncu --set full` reports (read here, no GPU):
    python tools/ncu_traffic.py profiles/r01_traffic.json fleet32=rep1.ncu-rep traj64=rep2.ncu-rep ...
Writes {label: {kernel, launches, dram_bytes_per_launch, dram_read, dram_write, time_us}} (medians over the captured
launches).  bench.py copies dram_bytes_per_launch into roofline.traffic for the matching workload."""
import csv, io, json, statistics, subprocess, sys

SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12,
         "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3, "second": 1e6}


def main():
    res = {}
    for spec in sys.argv[2:]:
        label, rep = spec.split("=", 1)
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(txt)))
        h, units = rows[0], rows[1]
        ik = h.index("Kernel Name")

        def col(name):
            i = h.index(name)
            return [float(r[i].replace(",", "")) * SCALE[units[i]] for r in rows[2:]]
        rd, wr, tm = col("dram__bytes_read.sum"), col("dram__bytes_write.sum"), col("gpu__time_duration.sum")
        res[label] = {"kernel": rows[2][ik], "launches": len(rows) - 2,
                      "dram_bytes_per_launch": statistics.median([a + b for a, b in zip(rd, wr)]),
                      "dram_read": statistics.median(rd), "dram_write": statistics.median(wr),
                      "time_us_under_ncu": statistics.median(tm), "report": rep.split("/")[-1]}
    json.dump(res, open(sys.argv[1], "w"), indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
