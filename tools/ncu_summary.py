#!/usr/bin/env python
"""Summarise `ncu -i rep --page raw --csv` files: python tools/ncu_summary.py out.txt out.json raw1.csv [raw2.csv ...]
Per kernel (first launch of each name): time, DRAM words per point, L2 words, GB/s, pipe / issue / stall figures."""
import csv
import json
import sys

NPTS3, NPTS2 = 1024 * 1024 * 64, 1024 * 1024
MUL = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
TU = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def main():
    out_txt, out_json, files = sys.argv[1], sys.argv[2], sys.argv[3:]
    lines = ["# ncu --set full --clock-control none, first captured launch of each kernel; bench.py --steps 1 --warmup 1 (B200, 1024x1024x64, T+S)",
             "# words = 8-byte words per grid point (3-D kernels: per 3-D point; k_step2d*: per 2-D point); dram%% = gpu__dram_throughput of the nominal peak (8.2 TB/s)",
             "%-44s %8s %6s %6s %6s %6s %5s %5s %5s %5s %5s %5s %4s %7s  %s" % ("kernel", "us", "rd_w", "wr_w", "l2_w", "GB/s", "dram%", "lts%", "l1%", "sm%", "ipc", "fp64%", "regs", "inst/pt", "top stalls (cycles per issued instruction)")]
    traffic = {}
    seen = set()
    for f in files:
        rows = list(csv.reader(open(f)))
        hdr, units, data = rows[0], rows[1], rows[2:]
        col = {n: i for i, n in enumerate(hdr)}

        def g(r, n):
            try:
                return float(r[col[n]].replace(",", ""))
            except Exception:
                return float("nan")
        for r in data:
            name = r[col["Kernel Name"]]
            short = name.replace("void ", "").split("(")[0]
            if short in seen:
                continue
            seen.add(short)
            npts = NPTS2 if "step2d" in short else NPTS3
            t = g(r, "gpu__time_duration.sum") * TU.get(units[col["gpu__time_duration.sum"]], 1.0)
            rd = g(r, "dram__bytes_read.sum") * MUL.get(units[col["dram__bytes_read.sum"]], 1)
            wr = g(r, "dram__bytes_write.sum") * MUL.get(units[col["dram__bytes_write.sum"]], 1)
            l2 = g(r, "lts__t_sectors.sum") * 32
            st = []
            for h in hdr:
                if "issue_stalled" in h and h.endswith("per_issue_active.ratio") and "not_issued" not in h:
                    try:
                        st.append((float(r[col[h]]), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))
                    except Exception:
                        pass
            st.sort(reverse=True)
            stalls = " ".join("%s=%.1f" % (n, v) for v, n in st[:4] if n != "selected")
            lines.append("%-44s %8.1f %6.2f %6.2f %6.2f %6.0f %5.1f %5.1f %5.1f %5.1f %5.2f %5.1f %4.0f %7.0f  %s" % (
                short[:44], t, rd / 8 / npts, wr / 8 / npts, l2 / 8 / npts, (rd + wr) / t / 1e3, g(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
                g(r, "lts__throughput.avg.pct_of_peak_sustained_elapsed"), g(r, "l1tex__throughput.avg.pct_of_peak_sustained_active"),
                g(r, "sm__throughput.avg.pct_of_peak_sustained_elapsed"), g(r, "sm__inst_issued.avg.per_cycle_active"),
                g(r, "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"), g(r, "launch__registers_per_thread"),
                g(r, "smsp__inst_executed.sum") * 32 / npts, stalls))
            traffic[short] = {"dram_bytes": rd + wr, "us": t}
    open(out_txt, "w").write("\n".join(lines) + "\n")
    json.dump(traffic, open(out_json, "w"), indent=1)


if __name__ == "__main__":
    main()
