"""Markdown table from an `ncu --set full` report: one row per captured launch.
    python tools/summarize_ncu.py gpurun_out/census.ncu-rep > profiles/rNN_ncu_census.md
"""
import csv
import io
import subprocess
import sys

COLS = [("gpu__time_duration.sum", "us", 1.0), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe %", 1.0),
        ("dram__bytes_read.sum", "DRAM rd MB", None), ("dram__bytes_write.sum", "DRAM wr MB", None),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %", 1.0),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 %", 1.0),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %", 1.0),
        ("launch__grid_size", "grid", 1.0), ("launch__registers_per_thread", "regs", 1.0),
        ("launch__shared_mem_per_block_dynamic", "dyn smem KB", None)]
TO_MB = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
TO_US = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def main(path):
    if path.endswith(".csv"):          # already exported with `ncu -i X --page raw --csv`
        out = open(path).read()
    else:
        out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    idx = {}
    for want, _, _ in COLS:
        for i, h in enumerate(hdr):
            if h == want:
                idx[want] = i
    kname = hdr.index("Kernel Name")
    print("| # | kernel | " + " | ".join(lbl for _, lbl, _ in COLS) + " |")
    print("|---|---|" + "---|" * len(COLS))
    for n, r in enumerate(body):
        cells = []
        for want, lbl, _ in COLS:
            i = idx.get(want)
            if i is None or r[i] == "":
                cells.append("-")
                continue
            v = float(r[i].replace(",", ""))
            u = units[i]
            if "MB" in lbl:
                v *= TO_MB.get(u, 1.0)
            elif "KB" in lbl:
                v *= {"byte": 1e-3, "Kbyte": 1.0, "Mbyte": 1e3}.get(u.split("/")[0], 1.0)
            elif lbl == "us":
                v *= TO_US.get(u, 1.0)
            cells.append(f"{v:.1f}" if v < 1e6 else f"{v:.3g}")
        name = r[kname].split("(")[0].replace("void ", "").replace("b200::", "")
        print(f"| {n} | `{name}` | " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
