#!/usr/bin/env python
"""Condense `ncu -i X.ncu-rep --page raw --csv` into the table kept under profiles/ (one line per kernel):
device time, grid, registers, DRAM bytes and throughput, tensor-pipe activity, achieved warps.

    python tools/ncu_summary.py gpurun_out/prof_steps_raw.csv > profiles/ncu_steps_summary.md
"""
import csv
import sys

COLS = [
    ("gpu__time_duration.sum", "time us", 1e-3),
    ("launch__grid_size", "grid", 1),
    ("launch__block_size", "block", 1),
    ("launch__registers_per_thread", "regs", 1),
    ("dram__bytes_read.sum", "DRAM rd MB", 1e-6),
    ("dram__bytes_write.sum", "DRAM wr MB", 1e-6),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %peak", 1),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %act", 1),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps %", 1),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %", 1),
    ("l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum", "TMA ld MB", 1e-6),
]


def main() -> int:
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    cols = [(k, n, s) for k, n, s in COLS if k in idx]
    print("| # | kernel | " + " | ".join(n for _, n, _ in cols) + " |")
    print("|---|---|" + "---|" * len(cols))
    for j, r in enumerate(data):
        name = r[idx["Kernel Name"]].split("(")[0].replace("ttd::", "")
        cells = []
        for k, n, s in cols:
            v = r[idx[k]].replace(",", "")
            try:
                f = float(v)
                u = units[idx[k]]
                if k == "gpu__time_duration.sum":
                    f = f / 1e3 if u in ("ns", "nsecond") else (f if u in ("us", "usecond") else f * 1e3 if u in ("ms", "msecond") else f)
                    cells.append(f"{f:.2f}")
                elif s != 1:
                    mult = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, s)
                    cells.append(f"{f * mult:.3f}")
                else:
                    cells.append(f"{f:.1f}" if f != int(f) else str(int(f)))
            except ValueError:
                cells.append(v or "-")
        print(f"| {j} | `{name}` | " + " | ".join(cells) + " |")
    return 0


if __name__ == "__main__":
    sys.exit(main())
