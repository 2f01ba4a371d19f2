"""Markdown table from an `ncu --page raw --csv` export: python tools/ncu_table.py raw.csv [first_id]
One row per launch (from launch ID first_id on): duration, pipe utilisations, occupancy, DRAM bytes, L2 hit rate."""
import csv
import re
import sys

COLS = [("gpu__time_duration.sum", "us"), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor %"),
        ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "issue %"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "XU %"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "FMA %"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps %"), ("launch__registers_per_thread", "regs"),
        ("dram__bytes_read.sum", "dram rd MB"), ("dram__bytes_write.sum", "dram wr MB"),
        ("lts__t_sector_hit_rate.pct", "L2 hit %"), ("smsp__inst_executed.sum", "warp inst")]
MULT = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
TIME = {"ns": 1e-3, "us": 1.0, "ms": 1e3}


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    first = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    h, u = rows[0], rows[1]
    ix = {n: h.index(n) for n, _ in COLS if n in h}
    print("| id | kernel | grid x block | " + " | ".join(l for _, l in COLS) + " |")
    print("|---|---|---|" + "---|" * len(COLS))
    for r in rows[2:]:
        if int(r[h.index("ID")]) < first:
            continue
        name = re.sub(r"\(.*", "", r[h.index("Kernel Name")]).replace("<unnamed>::", "").replace("void ", "")
        cells = []
        for n, _ in COLS:
            if n not in ix:
                cells.append("n/a")
                continue
            v, unit = r[ix[n]].replace(",", ""), u[ix[n]]
            try:
                f = float(v)
            except ValueError:
                cells.append(v)
                continue
            if "bytes" in n:
                f *= MULT.get(unit, 1.0)
                cells.append(f"{f:.2f}")
            elif n.startswith("gpu__time"):
                cells.append(f"{f * TIME.get(unit, 1.0):.1f}")
            elif n.endswith("inst_executed.sum") or n.startswith("launch__"):
                cells.append(f"{int(f)}")
            else:
                cells.append(f"{f:.1f}")
        grid = r[h.index("launch__grid_size")] if "launch__grid_size" in h else "?"
        block = r[h.index("launch__block_size")] if "launch__block_size" in h else "?"
        print(f"| {r[h.index('ID')]} | `{name}` | {grid} x {block} | " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main()
