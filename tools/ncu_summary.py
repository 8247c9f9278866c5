"""Condenses an `ncu --set full --csv --page raw` capture into the per-launch table kept under profiles/."""
import csv
import sys

COLS = [("gpu__time_duration.sum", "dur_us", 1e-3), ("dram__bytes_read.sum", "dram_rd_MB", 1e-6), ("dram__bytes_write.sum", "dram_wr_MB", 1e-6),
        ("lts__t_bytes.sum", "l2_MB", 1e-6), ("smsp__inst_executed.sum", "warp_inst_M", 1e-6),
        ("smsp__inst_executed_op_shared_atom.sum", "atoms_inst_M", 1e-6),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum", "atom_wavefronts_M", 1e-6),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_atom.sum", "atom_bank_conflicts_M", 1e-6),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct", 1), ("l1tex__throughput.avg.pct_of_peak_sustained_active", "l1tex_pct", 1),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2_pct", 1), ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct", 1),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_pct", 1), ("launch__registers_per_thread", "regs", 1),
        ("launch__grid_size", "grid", 1), ("launch__block_size", "block", 1), ("launch__shared_mem_per_block_dynamic", "dyn_smem_B", 1)]


def main(path, note=""):
    rows = list(csv.reader(open(path, errors="replace")))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    hdr = rows[hi]
    idx = {h: i for i, h in enumerate(hdr)}
    units = rows[hi + 1]
    # ncu picks the unit per column (ns/us/ms, byte/Kbyte/Mbyte/Gbyte): bring durations to ns and sizes to bytes first
    UNIT = {"ns": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6, "s": 1e9, "second": 1e9, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    if note:
        print(note)
    print("kernel | " + " | ".join(n for _, n, _ in COLS))
    for r in rows[hi + 2:]:
        if len(r) < len(hdr):
            continue
        vals = []
        for key, _, sc in COLS:
            v = r[idx[key]] if key in idx else ""
            try:
                v = float(v.replace(",", "")) * sc * (UNIT.get(units[idx[key]], 1.0) if key in idx and idx[key] < len(units) else 1.0)
                v = f"{v:.2f}" if sc != 1 or v != int(v) else f"{int(v)}"
            except ValueError:
                pass
            vals.append(v)
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "")
        print(name + " | " + " | ".join(vals))


if __name__ == "__main__":
    main(sys.argv[1], " ".join(sys.argv[2:]))
