"""Per-kernel summary of an `ncu --page raw --csv` export: duration, DRAM bytes and throughput, tensor-pipe / issue /
warp-occupancy counters (the metric names of /opt/skills/guides/B200_PROFILING.md).  Usage: ncu_summary.py file.raw.csv"""
import csv
import sys

WANT = [
    ("gpu__time_duration.sum", "dur"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
    ("sm__inst_executed_pipe_tensor.sum", "tensor_inst"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pct"),
    ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "hmma_pct"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_pct"),
    ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue_pct"),
    ("smsp__inst_executed_pipe_xu.sum", "xu_inst"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu_pct"),
    ("lts__t_bytes.sum", "l2_bytes"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem_wavefronts"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
]


def main(path):
    rows = list(csv.reader(open(path, newline="")))
    # the header is the first row that starts with "ID"; the row after it holds units
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr, units = rows[hi], rows[hi + 1]
    col = {n: i for i, n in enumerate(hdr)}
    name_i = col.get("Kernel Name")
    for r in rows[hi + 2:]:
        if len(r) < len(hdr):
            continue
        out = [r[name_i][:60]]
        for metric, short in WANT:
            i = col.get(metric)
            if i is not None and r[i] != "":
                out.append(f"{short}={r[i]}{units[i] if units[i] not in ('', '%') else ''}")
        print(" | ".join(out))


if __name__ == "__main__":
    main(sys.argv[1])
