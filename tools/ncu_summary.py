"""Turns an .ncu-rep (ncu --set full) into the short text summary kept under profiles/.

usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [out.txt]
Runs `ncu -i ... --page raw --csv` (no GPU needed) and keeps the metrics the roofline argument uses."""
import csv
import io
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum",
    "sm__cycles_active.avg",
    "gpc__cycles_elapsed.avg.per_second",
    "launch__registers_per_thread",
    "launch__grid_size",
    "launch__block_size",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "lts__t_sectors_srcunit_tex_op_read.sum",
    "lts__t_sectors_op_red.sum",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
]
KEEP_SUFFIX = [
    "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_alu_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fma_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fmaheavy_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_lsu_realtime.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_requests_pipe_lsu_mem_local_op_ld.sum",
    "l1tex__t_requests_pipe_lsu_mem_local_op_st.sum",
    "sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
    "sm__ops_path_tensor_op_utchmma_src_tf32_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    lines = []
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
        lines.append("kernel: %s" % name[:150])
        for i, n in enumerate(hdr):
            if n in KEEP or any(n.endswith(s) for s in KEEP_SUFFIX):
                lines.append("  %-100s %-12s %s" % (n, units[i], r[i]))
        try:
            rd = float(r[hdr.index("dram__bytes_read.sum")])
            wr = float(r[hdr.index("dram__bytes_write.sum")])
            scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
            t = rd * scale.get(units[hdr.index("dram__bytes_read.sum")], 1.0) + \
                wr * scale.get(units[hdr.index("dram__bytes_write.sum")], 1.0)
            lines.append("  traffic (dram read + write) bytes per launch: %.6e" % t)
        except Exception:
            pass
    text = "\n".join(lines) + "\n"
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(text)
    print(text)


if __name__ == "__main__":
    main()
