"""Per-kernel figures of an `ncu --set full` report as markdown (same fields as profiles/r1_ncu_summary.md).
Usage: python tools/summarize_ncu_rep.py <report.ncu-rep> <out.md> "<title>" """
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from summarize_profiles import full_metrics, short  # noqa: E402


def main():
    rep, out, title = sys.argv[1:4]
    lines = [f"# {title}", "", f"`ncu --set full --clock-control none --import-source on` (`gpurun_out/{os.path.basename(rep)}`), one launch per kernel.", ""]
    for m in full_metrics(rep):
        g = m["get"]
        rd, wr = g("dram__bytes_read.sum"), g("dram__bytes_write.sum")
        lines += [f"### `{short(m['name'])}`", "",
                  f"* time: {g('gpu__time_duration.sum')[0]} {g('gpu__time_duration.sum')[1]}; grid {m['grid']} x block {m['block']}; "
                  f"{g('launch__registers_per_thread')[0]} registers/thread",
                  f"* DRAM read {rd[0]} {rd[1]}, write {wr[0]} {wr[1]}; DRAM throughput {g('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed')[0]} % of peak",
                  f"* issue slots busy (active cycles): {g('smsp__issue_active.avg.pct_of_peak_sustained_active')[0]} %; achieved occupancy "
                  f"{g('sm__warps_active.avg.pct_of_peak_sustained_active')[0]} %; executed warp instructions {g('smsp__inst_executed.sum')[0]}",
                  f"* warp latency per instruction: {g('smsp__average_warp_latency_per_inst_issued.ratio')[0]} cycles; shared-memory bank conflicts "
                  f"{g('l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum')[0]}",
                  "* top stalls (warps per issue): " + ", ".join(f"{n} {v:.2f}" for v, n in m["stalls"]), ""]
    open(out, "w").write("\n".join(lines))
    print(f"wrote {out}")


if __name__ == "__main__":
    main()
