"""Headline metrics of every kernel in an .ncu-rep (one `ncu --set full` capture) as a small JSON for profiles/.
usage: python tools/ncu_summary.py <file.ncu-rep> [<file.ncu-rep> ...] > profiles/<name>.json"""
import csv
import json
import subprocess
import sys

WANT = {
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "lts__t_bytes.sum": "l2_bytes",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed": "mem_throughput_pct",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
    "sm__inst_executed.sum": "warp_instructions",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
    "smsp__thread_inst_executed_per_inst_executed.ratio": "threads_per_instruction",
    "launch__registers_per_thread": "registers",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "launch__occupancy_limit_registers": "occ_limit_regs",
    "launch__occupancy_limit_shared_mem": "occ_limit_smem",
    "sm__inst_executed_pipe_tensor.sum": "tensor_pipe_inst",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_active_pct",
    "l1tex__t_sector_hit_rate.pct": "l1_hit_pct",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct",
}


def summarize(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    head, units = rows[0], rows[1]
    out = []
    for row in rows[2:]:
        rec = {"kernel": row[head.index("Kernel Name")][:120]}
        for i, h in enumerate(head):
            if h in WANT and row[i] != "":
                try:
                    v = float(row[i].replace(",", ""))
                except ValueError:
                    continue
                rec[WANT[h]] = v
                if units[i]:
                    rec[WANT[h] + "_unit"] = units[i]
        # top stall reasons (warp-state samples)
        stalls = {}
        for i, h in enumerate(head):
            if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued") and row[i]:
                try:
                    stalls[h.replace("smsp__pcsamp_warps_issue_stalled_", "")] = float(row[i].replace(",", ""))
                except ValueError:
                    pass
        tot = sum(stalls.values()) or 1.0
        rec["stall_share"] = {k: round(v / tot, 3) for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:5]}
        out.append(rec)
    return out


if __name__ == "__main__":
    print(json.dumps({rep.split("/")[-1]: summarize(rep) for rep in sys.argv[1:]}, indent=1))
