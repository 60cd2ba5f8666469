"""ncu report -> the small JSON summaries kept under profiles/.

  ncu -i X.ncu-rep --page raw --csv > X_raw.csv ; ncu -i X.ncu-rep --page source --csv > X_src.csv
  python profiles/ncu_summary.py X_raw.csv [X_src.csv] [arena_bytes] > profiles/rN_ncu_<kernel>_summary.json

Keeps: duration, DRAM bytes, L1/L2 hit rates, LSU wavefronts and bank conflicts, issue activity, warps, registers, shared
memory, the per-issue stall reasons (warp cycles per issued instruction), and - with the source page - the instruction
mix per 32 arena bytes and the SASS mnemonics that show how the bytes arrive (UTMALDG = TMA tensor load)."""
import collections
import csv
import json
import re
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "lts__t_sectors_srcunit_tex_op_read.sum",
        "launch__block_size", "launch__grid_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "sm__cycles_elapsed.max", "sm__cycles_active.avg", "sm__cycles_active.min", "sm__cycles_active.max", "smsp__cycles_active.avg",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.per_cycle_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.per_cycle_active", "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"]


def main():
    raw = list(csv.reader(open(sys.argv[1])))
    hdr, units, vals = raw[0], raw[1], raw[-1]
    out = {}
    stalls = {}
    for h, u, v in zip(hdr, units, vals):
        if h == "Kernel Name":
            out[h] = v
        elif h in KEEP and h not in out:
            out[h] = [v, u]
        m = re.match(r"smsp__average_warps_issue_stalled_(\w+)_per_issue_active\.ratio", h)
        if m and m.group(1) not in stalls:
            try:
                if float(v) >= 0.05:
                    stalls[m.group(1)] = round(float(v), 3)
            except ValueError:
                pass
    out["warp_cycles_per_issued_instruction_by_stall_reason"] = dict(sorted(stalls.items(), key=lambda kv: -kv[1]))
    if len(sys.argv) > 2 and sys.argv[2] != "-":
        src = list(csv.reader(open(sys.argv[2])))
        h = src[1]
        ia, isrc = h.index("Instructions Executed"), h.index("Source")
        mix = collections.Counter()
        present = collections.Counter()
        total = 0
        for r in src[2:]:
            if len(r) <= ia:
                continue
            ins = r[isrc].strip()
            op = ins.split()[1] if ins.startswith("@") and len(ins.split()) > 1 else (ins.split() or ["?"])[0]
            n = int(r[ia] or 0)
            mix[op.split(".")[0]] += n
            total += n
            for key in ("UTMALDG", "UBLKCP", "SYNCS", "LDS.U8", "LDS.U16", "LDS.128", "IDP.4A", "ATOMS", "ATOMG", "REDUX", "VIMNMX3", "POPC", "LDG"):
                if op.startswith(key):
                    present[key] += 1
        arena = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
        if arena:
            per = arena / 32.0
            out["warp_instructions_per_32_arena_bytes"] = {k: round(v / per, 3) for k, v in mix.most_common(14)} | {"total": round(total / per, 3)}
        out["sass_mnemonics_present"] = dict(present)
    json.dump(out, sys.stdout, indent=1)
    print()


main()
