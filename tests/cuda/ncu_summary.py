"""Summarise an `ncu --set full` capture for profiles/: one block of headline counters per kernel launch,
the instruction mix (share of executed warp instructions per opcode) and, with --traffic, the per-launch DRAM
bytes bench.py reports as roofline.traffic.

    ncu -i gpurun_out/r01_stack.ncu-rep --page raw --csv > /tmp/raw.csv
    ncu -i gpurun_out/r01_stack.ncu-rep --page source --csv --print-source sass > /tmp/src.csv
    python tests/cuda/ncu_summary.py /tmp/raw.csv /tmp/src.csv [--traffic profiles/traffic.json]
"""
import collections
import csv
import json
import re
import sys

ROWS = [
    ("duration", "gpu__time_duration.sum"),
    ("dram read", "dram__bytes_read.sum"),
    ("dram write", "dram__bytes_write.sum"),
    ("tensor pipe active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
    ("issue slots busy", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    ("warp instructions issued", "smsp__inst_issued.sum"),
    ("SM throughput", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("L2 throughput", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("L2->SM bytes (bulk copies)", "l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum"),
    ("LSU shared wavefronts", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"),
    ("registers/thread", "launch__registers_per_thread"),
    ("CTAs/SM (smem limit)", "launch__occupancy_limit_shared_mem"),
    ("stall long_scoreboard / issue", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
    ("stall wait / issue", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"),
    ("stall short_scoreboard / issue", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio"),
    ("stall sleeping / issue", "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio"),
    ("stall barrier / issue", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"),
]
# kernel name fragment -> bench.py launch label (for traffic.json)
LABELS = [("tc_stack_kernel<192", "bf16.resstack.fused@4"), ("tc_stack_kernel<96", "bf16.resstack.fused+convT@7"),
          ("tc_stack_kernel<48", "bf16.resstack.fused+convT@10"), ("tc_stack_kernel<32", "bf16.resstack.fused+head@13")]


def to_bytes(value, unit):
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
    return int(float(value.replace(",", "")) * scale)


def main():
    raw, src = sys.argv[1], sys.argv[2]
    traffic_path = sys.argv[sys.argv.index("--traffic") + 1] if "--traffic" in sys.argv else None
    rows = list(csv.reader(open(raw)))
    H, U = rows[0], rows[1]
    col = {h: i for i, h in enumerate(H)}
    traffic = {}
    mixes = {}
    cur = None
    for r in csv.reader(open(src)):
        if r and r[0] == "Kernel Name":
            cur = mixes.setdefault(r[1], {"hdr": None, "ops": collections.Counter()})
            cur["fresh"] = not cur["ops"]
        elif r and r[0] == "Address":
            cur["hdr"] = r
        elif cur is not None and r and cur["fresh"]:
            h = cur["hdr"]
            m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[h.index("Source")])
            op = m.group(2) if m else "?"
            op = ".".join(op.split(".")[:2]) if op.startswith("SYNCS") else op.split(".")[0]
            cur["ops"][op] += int(r[h.index("Instructions Executed")])
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        label = next((lab for frag, lab in LABELS if frag in name), "")
        print(f"== {name}  [{label}]  grid {r[col['Grid Size']]} block {r[col['Block Size']]}")
        for title, metric in ROWS:
            if metric in col:
                print(f"   {title:<36}{r[col[metric]]:>16} {U[col[metric]]}")
        if label:
            traffic["bf16:" + label] = (to_bytes(r[col["dram__bytes_read.sum"]], U[col["dram__bytes_read.sum"]]) +
                                        to_bytes(r[col["dram__bytes_write.sum"]], U[col["dram__bytes_write.sum"]]))
        for full, mix in mixes.items():
            if name.replace("unnamed>", "<unnamed>").split("(")[0].replace("void ", "") in full.replace("(int)", ""):
                tot = sum(mix["ops"].values())
                top = ", ".join(f"{o} {100 * c / tot:.1f}%" for o, c in mix["ops"].most_common(14))
                print(f"   executed warp instructions by opcode: {top}")
                break
        print()
    if traffic_path:
        json.dump(traffic, open(traffic_path, "w"), indent=1)


if __name__ == "__main__":
    main()
