"""Text summary of one `ncu --set full` capture: ncu_summary.py <file.ncu-rep> > profiles/<name>.txt"""
import csv
import io
import re
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
KEEP = [r"^Kernel Name$", r"gpu__time_duration\.sum$", r"^launch__grid_size$", r"^launch__block_size$", r"launch__registers_per_thread$",
        r"launch__shared_mem_per_block_dynamic$", r"^dram__bytes_read\.sum$", r"^dram__bytes_write\.sum$", r"dram__throughput\.avg\.pct_of_peak_sustained_elapsed",
        r"sm__pipe_tensor_cycles_active\.avg\.pct_of_peak_sustained_(active|elapsed)", r"sm__cycles_elapsed\.avg\.per_second",
        r"smsp__issue_active\.avg\.pct_of_peak_sustained_active", r"sm__warps_active\.avg\.pct_of_peak_sustained_active",
        r"sm__inst_executed_pipe_(alu|fma|lsu|xu|tc|tma|tmem)\.avg\.pct_of_peak_sustained_active",
        r"l1tex__data_pipe_lsu_wavefronts\.avg\.pct_of_peak_sustained_elapsed", r"l1tex__data_pipe_tc_wavefronts_mem_shared\.sum\.pct",
        r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum$", r"lts__t_sectors_srcunit_tex_op_read\.sum$",
        r"smsp__average_warps_issue_stalled_.*_per_issue_active\.ratio", r"l1tex__m_xbar2l1tex_read_bytes_mem_dshared\.sum$"]
for vals in rows[2:]:
    print("=" * 100)
    for h, u, v in zip(hdr, units, vals):
        if any(re.search(k, h) for k in KEEP) and v not in ("", "0", "0.000000"):
            print(f"{h:95s} {v} {u}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
if len(rows) > 3:
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[2:] if len(r) == len(hdr)]
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = {h: sum(int(r[ix[h]] or 0) for r in data) for h in stalls}
    tot = sum(int(r[ix["# Samples"]] or 0) for r in data)
    print("=" * 100)
    print(f"warp-state samples: {tot}")
    for h, v in sorted(agg.items(), key=lambda x: -x[1])[:10]:
        print(f"  {h:28s} {v:9d} {100.0 * v / max(tot, 1):5.1f}%")
    print("hottest instructions (samples, executed, SASS, top stall):")
    for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]] or 0))[:25]:
        st = sorted(((h, int(r[ix[h]] or 0)) for h in stalls), key=lambda x: -x[1])[0]
        print(f"  {int(r[ix['# Samples']] or 0):7d} {r[ix['Instructions Executed']]:>10s}  {r[ix['Source']][:70]:70s} {st[0]}={st[1]}")
