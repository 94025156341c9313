"""dev tool: key counters and a per-block instruction summary from an .ncu-rep (run where ncu is installed).
   python tools/ncu_summary.py gpurun_out/x.ncu-rep [tiles]"""
import collections, csv, subprocess, sys
rep = sys.argv[1]
tiles = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_shared_mem", "sm__cycles_elapsed.avg", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct"]
print("| metric | unit | value |\n|---|---|---|")
for w in want:
    if w in hdr:
        i = hdr.index(w)
        print(f"| {w} | {units[i]} | {vals[i]} |")
stalls = {h.split("smsp__pcsamp_warps_issue_stalled_")[1]: int(vals[i]) for i, h in enumerate(hdr)
          if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued") and vals[i].isdigit()}
print("warp stall samples:", ", ".join(f"{k} {v}" for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:9]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
h = rows[1]
iS, iI, iW = h.index("Source"), h.index("Instructions Executed"), h.index("L1 Wavefronts Shared")
byop, tot, blocks, cur = collections.Counter(), 0, [], None
for r in rows[2:]:
    try:
        n = int(r[iI])
    except (ValueError, IndexError):
        continue
    t = r[iS].split()
    op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    byop[op] += n
    tot += n
    if cur is None or cur[0] != n:
        cur = [n, collections.Counter(), 0]
        blocks.append(cur)
    cur[1][op] += 1
    cur[2] += int(r[iW] or 0)
print(f"\ninstructions per tile {tot / tiles:.0f}; by opcode:", ", ".join(f"{o} {n / tiles:.0f}" for o, n in byop.most_common(14)))
print("blocks (executions per tile x instructions = warp-instructions per tile; shared wavefronts per tile):")
for n, c, w in blocks:
    if n * sum(c.values()) / tiles > 400:
        print(f"  {n / tiles:7.1f} x {sum(c.values()):4d} = {n * sum(c.values()) / tiles:7.0f}  wf {w / tiles:6.0f}  {dict(c.most_common(5))}")
