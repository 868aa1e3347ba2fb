"""Summarise an .ncu-rep (read here, no GPU needed): key metrics + instruction mix + hottest SASS lines."""
import csv, re, subprocess, sys, collections, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ['Kernel Name','gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'lts__t_bytes.sum', 'lts__t_sectors_srcunit_tex_op_read.sum', 'sm__cycles_active.avg', 'lts__t_sector_hit_rate.pct',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum','l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum']
for h, u, v in zip(hdr, units, vals):
    if h in want or h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio') and float(v or 0) > 0.2:
        print(f"{h} [{u}] = {v}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]; data = rows[2:]
ia, isrc, isamp = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
tot = sum(int(r[ia]) for r in data if r[ia].isdigit()); tots = sum(int(r[isamp] or 0) for r in data if r[ia].isdigit())
print("total warp instrs", tot, "samples", tots)
ops = collections.Counter(); samp = collections.Counter()
for r in data:
    if not r[ia].isdigit(): continue
    m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[isrc]); op = m.group(2).split('.')[0] if m else '?'
    ops[op] += int(r[ia]); samp[op] += int(r[isamp] or 0)
for op, c in ops.most_common(18): print(f"  {op:10s} {c:12d} {100*c/tot:5.1f}%  samples {100*samp[op]/max(tots,1):5.1f}%")
print("hottest SASS by samples:")
best = sorted([r for r in data if r[ia].isdigit()], key=lambda r: -int(r[isamp] or 0))[:int(sys.argv[2]) if len(sys.argv) > 2 else 25]
for r in best: print(f"  {int(r[isamp]):6d} {int(r[ia]):10d}  {r[isrc][:110]}")
