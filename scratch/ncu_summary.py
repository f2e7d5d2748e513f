"""Summarise an ncu report exported with --page raw --csv and --page source --csv: key metrics, shared-memory wavefronts by opcode, stall reasons."""
import collections, csv, sys
raw, src = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'smsp__inst_executed.sum', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__block_size', 'launch__shared_mem_per_block_dynamic', 'smsp__inst_executed_op_shared_atom.sum']
for w in want:
    for i, h in enumerate(hdr):
        if h == w:
            print(f"{w:86s} {units[i]:10s} {vals[i]}")
rows = list(csv.reader(open(src)))
hdr = rows[1]; data = rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
wf, cnt, samp = collections.Counter(), collections.Counter(), collections.Counter()
for r in data:
    t = r[ix['Source']].split()
    op = t[1] if t[0].startswith('@') else t[0]
    wf[op] += int(r[ix['L1 Wavefronts Shared']] or 0); cnt[op] += int(r[ix['Instructions Executed']] or 0); samp[op] += int(r[ix['# Samples']] or 0)
tot = sum(samp.values())
print("shared-memory wavefronts by opcode:")
for op, w in wf.most_common(8):
    if w: print(f"  {op:30s} insts {cnt[op]:>14d} wavefronts {w:>14d}  wf/inst {w / max(cnt[op], 1):.2f}")
print("warp-stall samples by opcode:")
for op, s in samp.most_common(10): print(f"  {op:30s} {100 * s / tot:.1f}%")
print("stall reasons:", {k[6:]: sum(int(r[ix[k]] or 0) for r in data) for k in hdr if k.startswith('stall_') and '(' not in k})
print("total instructions", sum(cnt.values()))
