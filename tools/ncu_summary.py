"""Summarise an ncu report exported with `--page raw --csv` and `--page source --csv` (run in the build container)."""
import csv, sys
raw, src = sys.argv[1], sys.argv[2]
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 25
rows = list(csv.reader(open(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sectors.sum', 'lts__throughput.avg.pct',
        'l1tex__throughput.avg.pct', 'sm__throughput.avg.pct', 'sm__inst_executed.sum', 'sm__warps_active.avg.pct',
        'smsp__issue_active.avg.pct', 'launch__registers_per_thread', 'launch__occupancy_limit', 'launch__grid_size', 'launch__block_size',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_atom.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum', 'lts__t_sector_hit_rate.pct',
        'l1tex__t_sector_hit_rate.pct', 'l1tex__data_pipe_lsu_wavefronts.sum', 'l1tex__lsu_writeback', 'sm__cycles_elapsed.avg ',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed_op_shared']
for i, h in enumerate(hdr):
    if any(h.startswith(w) for w in want) and 'per_second' not in h:
        print(f"{h} [{units[i]}] = {vals[i]}")
rows = list(csv.reader(open(src)))
print(rows[0][:2])
hdr = rows[1]; idx = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[2:] if len(r) >= len(hdr)]
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot = {s: sum(int(r[idx[s]] or 0) for r in data) for s in stalls}
ns = sum(int(r[idx['# Samples']] or 0) for r in data)
ni = sum(int(r[idx['Instructions Executed']] or 0) for r in data)
print('samples', ns, 'warp instructions', ni)
print('stalls:', ', '.join(f"{s[6:]} {100 * v / ns:.1f}%" for s, v in sorted(tot.items(), key=lambda x: -x[1])[:8]))
for r in sorted(data, key=lambda r: -int(r[idx['# Samples']] or 0))[:topn]:
    st = sorted(((s[6:], int(r[idx[s]] or 0)) for s in stalls), key=lambda x: -x[1])[:2]
    print(f"{int(r[idx['# Samples']]):7d} {100 * int(r[idx['# Samples']]) / ns:5.1f}%  x{r[idx['Instructions Executed']]:>9}  {r[idx['Source']][:70]:70s} {st}")
