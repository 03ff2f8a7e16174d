import csv, re, collections, sys, subprocess
rep=sys.argv[1]
raw=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines()))
hdr,units,vals=rows[0],rows[1],rows[2]
want=['gpu__time_duration.sum','dram__bytes_read.sum ','dram__bytes_write.sum ','launch__registers_per_thread','launch__occupancy_limit','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct','smsp__inst_executed.sum','l1tex__data_bank_conflicts_pipe_lsu_mem_shared','smsp__average_warps_issue_stalled','lts__t_sector_hit_rate.pct','l1tex__throughput.avg.pct_of_peak_sustained_active','launch__shared_mem_per_block_dynamic','launch__grid_size','sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active','sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active','sm__pipe_fmaheavy_cycles_active.avg','sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active','sm__throughput.avg','smsp__thread_inst_executed_per_inst','dram__throughput.avg.pct','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ','smsp__cycles_active.avg ','sm__cycles_elapsed.avg ']
want += ['l1tex__t_sector_hit_rate.pct', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum ', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum ',
         'l1tex__t_requests_pipe_lsu_mem_global_op_st.sum ', 'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum ',
         'l1tex__t_requests_pipe_lsu_mem_global_op_ldgsts.sum ', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ldgsts.sum ']
seen = {}
for h,u,v in zip(hdr,units,vals):
    if any((w in h+' ') for w in want):
        print(f"{h:88s} {v:>16s} {u}")
        seen[h] = v
def num(k):
    try: return float(seen[k].replace(',', ''))
    except Exception: return None
# sectors per request of the global accesses (north_star: "sectors per request"); 32-bit per lane = 4, 128-bit per lane = 16
for op, label in (('ld', 'global loads'), ('st', 'global stores'), ('ldgsts', 'cp.async copies')):
    r, t = num(f'l1tex__t_requests_pipe_lsu_mem_global_op_{op}.sum'), num(f'l1tex__t_sectors_pipe_lsu_mem_global_op_{op}.sum')
    if r and t:
        print(f"derived: sectors per request, {label:16s} {t / r:8.2f}")
dr, dw, ms = num('dram__bytes_read.sum'), num('dram__bytes_write.sum'), num('gpu__time_duration.sum')
src=subprocess.run(['ncu','-i',rep,'--page','source','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
hdr=rows[1]; idx={h:i for i,h in enumerate(hdr)}
si,ei,mi,ti=idx['Source'],idx['Instructions Executed'],idx['# Samples'],idx['Avg. Threads Executed']
tot=0; ops=collections.Counter(); smp=collections.Counter(); out=[]
for k,r in enumerate(rows[2:]):
    try: n=int(r[ei]); s=int(r[mi])
    except: continue
    m=re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[si]); op=m.group(2) if m else r[si][:20]
    key='.'.join(op.split('.')[:2]) if op.startswith(('LDG','STG','LDS','STS','LDC','BRA')) else op.split('.')[0]
    ops[key]+=n; smp[key]+=s; tot+=n
    out.append(f"{k:4d} {n:10d} {s:6d} {r[ti]:>5s}  {r[si][:100]}")
open(sys.argv[2],'w').write("\n".join(out))
print("total warp-instr",tot)
for op,n in ops.most_common(28): print(f"  {op:18s} {n:11d} {100*n/tot:5.1f}%  samples {smp[op]}")
