"""Summarise an ncu launch list (gpu__time_duration CSV) and/or a full .ncu-rep into text for profiles/."""
import csv, re, subprocess, sys, collections
def launches(path):
    lines=[l for l in open(path) if not l.startswith('==')]
    rows=[(r['Kernel Name'],float(r['Metric Value'].replace(',',''))) for r in csv.DictReader(lines)]
    ours=[(n,v) for n,v in rows if any(k in n for k in ['bgemm','contract','build_G','prep_unit','finalize','arena','assemble','delta_kernel','iso_project','iso_expand','lincomb','extract_E','rollout','small_unit','fid_','sparse_drive','aux_'])]
    idx=[i for i,(n,v) in enumerate(ours) if 'build_G' in n and True]
    # last evaluation = from the last norm pre-pass (first build_G of an eval) to the end
    pre=[i for i in idx if i+1 < len(ours) and 'build_G' in ours[i+1][0]]
    start=pre[-1] if pre else 0
    last=ours[start:]
    tot=sum(v for n,v in last)
    agg=collections.OrderedDict()
    for n,v in last:
        key=re.sub(r'\(.*','',n).replace('void ','').replace('qcb::','')
        a=agg.setdefault(key,[0,0.0]); a[0]+=1; a[1]+=v
    out=[f"last evaluation: {len(last)} launches, {tot/1e6:.3f} ms total (ncu: cold cache, serialised -> compare shares)"]
    for k,(c,v) in sorted(agg.items(), key=lambda kv:-kv[1][1]):
        out.append(f"  {k:48s} {c:4d} launches {v/1e6:9.3f} ms {100*v/tot:5.1f}%")
    return "\n".join(out)
KEYS=['gpu__time_duration.sum','launch__grid_size','launch__block_size','launch__registers_per_thread','launch__occupancy_limit_shared_mem','launch__occupancy_limit_registers','sm__warps_active.avg.pct_of_peak_sustained_active',
 'dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','lts__t_sector_hit_rate.pct','l1tex__t_sector_hit_rate.pct',
 'sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active','sm__ops_path_tensor_src_fp64.sum','sm__ops_path_tensor_src_fp64.sum.per_second','sm__ops_path_tensor_src_fp64.sum.pct_of_peak_sustained_elapsed',
 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__throughput.avg.pct_of_peak_sustained_elapsed',
 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','smsp__inst_executed.sum','smsp__warps_eligible.avg.per_cycle_active']
def full(path, maxk=3):
    raw=subprocess.run(['ncu','-i',path,'--page','raw','--csv'],capture_output=True,text=True).stdout
    rows=list(csv.reader(raw.splitlines())); hdr,units=rows[0],rows[1]; idx={k:i for i,k in enumerate(hdr)}
    out=[]
    for r in rows[2:2+maxk]:
        out.append(f"kernel: {r[idx['Kernel Name']]}")
        for k in KEYS:
            if k in idx: out.append(f"  {k:90s} {r[idx[k]]} {units[idx[k]]}")
        st=[(k,float(r[i])) for k,i in idx.items() if k.startswith('smsp__pcsamp_warps_issue_stalled_') and not k.endswith('_not_issued') and r[i] not in ('','n/a')]
        tot=sum(v for _,v in st) or 1
        out.append("  stall samples: "+", ".join(f"{k.replace('smsp__pcsamp_warps_issue_stalled_','')} {100*v/tot:.0f}%" for k,v in sorted(st,key=lambda kv:-kv[1])[:7]))
    return "\n".join(out)
if __name__=="__main__":
    for a in sys.argv[1:]:
        print(f"== {a}"); print(launches(a) if a.endswith('.csv') else full(a)); print()
