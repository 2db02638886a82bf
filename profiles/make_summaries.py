# builds profiles/r01_*.md|csv from gpurun_out (ncu reports) -- run on the CPU box
import csv, subprocess, sys, collections, json, os
os.chdir('/root/repo')
# 1. launch list
rows=list(csv.reader(l for l in open('gpurun_out/launches.csv') if not l.startswith('==')))
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value'); ii=hdr.index('ID')
out=[['id','kernel','gpu__time_duration_us']]
agg=collections.defaultdict(list)
for r in rows[1:]:
    if len(r)>vi:
        try: ns=float(r[vi].replace(',',''))
        except: continue
        out.append([r[ii], r[ki], '%.1f'%(ns/1e3)]); agg[r[ki]].append(ns/1e3)
with open('profiles/r01_launches.csv','w') as f:
    csv.writer(f).writerows(out)
tot=sum(sum(v) for v in agg.values())
lines=['# Round 1 -- ncu launch list (`profiles/prof.sh`, bench.py --batch 8 --steps 2 --warmup 3 --no-cpu --no-extras)','',
 'Per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes.','',
 '| kernel | launches | total us | share |','|---|---|---|---|']
for k,v in sorted(agg.items(), key=lambda kv:-sum(kv[1])):
    lines.append('| `%s` | %d | %.1f | %.1f %% |' % (k[:90], len(v), sum(v), 100*sum(v)/tot))
open('profiles/r01_launches.md','w').write('\n'.join(lines)+'\n')
# 2. per-kernel summaries
want=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
'launch__registers_per_thread','launch__grid_size','launch__block_size','launch__occupancy_limit_registers','sm__warps_active.avg.per_cycle_active',
'smsp__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__inst_issued.avg.per_cycle_active',
'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active','sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed','sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active','sm__throughput.avg.pct_of_peak_sustained_elapsed',
'smsp__average_warp_latency_per_inst_issued.ratio','sm__cycles_elapsed.max']
for name in ('fwd','bwd'):
    rep='gpurun_out/prof_%s.ncu-rep'%name
    o=subprocess.run(['ncu','-i',rep,'--page','raw','--csv'],capture_output=True,text=True).stdout
    rr=list(csv.reader(o.splitlines())); h=rr[0]; u=rr[1]; v=rr[2]
    d=dict(zip(h,zip(u,v)))
    kname=d['Kernel Name'][1]
    L=['# Round 1 -- ncu --set full, `%s`'%kname,'', 'bench.py --batch 8 (66.36 Mpx per launch), --clock-control none. Raw report: gpurun_out/prof_%s.ncu-rep (scratch).'%name,'','| metric | value | unit |','|---|---|---|']
    for w in want:
        if w in d: L.append('| %s | %s | %s |'%(w,d[w][1],d[w][0]))
    st=[(k,float(x[1].replace(',',''))) for k,x in d.items() if 'pcsamp_warps_issue_stalled' in k and 'not_issued' not in k]
    tot=sum(x for _,x in st)
    L+=['','Warp-state samples (smsp__pcsamp_warps_issue_stalled_*):','','| state | samples | share |','|---|---|---|']
    for k,x in sorted(st,key=lambda t:-t[1])[:10]:
        L.append('| %s | %d | %.1f %% |'%(k.replace('smsp__pcsamp_warps_issue_stalled_',''),x,100*x/tot))
    px=66.355200e6
    inst=float(d['smsp__inst_executed.sum'][1].replace(',',''))
    rd=d['dram__bytes_read.sum']; wr=d['dram__bytes_write.sum']
    def tobytes(t):
        val=float(t[1].replace(',','')); unit=t[0]
        return val*{'byte':1,'Kbyte':1e3,'Mbyte':1e6,'Gbyte':1e9}[unit]
    traffic=tobytes(rd)+tobytes(wr)
    L+=['','Derived: %.1f warp-instructions x32 / pixel = **%.0f thread-instructions per pixel**; DRAM traffic %.3f GB per launch = **%.1f B/px** (algorithmic: %d B/px).'%(inst*32/px/32*1, inst*32/px, traffic/1e9, traffic/px, 25 if name=='fwd' else 37)]
    open('profiles/r01_ncu_%s.md'%name,'w').write('\n'.join(L)+'\n')
    print(name, 'traffic per px', traffic/px, 'instr/px', inst*32/px)
