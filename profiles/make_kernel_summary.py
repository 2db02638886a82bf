# builds one profiles/<round>_ncu_<name>.md (round from $ROUND, default r02) from a gpurun_out/*.ncu-rep -- run on the CPU box:
#   python profiles/make_kernel_summary.py <report> <name> <pixels per launch> <algorithmic B/px> "<command>"
import csv, subprocess, sys, os
rep, name, px, alg, cmd = sys.argv[1], sys.argv[2], float(sys.argv[3]), sys.argv[4], sys.argv[5]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic',
        'sm__warps_active.avg.per_cycle_active', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_issued.avg.per_cycle_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__average_warp_latency_per_inst_issued.ratio', 'sm__cycles_elapsed.max']
o = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rr = list(csv.reader(o.splitlines())); h, u, v = rr[0], rr[1], rr[2]
d = dict(zip(h, zip(u, v)))
RND = os.environ.get('ROUND', 'r02')
L = ['# %s -- ncu --set full, `%s`' % (RND, d['Kernel Name'][1]), '',
     '`%s` (%.2f Mpx per launch), --clock-control none. Raw report: %s (scratch).' % (cmd, px / 1e6, rep), '',
     '| metric | value | unit |', '|---|---|---|']
for w in want:
    if w in d:
        L.append('| %s | %s | %s |' % (w, d[w][1], d[w][0]))
st = [(k, float(x[1].replace(',', ''))) for k, x in d.items() if 'pcsamp_warps_issue_stalled' in k and 'not_issued' not in k]
tot = sum(x for _, x in st)
L += ['', 'Warp-state samples (smsp__pcsamp_warps_issue_stalled_*):', '', '| state | samples | share |', '|---|---|---|']
for k, x in sorted(st, key=lambda t: -t[1])[:10]:
    L.append('| %s | %d | %.1f %% |' % (k.replace('smsp__pcsamp_warps_issue_stalled_', ''), x, 100 * x / tot))
def tobytes(t):
    return float(t[1].replace(',', '')) * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[t[0]]
inst = float(d['smsp__inst_executed.sum'][1].replace(',', ''))
traffic = tobytes(d['dram__bytes_read.sum']) + tobytes(d['dram__bytes_write.sum'])
line = 'Derived: **%.0f thread-instructions per pixel**; DRAM traffic %.3f GB per launch = **%.1f B/px** (algorithmic: %s B/px)' % (
    inst * 32 / px, traffic / 1e9, traffic / px, alg)
k = 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum'
if k in d:
    line += '; %.1f shared-memory wavefronts per pixel' % (float(d[k][1].replace(',', '')) / px)
L += ['', line + '.']
open(os.path.join(os.path.dirname(os.path.abspath(__file__)), '%s_ncu_%s.md' % (RND, name)), 'w').write('\n'.join(L) + '\n')
print('\n'.join(L))
