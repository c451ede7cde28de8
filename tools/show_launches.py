import csv, sys
from collections import defaultdict
rows=[r for r in csv.reader(open(sys.argv[1] if len(sys.argv)>1 else 'gpurun_out/wf_launches.csv')) if len(r)>10]
hdr=rows[0]; ci={h:i for i,h in enumerate(hdr)}
d=defaultdict(dict)
for r in rows[1:]:
    d[(int(r[ci['ID']]),r[ci['Kernel Name']][:34])][r[ci['Metric Name']]]=r[ci['Metric Value']]
short={'gpu__time_duration.sum':'us','smsp__thread_inst_executed_per_inst_executed.ratio':'thr/inst','smsp__inst_executed.sum':'Minst','sm__warps_active.avg.pct_of_peak_sustained_active':'warps%','smsp__issue_active.avg.pct_of_peak_sustained_active':'issue%'}
tot=0
for (i,k),m in d.items():
    us=float(m['gpu__time_duration.sum'])/1e3; tot+=us
    print(f"{i:3d} {k:36s} {us:9.1f} us  " + ' '.join(f"{short[a]}={float(b)/1e6:.1f}" if short[a]=='Minst' else f"{short[a]}={float(b):.1f}" for a,b in m.items() if a!='gpu__time_duration.sum'))
print('total us', tot)
