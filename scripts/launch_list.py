#!/usr/bin/env python
"""Per-kernel summary of one step from an ncu launch list (csv written by scripts/gpu_launches.sh):
duration and DRAM bytes of every launch from the last k_fscan on. Usage: python scripts/launch_list.py <launches.csv> [first kernel]"""
import csv,re,collections,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
hdr=rows[0]; idx={h:i for i,h in enumerate(hdr)}
L=collections.OrderedDict()
for r in rows[1:]:
    L.setdefault(r[idx['ID']],{'name':r[idx['Kernel Name']],'grid':r[idx['Grid Size']],'block':r[idx['Block Size']]})[r[idx['Metric Name']]]=float(r[idx['Metric Value']].replace(',',''))
ids=list(L); names=[L[i]['name'] for i in ids]
key=sys.argv[2] if len(sys.argv)>2 else 'k_fscan'
starts=[k for k,n in enumerate(names) if key in n]
s0=starts[-1]; tot=rd=wr=0
for i in ids[s0:]:
    d=L[i]; nm=re.sub(r'\(.*','',d['name']); nm=re.sub(r'.*::','',nm)[:48]
    t=d.get('gpu__time_duration.sum',0)/1000.0; r_=d.get('dram__bytes_read.sum',0)/1e6; w_=d.get('dram__bytes_write.sum',0)/1e6
    print("%-50s %8.1f us rd %7.2f wr %7.2f MB grid %s"%(nm,t,r_,w_,d['grid']))
    tot+=t; rd+=r_; wr+=w_
print("sum %.1f us, dram read %.1f MB write %.1f MB, launches %d"%(tot,rd,wr,len(ids)-s0))
