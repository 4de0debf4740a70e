import csv, re, collections, sys
f=sys.argv[1]
rows=[r for r in csv.reader(l for l in open(f) if l.startswith('"'))]
hdr=rows[0]; idx={h:i for i,h in enumerate(hdr)}
L=collections.OrderedDict()
for r in rows[1:]:
    if len(r)<len(hdr): continue
    id_=int(r[idx['ID']]); name=r[idx['Kernel Name']]; m=r[idx['Metric Name']]; v=float(r[idx['Metric Value']].replace(',','')); u=r[idx['Metric Unit']]
    d=L.setdefault(id_,{'name':name,'grid':r[idx['Grid Size']],'block':r[idx['Block Size']]})
    if m.startswith('gpu__time'): d['t']=v*{'ns':1e-6,'us':1e-3,'ms':1,'s':1e3}.get(u,1e-6)
    elif m.startswith('dram'):
        mul={'byte':1,'Kbyte':1e3,'Mbyte':1e6,'Gbyte':1e9}[u]
        d[m]=v*mul
    else: d[m]=v
ids=list(L)
last=[i for i in ids if 'work_key' in L[i]['name']][-1]
tot=0;tb=0
def short(n):
    n=n.replace('tk::','').replace('<unnamed>::','')
    m=re.match(r'(?:void )?(\w+)(<[^>]*>)?',n)
    return (m.group(1)+(m.group(2) or '')) if m else n[:50]
for i in ids:
    if i<last: continue
    d=L[i]
    rd=d.get('dram__bytes_read.sum',0); wr=d.get('dram__bytes_write.sum',0)
    tot+=d['t']; tb+=rd+wr
    print(f"{i:4d} {short(d['name']):48s} {d['grid']:>14s} {d['block']:>12s} {d['t']:8.3f} ms  rd {rd/1e6:9.1f} MB  wr {wr/1e6:9.1f} MB  winst {d.get('smsp__inst_executed.sum',0)/1e6:8.1f} M  thr/inst {d.get('smsp__thread_inst_executed_per_inst_executed.ratio',0):4.1f}")
print(f"sum {tot:.3f} ms serialised, DRAM {tb/1e9:.3f} GB per step ({len([i for i in ids if i>=last])} launches)")
