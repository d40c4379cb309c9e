# usage: ncu -i X.ncu-rep --page source --csv --print-source sass > src.csv; python scripts/ncu_source_buckets.py src.csv KERNEL_SUBSTRING BUCKET
# prints, per bucket of BUCKET consecutive SASS instructions of the first matching kernel: warp-stall samples, share, executed
# warp instructions, the three main stall reasons and the TMA / TMEM / barrier instructions inside (to recognise the region)
import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
want=sys.argv[2]; bs=int(sys.argv[3])
secs=[];cur=None
for r in rows:
    if r and r[0]=='Kernel Name': cur={'name':r[1],'rows':[]}; secs.append(cur); continue
    if cur is not None: cur['rows'].append(r)
for s in secs:
    if want not in s['name']: continue
    hdr=s['rows'][0]; body=[r for r in s['rows'][1:] if len(r)==len(hdr)]
    si=hdr.index('# Samples'); ie=hdr.index('Instructions Executed')
    stall_cols=[i for i,h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
    tot=sum(int(r[si]) for r in body)
    for b0 in range(0,len(body),bs):
        blk=body[b0:b0+bs]
        sm=sum(int(r[si]) for r in blk); ex=sum(int(r[ie]) for r in blk)
        agg={}
        for r in blk:
            for i in stall_cols: agg[hdr[i][6:]]=agg.get(hdr[i][6:],0)+int(r[i])
        st=sorted(agg.items(),key=lambda kv:-kv[1])[:3]
        key=[r[1].strip()[:28] for r in blk if any(t in r[1] for t in ('UTMA','LDTM','UTC','SYNCS','BAR','UBLK'))][:5]
        print(f'{b0:5d} samples={sm:5d} {100*sm/tot:5.1f}% exec={ex:9d} {st} {key}')
    break
