import csv, collections, sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=None; seq=[]
for r in rows:
    if "Kernel Name" in r: hdr=r; continue
    if hdr is None or len(r)<len(hdr): continue
    d=dict(zip(hdr,r))
    if d.get("Metric Name")!="gpu__time_duration.sum": continue
    v=float(d["Metric Value"].replace(",","")); u=d["Metric Unit"]
    v = v/1000 if u=="ns" else (v*1000 if u=="ms" else v)
    seq.append((d["Kernel Name"][:70], v))
starts=[i for i,(k,_) in enumerate(seq) if "stratified_z" in k]
fr=seq[starts[-1]:]
agg=collections.OrderedDict()
for k,v in fr:
    agg.setdefault(k,[0,0.0]); agg[k][0]+=1; agg[k][1]+=v
for k,(n,t) in sorted(agg.items(), key=lambda kv:-kv[1][1])[:9]: print(f"{t:9.1f} us {n:3d} {k}")
print("sum us", sum(v for _,v in fr), "launches", len(fr))
