import csv, collections, sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
hdr=rows[0]; i_name=hdr.index('Kernel Name'); i_m=hdr.index('Metric Name'); i_v=hdr.index('Metric Value'); i_id=hdr.index('ID')
d=collections.defaultdict(dict)
for r in rows[1:]:
    d[(r[i_id], r[i_name])][r[i_m]]=float(r[i_v].replace(',',''))
agg=collections.defaultdict(lambda:[0,0,0,0])
for (id_,name),m in d.items():
    a=agg[name.replace('void ','').replace('cmbpo::','')[:58]]; a[0]+=1; a[1]+=m.get('gpu__time_duration.sum',0); a[2]+=m.get('dram__bytes_read.sum',0); a[3]+=m.get('dram__bytes_write.sum',0)
tot=sum(a[1] for a in agg.values())
print(f"# {sys.argv[1]}: {len(d)} launches, {tot/1e6:.2f} ms total (ncu: cold-cache, serialised -- shares only)")
for k,a in sorted(agg.items(), key=lambda kv:-kv[1][1]):
    print(f"{k:58s} n={a[0]:3d} avg {a[1]/a[0]/1e3:9.1f} us share {100*a[1]/tot:5.1f}%  rd {a[2]/a[0]/1e6:8.1f} MB wr {a[3]/a[0]/1e6:8.1f} MB")
