import csv, collections, re, sys
path = sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/launches.csv'
frac = float(sys.argv[2]) if len(sys.argv) > 2 else 2/3
lines=[l for l in open(path) if not l.startswith('==')]
allr=list(csv.DictReader(lines))
last=allr[int(len(allr)*frac):]
agg=collections.defaultdict(lambda:[0,0.0])
for x in last:
    name=re.sub(r'\(.*','',x['Kernel Name'])[:72]
    v=float(x['Metric Value'].replace(',','')); u=x['Metric Unit']
    v = v/1000 if u=='ns' else (v*1000 if u=='ms' else v)
    agg[name][0]+=1; agg[name][1]+=v
tot=sum(v[1] for v in agg.values())
for k,v in sorted(agg.items(), key=lambda kv:-kv[1][1]):
    print(f"{v[1]:9.1f} us {v[0]:4d}x {v[1]/v[0]:8.1f} us/launch {100*v[1]/tot:5.1f}%  {k}")
print("total us %.1f launches %d" % (tot, sum(v[0] for v in agg.values())))
