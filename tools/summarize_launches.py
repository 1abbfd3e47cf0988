import csv, collections, re, sys
with open(sys.argv[1]) as f:
    lines=[l for l in f if not l.startswith('==')]
r=csv.DictReader(lines)
tot=collections.defaultdict(lambda:[0,0.0,0.0])
n=0
for row in r:
    name=row['Kernel Name']; v=float(row['Metric Value'].replace(',','')); unit=row['Metric Unit']
    if unit=='us': v*=1000
    elif unit=='ms': v*=1e6
    name=re.sub(r'\(.*','',name)
    name=name.replace('lsb::<unnamed>::','').replace('void ','')[:64]
    t=tot[name]; t[0]+=1; t[1]+=v; t[2]=max(t[2],v); n+=1
S=sum(t[1] for t in tot.values())
print(n,"launches, total %.1f us"%(S/1e3))
for k,t in sorted(tot.items(), key=lambda x:-x[1][1])[:int(sys.argv[2]) if len(sys.argv)>2 else 25]:
    print("%-66s n=%4d total=%9.1f us avg=%8.1f max=%8.1f share=%4.1f%%"%(k,t[0],t[1]/1e3,t[1]/1e3/t[0],t[2]/1e3,100*t[1]/S))
