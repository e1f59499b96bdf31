import csv,collections,sys
for f in sys.argv[1:]:
    with open(f) as fh:
        lines=[l for l in fh if not l.startswith('==')]
    r=csv.DictReader(lines)
    agg=collections.defaultdict(lambda:[0,0.0])
    for row in r:
        if row.get('Metric Name')!='gpu__time_duration.sum': continue
        name=row['Kernel Name']
        grid=row.get('Grid Size','')
        v=float(row['Metric Value'].replace(',',''))
        u=row['Metric Unit']
        if u=='ns': v/=1000
        elif u=='ms': v*=1000
        key=name[:80]+' grid='+grid
        agg[key][0]+=1; agg[key][1]+=v
    tot=sum(v[1] for v in agg.values())
    print(f, 'total us %.0f'%tot)
    for k,v in sorted(agg.items(), key=lambda kv:-kv[1][1])[:16]:
        print('  %-110s %5d %10.0f %5.1f%% %8.1f'%(k,v[0],v[1],100*v[1]/tot,v[1]/v[0]))
