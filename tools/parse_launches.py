import csv,sys,collections
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>5]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
agg=collections.OrderedDict()
for r in rows[1:]:
    try: v=float(r[vi].replace(',',''))
    except: continue
    k=r[ki].split('(')[0][-40:]
    agg.setdefault(k,[]).append(v/1e3)
for k,v in agg.items():
    print("%-42s n=%3d last=%.1f us  min=%.1f" % (k,len(v),v[-1],min(v)))
