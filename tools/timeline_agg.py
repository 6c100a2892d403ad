import re,collections,sys
rows=[]
for l in open(sys.argv[1]):
    m=re.match(r'\s*([\d.]+) us\s+([\d.]+) us\s+s(\d+)\s+(.*)',l)
    rows.append((float(m[1]),float(m[2]),int(m[3]),m[4]))
t0=[r for r in rows if r[0]>100][0][0]
main=[r for r in rows if r[2]==7 and r[0]>=t0]
d=collections.defaultdict(lambda:[0,0.0])
for r in main:
    k=re.sub(r'^void ','',r[3]); k=re.sub(r'\(anonymous namespace\)::','',k); k=k[:int(sys.argv[2]) if len(sys.argv)>2 else 70]
    d[k][0]+=1; d[k][1]+=r[1]
print("graph start",t0,"end",max(r[0]+r[1] for r in rows),"main busy",sum(r[1] for r in main), "n",len(main))
for k,v in sorted(d.items(),key=lambda kv:-kv[1][1])[:45]: print(f"{v[1]:9.1f} us x{v[0]:4d} {k}")
