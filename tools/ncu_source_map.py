import re,csv,collections,sys
dis,fn_pat,srccsv,srcfile=sys.argv[1:5]
cur=None; ins=[]; on=False
for l in open(dis):
    if l.startswith('\t.section') or l.startswith('//----'):
        on = ('.text.' in l and fn_pat in l)
    if not on: continue
    m=re.search(r'//## File "([^"]+)", line (\d+)(.*)',l)
    if m:
        cur=(m.group(1).split('/')[-1],int(m.group(2)),m.group(3)); continue
    m=re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);',l)
    if m: ins.append((cur,m.group(2)))
rows=list(csv.reader(open(srccsv)))
h=rows[1]; ie=h.index("Instructions Executed")
sass=[(r[1].strip(), int(r[ie])) for r in rows[2:] if len(r)>ie and r[ie].isdigit()]
print("disasm instr",len(ins),"profile instr",len(sass))
n=min(len(ins),len(sass))
agg=collections.Counter()
for (c,op),(s,cnt) in zip(ins[:n],sass[:n]):
    # use outermost location in our file if inlined info present
    key=(c[0],c[1]) if c else ('?',0)
    agg[key]+=cnt
tot=sum(agg.values()); scale=float(sys.argv[5]) if len(sys.argv)>5 else 1e6
print("total",tot,"per target",tot/scale)
src=open(srcfile).read().split('\n')
base=srcfile.split('/')[-1]
for (f,ln),c in sorted(agg.items(), key=lambda x:-x[1])[:60]:
    text = src[ln-1].strip()[:100] if f==base and ln-1<len(src) else ''
    print("%8.1f  %s:%d  %s"%(c/scale,f,ln,text))
