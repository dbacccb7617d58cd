"""Summarise an `ncu -i X.ncu-rep --page source --csv --kernel-name regex:K` export: stall reasons and the hottest instructions.
usage: python tools/ncu_stalls.py export.csv [...]"""
import csv,sys
def summarize(path, ntop=14):
    r=list(csv.reader(open(path)))
    # may contain several kernels; take first block
    h=r[1]; rows=[]
    for x in r[2:]:
        if len(x)!=len(h) or x[0]=="Address": break
        rows.append(x)
    ix={c:i for i,c in enumerate(h)}
    stalls=[c for c in h if c.startswith('stall_') and 'Not Issued' not in c]
    tot={c:0 for c in stalls}; S=0
    for x in rows:
        n=int(x[ix['# Samples']] or 0); S+=n
        for c in stalls: tot[c]+=int(x[ix[c]] or 0)
    out=[f"samples {S}, instructions {len(rows)}"]
    out.append("stall reasons: "+", ".join(f"{c[6:]} {100*v/S:.1f}%" for c,v in sorted(tot.items(), key=lambda kv:-kv[1])[:7]))
    top=sorted(rows,key=lambda x:-int(x[ix['# Samples']] or 0))[:ntop]
    for x in top:
        st={c:int(x[ix[c]] or 0) for c in stalls}
        dom=sorted(st.items(), key=lambda kv:-kv[1])[:2]
        out.append(f"  {x[ix['Address']][-4:]}  {x[ix['Source']].strip()[:52]:52s} {100*int(x[ix['# Samples']])/S:5.1f}%  "+", ".join(f"{c[6:]} {v}" for c,v in dom))
    return "\n".join(out)
for p in sys.argv[1:]:
    print("==",p); print(summarize(p))
