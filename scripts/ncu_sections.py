import csv
rows=list(csv.reader(open(__import__('sys').argv[1])))
out=[];k=0
for r in rows:
    if r and r[0]=="Kernel Name":
        k+=1
        if k>1: break
        continue
    if not r or r[0]=="Address": continue
    try: out.append((r[1].strip(), int(r[5]), int(r[2]), int(r[3])))
    except: pass
tot=sum(o[1] for o in out); ts=sum(o[2] for o in out); tn=sum(o[3] for o in out)
mx=max(o[1] for o in out if 'LDS.128' in o[0])
bars=[i for i,o in enumerate(out) if 'BAR.SYNC' in o[0] and o[1]>0.5*mx]
lds=[i for i,o in enumerate(out) if 'LDS.128' in o[0] and o[1]>0.9*mx]
print(bars,lds, 'inst',tot,'samples all',ts,'not-issued',tn)
cuts=sorted(set([0,lds[0]-14]+[b+1 for b in bars]+[len(out)]))
import os
wt=2766888890/int(os.environ.get("TILE","24576"))*int(os.environ.get("WARPS","12"))
for a,b in zip(cuts,cuts[1:]):
    i=sum(o[1] for o in out[a:b]); s=sum(o[2] for o in out[a:b]); n=sum(o[3] for o in out[a:b])
    print("%5d-%5d inst/warp-tile %6.1f (%5.1f%%)  samples %5.1f%%  not-issued %5.1f%%"%(a,b,i/wt,100*i/tot,100*s/ts,100*n/tn))
