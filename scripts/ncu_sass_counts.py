import csv,sys
rows=list(csv.reader(open(__import__('sys').argv[1])))
# first kernel only
out=[];k=0
for r in rows:
    if r and r[0]=="Kernel Name":
        k+=1
        if k>1: break
        continue
    if not r or r[0]=="Address": continue
    try: out.append((r[1].strip(), int(r[5]), int(r[2]), float(r[8] or 0)))
    except Exception as e: pass
tot=sum(o[1] for o in out); ts=sum(o[2] for o in out)
print("total inst",tot,"samples",ts, "n", len(out))
# per warp per tile: ntiles
ntiles = 2766888890/24576
wt = ntiles*12
acc=0
for i,o in enumerate(out):
    acc+=o[1]
    print("%4d %8.2f %6.2f%% s%5.2f%% a%4.1f  %s"%(i,o[1]/wt,100*o[1]/tot,100*o[2]/ts,o[3],o[0][:90]))
