import sys, csv, subprocess, collections
sys.path.insert(0,'/root/repo/tools')
import ncu_lines as nl
rep, pat = sys.argv[1], sys.argv[2]
lines = nl.sass_lines(pat)
raw = subprocess.run(["ncu","-i",rep,"--page","source","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines()))
hdr=next(r for r in rows if r and r[0]=="Address")
body=[r for r in rows if len(r)==len(hdr) and r[0].startswith("0x")]
col={h:i for i,h in enumerate(hdr)}
base=int(body[0][0],16)
tot=0; cls=collections.Counter(); n_static=collections.Counter()
import statistics
maxie=statistics.median([int(r[col["Instructions Executed"]]) for r in body if int(r[col["Instructions Executed"]])>0 and int(r[col["Thread Instructions Executed"]])/int(r[col["Instructions Executed"]])>28])
for r in body:
    ie=int(r[col["Instructions Executed"]]); te=int(r[col["Thread Instructions Executed"]])
    tot+=ie
    act=te/ie if ie else 0
    k = 'main(>28)' if act>28 else ('mid(12-28)' if act>12 else 'rare(<12)')
    cls[k]+=ie; n_static[k]+=1
print("static instrs", len(body), "max exec per instr", maxie, "=> warp-iterations")
for k in cls: print(k, "%.1f%%"%(100*cls[k]/tot), "dyn per warp-iter %.0f"%(cls[k]/maxie), "static", n_static[k])
print("total per warp-iteration %.0f"%(tot/maxie))
