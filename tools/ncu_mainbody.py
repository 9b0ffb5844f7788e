import sys, csv, subprocess, collections, statistics
sys.path.insert(0,'/root/repo/tools')
import ncu_lines as nl
rep, pat = sys.argv[1], sys.argv[2]
lines = nl.sass_lines(pat)
raw = (open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout)
rows=list(csv.reader(raw.splitlines()))
hdr=next(r for r in rows if r and r[0]=="Address")
body=[r for r in rows if len(r)==len(hdr) and r[0].startswith("0x")]
col={h:i for i,h in enumerate(hdr)}
base=int(body[0][0],16)
main=[r for r in body if int(r[col["Instructions Executed"]])>0 and int(r[col["Thread Instructions Executed"]])/int(r[col["Instructions Executed"]])>28]
norm=statistics.median(int(r[col["Instructions Executed"]]) for r in main)
ops=collections.Counter(); byline=collections.Counter()
for r in main:
    off=int(r[0],16)-base
    loc,ins=lines.get(off,(("?",0),r[1]))
    op=ins.split()[1] if ins.startswith("@") else ins.split()[0]
    ie=int(r[col["Instructions Executed"]])
    ops[op.split(".")[0]]+=ie/norm
    byline[loc]+=ie/norm
print("main body per iteration: %.0f"%sum(ops.values()))
for k,v in ops.most_common(30): print("%-8s %6.1f"%(k,v))
print()
for k,v in byline.most_common(45): print("%-24s:%-4d %6.1f"%(k[0][-24:],k[1],v))
