#!/usr/bin/env python3
"""Stall samples per CUDA source line of one `ncu --set full --import-source on` capture.

    ncu -i x.ncu-rep --page source --csv --print-source cuda,sass > /tmp/srcs.csv
    python tools/ncu_top_lines.py [N] [/tmp/srcs.csv]
Columns: share of all stall samples, share of executed warp instructions, file:line, source text, top stall reasons."""
import csv, collections, sys
rows=list(csv.reader(open(sys.argv[2] if len(sys.argv)>2 else '/tmp/srcs.csv')))
col={}; fname=None
lines=[]
for r in rows:
    if len(r)==2 and r[0]=="File Path": fname=r[1].split('/')[-1]; continue
    if r and r[0]=="Line No": col={k:i for i,k in enumerate(r)}; continue
    if not col or len(r)<len(col): continue
    if r[0] not in ("","Line No"):
        try:
            s=int(r[col["# Samples"]]); ie=int(r[col["Instructions Executed"]])
        except ValueError: continue
        st={k[6:]:int(r[i]) for k,i in col.items() if k.startswith("stall_") and "Not Issued" not in k and r[i].isdigit() and int(r[i])>0}
        lines.append((s,ie,fname,int(r[0]),r[1].strip()[:80],st))
tot=sum(l[0] for l in lines); toti=sum(l[1] for l in lines)
print("total samples",tot,"instr",toti)
n=int(sys.argv[1]) if len(sys.argv)>1 else 30
for s,ie,f,ln,t,st in sorted(lines,reverse=True)[:n]:
    top=sorted(st.items(),key=lambda kv:-kv[1])[:3]
    print(f"{100*s/tot:5.1f}% {100*ie/toti:5.1f}%i {f[:14]}:{ln:<4} {t:<80} {top}")
