#!/usr/bin/env python
"""Instructions with the most warp-stall samples in an `ncu --set full --import-source on` report (first launch):
`ncu_top_stalls.py report.ncu-rep [N]` prints index, samples, executions and the SASS text, in program order.  This view
found the per-k-block divisions of the GEMM producers, the MEMBAR/ERRBAR behind the epilogue's cluster-scope arrive and the
boundary mask of the attention kernel."""
import csv, subprocess, sys, io
rep=sys.argv[1]; n=int(sys.argv[2]) if len(sys.argv)>2 else 40
out=subprocess.run(["ncu","-i",rep,"--page","source","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(io.StringIO(out)))
# the source page of a multi-launch report is one section per launch: "Kernel Name" row, header row, instruction rows
sections=[]
for r in rows:
    if r and r[0]=="Kernel Name":
        sections.append({"name":r[1] if len(r)>1 else "", "hdr":None, "data":[]})
    elif sections and sections[-1]["hdr"] is None:
        sections[-1]["hdr"]=r
    elif sections and len(r)==len(sections[-1]["hdr"]):
        sections[-1]["data"].append(r)
for k,sec in enumerate(sections):
    hdr,data=sec["hdr"],sec["data"]
    isrc=hdr.index("Source"); isamp=hdr.index("# Samples"); iex=hdr.index("Instructions Executed")
    tot=sum(int(r[isamp]) for r in data)
    print(f"## launch {k}: {sec['name'][:100]}\ntotal samples",tot,"instrs",len(data))
    top=sorted(range(len(data)), key=lambda i:-int(data[i][isamp]))[:n]
    for i in sorted(top):
        r=data[i]; print(i, r[isamp], r[iex], r[isrc].strip()[:100])
