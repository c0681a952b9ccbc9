import re, subprocess, sys, collections
f, fn = sys.argv[1], sys.argv[2]
ranges = [tuple(int(x,16) for x in r.split('-')) for r in sys.argv[3:]]
txt = subprocess.run(["cuobjdump","-sass",f],capture_output=True,text=True).stdout
on=False
per=[collections.Counter() for _ in ranges]
for line in txt.splitlines():
    if "Function :" in line:
        on = fn in line
        continue
    if not on: continue
    m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        a=int(m.group(1),16)
        for i,(lo,hi) in enumerate(ranges):
            if lo<=a<hi: per[i][m.group(2)]+=1
for (lo,hi),c in zip(ranges,per):
    print("range %x-%x: %d instr"%(lo,hi,sum(c.values())))
    g=collections.Counter()
    for k,v in c.items():
        base=k.split('.')[0]
        if k.startswith("IMAD.WIDE"): base="IMAD.WIDE"
        elif k.startswith("IMAD.HI"): base="IMAD.HI"
        elif k.startswith("IMAD.MOV"): base="IMAD.MOV"
        elif k.startswith("IMAD.IADD"): base="IMAD.IADD"
        elif k.startswith("IMAD.X"): base="IMAD.X"
        elif k.startswith("IMAD.SHL"): base="IMAD.SHL"
        g[base]+=v
    print("  "+", ".join("%s=%d"%kv for kv in g.most_common()))
