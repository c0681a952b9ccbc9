import sys, pathlib
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent)); sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent/"tests"))
import numpy as np, zkevm_prover_b200 as zk
P=0xFFFFFFFF00000001
edge = [0, 1, 2, 7, 0xFFFFFFFF, 0x100000000, 0x100000001, P - 1, P, P + 1, 2**63, 0xFFFFFFFF00000000, 2**64 - 1, 0x123456789abcdef0]
a=np.array([x for x in edge for _ in edge],dtype=np.uint64); b=np.array([y for _ in edge for y in edge],dtype=np.uint64)
for op,f in (("canon",lambda x,y:x%P),("add",lambda x,y:(x+y)%P),("sub",lambda x,y:(x-y)%P),("mul",lambda x,y:x*y%P),("mul_nc",lambda x,y:x*y%P)):
    got=zk.field_op(op,a,b); bad=0
    for i in range(a.size):
        w=f(int(a[i]),int(b[i]))
        if int(got[i])!=w:
            bad+=1
            if bad<=6: print(op,"a=%016x b=%016x got=%016x want=%016x"%(int(a[i]),int(b[i]),int(got[i]),w))
    print(op,"bad",bad,"of",a.size)
