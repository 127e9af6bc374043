import sys; sys.path.insert(0,'/root/repo')
import numpy as np
import stepth_b200 as sb
from stepth_b200 import synth
from oracle import oracle
ctx = sb.Context(0)
w,h,rings=96,64,128
m,a = synth.make_pair("geo", w,h, seed=w+h); m=m.numpy(); a=a.numpy()
ref = oracle.search_dense(m,a,[36]*3,rings,threads=8)
try:
    got = ctx.depth_dense(m,a,[36]*3,rings,sb.RESAMPLE_OFF,want_raw=True)["raw"]
except sb.StepthPanic as e:
    got = e.partial["raw"]
bad = np.argwhere(got!=ref)
print("n bad", len(bad), "of", w*h)
for (y,x) in bad[:40]:
    ox,oy=(x+1)//2,(y+1)//2
    r = oracle.distance_dot_array(m[y,x,:3], a, (ox,oy), rings, [36]*3)
    print((x,y),"origin",(ox,oy),"ref",ref[y,x],"got",got[y,x],"oracle:",r)
