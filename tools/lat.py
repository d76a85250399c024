import sys, time; sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import brotli_lib_b200 as B, torch
from conftest import load_input
ctx=B.Context()
data=load_input("enc-ttf.bin")
for q,seg in ((11,0),(11,8192),(11,4096),(11,2048),(10,0),(5,0),(9,0)):
    ctx.set_tuning(segment_bytes=seg)
    for _ in range(2): out=ctx.encode(data,q,22,0)
    torch.cuda.synchronize(); t=time.perf_counter(); out=ctx.encode(data,q,22,0); dt=time.perf_counter()-t
    print(q,seg,f"{dt*1e3:.1f} ms",len(out),[(k,round(v,2)) for k,v in ctx.timings() if v>=0.3])
