import sys; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np, _oracle as O, brotli_lib_b200 as B
ctx=B.Context()
def bits(b,n): return ''.join(str((b[i>>3]>>(i&7))&1) for i in range(n))
for cmap,ncl in [(np.arange(64)%2,2),(np.arange(8)%2,2),(np.array([0,1]),2),(np.array([0,1,1,1]),2),(np.arange(9)%3,3),(np.arange(16)%4,4),(np.arange(25)%5,5),(np.array([0,0,0,1,1,1,0,0]),2)]:
    cmap=cmap.astype(np.uint32)
    g,gn=ctx.stage_context_map(cmap,ncl); o,on=O.encode_context_map(cmap,ncl)
    print(list(cmap[:10]),ncl,gn,on,g==o); 
    if g!=o: print(' gpu',bits(g,gn)); print(' ora',bits(o,on))
