import sys; sys.path.insert(0,'.')
import numpy as np
from blitzdg_b200.synthetic import structured_mesh, provision
from blitzdg_b200.initial_conditions import gaussian_hump
from blitzdg_b200.sw2d import SW2dSolver
for nx,ny in [(64,32),(250,125),(1000,500)]:
    ctx=provision(structured_mesh("tri",nx,ny),4).get_DGContext()
    H,h,hu,hv=gaussian_hump(ctx.x,ctx.y)
    sol=SW2dSolver(ctx,device=0)
    sol.set_state(h,hu,hv)
    sol.set_variant(1); Ra=sol.computeRHS()
    sol.set_variant(2); Rb=sol.computeRHS()
    import os
    os.environ["BDG_VOLUME_SCALAR"]="1"
    for a,b in zip(Ra,Rb):
        d=np.abs(a-b); i=np.unravel_index(d.argmax(),d.shape)
        print(nx,ny,"relerr",d.max()/np.abs(b).max(),"at",i,"a",a[i],"b",b[i], "max|b|",np.abs(b).max())
    sol.close()
