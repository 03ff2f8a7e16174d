import numpy as np, sys
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from thinvids_b200 import frames as F
from thinvids_b200.swscale import Scaler
from oracle import port
cases=[(0,2,135,0,457,44,0x20),(0,2,14,1,2,7,0x20),(0,2,290,0,2,145,0x20),(1,2,35,2,2,17,0x4),(1,3,281,0,2,140,0x2),(0,2,14,0,2,7,0x20),(0,4,14,0,4,7,0x20),(0,8,14,0,8,7,0x20),(0,2,64,0,2,32,4)]
for (sf,sw,sh,df,dw,dh,alg) in cases:
    src=F.random_frame(sf,sw,sh,5)
    want=port.PortScaler(sw,sh,sf,dw,dh,df,alg).scale(src)
    with Scaler(sw,sh,sf,dw,dh,df,alg) as s:
        try: s.set_path("slab")
        except Exception as e: print((sf,sw,sh,df,dw,dh,hex(alg)),"slab refused",e); continue
        got=s.scale(src)
    for pi,(a,b) in enumerate(zip(got,want)):
        bad=np.argwhere(a!=b)
        if len(bad):
            print((sf,sw,sh,df,dw,dh,hex(alg)),"plane",pi,a.shape,"nbad",len(bad),"first",bad[:6].tolist())
            print(" got ",a[tuple(bad[0])],a.reshape(a.shape[0],-1)[:4].tolist()); print(" want",b[tuple(bad[0])],b.reshape(b.shape[0],-1)[:4].tolist())
    print((sf,sw,sh,df,dw,dh,hex(alg)),"done")
