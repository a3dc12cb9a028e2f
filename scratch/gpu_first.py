import sys, numpy as np, time
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/oracle')
import qcb200
from qcb200 import workloads, Evaluator
import qc_oracle as O, qc_port as P
def rel(a,b): return float(np.abs(a-b).max()/max(np.abs(b).max(),1e-300))
for name,N,S,fp in [("cfg1_drift",12,None,0),("cfg2",9,None,0),("cfg3",6,None,0),("cfg3_ket",6,None,0),("cfg4",8,5,0),
                    ("cfg1_drift",12,None,2),("cfg3",5,None,2),("cfg4",6,3,2),("cfg5",5,None,0)]:
    w=workloads.make(name,N=N,S=S); L=w.layout
    OL=O.Layout(L.d,L.c,L.m,L.N,L.S,L.z_dim,L.x_off,L.u_off,L.dt_off,L.jac_mode)
    d0,j0,h0=P.evaluate(OL,w.G_drift,w.G_drives,w.Z,w.mu)
    ev=Evaluator(L,w.G_drift,w.G_drives,force_path=fp)
    t=time.time(); d1,j1,h1=ev.eval_all(w.Z,w.mu); dt=time.time()-t
    print(name,"path",fp,"d",L.d,"units",L.S*L.K,"delta",rel(d1,d0),"jac",rel(j1,j0),"hess",rel(h1,h0),"launches",ev.kernel_launches,"%.3fs"%dt, flush=True)
    r,c=ev.jacobian_structure(); ro,co=OL.jacobian_structure(); assert (r==ro).all() and (c==co).all()
    r,c=ev.hessian_structure(); ro,co=OL.hessian_structure(); assert (r==ro).all() and (c==co).all()
    ev.close()
print("OK")
