import sys, numpy as np, torch
sys.path.insert(0,'/root/repo')
import qcb200
from qcb200 import workloads, Evaluator
w=workloads.make("cfg4",N=13,S=5); L=w.layout
full=Evaluator(L,w.G_drift,w.G_drives); d,j,h=full.eval_all(w.Z,w.mu)
d=d.reshape(L.S,L.K,-1); j=j.reshape(L.S,L.K,-1); h=h.reshape(L.S,L.K,-1)
dZ=torch.from_numpy(w.Z).cuda(); dmu=torch.from_numpy(w.mu).cuda()
for k0,k1 in [(0,6),(6,12)]:
    ev=Evaluator(L,w.G_drift,w.G_drives,k_begin=k0,k_end=k1)
    dd=torch.empty(ev.defect_dim,dtype=torch.float64,device='cuda'); dj=torch.empty(ev.jacobian_nnz,dtype=torch.float64,device='cuda'); dh=torch.empty(ev.hessian_nnz,dtype=torch.float64,device='cuda')
    ev.eval_all_device(dZ,dmu,dd,dj,dh)
    print(k0,k1,np.abs(dd.cpu().numpy().reshape(L.S,k1-k0,-1)-d[:,k0:k1]).max(),np.abs(dj.cpu().numpy().reshape(L.S,k1-k0,-1)-j[:,k0:k1]).max(),np.abs(dh.cpu().numpy().reshape(L.S,k1-k0,-1)-h[:,k0:k1]).max())
    hd,hj,hh=ev.eval_all(w.Z,w.mu)
    print("  host",np.abs(hd.reshape(L.S,k1-k0,-1)-d[:,k0:k1]).max())
