import sys, numpy as np, time, torch
sys.path.insert(0,'/root/repo')
import qcb200
from qcb200 import workloads, Evaluator
def run(name,N=None,S=None,reps=3):
    w=workloads.make(name,N=N,S=S); L=w.layout
    ev=Evaluator(L,w.G_drift,w.G_drives)
    dZ=torch.from_numpy(w.Z).cuda(); dmu=torch.from_numpy(w.mu).cuda()
    dd=torch.empty(ev.defect_dim,dtype=torch.float64,device='cuda'); dj=torch.empty(ev.jacobian_nnz,dtype=torch.float64,device='cuda'); dh=torch.empty(ev.hessian_nnz,dtype=torch.float64,device='cuda')
    ev.eval_all_device(dZ,dmu,dd,dj,dh); torch.cuda.synchronize()
    ts=[]
    for _ in range(reps):
        t=time.perf_counter(); ev.eval_all_device(dZ,dmu,dd,dj,dh); torch.cuda.synchronize(); ts.append(time.perf_counter()-t)
    units=L.S*L.K; best=min(ts)
    print(f"{name}: units={units} d={L.d} best {best*1e3:.3f} ms -> {units/best:.0f} knot-evals/s  launches/eval={ev.kernel_launches//(reps+1)} jac_nnz={ev.jacobian_nnz}",flush=True)
    ev.close()
run("cfg1"); run("cfg2"); run("cfg3"); run("cfg3_ket"); run("cfg4"); run("cfg5",N=257); run("cfg5")
