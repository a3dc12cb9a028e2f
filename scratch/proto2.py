import numpy as np, sys, math
sys.path.insert(0,'/root/repo/oracle')
import qc_oracle as O
c=[1.0/math.factorial(k) for k in range(13)]
def algo(G0,Gd,X,Xn,M,u,dt,theta=0.25):
    m=len(u); d=G0.shape[0]; I=np.eye(d)
    G=G0+np.tensordot(u,Gd,axes=1)
    nrm=np.abs(dt*G).sum(axis=0).max()
    s=0 if nrm<=theta else int(math.ceil(math.log2(nrm/theta)))
    tau=dt/2**s; A=tau*G
    W=M@X.T
    A2=A@A; A3=A@A2; A4=A2@A2
    Z2=c[8]*I+c[9]*A+c[10]*A2+c[11]*A3+c[12]*A4
    Z1=c[4]*I+c[5]*A+c[6]*A2+c[7]*A3+A4@Z2
    E=[c[0]*I+c[1]*A+c[2]*A2+c[3]*A3+A4@Z1]
    for l in range(s): E.append(E[-1]@E[-1])
    Wl=[None]*(s+1); Wl[s]=W
    for l in range(s-1,-1,-1): Wl[l]=Wl[l+1]@E[l].T+E[l].T@Wl[l+1]
    W0=Wl[0]
    Zb1=A4.T@W0; Zb2=A4.T@Zb1
    Ab4=W0@Z1.T+Zb1@Z2.T+c[12]*Zb2
    Ab3=c[3]*W0+c[7]*Zb1+c[11]*Zb2
    Ab2=c[2]*W0+c[6]*Zb1+c[10]*Zb2+Ab4@A2.T+A2.T@Ab4+A.T@Ab3
    H=np.zeros((m,m)); Ls=[]; T=[]
    for j in range(m):
        B=Gd[j]
        dA2=B@A+A@B; dA3=B@A2+A@dA2; dA4=dA2@A2+A2@dA2
        dZ2=c[9]*B+c[10]*dA2+c[11]*dA3+c[12]*dA4
        dZ1=c[5]*B+c[6]*dA2+c[7]*dA3+dA4@Z2+A4@dZ2
        L0=c[1]*B+c[2]*dA2+c[3]*dA3+dA4@Z1+A4@dZ1
        T.append(dict(B=B,dA2=dA2,dA4=dA4,dZ1=dZ1,dZ2=dZ2))
        Ls.append(L0)
    for j in range(m):
        t=T[j]
        Ua=t['B'].T@Ab2; Ub=t['B'].T@Ab3+t['dA2'].T@Ab4; Uc=t['dA4'].T@Zb1; Ud=t['dA4'].T@W0
        for i in range(m):
            H[j,i]=np.sum(Ua*T[i]['B'])+np.sum(Ub*T[i]['dA2'])+np.sum(Uc*T[i]['dZ2'])+np.sum(Ud*T[i]['dZ1'])
    for l in range(s):
        S=[Ls[j].T@Wl[l+1] for j in range(m)]
        for j in range(m):
            for i in range(m): H[j,i]+=np.sum(S[j]*Ls[i])
        Ls=[E[l]@Ls[j]+Ls[j]@E[l] for j in range(m)]
    Ef=E[s]; Ls=[tau*L for L in Ls]; H*=tau*tau
    GE=G@Ef; WE=W@Ef.T; GW=G.T@W
    out=dict(E=Ef,delta=Xn-Ef@X,J_u=np.array([-(L@X) for L in Ls]),J_dt=-(GE@X),
      H_xu=np.array([-(L.T@M) for L in Ls]),H_xdt=-(GE.T@M),H_uu=-(H+H.T),
      H_udt=np.array([-(np.sum(WE*Gd[j])+np.sum(GW*Ls[j])) for j in range(m)]),H_dtdt=-np.sum(GW*GE),s=s)
    return out
rng=np.random.default_rng(1)
def randH(n):
    A=rng.normal(size=(n,n))+1j*rng.normal(size=(n,n)); return (A+A.conj().T)/2
for n,m,dt in [(2,2,0.2),(3,2,1.0),(4,3,3.0),(9,4,0.5)]:
    d=2*n;cc=n
    G0=O.G_of_H(randH(n)); Gd=np.array([O.G_of_H(randH(n)) for _ in range(m)])
    X=rng.normal(size=(d,cc)); Xn=rng.normal(size=(d,cc)); M=rng.normal(size=(d,cc)); u=0.3*rng.normal(size=m)
    a=algo(G0,Gd,X,Xn,M,u,dt); o=O.unit_eval(G0,Gd,X,Xn,M,u,dt)
    print(n,m,dt,"s=",a['s'],{k:float(np.abs(a[k]-o[k]).max()/max(1e-300,np.abs(o[k]).max())) for k in o})
    print("  asym", np.abs(a['H_uu']-a['H_uu'].T).max())
