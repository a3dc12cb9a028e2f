import torch, numpy as np, os
from multiprocessing import shared_memory
os.system("df -h /dev/shm | tail -1; ulimit -l")
torch.cuda.init()
for gb in (0.5, 4, 22):
    n=int(gb*(1<<30))
    try:
        shm=shared_memory.SharedMemory(create=True,size=n,name=f"probe{int(gb*10)}")
    except Exception as e:
        print(gb,"shm create failed",e); continue
    buf=np.ndarray((n//8,),dtype=np.float64,buffer=shm.buf)
    rc=torch.cuda.cudart().cudaHostRegister(buf.ctypes.data,n,0)
    print(gb,"GB register rc",int(rc),rc)
    if int(rc)==0: torch.cuda.cudart().cudaHostUnregister(buf.ctypes.data)
    del buf; shm.close(); shm.unlink()
