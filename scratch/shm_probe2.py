import torch, numpy as np, os, time
import torch.distributed as dist
from multiprocessing import shared_memory
rank=int(os.environ["RANK"]); local=int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev=torch.device("cuda",local)
dist.init_process_group("nccl",device_id=dev)
os.system("free -g | head -2") if rank==0 else None
for gb in (2, 21):
    n=int(gb*(1<<30)); name=f"probe2_{gb}"
    if rank==0: shm=shared_memory.SharedMemory(create=True,size=n,name=name)
    dist.barrier()
    if rank!=0: shm=shared_memory.SharedMemory(name=name)
    buf=np.ndarray((n//8,),dtype=np.float64,buffer=shm.buf)
    for attempt in ("whole","half"):
        if attempt=="whole": ptr,sz=buf.ctypes.data,n
        else: ptr,sz=buf.ctypes.data+rank*(n//2),n//2
        rc=torch.cuda.cudart().cudaHostRegister(ptr,sz,0)
        print(rank,gb,attempt,"rc",int(rc),flush=True)
        if int(rc)==0: torch.cuda.cudart().cudaHostUnregister(ptr)
        dist.barrier()
    del buf; shm.close(); dist.barrier()
    if rank==0: shm.unlink()
dist.destroy_process_group()
