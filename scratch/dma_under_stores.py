"""PCIe device->host bandwidth with idle host cores and beside the library's pool of streaming stores
(diagnostic for the e2e spread between boxes).  usage: python scratch/dma_under_stores.py [tag]"""
import ctypes
import os
import sys
import threading
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qcb200

tag = sys.argv[1] if len(sys.argv) > 1 else ""
lib = qcb200.load_library()
GB = 1 << 30
src = torch.empty(2 * GB // 8, dtype=torch.float64, device="cuda")
dst = torch.empty(2 * GB // 8, dtype=torch.float64, pin_memory=True)
fill = torch.empty(4 * GB // 8, dtype=torch.float64, pin_memory=True)
stream = torch.cuda.Stream()


def d2h(reps=4):
    with torch.cuda.stream(stream):
        dst.copy_(src, non_blocking=True)
        stream.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            dst.copy_(src, non_blocking=True)
        stream.synchronize()
    return reps * 2 * GB / (time.perf_counter() - t0) / 1e9


sec = ctypes.c_double(0.0)
lib.qcb_host_stream_fill(fill.data_ptr(), 4 * GB, ctypes.byref(sec))
lib.qcb_host_stream_fill(fill.data_ptr(), 4 * GB, ctypes.byref(sec))
print(tag, "threads", lib.qcb_host_threads(), "fill alone GB/s %.1f" % (4 * GB / sec.value / 1e9), flush=True)
print(tag, "D2H alone GB/s %.1f" % d2h(), flush=True)

stop = False
rates = []


def filler():
    s = ctypes.c_double(0.0)
    while not stop:
        lib.qcb_host_stream_fill(fill.data_ptr(), 4 * GB, ctypes.byref(s))
        rates.append(4 * GB / s.value / 1e9)


th = threading.Thread(target=filler)
th.start()
time.sleep(0.2)
r = d2h(8)
stop = True
th.join()
print(tag, "D2H beside the stores GB/s %.1f; fill meanwhile GB/s %s" % (r, " ".join("%.0f" % x for x in rates)), flush=True)
