"""How fast is cudaHostRegister on a populated MAP_SHARED mapping of a tmpfs file, and how fast is a D2H copy straight into it?
Decides whether the LOCO writer should DMA into the output file's pages instead of staging through page-locked buffers."""
import ctypes
import json
import mmap
import os
import sys
import time

import torch

path = sys.argv[1] if len(sys.argv) > 1 else "/dev/shm/ggrm_hostreg_probe"
n = int(float(sys.argv[2]) * (1 << 30)) if len(sys.argv) > 2 else 4 << 30
rt = ctypes.CDLL("libcudart.so.12")
fd = os.open(path, os.O_CREAT | os.O_RDWR | os.O_TRUNC, 0o644)
os.ftruncate(fd, n)
m = mmap.mmap(fd, n)
addr = ctypes.addressof(ctypes.c_char.from_buffer(m))
out = {"bytes": n}
t = time.time()
m.madvise(23)  # MADV_POPULATE_WRITE (single thread here; the engine uses several)
out["populate_s"] = time.time() - t
src = torch.full((n,), 7, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
t = time.time()
rc = rt.cudaHostRegister(ctypes.c_void_p(addr), ctypes.c_size_t(n), 0)
out["register_rc"], out["register_s"] = rc, time.time() - t
if rc == 0:
    for rep in range(2):
        t = time.time()
        rc2 = rt.cudaMemcpy(ctypes.c_void_p(addr), ctypes.c_void_p(src.data_ptr()), ctypes.c_size_t(n), 2)
        out[f"d2h_direct_s_{rep}"] = time.time() - t
    out["d2h_rc"] = rc2
    out["check"] = m[0] == 7 and m[n - 1] == 7
    t = time.time()
    rt.cudaHostUnregister(ctypes.c_void_p(addr))
    out["unregister_s"] = time.time() - t
# the staged path for comparison: D2H into a page-locked buffer, then memcpy into the (populated) mapping
pin = torch.empty(256 << 20, dtype=torch.uint8, pin_memory=True)
t = time.time()
for off in range(0, n, pin.numel()):
    c = min(pin.numel(), n - off)
    pin[:c].copy_(src[off:off + c])
    ctypes.memmove(addr + off, pin.data_ptr(), c)
out["d2h_staged_1thread_s"] = time.time() - t
print(json.dumps(out))
del m
os.close(fd)
os.unlink(path)
