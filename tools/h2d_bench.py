"""Host->device copy bandwidth for the e2e input (154 MB fp32 images): pinned, write-combined, split across streams."""
import ctypes, glob, os, sys, time
import torch
n = 256 * 3 * 224 * 224
dev = torch.device("cuda")
dst = torch.empty(n, dtype=torch.float32, device=dev)
def timeit(fn, reps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
src = torch.empty(n, dtype=torch.float32).pin_memory()
src.normal_()
ms = timeit(lambda: dst.copy_(src, non_blocking=True))
print(f"pinned, 1 stream: {ms:.3f} ms  {n*4/ms/1e6:.1f} GB/s")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
h = n // 2
def two():
    e = torch.cuda.Event(); e.record()
    s1.wait_event(e); s2.wait_event(e)
    with torch.cuda.stream(s1): dst[:h].copy_(src[:h], non_blocking=True)
    with torch.cuda.stream(s2): dst[h:].copy_(src[h:], non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
ms = timeit(two)
print(f"pinned, 2 streams: {ms:.3f} ms  {n*4/ms/1e6:.1f} GB/s")
# write-combined allocation through the CUDA runtime torch already loaded
libs = glob.glob(os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "cuda_runtime", "lib", "libcudart.so*"))
rt = ctypes.CDLL(libs[0])
p = ctypes.c_void_p()
rc = rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(n * 4), ctypes.c_uint(4))   # cudaHostAllocWriteCombined
print("cudaHostAlloc WC rc", rc)
if rc == 0:
    buf = (ctypes.c_float * n).from_address(p.value)
    wc = torch.frombuffer(buf, dtype=torch.float32)
    t0 = time.time(); wc.copy_(src); print(f"host fill of WC buffer: {time.time()-t0:.3f} s")
    ms = timeit(lambda: dst.copy_(wc, non_blocking=True))
    print(f"write-combined, 1 stream: {ms:.3f} ms  {n*4/ms/1e6:.1f} GB/s")
    print("equal", bool((dst.cpu() == src).all()))
# D2D for scale
d2 = torch.empty_like(dst)
ms = timeit(lambda: d2.copy_(dst))
print(f"d2d: {ms:.3f} ms")
os.system("nvidia-smi --query-gpu=pcie.link.gen.current,pcie.link.width.current,pcie.link.gen.max --format=csv")
os.system("nvidia-smi topo -m 2>/dev/null | head -n 8; nproc; lscpu | grep -i numa")
