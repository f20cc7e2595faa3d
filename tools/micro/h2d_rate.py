"""H2D copy rate of the bench's 797 MB input from pinned host memory: one copy, split over 2 / 4 streams, and from
write-combined pinned memory (cudaHostAllocWriteCombined)."""
import ctypes as C
import torch

dev = torch.device("cuda:0")
n = 384 * 721 * 1440
dst = torch.empty(n, dtype=torch.bfloat16, device=dev)
src = torch.empty(n, dtype=torch.bfloat16).pin_memory()
src.view(torch.int16).random_(0, 1000)


def timed(fn, reps=8):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    return ms, n * 2 / ms / 1e6


print("one copy            %.2f ms  %.1f GB/s" % timed(lambda: dst.copy_(src, non_blocking=True)))
for k in (2, 4):
    streams = [torch.cuda.Stream() for _ in range(k)]
    step = n // k

    def split():
        cur = torch.cuda.current_stream()
        for i, s in enumerate(streams):
            s.wait_stream(cur)
            with torch.cuda.stream(s):
                dst[i * step:(i + 1) * step].copy_(src[i * step:(i + 1) * step], non_blocking=True)
        for s in streams:
            cur.wait_stream(s)
    print("split over %d streams %.2f ms  %.1f GB/s" % ((k,) + timed(split)))

rt = C.CDLL("libcudart.so.12")
p = C.c_void_p()
rc = rt.cudaHostAlloc(C.byref(p), C.c_size_t(n * 2), C.c_uint(0x04))          # cudaHostAllocWriteCombined
print("cudaHostAlloc(WC) rc", rc)
if rc == 0:
    buf = (C.c_char * (n * 2)).from_address(p.value)
    wc = torch.frombuffer(buf, dtype=torch.bfloat16)
    wc.copy_(src)
    print("is_pinned", wc.is_pinned())
    print("write-combined      %.2f ms  %.1f GB/s" % timed(lambda: dst.copy_(wc, non_blocking=True)))
    print("equal", bool(torch.equal(dst.cpu(), src)))
