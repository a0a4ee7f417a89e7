import torch
dev=torch.device("cuda:0")
def t(fn, n=10):
    fn(); fn(); torch.cuda.synchronize()
    s=torch.cuda.Event(enable_timing=True); e=torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e)/n*1e-3
for mb in (256, 1024, 4096):
    buf=torch.empty(mb<<20, dtype=torch.uint8, device=dev)
    sec=t(lambda: buf.zero_())
    print("memset torch fill %5d MiB: %.1f GB/s" % (mb, buf.numel()/sec/1e9))
    f=buf.view(torch.float32)
    sec=t(lambda: f.fill_(1.5))
    print("fill_ f32         %5d MiB: %.1f GB/s" % (mb, buf.numel()/sec/1e9))
    import ctypes
    rt=ctypes.CDLL("libcudart.so")
    sec=t(lambda: rt.cudaMemsetAsync(ctypes.c_void_p(buf.data_ptr()), 0, ctypes.c_size_t(buf.numel()), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    print("cudaMemsetAsync   %5d MiB: %.1f GB/s" % (mb, buf.numel()/sec/1e9))
a=torch.empty(1<<30, dtype=torch.uint8, device=dev); b=torch.empty_like(a)
sec=t(lambda: b.copy_(a)); print("copy 1GiB read+write: %.1f GB/s" % (2*a.numel()/sec/1e9))
sec=t(lambda: a.sum(dtype=torch.int64) if False else a.view(torch.float32).sum()); print("read-only sum 1GiB: %.1f GB/s" % (a.numel()/sec/1e9))
# write-only with NON-constant data (constant fills may be compressed on the way to HBM): broadcast an L2-resident row
row = torch.randn(1, 1 << 16, device=dev)
for mb in (256, 1024):
    out = torch.empty((mb << 20) // (4 << 16), 1 << 16, device=dev)
    sec = t(lambda: out.copy_(row.expand_as(out)))
    print("broadcast-row write (random data) %5d MiB: %.1f GB/s" % (mb, out.numel() * 4 / sec / 1e9))
    src = torch.randn(1 << 20, device=dev)            # 4 MB, L2 resident
    o2 = out.view(-1, 1 << 20)
    sec = t(lambda: o2.copy_(src.expand_as(o2)))
    print("broadcast-4MB write (random data)  %5d MiB: %.1f GB/s" % (mb, out.numel() * 4 / sec / 1e9))
