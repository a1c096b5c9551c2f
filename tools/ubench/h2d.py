import torch, time
n = 280*1024*1024
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device='cuda')
for _ in range(3): d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): d.copy_(h, non_blocking=True)
e1.record(); torch.cuda.synchronize()
print("H2D one 280 MB copy: %.1f GB/s" % (10*n/ (e0.elapsed_time(e1)*1e-3)/1e9))
# chunks of 11 MB on 8 streams
ss = [torch.cuda.Stream() for _ in range(8)]
c = 11*1024*1024
e0.record()
for it in range(10):
    for i in range(24):
        with torch.cuda.stream(ss[i % 8]):
            d[i*c:(i+1)*c].copy_(h[i*c:(i+1)*c], non_blocking=True)
torch.cuda.synchronize()
e1.record(); torch.cuda.synchronize()
print("H2D 24 x 11 MB on 8 streams: %.1f GB/s" % (10*24*c/(e0.elapsed_time(e1)*1e-3)/1e9))
