"""PCIe probe on the GPU box: pinned H2D / D2H bandwidth alone and concurrently (sizes of the e2e legs)"""
import torch, time
dev = torch.device("cuda", 0)
def bw(nbytes_h2d, nbytes_d2h, reps=5):
    h_in = torch.empty(max(nbytes_h2d, 1), dtype=torch.uint8).pin_memory()
    d_in = torch.empty(max(nbytes_h2d, 1), dtype=torch.uint8, device=dev)
    h_out = torch.empty(max(nbytes_d2h, 1), dtype=torch.uint8).pin_memory()
    d_out = torch.empty(max(nbytes_d2h, 1), dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize(); t = time.time()
        if nbytes_h2d:
            with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
        if nbytes_d2h:
            with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize(); best = min(best, time.time() - t)
    return best
for a, b in ((240 << 20, 0), (0, 80 << 20), (240 << 20, 80 << 20), (140 << 20, 80 << 20), (100 << 20, 80 << 20), (4 << 20, 0), (0, 4 << 20), (1 << 20, 0)):
    t = bw(a, b)
    print("h2d %4d MiB  d2h %4d MiB : %.3f ms  h2d %.1f GB/s  d2h %.1f GB/s" % (a >> 20, b >> 20, t * 1e3, a / t / 1e9, b / t / 1e9), flush=True)
