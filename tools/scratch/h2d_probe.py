"""H2D bandwidth from pinned host memory with 1, 2, 4 concurrent copy streams and several copy sizes."""
import time
import torch

n = 149 << 20
h = torch.empty(n, dtype=torch.uint8).pin_memory()
h.fill_(3)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for streams in (1, 2, 4):
    for piece_mb in (0.45, 7.5, 149):
        piece = int(piece_mb * (1 << 20))
        ss = [torch.cuda.Stream() for _ in range(streams)]
        torch.cuda.synchronize()
        best = 1e9
        for rep in range(4):
            t0 = time.perf_counter()
            off, i = 0, 0
            while off < n:
                m = min(piece, n - off)
                with torch.cuda.stream(ss[i % streams]):
                    d[off:off + m].copy_(h[off:off + m], non_blocking=True)
                off += m
                i += 1
            torch.cuda.synchronize()
            best = min(best, time.perf_counter() - t0)
        print("streams %d piece %7.2f MB: %.2f ms  %.1f GB/s" % (streams, piece_mb, best * 1e3, n / best / 1e9))
