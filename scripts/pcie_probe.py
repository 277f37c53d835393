"""PCIe bound of the host-pointer calls: the copies of one 1 M x 31-mer batch (31 MB in, 16 MB out) alone,
together, and cut into the pipeline's chunks.  Prints GB/s and the queries/s each would allow."""
import torch, time
dev = torch.device("cuda:0")
Q, k = 1_000_000, 31
hin = torch.empty(Q * k, dtype=torch.uint8).pin_memory()
hout = torch.empty(2 * Q, dtype=torch.int64).pin_memory()
din = torch.empty(Q * k, dtype=torch.uint8, device=dev)
dout = torch.empty(2 * Q, dtype=torch.int64, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

def timed(fn, n=20):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n

def h2d(chunks=1):
    per = Q * k // chunks
    with torch.cuda.stream(s1):
        for c in range(chunks):
            din[c * per:(c + 1) * per].copy_(hin[c * per:(c + 1) * per], non_blocking=True)

def d2h(chunks=1):
    per = 2 * Q // chunks
    with torch.cuda.stream(s2):
        for c in range(chunks):
            hout[c * per:(c + 1) * per].copy_(dout[c * per:(c + 1) * per], non_blocking=True)

for ch in (1, 15, 30):
    a = timed(lambda: h2d(ch)); b = timed(lambda: d2h(ch)); c = timed(lambda: (h2d(ch), d2h(ch)))
    print(f"chunks {ch:2d}: H2D {Q*k/a/1e9:5.1f} GB/s ({a*1e6:5.0f} us)  D2H {16*Q/b/1e9:5.1f} GB/s ({b*1e6:5.0f} us)  "
          f"both {c*1e6:5.0f} us -> {Q/c/1e6:6.0f} M q/s")
