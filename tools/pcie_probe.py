"""PCIe probe (GPU box): device->host and host->device copy rates with pinned memory, in 43 MB pieces like the library's
chunk results, alone and while a DRAM-heavy kernel runs.  Prints ms per piece so a rate change along the buffer shows."""
import sys, time, torch
dev = torch.device("cuda", 0)
MB = 1 << 20
piece = 43 * MB
n = 24
h = torch.empty(n * piece, dtype=torch.uint8, pin_memory=True)
d = torch.empty(n * piece, dtype=torch.uint8, device=dev)
big_a = torch.empty(2 << 30, dtype=torch.uint8, device=dev)
big_b = torch.empty(2 << 30, dtype=torch.uint8, device=dev)
s_copy, s_k = torch.cuda.Stream(), torch.cuda.Stream()


def run(direction, busy):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
    torch.cuda.synchronize()
    if busy:
        with torch.cuda.stream(s_k):
            for _ in range(40):
                big_b.copy_(big_a)
    with torch.cuda.stream(s_copy):
        ev[0].record()
        for i in range(n):
            a, b = i * piece, (i + 1) * piece
            if direction == "d2h":
                h[a:b].copy_(d[a:b], non_blocking=True)
            else:
                d[a:b].copy_(h[a:b], non_blocking=True)
            ev[i + 1].record()
    torch.cuda.synchronize()
    ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(n)]
    tot = ev[0].elapsed_time(ev[n])
    print(f"{direction} busy={busy}: total {tot:.2f} ms = {n * piece / tot / 1e6:.1f} GB/s; per piece ms:", " ".join(f"{x:.2f}" for x in ms), flush=True)


for rep in range(2):
    for direction in ("d2h", "h2d"):
        for busy in (False, True):
            run(direction, busy)
# both directions at once
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
s2 = torch.cuda.Stream()
h2 = torch.empty(n * piece, dtype=torch.uint8, pin_memory=True)
d2 = torch.empty(n * piece, dtype=torch.uint8, device=dev)
torch.cuda.synchronize()
ev[0].record()
with torch.cuda.stream(s_copy):
    h.copy_(d, non_blocking=True)
with torch.cuda.stream(s2):
    d2.copy_(h2, non_blocking=True)
torch.cuda.synchronize()
ev[1].record()
torch.cuda.synchronize()
t = ev[0].elapsed_time(ev[1])
print(f"duplex: {n * piece / 1e6:.0f} MB each way in {t:.2f} ms = {n * piece / t / 1e6:.1f} GB/s per direction")
