"""Raw pinned-memory copy bandwidth of the box: H2D alone, D2H alone, both directions at once."""
import time, torch
n = 373 * 10**6 // 4
h = torch.empty(n, dtype=torch.float32).pin_memory(); d = torch.empty(n, dtype=torch.float32, device="cuda")
h2 = torch.empty(n // 6, dtype=torch.float32).pin_memory(); d2 = torch.empty(n // 6, dtype=torch.float32, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(up, down, reps=10):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps):
        if up:
            with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
        if down:
            with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps
for _ in range(2): run(True, True)
t = run(True, False); print(f"H2D alone: {n*4/t/1e9:.1f} GB/s ({t*1e3:.2f} ms per 373 MB)")
t = run(False, True); print(f"D2H alone: {n//6*4/t/1e9:.1f} GB/s")
t = run(True, True); print(f"H2D 373 MB + D2H 62 MB concurrently: {t*1e3:.2f} ms -> H2D {n*4/t/1e9:.1f} GB/s")
# chunked like the pipeline: 14 chunks, 4 arrays each
chunks = torch.chunk(h, 56); dch = torch.chunk(d, 56)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(10):
    for a, b in zip(chunks, dch):
        with torch.cuda.stream(s1): b.copy_(a, non_blocking=True)
torch.cuda.synchronize(); t = (time.perf_counter() - t0) / 10
print(f"H2D in 56 pieces: {n*4/t/1e9:.1f} GB/s")
