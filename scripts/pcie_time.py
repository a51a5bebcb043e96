"""Host link check for the e2e number: pinned H2D / D2H copy rates alone and together, at the e2e step's sizes."""
import torch
B, S = 64, 512
h = torch.empty(B * S * S * 3, dtype=torch.uint8).pin_memory(); d = torch.empty_like(h, device="cuda")
m = torch.empty(B * S * S, dtype=torch.uint8, device="cuda"); mh = torch.empty(B * S * S, dtype=torch.uint8).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, n=20):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    torch.cuda.synchronize(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
def up(): 
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
def down():
    with torch.cuda.stream(s2): mh.copy_(m, non_blocking=True)
def both(): up(); down()
a = t(up); b = t(down); c = t(both)
print(f"H2D {h.numel()/1e6:.1f} MB alone: {a*1e3:.0f} us = {h.numel()/a/1e6:.1f} GB/s")
print(f"D2H {m.numel()/1e6:.1f} MB alone: {b*1e3:.0f} us = {m.numel()/b/1e6:.1f} GB/s")
print(f"both concurrently: {c*1e3:.0f} us per pair -> H2D {h.numel()/c/1e6:.1f} GB/s")
