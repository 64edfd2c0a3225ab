"""Does a concurrent copy slow the step kernels?  Time life.step while a side stream runs copies."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gol_b200 import Life
W = 16384
mode = sys.argv[1] if len(sys.argv) > 1 else "auto"
life = Life(W, W, mode=mode)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); life.world.set_stream(0, stream.cuda_stream)
life.init(6, seed=1, density=0.5).sync()
side = torch.cuda.Stream()
n = 32 << 20
hbuf = torch.empty(n, dtype=torch.uint8).pin_memory()
dbuf = torch.empty(n, dtype=torch.uint8, device="cuda")
dbuf2 = torch.empty(n, dtype=torch.uint8, device="cuda")
big1 = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
big2 = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3): life.step(1024)
torch.cuda.synchronize()

def timed(label, side_fn, reps=4):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record(stream)
    for _ in range(reps):
        life.step(1024)
    b.record(stream)
    if side_fn:
        with torch.cuda.stream(side):
            side_fn()
    torch.cuda.synchronize()
    print(f"{mode:10s} {label:34s} {a.elapsed_time(b) / reps:8.3f} ms/step", flush=True)

timed("alone", None)
timed("H2D 32MiB x12", lambda: [dbuf.copy_(hbuf, non_blocking=True) for _ in range(12)])
timed("D2H 32MiB x12", lambda: [hbuf.copy_(dbuf, non_blocking=True) for _ in range(12)])
timed("D2D 512MiB x12 (cudaMemcpy)", lambda: [big2.copy_(big1, non_blocking=True) for _ in range(12)])
timed("D2D 32MiB x100", lambda: [dbuf2.copy_(dbuf, non_blocking=True) for _ in range(100)])
timed("alone again", None)
# control: is a plain ALU-bound torch kernel slowed as well?
x = torch.randint(0, 2**31 - 1, (64 << 20,), dtype=torch.int32, device="cuda")
def alu(reps=40):
    y = x
    for _ in range(reps):
        y = (y ^ (y << 3)) + 12345
    return y
def timed2(label, side_fn):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record(stream); alu(); b.record(stream)
    if side_fn:
        with torch.cuda.stream(side): side_fn()
    torch.cuda.synchronize()
    print(f"torch elementwise  {label:34s} {a.elapsed_time(b):8.3f} ms", flush=True)
timed2("alone", None)
timed2("H2D 32MiB x12", lambda: [dbuf.copy_(hbuf, non_blocking=True) for _ in range(12)])
timed2("alone", None)
