"""Where does the end-to-end pipeline lose its overlap?  GOL_PIPE_FLAGS: 1 no H2D, 2 no D2H, 4 no D2D."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gol_b200 import Life
from gol_b200 import _native as nat
W = 16384
life = Life(W, W)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); life.world.set_stream(0, stream.cuda_stream)
life.init(6, seed=1, density=0.5).sync()
stride = life.world.stride
hin = torch.empty((W, stride), dtype=torch.int64).pin_memory()
nat.check(nat.lib.gol_world_read_packed(life.world._h, 0, hin.data_ptr()))
hout = [torch.empty((W, stride), dtype=torch.int64).pin_memory() for _ in range(2)]
for _ in range(3): life.step(1024)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10): life.step(1024)
torch.cuda.synchronize()
print("kernel only ms/step", (time.perf_counter() - t0) * 100)
for _ in range(2): life.wait_result(life.run_from_host(hin, 1024, hout[0]))
torch.cuda.synchronize()
t0 = time.perf_counter(); prev = None
for i in range(10):
    slot = life.run_from_host(hin, 1024, hout[i & 1])
    if prev is not None: life.wait_result(prev)
    prev = slot
life.wait_result(prev); torch.cuda.synchronize()
print("flags", os.environ.get("GOL_PIPE_FLAGS", "0"), "pipelined ms/step", (time.perf_counter() - t0) * 100)
