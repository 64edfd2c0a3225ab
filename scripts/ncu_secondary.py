"""Launches every secondary kernel once on moderate sizes (for the ncu capture of profiles/)."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from gol_b200 import World, ops

W, H = 8192, 4096
rng = np.random.default_rng(1)
cells = torch.from_numpy((rng.random((H, W)) < 0.5).astype(np.uint8)).cuda()
words = ops.pack(cells)                 # pack_kernel
back = ops.unpack(words, W)             # unpack_kernel
pop = ops.population(words, W)          # population_partial_kernel + population_final_kernel
nxt = ops.step_torus(words, W)          # step_rows_kernel
torch.cuda.synchronize()
assert int(pop.item()) == int(cells.sum().item()) and bool((back == cells).all())

w = World(8200, 2048, nranks=2, backend="cuda", mode="generic", ndevices=1)   # unaligned width: generic kernel
w.connect()
w.init_pattern(6, 3, 0.5)               # pattern_init_kernel
w.step(3)                               # halo_prime_kernel + life_generic_kernel
w.sync()
w.close()
w = World(4096, 1024, nranks=1, backend="cuda", mode="fused", ndevices=1)
w.connect()
w.init_pattern(6, 3, 0.5)
w.step(8)
w.sync()
with tempfile.TemporaryDirectory() as d:
    w.dump_text(0, d)                   # format_rows_kernel
w.close()
print("secondary kernels ok")
