import sys, torch, ctypes as C
import os; sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from paddlebox_b200 import _native as N
from paddlebox_b200.ops.sparse_table import SparseTable, synthetic_key
from paddlebox_b200.parallel.sparse_engine import SparseEngine
dev = torch.device("cuda", 0)
B, S, dim, max_bag = 4096, 26, 16, 5
cfg = N.SparseOptConfig()
kpg = 20_000_000
t = SparseTable(int(kpg / 0.75) + 1024, dim, "adagrad", cfg)
t.fill_synthetic(kpg, 1, 0, 1, True)
eng = SparseEngine(t, B * S * max_bag)
g = torch.Generator().manual_seed(0)
lens = torch.randint(1, max_bag + 1, (S * B,), generator=g)
lod = torch.zeros(S * B + 1, dtype=torch.int32); lod[1:] = torch.cumsum(lens, 0)
n = int(lod[-1])
u = torch.rand(n, generator=g, dtype=torch.float64)
keys = synthetic_key((u * u * kpg).long().clamp_(max=kpg - 1), 1).to(dev)
lod = lod.to(dev)
H = 3 + dim
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
dout = torch.randn(B, S * H, device=dev); cvm = torch.ones(B, 2, device=dev)
r0 = eng.pull(keys)
def run(Bx, tag, do_flush=True):
    for _ in range(3):
        if do_flush: flush.zero_()
        eng.pull(keys, r0)
        # touch dout like the dense backward would
        dout.add_(0.0)
        eng.push_pooled(r0, dout[:Bx] if Bx else dout[:0], cvm, lod if Bx else lod[:1] * 0, Bx, S, H, 2, 0, -float(B), ld_dout=S * H)
    torch.cuda.synchronize()
    print(tag, eng.phase_times()["push"], flush=True)
run(B, "full push          ")
run(B, "full push, no flush", False)
run(0, "empty Q0/Q1 (B=0)  ")
