"""cProfile of the Python side of one train step (host enqueue only; the GPU queue is drained outside the profile)."""
import os, sys, time, cProfile, pstats, io
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
dev = torch.device('cuda', 0)
torch.cuda.set_device(0)
nets, graph, opts = bench.build_workload(dev)
T, B = 256, 528
inp, _ = bench.synthetic_minibatch(T, B, 1)
x = inp.to(dev)
for _ in range(8):
    bench.train_step(graph, opts, x, T, B)
torch.cuda.synchronize()
K = 10
pr = cProfile.Profile()
t0 = time.perf_counter()
pr.enable()
for _ in range(K):
    bench.train_step(graph, opts, x, T, B)
pr.disable()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print('host enqueue %.2f ms per step (under cProfile), until drained %.2f ms per step' % ((t1 - t0) / K * 1e3, (t2 - t0) / K * 1e3))
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats('tottime').print_stats(28)
print(s.getvalue()[:6000])
