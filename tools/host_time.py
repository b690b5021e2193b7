"""Host (Python + launch) time per train step vs GPU time: enqueue K steps without synchronising, then drain."""
import os, sys, time
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
t0 = time.perf_counter()
for _ in range(K):
    bench.train_step(graph, opts, x, T, B)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print('host enqueue %.2f ms per step, until drained %.2f ms per step' % ((t1 - t0) / K * 1e3, (t2 - t0) / K * 1e3))
for n in nets.values():
    pf = n.__dict__.get('_drop_prefetch')
    if pf is not None:
        print('drop-mask prefetch: %d hits, %d misses' % (pf.hits, pf.misses))
