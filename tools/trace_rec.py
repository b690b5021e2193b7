"""Per-phase clock trace of the tensor-core recurrent kernels (HCGS_TC_DEBUG=64)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'pytorch-kaldi-cgs_b200')); sys.path.insert(0, ROOT)
import numpy as np, torch
from b200 import native
from oracle import hcgs_oracle
H, T = 550, 256
d = torch.device('cuda', 0)
np.random.seed(3)
mask = torch.from_numpy(hcgs_oracle.conn_mat(H, H, [32, 4], [75, 62.5]).astype(np.float32))
plan = native.Plan(mask, recurrent=True)
us = [(torch.randn(H, H) * 0.1 * mask).to(d) for _ in range(4)]
FWD = ['loop top', 'pre wait', 'mbar wait', 'mma', 'B1 sync', 'issue prefetch', 'gate math', 'B2 sync', 'push',
       'B3 sync', 'fence+arrive', 'global stores']
BWD = ['loop top', 'input waits', 'item phase', 'items barrier', 'mma + push', 'slots-free wait', '-', '-', '-']
for B, cl in [(528, 4), (128, 8)]:
    cfg = native.TcConfig(plan, cl, 16)
    ldH = native.pad8(H)
    pre = torch.randn(T * B, 4 * ldH, device=d); drop = torch.bernoulli(torch.full((B, H), 0.8, device=d))
    h = torch.empty(T * B, native.pad8(H), device=d); c = torch.empty(T * B, ldH, device=d); gates = torch.empty(T * B, 4 * ldH, device=d)
    hb = torch.empty(T * B, ldH, device=d, dtype=torch.bfloat16)
    fp = torch.empty(cfg.fwd_pack_u32, device=d, dtype=torch.int32)
    bp = torch.empty(cfg.bwd_pack_u32, device=d, dtype=torch.int32)
    native.lstm_tc_pack(plan, cfg, us, fp, bp)
    dh = torch.randn(T * B, native.pad8(H), device=d); dpre = torch.empty(T * B, 4 * ldH, device=d)
    dpb = torch.empty(T * B, 4 * ldH, device=d, dtype=torch.bfloat16)
    os.environ['HCGS_TC_DEBUG'] = '64'
    for rep in range(2):
        native.lstm_tc_fwd(plan, cfg, fp, pre, drop, 1.0, 2, T, B, h, hb, c, gates)
        torch.cuda.synchronize()
    tr = native.tc_trace_read()
    tot = sum(tr[:12])
    print('fwd B=%d cl=%d: %.0f cycles/step' % (B, cl, tot / T))
    for n, v in zip(FWD, tr):
        print('   %-16s %7.0f cyc/step %5.1f%%' % (n, v / T, 100.0 * v / tot))
    for rep in range(2):
        native.lstm_tc_bwd(plan, cfg, bp, dh, gates, c, drop, 1.0, 2, T, B, dpre, dpb)
        torch.cuda.synchronize()
    tr = native.tc_trace_read()
    tot = sum(tr[:9])
    print('bwd B=%d cl=%d: %.0f cycles/step' % (B, cl, tot / T))
    for n, v in zip(BWD, tr):
        print('   %-16s %7.0f cyc/step %5.1f%%' % (n, v / T, 100.0 * v / tot))
os.environ['HCGS_TC_DEBUG'] = '0'
