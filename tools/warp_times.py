"""Diagnosis: clock stamps of every warp of CTA 0 inside one time step of lstm_tc_bwd (HCGS_TC_DEBUG=128)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'pytorch-kaldi-cgs_b200')); sys.path.insert(0, ROOT)
import numpy as np, torch
from b200 import native
from oracle import hcgs_oracle
H, T, B, cl = 550, 256, 528, 4
d = torch.device('cuda', 0)
np.random.seed(3)
mask = torch.from_numpy(hcgs_oracle.conn_mat(H, H, [32, 4], [75, 62.5]).astype(np.float32))
plan = native.Plan(mask, recurrent=True)
us = [(torch.randn(H, H) * 0.1 * mask).to(d) for _ in range(4)]
cfg = native.TcConfig(plan, cl, 16)
ldH = native.pad8(H)
pre = torch.randn(T * B, 4 * ldH, device=d); drop = torch.bernoulli(torch.full((B, H), 0.8, device=d))
h = torch.empty(T * B, ldH, device=d); c = torch.empty(T * B, ldH, device=d); gates = torch.empty(T * B, 4 * ldH, device=d)
hb = torch.empty(T * B, ldH, device=d, dtype=torch.bfloat16)
fp = torch.empty(cfg.fwd_pack_u32, device=d, dtype=torch.int32); bp = torch.empty(cfg.bwd_pack_u32, device=d, dtype=torch.int32)
native.lstm_tc_pack(plan, cfg, us, fp, bp)
native.lstm_tc_fwd(plan, cfg, fp, pre, drop, 1.0, 2, T, B, h, hb, c, gates)
dh = torch.randn(T * B, ldH, device=d); dpre = torch.empty(T * B, 4 * ldH, device=d)
dpb = torch.empty(T * B, 4 * ldH, device=d, dtype=torch.bfloat16)
torch.cuda.synchronize()
os.environ['HCGS_TC_DEBUG'] = '128'
for rep in range(2):
    native.lstm_tc_bwd(plan, cfg, bp, dh, gates, c, drop, 1.0, 2, T, B, dpre, dpb)
    torch.cuda.synchronize()
tl = native.tc_timeline_read()
t0 = min(r[0] for r in tl[:18] if r[0])
names = ['at bar1', 'after bar1', 'items done', 'after bar2', 'tasks start', 'task0 mma done', 'task0 push start', 'tasks done']
print('warp ' + ' '.join('%-16s' % n for n in names))
for w in range(18):
    print('%4d ' % w + ' '.join('%-16d' % (v - t0 if v else -1) for v in tl[w][:8]))
os.environ['HCGS_TC_DEBUG'] = '0'
