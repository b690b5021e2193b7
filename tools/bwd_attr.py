import os, sys
ROOT = '/root/repo'
sys.path.insert(0, os.path.join(ROOT, 'pytorch-kaldi-cgs_b200')); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tools'))
import numpy as np, torch
from b200 import native
from oracle import hcgs_oracle
from bench_rec import timed
H, T = 550, 256
d = torch.device('cuda', 0)
np.random.seed(3)
mask = torch.from_numpy(hcgs_oracle.conn_mat(H, H, [32, 4], [75, 62.5]).astype(np.float32))
plan = native.Plan(mask, recurrent=True)
us = [(torch.randn(H, H) * 0.1 * mask).to(d) for _ in range(4)]
for B, cl in [(528, 4)]:
    cfg = native.TcConfig(plan, cl, 16)
    ldH = native.pad8(H)
    pre = torch.randn(T * B, 4 * ldH, device=d); drop = torch.bernoulli(torch.full((B, H), 0.8, device=d))
    h = torch.empty(T * B, ldH, device=d); c = torch.empty(T * B, ldH, device=d); gates = torch.empty(T * B, 4 * ldH, device=d)
    hb = torch.empty(T * B, ldH, device=d, dtype=torch.bfloat16)
    fp = torch.empty(cfg.fwd_pack_u32, device=d, dtype=torch.int32)
    bp = torch.empty(cfg.bwd_pack_u32, device=d, dtype=torch.int32)
    native.lstm_tc_pack(plan, cfg, us, fp, bp)
    os.environ['HCGS_TC_DEBUG'] = '0'
    native.lstm_tc_fwd(plan, cfg, fp, pre, drop, 1.0, 2, T, B, h, hb, c, gates)
    dh = torch.randn(T * B, ldH, device=d); dpre = torch.empty(T * B, 4 * ldH, device=d)
    dpb = torch.empty(T * B, 4 * ldH, device=d, dtype=torch.bfloat16)
    for flags in [0, 4]:
        os.environ['HCGS_TC_DEBUG'] = str(flags)
        t = timed(lambda: native.lstm_tc_bwd(plan, cfg, bp, dh, gates, c, drop, 1.0, 2, T, B, dpre, dpb))
        print('bwd B=%d cl=%d flags=%d: %.3f ms  %.2f us/step' % (B, cl, flags, t, t * 1e3 / T), flush=True)
os.environ['HCGS_TC_DEBUG'] = '0'
