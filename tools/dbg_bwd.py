import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'pytorch-kaldi-cgs_b200')); sys.path.insert(0, ROOT)
import numpy as np, torch
from b200 import native
from oracle import hcgs_oracle
H, T, B, cl = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
blocks = [32, 4] if H >= 128 else [8, 4]
d = torch.device('cuda', 0)
np.random.seed(3)
mask = torch.from_numpy(hcgs_oracle.conn_mat(H, H, blocks, [50, 50]).astype(np.float32))
plan = native.Plan(mask, recurrent=True)
us = [(torch.randn(H, H) * 0.1 * mask).to(d) for _ in range(4)]
cfg = native.TcConfig(plan, cl, 16); ldH = native.pad8(H)
print('cfg', cfg.cluster_size, cfg.smem_bwd, cfg.strips_per_cta)
gates = torch.rand(T * B, 4 * ldH, device=d); c = torch.randn(T * B, ldH, device=d)
bp = torch.empty(cfg.bwd_pack_u32, device=d, dtype=torch.int32)
native.lstm_tc_pack(plan, cfg, us, None, bp)
dh = torch.randn(T * B, ldH, device=d); dpre = torch.empty(T * B, 4 * ldH, device=d)
dpb = torch.empty(T * B, 4 * ldH, device=d, dtype=torch.bfloat16)
os.environ['HCGS_TC_DEBUG'] = str(512 + 64 + int(sys.argv[5]) if len(sys.argv) > 5 else 576)
native.lstm_tc_bwd(plan, cfg, bp, dh, gates, c, None, 1.0, 2, T, B, dpre, dpb)
torch.cuda.synchronize()
tr = native.tc_trace_read()
print('trace', tr)
if tr[15] < 0:
    v = -tr[15]
    print('TIMEOUT tag %d block %d warp %d parity %d step %d' % (v // 100000000, v // 1000000 % 100, v // 10000 % 100, v // 1000 % 10, v % 1000))
