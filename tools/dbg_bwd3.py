"""Small-shape debug run of the tensor-core backward (HCGS_TC_DEBUG=512 reports the first wait that timed out)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'pytorch-kaldi-cgs_b200')); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np, torch
from b200 import native
from oracle import hcgs_oracle
H, B, T = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
blocks = [int(sys.argv[4]), int(sys.argv[5])]
cl = int(sys.argv[6])
d = torch.device('cuda', 0)
np.random.seed(3)
mask = torch.from_numpy(hcgs_oracle.conn_mat(H, H, blocks, [50, 50]).astype(np.float32))
plan = native.Plan(mask, recurrent=True)
cfg = native.TcConfig(plan, cl, 16)
print('cluster', cfg.cluster_size, 'smem bwd', cfg.smem_bwd, 'blocks', cfg.n_blocks, flush=True)
us = [(torch.randn(H, H) * 0.1 * mask).to(d) for _ in range(4)]
ldH = native.pad8(H)
pre = torch.randn(T * B, 4 * ldH, device=d); drop = torch.bernoulli(torch.full((B, H), 0.8, device=d))
h = torch.empty(T * B, ldH, device=d); c = torch.empty(T * B, ldH, device=d); gates = torch.empty(T * B, 4 * ldH, device=d)
hb = torch.empty(T * B, ldH, device=d, dtype=torch.bfloat16)
fp = torch.empty(cfg.fwd_pack_u32, device=d, dtype=torch.int32)
bp = torch.empty(max(cfg.bwd_pack_u32, 1), device=d, dtype=torch.int32)
native.lstm_tc_pack(plan, cfg, us, fp, bp)
native.lstm_tc_fwd(plan, cfg, fp, pre, drop, 1.0, 2, T, B, h, hb, c, gates)
torch.cuda.synchronize(); print('fwd ok', flush=True)
dh = torch.randn(T * B, ldH, device=d); dpre = torch.empty(T * B, 4 * ldH, device=d)
dpb = torch.empty(T * B, 4 * ldH, device=d, dtype=torch.bfloat16)
os.environ['HCGS_TC_DEBUG'] = os.environ.get('DBG', '512')
native.lstm_tc_bwd(plan, cfg, bp, dh, gates, c, drop, 1.0, 2, T, B, dpre, dpb)
torch.cuda.synchronize(); print('bwd ok', flush=True)
tr = native.tc_trace_read()
print('trace[15] (negative = {tag, block, warp, parity, step} of the first timed-out wait):', tr[15])
