"""Micro-benchmark of the fused output layer (hcgs_out_logsoftmax_prior_fwd) against the unfused pair
(tc_gemm_nt + logsoftmax_fwd): CUDA events, 3 warm-ups, algorithmic HBM bytes = rows * (ldx*2 + P*4).
Usage: python tools/bench_out.py [P] [H] [rows]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'pytorch-kaldi-cgs_b200'))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from b200 import native
from oracle import hcgs_oracle


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def run(P, H, M):
    d = torch.device('cuda', 0)
    np.random.seed(5)
    mask = torch.from_numpy(hcgs_oracle.conn_mat(P, H, [64, 4], [50, 25]).astype(np.float32))
    plan = native.Plan(mask)
    info = native.out_info(plan)
    w = (torch.randn(P, H) * mask * 0.05).to(d)
    bias = torch.randn(P, device=d)
    lp = torch.log_softmax(torch.randn(P, device=d), 0)
    ldx = native.pad8(H)
    xb = torch.randn(M, ldx, device=d).to(torch.bfloat16)
    out = torch.empty(M, P, device=d)
    opack = torch.empty(info['pack_bytes'] // 2, device=d, dtype=torch.bfloat16)
    native.out_pack(plan, w, opack)
    t_f = timed(lambda: native.out_logsoftmax_prior_fwd(plan, opack, xb, bias, lp, out))
    wpack = torch.empty(plan.nz_tiles * 32 * 32, device=d, dtype=torch.bfloat16)
    native.tc_pack_nt(plan, [w], wpack)
    z = torch.empty(M, P, device=d)
    out2 = torch.empty(M, P, device=d)

    def unfused():
        native.tc_gemm_nt(plan, 1, wpack, xb, z, [bias])
        native.logsoftmax_fwd(z, out2, lp)
    t_u = timed(unfused)
    bytes_alg = M * (ldx * 2 + P * 4)
    res = {'P': P, 'H': H, 'rows': M, 'out_info': info, 'fused_ms': t_f, 'unfused_ms': t_u,
           'fused_GBps': bytes_alg / t_f / 1e6, 'unfused_GBps_same_bytes': bytes_alg / t_u / 1e6,
           'max_abs_diff': (out - out2).abs().max().item()}
    print(json.dumps(res))
    plan.destroy()


if __name__ == '__main__':
    if len(sys.argv) > 1:
        run(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]))
    else:
        run(1944, 550, 131072)
        run(3480, 1024, 65536)
        run(48, 550, 131072)
