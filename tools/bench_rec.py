"""Micro-benchmark of the recurrent kernels alone (CUDA events, 3 warm-ups): fp32 cluster kernels vs
tensor-core kernels at the bench shapes.  Usage: python tools/bench_rec.py [H] [T]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'pytorch-kaldi-cgs_b200'))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from b200 import native
from oracle import hcgs_oracle


def timed(fn, n=5):
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


def main():
    H = int(sys.argv[1]) if len(sys.argv) > 1 else 550
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    d = torch.device('cuda', 0)
    np.random.seed(3)
    mask = torch.from_numpy(hcgs_oracle.conn_mat(H, H, [32, 4], [75, 62.5]).astype(np.float32))
    plan = native.Plan(mask, recurrent=True)
    us = [(torch.randn(H, H) * 0.1 * mask).to(d) for _ in range(4)]
    print('H=%d T=%d fp32 plan: cluster %d, %d threads' % (H, T, plan.cluster_size, plan.threads))
    for B, cl, nb in [(128, 8, 16), (256, 8, 16), (16, 8, 16), (128, 4, 16), (256, 4, 16), (512, 4, 16), (1024, 4, 16), (16, 4, 16)]:
        try:
            cfg = native.TcConfig(plan, cl, nb)
        except native.NativeError as e:
            print('B=%d cl=%d nb=%d: %s' % (B, cl, nb, e))
            continue
        ldH = native.pad8(H)
        print('   clusters co-resident: fwd %d bwd %d, smem fwd %d bwd %d' % (cfg.max_clusters_fwd, cfg.max_clusters_bwd, cfg.smem_fwd, cfg.smem_bwd))
        pre = torch.randn(T * B, 4 * ldH, device=d)
        drop = torch.bernoulli(torch.full((B, H), 0.8, device=d))
        h = torch.empty(T * B, native.pad8(H), device=d)
        hb = torch.empty(T * B, ldH, device=d, dtype=torch.bfloat16)
        c = torch.empty(T * B, ldH, device=d)
        gates = torch.empty(T * B, 4 * ldH, device=d)
        dh = torch.randn(T * B, native.pad8(H), device=d)
        dpre = torch.empty(T * B, 4 * ldH, device=d)
        dpb = torch.empty(T * B, 4 * ldH, device=d, dtype=torch.bfloat16)
        fp = torch.empty(cfg.fwd_pack_u32, device=d, dtype=torch.int32)
        bp = torch.empty(max(cfg.bwd_pack_u32, 1), device=d, dtype=torch.int32)
        native.lstm_tc_pack(plan, cfg, us, fp, bp)
        tf = timed(lambda: native.lstm_tc_fwd(plan, cfg, fp, pre, drop, 1.0, 2, T, B, h, hb, c, gates))
        tfe = timed(lambda: native.lstm_tc_fwd(plan, cfg, fp, pre, None, 0.8, 2, T, B, None, hb))
        tb = timed(lambda: native.lstm_tc_bwd(plan, cfg, bp, dh, gates, c, drop, 1.0, 2, T, B, dpre, dpb))
        line = 'B=%4d cl=%2d nb=%2d | tc fwd %.3f ms (%.2f us/step, %.1f Mfr/s) eval %.3f ms | bwd %.3f ms (%.2f us/step)' % (
            B, cfg.cluster_size, cfg.nb, tf, tf * 1e3 / T, T * B / tf / 1e3, tfe, tb, tb * 1e3 / T)
        if B <= 256 and cl == 8 and nb == 16:
            f32 = torch.empty(plan.fwd_pack_floats, device=d)
            b32 = torch.empty(max(plan.bwd_pack_floats, 1), device=d)
            native.lstm_pack(plan, us, f32, b32)
            pre3 = torch.randn(T, B, 4 * H, device=d)
            h3, c3, g3 = torch.empty(T, B, H, device=d), torch.empty(T, B, H, device=d), torch.empty(T, B, 4 * H, device=d)
            dp3 = torch.empty(T, B, 4 * H, device=d)
            t0 = timed(lambda: native.lstm_fwd(plan, f32, pre3, drop, 1.0, 2, h3, c3, g3))
            t1 = timed(lambda: native.lstm_bwd(plan, b32, dh.view(T, B, H), g3, c3, drop, 1.0, 2, dp3))
            line += ' | fp32 fwd %.3f bwd %.3f ms' % (t0, t1)
        print(line, flush=True)


if __name__ == '__main__':
    main()
