for f in 64; do
HCGS_TC_DEBUG_OVERRIDE=$f timeout 120 python - <<PY
import os, sys
sys.path.insert(0,'pytorch-kaldi-cgs_b200'); sys.path.insert(0,'.'); sys.path.insert(0,'tools')
import numpy as np, torch
from b200 import native
from oracle import hcgs_oracle
from bench_rec import timed
H,T=550,256
d=torch.device('cuda',0)
np.random.seed(3)
mask=torch.from_numpy(hcgs_oracle.conn_mat(H,H,[32,4],[75,62.5]).astype(np.float32))
plan=native.Plan(mask,recurrent=True)
us=[(torch.randn(H,H)*0.1*mask).to(d) for _ in range(4)]
flag=int(os.environ['HCGS_TC_DEBUG_OVERRIDE'])
for B,cl in [(512,4),(128,8)]:
    cfg=native.TcConfig(plan,cl,16); ldH=native.pad8(H)
    pre=torch.randn(T*B,4*ldH,device=d); h=torch.empty(T*B,ldH,device=d); c=torch.empty(T*B,ldH,device=d); g=torch.empty(T*B,4*ldH,device=d)
    hb=torch.empty(T*B,ldH,device=d,dtype=torch.bfloat16)
    fp=torch.empty(cfg.fwd_pack_u32,device=d,dtype=torch.int32); native.lstm_tc_pack(plan,cfg,us,fp,None)
    os.environ['HCGS_TC_DEBUG']='0'
    native.lstm_tc_fwd(plan,cfg,fp,pre,None,1.0,2,T,B,h,hb,c,g); ref=h.clone()
    os.environ['HCGS_TC_DEBUG']=str(flag & ~64)
    native.lstm_tc_fwd(plan,cfg,fp,pre,None,1.0,2,T,B,h,hb,c,g); torch.cuda.synchronize()
    print('flag',flag,'B',B,'cl',cl,'equal',torch.equal(h,ref))
    t=timed(lambda: native.lstm_tc_fwd(plan,cfg,fp,pre,None,1.0,2,T,B,h,hb,c,g))
    print('   %.3f ms %.2f us/step'%(t,t*1e3/T))
    os.environ['HCGS_TC_DEBUG']=str(flag)
    native.lstm_tc_fwd(plan,cfg,fp,pre,None,1.0,2,T,B,h,hb,c,g); torch.cuda.synchronize()
    tr=native.tc_trace_read()
    names=['top','prewait','mbarwait','mma','B1','issue','gate','B2','push','-','-','stores']
    print('   '+' '.join('%s=%d'%(n,v/T) for n,v in zip(names,tr) if n!='-'))
PY
done
