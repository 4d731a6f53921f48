import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, 'tests')
import numpy as np, torch
from flexflow_b200 import ops
for (M,K,N) in [(16,1024,512),(16,2048,512),(16,2048,1536),(16,1024,1536),(16,4096,256)]:
    torch.manual_seed(0)
    x = torch.randn(M,K,device='cuda').half(); w=(torch.randn(N,K,device='cuda')*0.05).half()
    h = ops.FFHandler('cuda:0'); m = ops.LinearMeta(h,K,N,max_tokens_per_batch=16)
    out = torch.zeros(M,N,device='cuda',dtype=torch.float16)
    ops.Linear.forward_kernel_wrapper(m,x,out,w,None,K,N,M); torch.cuda.synchronize()
    ref = (x.float()@w.float().t())
    err = (out.float()-ref).abs()
    bad = (err > 0.05).nonzero()
    print(M,K,N,'maxerr',err.max().item(),'nbad',len(bad), 'bad cols', sorted(set((bad[:,1]//8).tolist()))[:12], 'bad rows', sorted(set(bad[:,0].tolist()))[:16])
    # partial-k hypothesis: compare against only first 1024 k
    if len(bad):
        for kb in range(K//1024):
            refk = x[:, :1024*(kb+1)].float()@w[:, :1024*(kb+1)].float().t()
            print('   vs first',kb+1,'kblocks:', (out.float()-refk).abs().max().item())
