import sys; sys.path.insert(0,".")
import torch, numpy as np
import sparse_coding__b200 as sc
from sparse_coding__b200 import _native as nat
M,d,N,B=4,256,512,4096
torch.manual_seed(1)
models=[sc.FunctionalTiedSAE.init(d,N,float(l1),device="cuda") for l1 in np.logspace(-4,-2,M)]
ens=sc.FunctionalEnsemble(models, sc.FunctionalTiedSAE, sc.torchopt_compat.adam, {"lr":1e-3}, device="cuda")
x=torch.randn(4*B,d,device="cuda")
for i in range(10): ens.step_batch(x[(i%4)*B:(i%4+1)*B])
nat.profile_begin()
for i in range(20): ens.step_batch(x[(i%4)*B:(i%4+1)*B])
marks=nat.profile_end()
per={}
for n,ms in marks: per.setdefault(n,[]).append(ms)
print({k:round(float(np.mean(v))*1e3,1) for k,v in per.items()})
