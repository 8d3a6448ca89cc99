import torch
p=torch.nn.Parameter(torch.randn(4,device="cuda"))
opt=torch.optim.Adam([p],lr=0.1,fused=True)
(p*p).sum().backward(); print("before", p._version, p.data_ptr()); opt.step(); print("after fused adam", p._version, p.data_ptr())
