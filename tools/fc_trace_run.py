import os, sys, torch
sys.path.insert(0, os.getcwd())
import b200ctc
T,B,H,C=1500,64,800,30
g=torch.Generator().manual_seed(1)
x=torch.randn(T,B,H,generator=g).cuda(); w=(torch.rand(C,H,generator=g)/H**.5).cuda()
for i in range(3):
    b200ctc.fc.fc_forward(x,w)
torch.cuda.synchronize()
