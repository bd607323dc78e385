"""Diagnostic: host enqueue time vs wall time per training view, and a cProfile of the host side."""
import sys, time, torch, cProfile, pstats
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussctrl_exp_b200 import scene
from gaussctrl_exp_b200.model import GaussianModel
dev=torch.device('cuda:0')
NG=int(os.environ.get('NG','1000000'))
m=GaussianModel(scene.make_scene(NG,0),background_color=torch.tensor([0.1,0.2,0.3])).to(dev); m.train()
cams=scene.make_cameras(40,0); v=torch.randn(512,512,3,device=dev)/(512*512)
params=list(m.parameters())
def step(i):
    for p in params: p.grad=None
    out=m(cams[i%40]); out['rgb'].backward(v)
for i in range(20): step(i)
torch.cuda.synchronize()
t=time.perf_counter()
for i in range(100): step(i)
h=time.perf_counter()-t; torch.cuda.synchronize(); w=time.perf_counter()-t
print('host enqueue ms/step',h*10,'wall ms/step',w*10)
pr=cProfile.Profile(); pr.enable()
for i in range(100): step(i)
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats('cumulative').print_stats(25)
