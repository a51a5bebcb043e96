import copy, sys, torch
sys.path.insert(0, ".")
from instancesegmentation_b200.segment import _Unit
torch.backends.cudnn.allow_tf32 = False
torch.manual_seed(7)
u = _Unit(12, 24, 3, act=None)
with torch.no_grad():
    u.conv.bias.normal_(0, 2.0); u.bn.weight.uniform_(0.5, 1.5); u.bn.bias.normal_()
ref = copy.deepcopy(u).double().train()
x = torch.randn(4, 12, 24, 40); dy = torch.randn(4, 24, 24, 40)
xr = x.double().requires_grad_(True); yr = ref(xr); yr.backward(dy.double())
g = u.cuda().train(); xg = x.cuda().requires_grad_(True); yg = g(xg); yg.backward(dy.cuda())
d = (yg.double().cpu() - yr.detach()).abs()
print("y max err", float(d.max()), "at", torch.nonzero(d == d.max())[0].tolist(), "count>1e-3", int((d > 1e-3).sum()))
idx = torch.nonzero(d > 1e-3)
print(idx[:20].tolist())
d2 = (xg.grad.double().cpu() - xr.grad).abs(); print("dx max err", float(d2.max()))
print("rm err", float((g.bn.running_mean.double().cpu() - ref.bn.running_mean).abs().max()), "rv err", float((g.bn.running_var.double().cpu() - ref.bn.running_var).abs().max()))
