import sys, torch
sys.path.insert(0, ".")
from instancesegmentation_b200.dwconv import ConvTranspose2dK2S
torch.manual_seed(0)
m = ConvTranspose2dK2S(16, 4, 8, stride=4, padding=2).cuda().train()
x = torch.randn(2, 16, 9, 13, device="cuda", requires_grad=True)
y = m(x); torch.cuda.synchronize(); print("fwd ok", y.shape)
ref = torch.nn.functional.conv_transpose2d(x, m.weight, m.bias, 4, 2)
print("fwd err", float((y - ref).abs().max()))
y.sum().backward(); torch.cuda.synchronize(); print("bwd ok", float(x.grad.abs().max()), float(m.weight.grad.abs().max()))
