import sys, torch
sys.path.insert(0, ".")
from instancesegmentation_b200 import Segment
from instancesegmentation_b200.dwconv import DepthwiseConv2d
torch.manual_seed(3)
m = Segment(3).cuda().train()
shapes = {}
hooks = [mod.register_forward_hook(lambda mod, inp, out, n=n: shapes.__setitem__(n, tuple(inp[0].shape)))
         for n, mod in m.named_modules() if isinstance(mod, DepthwiseConv2d)]
import os
os.environ["ISG_DW_LIBRARY"] = "1"
with torch.no_grad():
    m(torch.randn(2, 3, 128, 128, device="cuda"))
del os.environ["ISG_DW_LIBRARY"]
for h in hooks: h.remove()
mods = dict(m.named_modules())
for n, shp in shapes.items():
    mod = mods[n]
    x = torch.randn(shp, device="cuda", requires_grad=True)
    print(n, shp, mod.kernel_size, mod.padding, mod.dilation, mod.bias is not None, flush=True)
    y = mod(x); torch.cuda.synchronize()
    y.sum().backward(); torch.cuda.synchronize()
print("all ok")
