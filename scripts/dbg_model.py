import sys, torch
sys.path.insert(0, ".")
from instancesegmentation_b200 import Segment
torch.manual_seed(3)
m = Segment(3).cuda().train()
x = torch.randn(2, 3, 128, 128, device="cuda")
y = m(x); y.sum().backward(); torch.cuda.synchronize(); print("ok")
