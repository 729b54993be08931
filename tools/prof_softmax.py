import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_dot_product.ops.softmax import softmax_forward
x = torch.randn(1, 16, 512, 65536, device='cuda', dtype=torch.bfloat16)
mask = torch.zeros(1, 512, 65536, device='cuda').bool()
for _ in range(3):
    softmax_forward(x, mask, 0.125, inplace=True)
torch.cuda.synchronize()
print('ok')
