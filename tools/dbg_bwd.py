import torch, sys, ctypes
sys.path.insert(0, '/root/repo')
from distributed_dot_product import _native
import distributed_dot_product.ops.gemm as G
lib = _native.load()
lib.ddp_debug_timeout_site.restype = ctypes.c_int
torch.manual_seed(9)
dev = torch.device('cuda')
for shape in [(1, 1, 128, 64), (1, 1, 256, 64), (2, 3, 200, 64), (1, 2, 130, 40), (1, 16, 1024, 64)]:
    b, h, t, d = shape
    dtype = torch.bfloat16
    g = torch.randn(b, h, t, d, device=dev).to(dtype)
    v = torch.randn(b, h, t, d, device=dev).to(dtype)
    tpad = (t + 7) // 8 * 8
    probs = torch.zeros(b, h, t, tpad, device=dev, dtype=dtype)[..., :t]
    probs.copy_(torch.softmax(torch.randn(b, h, t, t, device=dev), -1))
    delta = torch.randn(b, h, t, device=dev)
    got = G.gemm(_native.OP_NT, g, v, softmax_bwd=(probs, delta, 0.3))
    torch.cuda.synchronize()
    site = lib.ddp_debug_timeout_site()
    want = 0.3 * probs.float() * (g.float() @ v.float().transpose(-1, -2) - delta.unsqueeze(-1))
    print(shape, 'timeout site', site, 'max err', (got.float() - want).abs().max().item(), flush=True)
