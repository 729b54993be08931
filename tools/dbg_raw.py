import torch, sys, os
sys.path.insert(0, '/root/repo')
os.environ['DDP_GEMM_DBG'] = str(32 + 128)   # raw scores through the 8-warp apply kernel
from distributed_dot_product.ops.attention import fused_softmax_nt
torch.manual_seed(0)
dev = torch.device('cuda')
for d in (64, 128, 256, 512):
    for t in (256, 1024):
        k = torch.randn(1, 2, t, d, device=dev).bfloat16()
        q = torch.randn(1, 2, t, d, device=dev).bfloat16()
        bad = 0
        for rep in range(3):
            got = fused_softmax_nt(k, q, None, 0.2, 64).float()
            want = (k.float() @ q.float().transpose(-1, -2))
            err = (got - want).abs().view(2, t // 32, 32, t // 32, 32).amax(dim=(2, 4))
            bad += int((err > 0.5).sum())
        print('d', d, 't', t, 'bad 32x32 blocks over 3 reps:', bad, 'of', 3 * 2 * (t // 32) ** 2)
