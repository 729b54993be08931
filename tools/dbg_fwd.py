import torch, sys
sys.path.insert(0, '/root/repo')
from distributed_dot_product.ops.attention import fused_softmax_nt
torch.manual_seed(0)
dev = torch.device('cuda')
dtype = torch.bfloat16
b, h, t, d = 1, 1, 256, 64
k = torch.randn(b, h, t, d, device=dev).to(dtype)
q = torch.randn(b, h, t, d, device=dev).to(dtype)
got = fused_softmax_nt(k, q, None, 0.2, 64).float()[0, 0]
want = torch.softmax((k.float() @ q.float().transpose(-1, -2)) * 0.2, -1)[0, 0]
err = (got - want).abs()
blk = err.view(8, 32, 8, 32).amax(dim=(1, 3))
print((blk * 1000).round().int())
rb, cb = [int(x) for x in (blk == blk.max()).nonzero()[0]]
print('worst block', rb, cb, 'got', got[rb*32, cb*32:cb*32+8].tolist(), 'want', want[rb*32, cb*32:cb*32+8].tolist())
g4 = got.view(8, 32, 8, 32).permute(0, 2, 1, 3)
w4 = want.view(8, 32, 8, 32).permute(0, 2, 1, 3)
for rb in range(8):
    for cb in range(8):
        if blk[rb, cb] > 0.01:
            # which rows inside the block are wrong, and does the block match another reference block?
            e = (g4[rb, cb] - w4[rb, cb]).abs().amax(dim=1)
            bad_rows = (e > 0.01).nonzero().flatten().tolist()
            best = min(((g4[rb, cb] - w4[rb, c2]).abs().max().item(), c2) for c2 in range(8))
            bad_cols = ((g4[rb, cb] - w4[rb, cb]).abs().amax(dim=0) > 0.01).nonzero().flatten().tolist()
            print('block', rb, cb, 'bad rows', bad_rows, 'bad cols', bad_cols, 'closest ref col-block', best)
