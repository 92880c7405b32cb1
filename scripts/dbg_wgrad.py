import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200 import ops
from dcnn_b200.ops import cpu as ref
torch.manual_seed(0)
DEV = "cuda"
def cl(t): return t.to(DEV).contiguous(memory_format=torch.channels_last)
def run(tag, N, C, H, OC, k, s, p, structured=False):
    oh = (H + 2*p - k)//s + 1
    if structured:
        x = torch.zeros(N, C, H, H); dy = torch.zeros(N, OC, oh, oh)
        x[0, :, 0, 0] = torch.arange(1, C+1).float()
        dy[0, :, 0, 0] = torch.arange(1, OC+1).float() * 100
    else:
        x = torch.randn(N, C, H, H); dy = torch.randn(N, OC, oh, oh)
    dw_ref = torch.zeros(OC, C, k, k); ref.conv2d_backward_weight(dy, x, dw_ref, (s, s), (p, p))
    for init in (0.0, 1.0):
        dw = cl(torch.full((OC, C, k, k), init))
        ops.conv2d_backward_weight(cl(dy), cl(x), dw, (s, s), (p, p))
        torch.cuda.synchronize()
        d = dw.cpu() - init
        err = float((d - dw_ref).abs().max() / (dw_ref.abs().max() + 1e-12))
        print(f"[{tag}] init={init} relerr={err:.3e} got[min,max,absmean]=({float(d.min()):.4g},{float(d.max()):.4g},{float(d.abs().mean()):.4g}) "
              f"ref absmean={float(dw_ref.abs().mean()):.4g} nonzero={int((d != 0).sum())}/{d.numel()}")
    if structured:
        print("got[:4,:8,0,0]\n", d[:4, :8, 0, 0]); print("ref[:4,:8,0,0]\n", dw_ref[:4, :8, 0, 0])
        nz = (d[:, :, 0, 0] != 0).nonzero()
        print("nonzero idx sample", nz[:10].tolist())
print("DBG", os.environ.get("DCNN_WGRAD_DBG"))
run("1x1 struct", 1, 64, 8, 128, 1, 1, 0, True)
run("1x1 rand", 2, 64, 8, 128, 1, 1, 0)
run("3x3 rand", 4, 64, 16, 128, 3, 1, 1)
