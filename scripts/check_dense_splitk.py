"""Dense forward with a very long reduction and few output tiles (routed through the split-K weight-gradient kernel)
against torch.matmul: numerics (TF32 bound) and time."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dcnn_b200 import ops  # noqa: E402

torch.backends.cuda.matmul.allow_tf32 = False
for M, N, K in ((8, 12, 65536), (64, 128, 2097152)):
    x = torch.randn(M, K, device="cuda")
    w = torch.randn(N, K, device="cuda") / K ** 0.5
    b = torch.randn(N, device="cuda")
    y = ops.dense_forward(x, w, b, relu=True)
    ref = torch.relu(x @ w.t() + b)
    err = float((y - ref).abs().max() / ref.abs().max())
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(5):
        ops.dense_forward(x, w, None)
    e1.record()
    e1.synchronize()
    print(f"{M} x {N} x {K}: rel-max error {err:.2e}  {e0.elapsed_time(e1) / 5:.3f} ms", flush=True)
    assert err < 4e-3, err
print("ok")
