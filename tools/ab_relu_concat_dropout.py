"""A/B on one B200: fused relu+concat+dropout (csrc/relu_concat_dropout.cu) vs the stock chain, level-0 classifier
shapes of a 64-block step.  usage: python tools/ab_relu_concat_dropout.py [M]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.nn.functional as F

from pointcloudsegmentation_b200 import fused_ops


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / n


M = int(sys.argv[1]) if len(sys.argv) > 1 else 515000
for Na, Nb in ((512, 131), (256, 131)):
    a = torch.randn(M, Na, device="cuda")
    b = torch.randn(M, Nb, device="cuda")
    g = torch.randn(M, Na + Nb, device="cuda")
    for keep in (0.7, 1.0):
        out = fused_ops.relu_concat_dropout_forward(a, b, keep, 5)
        tf = timeit(lambda: fused_ops.relu_concat_dropout_forward(a, b, keep, 5))
        tb = timeit(lambda: fused_ops.relu_concat_dropout_backward(g, out, Na, keep, 5))
        ts = timeit(lambda: F.dropout(torch.cat([F.relu(a), b], 1), 1 - keep, True))
        fb = 4.0 * M * (Na + Nb) * 2
        bb = 4.0 * M * (3 * Na + 2 * Nb)
        print(f"M={M} Na={Na} Nb={Nb} keep={keep}: fused fwd {tf:7.1f} us ({fb / tf / 1e6:6.2f} TB/s)  bwd {tb:7.1f} us ({bb / tb / 1e6:6.2f} TB/s)"
              f"  stock fwd chain {ts:7.1f} us")
