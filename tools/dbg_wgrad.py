import sys, torch
sys.path.insert(0, "/root/repo")
from pointcloudsegmentation_b200._lib import lib, check
from pointcloudsegmentation_b200 import fused_ops
for (M, N, K) in [(64, 128, 32), (300, 32, 32), (5000, 128, 64)]:
    g = torch.randn(M, N, device="cuda"); x = torch.randn(M, K, device="cuda")
    S = int(lib.pcs_linear_tc_wgrad_partials(M, N, K, 0))
    part = torch.full((S, N, K), 7.0, device="cuda")
    check(lib.pcs_linear_tc_wgrad(g.data_ptr(), M, N, N, x.data_ptr(), K, K, 0, part.data_ptr(), fused_ops._stream()), "w")
    torch.cuda.synchronize()
    ref = g.double().t() @ x.double()
    out = part.sum(0)
    print(M, N, K, "S", S, "part absmax", float(part.abs().max()), "untouched(7.0) frac", float((part == 7.0).float().mean()))
    print(" out[0,:6]", out[0, :6].tolist())
    print(" ref[0,:6]", ref[0, :6].tolist())
    print(" out[:6,0]", out[:6, 0].tolist())
    print(" ref[:6,0]", ref[:6, 0].tolist())
