#!/usr/bin/env python
"""A/B: fp32 GEMMs of the S3DIS classifier shapes on stock cuBLAS 12.8 (torch's) vs system cuBLAS 12.9 with
BF16x9 fp32 emulation (CUBLAS_EMULATE_SINGLE_PRECISION=1).  Development aid."""
import ctypes
import os
import subprocess
import sys

if len(sys.argv) > 1 and sys.argv[1] == "child":
    if os.environ.get("PCS_PRELOAD_CUBLAS"):
        ctypes.CDLL("/usr/local/cuda/lib64/libcublasLt.so.12", mode=ctypes.RTLD_GLOBAL)
        ctypes.CDLL("/usr/local/cuda/lib64/libcublas.so.12", mode=ctypes.RTLD_GLOBAL)
    import torch

    print("cublas", torch.backends.cuda.preferred_blas_library(), "ver", torch._C._cuda_getCompiledVersion() if hasattr(torch._C, "_cuda_getCompiledVersion") else "?")
    with open("/proc/self/maps") as f:
        print(sorted({l.split()[-1] for l in f if "cublas" in l}))
    M = 133399
    shapes = [("fwd  X[M,643] W^T[643,256]", (M, 643), (643, 256), False),
              ("dX   dY[M,256] W[256,643]", (M, 256), (256, 643), False),
              ("dW   dY^T[256,M] X[M,643]", (M, 256), (M, 643), True),
              ("fwd  X[M,131] W^T[131,512]", (M, 131), (131, 512), False),
              ("dW   dY^T[512,M] X[M,131]", (M, 512), (M, 131), True),
              ("fwd  X[M,387] W^T[387,13]", (M, 387), (387, 13), False),
              ("pab  X[M,35]  W[35,128]", (M, 35), (35, 128), False)]
    g = torch.Generator(device="cuda").manual_seed(0)
    for name, sa, sb, ta in shapes:
        a = torch.randn(*sa, device="cuda", generator=g)
        b = torch.randn(*sb, device="cuda", generator=g)
        f = (lambda: a.t() @ b) if ta else (lambda: a @ b)
        for _ in range(3):
            y = f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            y = f()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        if ta:
            ref = a[:, :64].double().t() @ b.double()
            err = float((y[:64].double() - ref).abs().max() / ref.abs().max())
        else:
            ref = a[:4096].double() @ b.double()
            err = float((y[:4096].double() - ref).abs().max() / ref.abs().max())
        fl = 2.0 * sa[0] * sa[1] * (sb[1])
        print(f"{name:34s} {ms*1e3:8.1f} us  {fl/ms/1e9:7.1f} TFLOP/s  max rel err {err:.2e}")
    sys.exit(0)

for tag, env in (("stock cuBLAS (torch bundled), fp32", {}),
                 ("system cuBLAS 12.9, fp32", {"PCS_PRELOAD_CUBLAS": "1"}),
                 ("system cuBLAS 12.9, BF16x9 emulation", {"PCS_PRELOAD_CUBLAS": "1", "CUBLAS_EMULATE_SINGLE_PRECISION": "1"}),
                 ("system cuBLAS 12.9, BF16x9 eager", {"PCS_PRELOAD_CUBLAS": "1", "CUBLAS_EMULATE_SINGLE_PRECISION": "1",
                                                       "CUBLAS_EMULATION_STRATEGY": "eager"})):
    print("==", tag, flush=True)
    subprocess.run([sys.executable, __file__, "child"], env={**os.environ, **env})
