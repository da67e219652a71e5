"""What a SIMT (non-tensor-core) fp32 GEMM reaches on this GPU: torch.matmul with TF32 off at the C2 shape.
The dense-tile kernels (sddmm_tiled2, spmm_staged at density 0.5) do the same arithmetic."""
import torch
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
torch.set_float32_matmul_precision("highest")
dev = torch.device("cuda:0")
for n in (4096, 8192):
    a = torch.randn(n, n, device=dev)
    b = torch.randn(n, n, device=dev)
    for name, fn in (("A @ B", lambda: a @ b), ("A @ B^T", lambda: a @ b.t())):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"{n}^3 {name}: {ms:.3f} ms  {2 * n ** 3 / ms / 1e9:.1f} TFLOP/s")
