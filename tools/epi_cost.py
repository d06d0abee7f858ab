"""How much of the fused GEMMs' time is epilogue? Same shape, plain bf16 store vs fused epilogue."""
import os, sys, json, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import ops
from tools.microbench import timeit
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for s, (C, H) in enumerate([(128, 56), (256, 28), (512, 14), (1024, 7)]):
    M = 128 * H * H
    x = torch.randn(M, C, device=dev).bfloat16()
    w1 = (torch.randn(4 * C, C, device=dev) * 0.05).bfloat16()
    b1 = torch.randn(4 * C, device=dev)
    z = torch.randn(M, 4 * C, device=dev).bfloat16()
    cs = torch.zeros(4 * C, device=dev)
    r = dict(stage=s + 1)
    r["plain_N4C"] = round(timeit(lambda: ops.gemm_bf16(x, w1), flush=flush) * 1e3, 1)
    r["bias_gelu"] = round(timeit(lambda: ops.gemm_bf16(x, w1, epilogue=ops.EPI_BIAS_GELU, bias=b1), flush=flush) * 1e3, 1)
    r["gelu_grad"] = round(timeit(lambda: ops.gemm_bf16(x, w1, epilogue=ops.EPI_GELU_GRAD, R=z, colsum=cs), flush=flush) * 1e3, 1)
    r["gelu_grad_nocolsum"] = round(timeit(lambda: ops.gemm_bf16(x, w1, epilogue=ops.EPI_GELU_GRAD, R=z), flush=flush) * 1e3, 1)
    r["cublas"] = round(timeit(lambda: torch.matmul(x, w1.t()), flush=flush) * 1e3, 1)
    print(json.dumps(r), flush=True)
