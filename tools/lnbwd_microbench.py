"""Per-stage cost of the two LayerNorm-backward routes (CUDA events, L2 flushed):
old = dz GEMM, dxhat GEMM, row kernel + conv-bwd;  new = dz GEMM with row sums, dconv GEMM, conv-bwd."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_classification_b200 import ops
from tools.microbench import timeit
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
for C, H in ((128, 56), (256, 28), (512, 14), (1024, 7)):
    M = B * H * H
    torch.manual_seed(C)
    params = [torch.randn(C, 1, 7, 7, device=dev) * 0.1, torch.rand(C, device=dev) + 0.5, torch.randn(C, device=dev) * 0.1,
              torch.randn(4 * C, C, device=dev) * 0.05, torch.randn(4 * C, device=dev) * 0.1,
              torch.randn(C, 4 * C, device=dev) * 0.05, torch.randn(C, device=dev) * 0.1, torch.rand(C, device=dev)]
    fw = ops.prep_block_weights(*params)
    x = torch.randn(B, H, H, C, device=dev).bfloat16()
    xhat, mean, rstd = ops.dwconv_ln_fwd(x, fw.wt_dw, params[2])
    z, h = ops.fc1_gelu_fwd(xhat.view(M, C), fw.w1f, fw.b1f)
    dy = torch.randn(M, C, device=dev).bfloat16()
    dz, _ = ops.mlp_bwd_dz(dy, fw.w2fT, z)
    _, _, stats = ops.mlp_bwd_dz_stats(dy, fw.w2fT, z, fw.rb)
    dxhat = ops.mlp_bwd_dxhat(dz, fw.w1fT)
    dconv = ops.mlp_bwd_dconv(dz, fw.w1fT, xhat.view(M, C), stats, rstd)
    t = lambda fn: round(timeit(fn, flush=flush) * 1e3, 1)
    row = {"C": C, "M": M,
           "dz": t(lambda: ops.mlp_bwd_dz(dy, fw.w2fT, z)),
           "dz_stats(+fill)": t(lambda: ops.mlp_bwd_dz_stats(dy, fw.w2fT, z, fw.rb)),
           "dxhat": t(lambda: ops.mlp_bwd_dxhat(dz, fw.w1fT)),
           "dconv_gemm": t(lambda: ops.mlp_bwd_dconv(dz, fw.w1fT, xhat.view(M, C), stats, rstd)),
           "ln_bwd_rows+conv_bwd": t(lambda: ops.dwconv_ln_bwd(dxhat.view(B, H, H, C), xhat, rstd, x, dy.view(B, H, H, C), fw.wt_dw)),
           "conv_bwd": t(lambda: ops.dwconv_bwd(dconv.view(B, H, H, C), x, dy.view(B, H, H, C), fw.wt_dw))}
    row["old_total"] = round(row["dz"] + row["dxhat"] + row["ln_bwd_rows+conv_bwd"], 1)
    row["new_total"] = round(row["dz_stats(+fill)"] + row["dconv_gemm"] + row["conv_bwd"], 1)
    print(json.dumps(row), flush=True)
