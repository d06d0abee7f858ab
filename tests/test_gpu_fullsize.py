"""Full-size parity: one ConvNeXt block forward + backward at every stage shape of BASELINE config 2
(Base, B=128, 224x224), at the 384x384 shapes of config 4 and at the micro-batch-64 shape of
config 3, against the fp64 oracle evaluated on the same device (it needs seconds there, hours on
the host).  Same tolerances as the small-shape tests (bf16: 2e-2 of the tensor's max)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from test_gpu_block import BF16_REL, _bf, _dev, _make_pair, _rel  # noqa: E402  (tests/ is on sys.path)


FULL_SHAPES = [
    (128, 128, 56, 56), (128, 256, 28, 28), (128, 512, 14, 14), (128, 1024, 7, 7),   # cfg 2
    (64, 128, 96, 96), (64, 512, 24, 24), (64, 1024, 12, 12),                       # cfg 4 (384^2)
    (64, 512, 14, 14),                                                              # cfg 3, 8 GPUs
]


@pytest.mark.parametrize("shape", FULL_SHAPES)
def test_block_full_size_vs_fp64_oracle(cnx_lib, shape):
    B, C, H, W = shape
    dev = _dev()
    ref, blk = _make_pair(C, "o1", dev)
    torch.manual_seed(B + C + H)
    x32 = _bf(torch.randn(B, C, H, W, device=dev))
    gy = _bf(torch.randn(B, C, H, W, device=dev))
    xb = x32.bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_()
    yb = blk(xb)
    yb.backward(gy.bfloat16().contiguous(memory_format=torch.channels_last))
    torch.cuda.synchronize()
    from image_classification_b200 import _lib
    assert _lib.load().cnx_debug_timeout_flag() == 0

    ref = ref.double()
    xr = x32.double().requires_grad_()
    yr = ref(xr)
    yr.backward(gy.double())
    branch, branch_ref = yb.double() - x32.double(), (yr - xr).detach()
    assert _rel(branch, branch_ref) < BF16_REL + 2.0 ** -8 * float(yr.abs().max() / branch_ref.abs().max())
    assert _rel(yb, yr.detach()) < BF16_REL
    assert _rel(xb.grad, xr.grad) < BF16_REL
    for (n, p), (_, q) in zip(blk.named_parameters(), ref.named_parameters()):
        assert p.grad is not None and p.grad.dtype == torch.float32, n
        assert _rel(p.grad, q.grad) < BF16_REL, (n, _rel(p.grad, q.grad))
    del ref, xr, yr
    torch.cuda.empty_cache()
