"""cnx_optim_grad_norm_noalloc (include/cnx_b200.h): the gradient-norm launch with its per-chunk
partial sums in a file-scope `static __device__` array of the library instead of a caller buffer —
no counterpart in the reference; added for callers that cannot call cudaMalloc.

Checked through the C-ABI against the existing cnx_optim_grad_norm on the same gradients, twice in
a row (the second call on a DIFFERENT, smaller set of tensors, so stale contents of the shared array
must not leak in).  Both entry points launch the same kernel with the same fixed-order fold, so the
results are required to be bit-equal (the library's other optimizer tests allow 1e-6 relative
against torch; bit-equal is inside that).  Collected last on purpose (file name): a failure here
must not hide the tests before it under `pytest -x`."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _table(grads, chunk, dev):
    numel = np.array([g.numel() for g in grads], dtype=np.int64)
    prefix = np.zeros(len(grads) + 1, dtype=np.int64)
    np.cumsum((numel + chunk - 1) // chunk, out=prefix[1:])
    ptrs = np.array([g.data_ptr() for g in grads], dtype=np.int64)
    to = lambda a: torch.from_numpy(a).to(dev)
    return to(ptrs), to(numel), to(prefix), len(grads), int(prefix[-1])


def _both(lib_call, grads, scale, max_norm, dev, step0):
    from image_classification_b200 import _lib
    chunk = _lib.load().cnx_ema_chunk_elems()
    ptrs, numel, prefix, T, total = _table(grads, chunk, dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    partials = torch.full((total,), float("nan"), device=dev)
    st_a = torch.zeros(8, device=dev)
    st_b = torch.zeros(8, device=dev)
    st_a[3] = st_b[3] = step0
    sp = scale.data_ptr() if scale is not None else 0
    lib_call("cnx_optim_grad_norm", ptrs.data_ptr(), numel.data_ptr(), prefix.data_ptr(), T, total,
             sp, max_norm, partials.data_ptr(), st_a.data_ptr(), stream)
    lib_call("cnx_optim_grad_norm_noalloc", ptrs.data_ptr(), numel.data_ptr(), prefix.data_ptr(), T,
             total, sp, max_norm, st_b.data_ptr(), stream)
    torch.cuda.synchronize()
    return st_a.cpu(), st_b.cpu(), total


def test_grad_norm_noalloc_matches_grad_norm_twice_in_a_row(cnx_lib):
    from image_classification_b200 import _lib
    assert torch.cuda.is_available(), "GPU tests need a B200"
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(123)
    shapes1 = [(1024, 4096), (4096,), (1000, 37), (3,), (128, 7, 7), (8192,), (8193,)]
    grads1 = [torch.randn(*s, generator=g).to(dev) for s in shapes1]
    scale = torch.tensor(65536.0, device=dev)
    a, b, total1 = _both(_lib.call, [x * 65536.0 for x in grads1], scale, 1.0, dev, 4.0)
    assert total1 > 500
    assert torch.equal(a, b), (a, b)
    ref = float(torch.sqrt(sum((x.double() ** 2).sum() for x in grads1)))
    assert abs(float(b[0]) - ref) < 1e-5 * ref
    assert float(b[2]) == 0.0 and float(b[3]) == 5.0
    assert abs(float(b[5]) - min(1.0, 1.0 / (ref + 1e-6))) < 1e-6
    # second call right after: fewer chunks (the tail of the array still holds call 1's partials),
    # no loss scale, clipping off, and a non-finite gradient in one tensor for the flag
    grads2 = [torch.randn(77, 129, generator=g).to(dev), torch.randn(20000, generator=g).to(dev)]
    a, b, total2 = _both(_lib.call, grads2, None, 0.0, dev, 0.0)
    assert total2 < total1
    assert torch.equal(a, b), (a, b)
    ref2 = float(torch.sqrt(sum((x.double() ** 2).sum() for x in grads2)))
    assert abs(float(b[0]) - ref2) < 1e-5 * ref2 and float(b[5]) == 1.0 and float(b[3]) == 1.0
    grads2[1][12345] = float("inf")
    a, b, _ = _both(_lib.call, grads2, None, 1.0, dev, 2.0)
    assert float(a[2]) == 1.0 and float(b[2]) == 1.0 and float(b[3]) == 2.0   # step not counted
    # and once more on the first set: the array is reusable after a flagged call
    a, b, _ = _both(_lib.call, grads1, None, 1.0, dev, 0.0)
    assert torch.equal(a, b) and abs(float(b[0]) - ref) < 1e-5 * ref


def test_grad_norm_noalloc_rejects_sizes_beyond_the_static_scratch(cnx_lib):
    from image_classification_b200 import _lib
    assert torch.cuda.is_available(), "GPU tests need a B200"
    dev = torch.device("cuda:0")
    lib = _lib.load()
    limit = lib.cnx_optim_grad_norm_noalloc_max_chunks()
    assert limit == 65536
    x = torch.ones(16, device=dev)
    ptrs, numel, prefix, T, _ = _table([x], lib.cnx_ema_chunk_elems(), dev)
    state = torch.zeros(8, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    with pytest.raises(_lib.CnxError) as ei:
        # the size check is on the host and comes before any launch: the tables are never read
        _lib.call("cnx_optim_grad_norm_noalloc", ptrs.data_ptr(), numel.data_ptr(),
                  prefix.data_ptr(), T, limit + 1, 0, 1.0, state.data_ptr(), stream)
    assert "exceed the static scratch" in str(ei.value) and "65536" in str(ei.value)
    torch.cuda.synchronize()
    assert float(state.abs().sum()) == 0.0
