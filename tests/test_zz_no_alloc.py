"""GPU test of tnt_dense_forward_noalloc (include/tnt.h), the entry point that allocates nothing: its scratch is a static
__device__ array inside libtnt.so.  No counterpart in the reference, so
the check is against the existing entry point tnt_dense_forward(training = 0) through the same C ABI -- same kernels and
summation order, so the features must be bit-identical (the dense-encoder tests' own tolerance is rel 1e-4) -- on two
calls in a row with different inputs and batch sizes (the array keeps the previous call's contents), and against the
oracle.  Named test_zz_* so that pytest collects it last."""
import ctypes
import numpy as np
import pytest
import torch

from oracle import dense_encoder as D
from tests.util import pkg, rel_err

pytestmark = pytest.mark.gpu


def _encoder(K, E, seed):
    rng = np.random.default_rng(seed)
    P = D.init_params(K, E, seed=seed)
    P["b"] = (0.05 * rng.standard_normal(E)).astype(np.float32)
    P["gamma"] = (1.0 + 0.1 * rng.standard_normal(E)).astype(np.float32)
    P["beta"] = (0.1 * rng.standard_normal(E)).astype(np.float32)
    P["mean"] = (0.1 * rng.standard_normal(E)).astype(np.float32)
    P["var"] = (1.0 + 0.2 * rng.random(E)).astype(np.float32)
    enc = pkg.DenseEncoder(K, E, 0.0, 0.2, 0.01, batch_max=64)
    enc.compile(pkg.Adam(1e-4, beta_1=0.9, beta_2=0.98, epsilon=10.0e-9, clipnorm=0.1))
    enc.set_weights([P["W"], P["b"], P["gamma"], P["beta"], P["mean"], P["var"]])
    return enc, P, rng


def _call_both(enc, x):
    """(existing entry point, no-alloc entry point) on the same device input, both through the C ABI."""
    lib, ctx = enc._libh, enc._ctx
    B = x.shape[0]
    st = torch.cuda.current_stream().cuda_stream
    ref = torch.full((B, enc.E), float("nan"), device="cuda")
    got = torch.full((B, enc.E), float("nan"), device="cuda")
    assert lib.tnt_dense_forward(ctx, x.data_ptr(), B, 0, 0, ref.data_ptr(), st) == 0
    rc = lib.tnt_dense_forward_noalloc(ctx, x.data_ptr(), B, got.data_ptr(), st)
    assert rc == 0, lib.tnt_dense_last_error(ctx)
    torch.cuda.synchronize()
    return ref.cpu().numpy(), got.cpu().numpy()


@pytest.mark.parametrize("K,E", [(333, 128), (62756, 512)])        # ragged tiny case; NIC.py:27 visual-cortex size
def test_noalloc_matches_existing_entry_point_two_calls_in_a_row(K, E):
    enc, P, rng = _encoder(K, E, seed=11)
    Pt = {k: torch.as_tensor(v) for k, v in P.items()}
    for B in (64, 5):                 # the second call is smaller: stale rows of the first stay in the shared scratch
        x = torch.from_numpy(rng.standard_normal((B, K)).astype(np.float32)).cuda()
        ref, got = _call_both(enc, x)
        assert np.isfinite(got).all()
        assert np.array_equal(ref, got), ("not bit-identical", K, E, B, np.abs(ref - got).max())
        f_o = D.forward(Pt, x.cpu(), False).numpy().reshape(B, E)
        assert rel_err(got, f_o) < 1e-4, (K, E, B, rel_err(got, f_o))


def test_noalloc_allocates_nothing_and_rejects_sizes_beyond_the_static_scratch():
    enc, P, rng = _encoder(4096, 128, seed=13)
    x = torch.from_numpy(rng.standard_normal((32, 4096)).astype(np.float32)).cuda()
    out = torch.empty(32, 128, device="cuda")
    enc.forward_noalloc(x, out)       # warm (module load, first-use set-up of the kernels)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    for _ in range(3):
        enc.forward_noalloc(x, out)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    # torch's allocator untouched, and the driver's free memory unchanged up to noise from outside this process (the
    # buffers the allocating entry point owns for this size are 1 MiB + 14.5 MiB: a cudaMalloc on the call would show)
    assert torch.cuda.memory_reserved() == reserved0 and free0 - free1 < (4 << 20), (free0, free1)
    ref, got = _call_both(enc, x)
    assert np.array_equal(ref, got) and np.array_equal(got, out.cpu().numpy())
    # beyond the stated maximum (K <= 65536): a clear error, nothing launched
    big = pkg.DenseEncoder(65537, 128, 0.0, 0.0, 0.0, batch_max=4)
    big.compile(pkg.Adam(1e-4))
    xb = torch.zeros(2, 65537, device="cuda")
    ob = torch.empty(2, 128, device="cuda")
    n0 = big.launch_count()
    rc = big._libh.tnt_dense_forward_noalloc(big._ctx, xb.data_ptr(), 2, ob.data_ptr(), None)
    assert rc != 0 and big.launch_count() == n0
    msg = big._libh.tnt_dense_last_error(big._ctx)
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "65536" in msg and "tnt_dense_forward" in msg, msg
