"""uz_colsum_static: the allocation-free variant of uz_colsum (per-block partial sums in a file-scope device array inside
the library instead of a caller workspace - for callers under CUDA graph capture or in loops that may not cudaMalloc).
Called through the C ABI and checked against the existing uz_colsum entry point, two calls in a row (the second on
DIFFERENT data and a smaller shape, so a stale scratch row or a stale result would show).  Collected last on purpose
(the file name sorts after every other test file)."""
import pytest
import torch

from unet_zoo_b200 import _lib as L

pytestmark = pytest.mark.gpu


def _colsum_ws(lib, x, C, out, accumulate):
    need = lib.uz_colsum_workspace_bytes(C)
    ws = torch.empty(need, dtype=torch.uint8, device=x.device)
    L.check(lib.uz_colsum(x.data_ptr(), x.shape[-1], x.numel() // x.shape[-1], C, out.data_ptr(), accumulate,
                          ws.data_ptr(), need, None), "uz_colsum")


def _colsum_static(lib, x, C, out, accumulate):
    L.check(lib.uz_colsum_static(x.data_ptr(), x.shape[-1], x.numel() // x.shape[-1], C, out.data_ptr(), accumulate,
                                 None), "uz_colsum_static")


def test_colsum_static_matches_colsum_on_two_calls_in_a_row():
    lib = L.load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    g = torch.Generator(device="cpu").manual_seed(5)
    assert lib.uz_colsum_static_max_channels() == 1024
    # (pixels as N,H,W; row pitch ld; channels used).  First the widest supported slice on many rows, then a small one.
    cases = [((2, 96, 96), 1024, 1024, 0), ((1, 7, 5), 80, 64, 0), ((2, 33, 17), 256, 256, 1)]
    warm = torch.zeros(1, 2, 2, 64, dtype=torch.bfloat16, device=dev)      # first launch loads the kernels (lazy module
    _colsum_static(lib, warm, 64, torch.zeros(64, device=dev), 0)           # loading takes device memory once)
    _colsum_ws(lib, warm, 64, torch.zeros(64, device=dev), 0)
    for (n, h, w), ld, C, accumulate in cases:
        x = (torch.randn(n, h, w, ld, generator=g) + 0.25).to(torch.bfloat16).to(dev)
        ref = torch.full((C,), 3.0, device=dev)
        got = torch.full((C,), 3.0, device=dev)
        # everything the static call needs exists before it runs; it must not allocate device memory itself
        torch.cuda.synchronize()
        before = torch.cuda.mem_get_info()[0]
        _colsum_static(lib, x, C, got, accumulate)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == before
        _colsum_ws(lib, x, C, ref, accumulate)
        torch.cuda.synchronize()
        exact = x[..., :C].double().sum(dim=(0, 1, 2)) + (3.0 if accumulate else 0.0)
        # same kernels, same grid on a 148-SM part -> the same fp32 sums; 1e-5 relative leaves room for a clamped grid
        assert float((got.double() - ref.double()).abs().max() / ref.double().abs().max()) < 1e-5
        assert float((got.double() - exact).abs().max() / exact.abs().max()) < 1e-4


def test_colsum_static_rejects_what_the_scratch_cannot_hold():
    lib = L.load()
    dev = torch.device("cuda", 0)
    x = torch.zeros(1, 2, 2, 2048, dtype=torch.bfloat16, device=dev)
    out = torch.zeros(2048, device=dev)
    rc = lib.uz_colsum_static(x.data_ptr(), 2048, 4, 1032, out.data_ptr(), 0, None)
    assert rc == -4 and b"1024" in lib.uz_last_error()                  # UZ_ERR_WORKSPACE, names the limit
    assert lib.uz_colsum_static(x.data_ptr(), 2048, 4, 12, out.data_ptr(), 0, None) == -1     # C not a multiple of 8
    assert lib.uz_colsum_static(x.data_ptr(), 2048, 4, 64, None, 0, None) == -1               # no output
