"""odex_bn_relu_fwd_static (include/odex.h): the BatchNorm + ReLU forward with its scratch in a `static __device__` array inside
libodex.so, for callers that cannot allocate.  No counterpart in the reference.
Called through the C-ABI and compared with the existing entry point odex_bn_relu_fwd (caller-provided scratch) at the
tolerance of tests/test_gpu_kernels.py::test_bn_relu_matches_torch.  Collected last on purpose (file name)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-6          # the forward tolerance of test_bn_relu_matches_torch


def rel_err(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _buffers(M, C, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(M, C, device="cuda", generator=g) * 3 + 10.0           # |mean| >> std
    gamma = torch.rand(C, device="cuda", generator=g) + 0.5
    beta = torch.randn(C, device="cuda", generator=g) * 0.2
    return x, gamma, beta


def _existing(lib, _lib, x, gamma, beta, eps, relu):
    M, C = x.shape
    y, stats = torch.empty_like(x), torch.empty(3, C, device="cuda")
    n = int(lib.odex_bn_scratch_floats(M, C))
    scratch = torch.empty(n, device="cuda")
    _lib.check(lib.odex_bn_relu_fwd(x.data_ptr(), M, C, gamma.data_ptr(), beta.data_ptr(), eps, relu, y.data_ptr(), stats.data_ptr(),
                                    scratch.data_ptr(), n, torch.cuda.current_stream().cuda_stream), "odex_bn_relu_fwd")
    return y, stats


def test_bn_relu_fwd_static_matches_the_existing_entry_on_two_calls_in_a_row():
    from ode_rl_b200 import _lib
    lib = _lib.lib()
    eps = 1e-5
    cap = int(lib.odex_bn_static_scratch_floats())
    assert cap == 2 * 592 * 128
    # call 1 fills the whole array (592 slabs x 128 channels); call 2 is smaller and of another shape, so what call 1 left behind
    # lies in and beyond its slots; call 3 repeats call 1's shape with other data
    cases = [(48000, 128, 1, 11), (6 * 16 * 24, 64, 1, 12), (48000, 128, 0, 13), (18944, 256, 1, 14)]
    prepared = []
    for M, C, relu, seed in cases:
        x, gamma, beta = _buffers(M, C, seed)
        y_ref, st_ref = _existing(lib, _lib, x, gamma, beta, eps, relu)
        prepared.append((x, gamma, beta, relu, y_ref, st_ref, torch.empty_like(x), torch.empty(3, C, device="cuda")))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    stream = torch.cuda.current_stream().cuda_stream
    for x, gamma, beta, relu, _, _, y, stats in prepared:            # back to back on one stream, nothing allocated in between
        M, C = x.shape
        _lib.check(lib.odex_bn_relu_fwd_static(x.data_ptr(), M, C, gamma.data_ptr(), beta.data_ptr(), eps, relu, y.data_ptr(),
                                               stats.data_ptr(), stream), "odex_bn_relu_fwd_static")
    assert torch.cuda.memory_allocated() == before
    torch.cuda.synchronize()
    for x, gamma, beta, relu, y_ref, st_ref, y, stats in prepared:
        assert torch.isfinite(y).all()
        assert rel_err(y, y_ref) < TOL, (tuple(x.shape), rel_err(y, y_ref))
        assert rel_err(stats, st_ref) < TOL, (tuple(x.shape), rel_err(stats, st_ref))
        if relu:
            assert float(y.min()) >= 0.0
    # and against torch, as the existing entry's own test does
    x, gamma, beta, relu, _, _, y, _ = prepared[1]
    want = torch.relu(torch.nn.functional.batch_norm(x, None, None, gamma, beta, True, 0.0, eps))
    assert rel_err(y, want) < TOL


def test_bn_relu_fwd_static_rejects_sizes_beyond_its_array():
    from ode_rl_b200 import _lib
    lib = _lib.lib()
    M, C = 18944 + 64, 256                      # 297 slabs x 256 channels x 2 > 151 552 floats
    x, gamma, beta = _buffers(M, C, 3)
    y, stats = torch.empty_like(x), torch.empty(3, C, device="cuda")
    rc = lib.odex_bn_relu_fwd_static(x.data_ptr(), M, C, gamma.data_ptr(), beta.data_ptr(), 1e-5, 1, y.data_ptr(), stats.data_ptr(),
                                     torch.cuda.current_stream().cuda_stream)
    assert rc == 4                              # ODEX_E_UNSUPPORTED
    msg = lib.odex_last_error().decode()
    assert "built-in array" in msg and "odex_bn_relu_fwd" in msg
    with pytest.raises(_lib.OdexError):
        _lib.check(rc, "odex_bn_relu_fwd_static")
    torch.cuda.synchronize()
