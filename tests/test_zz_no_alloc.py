"""srn_bn_train_forward_noalloc: the BatchNormalizationF16 forward for callers that cannot allocate GPU memory.

Some users of this library run inside a framework that has already handed the whole GPU to its own
allocator, or sit in a loop where the device-wide synchronisation of cudaMalloc/cudaFree is not
acceptable.  For them the library offers a variant whose working storage is a fixed array inside the
library image.  This test drives it through the C ABI on device buffers that the *test* allocates with
torch, twice in a row with different shapes (the second call sees the first call's leftovers in the
shared array), and checks both results against the ordinary entry point, which allocates its own scratch.
The file name keeps it last in collection order.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run_noalloc(lib, x, gamma, beta, mm, mv, stream=None):
    import torch
    dev = torch.device('cuda', 0)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.int16)).to(dev)
    d_x, d_g, d_b = t(x), t(gamma), t(beta)
    d_mov = t(np.concatenate([mm, mv]))                      # moving mean then moving variance, one array
    d_y = torch.empty_like(d_x)
    torch.cuda.synchronize()
    rows, ch = x.shape
    rc = lib.srn_bn_train_forward_noalloc(0, stream, d_x.data_ptr(), rows, ch, d_g.data_ptr(), d_b.data_ptr(),
                                          d_mov.data_ptr(), d_mov.data_ptr() + 2 * ch, d_y.data_ptr())
    assert rc == 0, lib.srn_last_error().decode()
    y = d_y.cpu().numpy().view(np.float16)
    mov = d_mov.cpu().numpy().view(np.float16)
    return y, mov[:ch], mov[ch:]


def test_noalloc_matches_allocating_entry_point_twice_in_a_row():
    from serann_b200.engine import bn_train_forward, load_library
    lib = load_library()
    r = np.random.default_rng(123)
    for rows, ch in [(6000, 48), (750, 7)]:                  # second call: smaller, reuses the dirty scratch
        x = (r.standard_normal((rows, ch)) * 1.5 + 0.5).astype(np.float16)
        gamma = (1 + 0.1 * r.standard_normal(ch)).astype(np.float16)
        beta = (0.1 * r.standard_normal(ch)).astype(np.float16)
        mm = (0.1 * r.standard_normal(ch)).astype(np.float16)
        mv = np.abs(1 + 0.1 * r.standard_normal(ch)).astype(np.float16)
        y_ref, mm_ref, mv_ref = bn_train_forward(x, gamma, beta, mm, mv)
        y, mm2, mv2 = _run_noalloc(lib, x, gamma, beta, mm, mv)
        # same kernels, same arithmetic: the tolerance of test_bn_train_forward_kernel (2e-3) is met exactly
        scale = float(np.abs(y_ref.astype(np.float32)).max())
        assert float(np.abs(y.astype(np.float32) - y_ref.astype(np.float32)).max()) <= 2e-3 * scale
        assert np.array_equal(y, y_ref)
        assert np.array_equal(mm2, mm_ref) and np.array_equal(mv2, mv_ref)
        # and against the CPU oracle directly, so that this file is a parity test in its own right
        import torch
        from oracle import serann_oracle as O
        t = lambda a: torch.tensor(a.astype(np.float32))
        tm, tv = t(mm), t(mv)
        ref = O._bn_train(t(x), t(gamma), t(beta), tm, tv).numpy()
        assert float(np.abs(y.astype(np.float32) - ref).max()) <= 2e-3 * float(np.abs(ref).max())
        assert np.allclose(mm2.astype(np.float32), tm.numpy(), rtol=2e-3, atol=1e-4)
        assert np.allclose(mv2.astype(np.float32), tv.numpy(), rtol=2e-3, atol=1e-4)


def test_noalloc_on_a_caller_stream_and_current_device_untouched():
    """A caller inside a framework passes the framework's own stream; the call must run there and must not
    change which device is current for the calling thread."""
    import torch
    from serann_b200.engine import bn_train_forward, load_library
    lib = load_library()
    r = np.random.default_rng(5)
    x = r.standard_normal((3000, 16)).astype(np.float16)
    gamma, beta = np.ones(16, np.float16), np.zeros(16, np.float16)
    mm, mv = np.zeros(16, np.float16), np.ones(16, np.float16)
    y_ref, mm_ref, mv_ref = bn_train_forward(x, gamma, beta, mm, mv)
    st = torch.cuda.Stream(device=0)
    before = torch.cuda.current_device()
    y, mm2, mv2 = _run_noalloc(lib, x, gamma, beta, mm, mv, stream=C.c_void_p(st.cuda_stream))
    assert torch.cuda.current_device() == before
    assert np.array_equal(y, y_ref) and np.array_equal(mm2, mm_ref) and np.array_equal(mv2, mv_ref)


def test_noalloc_refuses_a_misaligned_tensor_slice():
    """With C % 8 == 0 the kernels use 16-byte accesses: a view that starts at an odd element offset is turned
    down with SRN_ERR_ARG before any kernel runs (a fault would poison the caller's CUDA context)."""
    import torch
    from serann_b200.engine import load_library
    lib = load_library()
    buf = torch.zeros(4096, dtype=torch.int16, device='cuda')
    par = torch.zeros(64, dtype=torch.int16, device='cuda')
    p = buf.data_ptr()
    rc = lib.srn_bn_train_forward_noalloc(0, None, p + 2, 64, 16, par.data_ptr(), par.data_ptr(), par.data_ptr(),
                                          par.data_ptr() + 32, p + 2048)
    assert rc == -1 and b'16-byte aligned' in lib.srn_last_error()
    # C not a multiple of 8 runs the scalar path: the same offset is fine there
    rc = lib.srn_bn_train_forward_noalloc(0, None, p + 2, 64, 7, par.data_ptr(), par.data_ptr(), par.data_ptr(),
                                          par.data_ptr() + 14, p + 2048)
    assert rc == 0, lib.srn_last_error().decode()
    torch.cuda.synchronize()


def test_noalloc_rejects_sizes_beyond_the_builtin_scratch():
    import torch
    from serann_b200.engine import load_library
    lib = load_library()
    buf = torch.zeros(16, dtype=torch.int16, device='cuda')
    p = buf.data_ptr()
    # 2^24 rows x 64 channels would need 64 * (4 + 2 * 16384) floats > the 4 MiB array
    rc = lib.srn_bn_train_forward_noalloc(0, None, p, 1 << 24, 64, p, p, p, p + 128, p)
    assert rc == -5
    assert b'exceeds the built-in scratch' in lib.srn_last_error()
