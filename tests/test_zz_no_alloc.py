"""The allocation-free FFT entry point (sgz_fft_forward_noalloc) against the handle-based one.

Some callers of this library may not allocate GPU memory at the moment they call it: a real-time
audio or render callback (cudaMalloc can take a driver-wide lock and stall the device), or a host
application that pre-allocates everything and forbids later allocation.  For them the library
carries a fixed working buffer inside its own image.  This test checks, through the C ABI, that the
entry point that uses that buffer produces what sgz_spectrum_transform produces, on two calls in a
row (the second call sees the first call's leftovers in the shared buffer and must not care), and
that it refuses sizes the fixed buffer cannot hold.

Collected last on purpose (file name), so a run that stops at the first failure has already run
every other test; it does not depend on any of them and also passes alone in a fresh process."""
import numpy as np
import pytest

from signalizer_b200 import _abi, _lib
from signalizer_b200.spectrum import TransformPairBatch, fft_forward_noalloc

pytestmark = pytest.mark.gpu


def reference_fft(x):
    """x [batch, n, 2] float32 -> raw FFT through the existing handle entry points
    (Complex configuration + rectangular window = the transform of re + i im as is)."""
    batch, n, _ = x.shape
    cfg = _abi.default_config(window_size=n, window_type=_abi.WIN_RECTANGULAR, configuration=_abi.CH_COMPLEX, axis_points=16)
    h = TransformPairBatch(cfg, batch)
    assert h.N == n
    assert h.prepare_transform(x)
    h.do_transform()
    return h.get_raw_fft()


def close_enough(got, want):
    # same bound as the raw-FFT parity tests in test_spectrum_gpu.py: 2e-6 of the value plus 2e-6 of
    # the largest bin of the sequence
    g = got[..., 0].astype(np.float64) + 1j * got[..., 1]
    w = want[..., 0].astype(np.float64) + 1j * want[..., 1]
    peak = np.abs(w).max(axis=1, keepdims=True)
    assert np.all(np.abs(g - w) <= 2e-6 * np.abs(w) + 2e-6 * peak), float(np.max(np.abs(g - w) / peak))


@pytest.mark.parametrize("n,batch", [(4096, 8), (65536, 2), (2, 3), (1 << 20, 1)])
def test_noalloc_fft_matches_handle_fft_twice_in_a_row(n, batch):
    import torch
    rng = np.random.default_rng(n + batch)
    d_in = torch.empty((batch, n, 2), dtype=torch.float32, device="cuda")
    d_out = torch.empty_like(d_in)
    stream = torch.cuda.current_stream().cuda_stream
    for call in range(2):
        x = rng.standard_normal((batch, n, 2)).astype(np.float32) * np.float32(1 + 3 * call)
        want = reference_fft(x) if n >= 32 else None
        d_in.copy_(torch.from_numpy(x))
        d_out.fill_(float("nan"))
        fft_forward_noalloc(d_in, d_out, n, batch, stream)
        torch.cuda.synchronize()
        got = d_out.cpu().numpy()
        # an independent answer at every size (numpy, float64) ...
        z = np.fft.fft(x[..., 0].astype(np.float64) + 1j * x[..., 1], axis=1)
        close_enough(got, np.stack([z.real, z.imag], axis=-1))
        # ... and the existing entry point's (the handle API, which starts at 32 points)
        if want is not None:
            close_enough(got, want)


def test_noalloc_fft_rejects_what_the_fixed_buffer_cannot_hold():
    import torch
    lib = _lib.load()
    d = torch.zeros(64, dtype=torch.float32, device="cuda")
    e = torch.zeros(64, dtype=torch.float32, device="cuda")
    for n, batch in (((1 << 20), 2), ((1 << 21), 1), (4096, 257)):
        with pytest.raises(_lib.SgzError) as err:
            _lib.check(lib.sgz_fft_forward_noalloc(d.data_ptr(), e.data_ptr(), n, batch, None))
        assert err.value.code == _abi.EINVAL and "exceeds the built-in scratch" in str(err.value)
    for n in (0, 1, 3, 1000):
        with pytest.raises(_lib.SgzError) as err:
            _lib.check(lib.sgz_fft_forward_noalloc(d.data_ptr(), e.data_ptr(), n, 1, None))
        assert err.value.code == _abi.EINVAL
    with pytest.raises(_lib.SgzError):
        _lib.check(lib.sgz_fft_forward_noalloc(d.data_ptr(), d.data_ptr(), 8, 1, None))  # in place
    with pytest.raises(_lib.SgzError) as err:
        _lib.check(lib.sgz_fft_forward_noalloc(d.data_ptr(), d.data_ptr() + 32, 8, 1, None))  # 8 points = 64 bytes: the ranges share 32 bytes
    assert "overlap" in str(err.value)
