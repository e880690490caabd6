"""Host logic of the product (no GPU needed): the C-ABI library loads, exports every symbol of
include/sgz.h, refuses to compute without a device, and derives tables that are bit-identical to
the oracle's (integer bin maps, float32 window / frequency / slope tables)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from oracle.binding import OracleSpectrum
from signalizer_b200 import _abi, _lib
from signalizer_b200.spectrum import TransformPairBatch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "sgz.h")).read()
    declared = set(re.findall(r"\b(sgz_[a-z_0-9]+)\s*\(", header))
    assert declared, "no declarations parsed"
    lib = _lib.load()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/sgz.h but not exported"
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)


def test_struct_layout_matches_header_defaults():
    lib = _lib.load()
    cfg = _abi.SpectrumConfig()
    lib.sgz_spectrum_default_config(C.byref(cfg))
    ref = _abi.default_config()
    for name, _ in _abi.SpectrumConfig._fields_:
        a, b = getattr(cfg, name), getattr(ref, name)
        if name == "pole":
            assert list(a) == list(b)
        else:
            assert a == b, name


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_lib.SgzError) as e:
        TransformPairBatch(_abi.default_config(), 1, device=0)
    assert e.value.code == _abi.ECUDA
    t = TransformPairBatch(_abi.default_config(), 1, device=-1)  # tables-only handle
    audio = np.zeros((1, 8192, 2), np.float32)
    with pytest.raises(_lib.SgzError) as e:
        t.prepare_transform(audio)
    assert e.value.code == _abi.ECUDA
    with pytest.raises(_lib.SgzError):
        t.process_frames(audio, 2048, 1)


def test_create_rejects_bad_configs():
    for bad in (dict(axis_points=1), dict(window_size=0), dict(configuration=9), dict(bin_polation=5),
                dict(sample_rate=0.0), dict(abi_version=99), dict(window_type=42)):
        with pytest.raises(_lib.SgzError) as e:
            TransformPairBatch(_abi.default_config(**bad), 1, device=-1)
        assert e.value.code == _abi.EINVAL, bad


SIZES = [(4096, _abi.WIN_HANN, 1024), (8192, _abi.WIN_BLACKMAN_HARRIS, 1024), (1000, _abi.WIN_HAMMING, 300),
         (32, _abi.WIN_RECTANGULAR, 16), (5, _abi.WIN_HANN, 2), (65536, _abi.WIN_NUTTALL, 2000), (2048, _abi.WIN_FLATTOP, 777)]


@pytest.mark.parametrize("size", SIZES)
@pytest.mark.parametrize("conf", range(8))
@pytest.mark.parametrize("scaling", [_abi.VIEW_LINEAR, _abi.VIEW_LOGARITHMIC])
def test_tables_bit_exact_against_oracle(size, conf, scaling):
    for bp in (_abi.BIN_NONE, _abi.BIN_LINEAR, _abi.BIN_LANCZOS):
        for view in ((0.0, 1.0), (0.25, 0.6)):
            cfg = _abi.default_config(window_size=size[0], window_type=size[1], axis_points=size[2], configuration=conf,
                                      bin_polation=bp, view_scaling=scaling, view_left=view[0], view_right=view[1],
                                      slope_a=0.5, slope_b=2.0, window_shape=(size[0] % 2))
            t = TransformPairBatch(cfg, 1, device=-1)
            o = OracleSpectrum(cfg, 1)
            assert (t.N, t.P, t.channels) == (o.N, o.P, o.channels)
            np.testing.assert_array_equal(t.window_kernel(), o.window_kernel().astype(np.float32))
            np.testing.assert_array_equal(t.mapped_frequencies(), o.mapped_frequencies().astype(np.float32))
            np.testing.assert_array_equal(t.slope_map(), o.slope_map().astype(np.float32))
            assert t.info.window_kernel_scale == o.info.window_kernel_scale
            assert t.info.inv_size == o.info.inv_size
            assert t.info.freq_to_bin == o.info.freq_to_bin
            assert t.info.resonator_vectors == o.info.resonator_vectors
            for got, want, name in zip(t.bin_map(), o.bin_map(), ("mode", "first", "count", "fallback")):
                np.testing.assert_array_equal(got, want, err_msg=f"{name} bp={bp} view={view}")
            if conf != _abi.CH_COMPLEX:
                assert t.info.break_point == o.info.break_point


def test_fused_eligibility():
    def fused(**kw):
        return TransformPairBatch(_abi.default_config(**kw), 1, device=-1).info.fused_available
    assert fused() == 1                                             # C3: 8192 BH Separate Lanczos
    assert fused(window_size=4096, window_type=_abi.WIN_HANN) == 1  # C1
    assert fused(window_size=6000) == 1                             # zero-padded to 8192
    assert fused(configuration=_abi.CH_LEFT) == 1
    assert fused(configuration=_abi.CH_PHASE) == 1
    assert fused(configuration=_abi.CH_COMPLEX) == 1
    assert fused(window_size=2048) == 1                             # k_fused8<2>
    assert fused(window_size=1024) == 1                             # k_fused8<1>
    assert fused(window_size=512) == 0
    assert fused(window_size=16384) == 1                            # k_fused8<16>
    assert fused(window_size=20000) == 0
    assert fused(algo=_abi.ALGO_RSNT) == 0
