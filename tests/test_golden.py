"""Committed golden vectors (tests/golden/spectrum_golden.npz, made by tests/golden/make_golden.py from oracle/_ref, i.e.
the reference's own TransformDSP.inl run in the build container): the restated oracle - and oracle/_ref where it is
present - must reproduce them bit for bit, and the CUDA path must match them within the stated tolerance without
either in the loop."""
import importlib.util
import json
import os

import numpy as np
import pytest

from signalizer_b200 import _abi, signals

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
make_golden = importlib.util.module_from_spec(spec)
spec.loader.exec_module(make_golden)
GOLD = np.load(os.path.join(HERE, "golden", "spectrum_golden.npz"))


def test_golden_metadata_matches_generator():
    meta = json.loads(bytes(GOLD["meta"]).decode())
    assert meta["cases"] == json.loads(json.dumps(make_golden.CASES))
    assert (meta["streams"], meta["frames"]) == (make_golden.S, make_golden.F)


@pytest.mark.parametrize("name", list(make_golden.CASES))
def test_oracle_reproduces_golden_bit_exact(name):
    res, st = make_golden.run(make_golden.CASES[name])
    np.testing.assert_array_equal(res, GOLD[name + "/results"])
    np.testing.assert_array_equal(st, GOLD[name + "/states"])


@pytest.mark.parametrize("name", list(make_golden.CASES))
def test_reference_build_reproduces_golden_bit_exact(name):
    from oracle import binding
    binding.build()
    if not binding.ref_available():
        pytest.skip("oracle/_ref not built (no reference checkout)")
    res, st = make_golden.run(make_golden.CASES[name], ref=True)
    np.testing.assert_array_equal(res, GOLD[name + "/results"])
    np.testing.assert_array_equal(st, GOLD[name + "/states"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(make_golden.CASES))
def test_cuda_matches_golden(name):
    from signalizer_b200.spectrum import TransformPairBatch
    from tolerances import assert_linear_close, assert_log_close
    ov = make_golden.CASES[name]
    cfg = _abi.default_config(**ov)
    S, F = make_golden.S, make_golden.F
    hop = cfg.window_size // 4
    audio = signals.batch(S, signals.frames_length(cfg.window_size, hop, F), first_stream=1)
    g = TransformPairBatch(cfg, S)
    got = g.process_frames(audio, hop, F)
    want = GOLD[name + "/results"]
    conf = cfg.configuration
    used = 2 if conf in (_abi.CH_SEPARATE, _abi.CH_MIDSIDE) else 1
    dy = 1.0 / np.log(1e6)
    clipped = want <= -300
    lin = np.where(clipped, 0.0, 1e-6 * np.exp(np.where(clipped, 0.0, want.astype(np.float64)) / dy))
    for s in range(S):
        for f in range(F):
            w, gg, l = want[s, f, ..., :used], got[s, f, ..., :used], lin[s, f, ..., :used]
            m = ~clipped[s, f, ..., :used]
            assert np.all(gg[~m] <= -300) or not (~m).any()
            assert_log_close(gg[m], w[m], lin=l[m], peak=l.max(), dy=dy, what=f"{name} stream {s} frame {f}")
    st = GOLD[name + "/states"]
    for k in range(2):
        assert_linear_close(g.get_states(k)[..., 0], st[k][..., 0], frame_axes=(1,), what=f"{name} states {k}")
