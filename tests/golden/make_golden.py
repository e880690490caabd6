#!/usr/bin/env python
"""Regenerates tests/golden/spectrum_golden.npz from oracle/_ref: the REFERENCE's own Source/Spectrum transform code
(TransformDSP.inl, TransformPair.h, TransformConstant.h) compiled in this container against stand-ins for its absent cpl
library (oracle/build_ref.sh).  The reference ships no golden vectors of its own; these are outputs of
its in-tree source run here, and the restated oracle must reproduce them bit for bit (tests/test_golden.py), as must
oracle/_ref wherever it is present.  The cpl primitives underneath (FFT, filters, window) are this repo's decisions.
Where the reference checkout is missing the restated oracle generates the same bits (tests/test_oracle_ref.py).
Inputs are the seeded synthetic signals of signalizer_b200/signals.py; every case records its configuration overrides.

    python tests/golden/make_golden.py"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import binding  # noqa: E402
from oracle.binding import OracleSpectrum  # noqa: E402
from signalizer_b200 import _abi, signals  # noqa: E402

CASES = {
    "c1_hann4096_separate_lanczos": dict(window_size=4096, window_type=_abi.WIN_HANN),
    "c3_bh8192_separate_lanczos": dict(),
    "bh8192_left_linear": dict(configuration=_abi.CH_LEFT, bin_polation=_abi.BIN_LINEAR),
    "hamming6000_midside_none": dict(window_size=6000, window_type=_abi.WIN_HAMMING, configuration=_abi.CH_MIDSIDE, bin_polation=_abi.BIN_NONE),
    "phase2048_lanczos_p256": dict(window_size=2048, configuration=_abi.CH_PHASE, axis_points=256),
    "complex2048_lanczos": dict(window_size=2048, configuration=_abi.CH_COMPLEX),
    "hann1024_separate_linear_p512": dict(window_size=1024, window_type=_abi.WIN_HANN, bin_polation=_abi.BIN_LINEAR, axis_points=512),
    "bh16384_side_lanczos_p700": dict(window_size=16384, configuration=_abi.CH_SIDE, axis_points=700),
}
S, F = 3, 3


def run(overrides, ref=False):
    cfg = _abi.default_config(**overrides)
    hop = cfg.window_size // 4
    audio = signals.batch(S, signals.frames_length(cfg.window_size, hop, F), first_stream=1)
    o = OracleSpectrum(cfg, S, ref=ref)
    res = o.process_frames(audio, hop, F)
    return res, np.stack([o.states(0), o.states(1)])


if __name__ == "__main__":
    out = {}
    binding.build()
    ref = binding.ref_available()
    for name, ov in CASES.items():
        res, st = run(ov, ref=ref)
        out[name + "/results"] = res
        out[name + "/states"] = st
    out["meta"] = np.frombuffer(json.dumps({"streams": S, "frames": F, "first_stream": 1, "cases": CASES,
                                                 "source": "oracle/_ref (reference TransformDSP.inl)" if ref else "oracle (restatement)"}).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "spectrum_golden.npz"), **out)
    print("wrote", len(CASES), "cases from", "oracle/_ref" if ref else "the restated oracle")
