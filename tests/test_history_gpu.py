"""Device-resident audio history: the ring keeps the reference's history semantics
(the last n samples, oldest first, silence before time zero — what cpl::AudioStream's two-segment
views deliver at TransformDSP.inl:49,65-88), and streaming appends + history_process emit exactly
the frames audioEntryPoint (TransformDSP.inl:1165-1211) schedules, equal to the oracle's frames on
the same samples."""
import numpy as np
import pytest

from oracle.binding import OracleSpectrum
from signalizer_b200 import _abi, _lib, signals
from signalizer_b200.spectrum import TransformPairBatch
from test_spectrum_gpu import compare_frames

pytestmark = pytest.mark.gpu


def history_model(chunks, capacity, n_last):
    """numpy statement of the ring: zeros, then everything appended; the last n_last samples"""
    S = chunks[0].shape[0]
    total = np.concatenate([np.zeros((S, capacity, 2), np.float32)] + chunks, axis=1)
    return total[:, total.shape[1] - n_last:]


@pytest.mark.parametrize("layout", [_abi.LAYOUT_INTERLEAVED, _abi.LAYOUT_PLANAR])
@pytest.mark.parametrize("device_blocks", [False, True])
def test_history_ring_semantics(layout, device_blocks):
    import torch
    cfg = _abi.default_config(window_size=1024)
    S, C = 3, 1500
    g = TransformPairBatch(cfg, S)
    g.history_reserve(C)
    assert g.history_info() == (C, 0, 0)
    assert np.all(g.history_read(C) == 0)
    audio = signals.batch(S, 9000)
    chunks, pos = [], 0
    for n in (1, 499, 1000, 1, 1500, 37, 4000, 1962):   # wraps several times; 4000 > capacity
        blk = np.ascontiguousarray(audio[:, pos:pos + n]); pos += n
        chunks.append(blk)
        send = blk if layout == _abi.LAYOUT_INTERLEAVED else np.ascontiguousarray(blk.transpose(0, 2, 1))
        g.history_append(torch.from_numpy(send).cuda() if device_blocks else send, layout=layout)
        for n_last in (C, 1024, 1, 0, 777):
            np.testing.assert_array_equal(g.history_read(n_last), history_model(chunks, C, n_last))
    assert g.history_info() == (C, pos, 0)
    g.history_reserve(C)   # reserving again clears
    assert np.all(g.history_read(C) == 0) and g.history_info() == (C, 0, 0)


@pytest.mark.parametrize("size,hop", [(8192, 2048), (4096, 1024), (3000, 750), (8192, 1023)])
def test_streaming_frames_match_oracle_and_batched_call(size, hop):
    cfg = _abi.default_config(window_size=size)
    S, W = 4, size
    total = 9 * hop + 517
    audio = signals.batch(S, total)
    g = TransformPairBatch(cfg, S)
    g.history_reserve(W + 4 * hop)
    frames, pos, rng = [], 0, np.random.default_rng(7)
    while pos < total:
        n = int(min(total - pos, rng.integers(1, 3 * hop)))
        g.history_append(np.ascontiguousarray(audio[:, pos:pos + n])); pos += n
        got, k = g.history_process(hop, 8)
        assert k == (pos // hop) - sum(f.shape[1] for f in frames)
        if k:
            frames.append(got.copy())
    got = np.concatenate(frames, axis=1)
    F = total // hop
    assert got.shape[1] == F and g.history_info()[2] == F * hop
    # the frame at boundary b sees [b - W, b): the batched call on the signal prefixed by W - hop samples of silence
    padded = np.concatenate([np.zeros((S, W - hop, 2), np.float32), audio], axis=1)
    want = OracleSpectrum(cfg, S).process_frames(padded, hop, F)
    compare_frames(got, want, cfg, "streaming vs oracle")
    same = TransformPairBatch(cfg, S).process_frames(padded, hop, F)
    np.testing.assert_array_equal(got, same)   # same kernel, same arithmetic, state carried across the calls


def test_history_max_frames_leaves_the_rest_pending_and_overrun_is_an_error():
    cfg = _abi.default_config(window_size=1024)
    S, hop = 2, 256
    g = TransformPairBatch(cfg, S)
    g.history_reserve(4096)
    audio = signals.batch(S, 3000)
    g.history_append(audio[:, :2000].copy())
    a, k = g.history_process(hop, 3)
    assert k == 3 and g.history_info()[2] == 3 * hop
    b, k = g.history_process(hop, 100)
    assert k == 2000 // hop - 3
    ref = TransformPairBatch(cfg, S)
    padded = np.concatenate([np.zeros((S, 1024 - hop, 2), np.float32), audio], axis=1)
    np.testing.assert_array_equal(np.concatenate([a, b], axis=1), ref.process_frames(padded, hop, 2000 // hop))
    # nothing due: zero frames, not an error
    _, k = g.history_process(hop, 4)
    assert k == 0
    # overrun: 4096-sample ring, the oldest pending window starts more than 4096 samples back
    g.history_append(signals.batch(S, 4000))
    with pytest.raises(_lib.SgzError) as e:
        g.history_process(hop, 4)
    assert e.value.code == _abi.ESTATE and "overrun" in str(e.value)
    # errors before reserve / bad capacity
    h = TransformPairBatch(cfg, S)
    with pytest.raises(_lib.SgzError):
        h.history_append(audio[:, :10].copy())
    with pytest.raises(_lib.SgzError):
        h.history_reserve(1000)   # smaller than the window


def test_history_view_feeds_the_staged_calls():
    import torch
    cfg = _abi.default_config(window_size=2048)
    S = 3
    g = TransformPairBatch(cfg, S)
    g.history_reserve(5000)
    audio = signals.batch(S, 7000)
    g.history_append(audio[:, :3000].copy()); g.history_append(audio[:, 3000:].copy())
    tail = g.history_read(2048)
    np.testing.assert_array_equal(tail, audio[:, -2048:])
    ref = TransformPairBatch(cfg, S)
    assert ref.prepare_transform(tail)
    ref.do_transform()
    want = ref.get_raw_fft()
    ptr, stride = g.history_view(2048)
    assert g.prepare_transform_device_view(ptr, 2048, stride)
    g.do_transform()
    np.testing.assert_array_equal(g.get_raw_fft(), want)
