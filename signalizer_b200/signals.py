"""Synthetic audio for the offline harness: sines, log chirps and
counter-based uniform noise, float32, 48 kHz, interleaved stereo [stream][sample][2].

Signal type is `stream mod 3`; the noise is a pure function of (stream, channel, sample) so any
rank or language regenerates identical bits."""
import numpy as np

_MASK = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = (x + np.uint64(0x9E3779B97F4A7C15)) & _MASK
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _MASK
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _MASK
        return z ^ (z >> np.uint64(31))


def noise(stream: int, channel: int, n: int) -> np.ndarray:
    idx = np.arange(n, dtype=np.uint64)
    key = np.uint64(0x5167A11E) ^ (np.uint64(stream) << np.uint64(33)) ^ (np.uint64(channel) << np.uint64(32))
    u = _splitmix64(key ^ idx) >> np.uint64(40)
    return (u.astype(np.float64) * 2.0 ** -23 - 1.0).astype(np.float32)


def stream_audio(stream: int, n: int, sample_rate: float = 48000.0) -> np.ndarray:
    """One stereo stream, shape [n, 2] float32."""
    t = np.arange(n, dtype=np.float64)
    kind = stream % 3
    out = np.empty((n, 2), dtype=np.float32)
    if kind == 0:
        f = 55.0 * 2.0 ** ((stream % 96) / 12.0)
        ph = 2.0 * np.pi * f * t / sample_rate
        out[:, 0] = (0.5 * np.sin(ph)).astype(np.float32)
        out[:, 1] = (0.25 * np.sin(ph + np.pi / 3.0)).astype(np.float32)
    elif kind == 1:
        f0, f1 = 20.0, 20000.0
        T = n / sample_rate
        k = np.log(f1 / f0)
        ph = 2.0 * np.pi * f0 * T / k * (np.exp(t / sample_rate / T * k) - 1.0)
        out[:, 0] = (0.5 * np.sin(ph)).astype(np.float32)
        out[:, 1] = (0.5 * np.cos(ph)).astype(np.float32)
    else:
        out[:, 0] = noise(stream, 0, n)
        out[:, 1] = noise(stream, 1, n)
    return out


def batch(n_streams: int, n: int, sample_rate: float = 48000.0, first_stream: int = 0) -> np.ndarray:
    """[n_streams, n, 2] float32, interleaved stereo."""
    out = np.empty((n_streams, n, 2), dtype=np.float32)
    for s in range(n_streams):
        out[s] = stream_audio(first_stream + s, n, sample_rate)
    return out


def frames_length(window: int, hop: int, n_frames: int) -> int:
    return window + (n_frames - 1) * hop
