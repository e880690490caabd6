"""TriggeringProcessor::update / processMutating (Source/Oscilloscope/StreamPreprocessing.h:54-206): the window
scheduler that consumes the trigger positions.  Integer bookkeeping, so everything is compared exactly: events
(kind, origin, size, buffered), the carried state and the peaks left in the queue, product against the oracle's
restatement, over several callback blocks.  The scheduler runs on the host, so the parity cases need no GPU; the GPU
test chains it behind sgz_scope_trigger."""
import ctypes as C

import numpy as np
import pytest

from oracle import binding
from signalizer_b200 import _abi, _lib, signals
from signalizer_b200.scope import scope_trigger_windows

binding.build()
PROCESS, SWAP = _abi.WINDOW_PROCESS, _abi.WINDOW_SWAP


def oracle_lib():
    lib = binding.load()
    lib.sgzo_scope_trigger_windows.restype = C.c_int
    lib.sgzo_scope_trigger_windows.argtypes = [C.c_uint32, C.POINTER(_abi.ScopeWindowState), C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.c_uint32,
                                               C.c_uint64, C.c_uint64, C.POINTER(_abi.ScopeWindowEvent), C.POINTER(C.c_uint32), C.c_uint32]
    return lib


def fresh_states(S, window_size):
    st = (_abi.ScopeWindowState * S)()
    for s in range(S):
        st[s].window_size = window_size
        st[s].window_changed = 1   # setSettings on a new processor: ceil(new) != ceil(old)
    return st


def state_tuple(st):
    return [(x.window_size, x.old_peak, x.current_peak, x.buffered_samples, x.front_origin, x.steady_clock, x.is_working_on_peak, x.window_changed) for x in st]


def run_blocks(lib, S, window_size, blocks, peak_lists, max_peaks=64):
    """blocks: [(steady_clock, n_samples)], peak_lists[b][s]: positions found in block b for stream s"""
    states = fresh_states(S, window_size)
    peaks = np.zeros((S, max_peaks), dtype=np.uint64)
    counts = np.zeros(S, dtype=np.uint32)
    log = []
    for b, (clock, n) in enumerate(blocks):
        for s in range(S):
            new = peak_lists[b][s]
            peaks[s, counts[s]:counts[s] + len(new)] = new
            counts[s] += len(new)
        ev = scope_trigger_windows(lib, states, peaks, counts, clock, n)
        log.append((ev, state_tuple(states), [peaks[s, :counts[s]].tolist() for s in range(S)]))
    return log


def test_periodic_triggers_give_centred_windows():
    """peaks every 250 samples, window 100: each swap shows the samples around one peak"""
    lib = _lib.load()
    blocks = [(1000 + 512 * b, 512) for b in range(6)]
    peak_lists = [[[p for p in range(1100, 5000, 250) if clock <= p < clock + n]] for clock, n in blocks]
    log = run_blocks(lib, 1, 100.0, blocks, peak_lists)
    swaps = [e for ev, _, _ in log for e in ev[0] if e[0] == SWAP]
    consumed = [e for ev, _, _ in log for e in ev[0] if e[0] == PROCESS]
    assert sum(e[2] for e in consumed) == 6 * 512                 # every sample went into the back buffer exactly once
    assert len(swaps) >= 8
    peaks = list(range(1100, 5000, 250))
    for (kind, origin, size, buffered), p in zip(swaps, peaks):
        assert origin <= p < origin + size                         # the window holds its trigger position
        assert size <= 102 and size <= buffered                    # ceil(halfSize + needed + 1) capped by what is buffered
        assert abs((origin + size / 2) - p) <= 52


@pytest.mark.parametrize("window_size", [0.0, 1.0, 7.5, 100.0, 333.3, 4096.0])
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_product_equals_oracle_on_random_queues(window_size, seed):
    rng = np.random.default_rng(seed)
    S, B = 5, 12
    blocks, clock = [], 5000
    for b in range(B):
        n = int(rng.integers(1, 700))
        if b == 7:
            clock += 1000   # a gap in the steady clock (the host skipped audio): frontOrigin is re-based (:80-84)
        blocks.append((clock, n))
        clock += n
    peak_lists = []
    for clock, n in blocks:
        per_stream = []
        for s in range(S):
            k = int(rng.integers(0, 4)) if s else 0            # stream 0 never triggers
            per_stream.append(sorted(int(clock + x) for x in rng.integers(0, n, size=k)))
        peak_lists.append(per_stream)
    got = run_blocks(_lib.load(), S, window_size, blocks, peak_lists)
    want = run_blocks(oracle_lib(), S, window_size, blocks, peak_lists)
    assert got == want
    assert any(e[0] == SWAP for ev, _, _ in got for stream in ev for e in stream) or window_size == 0.0


def test_window_change_drops_stale_peaks_and_event_overflow_is_reported():
    lib, olib = _lib.load(), oracle_lib()
    for L in (lib, olib):
        states = fresh_states(1, 64.0)
        peaks = np.zeros((1, 16), dtype=np.uint64)
        peaks[0, :3] = [10, 20, 5000]
        counts = np.array([3], dtype=np.uint32)
        states[0].is_working_on_peak = 1                            # update(): pops the peak being worked on, then everything before the clock
        ev = scope_trigger_windows(L, states, peaks, counts, 4000, 100)
        assert counts[0] == 1 and peaks[0, 0] == 5000
        assert ev[0] == [(PROCESS, 4000, 100, 0)]
        assert states[0].front_origin == 4100 - 65 and states[0].buffered_samples == 65   # capped at ceil(window) + 1
    states = fresh_states(1, 8.0)
    peaks = np.arange(100, 100 + 16 * 40, 40, dtype=np.uint64).reshape(1, 16)
    counts = np.array([16], dtype=np.uint32)
    with pytest.raises(ValueError):
        scope_trigger_windows(lib, states, peaks, counts, 0, 2000, max_events=4)
    assert counts[0] == 16 and states[0].window_changed == 1        # nothing written back


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [0, 1])
def test_trigger_positions_feed_the_window_scheduler(mode):
    """sgz_scope_trigger (GPU) -> sgz_scope_trigger_windows, block by block, against the oracle doing both"""
    from oracle.binding import OracleScope
    from signalizer_b200.scope import Context
    S, n, B = 6, 4096, 4
    audio = signals.batch(S, n * B, first_stream=3)
    ctx, orc = Context(), OracleScope()
    logs = []
    for impl, lib in ((ctx, _lib.load()), (orc, oracle_lib())):
        trig = (_abi.ScopeTriggerState * S)()
        states = fresh_states(S, 600.0)
        peaks = np.zeros((S, 2048), dtype=np.uint64)
        counts = np.zeros(S, dtype=np.uint32)
        log = []
        for b in range(B):
            block = np.ascontiguousarray(audio[:, b * n:(b + 1) * n])
            clock = 10_000 + b * n
            found, fc = impl.scope_trigger(mode, _abi.OSC_LEFT, block, trig, 0.05, 0.1, steady_clock=clock, max_peaks=256)
            for s in range(S):
                k = min(len(found[s]), peaks.shape[1] - int(counts[s]))   # (noise streams trigger far faster than windows consume)
                peaks[s, counts[s]:counts[s] + k] = found[s][:k]
                counts[s] += k
            ev = scope_trigger_windows(lib, states, peaks, counts, clock, n, max_events=1024)
            log.append((ev, state_tuple(states), counts.tolist()))
        logs.append(log)
    assert logs[0] == logs[1]
    assert any(e[0] == SWAP for ev, _, _ in logs[0] for stream in ev for e in stream)
