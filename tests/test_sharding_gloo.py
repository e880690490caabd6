"""N > 1 host logic on CPU: two gloo ranks shard the streams, each computes its shard (the oracle
stands in for the GPU here — tests may), rank 0 gathers, and the result equals the single-process
run.  Mirrors how bench.py / a harness shards streams over GPUs with no data-path collective."""
import os
import socket

import numpy as np
import pytest

from signalizer_b200.sharding import stream_range


def test_stream_range_partitions_exactly():
    for n in (0, 1, 7, 1024, 1025):
        for world in (1, 2, 3, 8):
            spans = [stream_range(r, world, n) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        stream_range(2, 2, 10)


def _worker(rank, world, port, n_streams, hop, frames, out_path):
    import torch
    import torch.distributed as dist
    from oracle.binding import OracleSpectrum
    from signalizer_b200 import _abi, signals
    from signalizer_b200.sharding import gather_results, stream_range
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg = _abi.default_config(window_size=1024, axis_points=128)
        lo, hi = stream_range(rank, world, n_streams)
        n = signals.frames_length(cfg.window_size, hop, frames)
        audio = signals.batch(hi - lo, n, first_stream=lo)
        local = OracleSpectrum(cfg, hi - lo).process_frames(audio, hop, frames)
        full = gather_results(torch.from_numpy(local), n_streams, dst=0)
        if rank == 0:
            np.save(out_path, full.numpy())
        else:
            assert full is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_rank_sharding_matches_single_process(tmp_path):
    import torch.multiprocessing as mp
    from oracle.binding import OracleSpectrum
    from signalizer_b200 import _abi, signals
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    n_streams, hop, frames = 5, 256, 3  # odd count: ranks own 3 and 2 streams
    out = str(tmp_path / "gathered.npy")
    mp.spawn(_worker, args=(2, port, n_streams, hop, frames, out), nprocs=2, join=True)
    got = np.load(out)
    cfg = _abi.default_config(window_size=1024, axis_points=128)
    audio = signals.batch(n_streams, signals.frames_length(1024, hop, frames))
    want = OracleSpectrum(cfg, n_streams).process_frames(audio, hop, frames)
    np.testing.assert_array_equal(got, want)
