"""N>1 host logic on CPU: world_size-2 gloo process group, stream partition + host gather of segments
(the multi-GPU path has no data-path collective).  Each rank runs the ORACLE as the stand-in
checker for its streams (no GPU here) and rank 0 compares the gathered result with a single-rank run."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_covers_every_stream_once(fv):
    for n in (1, 7, 20, 160):
        for w in (1, 2, 3, 4, 8):
            seen = []
            for r in range(w):
                seen += list(fv.sharding.partition_streams(n, w, r))
            assert seen == list(range(n))
            sizes = [len(fv.sharding.partition_streams(n, w, r)) for r in range(w)]
            assert max(sizes) - min(sizes) <= 1
    assert list(fv.sharding.partition_streams(160, 8, 3)) == list(range(60, 80))  # BASELINE config 4: 20 per GPU
    with pytest.raises(ValueError):
        fv.sharding.partition_streams(4, 2, 2)


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    import fvad_b200
    from oracle import pyoracle as po
    fv = fvad_b200.load()
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        wd = fv.weights.synthetic(seed=1234)
        local = {}
        for s in fv.sharding.partition_streams(3, world, rank):
            x = _stream(s)
            p = po.Pipeline(wd, 16000, 1)
            p.push_samples(x)
            local[s] = p.segments(0)
        merged = fv.sharding.gather_segments(local, dst=0)
        if rank == 0:
            q.put(merged)
        else:
            assert merged is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _stream(s):
    rng = np.random.default_rng(500 + s)
    n = 16000 * 8
    t = np.arange(n) / 16000.0
    x = rng.normal(0, 0.0015, n)
    m = (t >= 1.0 + 0.3 * s) & (t < 2.6 + 0.3 * s)
    x[m] += 0.25 * np.sin(2 * np.pi * (700 + 150 * s) * t[m]) + 0.2 * np.sin(2 * np.pi * 1500 * t[m])
    return x.astype(np.float32)[None, :]


def test_two_rank_gloo_gather_equals_single_rank(fv, po, wd):
    import torch.multiprocessing as mp
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    merged = q.get(timeout=240)
    for p in procs:
        p.join(timeout=240)
        assert p.exitcode == 0
    expect = {}
    for s in range(3):
        p = po.Pipeline(wd, 16000, 1)
        p.push_samples(_stream(s))
        expect[s] = p.segments(0)
    assert merged == expect
    assert sum(len(v) for v in merged.values()) >= 2  # the streams do contain detected speech
