"""CPU: the N > 1 host logic (stream sharding + manifest gather) under a world_size-2 gloo group.
The data path itself has no collective: each rank encodes its own streams."""
import os
import socket
import sys

import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load_engine():
    """engine.py on its own: its module-level import of the CUDA binding is satisfied by a
    throw-away stub (there is no GPU here and the sharding helpers never touch it)."""
    import importlib.util
    import types
    saved = {k: sys.modules.get(k) for k in ("audio_codec_b200", "audio_codec_b200._lib")}
    pkg = types.ModuleType("audio_codec_b200")
    pkg.__path__ = [os.path.join(ROOT, "audio_codec_b200")]
    pkg._lib = types.ModuleType("audio_codec_b200._lib")
    sys.modules["audio_codec_b200"] = pkg
    sys.modules["audio_codec_b200._lib"] = pkg._lib
    try:
        spec = importlib.util.spec_from_file_location("audio_codec_b200._engine_host",
                                                      os.path.join(ROOT, "audio_codec_b200", "engine.py"))
        m = importlib.util.module_from_spec(spec)
        m.__package__ = "audio_codec_b200"
        spec.loader.exec_module(m)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return m


def test_shard_streams_is_a_balanced_partition():
    eng = _load_engine()
    dur = [20.0, 3.5, 16.4, 16.4, 8.0, 1.0, 30.0, 2.0, 2.0, 12.5]
    for world in (1, 2, 4, 8):
        parts = eng.shard_streams(dur, world)
        assert sorted(i for p in parts for i in p) == list(range(len(dur)))
        loads = [sum(dur[i] for i in p) for p in parts]
        assert max(loads) - min(loads) <= max(dur)          # LPT bound
    assert eng.shard_streams([], 2) == [[], []]


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    eng = _load_engine()
    dur = [5.0, 1.0, 4.0, 2.0, 3.0]
    mine = eng.shard_streams(dur, world)[rank]
    rows = [(i, 1000 * i + 7, 0) for i in mine]             # stand-in for (stream, n_bytes, err) of the local encode
    table = eng.gather_manifest(rows)
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, mine, table))


def test_world2_gloo_manifest():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res.sort()
    assert sorted(res[0][1] + res[1][1]) == [0, 1, 2, 3, 4] and not set(res[0][1]) & set(res[1][1])
    want = [(i, 1000 * i + 7, 0) for i in range(5)]
    assert res[0][2] == want and res[1][2] == want
