"""Host-side logic of the N>1 path (prompt sharding + result gather + max-over-ranks timing) with world_size 2 over
gloo on CPU.  The data path itself has no collective."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from echo_tts_b200 import serving


def test_shard_indices_cover_exactly_once():
    for n in (0, 1, 7, 64):
        for world in (1, 2, 3, 8):
            parts = [serving.shard_indices(n, r, world) for r in range(world)]
            assert sorted(i for p in parts for i in p) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    with pytest.raises(ValueError):
        serving.shard_indices(4, 2, 2)


def test_branch_rows_follow_reference_mask_layout():
    """cond | text-dropped | speaker-dropped blocks of B rows (reference inference.py:494-497)."""
    assert [serving.branch_rows(r, 4) for r in range(3)] == [slice(0, 4), slice(4, 8), slice(8, 12)]
    with pytest.raises(ValueError):
        serving.branch_rows(3, 4)


def test_run_sharded_without_process_group():
    assert serving.run_sharded(["a", "bb"], lambda s, i: (i, len(s))) == [(0, 1), (1, 2)]


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        prompts = [f"prompt {i}" for i in range(7)]
        seen = []

        def generate(p, i):
            seen.append(i)
            return torch.full((i + 1,), float(i))            # ragged "waveforms"

        out = serving.run_sharded(prompts, generate)
        ms = serving.max_over_ranks_ms(10.0 + rank)
        q.put((rank, seen, None if out is None else [t.tolist() for t in out], ms))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, seen0, out0, ms0), (r1, seen1, out1, ms1) = res
    assert seen0 == [0, 2, 4, 6] and seen1 == [1, 3, 5]       # disjoint round-robin shares
    assert out1 is None and out0 == [[float(i)] * (i + 1) for i in range(7)]   # original order, on rank 0 only
    assert ms0 == ms1 == 11.0                                  # max over ranks
