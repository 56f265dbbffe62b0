"""Host-side batch sharding (styletts2_b200/shard.py) on CPU: ranges, and a world-size-2 gloo run
where every rank decodes its rows with a stand-in per-item function and the gathered result must
equal the single-process one bit for bit (the property the GPU test proves for the real kernels)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from styletts2_b200.shard import shard_range, decode_sharded


def test_shard_ranges_cover_the_batch():
    for B in (0, 1, 5, 32, 513):
        for world in (1, 2, 3, 8):
            got = [shard_range(B, world, r) for r in range(world)]
            assert got[0][0] == 0 and got[-1][1] == B
            assert all(a[1] == b[0] for a, b in zip(got, got[1:]))
            sizes = [h - l for l, h in got]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _fake_decoder(asr, F0, N, s, noise=None):
    # per-item, deterministic, depends on every input of the item only
    w = asr.mean(dim=(1, 2)) + F0.sum(1) * 1e-3 + N.mean(1) + s.pow(2).mean(1)
    base = torch.arange(300 * F0.shape[1], dtype=torch.float32)[None] * 1e-3
    out = torch.sin(base * w[:, None])
    if noise is not None:
        out = out + noise
    return out[:, None, :]


def _inputs(B, T=4):
    g = torch.Generator().manual_seed(3)
    return (torch.randn(B, 512, T // 2, generator=g), torch.rand(B, T, generator=g) * 200,
            torch.randn(B, T, generator=g), torch.randn(B, 128, generator=g), torch.randn(B, 300 * T, generator=g))


def _worker(rank, world, port, B, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    asr, F0, N, s, noise = _inputs(B)
    full = decode_sharded(_fake_decoder, asr, F0, N, s, gather=True, noise=noise)
    mine = decode_sharded(_fake_decoder, asr, F0, N, s, noise=noise)
    lo, hi = shard_range(B, world, rank)
    ok = torch.equal(full, _fake_decoder(asr, F0, N, s, noise)) and torch.equal(mine, full[lo:hi])
    q.put((rank, bool(ok), tuple(full.shape)))
    dist.destroy_process_group()


@pytest.mark.parametrize("B", [4, 5])
def test_world_size_2_gloo_matches_single_process(B):
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, B, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res), res
    assert all(shape == (B, 1, 1200) for _, _, shape in res)
