"""The N>1 path on CPU: two gloo ranks shard a batch of independent images (no data-path collective), process
their shards with the oracle standing in for the device, and agree on timing and statuses the way bench.py does."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from slimg_b200 import _shard


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 128, 1024, 1025):
        for world in (1, 2, 3, 4, 8):
            spans = [_shard.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    assert [_shard.shard_range(1024, r, 8) for r in (0, 7)] == [(0, 128), (896, 1024)]     # config 3 at 8 GPUs
    with pytest.raises(ValueError):
        _shard.shard_range(4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n_images, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        lo, hi = _shard.shard_range(n_images, rank, world)
        outs, statuses = [], []
        for i in range(lo, hi):
            rng = np.random.default_rng(0x51136 + i)          # image i is the same whichever rank owns it
            img = rng.integers(0, 256, (30, 40, 4), dtype=np.uint8)
            if i == 3:                                        # one bad image: must fail alone (ErrorCollector)
                statuses.append(7)
                continue
            r = oracle.resize(img, oracle.ResizeMode.width(20))
            r = oracle.crop(r, oracle.CropMode.aspect_ratio(16, 9))
            outs.append((i, oracle.extend(r, oracle.ExtendMode.aspect_ratio(1, 1), (255, 255, 255, 255))))
            statuses.append(0)
        dist.barrier()
        ms = _shard.max_over_ranks(10.0 + 5.0 * rank)         # rank 1 is the slow one
        all_status = _shard.gather_statuses(statuses, n_images)
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), ms=ms, status=np.array(all_status), lo=lo, hi=hi,
                 **{f"img{i}": o for i, o in outs})
    finally:
        dist.destroy_process_group()


def test_two_ranks_shard_images_without_collective(tmp_path):
    world, n_images = 2, 7
    mp.spawn(_worker, args=(world, _free_port(), n_images, str(tmp_path)), nprocs=world, join=True)
    import oracle
    got = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    # every rank agrees on the slowest rank's time and on every image's status
    assert [float(g["ms"]) for g in got] == [15.0, 15.0]
    want_status = [0, 0, 0, 7, 0, 0, 0]
    assert all(g["status"].tolist() == want_status for g in got)
    # shards are disjoint and cover the job; each image equals the single-process result
    assert (int(got[0]["lo"]), int(got[0]["hi"]), int(got[1]["lo"]), int(got[1]["hi"])) == (0, 4, 4, 7)
    seen = set()
    for g in got:
        for key in g.files:
            if not key.startswith("img"):
                continue
            i = int(key[3:])
            assert i not in seen
            seen.add(i)
            rng = np.random.default_rng(0x51136 + i)
            img = rng.integers(0, 256, (30, 40, 4), dtype=np.uint8)
            r = oracle.crop(oracle.resize(img, oracle.ResizeMode.width(20)), oracle.CropMode.aspect_ratio(16, 9))
            want = oracle.extend(r, oracle.ExtendMode.aspect_ratio(1, 1), (255, 255, 255, 255))
            assert np.array_equal(g[key], want)
    assert seen == {0, 1, 2, 4, 5, 6}
    assert _shard.whole_job_throughput(24.0 * 128, 2, 0.015) == pytest.approx(24.0 * 256 / 0.015)
