"""World-size-2 gloo test of the multi-GPU host logic (sharding, max-over-ranks timing, gather).

The transform passed in here is the CPU oracle expression -- this test is about the host-side
plumbing that bench.py --gpus N uses, not about the kernels (those are covered by -m gpu tests)."""
import os
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
    sys.path.insert(0, str(ROOT))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as O
    from trt_dft_plugins import multi_gpu as mg
    torch.manual_seed(1)
    full = torch.randn(7, 6, 10)                      # 7 images: uneven split 4 + 3
    y = mg.sharded_apply(full.shape, 2, lambda x: O.torch_rfftn(x, 2), lambda s, c: full[s:s + c], gather=True)
    slow = mg.max_over_ranks(10.0 + rank)             # rank 1 is "slower"
    if rank == 0:
        torch.save({"y": y, "slow": slow, "want": O.torch_rfftn(full, 2)}, os.path.join(out_dir, "r0.pt"))
    else:
        assert y is None and slow == 11.0
    dist.barrier()
    dist.destroy_process_group()


def test_shard_batch_partitions():
    sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
    from trt_dft_plugins.multi_gpu import shard_batch
    for batch in (0, 1, 7, 1280, 6144):
        for world in (1, 2, 4, 8):
            spans = [shard_batch(batch, r, world) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == batch
            for (s0, c0), (s1, _) in zip(spans, spans[1:]):
                assert s0 + c0 == s1
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
    with pytest.raises(ValueError):
        shard_batch(4, 2, 2)


def test_two_rank_gloo_sharding(tmp_path):
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r = torch.load(tmp_path / "r0.pt")
    assert r["slow"] == 11.0                          # max over ranks
    assert torch.equal(r["y"], r["want"])             # gathered shards == full-batch transform
