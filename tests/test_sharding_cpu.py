"""CPU tests of the multi-GPU host logic: contiguous unit ranges per rank and the per-rank base
offsets of a sharded BYTES tensor, with a real world_size-2 gloo process group. The data path has
no collective; what is checked is that concatenating per-rank results in rank order
equals the unsharded oracle result."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from oracle import bytes_codec as bc
from tests.golden import cases
from triton_client_b200.sharding import exclusive_offsets, shard_range


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 8, 255, 256, 8192, 1000003):
        for world in (1, 2, 3, 4, 8):
            ranges = [shard_range(n, r, world) for r in range(world)]
            assert ranges[0][0] == 0 and ranges[-1][1] == n
            for (a, b), (c, d) in zip(ranges, ranges[1:]):
                assert b == c and a <= b
            sizes = [b - a for a, b in ranges]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)
    assert exclusive_offsets([5, 0, 7]) == ([0, 5, 5], 12)


def test_serialization_is_concatenative():
    blob, lens = cases.gbytes(1000)
    arr = cases.to_object_array(blob, lens)
    whole = bc.serialize_elements(arr)
    for world in (2, 3, 8):
        parts = [bc.serialize_elements(arr[slice(*shard_range(len(arr), r, world))]) for r in range(world)]
        assert b"".join(parts) == whole
        offs, total = exclusive_offsets([len(p) for p in parts])
        assert total == len(whole)
        for r in range(world):
            assert whole[offs[r]: offs[r] + len(parts[r])] == parts[r]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist

    from triton_client_b200.sharding import bytes_base_offset, gather_sizes

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    blob, lens = cases.gbytes(1000)
    arr = cases.to_object_array(blob, lens)
    lo, hi = shard_range(len(arr), rank, world)
    mine = bc.serialize_elements(arr[lo:hi])   # each rank serialises its own element range
    base, total = bytes_base_offset(len(mine))
    sizes = gather_sizes(len(mine))
    np.save(os.path.join(out_dir, "r%d.npy" % rank), np.array([lo, hi, base, total, len(mine)] + sizes))
    with open(os.path.join(out_dir, "r%d.bin" % rank), "wb") as f:
        f.write(mine)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_sharded_bytes(tmp_path, monkeypatch):
    world = 2
    port = _free_port()
    # spawned ranks must be able to import this module (tests.*) whatever pytest's cwd / import mode
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    monkeypatch.setenv("PYTHONPATH", root + os.pathsep + os.environ.get("PYTHONPATH", ""))
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    blob, lens = cases.gbytes(1000)
    whole = bc.serialize_elements(cases.to_object_array(blob, lens))
    logical = bytearray(len(whole))
    for r in range(world):
        meta = np.load(tmp_path / ("r%d.npy" % r))
        lo, hi, base, total, mine = (int(x) for x in meta[:5])
        assert total == len(whole) and list(meta[5:]) == [int(np.load(tmp_path / ("r%d.npy" % q))[4]) for q in range(world)]
        assert (lo, hi) == shard_range(1000, r, world)
        piece = open(tmp_path / ("r%d.bin" % r), "rb").read()
        assert len(piece) == mine
        logical[base: base + mine] = piece     # what each rank's region holds, placed at its base offset
    assert bytes(logical) == whole
