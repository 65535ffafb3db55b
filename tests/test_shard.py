"""CPU tier: multi-GPU host logic (block sharding + size all-gather) with
world_size 2 over gloo.  The per-block encoder is stood in by the oracle
(checker) because there is no GPU here; the GPU tier runs the same assembly
with the CUDA encoder."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

from salz_b200 import shard


def test_block_range_partition():
    for nb in (1, 2, 7, 257, 8193):
        for w in (1, 2, 4, 8):
            got = []
            for r in range(w):
                lo, hi = shard.block_range(nb, r, w)
                got += list(range(lo, hi))
            assert got == list(range(nb))


def test_record_offsets():
    offs = shard.record_offsets([10, 20, 5])
    assert offs.tolist() == [8, 22, 46, 55]


def _worker(rank, world, port, data_bytes, block_len, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle.binding import Port
    orc = Port()
    data = np.frombuffer(data_bytes, np.uint8)
    nblocks = (len(data) + block_len - 1) // block_len
    lo, hi = shard.block_range(nblocks, rank, world)
    recs = [orc.encode(data[b * block_len: (b + 1) * block_len]) for b in range(lo, hi)]
    sizes = shard.allgather_sizes([len(r) for r in recs], nblocks, rank, world)
    offs = shard.record_offsets(sizes)
    q.put((rank, lo, [(int(offs[lo + i]), len(r).to_bytes(4, "little") + r) for i, r in enumerate(recs)],
           int(offs[-1])))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_assembly_matches_single_process():
    from oracle.binding import Port
    from salz_b200 import synth
    block_len = 1 << 15
    data = synth.english_like(5 * block_len + 4096, seed=8)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, data.tobytes(), block_len, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    total = results[0][3]
    out = bytearray(total)
    out[0:8] = (0x53414C5A).to_bytes(4, "little") + block_len.to_bytes(4, "little")
    for _, _, recs, _ in results:
        for off, rec in recs:
            out[off: off + len(rec)] = rec
    orc = Port()
    want = bytearray((0x53414C5A).to_bytes(4, "little") + block_len.to_bytes(4, "little"))
    for off in range(0, len(data), block_len):
        r = orc.encode(data[off: off + block_len])
        want += len(r).to_bytes(4, "little") + r
    assert bytes(out) == bytes(want)


def test_block_range_matches_the_c_library():
    """salz_b200/shard.py (torch.distributed tests) and salz_cuda_block_range (the C path) cut
    the block index space the same way."""
    from salz_b200 import api, shard
    for nblocks in (1, 2, 7, 8, 9, 257, 8193):
        for world in (1, 2, 3, 4, 8):
            cover = []
            for r in range(world):
                assert api.block_range(nblocks, r, world) == shard.block_range(nblocks, r, world)
                lo, hi = api.block_range(nblocks, r, world)
                cover += list(range(lo, hi))
            assert cover == list(range(nblocks))
