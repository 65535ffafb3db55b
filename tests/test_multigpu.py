"""GPU tier, >= 2 devices: blocks sharded over two processes, one per GPU, sizes
all-gathered over NCCL, assembled file identical to the single-process reference
stream.  Skipped on a one-GPU box; the host logic is covered on CPU by
tests/test_shard.py (gloo)."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, data_bytes, block_len, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    import salz_b200
    from salz_b200 import shard
    data = np.frombuffer(data_bytes, np.uint8)
    nblocks = (len(data) + block_len - 1) // block_len
    lo, hi = shard.block_range(nblocks, rank, world)
    enc = salz_b200.SalzCuda(device=rank, max_block_len=block_len, max_batch_blocks=max(1, hi - lo))
    mine = data[lo * block_len: min(len(data), hi * block_len)]
    stream, local_sizes = enc.encode_stream(mine, block_len)
    enc.close()
    sizes = shard.allgather_sizes(local_sizes, nblocks, rank, world, device=torch.device("cuda", rank))
    offs = shard.record_offsets(sizes)
    q.put((rank, int(offs[lo]), stream, int(offs[-1])))
    dist.barrier()
    dist.destroy_process_group()


def test_two_gpu_assembly_matches_reference_stream(port):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from oracle.binding import Ref
    from salz_b200 import synth
    orc = Ref() if Ref.available() else port
    block_len = 1 << 17
    data = np.concatenate([synth.english_like(11 * block_len, seed=8), synth.uniform_random(2 * block_len, seed=9),
                           synth.dna_like(4 * block_len + 4096, seed=10, block=block_len)])
    s = socket.socket(); s.bind(("127.0.0.1", 0)); prt = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, prt, data.tobytes(), block_len, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in range(2)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    total = results[0][3]
    out = bytearray(total)
    out[0:8] = (0x53414C5A).to_bytes(4, "little") + block_len.to_bytes(4, "little")
    for _, off, stream, _ in results:
        out[off: off + len(stream)] = stream
    want = bytearray((0x53414C5A).to_bytes(4, "little") + block_len.to_bytes(4, "little"))
    for off in range(0, len(data), block_len):
        r = orc.encode(data[off: off + block_len])
        want += len(r).to_bytes(4, "little") + r
    assert bytes(out) == bytes(want)
