"""GPU tier, multi-GPU path: blocks sharded over the devices of the box, sizes
all-gathered over NCCL, assembled stream identical to the single-process reference stream.

The C entry points (salz_b200/csrc/multi_gpu.cu) run on however many GPUs the box has - one
on the driver's test box (NCCL world of 1), eight under the scaling run; the process-per-GPU
torch.distributed variant and the two-contexts-in-one-process test need two.  The host logic
is covered on CPU by tests/test_shard.py (gloo)."""
import os
import socket
import subprocess

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


def test_two_gpu_assembly_matches_reference_stream(orc):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from salz_b200 import synth
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


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _reference_stream(orc, data, bl):
    out = b""
    sizes = []
    for off in range(0, len(data), bl):
        r = orc.encode(data[off: off + bl])
        sizes.append(len(r))
        out += len(r).to_bytes(4, "little") + r
    return out, sizes


def _mixed(bl, nblocks, tail=4321):
    from salz_b200 import synth
    if nblocks < 4:
        return synth.english_like((nblocks - 1) * bl + tail, seed=40)
    k = nblocks // 4
    return np.concatenate([synth.english_like(2 * k * bl, seed=41), synth.uniform_random(k * bl, seed=42),
                           synth.dna_like((nblocks - 3 * k - 1) * bl + tail, seed=43, block=bl)])


def test_c_multi_entry_on_all_visible_gpus(orc):
    """salz_cuda_multi_encode_stream (one host thread, context and NCCL rank per device): the
    assembled stream equals the reference's records, for streams longer and shorter than
    devices x batch, and for fewer blocks than devices."""
    import torch  # noqa: F401  (torch's NCCL first: one copy per process)
    import salz_b200
    bl = 1 << 16
    m = salz_b200.SalzCudaMulti(ndev=0, max_block_len=bl, max_batch_blocks=6)
    try:
        G = m.ndev
        assert G == torch.cuda.device_count()
        for nblocks in (1, max(1, G - 1), 5 * G + 3, 13 * G + 1):
            data = _mixed(bl, nblocks)
            want, wsizes = _reference_stream(orc, data, bl)
            stream, sizes = m.encode_stream(data, bl)
            assert sizes.tolist() == wsizes, nblocks
            assert stream == want, nblocks
    finally:
        m.close()


def test_c_comm_allgather_offsets(orc):
    """salz_cuda_comm_* with a thread per visible device: every rank receives all sizes and
    the record offsets off_b = sum_{j<b} (4 + len_j) (ref programs/salzcli.c:162-169)."""
    import threading
    import torch
    import salz_b200
    from salz_b200 import api
    G = torch.cuda.device_count()
    uid = api.unique_id()
    nblocks = 5 * G + 2
    rng = np.random.default_rng(5)
    sizes = rng.integers(13, 1 << 20, nblocks).astype(np.uint32)
    res = [None] * G

    def rank_main(r):
        comm = salz_b200.ShardComm(r, uid, r, G)
        try:
            lo, hi = api.block_range(nblocks, r, G)
            res[r] = comm.allgather_sizes(sizes[lo:hi], nblocks) + (comm.gathers(),)
        finally:
            comm.close()

    th = [threading.Thread(target=rank_main, args=(r,)) for r in range(G)]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=300)
    want_offs = np.concatenate([[0], np.cumsum(sizes.astype(np.uint64) + 4)])
    for r in range(G):
        assert res[r] is not None, r
        got, offs, gathers = res[r]
        assert np.array_equal(got, sizes) and np.array_equal(offs, want_offs) and gathers == 1


def test_two_contexts_two_devices_one_process(orc):
    """Kernel attributes (opt-in shared memory of k_scatter / k_parse) are per device: a
    context on device 1 created after one on device 0, in the same process, must work."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import salz_b200
    from salz_b200 import synth
    bl = 1 << 18
    data = np.concatenate([synth.english_like(3 * bl, seed=3), synth.dna_like(bl + 999, seed=4, block=bl)])
    want, _ = _reference_stream(orc, data, bl)
    e0 = salz_b200.SalzCuda(0, bl, 4)
    e1 = salz_b200.SalzCuda(1, bl, 4)
    try:
        assert e0.encode_stream(data, bl)[0] == want
        assert e1.encode_stream(data, bl)[0] == want
        assert e0.encode_stream(data, bl)[0] == want
    finally:
        e0.close()
        e1.close()


def test_cli_gpus_option_matches_reference_cli(tmp_path):
    """`salz --gpus 0 -k -5 file` (all visible GPUs, pipelined reader/encoder/writer, several
    slots) writes the reference CLI's file."""
    ref_cli = os.path.join(ROOT, "oracle", "_ref", "salz_ref")
    assert os.path.exists(ref_cli), "prebuilt reference CLI absent (oracle/_ref/salz_ref)"
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "lib")], check=True)
    from salz_b200 import synth
    data = np.concatenate([synth.english_like(9 << 20, seed=2), synth.uniform_random((2 << 20) + 4096, seed=3),
                           synth.dna_like((5 << 20) + 777, seed=4)])
    a = tmp_path / "a.bin"; b = tmp_path / "b.bin"
    a.write_bytes(data.tobytes()); b.write_bytes(data.tobytes())
    env = dict(os.environ, SALZ_BATCH_MIB="2")
    subprocess.run([os.path.join(ROOT, "programs", "salz"), "--gpus", "0", "-k", "-q", "-5", str(a)], check=True, env=env)
    subprocess.run([ref_cli, "-k", "-q", "-5", str(b)], check=True)
    assert (tmp_path / "a.bin.salz").read_bytes() == (tmp_path / "b.bin.salz").read_bytes()
