"""Which blocks of the 1 GiB mixed stream differ between one batch of 1025 blocks and batches of 512?"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import salz_b200
from oracle.binding import Ref

os.environ.setdefault("SALZ_BENCH_CACHE", "/tmp/c")


def main():
    bl = 1 << 20
    size = (1024 << 20) + 4096
    data = bench.make_input("mixed", size, 1, bl)
    nb = (size + bl - 1) // bl
    out = {}
    import torch
    d_src = torch.from_numpy(data).to("cuda:0")
    cap = salz_b200.api.stream_len_max(size, bl)
    d_dst = torch.empty(cap, dtype=torch.uint8, device="cuda:0")
    for batch in (512, nb):
        e = salz_b200.SalzCuda(0, bl, batch)
        sizes = np.zeros(nb, np.uint32)
        n = e.encode_stream_ptr(d_src.data_ptr(), size, bl, d_dst.data_ptr(), cap, sizes, device=True)   # whole batches, no split
        torch.cuda.synchronize()
        stream = d_dst[:n].cpu().numpy().tobytes()
        out[batch] = (stream, sizes)
        st = e.stats()
        print("batch", batch, "bytes", len(stream), "rounds", st["doubling_rounds"], flush=True)
        e.close()
    a, b = out[512], out[nb]
    diff = np.flatnonzero(a[1] != b[1])
    print("blocks with different sizes:", len(diff), diff[:20])
    offa = np.concatenate([[0], np.cumsum(a[1].astype(np.int64) + 4)])
    offb = np.concatenate([[0], np.cumsum(b[1].astype(np.int64) + 4)])
    bad = [i for i in range(nb) if a[0][offa[i]:offa[i + 1]] != b[0][offb[i]:offb[i + 1]]]
    print("blocks with different bytes:", len(bad), bad[:40])
    orc = Ref()
    for i in bad[:3]:
        want = orc.encode(data[i * bl:(i + 1) * bl])
        ra = a[0][offa[i] + 4:offa[i + 1]]
        rb = b[0][offb[i] + 4:offb[i + 1]]
        print("block", i, "section kind", (i // 8) % 3, "ref len", len(want), "batch512 ok", ra == want, "batch%d ok" % nb, rb == want, "len", len(rb),
              "first diff", next((k for k in range(min(len(rb), len(want))) if rb[k] != want[k]), None))


if __name__ == "__main__":
    main()
