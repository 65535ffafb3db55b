"""GPU tier, collected last on purpose: the no-allocation
entry point salz_cuda_lcp_noalloc (include/salz_cuda.h; scratch in a static device array,
csrc/lcp.cu) against the existing entry point salz_cuda_debug_stages, through the C ABI.
Integer work: the bar is bit-exact equality, as everywhere in tests/test_gpu.py.  The calls
share one scratch array that keeps the last call's contents, so the test makes several calls
in a row - a large text, a small one, the large one again - and must pass both alone in a
fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

import salz_b200
from salz_b200 import api, synth

pytestmark = pytest.mark.gpu


def _noalloc(lib, torch, text, sa):
    """PLCP / LCP of `text` (numpy u8) from `sa` (numpy i32) through salz_cuda_lcp_noalloc.
    torch only owns the caller-side device buffers; the entry point itself allocates nothing."""
    n = len(text)
    dev = torch.device("cuda", 0)
    d_text = torch.from_numpy(np.array(text, dtype=np.uint8)).to(dev)   # a writable copy: torch warns on read-only views
    d_sa = torch.from_numpy(np.array(sa, dtype=np.int32)).to(dev)
    d_plcp = torch.full((n,), -7, dtype=torch.int32, device=dev)
    d_lcp = torch.full((n,), -7, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    rc = lib.salz_cuda_lcp_noalloc(0, d_text.data_ptr(), n, d_sa.data_ptr(), d_plcp.data_ptr(), d_lcp.data_ptr(), None)
    assert rc == 0
    torch.cuda.synchronize()
    return d_plcp.cpu().numpy(), d_lcp.cpu().numpy()


def test_lcp_noalloc_matches_debug_stages():
    import torch
    lib = salz_b200.load_library()
    assert api.NOALLOC_MAX_N == 1 << 18
    enc = salz_b200.SalzCuda(device=0, max_block_len=1 << 18, max_batch_blocks=1)
    try:
        cases = [("eng250k", synth.english_like(250000, seed=61)),
                 ("dna5k", synth.dna_like(5000, seed=62, block=5000)),
                 ("a70k", np.full(70000, 0x61, np.uint8)),
                 ("eng250k", None),                                   # the first text again
                 ("max", synth.english_like((1 << 18) - 8, seed=63)),  # the largest text a 2^18 context takes
                 ("tiny", np.frombuffer(b"abracadabra!", np.uint8))]
        keep = {}
        for name, data in cases:
            if data is None:
                data = keep[name]
            keep[name] = data
            st = enc.stages(np.concatenate([data, np.zeros(8, np.uint8)]))   # the existing entry point: n = N - 8
            n = len(data)
            assert len(st["sa"]) == n
            for rep in range(2):                                             # two calls in a row
                plcp, lcp = _noalloc(lib, torch, data, st["sa"])
                assert np.array_equal(plcp, st["plcp"]), (name, rep, "plcp")
                assert np.array_equal(lcp, st["lcp"]), (name, rep, "lcp")
    finally:
        enc.close()


def test_lcp_noalloc_rejects_oversize_and_sets_up_with_zero():
    import torch
    lib = salz_b200.load_library()
    assert lib.salz_cuda_lcp_noalloc(0, None, 0, None, None, None, None) == 0          # set-up only
    n = api.NOALLOC_MAX_N + 1
    d = torch.zeros(n, dtype=torch.uint8, device="cuda:0")
    s = torch.zeros(n, dtype=torch.int32, device="cuda:0")
    o = torch.zeros(n, dtype=torch.int32, device="cuda:0")
    assert lib.salz_cuda_lcp_noalloc(0, d.data_ptr(), n, s.data_ptr(), None, o.data_ptr(), None) == -2
    assert lib.salz_cuda_lcp_noalloc(0, None, 10, None, None, None, None) == -1          # NULL buffers
    assert int(o.abs().sum().item()) == 0                                                # nothing was written
