"""CPU tier: the oracle restatement against golden vectors and the compiled
reference, the kernel model (design check) against the oracle."""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

from salz_b200 import synth
import kernel_model as km

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))


def test_known_answers(port):
    # SURVEY.md Appendix A (produced by the compiled reference)
    assert port.encode(b"abcdefghi").hex() == "09000000616263646566676869"
    assert port.encode(b"a" * 64).hex() == "120000010000000000000c6061006161616161616161"
    assert port.encode(b"abcabcabcabcabcabcabcabcXYZ01234").hex() == \
        "1400000100000000000028186162630258595a3031323334"
    for n in range(0, 9):
        assert port.encode(b"x" * n) is None


@pytest.mark.parametrize("vec", GOLDEN["vectors"], ids=lambda v: v["name"])
def test_golden_vectors(port, vec):
    data = bytes.fromhex(vec["input"])
    st = port.stages(data)
    assert st["out"].hex() == vec["output"]
    if "sa" in vec:
        assert st["sa"].tolist() == vec["sa"]
        assert st["lcp"].tolist() == vec["lcp"]
        assert st["cand"][1:].reshape(-1).tolist() == vec["cand"]
        assert st["parse"][1:].reshape(-1).tolist() == vec["parse"]
    assert port.decode(st["out"], len(data)) == data


def test_golden_digests(port):
    gens = {"english_1MiB": lambda: synth.english_like(1 << 20, seed=5),
            "dna_1MiB": lambda: synth.dna_like(1 << 20, seed=7),
            "markov_300k": lambda: synth.markov_text(300000, seed=42)}
    for d in GOLDEN["digests"]:
        if d["name"] in gens:
            out = port.encode(gens[d["name"]]())
            assert len(out) == d["len"] and hashlib.sha256(out).hexdigest() == d["sha256"], d["name"]


def test_vnibble_known_answers(port):
    for v in GOLDEN["vnibble"]:
        assert port.lib.oracle_vnibble_size(v["v"]) == v["size"]


def test_port_matches_reference(port, ref):
    for name, data in synth.small_cases(seed=11):
        a, b = port.stages(data), ref.stages(data)
        for k in ("sa", "plcp", "lcp"):
            assert np.array_equal(a[k], b[k]), (name, k)
        assert np.array_equal(a["cand"][1:], b["cand"][1:]), name
        assert np.array_equal(a["parse"][1:], b["parse"][1:]), name
        assert a["out"] == b["out"], name
        assert ref.decode(a["out"], len(data)) == data.tobytes()
        assert port.decode(b["out"], len(data)) == data.tobytes()


def test_port_matches_reference_1mib(port, ref):
    for data in (synth.english_like((1 << 20), seed=21), synth.dna_like((1 << 20) - 77, seed=22)):
        assert port.encode(data) == ref.encode(data)


def test_capacity_semantics(port, ref):
    data = synth.english_like(5000, seed=3)
    full = ref.encode(data)
    for cap in (3, 4, len(full) - 1, len(full), len(full) + 1):
        assert port.encode(data, cap=cap) == ref.encode(data, cap=cap), cap
    rnd = synth.uniform_random(5000, seed=3)
    for cap in (5004, 5003, 5004 + 8 * 79 - 1, 5004 + 8 * 79):
        assert port.encode(rnd, cap=cap) == ref.encode(rnd, cap=cap), cap


def test_kernel_model_matches_oracle(port):
    """Design check of the device algorithms (tests/kernel_model.py)."""
    cases = synth.small_cases(seed=5)[:15]
    cases += [("zeros", np.zeros(100, np.uint8)),
              ("zero_tail", np.concatenate([np.arange(50, dtype=np.uint8) + 1, np.zeros(20, np.uint8)])),
              ("a3500", np.full(3500, 97, np.uint8))]
    texts = [c[1][: len(c[1]) - 8].tobytes() for c in cases]
    sas, _ = km.sa_model(texts)
    for (name, a), sa in zip(cases, sas):
        st = port.stages(a)
        n = len(a) - 8
        assert np.array_equal(sa, st["sa"]), name
        plcp, lcp = km.plcp_model(a[:n], st["sa"])
        assert np.array_equal(plcp, st["plcp"]) and np.array_equal(lcp, st["lcp"]), name
        cand = km.ansv_model(st["sa"], st["lcp"])
        assert np.array_equal(cand[1:], st["cand"][1:]), name
        cost, choice, jump, _ = km.parse_model(st["cand"], len(a))
        assert np.array_equal(choice[1:n], st["parse"][1:n, :2]), name
        assert np.array_equal(cost[1:n], st["parse"][1:n, 2]), name
        assert km.emit_model(a, len(a), choice, jump) == st["out"], name
        # the shortened first pass (left end of every segment only) changes nothing in the result
        cost2, choice2, _, _ = km.parse_model(st["cand"], len(a), seg_len=128, warm=32)
        assert np.array_equal(choice2[1:n], choice[1:n]) and np.array_equal(cost2[1:n], cost[1:n]), name


def test_chunk_entry_speculation_is_exact():
    """The speculative chunk-entry chain (emit.cu k_chunk_entries) against the plain chain on
    random step functions: short steps (paths merge), long jumps over ranges, both mixed."""
    rng = np.random.default_rng(5)
    for trial in range(60):
        chunk = int(rng.choice([4, 8, 16]))
        lanes = int(rng.choice([2, 4, 8]))
        N = int(rng.integers(1, 600))
        kind = trial % 3
        step = rng.integers(1, 4, N) if kind == 0 else rng.integers(1, 3 * chunk * lanes, N) if kind == 1 else \
            np.where(rng.random(N) < 0.05, rng.integers(1, 200, N), rng.integers(1, 3, N))
        # jump[p] = first position >= the end of p's chunk on the path p, p + step[p], ...
        jump = np.zeros(N, np.int64)
        for p in range(N - 1, -1, -1):
            t = min(N, p + int(step[p]))
            end = min(N, (p // chunk + 1) * chunk)
            jump[p] = t if t >= end else jump[t]
        want = [-1] * ((N + chunk - 1) // chunk)
        e = 0
        while e < N:
            want[e // chunk] = e
            e = int(jump[e])
        assert km.chunk_entries_model(jump, N, chunk, lanes) == want, (trial, chunk, lanes, N)


def test_radix_pass_model_order_check_and_ticket_schedule():
    """Design check of the default radix pass (sa.cu: k_scatter_tma<5>, k_sched_interleave) without a
    GPU: slots handed out in lane order give a stable pass - five passes sort - and the order check
    stays silent; slots handed out in any other order are caught by the check.  The interleaved
    ticket schedule is a permutation in which every tile comes after its block's earlier tiles."""
    rng = np.random.default_rng(7)
    for kind in ("uniform", "skewed", "runs"):
        n = 1000
        if kind == "uniform":
            keys = rng.integers(0, 1 << 40, n, dtype=np.uint64)
        elif kind == "skewed":
            keys = (rng.integers(0, 6, n, dtype=np.uint64) << np.uint64(32)) | (rng.integers(0, 3, n, dtype=np.uint64) << np.uint64(8)) | rng.integers(0, 2, n, dtype=np.uint64)
        else:
            keys = np.repeat(rng.integers(0, 1 << 40, n // 50, dtype=np.uint64), 50)
        keys = (keys << np.uint64(20)) | np.arange(n, dtype=np.uint64)        # unique keys, like (group, second, p)
        cur = keys.copy()
        for ps in range(5):
            cur, bad = km.scatter_pass_model(cur, 20 + 8 * ps)
            assert bad == 0, (kind, ps)
        assert np.array_equal(cur, np.sort(keys)), kind
        # hardware that served the lanes of a step in another order: detected, tile by tile
        rev, bad = km.scatter_pass_model(keys, 20, lane_order=range(31, -1, -1))
        if kind != "uniform":      # equal digits inside a step are certain for these inputs
            assert bad > 0, kind
        shuf = list(rng.permutation(32))
        _, bad2 = km.scatter_pass_model(keys, 20, lane_order=shuf)
        if kind == "runs":
            assert bad2 > 0
    for counts, group in (([5, 5, 5, 5], 2), ([3, 0, 7, 1, 4], 4), ([6], 64), ([2, 9, 9, 1, 0, 3, 8], 3), ([4] * 10, 64)):
        sched = km.sched_interleave_model(counts, group)
        assert sorted(sched) == list(range(sum(counts)))
        base = np.concatenate([[0], np.cumsum(counts)])
        pos = np.empty(len(sched), np.int64)
        pos[sched] = np.arange(len(sched))
        for b, c in enumerate(counts):
            for lt in range(1, c):
                assert pos[base[b] + lt - 1] < pos[base[b] + lt]      # look-back predecessors claimed first
        # groups stay in block order: every ticket of a group precedes every ticket of the next one
        for g0 in range(group, len(counts), group):
            lo, hi = pos[:base[g0]], pos[base[g0]:]
            if len(lo) and len(hi):
                assert lo.max() < hi.min()
