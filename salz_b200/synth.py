"""Seeded synthetic inputs of the shapes BASELINE.json names.

All generators are numpy-only and deterministic for a given (size, seed).
Sizes follow the reference CLI's legality rule: callers should use
k*block + tail with tail >= 9 bytes.
"""
import numpy as np


def markov_text(size, seed=42, order=2):
    """Order-2 Markov text over 27 symbols (a-z, space), Dirichlet(0.15) rows."""
    rng = np.random.default_rng(seed)
    k = 27
    alpha = np.frombuffer(b"abcdefghijklmnopqrstuvwxyz ", dtype=np.uint8)
    rows = rng.dirichlet(np.full(k, 0.15), size=k * k)
    cdf = np.cumsum(rows, axis=1)
    u = rng.random(size)
    out = np.empty(size, dtype=np.int64)
    a = b = 0
    for i in range(size):
        c = int(np.searchsorted(cdf[a * k + b], u[i]))
        if c >= k:
            c = k - 1
        out[i] = c
        a, b = b, c
    return alpha[out]


def english_like(size, seed=1, vocab=4096, zipf_s=1.3):
    """Zipf-distributed draws from a fixed vocabulary joined by spaces/punctuation."""
    rng = np.random.default_rng(seed)
    letters = np.frombuffer(b"etaoinshrdlcumwfgypbvkjxqz", dtype=np.uint8)
    lw = 1.0 / np.arange(1, 27) ** 0.8
    lw /= lw.sum()
    wl = np.clip(rng.poisson(4.2, vocab) + 1, 1, 14)
    words = [letters[rng.choice(26, size=int(l), p=lw)].tobytes() for l in wl]
    pw = 1.0 / np.arange(1, vocab + 1) ** zipf_s
    pw /= pw.sum()
    # draw enough words for `size` bytes (mean word+sep ~ 6.5 bytes)
    nw = int(size / 4.0) + 64
    idx = rng.choice(vocab, size=nw, p=pw)
    seps = rng.choice(np.array([0, 0, 0, 0, 0, 0, 0, 1, 2, 3]), size=nw)
    septab = [b" ", b", ", b". ", b"\n"]
    # token table: row = word * 4 + separator, padded to 16 bytes
    tab = np.zeros((vocab * 4, 16), np.uint8)
    tlen = np.zeros(vocab * 4, np.int64)
    for w in range(vocab):
        for s in range(4):
            t = words[w] + septab[s]
            tab[w * 4 + s, : len(t)] = np.frombuffer(t, np.uint8)
            tlen[w * 4 + s] = len(t)
    tok = idx.astype(np.int64) * 4 + seps
    lens = tlen[tok]
    ends = np.cumsum(lens)
    k = min(int(np.searchsorted(ends, size, side="left")) + 1, nw)  # tokens until >= size bytes
    total = int(ends[k - 1])
    out = np.empty(max(total, size), dtype=np.uint8)
    step = 1 << 21
    for lo in range(0, k, step):
        hi = min(k, lo + step)
        ln = lens[lo:hi]
        first = int(ends[lo] - lens[lo])
        nbytes = int(ends[hi - 1]) - first
        tokid = np.repeat(np.arange(lo, hi), ln)
        within = np.arange(nbytes) - np.repeat(ends[lo:hi] - ln - first, ln)
        out[first: first + nbytes] = tab[tok[tokid], within]
    while total < size:
        ln = min(total, size - total)
        out[total: total + ln] = out[:ln]
        total += ln
    return out[:size].copy()


def dna_like(size, seed=2, block=1 << 20, repeats_per_block=3, kmin=6, kmax=19):
    """Uniform ACGT background with planted repeats of length 2^k copied from
    earlier positions of the same block (so they survive blocking)."""
    rng = np.random.default_rng(seed)
    out = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=size)].copy()
    nblocks = (size + block - 1) // block
    for b in range(nblocks):
        lo = b * block
        hi = min(size, lo + block)
        span = hi - lo
        for _ in range(repeats_per_block):
            k = int(rng.integers(kmin, kmax + 1))
            L = min(1 << k, span // 2 - 1)
            if L < 16:
                continue
            src = int(rng.integers(0, span - 2 * L + 1))
            dst = int(rng.integers(src + L, span - L + 1))
            out[lo + dst: lo + dst + L] = out[lo + src: lo + src + L]
    return out


def single_byte(size, byte=0x61):
    return np.full(size, byte, dtype=np.uint8)


def uniform_random(size, seed=4):
    return np.random.default_rng(seed).integers(0, 256, size=size, dtype=np.uint8)


MIXED_SECTION = 8 << 20


def mixed_section(k, ln, seed=3):
    """Section k of the mixed stream (config 4): english-like, dna-like, random in turn."""
    kind = k % 3
    if kind == 0:
        return english_like(ln, seed=seed * 1000 + k)
    if kind == 1:
        return dna_like(ln, seed=seed * 1000 + k)
    return uniform_random(ln, seed=seed * 1000 + k)


def _mixed_job(job):
    k, ln, seed, shm_name, off = job
    from multiprocessing import shared_memory
    shm = shared_memory.SharedMemory(name=shm_name)
    try:
        np.frombuffer(shm.buf, np.uint8, ln, off)[:] = mixed_section(k, ln, seed)
    finally:
        shm.close()
    return k


def mixed_stream(size, seed=3, section=MIXED_SECTION, lo=0, hi=None, workers=1):
    """Bytes [lo, hi) of the concatenation of english-like, dna-like and random sections
    (config 4) of `size` bytes in total.  Section k is generated independently (seed
    seed*1000+k), so any range costs only its own sections; `workers` > 1 spreads them over
    processes (spawn: safe after CUDA is initialised in the caller)."""
    hi = size if hi is None else hi
    k0, k1 = lo // section, (hi + section - 1) // section
    out = np.empty(hi - lo, dtype=np.uint8)
    jobs = []
    for k in range(k0, k1):
        ln = min(section, size - k * section)
        jobs.append((k, ln))
    if workers > 1 and len(jobs) > 1:
        import multiprocessing as mp
        from multiprocessing import shared_memory
        total = sum(ln for _, ln in jobs)
        shm = shared_memory.SharedMemory(create=True, size=total)
        try:
            offs, o = [], 0
            for _, ln in jobs:
                offs.append(o)
                o += ln
            with mp.get_context("spawn").Pool(min(workers, len(jobs))) as pool:
                # english sections are the slow ones: hand them out first
                order = sorted(range(len(jobs)), key=lambda i: (jobs[i][0] % 3, i))
                list(pool.imap_unordered(_mixed_job, [(jobs[i][0], jobs[i][1], seed, shm.name, offs[i]) for i in order],
                                         chunksize=1))
            full = np.frombuffer(shm.buf, np.uint8, total)
            a = lo - k0 * section
            out[:] = full[a: a + (hi - lo)]
            del full
        finally:
            shm.close()
            shm.unlink()
        return out
    pos = 0
    for k, ln in jobs:
        sec = mixed_section(k, ln, seed)
        a = max(lo, k * section) - k * section
        b = min(hi, k * section + ln) - k * section
        out[pos: pos + (b - a)] = sec[a:b]
        pos += b - a
    return out


def markov_text_lanes(size, seed=42, lanes=4096):
    """Order-2 Markov text like markov_text (27 symbols, Dirichlet(0.15) rows), generated as
    `lanes` independent chains side by side so that it vectorises: 16 MiB in about a second
    instead of a minute.  Not the same bytes as markov_text; same statistics."""
    rng = np.random.default_rng(seed)
    k = 27
    alpha = np.frombuffer(b"abcdefghijklmnopqrstuvwxyz ", dtype=np.uint8)
    cdf = np.cumsum(rng.dirichlet(np.full(k, 0.15), size=k * k), axis=1)
    steps = (size + lanes - 1) // lanes
    out = np.empty((lanes, steps), dtype=np.uint8)
    a = rng.integers(0, k, lanes)
    b = rng.integers(0, k, lanes)
    for i in range(steps):
        u = rng.random(lanes)
        c = np.minimum((cdf[a * k + b] < u[:, None]).sum(axis=1), k - 1)
        out[:, i] = c
        a, b = b, c
    return alpha[out.reshape(-1)[:size]]


def small_cases(seed=0):
    """The five small-input kinds the survey used (Appendix B), 9..3000 bytes."""
    rng = np.random.default_rng(seed)
    cases = []
    for i in range(8):
        n = int(rng.integers(9, 3000))
        cases.append(("binary%d" % i, rng.integers(0, 2, n).astype(np.uint8) + 97))
        cases.append(("random%d" % i, rng.integers(0, 256, n).astype(np.uint8)))
        cases.append(("single%d" % i, np.full(n, int(rng.integers(0, 256)), np.uint8)))
        per = rng.integers(0, 256, int(rng.integers(1, 40))).astype(np.uint8)
        cases.append(("periodic%d" % i, np.resize(per, n)))
        base = rng.integers(0, 4, n).astype(np.uint8) + 65
        if n > 64:
            L = int(rng.integers(8, n // 3))
            s = int(rng.integers(0, n - 2 * L))
            d = int(rng.integers(s + L, n - L + 1))
            base[d: d + L] = base[s: s + L]
        cases.append(("planted%d" % i, base))
    return cases


def fibonacci_word(n):
    """Prefix of the infinite Fibonacci word over {a, b}: the classic worst case of repeat-rich text
    (every prefix doubling round keeps almost all suffixes tied)."""
    a, b = b"a", b"ab"
    while len(b) < n:
        a, b = b, b + a
    return np.frombuffer(b[:n], np.uint8).copy()


def thue_morse(n):
    """Thue-Morse word (cube-free, but full of squares at every scale)."""
    x = np.arange(n, dtype=np.uint64)
    par = np.zeros(n, np.uint8)
    while x.any():
        par ^= (x & 1).astype(np.uint8)
        x >>= 1
    return par + 97


def de_bruijn(k, n):
    """de Bruijn sequence B(k, n): every n-gram over k symbols exactly once (no repeat longer than n - 1)."""
    a = [0] * (k * n)
    seq = []

    def db(t, p):
        if t > n:
            if n % p == 0:
                seq.extend(a[1:p + 1])
        else:
            a[t] = a[t - p]
            db(t + 1, p)
            for j in range(a[t - p] + 1, k):
                a[t] = j
                db(t + 1, t)

    db(1, 1)
    return np.array(seq, np.uint8) + 65


def growing_runs(n):
    """a b aa b aaa b ...: runs of every length, each a prefix of the next."""
    out, k, tot = [], 1, 0
    while tot < n:
        out.append(np.full(k, 97, np.uint8))
        out.append(np.array([98], np.uint8))
        tot += k + 1
        k += 1
    return np.concatenate(out)[:n]


def adversarial_cases(scale=1):
    """Inputs with suffix structure that prefix doubling finds hard; sizes multiply with `scale`."""
    defect = np.resize(np.frombuffer(b"abcabcabd", np.uint8), 250000 * scale).copy()
    defect[len(defect) // 2] = 120          # one defect in the middle of a long period
    two = np.concatenate([fibonacci_word(120000 * scale), thue_morse(90000 * scale), fibonacci_word(100000 * scale)])
    return [("fibonacci", fibonacci_word(300000 * scale)), ("thue_morse", thue_morse(300000 * scale)),
            ("de_bruijn_4_9", de_bruijn(4, 9)), ("growing_runs", growing_runs(200000 * scale)),
            ("sorted_bytes", np.resize(np.arange(256, dtype=np.uint8), 150000 * scale)),
            ("period_defect", defect), ("fib_thue_fib", two)]
