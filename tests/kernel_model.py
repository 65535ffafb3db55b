"""Pure-Python model of the device algorithms in salz_b200/csrc/*.cu.

Not the product and not the oracle: a slow, thread-by-thread transcription of
what the kernels compute, used by the CPU test tier to check the *design*
(prefix doubling with compaction, irreducible-PLCP fill, pyramid ANSV walk,
windowed parse fixpoint, closed-form emit) against the oracle without a GPU.
Only suitable for inputs of a few thousand bytes.
"""
import numpy as np

INF = 0x3FFFFFFF
IMAX = 0x7FFFFFFF
CHUNK = 2048


# ---------------------------------------------------------------- sa.cu ----
def sa_model(blocks, logS=12):
    """blocks: list of bytes objects (the n-byte texts to sort).  Returns list of SA arrays."""
    S = 1 << logS
    B = len(blocks)
    bbits = max(0, (B - 1).bit_length())
    h0 = min(8, (64 - bbits) // 8)
    ns = [len(t) for t in blocks]
    keys, vals, slots = [], [], []
    for b, t in enumerate(blocks):
        n = ns[b]
        for r in range(n):          # descending position order inside a block
            p = n - 1 - r
            chunk = t[p:p + h0]
            chunk = chunk + b"\0" * (h0 - len(chunk))
            k = int.from_bytes(chunk, "big") | ((b << (8 * h0)) if h0 < 8 else 0)
            keys.append(k); vals.append((b << logS) | p); slots.append((b << logS) | r)
    rank = {}
    SA = {}
    rounds = 0

    def rerank(keys, vals, slots, sec_mask):
        cnt = len(keys)
        head = [k == 0 or keys[k] != keys[k - 1] or (sec_mask and (keys[k] & sec_mask) == 0)
                for k in range(cnt)] + [True]
        nv, nsl, npr = [], [], []
        cur = None
        for k in range(cnt):
            if head[k]:
                cur = slots[k]
            rank[vals[k]] = cur
            if head[k] and head[k + 1]:
                SA[slots[k]] = vals[k] & (S - 1)
            else:
                nv.append(vals[k]); nsl.append(slots[k]); npr.append(cur)
        return nv, nsl, npr

    order = sorted(range(len(keys)), key=lambda i: keys[i])  # python sort is stable
    keys = [keys[i] for i in order]; vals = [vals[i] for i in order]
    vals, slots, prim = rerank(keys, vals, slots, 0)
    h = h0
    while vals:
        assert h < S, "no convergence"
        keys = []
        for k, g in enumerate(vals):
            b, p = g >> logS, g & (S - 1)
            sec = (rank[g + h] & (S - 1)) + 1 if p + h < ns[b] else 0
            keys.append((prim[k] << logS) | sec)
        order = sorted(range(len(keys)), key=lambda i: keys[i])
        keys = [keys[i] for i in order]; vals = [vals[i] for i in order]
        vals, slots, prim = rerank(keys, vals, slots, S - 1)
        h *= 2
        rounds += 1
    out = []
    for b in range(B):
        out.append(np.array([SA[(b << logS) | r] for r in range(ns[b])], dtype=np.int32))
    return out, rounds


# --------------------------------------------------------------- lcp.cu ----
def plcp_model(T, SA):
    n = len(SA)
    isa = np.zeros(n, np.int64); isa[SA] = np.arange(n)
    V = np.zeros(n, np.int64)
    for p in range(n):
        r = isa[p]
        if r == 0:
            V[p] = p
            continue
        q = int(SA[r - 1])
        if p == 0 or q == 0 or T[p - 1] != T[q - 1]:
            cap = n - max(p, q)
            l = 0
            while l < cap and T[p + l] == T[q + l]:
                l += 1
            V[p] = l + p
    plcp = np.maximum.accumulate(V) - np.arange(n)
    lcp = np.zeros(n, np.int64); lcp[isa] = plcp
    return plcp.astype(np.int32), lcp.astype(np.int32)


# ------------------------------------------------------------- parse.cu ----
def build_pyramids(SA, LCP, logS):
    S = 1 << logS
    n = len(SA)
    sa = [np.full(S, IMAX, np.int64)]; lc = [np.full(S, IMAX, np.int64)]
    sa[0][:n] = SA; lc[0][:n] = LCP
    l = 0
    while (S >> (5 * l)) > 32:
        sa.append(sa[l].reshape(-1, 32).min(axis=1)); lc.append(lc[l].reshape(-1, 32).min(axis=1))
        l += 1
    return sa, lc, l


def ansv_model(SA, LCP, logS=12):
    n = len(SA)
    sa, lc, top = build_pyramids(SA, LCP, logS)
    cand = np.zeros((n, 4), np.int32)
    for r in range(n):
        v = int(SA[r])
        # PSV
        m = int(lc[0][r]); lvl = 0; pos = r; fl = -1; fj = 0
        while True:
            start = pos & ~31
            j = pos - 1
            while j >= start:
                if sa[lvl][j] < v:
                    fl, fj = lvl, j
                    break
                m = min(m, int(lc[lvl][j])); j -= 1
            if fl >= 0 or lvl == top:
                break
            pos >>= 5; lvl += 1
        psv, pl = -1, 0
        if fl >= 0:
            while fl > 0:
                fl -= 1
                c0 = fj << 5
                for c in range(c0 + 31, c0 - 1, -1):
                    if sa[fl][c] < v:
                        fj = c
                        break
                    m = min(m, int(lc[fl][c]))
            psv, pl = int(sa[0][fj]), m
        # NSV
        m = IMAX; lvl = 0; pos = r; fl = -1; fj = 0
        while True:
            size = 1 << (logS - 5 * lvl)
            end = min(pos | 31, size - 1)
            j = pos + 1
            while j <= end:
                if sa[lvl][j] < v:
                    fl, fj = lvl, j
                    break
                m = min(m, int(lc[lvl][j])); j += 1
            if fl >= 0 or lvl == top:
                break
            pos >>= 5; lvl += 1
        nsv, nl = -1, 0
        if fl >= 0:
            while fl > 0:
                fl -= 1
                c0 = fj << 5
                for c in range(c0, c0 + 32):
                    if sa[fl][c] < v:
                        fj = c
                        break
                    m = min(m, int(lc[fl][c]))
            nsv, nl = int(sa[0][fj]), min(m, int(lc[0][fj]))
        cand[v] = (-1, 1, -1, 1) if v == 0 else (v - psv, pl, v - nsv, nl)
    return cand


def vn_count(v):
    k = 1
    for lim in (8, 72, 584, 4680, 37448, 299592, 2396744):
        if v >= lim:
            k += 1
    return k


def pre_bits(o):
    return 1 + 4 * vn_count((o - 1) >> 8)


def post_bits(l):
    return ((l - 3) >> 3) + 4


def parse_model(cand, N):
    """Window/lane transcription of k_parse.  Returns cost[n+1], choice[N,2], jump[N]."""
    n = N - 8
    cost = np.zeros(N + 64, np.int64)
    choice = np.zeros((N, 2), np.int64)
    jump = np.zeros(N, np.int64)
    ring = np.zeros(CHUNK, np.int64)
    cost_right = INF
    iters_max = 0
    for w in range((N - 1) >> 5, -1, -1):
        q = (w << 5) + 31
        P = [q - l for l in range(32)]
        chunk_end = ((P[0] >> 11) + 1) << 11
        a1 = [INF] * 32; a2 = [INF] * 32; l1 = [0] * 32; l2 = [0] * 32; o1 = [0] * 32; o2 = [0] * 32
        f1 = [INF] * 32; f2 = [INF] * 32; n1 = [False] * 32; n2 = [False] * 32
        x = [INF] * 32
        for l, p in enumerate(P):
            if 1 <= p < n:
                c = cand[p]
                if c[1] >= 3:
                    o1[l], l1[l] = int(c[0]), int(c[1]); a1[l] = pre_bits(o1[l]) + 8 + post_bits(l1[l])
                    if l1[l] <= l: n1[l] = True
                    else: f1[l] = a1[l] + int(cost[p + l1[l]])
                if c[3] >= 3:
                    o2[l], l2[l] = int(c[2]), int(c[3]); a2[l] = pre_bits(o2[l]) + 8 + post_bits(l2[l])
                    if l2[l] <= l: n2[l] = True
                    else: f2[l] = a2[l] + int(cost[p + l2[l]])
            x[l] = 0 if p == n else min(f1[l], f2[l])
        xfar = list(x)
        any_near = any(n1) or any(n2)
        it = 0
        while True:
            it += 1
            y = [x[l] - 9 * l for l in range(32)]
            for l in range(1, 32):
                y[l] = min(y[l], y[l - 1])
            cst = [min(y[l] + 9 * l, cost_right + 9 * (l + 1)) for l in range(32)]
            for l, p in enumerate(P):
                if p > n: cst[l] = INF
            if not any_near:
                break
            changed = False
            for l in range(32):
                nx = xfar[l]
                if n1[l]: nx = min(nx, a1[l] + cst[l - l1[l]])
                if n2[l]: nx = min(nx, a2[l] + cst[l - l2[l]])
                if nx < x[l]: changed = True
                x[l] = min(x[l], nx)
            if not changed:
                break
        iters_max = max(iters_max, it)
        t = [0] * 32
        for l, p in enumerate(P):
            right = cost_right if l == 0 else cst[l - 1]
            c1 = a1[l] + cst[l - l1[l]] if n1[l] else f1[l]
            c2 = a2[l] + cst[l - l2[l]] if n2[l] else f2[l]
            best, bo, bl = 9 + right, 0, 1
            if l1[l] >= 3 and c1 < best: best, bo, bl = c1, o1[l], l1[l]
            if l2[l] >= 3 and c2 < best: best, bo, bl = c2, o2[l], l2[l]
            if p == 0 or p >= n: bo, bl = 0, 1
            t[l] = p + bl
            if p < N:
                if p <= n: cost[p] = cst[l]
                choice[p] = (bo, bl)
        for _ in range(5):
            t = [t[q - t[l]] if (t[l] <= q and t[l] < N) else t[l] for l in range(32)]
        for l, p in enumerate(P):
            j = t[l]
            if t[l] < chunk_end and t[l] < N:
                j = ring[t[l] & (CHUNK - 1)]
            t[l] = j
        for l, p in enumerate(P):
            ring[p & (CHUNK - 1)] = t[l]
            if p < N: jump[p] = t[l]
        cost_right = cst[31]
    return cost[: n + 1], choice, jump, iters_max


# -------------------------------------------------------------- emit.cu ----
def emit_model(T, N, choice, jump):
    """Chunk walk + closed-form layout.  Returns the encoded block (bytes)."""
    NONE = -1
    nch = (N + CHUNK - 1) // CHUNK
    entry = [NONE] * nch
    e = 0
    while e < N:
        entry[e >> 11] = e
        e = int(jump[e])

    def tok_bits(o, l):
        return (1, 0) if l == 1 else (pre_bits(o), post_bits(l))

    ctok = [0] * nch; cbit = [0] * nch
    for c in range(nch):
        if entry[c] == NONE: continue
        p = entry[c]; end = min(N, ((p >> 11) + 1) << 11)
        while p < end:
            o, l = int(choice[p][0]), int(choice[p][1])
            pr, po = tok_bits(o, l)
            ctok[c] += 1; cbit[c] += pr + po; p += l
    m = sum(ctok); Bm = sum(cbit)
    tb = np.cumsum([0] + ctok[:-1]); bb = np.cumsum([0] + cbit[:-1])
    salz = 4 + m + 8 * ((Bm + 63) // 64)
    if salz > N + 4:
        hdr = N & 0xFFFFFF
        return hdr.to_bytes(4, "little") + bytes(T[:N])
    out = bytearray(salz)
    written = np.zeros(salz, np.int32)
    tokPos = [0] * m; tokBit = [0] * m; first = [0] * ((Bm + 63) // 64)
    for c in range(nch):
        if entry[c] == NONE: continue
        p = entry[c]; end = min(N, ((p >> 11) + 1) << 11)
        t, Bt = int(tb[c]), int(bb[c])
        while p < end:
            o, l = int(choice[p][0]), int(choice[p][1])
            pr, po = tok_bits(o, l)
            tokPos[t] = p; tokBit[t] = Bt
            bp = 4 + t + 8 * ((Bt + pr + 63) >> 6)
            out[bp] = T[p] if l == 1 else ((o - 1) & 0xFF); written[bp] += 1
            Bn = Bt + pr + po
            for k in range((Bt + 63) >> 6, (Bn + 63) >> 6): first[k] = t
            Bt = Bn; t += 1; p += l

    def place(w, w0, s, nb, val):
        rel = s + nb - 1 - w0
        if rel < 0: return w
        pos = 63 - rel
        if pos >= 0: return (w | (val << pos)) & 0xFFFFFFFFFFFFFFFF
        if -pos < nb: return w | (val >> (-pos))
        return w

    for k in range((Bm + 63) // 64):
        w0 = k << 6; t0 = first[k]; w = 0; y = t0
        t = t0
        while t < m and tokBit[t] < w0 + 64:
            Bt = tokBit[t]; o, l = int(choice[tokPos[t]][0]), int(choice[tokPos[t]][1])
            if l != 1:
                v = (o - 1) >> 8
                pv = (v & 7) | 8; nn = 1; v = (v >> 3) - 1
                while v >= 0:
                    pv |= (v & 7) << (4 * nn); nn += 1; v = (v >> 3) - 1
                pr = 1 + 4 * nn; pv |= 1 << (4 * nn)
                w = place(w, w0, Bt, pr, pv)
                if t == t0 and w0 >= Bt + pr: y = t0 + 1
                qq = (l - 3) >> 3
                w = place(w, w0, Bt + pr + qq, 4, 8 | ((l - 3) & 7))
            t += 1
        pos = 4 + 8 * k + y
        out[pos:pos + 8] = w.to_bytes(8, "little"); written[pos:pos + 8] += 1
    hdr = (1 << 24) | ((salz - 4) & 0xFFFFFF)
    out[0:4] = hdr.to_bytes(4, "little"); written[0:4] += 1
    assert (written == 1).all(), "every output byte must be written exactly once"
    return bytes(out)
