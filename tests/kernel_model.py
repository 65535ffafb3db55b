"""Pure-Python model of the device algorithms in salz_b200/csrc/*.cu.

Not the product and not the oracle: a slow, thread-by-thread transcription of
what the kernels compute, used by the CPU test tier to check the *design*
(prefix doubling with compaction, irreducible-PLCP fill, pyramid ANSV walk,
segmented shifted-cost parse, closed-form emit) against the oracle without a GPU.
Only suitable for inputs of a few thousand bytes.
"""
import numpy as np

INF = 0x3FFFFFFF
IMAX = 0x7FFFFFFF
CHUNK = 2048


# ---------------------------------------------------------------- sa.cu ----
def sa_model(blocks, logS=12, window=48, seed=0):
    """Model of sa.cu for the packed format: dense codes from 1 (k_alphabet), round 0
    on the first h0 characters, then rounds of k_local over windows of `window` list
    elements (groups inside a window are sorted and re-ranked locally, groups that
    cross a window boundary take the global path), u64-style stamped ranks read and
    written in the same launch, and a bag whose chunks arrive in arbitrary order.
    blocks: list of bytes objects (the n-byte texts).  Returns (list of SA arrays, rounds)."""
    import random
    rnd = random.Random(seed)
    out, rounds_max = [], 0
    for t in blocks:
        n = len(t)
        if n == 0:
            out.append(np.zeros(0, np.int32)); continue
        syms = sorted(set(t))
        code = {c: i + 1 for i, c in enumerate(syms)}          # 0 = past the end
        bits = max(1, len(syms).bit_length())
        h0 = min(40, (2 * logS) // bits)
        SA = [None] * n
        rank = [None] * n                                        # (stamp, old, new)

        def rerank(sorted_elems, rnd_no):
            """sorted_elems: list of (group, sec, p) sorted by (group, sec); returns survivors in order."""
            cnt = len(sorted_elems)
            surv = []
            k = 0
            while k < cnt:
                g = sorted_elems[k][0]
                e = k
                while e < cnt and sorted_elems[e][0] == g:
                    e += 1
                # old group [k, e): split by sec (sec == 0 is always its own head)
                i = k
                while i < e:
                    j = i + 1
                    if sorted_elems[i][1] != 0:
                        while j < e and sorted_elems[j][1] == sorted_elems[i][1]:
                            j += 1
                    r = g + (i - k)
                    for q in range(i, j):
                        pp = sorted_elems[q][2]
                        rank[pp] = (rnd_no, g, r)
                        if j - i == 1:
                            SA[r] = pp
                        else:
                            surv.append((r, pp))
                    i = j
                k = e
            return surv

        # round 0: global sort on the code key
        def key0(p):
            k = 0
            for c in range(h0):
                k = (k << bits) | (code[t[p + c]] if p + c < n else 0)
            return k
        elems = sorted(((0, key0(p) + 1, p) for p in range(n)), key=lambda x: (x[0], x[1]))
        # (+1 keeps "sec == 0 is a head" out of round 0: its keys are plain characters)
        bag = rerank(elems, 0)
        h, rnd_no = h0, 1
        while bag:
            assert h < (1 << logS), "no convergence"
            def read(q):
                st, old, new = rank[q]
                return old if st == rnd_no else new
            tiles = [bag[i:i + window] for i in range(0, len(bag), window)]
            cross = []          # global-path list, pieces in arbitrary order
            nxt_chunks = []
            order = list(range(len(tiles)))
            rnd.shuffle(order)  # CTAs run in any order: reads see a mix of refined and unrefined ranks
            for ti in order:
                tile = tiles[ti]
                prevG = tiles[ti - 1][-1][0] if ti > 0 else None
                nextG = tiles[ti + 1][0][0] if ti + 1 < len(tiles) else None
                gF, gL = tile[0][0], tile[-1][0]
                keyed = [(g, (read(pp + h) + 1) if pp + h < n else 0, pp) for g, pp in tile]
                own, fwd = [], []
                for e in keyed:
                    if (prevG == gF and e[0] == gF) or (nextG == gL and e[0] == gL):
                        fwd.append(e)
                    else:
                        own.append(e)
                cross.append(fwd)
                own.sort(key=lambda x: (x[0], x[1], x[2]))
                nxt_chunks.append(rerank(own, rnd_no))
            rnd.shuffle(cross)
            glob = sorted((e for piece in cross for e in piece), key=lambda x: (x[0], x[1]))
            nxt_chunks.append(rerank(glob, rnd_no))
            rnd.shuffle(nxt_chunks)   # groups stay contiguous, their order is arbitrary
            # the global path's survivors form ONE chunk: groups crossing rerank tiles stay contiguous
            bag = [e for ch in nxt_chunks for e in ch]
            h *= 2
            rnd_no += 1
        rounds_max = max(rounds_max, rnd_no - 1)
        out.append(np.array(SA, dtype=np.int32))
    return out, rounds_max


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


def parse_model(cand, N, seg_len=256, max_passes=3, warm=0):
    """Model of parse.cu: the chain in the shifted space d[p] = cost[p] + 9p (one min per
    position), the block cut into segments of `seg_len` positions that keep d relative to
    their right end and read everything to their right from the PREVIOUS pass's array
    (k_parse + k_seg_scan), the parallel check that the recurrence holds everywhere
    (k_choose) from the second pass on, and after `max_passes` the plain single chain.
    warm > 0: the first pass walks only the leftmost `warm` positions of every segment
    (values relative to the end of that window, the rest of the array stays 0).
    Returns cost[n+1], choice[N,2], jump[N], passes (0 = fell back to one chain)."""
    n = N - 8
    w = {}
    for p in range(1, n):
        c = cand[p]
        for (o, l) in ((int(c[0]), int(c[1])), (int(c[2]), int(c[3]))):
            if l >= 3:
                w[(p, l, o)] = pre_bits(o) + 8 + post_bits(l) - 9 * l

    def edges(p):
        c = cand[p]
        return [(o, l) for (o, l) in ((int(c[0]), int(c[1])), (int(c[2]), int(c[3]))) if l >= 3]

    def run_pass(seg, old, base, window=0):
        K = (N + seg - 1) // seg
        new = np.zeros(N + 1, np.int64)
        for j in range(K):                       # independent CTAs
            L, R = j * seg, min(N, (j + 1) * seg)
            if window:
                R = min(R, L + window)
            d = INF if R == N else 0
            for p in range(R - 1, L - 1, -1):
                if p > n:
                    continue
                m = 0 if p == n else INF
                if 1 <= p < n:
                    for (o, l) in edges(p):
                        t = p + l
                        v = new[t] if t < R else old[t] + base[t // seg] - base[j]
                        m = min(m, v + w[(p, l, o)])
                d = min(d, m)
                new[p] = d
        nb = np.zeros(K, np.int64)               # k_seg_scan
        b = 9 * n
        for j in range(K - 1, -1, -1):
            nb[j] = b
            b += new[j * seg]
        return new, nb

    def choose(seg, dr, base):
        """k_choose: choices + number of positions whose recurrence fails."""
        choice = np.zeros((N, 2), np.int64)
        choice[:, 1] = 1
        bad = 0
        dabs = lambda t: dr[t] + base[t // seg]
        for p in range(0, n + 1):
            best, bo, bl = 9 * n, 0, 1
            if p < n:
                best = dabs(p + 1)
                if p >= 1:
                    for (o, l) in edges(p):
                        v = dabs(p + l) + w[(p, l, o)]
                        if v < best:
                            best, bo, bl = v, o, l
            bad += dabs(p) != best
            choice[p] = (bo, bl)
        return choice, bad

    K = (N + seg_len - 1) // seg_len
    passes, done = 0, False
    seg = seg_len
    if K > 1:
        old = np.zeros(N + 1, np.int64)          # all-literal estimate: d = base everywhere
        base = np.full(K, 9 * n, np.int64)
        for passes in range(1, max_passes + (2 if warm else 1)):
            dr, base = run_pass(seg, old, base, warm if passes == 1 else 0)
            if passes >= 2:
                choice, bad = choose(seg, dr, base)
                if bad == 0:
                    done = True
                    break
            old = dr
    if not done:
        passes, seg = 0, 1 << 30
        dr, base = run_pass(seg, np.zeros(N + 1, np.int64), np.full(1, 9 * n, np.int64))
        choice, bad = choose(seg, dr, base)
        assert bad == 0
    cost = np.array([dr[p] + base[p // seg] - 9 * p for p in range(n + 1)], np.int64)
    jump = np.zeros(N, np.int64)
    for p in range(N - 1, -1, -1):
        end = min(N, ((p >> 11) + 1) << 11)
        t = p + int(choice[p][1])
        jump[p] = t if t >= end else jump[t]
    return cost, choice, jump, passes


# -------------------------------------------------------------- emit.cu ----
def chunk_entries_model(jump, N, chunk=CHUNK, lanes=32):
    """k_chunk_entries: the chain e_0 = 0, e_{i+1} = jump[e_i] by exact speculation.  Lane j
    walks range j of the chunks from the range's first position; the true chain takes a
    lane's walk over where it steps on a position that lane has visited and walks (and
    erases) by itself until then.  Returns entry[chunk] (-1 = no token starts there)."""
    NONE = -1
    nch = (N + chunk - 1) // chunk
    per = (nch + lanes - 1) // lanes
    ce = [NONE] * nch
    ex = [0] * lanes
    for j in range(lanes):                      # phase 1, all lanes at once
        e, lim = j * per * chunk, min(N, (j + 1) * per * chunk)
        while e < lim:
            ce[e // chunk] = e
            e = int(jump[e])
        ex[j] = e
    x, clean = ex[0], min(nch, per)             # phase 2, lane 0
    while x < N:
        c = x // chunk
        j = c // per
        rend = min(nch, (j + 1) * per)
        for k in range(clean, c):
            ce[k] = NONE
        while True:
            if ce[c] == x:
                x = ex[j]
                break
            nx = int(jump[x])
            c2 = nx // chunk if nx < N else nch
            ce[c] = x
            for k in range(c + 1, min(c2, rend)):
                ce[k] = NONE
            x, c = nx, c2
            if c >= rend:
                break
        clean = rend
    for k in range(clean, nch):
        ce[k] = NONE
    return ce


def emit_model(T, N, choice, jump):
    """Chunk walk + closed-form layout.  Returns the encoded block (bytes)."""
    NONE = -1
    nch = (N + CHUNK - 1) // CHUNK
    entry = [NONE] * nch
    e = 0
    while e < N:
        entry[e >> 11] = e
        e = int(jump[e])
    assert entry == chunk_entries_model(jump, N)

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


# ---------------------------------------------------- sa.cu: radix pass ----
def sched_interleave_model(tile_counts, group):
    """k_sched_interleave: ticket order of a single-pass launch.  tile_counts[b] = tiles of block b
    (tile ids are block-major).  Inside a group of `group` consecutive blocks tickets run over
    (tile index in its block, block).  Returns sched (ticket -> tile id)."""
    base = np.concatenate([[0], np.cumsum(tile_counts)]).astype(np.int64)
    n = int(base[-1])
    sched = np.full(n, -1, np.int64)
    B = len(tile_counts)
    for b in range(B):
        g0 = b - b % group
        g1 = min(g0 + group, B)
        for lt in range(tile_counts[b]):
            t = base[g0]
            for bb in range(g0, g1):
                cnt = tile_counts[bb]
                t += min(cnt, lt) + (1 if (bb < b and cnt > lt) else 0)
            assert sched[t] == -1
            sched[t] = base[b] + lt
    return sched


def scatter_pass_model(keys, shift, tile=96, warps=3, lane_order=None):
    """One stable 8-bit LSD pass over ONE block's list the way k_scatter_tma<5> does it (optimistic,
    verified ranks).  Per tile: warp w holds tile indices w*W + step*32 + lane (W = tile / warps);
    the counting atomicAdd of a step hands out slots of (warp, digit) in `lane_order` (None = lane
    order, what the hardware does; anything else models hardware that serialises differently);
    a key is staged as its tile index at digit start + warp prefix + slot; the write-out fetches
    the keys through the staged indices and checks that the indices of every digit run increase.
    Returns (keys after the pass, number of tiles whose order check failed)."""
    keys = np.asarray(keys, np.uint64)
    n = len(keys)
    out = np.zeros(n, np.uint64)
    dig_all = ((keys >> np.uint64(shift)) & np.uint64(255)).astype(np.int64)
    digit_base = np.concatenate([[0], np.cumsum(np.bincount(dig_all, minlength=256))])[:256]
    prev = np.zeros(256, np.int64)            # what the decoupled look-back delivers: earlier tiles' counts
    per_warp = tile // warps
    steps = per_warp // 32
    unordered = 0
    for t0 in range(0, n, tile):
        valid = min(tile, n - t0)
        tk = np.full(tile, np.uint64(2**64 - 1), np.uint64)    # padding keys: digit 255, after every real key
        tk[:valid] = keys[t0:t0 + valid]
        d = ((tk >> np.uint64(shift)) & np.uint64(255)).astype(np.int64)
        cnt = np.zeros((warps, 256), np.int64)
        slot = np.zeros(tile, np.int64)
        for w in range(warps):
            for s in range(steps):
                lanes = list(range(32)) if lane_order is None else list(lane_order)
                for lane in lanes:                              # the order in which the atomic unit serves the lanes
                    i = w * per_warp + s * 32 + lane
                    slot[i] = cnt[w, d[i]]
                    cnt[w, d[i]] += 1
        total = cnt.sum(axis=0)
        dstart = np.concatenate([[0], np.cumsum(total)])[:256]
        first = dstart[None, :] + np.cumsum(cnt, axis=0) - cnt   # digit start + warps before mine
        sidx = np.full(tile, -1, np.int64)
        for i in range(tile):
            pos = first[i // per_warp, d[i]] + slot[i]
            assert sidx[pos] == -1
            sidx[pos] = i
        bad = False
        for j in range(valid):
            si = sidx[j]
            if si >= valid or (j > 0 and d[sidx[j - 1]] == d[si] and sidx[j - 1] >= si):
                bad = True
            dj = d[si]
            out[digit_base[dj] + prev[dj] + (j - dstart[dj])] = tk[si]
        unordered += bad
        total[255] -= tile - valid
        prev += total
    return out, unordered
