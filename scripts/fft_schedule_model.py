# Model of the register-blocked kissfft-DAG schedule: validates digits/twiddles/padding and bank conflicts.
import numpy as np, itertools, sys

def radices(sf):
    r = [4]*(sf//2)
    if sf & 1: r.append(2)
    return r

PASSES = {  # stage groups inner->outer expressed as lists of stage index ranges [a,b]
}
def passes_for(sf):
    r = radices(sf); L = len(r)
    # innermost pass: last two stages; then pairs of radix-4 stages; a final single stage if left over
    out = []
    b = L-1
    while b >= 0:
        a = max(b-1, 0)
        out.append((a, b))
        b = a-1
    return out

def run(sf, pad):
    N = 1 << sf; r = radices(sf); L = len(r)
    m = [N // int(np.prod(r[:s+1])) for s in range(L)]
    F = [int(np.prod(r[:s])) for s in range(L)]
    tw = np.exp(-2j*np.pi*np.arange(N)/N)
    x = np.random.randn(N) + 1j*np.random.randn(N)
    T = N // 16
    smem = np.zeros(4*N, complex)
    ps = passes_for(sf)
    regs = np.zeros((T, 16), complex)
    conflicts = {}
    def bank_conf(addrs, tag):
        # addrs: per-lane float2 index for one LDS.64/STS.64 by up to 32 lanes; phases of 16 lanes
        worst = 1
        for h in range(0, len(addrs), 16):
            a = addrs[h:h+16]
            banks = {}
            for v in set(a): banks.setdefault(v % 16, set()).add(v)
            worst = max(worst, max(len(s) for s in banks.values()))
        conflicts[tag] = max(conflicts.get(tag, 1), worst)
    for pi, (a, b) in enumerate(ps):
        R = int(np.prod(r[a:b+1])); G = 16 // R; mb = m[b]
        first, last = pi == 0, pi == len(ps)-1
        # element digit enumeration in position order: e = sum d_s * m_s/m_b
        def digits_of_e(e):
            d = {}
            for s in range(a, b+1):
                w = m[s] // mb
                d[s] = (e // w) % r[s]
            return d
        xbase = np.zeros((T, G), int)
        for t in range(T):
            for g in range(G):
                if first:
                    gin = t + T*g  # input base index, digits d_0..d_{a-1} with F weights
                    X = 0; rem = gin
                    for s in range(a):
                        X += (rem % r[s]) * m[s]; rem //= r[s]
                    assert rem == 0
                    xbase[t, g] = X
                    Fa = F[a]
                    for e in range(R):
                        d = digits_of_e(e)
                        jj = sum(d[s]*F[s] for s in range(a, b+1)) // Fa
                        i = gin + Fa*jj
                        regs[t, g*R+e] = x[i]
                else:
                    gam = t + T*g
                    klow = gam % mb; H = gam // mb
                    xbase[t, g] = H*R*mb + klow
        if not first:
            for g in range(G):
                for e in range(R):
                    addrs = [pad(xbase[t, g] + e*mb) for t in range(T)]
                    for w in range(0, T, 32): bank_conf(addrs[w:w+32], f"sf{sf} pass{pi} load")
                    for t in range(T): regs[t, g*R+e] = smem[addrs[t]]
        # stages innermost first
        for s in range(b, a-1, -1):
            w = m[s] // mb
            for t in range(T):
                for g in range(G):
                    klow = xbase[t, g] % mb if not first else 0
                    for e0 in range(R):
                        d = digits_of_e(e0)
                        if d[s] != 0: continue
                        cpart = sum(d[u]*m[u] for u in range(s+1, b+1))
                        k = cpart + klow
                        idx = [g*R + e0 + q*w for q in range(r[s])]
                        v = [regs[t, i]*tw[k*F[s]*q] for q, i in enumerate(idx)]
                        if r[s] == 2:
                            regs[t, idx[0]], regs[t, idx[1]] = v[0]+v[1], v[0]-v[1]
                        else:
                            s5 = v[0]-v[2]; g0 = v[0]+v[2]; s3 = v[1]+v[3]; dd = v[1]-v[3]; s4 = -1j*dd
                            regs[t, idx[0]] = g0+s3; regs[t, idx[2]] = g0-s3; regs[t, idx[1]] = s5+s4; regs[t, idx[3]] = s5-s4
        if not last:
            for g in range(G):
                for e in range(R):
                    addrs = [pad(xbase[t, g] + e*mb) for t in range(T)]
                    for w in range(0, T, 32): bank_conf(addrs[w:w+32], f"sf{sf} pass{pi} store")
                    for t in range(T): smem[addrs[t]] = regs[t, g*R+e]
        else:
            X = np.zeros(N, complex)
            for t in range(T):
                for g in range(G):
                    for e in range(R): X[xbase[t, g] + e*mb] = regs[t, g*R+e]
            err = np.abs(X - np.fft.fft(x)).max()
            return err, conflicts, ps

if __name__ == "__main__":
    for sf in range(7, 13):
        for name, pad in (("none", lambda X: X), ("x>>4", lambda X: X + (X >> 4)), ("x>>8", lambda X: X + (X >> 8)), ("x>>4+x>>8", lambda X: X + (X >> 4) + (X >> 8))):
            err, conf, ps = run(sf, pad)
            print(sf, ps, name, f"err={err:.1e}", {k.split(' ',1)[1]: v for k, v in conf.items()})

def conflicts_multi(sf, pad, woff):
    """bank conflicts when 32/T windows share a warp (T<32); lanes = (win, t)"""
    N = 1 << sf; r = radices(sf); L = len(r)
    m = [N // int(np.prod(r[:s+1])) for s in range(L)]
    T = N // 16; W = max(1, 32 // T)
    ps = passes_for(sf); worst = {}
    size = pad(N-1) + 1
    wstride = size + ((woff - size) % 16)
    for pi, (a, b) in enumerate(ps):
        R = int(np.prod(r[a:b+1])); G = 16 // R; mb = m[b]
        for g in range(G):
            for e in range(R):
                for kind in ("store", "load"):
                    if kind == "store" and pi == len(ps)-1: continue
                    if kind == "load" and pi == 0: continue
                    use_pi = pi
                    addrs = []
                    for w in range(W):
                        for t in range(T):
                            if pi == 0:
                                gin = t + T*g; X = 0; rem = gin
                                for s in range(a): X += (rem % r[s]) * m[s]; rem //= r[s]
                            else:
                                gam = t + T*g; X = (gam // mb)*R*mb + gam % mb
                            addrs.append(w*wstride + pad(X + e*mb))
                    for h in range(0, len(addrs), 16):
                        aa = addrs[h:h+16]; banks = {}
                        for v in set(aa): banks.setdefault(v % 16, set()).add(v)
                        c = max(len(s) for s in banks.values())
                        key = f"p{pi}{kind}"; worst[key] = max(worst.get(key, 1), c)
    return worst, wstride

if __name__ == "__main__":
    print("multi-window search")
    for sf in (7, 8):
        best = None
        for c3, c4, c5, c8 in itertools.product(range(0, 3), range(0, 3), range(0, 2), range(0, 2)):
            pad = lambda X: X + c3*(X >> 3) + c4*(X >> 4) + c5*(X >> 5) + c8*(X >> 8)
            for woff in range(16):
                w, ws = conflicts_multi(sf, pad, woff)
                score = sum(w.values())
                if best is None or score < best[0] or (score == best[0] and pad((1<<sf)-1) < best[5]):
                    best = (score, (c3, c4, c5, c8), woff, w, ws, pad((1<<sf)-1))
        print(sf, best)
