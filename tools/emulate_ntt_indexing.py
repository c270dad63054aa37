"""Index-math emulation of csrc/ntt.cu (tile / pass / twiddle mapping) checked against the oracle NTT.
Development aid only (no GPU in the build container); not part of the product or the tests."""
import sys
sys.path.insert(0, '/root/repo')
import numpy as np
from oracle.ckks_oracle import lib, P

TB = 11

def run(logn, inverse):
    L = lib()
    n = 1 << logn
    q = L.orc_first_prime(30, 2 * n)   # small prime so products fit in uint64? no: use python ints
    psi = L.orc_min_root_of_unity(2 * n, q)
    tab = np.zeros(n, dtype=np.uint64)
    L.orc_make_psi_table(P(tab), logn, psi if not inverse else L.orc_invmod(psi, q), q)
    rng = np.random.default_rng(0)
    a = rng.integers(0, q, n, dtype=np.uint64)
    ref = a.copy()
    (L.orc_ntt_inv if inverse else L.orc_ntt_fwd)(P(ref), logn, q, P(tab))
    W = tab.astype(object)
    x = a.astype(object)
    tb = min(TB, logn)
    s1 = logn - tb
    phases = []
    if s1 > 0:
        phases.append(dict(strided=True, s0=0, ns=s1))
    phases.append(dict(strided=False, s0=s1, ns=tb))
    if inverse:
        phases = phases[::-1]
    for ph in phases:
        ntiles = n >> tb
        for tile in range(ntiles):
            e = np.arange(1 << tb)
            if ph['strided']:
                cb = tb - s1
                g = ((e >> cb) << (logn - s1)) | (tile * (1 << cb) + (e & ((1 << cb) - 1)))
                pre = 0
            else:
                g = tile * (1 << tb) + e
                pre = tile
            v = x[g].copy()
            # passes
            groups = []
            sl = 0
            while sl < ph['ns']:
                S = min(3, ph['ns'] - sl)
                groups.append((sl, S))
                sl += S
            if inverse:
                groups = groups[::-1]
            for (sl_lo, S) in groups:
                lowbits = tb - sl_lo - S
                nu = (1 << tb) >> S
                for u in range(nu):
                    e0 = ((u >> lowbits) << (lowbits + S)) | (u & ((1 << lowbits) - 1))
                    idx = [e0 + (k << lowbits) for k in range(1 << S)]
                    r = [v[i] for i in idx]
                    jr = range(S) if not inverse else range(S - 1, -1, -1)
                    for j in jr:
                        sl_ = sl_lo + j
                        s = ph['s0'] + sl_
                        bit = 1 << (S - 1 - j)
                        for k in range(1 << S):
                            if k & bit:
                                continue
                            ti = (1 << s) + ((pre << sl_) | ((e0 >> (tb - sl_lo)) << j) | (k >> (S - j)))
                            w = int(W[ti])
                            A, B = int(r[k]), int(r[k | bit])
                            if not inverse:
                                T = B * w % q
                                r[k], r[k | bit] = (A + T) % q, (A - T) % q
                            else:
                                r[k], r[k | bit] = (A + B) % q, (A - B) * w % q
                    for i, val in zip(idx, r):
                        v[i] = val
            x[g] = v
    if inverse:
        ninv = pow(n, -1, q)
        x = np.array([int(t) * ninv % q for t in x], dtype=object)
    ok = all(int(p) == int(r) for p, r in zip(x, ref))
    print("logn", logn, "inverse", inverse, "OK" if ok else "MISMATCH")
    return ok

if __name__ == "__main__":
    res = [run(l, inv) for l in (8, 11, 12, 13) for inv in (False, True)]
    sys.exit(0 if all(res) else 1)
