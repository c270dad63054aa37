"""oracle/bootstrap.py -- TEST INFRASTRUCTURE ONLY (CPU oracle).

CKKS bootstrapping behind OpenFHE's EvalBootstrapSetup / EvalBootstrapKeyGen / EvalBootstrap, as the reference calls them
(src/k-way/EvalUtils.cpp:57-86, src/sign.cpp:164-171, src/kway_adapter.h:55-61; fixtures tests/k-way/KWaySort2Test.cpp:47-49,
tests/k-way/EvalUtilsTest.cpp:34-35,132-158).  OpenFHE 1.1.4's FHECKKSRNS is absent offline, so this restates the published
algorithm [CHKKS18, HK20, BMTH21] with OpenFHE's parameterisation for UNIFORM_TERNARY secrets (K = 512, R = 6 double-angle
steps) and sparse packing (slots <= ring/4):

  adjust     bring the ciphertext to its last limb q0 with plaintext integer m * q0 * 2^-corr (one constant mult + rescale)
  ModRaise   read the q0 residues as centred integers in every limb of level `raise_level`: plaintext m q0 2^-c + q0 I(X)
  SubSum     log2(ring / 2 slots) rotations by slots * 2^j: projects I(X) onto the packing subring
  C2S        U^-1 as `levelBudget[0]` groups of inverse special-FFT layers (generalised diagonals, hoisted rotations); the
             last group is evaluated twice (G and -i G) so that u + conj(u) gives Re and Im of the coefficient pairs / K
  EvalMod    per part: Chebyshev series (degree 119) of cos((2 pi K u - pi/2) / 2^R), then R times c <- 2 c^2 - 1
  S2C        U as `levelBudget[1]` groups; the first one is applied to both parts (G and i G) and summed

The engine twin is sorting-fhe_b200/csrc/bootstrap.cu: same op sequence, same constants (the FFT-layer products are
formed with the same elementary double operations in the same order, so the encoded diagonals agree bit for bit).
"""
import math

import numpy as np


K_UNIFORM = 512          # bound on |I| + 1 for UNIFORM_TERNARY secrets (OpenFHE K_UNIFORM)
R_UNIFORM = 6            # double-angle iterations (OpenFHE R_UNIFORM)
CHEB_DEGREE = 119        # same Paterson-Stockmeyer depth (8 - 1) as OpenFHE's degree-89 table
CORRECTION = 10          # message scaled by 2^-10 relative to q0 before the modular reduction


def boot_cheb_coefficients():
    """c_i = (2 / n) sum_j f(u_j) cos(step i (j + 1/2)), u_j = cos(step (j + 1/2)), f(u) = cos((2 pi K u - pi/2) / 2^R):
    libm calls and plain left-to-right sums, the same operations in the same order as csrc/bootstrap.cu"""
    n = CHEB_DEGREE + 1
    step = math.pi / float(n)
    f = []
    for j in range(n):
        u = math.cos(step * (float(j) + 0.5))
        f.append(math.cos((2.0 * math.pi * float(K_UNIFORM) * u - 0.5 * math.pi) / float(1 << R_UNIFORM)))
    c = []
    for i in range(n):
        acc = 0.0
        for j in range(n):
            acc += f[j] * math.cos(step * float(i) * (float(j) + 0.5))
        c.append(acc * 2.0 / float(n))
    return c


def layer_groups(log_s, budget):
    """sizes of `budget` groups of consecutive FFT layers, as even as possible, larger groups first"""
    b = max(1, min(budget, log_s))
    base, extra = divmod(log_s, b)
    return [base + (1 if i < extra else 0) for i in range(b)]


def _layer_pairs(s, length, ksi, rot, M, inverse):
    lenh, lenq = length >> 1, length << 2
    gap = M // lenq
    p, q, wr, wi = [], [], [], []
    for i in range(0, s, length):
        for j in range(lenh):
            idx = ((lenq - (int(rot[j]) % lenq)) if inverse else (int(rot[j]) % lenq)) * gap
            p.append(i + j); q.append(i + j + lenh); wr.append(ksi[2 * idx]); wi.append(ksi[2 * idx + 1])
    return np.array(p), np.array(q), np.array(wr)[:, None], np.array(wi)[:, None]


def group_matrix(s, lengths, ksi, rot, M, inverse):
    """dense (re, im) of the product of the special-FFT butterfly layers `lengths` (application order).  Forward layer:
    [u; v] -> [u + w v; u - w v]; inverse layer: [a; b] -> [(a + b) / 2; (a - b) w' / 2] (the 1/size of the inverse
    transform spread over its layers).  Same elementary operations, in the same order, as csrc/bootstrap.cu."""
    re, im = np.eye(s), np.zeros((s, s))
    for length in lengths:
        p, q, wr, wi = _layer_pairs(s, length, ksi, rot, M, inverse)
        ar, ai, br, bi = re[p], im[p], re[q], im[q]
        if not inverse:
            vr = br * wr - bi * wi
            vi = br * wi + bi * wr
            re[p], im[p], re[q], im[q] = ar + vr, ai + vi, ar - vr, ai - vi
        else:
            ur, ui, dr, di = ar + br, ai + bi, ar - br, ai - bi
            re[p], im[p] = ur * 0.5, ui * 0.5
            re[q], im[q] = (dr * wr - di * wi) * 0.5, (dr * wi + di * wr) * 0.5
    return re, im


def diagonals(re, im, scalar_re, scalar_im):
    """generalised diagonals d -> (re[s], im[s]) with diag_d[i] = G[i, (i + d) mod s] * scalar; zero diagonals dropped"""
    s = re.shape[0]
    idx = np.arange(s)
    out = {}
    for d in range(s):
        cr, ci = re[idx, (idx + d) % s], im[idx, (idx + d) % s]
        if not (np.any(cr != 0.0) or np.any(ci != 0.0)):
            continue
        out[d] = (cr * scalar_re - ci * scalar_im, cr * scalar_im + ci * scalar_re)
    return out


class Bootstrapper:
    """EvalBootstrapSetup + EvalBootstrapKeyGen + EvalBootstrap for one (context, slots) pair"""

    def __init__(self, cc, level_budget, slots):
        c = cc.ctx
        if slots < 2 or slots & (slots - 1) or slots > c.n // 4:
            raise ValueError("EvalBootstrapSetup: sparse packing only (2 <= slots <= ring/4, power of two)")
        self.cc, self.s, self.budget = cc, slots, tuple(level_budget)
        self.log_s = slots.bit_length() - 1
        self.raise_level = 1 if c.p.ext else 0
        ksi, rot, M = c.fft_tables()
        lens_fwd = [2 << i for i in range(self.log_s)]                  # 2, 4, ..., s (application order of the forward FFT)
        gC, gS = layer_groups(self.log_s, self.budget[0]), layer_groups(self.log_s, self.budget[1])
        g = c.n // (2 * slots)
        q0 = float(c.q[0])
        sf_r = c.sf_real[self.raise_level]
        tot_c = sf_r / (g * q0 * K_UNIFORM) * 0.5                       # C2S: normalise to (coefficient pairs) / K, and the 1/2 of Re / Im
        tot_s = float(1 << CORRECTION) / (2.0 * math.pi)                # S2C: undo 2^-corr and the 2 pi of sin(2 pi t)
        sc, ss = math.pow(tot_c, 1.0 / len(gC)), math.pow(tot_s, 1.0 / len(gS))
        self.c2s, self.s2c = [], []
        pos = self.log_s
        for k, cnt in enumerate(gC):                                    # inverse layers: len = s first
            lens = [lens_fwd[pos - 1 - t] for t in range(cnt)]
            pos -= cnt
            re, im = group_matrix(slots, lens, ksi, rot, M, True)
            grp = {"main": diagonals(re, im, sc, 0.0)}
            if k == len(gC) - 1:
                grp["alt"] = diagonals(re, im, 0.0, -sc)                # -i G: its Re-extraction yields the imaginary parts
            self.c2s.append(grp)
        pos = 0
        for k, cnt in enumerate(gS):
            lens = [lens_fwd[pos + t] for t in range(cnt)]
            pos += cnt
            re, im = group_matrix(slots, lens, ksi, rot, M, False)
            grp = {"main": diagonals(re, im, ss, 0.0)}
            if k == 0:
                grp["alt"] = diagonals(re, im, 0.0, ss)                 # i G: applied to the imaginary-part ciphertext
            self.s2c.append(grp)
        self.cheb = boot_cheb_coefficients()
        self._pt = {}

    # ---- EvalBootstrapKeyGen
    def rotation_indices(self):
        idx = set()
        g = self.cc.ctx.n // (2 * self.s)
        j = 1
        while j < g:
            idx.add(self.s * j); j <<= 1
        for grp in self.c2s + self.s2c:
            for d in grp["main"]:
                if d:
                    idx.add(d)
        return sorted(idx)

    def keygen(self):
        self.cc.rot_keygen(self.rotation_indices())
        self.cc.conj_keygen()

    # ---- pieces
    def _plain(self, key, diag, level):
        k = (key, level)
        if k not in self._pt:
            re, im = diag
            self._pt[k] = self.cc.encode(re + 1j * im, 1, level, self.s)
        return self._pt[k]

    def _linear(self, ct, grp_id, which, diags):
        """sum_d diag_d (.) Rot(ct, d), one level; the rotations of one input share its ModUp in the engine"""
        cc = self.cc
        if ct.deg == 2:
            ct = ct.clone(); cc.rescale(ct)
        terms = []
        for d in sorted(diags):
            r = ct if d == 0 else cc.EvalRotate(ct, d)
            terms.append(cc.EvalMult(r, self._plain((grp_id, which, d), diags[d], ct.level)))
        return cc.EvalAddMany(terms)

    def _eval_mod(self, ct):
        cc = self.cc
        c = cc.EvalChebyshevSeriesPS(ct, self.cheb, -1.0, 1.0)
        for _ in range(R_UNIFORM):
            sq = cc.EvalSquare(c)
            c = cc.EvalSub(cc.EvalAdd(sq, sq), 1.0)
        return c

    # ---- EvalBootstrap
    def bootstrap(self, ct_in):
        cc, c = self.cc, self.cc.ctx
        ct = ct_in.clone()
        if ct.deg == 2:
            cc.rescale(ct)
        if ct.nlimbs < 2:
            raise ValueError("EvalBootstrap: the ciphertext has no limb left to drop (level == multiplicative depth)")
        cc.level_reduce(ct, ct.nlimbs - 2)
        last = c.size_q - 1
        cc.mult_core_const(ct, float(c.q[0]) / float(1 << CORRECTION) / c.sf_real[last])
        cc.rescale(ct)                                   # one limb: plaintext integer m * q0 * 2^-corr
        ct = cc.mod_raise(ct, self.raise_level)
        ct.slots = self.s
        j, g = 1, c.n // (2 * self.s)
        while j < g:                                     # SubSum
            ct = cc.EvalAdd(ct, cc.EvalRotate(ct, self.s * j)); j <<= 1
        for k, grp in enumerate(self.c2s[:-1]):
            ct = self._linear(ct, ("c", k), "main", grp["main"])
        k = len(self.c2s) - 1
        u = self._linear(ct, ("c", k), "main", self.c2s[k]["main"])
        v = self._linear(ct, ("c", k), "alt", self.c2s[k]["alt"])
        re = cc.EvalAdd(u, cc.EvalConjugate(u))
        im = cc.EvalAdd(v, cc.EvalConjugate(v))
        re, im = self._eval_mod(re), self._eval_mod(im)
        a = self._linear(re, ("s", 0), "main", self.s2c[0]["main"])
        b = self._linear(im, ("s", 0), "alt", self.s2c[0]["alt"])
        ct = cc.EvalAdd(a, b)
        for k, grp in enumerate(self.s2c[1:], start=1):
            ct = self._linear(ct, ("s", k), "main", grp["main"])
        ct.slots = ct_in.slots
        return ct
