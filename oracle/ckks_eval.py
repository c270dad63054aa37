"""oracle/ckks_eval.py -- TEST INFRASTRUCTURE ONLY (CPU oracle, evaluator half).

Restates, on top of oracle/ckks_oracle.py + ckks_prims.c, the OpenFHE 1.1.4 CKKS evaluator
semantics the reference relies on:
key generation, encode/encrypt/decrypt, EvalAdd/Sub/Mult/Square/Rotate with the
FLEXIBLEAUTO(EXT) level/scale bookkeeping (leveledshe-ckksrns.cpp AdjustLevelsAndDepthInPlace),
hybrid key switching (keyswitch-hybrid.cpp) and rescale (DropLastElementAndScale).
Parity unpinned vs OpenFHE at limb level (see ckks_oracle.py header).
"""
import math

import numpy as np

from .ckks_oracle import (P, PD, STREAM_PK_A, STREAM_PK_E, STREAM_SECRET, Context, lib, stream_enc,
                          stream_evk, u32p)


class Ct:
    """Ciphertext or (parts == 1) plaintext: limb-major uint64 array [parts][nlimbs][n] in
    EVALUATION format + the metadata OpenFHE keeps (level, noiseScaleDeg, scalingFactor, slots)."""

    def __init__(self, d, level, deg, sf, slots):
        self.d, self.level, self.deg, self.sf, self.slots = d, level, deg, sf, slots

    @property
    def nlimbs(self):
        return self.d.shape[1]

    def clone(self):
        return Ct(self.d.copy(), self.level, self.deg, self.sf, self.slots)


class Evaluator:
    def __init__(self, ctx: Context):
        self.ctx = ctx
        self.L = lib()
        self.secret = None  # [size_q + size_p][n] EVALUATION
        self.pk = None
        self.relin = None
        self.rot = {}  # automorphism index -> evk
        self.stats = {}

    # ------------------------------------------------------------------ sampling
    def _small(self, stream, kind, mod_ids):
        """small integer polynomial (ternary kind 0 / error kind 1), NTT'd, on the given moduli"""
        c, n = self.ctx, self.ctx.n
        out = np.empty((len(mod_ids), n), dtype=np.uint64)
        for r, mi in enumerate(mod_ids):
            self.L.orc_sample(P(out[r]), n, c.moduli[mi], c.seed, stream, kind)
            c.ntt(out[r], mi)
        return out

    def _uniform(self, stream, mod_ids):
        """uniform residues, EVALUATION format: hash(seed, stream, ((limb+1) << 32) | i) % q"""
        c, n = self.ctx, self.ctx.n
        out = np.empty((len(mod_ids), n), dtype=np.uint64)
        for r, mi in enumerate(mod_ids):
            self._uniform_limb(out[r], stream, (mi + 1) << 32, c.moduli[mi])
        return out

    def _uniform_limb(self, out, stream, base, q):
        # hash(seed, stream, base + i) % q ; done in C via orc_hash would be slow per element,
        # so reproduce the hash with numpy uint64 arithmetic (wraps mod 2^64).
        n = out.shape[0]
        seed = np.uint64(self.ctx.seed)
        with np.errstate(over="ignore"):
            def mix(x):
                x = x ^ (x >> np.uint64(30)); x = x * np.uint64(0xbf58476d1ce4e5b9)
                x = x ^ (x >> np.uint64(27)); x = x * np.uint64(0x94d049bb133111eb)
                return x ^ (x >> np.uint64(31))
            a = mix(np.array([seed + np.uint64(0x9E3779B97F4A7C15) * np.uint64(stream + 1)], dtype=np.uint64))
            idx = np.arange(n, dtype=np.uint64) + np.uint64(base)
            b = mix(idx + np.uint64(0xD1B54A32D192ED03))
            h = mix(a ^ b)
        out[:] = h % np.uint64(q)

    # ------------------------------------------------------------------ keys
    def keygen(self):
        """KeyGen (pke-rns): s ternary, pk = (-a*s + e, a).  Secret kept on Q and P limbs."""
        c = self.ctx
        all_ids = list(range(c.size_q + c.size_p))
        qids = list(range(c.size_q))
        self.secret = self._small(STREAM_SECRET, 0, all_ids)
        a = self._uniform(STREAM_PK_A, qids)
        e = self._small(STREAM_PK_E, 1, qids)
        b = np.empty_like(a)
        for i in qids:
            q = c.q[i]
            self.L.orc_vmul(P(b[i]), P(a[i]), P(self.secret[i]), c.n, q)
            self.L.orc_vsub(P(b[i]), P(e[i]), P(b[i]), c.n, q)
        self.pk = (b, a)

    def _switch_key_gen(self, s_old, s_new, key_tag):
        """KeySwitchHYBRID::KeySwitchGenInternal: b_j = -a_j*sNew + e_j + [limb in digit j] P*sOld"""
        c = self.ctx
        ids = list(range(c.size_q + c.size_p))
        av, bv = [], []
        for part in range(c.num_part_q):
            a = self._uniform(stream_evk(key_tag, part, 0), ids)
            e = self._small(stream_evk(key_tag, part, 1), 1, ids)
            b = np.empty_like(a)
            lo, hi = c.alpha * part, min(c.alpha * (part + 1), c.size_q)
            for i in ids:
                q = c.moduli[i]
                self.L.orc_vmul(P(b[i]), P(a[i]), P(s_new[i]), c.n, q)
                self.L.orc_vsub(P(b[i]), P(e[i]), P(b[i]), c.n, q)
                if lo <= i < hi:
                    tmp = np.empty(c.n, dtype=np.uint64)
                    self.L.orc_vmuls(P(tmp), P(s_old[i]), c.Pprod % q, c.n, q)
                    self.L.orc_vadd(P(b[i]), P(b[i]), P(tmp), c.n, q)
            av.append(a)
            bv.append(b)
        return (av, bv)

    def relin_keygen(self):
        c = self.ctx
        s2 = np.empty_like(self.secret)
        for i in range(s2.shape[0]):
            self.L.orc_vmul(P(s2[i]), P(self.secret[i]), P(self.secret[i]), c.n, c.moduli[i])
        self.relin = self._switch_key_gen(s2, self.secret, 0)

    def auto_index(self, r):
        """FindAutomorphismIndex2nComplex: 5^r mod 2n (r>0 left rotation), inverse for r<0"""
        m = 2 * self.ctx.n
        if r == 0:
            return 1
        g = 5 if r > 0 else pow(5, -1, m)
        return pow(g, abs(r), m)

    def rot_keygen(self, indices):
        """EvalAutomorphismKeyGen: key switches from s to sigma_{k^-1}(s)"""
        c = self.ctx
        for r in indices:
            k = self.auto_index(r)
            if k in self.rot:
                continue
            kinv = pow(k, -1, 2 * c.n)
            sp = np.empty_like(self.secret)
            for i in range(sp.shape[0]):
                self.L.orc_automorph_eval(P(sp[i]), P(self.secret[i]), c.log_n, kinv)
            self.rot[k] = self._switch_key_gen(self.secret, sp, k)

    def conj_keygen(self):
        """key for the conjugation automorphism X -> X^(2n-1) (its own inverse); EvalBootstrapKeyGen generates it"""
        c = self.ctx
        k = 2 * c.n - 1
        if k in self.rot:
            return
        sp = np.empty_like(self.secret)
        for i in range(sp.shape[0]):
            self.L.orc_automorph_eval(P(sp[i]), P(self.secret[i]), c.log_n, k)
        self.rot[k] = self._switch_key_gen(self.secret, sp, k)

    # ------------------------------------------------------------------ encode / decode
    def encode(self, values, scale_deg=1, level=0, slots=None):
        """MakeCKKSPackedPlaintext(values, scaleDeg, level, nullptr, slots): special inverse FFT,
        scale, round, RNS, NTT.  EXT at level 0: scale = sf_big[0], noiseScaleDeg 2."""
        c = self.ctx
        if slots is None:
            slots = 1
            while slots < len(values):
                slots <<= 1
        assert len(values) <= slots <= c.n // 2
        if c.p.ext and level == 0:
            sf, deg = c.sf_big[0], 2
        else:
            sf, deg = c.sf_real[level] ** scale_deg, scale_deg
        v = np.zeros(2 * slots, dtype=np.float64)
        vals = np.asarray(values)
        if np.iscomplexobj(vals):                      # complex slot values (bootstrapping's linear-transform diagonals)
            v[0:2 * len(values):2] = vals.real
            v[1:2 * len(values):2] = vals.imag
        else:
            v[0:2 * len(values):2] = np.asarray(values, dtype=np.float64)
        ksi, rot, M = c.fft_tables()
        self.L.orc_fft_special_inv(PD(v), slots, PD(ksi), rot.ctypes.data_as(u32p), M)
        nh, gap = c.n // 2, (c.n // 2) // slots
        vs = v * sf
        # CKKSPackedEncoding::Encode (64-bit native): values beyond 61 bits are scaled down by 2^logApprox
        # before rounding and the residues multiplied back by 2^logApprox mod q_i
        mx = float(np.max(np.abs(vs)))
        log_approx = 0
        if mx > 0.0:
            f, e = math.frexp(mx)
            logc = e - 1 if f == 0.5 else e
            log_approx = max(0, logc - 61)
        vs = np.ldexp(vs, -log_approx)
        coef = np.zeros(c.n, dtype=np.float64)
        coef[0:nh:gap] = vs[0::2]
        coef[nh::gap] = vs[1::2]
        nl = c.size_q - level
        d = np.empty((1, nl, c.n), dtype=np.uint64)
        for i in range(nl):
            self.L.orc_round_to_rns(P(d[0, i]), PD(coef), c.n, c.q[i])
            if log_approx:
                self.L.orc_vmuls(P(d[0, i]), P(d[0, i]), (1 << log_approx) % c.q[i], c.n, c.q[i])
        c.ntt_rows(d[0], list(range(nl)))
        return Ct(d, level, deg, sf, slots)

    def encrypt(self, pt: Ct):
        """Encrypt(pk, pt): c0 = b*v + e0 + m, c1 = a*v + e1 on the plaintext's limbs"""
        c = self.ctx
        nl = pt.nlimbs
        ids = list(range(nl))
        cnt = c.enc_counter
        c.enc_counter += 1
        v = self._small(stream_enc(cnt, 0), 0, ids)
        e0 = self._small(stream_enc(cnt, 1), 1, ids)
        e1 = self._small(stream_enc(cnt, 2), 1, ids)
        d = np.empty((2, nl, c.n), dtype=np.uint64)
        for i in ids:
            q = c.q[i]
            self.L.orc_vmul(P(d[0, i]), P(self.pk[0][i]), P(v[i]), c.n, q)
            self.L.orc_vadd(P(d[0, i]), P(d[0, i]), P(e0[i]), c.n, q)
            self.L.orc_vadd(P(d[0, i]), P(d[0, i]), P(pt.d[0, i]), c.n, q)
            self.L.orc_vmul(P(d[1, i]), P(self.pk[1][i]), P(v[i]), c.n, q)
            self.L.orc_vadd(P(d[1, i]), P(d[1, i]), P(e1[i]), c.n, q)
        return Ct(d, pt.level, pt.deg, pt.sf, pt.slots)

    def decrypt(self, ct: Ct, length=None):
        """Decrypt + CKKS decode (real parts).  Uses limbs 0,1 only: the centred value of
        c0 + c1*s is far below q0*q1/2.  (OpenFHE's decode-time noise flooding is omitted.)"""
        c = self.ctx
        use = min(2, ct.nlimbs)
        m = np.empty((use, c.n), dtype=np.uint64)
        for i in range(use):
            q = c.q[i]
            self.L.orc_vmul(P(m[i]), P(ct.d[1, i]), P(self.secret[i]), c.n, q)
            self.L.orc_vadd(P(m[i]), P(m[i]), P(ct.d[0, i]), c.n, q)
            if ct.d.shape[0] == 3:
                t = np.empty(c.n, dtype=np.uint64)
                self.L.orc_vmul(P(t), P(ct.d[2, i]), P(self.secret[i]), c.n, q)
                self.L.orc_vmul(P(t), P(t), P(self.secret[i]), c.n, q)
                self.L.orc_vadd(P(m[i]), P(m[i]), P(t), c.n, q)
            c.intt(m[i], i)
        return self.decode_coeffs(m, ct.sf, ct.slots, length)

    def decode_coeffs(self, m, sf, slots, length=None):
        c = self.ctx
        co = np.empty(c.n, dtype=np.float64)
        if m.shape[0] >= 2:
            self.L.orc_crt2_to_double(PD(co), P(m[0]), P(m[1]), c.n, c.q[0], c.q[1])
        else:
            self.L.orc_crt1_to_double(PD(co), P(m[0]), c.n, c.q[0])
        nh, gap = c.n // 2, (c.n // 2) // slots
        v = np.empty(2 * slots, dtype=np.float64)
        v[0::2] = co[0:nh:gap] / sf
        v[1::2] = co[nh::gap] / sf
        ksi, rot, M = c.fft_tables()
        self.L.orc_fft_special(PD(v), slots, PD(ksi), rot.ctypes.data_as(u32p), M)
        out = v[0::2].copy()
        return out if length is None else out[:length]

    def decode(self, pt: Ct, length=None):
        m = pt.d[0, :min(2, pt.nlimbs)].copy()
        for i in range(m.shape[0]):
            self.ctx.intt(m[i], i)
        return self.decode_coeffs(m, pt.sf, pt.slots, length)
