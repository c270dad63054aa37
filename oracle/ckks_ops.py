"""oracle/ckks_ops.py -- TEST INFRASTRUCTURE ONLY (CPU oracle, homomorphic ops).

`OracleCC` offers the `lbcrypto::CryptoContextImpl<DCRTPoly>` methods the reference calls
(SURVEY.md 8(b)); each method cites the reference call sites it serves and restates the
OpenFHE 1.1.4 semantics (leveledshe-ckksrns.cpp, keyswitch-hybrid.cpp, dcrtpoly-impl.h).
"""
import numpy as np

from .cheby import ChebyshevMixin
from .ckks_eval import Ct, Evaluator
from .ckks_oracle import P


class OracleCC(ChebyshevMixin, Evaluator):
    # ---------------------------------------------------------------- low level
    def _count(self, name, nl):
        self.stats[(name, nl)] = self.stats.get((name, nl), 0) + 1

    def _scalar_residues(self, v, nl):
        return [v % self.ctx.q[i] for i in range(nl)]

    def _mul_scalar_int(self, ct, v):
        c = self.ctx
        for p in range(ct.d.shape[0]):
            for i in range(ct.nlimbs):
                self.L.orc_vmuls(P(ct.d[p, i]), P(ct.d[p, i]), v % c.q[i], c.n, c.q[i])

    def mult_core_const(self, ct, x):
        """EvalMultCoreInPlace(ct, double): times (int128)(x*sf_real[level] + 0.5) (C truncation);
        noiseScaleDeg+1, scalingFactor *= sf_real[level]  (GetElementForEvalMult)."""
        sf = self.ctx.sf_real[ct.level]
        v = int(x * sf + 0.5)
        self._mul_scalar_int(ct, v)
        ct.deg += 1
        ct.sf *= sf

    def rescale(self, ct):
        """ModReduceInternalInPlace -> DCRTPoly::DropLastElementAndScale: drop the last limb q_l,
        c_i <- (c_i - [c_l]_centred) * q_l^-1 mod q_i; level+1, deg-1, sf /= q_l."""
        c = self.ctx
        nl = ct.nlimbs
        ql = c.q[nl - 1]
        self._count("rescale", nl)
        out = np.empty((ct.d.shape[0], nl - 1, c.n), dtype=np.uint64)
        for p in range(ct.d.shape[0]):
            last = ct.d[p, nl - 1].copy()
            c.intt(last, nl - 1)
            t = np.empty((nl - 1, c.n), dtype=np.uint64)
            for i in range(nl - 1):
                self.L.orc_switch_modulus(P(t[i]), P(last), c.n, ql, c.q[i])
            c.ntt_rows(t, list(range(nl - 1)))
            for i in range(nl - 1):
                qi = c.q[i]
                self.L.orc_vsub(P(out[p, i]), P(ct.d[p, i]), P(t[i]), c.n, qi)
                self.L.orc_vmuls(P(out[p, i]), P(out[p, i]), pow(ql % qi, -1, qi), c.n, qi)
        ct.d = out
        ct.level += 1
        ct.deg -= 1
        ct.sf /= float(ql)

    def level_reduce(self, ct, levels):
        """LevelReduceInternalInPlace: drop `levels` last limbs, no scaling"""
        if levels > 0:
            ct.d = np.ascontiguousarray(ct.d[:, :ct.nlimbs - levels])
            ct.level += levels

    def adjust_levels_and_depth(self, a, b):
        """LeveledSHECKKSRNS::AdjustLevelsAndDepthInPlace (FLEXIBLEAUTO / FLEXIBLEAUTOEXT)"""
        c = self.ctx
        if a.level == b.level:
            if a.deg < b.deg:
                self.mult_core_const(a, 1.0)
            elif b.deg < a.deg:
                self.mult_core_const(b, 1.0)
            return
        if a.level > b.level:
            a, b = b, a
        # now a is at the lower level (more limbs) and must be brought to b
        l1, l2 = a.level, b.level
        scf = c.sf_real[l1]
        if a.deg == 2:
            if b.deg == 2:
                q1 = float(c.q[a.nlimbs - 1])
                self.mult_core_const(a, b.sf / a.sf * q1 / scf)
                self.rescale(a)
                if l1 + 1 < l2:
                    self.level_reduce(a, l2 - l1 - 1)
                a.sf = b.sf
            else:
                if l1 + 1 == l2:
                    self.rescale(a)
                else:
                    q1 = float(c.q[a.nlimbs - 1])
                    self.mult_core_const(a, c.sf_big[l2 - 1] / a.sf * q1 / scf)
                    self.rescale(a)
                    if l1 + 2 < l2:
                        self.level_reduce(a, l2 - l1 - 2)
                    self.rescale(a)
                    a.sf = b.sf
        else:
            if b.deg == 2:
                self.mult_core_const(a, b.sf / a.sf / scf)
                self.level_reduce(a, l2 - l1)
                a.sf = b.sf
            else:
                self.mult_core_const(a, c.sf_big[l2 - 1] / a.sf / scf)
                if l1 + 1 < l2:
                    self.level_reduce(a, l2 - l1 - 1)
                self.rescale(a)
                a.sf = b.sf

    def adjust_to_one(self, a, b):
        """AdjustLevelsAndDepthToOneInPlace"""
        self.adjust_levels_and_depth(a, b)
        if a.deg == 2:
            self.rescale(a)
            self.rescale(b)

    # ---------------------------------------------------------------- key switching
    def key_switch_core(self, poly, evk):
        """KeySwitchHYBRID::KeySwitchCore on one EVALUATION polynomial [nl][n]:
        EvalKeySwitchPrecomputeCore (ModUp) -> EvalFastKeySwitchCoreExt (inner product with the
        evk limbs) -> ApproxModDown.  Returns (ct0, ct1) on the nl Q limbs."""
        c = self.ctx
        nl = poly.shape[0]
        self._count("keyswitch", nl)
        parts, md = c.ks_tables(nl)
        ext_ids = list(range(nl)) + [c.size_q + k for k in range(c.size_p)]
        acc = np.zeros((2, len(ext_ids), c.n), dtype=np.uint64)
        av, bv = evk
        for j, pt in enumerate(parts):
            dig = pt["dig"]
            coef = np.ascontiguousarray(poly[dig[0]:dig[-1] + 1].copy())
            c.ntt_rows(coef, dig, inverse=True)
            conv = np.empty((len(pt["compl"]), c.n), dtype=np.uint64)
            self.L.orc_approx_switch_crt_basis(P(conv), P(coef), c.n, len(dig), len(pt["compl"]),
                                               P(pt["qin"]), P(pt["qout"]), P(pt["hatinv"]), P(pt["hatmod"]))
            c.ntt_rows(conv, pt["compl"])
            pos = {mi: r for r, mi in enumerate(pt["compl"])}
            for e, mi in enumerate(ext_ids):
                dj = poly[mi] if mi in dig else conv[pos[mi]]
                q = c.moduli[mi]
                self.L.orc_vmuladd(P(acc[0, e]), P(dj), P(bv[j][mi]), c.n, q)
                self.L.orc_vmuladd(P(acc[1, e]), P(dj), P(av[j][mi]), c.n, q)
        out = np.empty((2, nl, c.n), dtype=np.uint64)
        for h in range(2):
            pp = np.ascontiguousarray(acc[h, nl:].copy())
            c.ntt_rows(pp, [c.size_q + k for k in range(c.size_p)], inverse=True)
            conv = np.empty((nl, c.n), dtype=np.uint64)
            self.L.orc_approx_switch_crt_basis(P(conv), P(pp), c.n, c.size_p, nl, P(md["pin"]), P(md["qout"]),
                                               P(md["phatinv"]), P(md["phatmod"]))
            c.ntt_rows(conv, list(range(nl)))
            for i in range(nl):
                q = c.q[i]
                self.L.orc_vsub(P(out[h, i]), P(acc[h, i]), P(conv[i]), c.n, q)
                self.L.orc_vmuls(P(out[h, i]), P(out[h, i]), md["pinv"][i], c.n, q)
        return out

    # ---------------------------------------------------------------- public API (OpenFHE names)
    def EvalAdd(self, a, b, sign=1):
        """EvalAdd/EvalSub(ct,ct|pt), EvalAdd/EvalSub(ct,double)  (sort_algo.h:358,454,504,708,720;
        comparison.cpp:13,19)"""
        if isinstance(b, (int, float)):
            r = a.clone()
            self._add_const(r, sign * float(b))
            return r
        x, y = a.clone(), b.clone()
        self.adjust_levels_and_depth(x, y)
        c = self.ctx
        ct, pt = (x, y) if x.d.shape[0] >= y.d.shape[0] else (y, x)
        out = ct.d.copy()
        if ct is y and sign < 0:  # pt - ct
            for p in range(out.shape[0]):
                for i in range(ct.nlimbs):
                    self.L.orc_vneg(P(out[p, i]), P(out[p, i]), c.n, c.q[i])
            for i in range(ct.nlimbs):
                self.L.orc_vadd(P(out[0, i]), P(out[0, i]), P(pt.d[0, i]), c.n, c.q[i])
        else:
            f = self.L.orc_vadd if sign > 0 else self.L.orc_vsub
            for p in range(pt.d.shape[0]):
                for i in range(ct.nlimbs):
                    f(P(out[p, i]), P(out[p, i]), P(pt.d[p, i]), c.n, c.q[i])
        self._count("add", ct.nlimbs)
        return Ct(out, ct.level, ct.deg, ct.sf, max(a.slots, b.slots))

    def EvalSub(self, a, b):
        if isinstance(a, (int, float)):  # EvalSub(double, ct)  (comparison.cpp:38)
            return self.EvalAdd(self.EvalNegate(b), float(a))
        return self.EvalAdd(a, b, sign=-1)

    add_const_mode = 0

    def _add_const(self, ct, x):
        """EvalAddInPlace(ct,double): add the scaled constant (sign applied) to c0.
        add_const_mode 0: round(|x| * ct.scalingFactor).
        add_const_mode 1: OpenFHE 1.1.4 LeveledSHECKKSRNS::GetElementForEvalAddOrSub as recalled (unverifiable offline,
        SURVEY 8c'-7): round(|x| * sf_level) multiplied by round(sf_level) once per extra noiseScaleDeg (CRTMult); at
        level 0 under FLEXIBLEAUTOEXT the big scaling factor is used and the degree loop is skipped."""
        c = self.ctx
        if self.add_const_mode == 0 or ct.deg <= 1 or (c.p.ext and ct.level == 0):
            v = int(abs(x) * ct.sf + 0.5)
        else:
            sf = c.sf_real[ct.level]
            v = int(abs(x) * sf + 0.5) * int(sf + 0.5) ** (ct.deg - 1)
        if x < 0:
            v = -v
        for i in range(ct.nlimbs):
            self.L.orc_vadds(P(ct.d[0, i]), P(ct.d[0, i]), v % c.q[i], c.n, c.q[i])

    def EvalNegate(self, a):
        c = self.ctx
        r = a.clone()
        for p in range(r.d.shape[0]):
            for i in range(r.nlimbs):
                self.L.orc_vneg(P(r.d[p, i]), P(r.d[p, i]), c.n, c.q[i])
        return r

    def EvalAddMany(self, cts):
        """EvalAddMany (sort_algo.h:364): binary-tree sum"""
        cts = list(cts)
        while len(cts) > 1:
            nxt = [self.EvalAdd(cts[i], cts[i + 1]) for i in range(0, len(cts) - 1, 2)]
            if len(cts) % 2:
                nxt.append(cts[-1])
            cts = nxt
        return cts[0]

    def EvalMult(self, a, b):
        """EvalMult(ct,ct) / EvalMultAndRelinearize / EvalMult(ct,pt) / EvalMult(ct,double)
        (sign.cpp:23-56, comparison.cpp:19, sort_algo.h:358,576,723,730)"""
        c = self.ctx
        if isinstance(b, (int, float)):
            r = a.clone()
            if r.deg == 2:
                self.rescale(r)
            self.mult_core_const(r, float(b))
            return r
        x, y = a.clone(), b.clone()
        self.adjust_to_one(x, y)
        nl = x.nlimbs
        if x.d.shape[0] == 2 and y.d.shape[0] == 2:
            self._count("mult", nl)
            d = np.empty((3, nl, c.n), dtype=np.uint64)
            t = np.empty(c.n, dtype=np.uint64)
            for i in range(nl):
                q = c.q[i]
                self.L.orc_vmul(P(d[0, i]), P(x.d[0, i]), P(y.d[0, i]), c.n, q)
                self.L.orc_vmul(P(d[1, i]), P(x.d[0, i]), P(y.d[1, i]), c.n, q)
                self.L.orc_vmul(P(t), P(x.d[1, i]), P(y.d[0, i]), c.n, q)
                self.L.orc_vadd(P(d[1, i]), P(d[1, i]), P(t), c.n, q)
                self.L.orc_vmul(P(d[2, i]), P(x.d[1, i]), P(y.d[1, i]), c.n, q)
            ks = self.key_switch_core(np.ascontiguousarray(d[2]), self.relin)
            out = np.empty((2, nl, c.n), dtype=np.uint64)
            for i in range(nl):
                self.L.orc_vadd(P(out[0, i]), P(d[0, i]), P(ks[0, i]), c.n, c.q[i])
                self.L.orc_vadd(P(out[1, i]), P(d[1, i]), P(ks[1, i]), c.n, c.q[i])
            slots = max(a.slots, b.slots)
        else:
            ct, pt = (x, y) if x.d.shape[0] == 2 else (y, x)
            self._count("mult_pt", nl)
            out = np.empty_like(ct.d)
            for p in range(2):
                for i in range(nl):
                    self.L.orc_vmul(P(out[p, i]), P(ct.d[p, i]), P(pt.d[0, i]), c.n, c.q[i])
            slots = ct.slots
        return Ct(out, x.level, x.deg + y.deg, x.sf * y.sf, slots)

    EvalMultAndRelinearize = EvalMult

    def EvalSquare(self, a):
        return self.EvalMult(a, a)

    def EvalConjugate(self, a):
        """complex conjugation of every slot: EvalAutomorphism with index 2n - 1"""
        return self.EvalRotate(a, None, k=2 * self.ctx.n - 1)

    def mod_raise(self, ct, level):
        """bootstrapping's ModRaise: the single limb of `ct` (mod q0), read as centred integers, in every limb of
        `level`; noiseScaleDeg 1, scaling factor := the level's"""
        c = self.ctx
        assert ct.nlimbs == 1 and ct.deg == 1
        nl = c.size_q - level
        out = np.empty((ct.d.shape[0], nl, c.n), dtype=np.uint64)
        for p in range(ct.d.shape[0]):
            coef = ct.d[p, 0].copy()
            c.intt(coef, 0)
            for i in range(nl):
                self.L.orc_switch_modulus(P(out[p, i]), P(coef), c.n, c.q[0], c.q[i])
            c.ntt_rows(out[p], list(range(nl)))
        return Ct(out, level, 1, c.sf_real[level], ct.slots)

    # ---- bootstrapping (oracle/bootstrap.py)
    def EvalBootstrapSetup(self, level_budget=(4, 4), dim1=(0, 0), slots=0):
        from .bootstrap import Bootstrapper
        if not hasattr(self, "_boot"):
            self._boot = {}
        slots = slots or self.ctx.n // 4
        self._boot[slots] = Bootstrapper(self, level_budget, slots)

    def EvalBootstrapKeyGen(self, slots):
        self._boot[slots].keygen()

    def EvalBootstrap(self, ct, iters=1, prec=0):
        return self._boot[ct.slots].bootstrap(ct)

    def EvalRotate(self, a, r, k=None):
        """EvalRotate -> EvalAutomorphism: key-switch c1 with the rotation key, THEN apply the
        automorphism index permutation to both polynomials (rotation.h:224,231)"""
        c = self.ctx
        if k is None:
            k = self.auto_index(r)
        if k not in self.rot:
            raise KeyError("EvalRotate: no rotation key for index %s" % r)
        self._count("rotate", a.nlimbs)
        ks = self.key_switch_core(np.ascontiguousarray(a.d[1]), self.rot[k])
        out = np.empty_like(a.d)
        t = np.empty(c.n, dtype=np.uint64)
        for i in range(a.nlimbs):
            self.L.orc_vadd(P(t), P(a.d[0, i]), P(ks[0, i]), c.n, c.q[i])
            self.L.orc_automorph_eval(P(out[0, i]), P(t), c.log_n, k)
            self.L.orc_automorph_eval(P(out[1, i]), P(ks[1, i]), c.log_n, k)
        return Ct(out, a.level, a.deg, a.sf, a.slots)

    def MakeCKKSPackedPlaintext(self, values, scale_deg=1, level=0, params=None, slots=None):
        return self.encode(values, scale_deg, level, slots)
