"""oracle/ckks_oracle.py -- TEST INFRASTRUCTURE ONLY (CPU oracle).

CPU restatement of the RNS-CKKS engine the reference (oksuman/sorting-fhe) calls through
`lbcrypto::CryptoContext<DCRTPoly>` (every `cc->Eval*` in /root/reference/src/sort_algo.h,
src/sign.cpp, src/comparison.cpp, src/rotation.h).  The arithmetic lives in OpenFHE 1.1.4
(README.md:25, CMakeLists.txt:35), an un-vendored dependency absent from /root/reference, so
this file restates OpenFHE's published algorithms (scheme/ckksrns/*, keyswitch/keyswitch-hybrid,
core/lattice/hal/default/dcrtpoly-impl.h) from knowledge of that source.

PARITY STATUS: "parity unpinned" at the limb level -- no reference test pins a ciphertext
limb, and OpenFHE cannot be built here.  What IS pinned: the decrypted known-answer vectors of
tests/CompareTest.cpp, tests/SignTest.cpp (see tests/test_golden.py) and the level budgets
of src/sort_algo.h:87-201.

Array primitives are in oracle/ckks_prims.c (plain C); this module holds the context tables,
the FLEXIBLEAUTO(EXT) bookkeeping and the evaluator.  The product (CUDA) path never imports it.
"""
import ctypes
import math
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
u64p = ctypes.POINTER(ctypes.c_uint64)
f64p = ctypes.POINTER(ctypes.c_double)
u32p = ctypes.POINTER(ctypes.c_uint32)


def build_lib(force=False):
    so = os.path.join(_HERE, "libckks_oracle.so")
    srcs = [os.path.join(_HERE, "ckks_prims.c"), os.path.join(_HERE, "ckks_fast.c")]
    if force or not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(x) for x in srcs):
        subprocess.check_call(["gcc", "-O3", "-fPIC", "-shared", "-ffp-contract=off", "-fopenmp",
                               "-o", so] + srcs + ["-lm"])
    return so


SLOW = bool(int(os.environ.get("ORACLE_SLOW", "0")))  # 1: use only the slow reference loops of ckks_prims.c
_FAST_TWINS = ["vadd", "vsub", "vneg", "vmul", "vmuladd", "vmuls", "vadds", "automorph_eval",
               "approx_switch_crt_basis", "switch_modulus"]


def lib():
    global _LIB
    if _LIB is None:
        L = ctypes.CDLL(build_lib())
        U, I, S = ctypes.c_uint64, ctypes.c_int, ctypes.c_size_t
        for name, res, args in [
            ("orc_powmod", U, [U, U, U]), ("orc_invmod", U, [U, U]), ("orc_is_prime", I, [U]),
            ("orc_first_prime", U, [I, U]), ("orc_prev_prime", U, [U, U]), ("orc_next_prime", U, [U, U]),
            ("orc_min_root_of_unity", U, [U, U]), ("orc_make_psi_table", None, [u64p, I, U, U]),
            ("orc_ntt_fwd", None, [u64p, I, U, u64p]), ("orc_ntt_inv", None, [u64p, I, U, u64p]),
            ("orc_vadd", None, [u64p, u64p, u64p, S, U]), ("orc_vsub", None, [u64p, u64p, u64p, S, U]),
            ("orc_vmul", None, [u64p, u64p, u64p, S, U]), ("orc_vmuladd", None, [u64p, u64p, u64p, S, U]),
            ("orc_vmuls", None, [u64p, u64p, U, S, U]), ("orc_vadds", None, [u64p, u64p, U, S, U]),
            ("orc_vneg", None, [u64p, u64p, S, U]), ("orc_automorph_eval", None, [u64p, u64p, I, U]),
            ("orc_approx_switch_crt_basis", None, [u64p, u64p, S, I, I, u64p, u64p, u64p, u64p]),
            ("orc_switch_modulus", None, [u64p, u64p, S, U, U]), ("orc_hash", U, [U, U, U]),
            ("orc_sample", None, [u64p, S, U, U, U, I]),
            ("orc_fft_special_inv", None, [f64p, ctypes.c_uint32, f64p, u32p, ctypes.c_uint32]),
            ("orc_fft_special", None, [f64p, ctypes.c_uint32, f64p, u32p, ctypes.c_uint32]),
            ("orc_round_to_rns", None, [u64p, f64p, S, U]),
            ("orc_crt2_to_double", None, [f64p, u64p, u64p, S, U, U]),
            ("orc_crt1_to_double", None, [f64p, u64p, S, U]),
            ("orf_shoup_table", None, [u64p, u64p, S, U]),
            ("orf_ntt_batch", None, [u64p, I, I, ctypes.POINTER(ctypes.c_int32), u64p, u64p, u64p, u64p, I]),
        ]:
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        for nm in _FAST_TWINS:  # fast twins share the slow functions' signatures
            slow, fast = getattr(L, "orc_" + nm), getattr(L, "orf_" + nm)
            fast.restype, fast.argtypes = slow.restype, slow.argtypes
            setattr(L, "slow_" + nm, slow)
            if not SLOW:
                setattr(L, "orc_" + nm, fast)
        _LIB = L
    return _LIB


def P(a):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u64p)


def PD(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(f64p)


def stream_evk(key_tag, part, which):
    """sampling stream id of an evaluation key component (which: 0 = a, 1 = e)"""
    return (1 << 32) + key_tag * 64 + part * 2 + which


STREAM_SECRET, STREAM_PK_A, STREAM_PK_E = 1, 2, 3


def stream_enc(counter, which):
    return (2 << 32) + 3 * counter + which


class Params:
    """CCParams<CryptoContextCKKSRNS> subset the reference sets (src/sort_algo.h:87-201,
    tests/DirectSortTest.cpp:29-44, src/kway_adapter.h:41-64)."""

    def __init__(self, log_n, mult_depth, scaling_mod_size=40, first_mod_size=60, num_large_digits=0,
                 ext=True, seed=1):
        self.log_n, self.mult_depth = log_n, mult_depth
        self.scaling_mod_size, self.first_mod_size = scaling_mod_size, first_mod_size
        self.num_large_digits, self.ext, self.seed = num_large_digits, ext, seed


class Context:
    """Moduli chain, scaling factors and key-switch tables, restating OpenFHE
    ParameterGenerationCKKSRNS::ParamsGenCKKSRNS and CryptoParametersCKKSRNS::PrecomputeCRTTables."""

    def __init__(self, p: Params):
        L = lib()
        self.p = p
        self.log_n, self.n = p.log_n, 1 << p.log_n
        self.seed = p.seed
        m = 2 * self.n
        num_primes = p.mult_depth + 1
        ext_bits = 20 if p.ext else 0
        size_q = num_primes + (1 if p.ext else 0)
        q = [0] * size_q
        q[num_primes - 1] = L.orc_first_prime(p.scaling_mod_size, m)
        if num_primes > 1:
            sf = float(q[num_primes - 1])
            cnt = 0
            for i in range(num_primes - 2, 0, -1):
                sf = (sf * sf) / float(q[i + 1])
                sf_int = int(round(sf))  # std::llround
                sf_rem = sf_int % m
                if cnt % 2 == 0:
                    c = sf_int - m - sf_rem + 1
                    while True:
                        c = L.orc_prev_prime(c, m)
                        if c not in q[i + 1:num_primes]:
                            break
                else:
                    c = sf_int + m - sf_rem + 1
                    while True:
                        c = L.orc_next_prime(c, m)
                        if c not in q[i + 1:num_primes]:
                            break
                q[i] = c
                cnt += 1
        if p.first_mod_size == p.scaling_mod_size:
            q_prev = min(x for x in q[1:num_primes]) if num_primes > 1 else q[0]
            q[0] = L.orc_prev_prime(q_prev, m)
        else:
            q[0] = L.orc_prev_prime(L.orc_first_prime(p.first_mod_size, m), m)
        if p.ext:
            if ext_bits in (p.scaling_mod_size, p.first_mod_size):
                q[num_primes] = L.orc_prev_prime(q[0], m)
            else:
                q[num_primes] = L.orc_prev_prime(L.orc_first_prime(ext_bits, m), m)
        self.q, self.size_q = q, size_q
        # hybrid key switching: digits and special primes
        dnum = p.num_large_digits
        if dnum == 0:
            dnum = 3 if p.mult_depth > 3 else (2 if p.mult_depth > 0 else 1)
        self.alpha = (size_q + dnum - 1) // dnum
        self.num_part_q = (size_q + self.alpha - 1) // self.alpha
        max_bits = 0
        for j in range(self.num_part_q):
            prod = 1
            for k in range(j * self.alpha, min((j + 1) * self.alpha, size_q)):
                prod *= q[k]
            max_bits = max(max_bits, prod.bit_length())
        self.size_p = (max_bits + 59) // 60
        pp, prev = [], L.orc_first_prime(60, m)
        for _ in range(self.size_p):
            while True:
                prev = L.orc_prev_prime(prev, m)
                if prev not in q:
                    break
            pp.append(prev)
        self.pmod = pp
        self.moduli = q + pp  # global modulus index space: Q then P
        # NTT tables
        M = len(self.moduli)
        self.psi = []
        self.psi_br = np.zeros((M, self.n), dtype=np.uint64)       # [modulus][bitrev(i)] = psi^i
        self.psi_inv_br = np.zeros((M, self.n), dtype=np.uint64)
        self.psi_br_sh = np.zeros((M, self.n), dtype=np.uint64)    # Shoup companions (fast path only)
        self.psi_inv_br_sh = np.zeros((M, self.n), dtype=np.uint64)
        for i, mod in enumerate(self.moduli):
            psi = L.orc_min_root_of_unity(m, mod)
            L.orc_make_psi_table(P(self.psi_br[i]), self.log_n, psi, mod)
            L.orc_make_psi_table(P(self.psi_inv_br[i]), self.log_n, L.orc_invmod(psi, mod), mod)
            L.orf_shoup_table(P(self.psi_br_sh[i]), P(self.psi_br[i]), self.n, mod)
            L.orf_shoup_table(P(self.psi_inv_br_sh[i]), P(self.psi_inv_br[i]), self.n, mod)
            self.psi.append(psi)
        self.mod_arr = np.array(self.moduli, dtype=np.uint64)
        self.ninv_arr = np.array([pow(self.n, -1, x) for x in self.moduli], dtype=np.uint64)
        # scaling factors (FLEXIBLEAUTO / FLEXIBLEAUTOEXT)
        sfr = [0.0] * size_q
        sfr[0] = float(q[size_q - 1])
        if p.ext and size_q > 1:
            sfr[1] = float(q[size_q - 2])
        npre = 2 if p.ext else 1
        for k in range(npre, size_q):
            sfr[k] = sfr[k - 1] * sfr[k - 1] / float(q[size_q - k])
        self.sf_real = sfr
        big = [0.0] * max(size_q - 1, 0)
        if big:
            big[0] = sfr[0] * sfr[1] if p.ext else sfr[0] * sfr[0]
            for k in range(1, size_q - 1):
                big[k] = sfr[k] * sfr[k]
        self.sf_big = big
        self.Pprod = 1
        for x in pp:
            self.Pprod *= x
        self._ks_cache = {}
        self._fft_cache = {}
        self.enc_counter = 0

    # ---- limb helpers -------------------------------------------------------------
    def ntt(self, a, mod_idx):
        self.ntt_rows(a.reshape(1, -1), [mod_idx])

    def intt(self, a, mod_idx):
        self.ntt_rows(a.reshape(1, -1), [mod_idx], inverse=True)

    def ntt_rows(self, a, mod_ids, inverse=False):
        """in-place (inverse) NTT of the rows of a[rows][n]; row r is a limb modulo moduli[mod_ids[r]]"""
        assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"] and a.shape == (len(mod_ids), self.n)
        L = lib()
        if SLOW:
            for r, mi in enumerate(mod_ids):
                if inverse:
                    L.orc_ntt_inv(P(a[r]), self.log_n, self.moduli[mi], P(self.psi_inv_br[mi]))
                else:
                    L.orc_ntt_fwd(P(a[r]), self.log_n, self.moduli[mi], P(self.psi_br[mi]))
            return
        ids = (ctypes.c_int32 * len(mod_ids))(*mod_ids)
        tab, tsh = (self.psi_inv_br, self.psi_inv_br_sh) if inverse else (self.psi_br, self.psi_br_sh)
        L.orf_ntt_batch(P(a), len(mod_ids), self.log_n, ids, P(self.mod_arr), P(tab), P(tsh), P(self.ninv_arr),
                        1 if inverse else 0)

    def ks_tables(self, size_ql):
        """tables of EvalKeySwitchPrecomputeCore / ApproxModDown at a given number of Q limbs"""
        if size_ql in self._ks_cache:
            return self._ks_cache[size_ql]
        q, pp, a = self.q, self.pmod, self.alpha
        nparts = min((size_ql + a - 1) // a, self.num_part_q)
        parts = []
        for part in range(nparts):
            dig = list(range(a * part, min(a * (part + 1), size_ql)))
            compl = [i for i in range(size_ql) if i not in dig] + [self.size_q + k for k in range(self.size_p)]
            qin = [q[i] for i in dig]
            qout = [self.moduli[i] for i in compl]
            hat = []
            for i in dig:
                h = 1
                for k in dig:
                    if k != i:
                        h *= q[k]
                hat.append(h)
            hatinv = np.array([pow(h % qi, -1, qi) for h, qi in zip(hat, qin)], dtype=np.uint64)
            hatmod = np.array([[h % t for t in qout] for h in hat], dtype=np.uint64)
            parts.append(dict(dig=dig, compl=compl, qin=np.array(qin, dtype=np.uint64),
                              qout=np.array(qout, dtype=np.uint64), hatinv=hatinv, hatmod=np.ascontiguousarray(hatmod)))
        # mod-down P -> Q_l
        phat = [self.Pprod // x for x in pp]
        md = dict(
            pin=np.array(pp, dtype=np.uint64), qout=np.array(q[:size_ql], dtype=np.uint64),
            phatinv=np.array([pow(h % x, -1, x) for h, x in zip(phat, pp)], dtype=np.uint64),
            phatmod=np.ascontiguousarray(np.array([[h % q[i] for i in range(size_ql)] for h in phat], dtype=np.uint64)),
            pinv=[pow(self.Pprod % q[i], -1, q[i]) for i in range(size_ql)])
        self._ks_cache[size_ql] = (parts, md)
        return parts, md

    def fft_tables(self):
        if "t" not in self._fft_cache:
            M = 2 * self.n
            k = np.arange(M + 1, dtype=np.float64)
            ang = 2.0 * math.pi * k / M
            ksi = np.empty(2 * (M + 1), dtype=np.float64)
            ksi[0::2] = [math.cos(x) for x in ang]
            ksi[1::2] = [math.sin(x) for x in ang]
            rot = np.empty(self.n // 2, dtype=np.uint32)
            f = 1
            for j in range(self.n // 2):
                rot[j] = f
                f = (f * 5) % M
            self._fft_cache["t"] = (ksi, rot, M)
        return self._fft_cache["t"]
