"""sorting-fhe_b200 -- B200-native RNS-CKKS engine behind the oksuman/sorting-fhe sorter API.

This module is only the ctypes binding of the C-ABI in include/ckks_b200.h (libckks_b200.so,
hand-written sm_100a CUDA) plus the binding of the host-side sorter library
(libsortfhe_host.so: the C++ re-creation of the reference's sort_algo.h / comparison.h / sign.h /
rotation.h).  There is no Python or CPU compute path: if the shared library is missing the import
fails loudly, and every compute call fails without a CUDA device (dry_run contexts carry metadata only).

The package directory name contains a hyphen, so import it with
    importlib.import_module("sorting-fhe_b200")
"""
import ctypes
import json
import os

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CKKS_B200_LIB") or os.path.join(_DIR, "libckks_b200.so")   # env: kernel A/B builds (tools/build_variant.sh)
if not os.path.exists(LIB_PATH):
    raise ImportError("libckks_b200.so is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(the CUDA extension is mandatory; there is no fallback path)")
_L = ctypes.CDLL(LIB_PATH)
_L.ckks_last_error.restype = ctypes.c_char_p
_vp, _i, _d, _u64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_uint64
_pp = ctypes.POINTER(ctypes.c_void_p)


class CkksError(RuntimeError):
    pass


def rng_block(key, stream, index):
    """ChaCha20 block (RFC 8439) of the production sampler, host instance (ckks_rng_block): key = 32 bytes -> 64 bytes"""
    import struct
    k = (ctypes.c_uint32 * 8)(*struct.unpack("<8I", bytes(key)))
    o = (ctypes.c_uint32 * 16)()
    _chk(_L.ckks_rng_block(k, ctypes.c_uint64(stream), ctypes.c_uint64(index), o))
    return struct.pack("<16I", *o)


def _chk(rc):
    if rc != 0:
        raise CkksError(_L.ckks_last_error().decode())


def _u64p(a):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64))


def _f64p(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


class Ciphertext:
    """Ciphertext<DCRTPoly> / Plaintext handle (value semantics of the shared_ptr are the caller's)"""

    def __init__(self, eng, h):
        self.eng, self.h = eng, h

    def __del__(self):
        try:
            if self.h:
                _L.ckks_ct_free(ctypes.c_void_p(self.h))
                self.h = None
        except Exception:
            pass

    def _get(self, what):
        o = ctypes.c_int64()
        _chk(_L.ckks_ct_get(_vp(self.h), what, ctypes.byref(o)))
        return o.value

    slots = property(lambda s: s._get(0), lambda s, v: _chk(_L.ckks_ct_set_slots(_vp(s.h), int(v))))
    level = property(lambda s: s._get(1))
    deg = property(lambda s: s._get(2))
    nlimbs = property(lambda s: s._get(3))
    parts = property(lambda s: s._get(4))

    @property
    def sf(self):
        o = ctypes.c_double()
        _chk(_L.ckks_ct_get_scaling_factor(_vp(self.h), ctypes.byref(o)))
        return o.value

    def clone(self):
        o = ctypes.c_void_p()
        _chk(_L.ckks_ct_clone(_vp(self.h), ctypes.byref(o)))
        return Ciphertext(self.eng, o.value)

    def export(self):
        """raw limbs [parts][nlimbs][n] (device -> host)"""
        a = np.empty((self.parts, self.nlimbs, self.eng.n), dtype=np.uint64)
        _chk(_L.ckks_ct_export(_vp(self.h), _u64p(a), a.size))
        return a

    # OpenFHE spellings used by the reference code
    def Clone(self):
        return self.clone()

    def SetSlots(self, s):
        self.slots = s

    def GetSlots(self):
        return self.slots

    def GetLevel(self):
        return self.level


class Engine:
    """CryptoContext<DCRTPoly> over the C-ABI"""

    def __init__(self, log_n, mult_depth, scaling_mod_size=40, first_mod_size=60, num_large_digits=0, ext=True,
                 seed=1, dry_run=False, stream=None):
        h = ctypes.c_void_p()
        _chk(_L.ckks_ctx_create(ctypes.byref(h), log_n, mult_depth, scaling_mod_size, first_mod_size, num_large_digits,
                                1 if ext else 0, _u64(seed), 1 if dry_run else 0))
        self.h = h.value
        self.log_n, self.n, self.dry = log_n, 1 << log_n, dry_run
        if stream is not None:
            _chk(_L.ckks_ctx_set_stream(_vp(self.h), _vp(stream)))
        self.size_q, self.size_p = self.info(1), self.info(2)
        self.alpha, self.num_part_q = self.info(3), self.info(4)

    ctx = property(lambda s: s)  # so algorithm code written against oracle.OracleCC (cc.ctx.n) also drives the engine

    def close(self):
        if self.h:
            _H.sortfhe_ctx_release(_vp(self.h))      # sorter-layer wrapper + its cached plaintext handles
            _L.ckks_ctx_destroy(_vp(self.h))
            self.h = None

    def set_option(self, what, value):
        """ckks_ctx_set_option: what = "add_const_mode" (0 | 1) or "modup_cache" (handles that keep hoisted ModUp digits)"""
        _chk(_L.ckks_ctx_set_option(_vp(self.h), {"add_const_mode": 0, "modup_cache": 1}[what], ctypes.c_int64(value)))

    def plaintext_cache_stats(self):
        """(entries, device bytes upper bound, hits, misses) of the sorter layer's encoded-plaintext cache"""
        v = [ctypes.c_int64() for _ in range(4)]
        _hchk(_H.sortfhe_plaintext_cache_stats(_vp(self.h), *[ctypes.byref(x) for x in v]))
        return tuple(x.value for x in v)

    def info(self, what):
        o = ctypes.c_int64()
        _chk(_L.ckks_ctx_info(_vp(self.h), what, ctypes.byref(o)))
        return o.value

    def sync(self):
        _chk(_L.ckks_ctx_sync(_vp(self.h)))

    def moduli(self):
        m = np.zeros(self.size_q + self.size_p, dtype=np.uint64)
        r = np.zeros_like(m)
        _chk(_L.ckks_ctx_get_moduli(_vp(self.h), _u64p(m), _u64p(r), m.size))
        return [int(x) for x in m], [int(x) for x in r]

    def scaling_factors(self):
        a = np.zeros(self.size_q, dtype=np.float64)
        b = np.zeros(self.size_q, dtype=np.float64)
        _chk(_L.ckks_ctx_get_scaling_factors(_vp(self.h), _f64p(a), _f64p(b), a.size))
        return a, b[:max(self.size_q - 1, 0)]

    def stats(self, reset=False):
        buf = ctypes.create_string_buffer(1 << 20)
        _chk(_L.ckks_stats_json(_vp(self.h), buf, len(buf)))
        if reset:
            _chk(_L.ckks_stats_reset(_vp(self.h)))
        return json.loads(buf.value.decode())

    KCLASSES = ["ntt", "baseconv", "ks_inner", "moddown", "pointwise", "rescale", "automorph", "encode", "copy"]

    def profile_enable(self, classes=None):
        """CUDA-event timing of kernel classes (None = all, [] = off)"""
        mask = sum(1 << self.KCLASSES.index(k) for k in (self.KCLASSES if classes is None else classes))
        _chk(_L.ckks_profile_enable(_vp(self.h), ctypes.c_uint(mask)))

    def profile(self):
        """{class: {"ms", "launches", "bytes"}} since the last call"""
        buf = ctypes.create_string_buffer(1 << 16)
        _chk(_L.ckks_profile_json(_vp(self.h), buf, len(buf)))
        return {k: {"ms": v[0], "launches": int(v[1]), "bytes": v[2]} for k, v in json.loads(buf.value.decode()).items()}

    # ---- keys
    def KeyGen(self):
        _chk(_L.ckks_keygen(_vp(self.h)))

    def EvalMultKeyGen(self):
        _chk(_L.ckks_eval_mult_keygen(_vp(self.h)))

    def EvalRotateKeyGen(self, indices):
        a = (ctypes.c_int32 * len(indices))(*indices)
        _chk(_L.ckks_eval_rotate_keygen(_vp(self.h), a, len(indices)))

    def export_secret(self):
        a = np.empty((self.size_q + self.size_p, self.n), dtype=np.uint64)
        _chk(_L.ckks_export_secret(_vp(self.h), _u64p(a), a.size))
        return a

    def export_public_key(self):
        a = np.empty((2, self.size_q, self.n), dtype=np.uint64)
        _chk(_L.ckks_export_public_key(_vp(self.h), _u64p(a), a.size))
        return a

    def export_eval_key(self, rotation=None):
        a = np.empty((self.num_part_q, 2, self.size_q + self.size_p, self.n), dtype=np.uint64)
        _chk(_L.ckks_export_eval_key(_vp(self.h), 0 if rotation is None else 1, 0 if rotation is None else rotation,
                                     _u64p(a), a.size))
        return a

    def import_eval_key(self, a, rotation=None):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        _chk(_L.ckks_import_eval_key(_vp(self.h), 0 if rotation is None else 1, 0 if rotation is None else rotation,
                                     _u64p(a), a.size))

    # ---- encode / encrypt / decrypt
    def _out(self, rc, o):
        _chk(rc)
        return Ciphertext(self, o.value)

    def MakeCKKSPackedPlaintext(self, values, scale_deg=1, level=0, params=None, slots=None):
        v = np.ascontiguousarray(values, dtype=np.float64)
        o = ctypes.c_void_p()
        return self._out(_L.ckks_make_packed_plaintext(_vp(self.h), _f64p(v), v.size, scale_deg, level, slots or 0,
                                                       ctypes.byref(o)), o)

    encode = lambda self, values, scale_deg=1, level=0, slots=None: self.MakeCKKSPackedPlaintext(values, scale_deg, level, None, slots)

    def encrypt(self, pt):
        o = ctypes.c_void_p()
        return self._out(_L.ckks_encrypt(_vp(self.h), _vp(pt.h), ctypes.byref(o)), o)

    def decrypt(self, ct, length=None):
        n = ct.slots if length is None else length
        out = np.zeros(n, dtype=np.float64)
        _chk(_L.ckks_decrypt(_vp(self.h), _vp(ct.h), _f64p(out), n))
        return out

    decode = decrypt

    def import_ct(self, arr, level, deg, sf, slots):
        arr = np.ascontiguousarray(arr, dtype=np.uint64)
        o = ctypes.c_void_p()
        return self._out(_L.ckks_ct_import(_vp(self.h), _u64p(arr), arr.shape[0], arr.shape[1], level, deg, _d(sf), slots,
                                           ctypes.byref(o)), o)

    def import_ct_ptr(self, host_ptr, polys, limbs, level, deg, sf, slots):
        """ckks_ct_import from a raw host pointer (e.g. a pinned torch tensor's data_ptr())"""
        o = ctypes.c_void_p()
        return self._out(_L.ckks_ct_import(_vp(self.h), ctypes.cast(host_ptr, ctypes.POINTER(ctypes.c_uint64)), polys, limbs,
                                           level, deg, _d(sf), slots, ctypes.byref(o)), o)

    def export_ct_ptr(self, ct, host_ptr, cap_words):
        _chk(_L.ckks_ct_export(_vp(ct.h), ctypes.cast(host_ptr, ctypes.POINTER(ctypes.c_uint64)), cap_words))

    # ---- evaluation (OpenFHE names)
    def _bin(self, f, a, b):
        o = ctypes.c_void_p()
        return self._out(f(_vp(self.h), _vp(a.h), _vp(b.h), ctypes.byref(o)), o)

    def _cst(self, f, a, x):
        o = ctypes.c_void_p()
        return self._out(f(_vp(self.h), _vp(a.h), _d(float(x)), ctypes.byref(o)), o)

    def EvalAdd(self, a, b):
        if isinstance(b, (int, float)):
            return self._cst(_L.ckks_eval_add_const, a, b)
        return self._bin(_L.ckks_eval_add, a, b)

    def EvalSub(self, a, b):
        if isinstance(a, (int, float)):
            o = ctypes.c_void_p()
            return self._out(_L.ckks_eval_const_sub(_vp(self.h), _d(float(a)), _vp(b.h), ctypes.byref(o)), o)
        if isinstance(b, (int, float)):
            return self._cst(_L.ckks_eval_sub_const, a, b)
        return self._bin(_L.ckks_eval_sub, a, b)

    def EvalNegate(self, a):
        o = ctypes.c_void_p()
        return self._out(_L.ckks_eval_negate(_vp(self.h), _vp(a.h), ctypes.byref(o)), o)

    def EvalAddMany(self, cts):
        arr = (ctypes.c_void_p * len(cts))(*[c.h for c in cts])
        o = ctypes.c_void_p()
        return self._out(_L.ckks_eval_add_many(_vp(self.h), arr, len(cts), ctypes.byref(o)), o)

    def EvalInnerProductPlain(self, init, cts, pts):
        """T = init; T = EvalAdd(T, EvalMult(cts[i], pts[i])) for every i, one engine call (same limbs as the chain)"""
        a = (ctypes.c_void_p * len(cts))(*[c.h for c in cts])
        p = (ctypes.c_void_p * len(pts))(*[c.h for c in pts])
        o = ctypes.c_void_p()
        return self._out(_L.ckks_eval_inner_product_plain(_vp(self.h), _vp(init.h), a, p, len(cts), ctypes.byref(o)), o)

    def EvalInnerProductPlainMulti(self, init, cts, pts_groups):
        """[EvalInnerProductPlain(init, cts, pts) for pts in pts_groups] as one engine call (ciphertexts read once per pass of four outputs)"""
        flat = [p for g in pts_groups for p in g]
        if any(len(g) != len(cts) for g in pts_groups):
            raise ValueError("EvalInnerProductPlainMulti: size mismatch")
        a = (ctypes.c_void_p * len(cts))(*[c.h for c in cts])
        p = (ctypes.c_void_p * len(flat))(*[c.h for c in flat])
        outs = (ctypes.c_void_p * len(pts_groups))()
        _chk(_L.ckks_eval_inner_product_plain_multi(_vp(self.h), _vp(init.h), a, p, len(cts), len(pts_groups), outs))
        return [Ciphertext(self, h) for h in outs]

    def EvalMult(self, a, b):
        if isinstance(b, (int, float)):
            return self._cst(_L.ckks_eval_mult_const, a, b)
        return self._bin(_L.ckks_eval_mult, a, b)

    EvalMultAndRelinearize = EvalMult

    def EvalSquare(self, a):
        o = ctypes.c_void_p()
        return self._out(_L.ckks_eval_square(_vp(self.h), _vp(a.h), ctypes.byref(o)), o)

    def EvalRotate(self, a, r):
        o = ctypes.c_void_p()
        return self._out(_L.ckks_eval_rotate(_vp(self.h), _vp(a.h), int(r), ctypes.byref(o)), o)

    def set_encryption_counter(self, counter):
        _chk(_L.ckks_ctx_set_encryption_counter(_vp(self.h), ctypes.c_uint64(counter)))

    def ct_device_ptr(self, ct):
        """(device pointer, words) of a ciphertext's limbs -- for collectives on the data where it lies"""
        p, w = ctypes.c_void_p(), ctypes.c_size_t()
        _chk(_L.ckks_ct_device_ptr(_vp(ct.h), ctypes.byref(p), ctypes.byref(w)))
        return p.value, w.value

    def ct_mod_reduce(self, ct):
        _chk(_L.ckks_ct_mod_reduce(_vp(self.h), _vp(ct.h)))

    # ---- bootstrapping (csrc/bootstrap.cu)
    def EvalBootstrapSetup(self, level_budget=(4, 4), dim1=(0, 0), slots=0):
        _chk(_L.ckks_eval_bootstrap_setup(_vp(self.h), int(level_budget[0]), int(level_budget[1]), int(dim1[0]), int(dim1[1]), int(slots)))

    def EvalBootstrapKeyGen(self, slots=0):
        _chk(_L.ckks_eval_bootstrap_keygen(_vp(self.h), int(slots)))

    def EvalBootstrap(self, ct, iters=1, prec=0):
        o = ctypes.c_void_p()
        return self._out(_L.ckks_eval_bootstrap(_vp(self.h), _vp(ct.h), ctypes.byref(o)), o)

    def EvalConjugate(self, ct):
        o = ctypes.c_void_p()
        return self._out(_L.ckks_eval_conjugate(_vp(self.h), _vp(ct.h), ctypes.byref(o)), o)

    def mod_raise(self, ct, level):
        o = ctypes.c_void_p()
        return self._out(_L.ckks_mod_raise(_vp(self.h), _vp(ct.h), int(level), ctypes.byref(o)), o)

    def encode_complex(self, values, scale_deg=1, level=0, slots=None):
        v = np.asarray(values, dtype=np.complex128)
        if slots is None:
            slots = 1
            while slots < len(v):
                slots <<= 1
        buf = np.empty(2 * len(v), dtype=np.float64)
        buf[0::2], buf[1::2] = v.real, v.imag
        o = ctypes.c_void_p()
        return self._out(_L.ckks_make_packed_plaintext_complex(_vp(self.h), _f64p(buf), len(v), scale_deg, level, slots, ctypes.byref(o)), o)

    def EvalFastRotationPrecompute(self, a):
        o = ctypes.c_void_p()
        return self._out(_L.ckks_fast_rotation_precompute(_vp(self.h), _vp(a.h), ctypes.byref(o)), o)

    def EvalFastRotation(self, a, r, m, digits):
        o = ctypes.c_void_p()
        return self._out(_L.ckks_fast_rotation(_vp(self.h), _vp(a.h), int(r), _vp(digits.h), ctypes.byref(o)), o)

    def EvalChebyshevSeriesPS(self, x, coeffs, a=-1.0, b=1.0):
        v = np.ascontiguousarray(coeffs, dtype=np.float64)
        o = ctypes.c_void_p()
        return self._out(_L.ckks_eval_chebyshev_series_ps(_vp(self.h), _vp(x.h), _f64p(v), v.size, _d(a), _d(b),
                                                          ctypes.byref(o)), o)

    def EvalChebyshevSeries(self, x, coeffs, a=-1.0, b=1.0):
        v = np.ascontiguousarray(coeffs, dtype=np.float64)
        o = ctypes.c_void_p()
        return self._out(_L.ckks_eval_chebyshev_series(_vp(self.h), _vp(x.h), _f64p(v), v.size, _d(a), _d(b),
                                                       ctypes.byref(o)), o)

    @staticmethod
    def EvalChebyshevCoefficients(func, a, b, degree):
        """interpolation at the degree+1 Chebyshev nodes of [a, b]; c_0 doubled (halved again by the evaluators)"""
        cnt = degree + 1
        j = np.arange(cnt) + 0.5
        f = np.array([func(v) for v in np.cos(np.pi / cnt * j) * 0.5 * (b - a) + 0.5 * (b + a)], dtype=np.float64)
        return np.array([np.sum(f * np.cos(np.pi / cnt * i * j)) for i in range(cnt)]) * 2.0 / cnt

    def EvalChebyshevFunction(self, func, x, a, b, degree):
        """mehp24_utils.cpp:109-163, sign.cpp:616-633"""
        return self.EvalChebyshevSeries(x, self.EvalChebyshevCoefficients(func, a, b, degree), a, b)

    def EvalPolyLinear(self, x, coeffs):
        v = np.ascontiguousarray(coeffs, dtype=np.float64)
        o = ctypes.c_void_p()
        return self._out(_L.ckks_eval_poly_linear(_vp(self.h), _vp(x.h), _f64p(v), v.size, ctypes.byref(o)), o)

    # ---- raw kernels
    def ntt_host(self, data, mod_ids, inverse=False):
        data = np.ascontiguousarray(data, dtype=np.uint64)
        ids = (ctypes.c_int32 * len(mod_ids))(*mod_ids)
        _chk(_L.ckks_ntt_host(_vp(self.h), _u64p(data), data.shape[0], ids, 1 if inverse else 0))
        return data

    def ntt_dev(self, dev_ptr, rows, first_mod=0, inverse=False):
        _chk(_L.ckks_ntt_dev(_vp(self.h), _vp(dev_ptr), rows, first_mod, 1 if inverse else 0))

    def key_switch_dev(self, src_ptr, limbs, out_ptr, rotation=None):
        _chk(_L.ckks_key_switch_dev(_vp(self.h), _vp(src_ptr), limbs, 0 if rotation is None else 1,
                                    0 if rotation is None else rotation, _vp(out_ptr)))

    def key_switch_dev_noalloc(self, src_ptr, limbs, out_ptr, rotation=None):
        """allocation-free variant over the library's static device scratch (include/ckks_b200.h; no reference counterpart)"""
        _chk(_L.ckks_key_switch_dev_noalloc(_vp(self.h), _vp(src_ptr), limbs, 0 if rotation is None else 1,
                                            0 if rotation is None else rotation, _vp(out_ptr)))


_L.ckks_eval_bootstrap_setup.argtypes = [_vp, _i, _i, _i, _i, _i]
_L.ckks_eval_bootstrap_keygen.argtypes = [_vp, _i]
_L.ckks_eval_bootstrap.argtypes = [_vp, _vp, _pp]
_L.ckks_eval_conjugate.argtypes = [_vp, _vp, _pp]
_L.ckks_mod_raise.argtypes = [_vp, _vp, _i, _pp]
_L.ckks_make_packed_plaintext_complex.argtypes = [_vp, ctypes.POINTER(ctypes.c_double), _i, _i, _i, _i, _pp]
_L.ckks_ctx_set_option.argtypes = [_vp, _i, ctypes.c_int64]
_L.ckks_ctx_create.argtypes = [_pp, _i, _i, _i, _i, _i, _i, _u64, _i]
_L.ckks_ct_import.argtypes = [_vp, ctypes.POINTER(ctypes.c_uint64), _i, _i, _i, _i, _d, _i, _pp]
_L.ckks_eval_add_const.argtypes = [_vp, _vp, _d, _pp]
_L.ckks_eval_sub_const.argtypes = [_vp, _vp, _d, _pp]
_L.ckks_eval_mult_const.argtypes = [_vp, _vp, _d, _pp]
_L.ckks_eval_const_sub.argtypes = [_vp, _d, _vp, _pp]
_L.ckks_eval_chebyshev_series.argtypes = [_vp, _vp, ctypes.POINTER(ctypes.c_double), _i, _d, _d, _pp]
_L.ckks_eval_chebyshev_series_ps.argtypes = [_vp, _vp, ctypes.POINTER(ctypes.c_double), _i, _d, _d, _pp]
_L.ckks_make_packed_plaintext.argtypes = [_vp, ctypes.POINTER(ctypes.c_double), _i, _i, _i, _i, _pp]
_L.ckks_ct_export.argtypes = [_vp, ctypes.POINTER(ctypes.c_uint64), ctypes.c_size_t]
_L.ckks_export_secret.argtypes = [_vp, ctypes.POINTER(ctypes.c_uint64), ctypes.c_size_t]
_L.ckks_export_public_key.argtypes = [_vp, ctypes.POINTER(ctypes.c_uint64), ctypes.c_size_t]
_L.ckks_export_eval_key.argtypes = [_vp, _i, ctypes.c_int32, ctypes.POINTER(ctypes.c_uint64), ctypes.c_size_t]
_L.ckks_import_eval_key.argtypes = [_vp, _i, ctypes.c_int32, ctypes.POINTER(ctypes.c_uint64), ctypes.c_size_t]
_L.ckks_profile_json.argtypes = [_vp, ctypes.c_char_p, ctypes.c_size_t]
_L.ckks_profile_enable.argtypes = [_vp, ctypes.c_uint]
_L.ckks_stats_json.argtypes = [_vp, ctypes.c_char_p, ctypes.c_size_t]


# ------------------------------------------------------------------------------------------------
# sorter layer (libsortfhe_host.so, include/sortfhe_host.h): the reference's sort_algo.h / comparison.h /
# sign.h / rotation.h API, re-created in host C++ on top of the engine's C-ABI
HOST_LIB_PATH = os.path.join(_DIR, "libsortfhe_host.so")
if not os.path.exists(HOST_LIB_PATH):
    raise ImportError("libsortfhe_host.so is not built: run `python -c 'import __graft_entry__ as g; g.build()'`")
_H = ctypes.CDLL(HOST_LIB_PATH)
_H.sortfhe_last_error.restype = ctypes.c_char_p
_i32p = ctypes.POINTER(ctypes.c_int32)
_ip = ctypes.POINTER(ctypes.c_int)
_H.sortfhe_ctx_release.argtypes = [_vp]
_H.sortfhe_plaintext_cache_stats.argtypes = [_vp] + [ctypes.POINTER(ctypes.c_int64)] * 4
_H.sortfhe_direct_sort_params.argtypes = [_i, _i, _ip, _i32p, _i, _ip]
_H.sortfhe_hybrid_sort_params.argtypes = [_i, _i, _i, _ip, _i32p, _i, _ip]
_H.sortfhe_hybrid_sort.argtypes = [_vp, _vp, _i, _i, _i, _i, _i, _i32p, _i, _pp]
_H.sortfhe_hybrid_index_check.argtypes = [_vp, _vp, _vp, _i, _i, _i32p, _i, _pp]
_H.sortfhe_direct_sort.argtypes = [_vp, _vp, _i, _i, _i, _i, _i32p, _i, _pp]
_H.sortfhe_construct_rank.argtypes = [_vp, _vp, _i, _i, _i, _i, _i32p, _i, _pp]
_H.sortfhe_rotation_index_check.argtypes = [_vp, _vp, _vp, _i, _i32p, _i, _pp]
_H.sortfhe_compare.argtypes = [_vp, _vp, _vp, _i, _i, _i, _pp]
_H.sortfhe_sign.argtypes = [_vp, _vp, _i, _i, _i, _pp]
_H.sortfhe_sign_func.argtypes = [_vp, _vp, _i, _i, _i, _i, _pp]
_H.sortfhe_indicator.argtypes = [_vp, _vp, _d, _i, _i, _i, _pp]
_H.sortfhe_rotate.argtypes = [_vp, _vp, _i, _i, _i32p, _i, _pp]
_H.sortfhe_tree_rotate.argtypes = [_vp, _vp, _i, _i, _i, _i, _i32p, _i, _pp]
_H.sortfhe_decompose.argtypes = [_i, _i, _i, _i, _i32p, _i, _i32p, _i, _ip]
_H.sortfhe_sinc_coefficients.argtypes = [_i, _i, ctypes.POINTER(ctypes.c_double), _i, _ip]
_H.sortfhe_num_batches.argtypes = [_i, _i, _ip]
_H.sortfhe_construct_rank_partial.argtypes = [_vp, _vp, _i, _i, _i, _i, _i32p, _i, _i, _i, _i, _pp]
_H.sortfhe_finish_rank.argtypes = [_vp, _vp, _i, _i32p, _i, _pp]
_H.sortfhe_index_check_partial.argtypes = [_vp, _vp, _vp, _i, _i32p, _i, _i, _i, _i, _pp]
_H.sortfhe_finish_index_check.argtypes = [_vp, _vp, _i, _i32p, _i, _pp]
_H.sortfhe_kway_params.argtypes = [_i, _ip, _ip, _ip, _i32p, _i, _ip]
_H.sortfhe_kway_sort_type.argtypes = [_i, _i, _i, _ip, _ip, _ip, _ip, ctypes.POINTER(ctypes.c_int64)]
_H.sortfhe_kway_gen_indices.argtypes = [_i, _i, _i, _i, _i, _i, _i32p]
_H.sortfhe_kway_sort.argtypes = [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _pp]
_H.sortfhe_mehp24_params.argtypes = [_i, _ip, _i32p, _i, _ip]
_H.sortfhe_mehp24_sort_fg.argtypes = [_vp, _vp, _i, _i, _i, _i, _i, _i, _pp]
_H.sortfhe_mehp24_sort_large.argtypes = [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _pp]
_H.sortfhe_mehp24_split.argtypes = [_vp, _vp, _i, _i, _pp]
_H.sortfhe_mehp24_combine.argtypes = [_vp, _pp, _i, _i, _pp]
_H.sortfhe_mehp24_block_compare_partial.argtypes = [_vp, _pp, _i, _i, _i, _i, _i, _i, _i, _pp, _pp]
_H.sortfhe_mehp24_block_ranks.argtypes = [_vp, _pp, _pp, _i, _i, _pp]
_H.sortfhe_mehp24_block_sorted.argtypes = [_vp, _pp, _pp, _i, _i, _i, _i, _i, _pp]
_H.sortfhe_mehp24_matrix_op.argtypes = [_vp, _vp, _i, _i, _i, _pp]
_H.sortfhe_mehp24_indicator_adv.argtypes = [_vp, _vp, _d, _i, _i, _pp]


def _hchk(rc):
    if rc != 0:
        raise CkksError(_H.sortfhe_last_error().decode())


def _idx(indices):
    return (ctypes.c_int32 * len(indices))(*indices), len(indices)


def direct_sort_params(n, ring_dim=1 << 17):
    """DirectSort<N>::getSizeParameters -> (multDepth, rotation key indices)"""
    depth, cnt = ctypes.c_int(), ctypes.c_int()
    buf = (ctypes.c_int32 * 4096)()
    _hchk(_H.sortfhe_direct_sort_params(n, ring_dim, ctypes.byref(depth), buf, 4096, ctypes.byref(cnt)))
    return depth.value, list(buf[:cnt.value])


HYBRID_VARIANTS = {"hybrid": 0, "hybrid1": 1, "hybrid2": 2}


def hybrid_sort_params(n, variant="hybrid1", ring_dim=1 << 17):
    """(multDepth, rotation key indices) of sort_hybrid / sort_hybrid1 / sort_hybrid2 (tests/DirectSortH*Test.cpp)"""
    depth, cnt = ctypes.c_int(), ctypes.c_int()
    buf = (ctypes.c_int32 * 4096)()
    _hchk(_H.sortfhe_hybrid_sort_params(n, HYBRID_VARIANTS[variant], ring_dim, ctypes.byref(depth), buf, 4096, ctypes.byref(cnt)))
    return depth.value, list(buf[:cnt.value])


def num_batches(n, ring_dim=1 << 17):
    """num_batch of DirectSort<N> at a ring dimension (src/sort_algo.h:378-380)"""
    v = ctypes.c_int()
    _hchk(_H.sortfhe_num_batches(n, ring_dim, ctypes.byref(v)))
    return v.value


class DirectSort:
    """DirectSort<N> over an Engine (reference src/sort_algo.h:61-774)"""

    def __init__(self, eng, n, rot_indices):
        self.eng, self.n, self.rots = eng, n, list(rot_indices)

    def _run(self, f, *args):
        a, c = _idx(self.rots)
        o = ctypes.c_void_p()
        _hchk(f(_vp(self.eng.h), *args, a, c, ctypes.byref(o)))
        return Ciphertext(self.eng, o.value)

    def sort(self, ct, sign_cfg=(3, 3, 2)):
        return self._run(_H.sortfhe_direct_sort, _vp(ct.h), self.n, *sign_cfg)

    def construct_rank(self, ct, sign_cfg=(3, 3, 2)):
        return self._run(_H.sortfhe_construct_rank, _vp(ct.h), self.n, *sign_cfg)

    def rotation_index_check(self, rank, ct):
        return self._run(_H.sortfhe_rotation_index_check, _vp(rank.h), _vp(ct.h), self.n)

    def sort_hybrid(self, ct, sign_cfg=(3, 3, 2), variant="hybrid1"):
        """sort_hybrid / sort_hybrid1 / sort_hybrid2 (src/sort_algo.h:1051-1065, 1214-1230, 1375-1391)"""
        return self._run(_H.sortfhe_hybrid_sort, _vp(ct.h), self.n, HYBRID_VARIANTS[variant], *sign_cfg)

    def hybrid_index_check(self, rank, ct, variant="hybrid1"):
        return self._run(_H.sortfhe_hybrid_index_check, _vp(rank.h), _vp(ct.h), self.n, HYBRID_VARIANTS[variant])

    # shardable halves (multi-GPU, see shard.py): batches [lo, hi) of the two batch loops, and the folds after the all-reduce
    def _run_tail(self, f, head, *tail):
        a, c = _idx(self.rots)
        o = ctypes.c_void_p()
        _hchk(f(_vp(self.eng.h), *head, a, c, *tail, ctypes.byref(o)))
        return Ciphertext(self.eng, o.value)

    def construct_rank_partial(self, ct, sign_cfg, lo, hi, seed=True):
        return self._run_tail(_H.sortfhe_construct_rank_partial, (_vp(ct.h), self.n, *sign_cfg), lo, hi, int(seed))

    def finish_rank(self, partial):
        return self._run_tail(_H.sortfhe_finish_rank, (_vp(partial.h), self.n))

    def index_check_partial(self, rank, ct, lo, hi, seed=True):
        return self._run_tail(_H.sortfhe_index_check_partial, (_vp(rank.h), _vp(ct.h), self.n), lo, hi, int(seed))

    def finish_index_check(self, partial):
        return self._run_tail(_H.sortfhe_finish_index_check, (_vp(partial.h), self.n))


def compare(eng, a, b, sign_cfg=(4, 3, 3)):
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_compare(_vp(eng.h), _vp(a.h), _vp(b.h), *sign_cfg, ctypes.byref(o)))
    return Ciphertext(eng, o.value)


SIGN_FUNCS = {"CompositeSign": 0, "SignumPolycircuit": 1, "Tanh": 2, "NaiveDiscrete": 3}


def sign(eng, x, sign_cfg=(3, 0, 1), func="CompositeSign"):
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_sign_func(_vp(eng.h), _vp(x.h), SIGN_FUNCS[func], *sign_cfg, ctypes.byref(o)))
    return Ciphertext(eng, o.value)


def indicator(eng, x, c, sign_cfg=(3, 3, 2)):
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_indicator(_vp(eng.h), _vp(x.h), _d(c), *sign_cfg, ctypes.byref(o)))
    return Ciphertext(eng, o.value)


def rotate(eng, x, n, rotation, rot_indices):
    a, c = _idx(rot_indices)
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_rotate(_vp(eng.h), _vp(x.h), n, rotation, a, c, ctypes.byref(o)))
    return Ciphertext(eng, o.value)


def tree_rotate(eng, x, n, start, end, rot_indices, algo="NAF"):
    """RotationTree<N>::buildTree(start, end) + treeRotate for every rotation of the range -> list of ciphertexts"""
    a, c = _idx(rot_indices)
    outs = (ctypes.c_void_p * (end - start + 1))()
    _hchk(_H.sortfhe_tree_rotate(_vp(eng.h), _vp(x.h), n, start, end, {"NAF": 0, "BNAF": 1, "BINARY": 2}[algo], a, c, outs))
    return [Ciphertext(eng, h) for h in outs]


def decompose(n, rotation, wrap, rot_indices, algo="BINARY"):
    a, c = _idx(rot_indices)
    steps, cnt = (ctypes.c_int32 * 256)(), ctypes.c_int()
    _hchk(_H.sortfhe_decompose(n, rotation, wrap, {"NAF": 0, "BNAF": 1, "BINARY": 2}[algo], a, c, steps, 256,
                               ctypes.byref(cnt)))
    return list(steps[:cnt.value])


def sinc_coefficients(n, doubled=True):
    cnt = ctypes.c_int()
    buf = np.zeros(16384, dtype=np.float64)
    _hchk(_H.sortfhe_sinc_coefficients(n, 1 if doubled else 0, _f64p(buf), buf.size, ctypes.byref(cnt)))
    return buf[:cnt.value].copy()


# ---------------------------------------------------------------- MEHP24 (reference src/mehp24/)
MEHP24_OPS = {"replicateRow": 0, "replicateColumn": 1, "sumRows": 2, "sumColumns": 3, "transposeRow": 4, "transposeColumn": 5}


def mehp24_params(n):
    """(fixture multDepth, getRotationIndices(n)) -- tests/mehp24/Mehp24SortTest.cpp:39-72, mehp24_utils.cpp:186-213"""
    depth, cnt = ctypes.c_int(), ctypes.c_int()
    buf = (ctypes.c_int32 * 256)()
    _hchk(_H.sortfhe_mehp24_params(n, ctypes.byref(depth), buf, 256, ctypes.byref(cnt)))
    return depth.value, list(buf[:cnt.value])


def mehp24_sort_fg(eng, ct, n, sign_cfg=(3, 2, 2), dg_i=None, df_i=2):
    """mehp24::sortFG, single ciphertext (mehp24_sort.cpp:251-282); dg_i defaults to (log2 n + 1) / 2 as in the fixture"""
    if dg_i is None:
        dg_i = (int(np.log2(n)) + 1) // 2
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_mehp24_sort_fg(_vp(eng.h), _vp(ct.h), n, *sign_cfg, dg_i, df_i, ctypes.byref(o)))
    return Ciphertext(eng, o.value)


def mehp24_sort_large(eng, ct, total_len, sub_len, sign_cfg, dg_i=None, df_i=2):
    """mehp24::sortLargeArrayFG (split -> multi-ciphertext sortFG -> combine), src/mehp24/mehp24_sort.cpp:445-645"""
    if dg_i is None:
        dg_i = int((np.log2(total_len) + 1) / 2)
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_mehp24_sort_large(_vp(eng.h), _vp(ct.h), total_len, sub_len, *sign_cfg, dg_i, df_i, ctypes.byref(o)))
    return Ciphertext(eng, o.value)


def mehp24_sort_fg_adv(eng, ct, n, dg_c, df_c, dg_i, df_i):
    """mehp24::sortFG, compareAdv overload, single ciphertext (mehp24_sort.cpp:219-249)"""
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_mehp24_sort_fg_adv(_vp(eng.h), _vp(ct.h), n, dg_c, df_c, dg_i, df_i, ctypes.byref(o)))
    return Ciphertext(eng, o.value)


def mehp24_sort_large_adv(eng, ct, total_len, sub_len, dg_c, df_c, dg_i, df_i):
    """mehp24::sortLargeArrayFG, compareAdv overload (mehp24_sort.cpp:285-443, 607-623)"""
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_mehp24_sort_large_adv(_vp(eng.h), _vp(ct.h), total_len, sub_len, dg_c, df_c, dg_i, df_i, ctypes.byref(o)))
    return Ciphertext(eng, o.value)


def mehp24_sort_cheb(eng, ct, n, left_bound_c, right_bound_c, degree_c, degree_i):
    """mehp24::sort, single ciphertext (mehp24_sort.cpp:23-47): sorted values in column 0 of the n x n matrix"""
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_mehp24_sort_cheb(_vp(eng.h), _vp(ct.h), n, ctypes.c_double(left_bound_c), ctypes.c_double(right_bound_c),
                                      degree_c, degree_i, ctypes.byref(o)))
    return Ciphertext(eng, o.value)


def mehp24_sort_cheb_multi(eng, parts, sub_len, left_bound_c, right_bound_c, degree_c, degree_i):
    """mehp24::sort over several ciphertexts (mehp24_sort.cpp:49-217)"""
    outs = (ctypes.c_void_p * len(parts))()
    _hchk(_H.sortfhe_mehp24_sort_cheb_multi(_vp(eng.h), _harr(parts), len(parts), sub_len, ctypes.c_double(left_bound_c),
                                            ctypes.c_double(right_bound_c), degree_c, degree_i, outs))
    return [Ciphertext(eng, h) for h in outs]


def _harr(cts):
    """array of ciphertext handles (NULL for None) for the `const ckks_ct_t* const*` parameters"""
    return (ctypes.c_void_p * len(cts))(*[c.h if c is not None else None for c in cts])


def _hout(eng, arr):
    return [Ciphertext(eng, h) if h else None for h in arr]


def mehp24_split(eng, ct, total_len, sub_len):
    """mehp24::splitCiphertext (mehp24_utils.cpp:266-287)"""
    out = (ctypes.c_void_p * (total_len // sub_len))()
    _hchk(_H.sortfhe_mehp24_split(_vp(eng.h), _vp(ct.h), total_len, sub_len, out))
    return _hout(eng, out)


def mehp24_combine(eng, parts, sub_len):
    """mehp24::combineCiphertext (mehp24_utils.cpp:289-302)"""
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_mehp24_combine(_vp(eng.h), _harr(parts), len(parts), sub_len, ctypes.byref(o)))
    return Ciphertext(eng, o.value)


def mehp24_compare_units(nc):
    """number of block comparisons of the multi-ciphertext sortFG: the upper triangle of the nc x nc block matrix"""
    return nc * (nc + 1) // 2


def mehp24_block_compare_partial(eng, parts, sub_len, sign_cfg, u_lo, u_hi):
    """comparison units [u_lo, u_hi) of mehp24_sort.cpp:484-522 -> (Cv[nc], Ch[nc] with Ch[0] = None), zero-filled"""
    nc = len(parts)
    cv, ch = (ctypes.c_void_p * nc)(), (ctypes.c_void_p * nc)()
    _hchk(_H.sortfhe_mehp24_block_compare_partial(_vp(eng.h), _harr(parts), nc, sub_len, *sign_cfg, u_lo, u_hi, cv, ch))
    return _hout(eng, cv), _hout(eng, ch)


def mehp24_block_ranks(eng, cv, ch, sub_len):
    """rank matrices of the nc blocks from the combined comparison sums (mehp24_sort.cpp:524-552)"""
    out = (ctypes.c_void_p * len(cv))()
    _hchk(_H.sortfhe_mehp24_block_ranks(_vp(eng.h), _harr(cv), _harr(ch), len(cv), sub_len, out))
    return _hout(eng, out)


def mehp24_block_sorted(eng, parts, s, sub_len, j, dg_i, df_i=2):
    """target block j of the multi-ciphertext sortFG (mehp24_sort.cpp:554-603)"""
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_mehp24_block_sorted(_vp(eng.h), _harr(parts), _harr(s), len(parts), sub_len, j, dg_i, df_i, ctypes.byref(o)))
    return Ciphertext(eng, o.value)


def mehp24_matrix_op(eng, ct, n, op, mask_output=False):
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_mehp24_matrix_op(_vp(eng.h), _vp(ct.h), n, MEHP24_OPS[op], int(mask_output), ctypes.byref(o)))
    return Ciphertext(eng, o.value)


def mehp24_indicator_adv(eng, ct, b, dg, df):
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_mehp24_indicator_adv(_vp(eng.h), _vp(ct.h), float(b), dg, df, ctypes.byref(o)))
    return Ciphertext(eng, o.value)


# ---------------------------------------------------------------- k-way network (reference src/k-way/)
def kway_params(n):
    """KWayAdapter<N>::getSizeParameters -> (multDepth, scalingModSize, firstModSize, rotations)"""
    d, sc, fm, cnt = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    buf = (ctypes.c_int32 * 64)()
    _hchk(_H.sortfhe_kway_params(n, ctypes.byref(d), ctypes.byref(sc), ctypes.byref(fm), buf, 64, ctypes.byref(cnt)))
    return d.value, sc.value, fm.value, list(buf[:cnt.value])


def kway_sort_type(k, M, stage):
    """sortType + stage count + getRotateDistance -> (m, logDist, slope, stage_count, shift)"""
    m, ld, sl, sc = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    rd = ctypes.c_int64()
    _hchk(_H.sortfhe_kway_sort_type(k, M, stage, ctypes.byref(m), ctypes.byref(ld), ctypes.byref(sl), ctypes.byref(sc), ctypes.byref(rd)))
    return m.value, ld.value, sl.value, sc.value, rd.value


def kway_gen_indices(num_slots, k, M, m, log_dist, slope):
    out = np.zeros((2, num_slots), dtype=np.int32)
    _hchk(_H.sortfhe_kway_gen_indices(num_slots, k, M, m, log_dist, slope, out.ctypes.data_as(_i32p)))
    return out


def kway_sort(eng, ct, num_slots, k, M, sign_cfg=(3, 2, 2), mult_depth=100, first_stage=0, last_stage=-1):
    """kwaySort::Sorter::sorter over stages [first_stage, last_stage) (Sorter.cpp:284-404)"""
    o = ctypes.c_void_p()
    _hchk(_H.sortfhe_kway_sort(_vp(eng.h), _vp(ct.h), num_slots, k, M, *sign_cfg, mult_depth, first_stage, last_stage, ctypes.byref(o)))
    return Ciphertext(eng, o.value)
