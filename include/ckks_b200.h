/* ckks_b200.h -- C-ABI of the B200-native RNS-CKKS evaluation engine (libckks_b200.so).
 *
 * Drop-in boundary for oksuman/sorting-fhe.  The reference has no FFI: its seam is the set of
 * lbcrypto::CryptoContextImpl<DCRTPoly> member functions its headers call after
 * `#include "openfhe.h"` (src/sort_algo.h:3-12, src/comparison.h:3-9, src/sign.h:3-4,
 * src/rotation.h:1-10, src/sort.h:1-13).  Every entry point below replaces one of those member
 * functions (cited per function); sorting-fhe_b200/host/openfhe_shim.h re-creates the lbcrypto
 * value types on top of this ABI so the sorter headers compile unchanged in spirit.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success, non-zero on
 * failure with the message available from ckks_last_error() (thread-local) -- OpenFHE's
 * OpenFHEException becomes a status code, the C++ shim rethrows std::runtime_error.
 * Ciphertexts/plaintexts are opaque handles owning device memory, limb-major uint64
 * [part][limb][ring_dim], EVALUATION (bit-reversed NTT) format; results are fresh handles,
 * inputs are never mutated except by the *_set_* metadata calls (SetSlots semantics,
 * src/sort_algo.h:429,434,447,501,710-711).  There is no CPU fallback: without a CUDA device
 * every compute call fails, except on a context created with dry_run=1 (metadata only).
 */
#ifndef CKKS_B200_H
#define CKKS_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct ckks_ctx ckks_ctx_t; /* CryptoContext<DCRTPoly> */
typedef struct ckks_ct ckks_ct_t;   /* Ciphertext<DCRTPoly> (2 polys) or Plaintext (1 poly) */

const char* ckks_last_error(void);

/* GenCryptoContext(CCParams<CryptoContextCKKSRNS>) -- tests/DirectSortTest.cpp:29-44,
 * src/sort_algo.h:87-201 (SetMultiplicativeDepth, SetScalingModSize, SetRingDim),
 * src/kway_adapter.h:45-46 (SetFirstModSize).  scaling_technique: 1 = FLEXIBLEAUTOEXT (OpenFHE
 * default), 0 = FLEXIBLEAUTO.  num_large_digits 0 = OpenFHE default (at most 4 digits are supported).  seed drives the
 * counter-based sampler for keys and encryption randomness: a non-zero seed makes every key and ciphertext
 * reproducible (parity tests; NOT secure -- anyone who knows it can recompute the secret key); seed 0 draws the
 * key of the sampler from OS entropy (getrandom).  See INTEGRATION.md "Randomness". */
int ckks_ctx_create(ckks_ctx_t** out, int log_ring_dim, int mult_depth, int scaling_mod_size, int first_mod_size,
                    int num_large_digits, int scaling_technique, uint64_t seed, int dry_run);
/* Context whose moduli (Q chain then P) and primitive 2n-th roots of unity come from an imported OpenFHE context
 * (reference src/sort.h:34 DeserializeFromFile(ccLocation, ...)) instead of being re-derived; count must equal the limb
 * counts the parameters imply, every modulus must be an NTT prime.  roots may be NULL (minimal roots are derived). */
int ckks_ctx_create_from_moduli(ckks_ctx_t** out, int log_ring_dim, int mult_depth, int scaling_mod_size, int first_mod_size,
                                int num_large_digits, int scaling_technique, uint64_t seed, int dry_run,
                                const uint64_t* moduli, const uint64_t* roots, int count);
int ckks_ctx_destroy(ckks_ctx_t* ctx); /* ClearEvalMultKeys/ClearEvalAutomorphismKeys/ReleaseAllContexts */
int ckks_ctx_set_stream(ckks_ctx_t* ctx, void* cuda_stream);
int ckks_ctx_sync(ckks_ctx_t* ctx);
/* what: 0 ring dim (GetRingDimension, sort_algo.h:80), 1 #Q limbs, 2 #P limbs, 3 digit size alpha, 4 #digits,
 * 5 cyclotomic order (rotation.h:267), 6 kernel launches so far, 7 device bytes of keys, 8 process-unique context id,
 * 9 the engine's cudaStream_t (as an integer: collectives are enqueued on it, sorting-fhe_b200/shard.py),
 * 10 free device memory in bytes (the sorter layer sizes its encoded-plaintext cache from it) */
int ckks_ctx_info(ckks_ctx_t* ctx, int what, int64_t* out);
/* the n-th Encrypt of a context draws its randomness from sampler stream n; resetting the counter replays it
 * (reproducible tests: e.g. two sorter objects built after the same reset hold the same encryption of zeros) */
/* options: 0 = EvalAdd(ct, double) convention on noiseScaleDeg > 1 (0: round(|c| * ciphertext scale), default;
 * 1: OpenFHE 1.1.4 GetElementForEvalAddOrSub form, round(|c| * sf_level) times round(sf_level) per extra degree);
 * 1 = number of ciphertext handles that may keep their hoisted ModUp digits (default 8, 0 disables the cache);
 * 2 = capacity in MiB of the temporaries arena (freed temporaries are kept in exact-size bins per stream and reused
 *     without a cudaMallocAsync / cudaFreeAsync call; default: an eighth of the free device memory, at most 16 GiB;
 *     0 hands every cached buffer back to the pool and disables the arena; env CKKS_B200_ARENA_MB sets the default) */
int ckks_ctx_set_option(ckks_ctx_t* ctx, int what, int64_t value);
int ckks_ctx_set_encryption_counter(ckks_ctx_t* ctx, uint64_t counter);
int ckks_ctx_get_moduli(ckks_ctx_t* ctx, uint64_t* moduli, uint64_t* roots, int cap);   /* Q then P */
int ckks_ctx_get_scaling_factors(ckks_ctx_t* ctx, double* sf_real, double* sf_big, int cap);

/* KeyGen / EvalMultKeyGen / EvalRotateKeyGen -- tests/DirectSortTest.cpp:40-57 */
int ckks_keygen(ckks_ctx_t* ctx);
int ckks_eval_mult_keygen(ckks_ctx_t* ctx);
int ckks_eval_rotate_keygen(ckks_ctx_t* ctx, const int32_t* indices, int count);
/* raw key material (host buffers), normal form: secret [Q+P][n]; pk [2][Q][n] (b,a);
 * evk (kind 0 = relin, else signed rotation index via is_rotation) [digit][2 (b,a)][Q+P][n].
 * This is the hook the OpenFHE BINARY importer (sort.h:31-74) feeds. */
int ckks_export_secret(ckks_ctx_t* ctx, uint64_t* host, size_t cap_words);
int ckks_export_public_key(ckks_ctx_t* ctx, uint64_t* host, size_t cap_words);
int ckks_import_secret(ckks_ctx_t* ctx, const uint64_t* host, size_t words);        /* [M][n] EVALUATION (tests; src/sort.h never loads one) */
int ckks_import_public_key(ckks_ctx_t* ctx, const uint64_t* host, size_t words);    /* [2 (b, a)][sizeQ][n], sort.h:40 */
int ckks_export_eval_key(ckks_ctx_t* ctx, int is_rotation, int32_t index, uint64_t* host, size_t cap_words);
/* is_rotation: 0 = the relinearisation key, 1 = `index` is a rotation amount, 2 = `index` is the automorphism index (5^r mod 2n,
 * or 2n - 1 for the conjugation) -- the key OpenFHE's EvalAutomorphismKey map uses (src/sort.h:63) */
int ckks_list_automorphism_keys(ckks_ctx_t* ctx, uint32_t* automorphism_indices, int cap, int* count);
int ckks_import_eval_key(ckks_ctx_t* ctx, int is_rotation, int32_t index, const uint64_t* host, size_t words);

/* MakeCKKSPackedPlaintext(vec, scaleDeg, level, nullptr, slots) -- sort_algo.h:341,356,573,706,715.
 * slots = 0 -> next power of two >= count. */
int ckks_make_packed_plaintext(ckks_ctx_t* ctx, const double* values, int count, int scale_deg, int level, int slots,
                               ckks_ct_t** out);
int ckks_encrypt(ckks_ctx_t* ctx, const ckks_ct_t* plaintext, ckks_ct_t** out);            /* encryption.cpp:10 */
int ckks_decrypt(ckks_ctx_t* ctx, const ckks_ct_t* ct, double* values, int count);        /* encryption.cpp:18,31 */

/* handle management + metadata (Clone, SetSlots/GetSlots, GetLevel, GetNoiseScaleDeg, GetScalingFactor) */
int ckks_ct_free(ckks_ct_t* ct);
/* ct->Clone() (src/sort_algo.h:330,433,563-566).  Copy-on-write: the clone shares the source's limbs -- and the rescaled /
 * level-adjusted twins and ModUp digits cached on them -- until ckks_ct_set_slots, ckks_ct_mod_reduce or ckks_ct_device_ptr is
 * applied to either handle; every evaluator entry point is out of place on its operands, so callers see value semantics. */
int ckks_ct_clone(const ckks_ct_t* ct, ckks_ct_t** out);
int ckks_ct_set_slots(ckks_ct_t* ct, int slots);
/* what: 0 slots, 1 level, 2 noiseScaleDeg, 3 #limbs, 4 #polys */
int ckks_ct_get(const ckks_ct_t* ct, int what, int64_t* out);
int ckks_ct_get_scaling_factor(const ckks_ct_t* ct, double* out);
/* host <-> device transfer of the raw limbs (the e2e path of bench.py; OpenFHE import format hook) */
int ckks_ct_export(const ckks_ct_t* ct, uint64_t* host, size_t cap_words);
int ckks_ct_import(ckks_ctx_t* ctx, const uint64_t* host, int polys, int limbs, int level, int noise_scale_deg,
                   double scaling_factor, int slots, ckks_ct_t** out);

/* multi-GPU exchange of partial ciphertext sums (replaces the `omp critical` + EvalAddInPlace accumulation of
 * sort_algo.h:471-472, 740-741 when the comparison batches are sharded across GPUs): the caller all-reduces (SUM) the
 * raw uint64 limbs in place through the device pointer (NCCL), then ckks_ct_mod_reduce folds every word back below
 * its limb modulus.  Exact for up to 8 addends (residues < 2^61). */
int ckks_ct_device_ptr(const ckks_ct_t* ct, void** dev_ptr, size_t* words);
int ckks_ct_mod_reduce(ckks_ctx_t* ctx, ckks_ct_t* ct);

/* EvalAdd / EvalSub over (ct,ct) (ct,pt) (pt,ct) -- ~130 call sites, e.g. sort_algo.h:358,454,708,720 */
int ckks_eval_add(ckks_ctx_t* ctx, const ckks_ct_t* a, const ckks_ct_t* b, ckks_ct_t** out);
int ckks_eval_sub(ckks_ctx_t* ctx, const ckks_ct_t* a, const ckks_ct_t* b, ckks_ct_t** out);
int ckks_eval_add_const(ckks_ctx_t* ctx, const ckks_ct_t* a, double c, ckks_ct_t** out);   /* comparison.cpp:19 */
int ckks_eval_sub_const(ckks_ctx_t* ctx, const ckks_ct_t* a, double c, ckks_ct_t** out);   /* sort_algo.h:504 */
int ckks_eval_const_sub(ckks_ctx_t* ctx, double c, const ckks_ct_t* a, ckks_ct_t** out);   /* T = init; for i < count: T = EvalAdd(T, EvalMult(cts[i], pts[i])) -- the masked baby-step sums of vecRotsOpt and
 * blindRotationOptN (src/sort_algo.h:336-345, 567-577) as one call.  Same limbs as the chain of calls; aligned operands
 * take a single fused pass over HBM. */
int ckks_eval_inner_product_plain(ckks_ctx_t* ctx, const ckks_ct_t* init, const ckks_ct_t* const* cts, const ckks_ct_t* const* pts,
                                  int count, ckks_ct_t** out);
/* `groups` such sums over the same ciphertexts and seed, outs[g] = init + sum_i cts[i] * pts[g * count + i]: the num_partition / np
 * mask sets of one vecRotsOpt / blindRotationOptN call (src/sort_algo.h:331-352, 563-580).  Same limbs as `groups` single calls;
 * the ciphertexts are read once per pass of (by default) four outputs. */
int ckks_eval_inner_product_plain_multi(ckks_ctx_t* ctx, const ckks_ct_t* init, const ckks_ct_t* const* cts, const ckks_ct_t* const* pts,
                                        int count, int groups, ckks_ct_t** outs);
/* comparison.cpp:38 */
int ckks_eval_negate(ckks_ctx_t* ctx, const ckks_ct_t* a, ckks_ct_t** out);
int ckks_eval_add_many(ckks_ctx_t* ctx, const ckks_ct_t* const* cts, int count, ckks_ct_t** out); /* sort_algo.h:364 */
/* EvalMult(ct,ct) == EvalMultAndRelinearize, EvalMult(ct,pt), EvalMult(ct,double), EvalSquare -- sign.cpp:23-56 */
int ckks_eval_mult(ckks_ctx_t* ctx, const ckks_ct_t* a, const ckks_ct_t* b, ckks_ct_t** out);
int ckks_eval_mult_const(ckks_ctx_t* ctx, const ckks_ct_t* a, double c, ckks_ct_t** out);
int ckks_eval_square(ckks_ctx_t* ctx, const ckks_ct_t* a, ckks_ct_t** out);
/* EvalRotate(ct, index) -- rotation.h:224,231; fails if no key for index (OpenFHE throws) */
int ckks_eval_rotate(ckks_ctx_t* ctx, const ckks_ct_t* a, int32_t index, ckks_ct_t** out);
int ckks_has_rotation_key(ckks_ctx_t* ctx, int32_t index);
/* EvalFastRotationPrecompute(ct) / EvalFastRotation(ct, index, M, digits) -- rotation.h:286,344-347 (RotationTree).
 * The precomputed ModUp digits are a handle of their own (free with ckks_ct_free); results are bit-identical to
 * ckks_eval_rotate.  ckks_eval_rotate itself keeps the digits of a ciphertext it has rotated once, so repeated
 * rotations of one handle (the sorters' baby steps, sort_algo.h:425-431,733-736) are hoisted transparently. */
int ckks_fast_rotation_precompute(ckks_ctx_t* ctx, const ckks_ct_t* ct, ckks_ct_t** precomp);
int ckks_fast_rotation(ckks_ctx_t* ctx, const ckks_ct_t* ct, int32_t index, const ckks_ct_t* precomp, ckks_ct_t** out);
/* EvalChebyshevSeriesPS(ct, coeffs, a, b) -- sort_algo.h:727-728, sign.cpp:76 */
int ckks_eval_chebyshev_series_ps(ckks_ctx_t* ctx, const ckks_ct_t* x, const double* coeffs, int count, double a, double b,
                                  ckks_ct_t** out);
/* EvalChebyshevSeries(ct, coeffs, a, b): the linear evaluator below degree 5, Paterson-Stockmeyer from there on; the
 * series step of EvalChebyshevFunction (mehp24_utils.cpp:109-163, sign.cpp:616-633 -- the coefficients of the
 * function are computed on the host, see host/openfhe_shim.h EvalChebyshevCoefficients) */
int ckks_eval_chebyshev_series(ckks_ctx_t* ctx, const ckks_ct_t* x, const double* coeffs, int count, double a, double b,
                               ckks_ct_t** out);
/* EvalPolyLinear(ct, coeffs) -- mehp24_utils.cpp:256-259 */
/* CKKS bootstrapping (reference src/k-way/EvalUtils.cpp:76, src/sign.cpp:168; set-up as tests/k-way/KWaySort2Test.cpp:47-49):
 * EvalBootstrapSetup(levelBudget, dim1, slots) precomputes the CoeffsToSlots / SlotsToCoeffs diagonals for `slots` (sparse
 * packing, slots <= ring / 4; dim1 is accepted and ignored -- the diagonals of a group are evaluated with hoisted rotations),
 * EvalBootstrapKeyGen generates the rotation and conjugation keys, EvalBootstrap refreshes a ciphertext whose level is
 * below the multiplicative depth.  ckks_make_packed_plaintext_complex / ckks_eval_conjugate / ckks_mod_raise are the three
 * primitives bootstrapping adds to the engine (exposed for the parity tests). */
int ckks_eval_bootstrap_setup(ckks_ctx_t* ctx, int budget_c2s, int budget_s2c, int dim1_c2s, int dim1_s2c, int slots);
int ckks_eval_bootstrap_keygen(ckks_ctx_t* ctx, int slots);
int ckks_eval_bootstrap(ckks_ctx_t* ctx, const ckks_ct_t* ct, ckks_ct_t** out);
int ckks_make_packed_plaintext_complex(ckks_ctx_t* ctx, const double* re_im, int count, int scale_deg, int level, int slots,
                                       ckks_ct_t** out);
int ckks_eval_conjugate(ckks_ctx_t* ctx, const ckks_ct_t* ct, ckks_ct_t** out);
int ckks_mod_raise(ckks_ctx_t* ctx, const ckks_ct_t* ct, int level, ckks_ct_t** out);

int ckks_eval_poly_linear(ckks_ctx_t* ctx, const ckks_ct_t* x, const double* coeffs, int count, ckks_ct_t** out);

/* op statistics (JSON) for roofline accounting; reset clears counters */
int ckks_stats_json(ckks_ctx_t* ctx, char* buf, size_t cap);
int ckks_stats_reset(ckks_ctx_t* ctx);

/* per-kernel-class device timing (CUDA events on the engine stream around every launch of the selected
 * classes).  class bits: 0 ntt, 1 base conversion, 2 evk inner product, 3 mod-down epilogue, 4 pointwise,
 * 5 rescale, 6 automorphism, 7 encode, 8 copies.  ckks_profile_json synchronises, returns the totals
 * {"class": [ms, launches, algorithmic_bytes], ...} accumulated since the last call, and clears them. */
int ckks_profile_enable(ckks_ctx_t* ctx, unsigned class_mask);
int ckks_profile_json(ckks_ctx_t* ctx, char* buf, size_t cap);

/* kernel-level entry points on raw buffers, for parity tests and kernel benchmarks.
 * *_host variants stage through host memory; *_dev variants take device pointers. */
/* the ChaCha20 block function (RFC 8439) behind the production sampler (contexts created with seed 0: key = 256 bits of
 * getrandom(2), counter = sample index, nonce = sampler stream), evaluated on the host: known-answer hook for the tests */
int ckks_rng_block(const uint32_t* key8, uint64_t stream, uint64_t index, uint32_t* out16);
int ckks_ntt_host(ckks_ctx_t* ctx, uint64_t* data, int rows, const int32_t* modulus_ids, int inverse);
int ckks_ntt_dev(ckks_ctx_t* ctx, void* dev_data, int rows, int first_modulus, int inverse);
int ckks_key_switch_dev(ckks_ctx_t* ctx, const void* dev_src, int limbs, int is_rotation, int32_t index, void* dev_out);
/* No counterpart in the reference: added for callers that cannot allocate (CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same arguments and the same, bit-identical
 * result as ckks_key_switch_dev (dev_src [limbs][n] -> dev_out [2][limbs][n], EVALUATION format), but every temporary
 * lives in one file-scope static device array of 16 MiB (2^21 words) per GPU, resolved once per device with
 * cudaGetSymbolAddress; the call performs no cudaMalloc / cudaFree / pool allocation.  It needs
 * limbs + sum_j (limbs + sizeP - |digit j|) + 2 (limbs + sizeP) + 2 sizeP + 2 limbs rows of n words; a request beyond
 * 2^21 words is rejected with an error that states both sizes (ring 2^13 fits up to 256 rows; a full-depth key switch
 * at ring 2^17 does not fit and stays with ckks_key_switch_dev).  Preconditions, each reported as an error:
 * the evaluation key exists; the constant tables of this limb count exist (they are allocated on first use -- one call of
 * ckks_key_switch_dev or any EvalMult / EvalRotate at this level during set-up); per-class profiling is off; the context's
 * device is current.  The first call on a device does the symbol lookup and must not run under stream capture.  All calls
 * on a device share the array (it keeps the previous call's contents; every region is written before it is read): calls
 * that overlap on the GPU -- two contexts or streams -- race and must be serialised by the caller. */
int ckks_key_switch_dev_noalloc(ckks_ctx_t* ctx, const void* dev_src, int limbs, int is_rotation, int32_t index, void* dev_out);

#ifdef __cplusplus
}
#endif
#endif
