/* sortfhe_host.h -- C entry points of libsortfhe_host.so: the sorter layer of oksuman/sorting-fhe
 * (sort_algo.h / comparison.h / sign.h / rotation.h, re-created in sorting-fhe_b200/host/ on top of the engine's
 * C-ABI) exposed with plain pointers so that non-C++ callers (the Python harness, bench.py, tests) drive exactly
 * the code a C++ user of the reference's headers would.  Handles are the ones of ckks_b200.h.
 * Every function returns 0 on success; on failure the message is in sortfhe_last_error().
 */
#ifndef SORTFHE_HOST_H
#define SORTFHE_HOST_H
#include "ckks_b200.h"
#ifdef __cplusplus
extern "C" {
#endif

const char* sortfhe_last_error(void);

/* DirectSort<N>::getSizeParameters (reference src/sort_algo.h:87-201): multiplicative depth and rotation key
 * set for array size n at the given ring dimension.  Returns the number of rotations through *count. */
int sortfhe_direct_sort_params(int n, int ring_dim, int* mult_depth, int32_t* rotations, int cap, int* count);

/* DirectSort<N>::sort / constructRank / rotationIndexCheckN (src/sort_algo.h:752-774, 368-506, 658-750) with
 * sign(CompositeSign, SignConfig(CompositeSignConfig(sign_n, dg, df))).  rot_indices = the rotation keys that
 * exist (RotationComposer decomposes the others, src/rotation.h:215-233). */
int sortfhe_direct_sort(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int sign_n, int dg, int df,
                        const int32_t* rot_indices, int rot_count, ckks_ct_t** out);
int sortfhe_construct_rank(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int sign_n, int dg, int df,
                           const int32_t* rot_indices, int rot_count, ckks_ct_t** out);
int sortfhe_rotation_index_check(ckks_ctx_t* ctx, const ckks_ct_t* rank, const ckks_ct_t* input, int n,
                                 const int32_t* rot_indices, int rot_count, ckks_ct_t** out);

/* Comparison::compare (src/comparison.cpp:4-22) and sign() (src/sign.cpp:635-651), CompositeSign only */
int sortfhe_compare(ckks_ctx_t* ctx, const ckks_ct_t* a, const ckks_ct_t* b, int sign_n, int dg, int df, ckks_ct_t** out);
int sortfhe_sign(ckks_ctx_t* ctx, const ckks_ct_t* x, int sign_n, int dg, int df, ckks_ct_t** out);
/* Comparison::indicator (src/comparison.cpp:24-40) */
int sortfhe_indicator(ckks_ctx_t* ctx, const ckks_ct_t* x, double c, int sign_n, int dg, int df, ckks_ct_t** out);

/* RotationComposer<N>::rotate (src/rotation.h:215-233) and Decomposer<N>::decompose (src/rotation.h:53-103);
 * algo: 0 NAF, 1 BNAF, 2 BINARY.  steps receives the signed step sizes. */
int sortfhe_rotate(ckks_ctx_t* ctx, const ckks_ct_t* x, int n, int rotation, const int32_t* rot_indices, int rot_count,
                   ckks_ct_t** out);
int sortfhe_decompose(int n, int rotation, int wrap, int algo, const int32_t* rot_indices, int rot_count, int32_t* steps,
                      int cap, int* count);

/* selectDoubledSincCoefficients<N>() / selectCoefficients<N>() (src/sort_algo.h:725-728, 627-629) */
int sortfhe_sinc_coefficients(int n, int doubled, double* coeffs, int cap, int* count);

#ifdef __cplusplus
}
#endif
#endif
