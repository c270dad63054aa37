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

/* The sorter layer keeps one CryptoContext wrapper per engine context; it owns the encoded-plaintext cache (the 0/1 masks
 * and index patterns of reference src/sort_algo.h:341,356,573,706,718 depend on (N, ring, level) only and are encoded once
 * per context instead of once per sort).  sortfhe_ctx_release drops that wrapper and its cached plaintext handles: call it
 * before ckks_ctx_destroy.  sortfhe_plaintext_cache_stats reports entries / device bytes (upper bound) / hits / misses. */
int sortfhe_ctx_release(ckks_ctx_t* ctx);
int sortfhe_plaintext_cache_stats(ckks_ctx_t* ctx, int64_t* entries, int64_t* bytes, int64_t* hits, int64_t* misses);

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

/* The shardable halves of DirectSort for shapes with num_batch > 1 (N >= 512 at ring 2^17; any N > ring/2N on small
 * rings): the comparison batches of constructRank (src/sort_algo.h:438-492) and the index-check batches
 * (:713-742) are independent.  *_partial sums the batches [b_lo, b_hi); exactly one shard passes seed = 1 (the
 * accumulator's cached encryption of zeros, :433); the partial sums are combined limb-wise (ckks_ct_device_ptr +
 * NCCL all-reduce + ckks_ct_mod_reduce) and handed to finish_* (rotate-and-sum, :494-504, :744-748).  With one shard
 * [0, num_batch) the result equals sortfhe_construct_rank / sortfhe_rotation_index_check bit for bit. */
int sortfhe_num_batches(int n, int ring_dim, int* num_batch);

/* Hybrid sorters (src/sort_algo.h:1051-1065 sort_hybrid, :1214-1230 sort_hybrid1, :1375-1391 sort_hybrid2): constructRank
 * followed by the matrix-style index check of :893-1049 / :1067-1211 / :1232-1372.  variant 0 = sort_hybrid (sinc below
 * N = 256, composite-sign indicator above), 1 = sort_hybrid1 (MEHP24 indicatorAdv), 2 = sort_hybrid2 (sinc).  The
 * reference's PrivateKey argument is debug-only and has no counterpart.  Parameters: multiplicative depth and rotation
 * key set of tests/DirectSortHTest.cpp:28-92, DirectSortH1Test.cpp:37-101, DirectSortH2Test.cpp:37-101. */
int sortfhe_hybrid_sort_params(int n, int variant, int ring_dim, int* mult_depth, int32_t* rotations, int cap, int* count);
int sortfhe_hybrid_sort(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int variant, int sign_n, int dg, int df,
                        const int32_t* rot_indices, int rot_count, ckks_ct_t** out);
int sortfhe_hybrid_index_check(ckks_ctx_t* ctx, const ckks_ct_t* rank, const ckks_ct_t* input, int n, int variant,
                               const int32_t* rot_indices, int rot_count, ckks_ct_t** out);
int sortfhe_construct_rank_partial(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int sign_n, int dg, int df,
                                   const int32_t* rot_indices, int rot_count, int b_lo, int b_hi, int seed, ckks_ct_t** out);
int sortfhe_finish_rank(ckks_ctx_t* ctx, const ckks_ct_t* partial, int n, const int32_t* rot_indices, int rot_count, ckks_ct_t** out);
int sortfhe_index_check_partial(ckks_ctx_t* ctx, const ckks_ct_t* rank, const ckks_ct_t* input, int n,
                                const int32_t* rot_indices, int rot_count, int b_lo, int b_hi, int seed, ckks_ct_t** out);
int sortfhe_finish_index_check(ckks_ctx_t* ctx, const ckks_ct_t* partial, int n, const int32_t* rot_indices, int rot_count,
                               ckks_ct_t** out);

/* Comparison::compare (src/comparison.cpp:4-22) and sign() (src/sign.cpp:635-651), CompositeSign only */
int sortfhe_compare(ckks_ctx_t* ctx, const ckks_ct_t* a, const ckks_ct_t* b, int sign_n, int dg, int df, ckks_ct_t** out);
int sortfhe_sign(ckks_ctx_t* ctx, const ckks_ct_t* x, int sign_n, int dg, int df, ckks_ct_t** out);
/* sign(x, cc, func, cfg) with func in the order of enum SignFunc (src/sign.h:6-11): 0 CompositeSign, 2 Tanh (degree-1006
 * Chebyshev, sign.cpp:625-633), 3 NaiveDiscrete (degree-119 step, sign.cpp:610-623); 1 SignumPolycircuit is refused */
int sortfhe_sign_func(ckks_ctx_t* ctx, const ckks_ct_t* x, int func, int sign_n, int dg, int df, ckks_ct_t** out);
/* Comparison::indicator (src/comparison.cpp:24-40) */
int sortfhe_indicator(ckks_ctx_t* ctx, const ckks_ct_t* x, double c, int sign_n, int dg, int df, ckks_ct_t** out);

/* RotationComposer<N>::rotate (src/rotation.h:215-233) and Decomposer<N>::decompose (src/rotation.h:53-103);
 * algo: 0 NAF, 1 BNAF, 2 BINARY.  steps receives the signed step sizes. */
int sortfhe_rotate(ckks_ctx_t* ctx, const ckks_ct_t* x, int n, int rotation, const int32_t* rot_indices, int rot_count,
                   ckks_ct_t** out);
/* RotationTree<N> (src/rotation.h:239-358): builds the tree for the rotations start..end and returns treeRotate(x, r)
 * for every r in that range (hoisted EvalFastRotation below the root and below steps wider than 2); outs receives
 * end - start + 1 fresh handles */
int sortfhe_tree_rotate(ckks_ctx_t* ctx, const ckks_ct_t* x, int n, int start, int end, int algo, const int32_t* rot_indices,
                        int rot_count, ckks_ct_t** outs);
int sortfhe_decompose(int n, int rotation, int wrap, int algo, const int32_t* rot_indices, int rot_count, int32_t* steps,
                      int cap, int* count);

/* selectDoubledSincCoefficients<N>() / selectCoefficients<N>() (src/sort_algo.h:725-728, 627-629) */
int sortfhe_sinc_coefficients(int n, int doubled, double* coeffs, int cap, int* count);

/* MEHP24 ranking sort (reference src/mehp24/): getRotationIndices (mehp24_utils.cpp:186-213) + the fixture's depth
 * table (tests/mehp24/Mehp24SortTest.cpp:39-72); sortFG single-ciphertext, Comparison-based overload
 * (mehp24_sort.cpp:251-282) on a length-n vector stored in an n x n slot matrix; the matrix utilities
 * (mehp24_utils.cpp:44-103; op: 0 replicateRow, 1 replicateColumn, 2 sumRows, 3 sumColumns, 4 transposeRow,
 * 5 transposeColumn); indicatorAdv (mehp24_utils.cpp:166-174). */
int sortfhe_mehp24_params(int n, int* mult_depth, int32_t* rotations, int cap, int* count);
int sortfhe_mehp24_sort_fg(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int sign_n, int dg, int df, int dg_i, int df_i,
                           ckks_ct_t** out);
/* mehp24::sortLargeArrayFG (src/mehp24/mehp24_sort.cpp:625-645): splitCiphertext -> multi-ciphertext sortFG (:445-605) ->
 * combineCiphertext, for total_len = m * sub_len values in one ciphertext of sub_len^2 slots (the reference uses sub_len = 256
 * for N = 512..2048, tests/mehp24/Mehp24SortTest.cpp:150-157).  Needs the rotation keys of sortfhe_mehp24_params(total_len). */
int sortfhe_mehp24_sort_large(ckks_ctx_t* ctx, const ckks_ct_t* input, int total_len, int sub_len, int sign_n, int dg, int df,
                              int dg_i, int df_i, ckks_ct_t** out);
/* The MEHP24 overloads that do not go through Comparison (SURVEY 8 A19 "other overloads"): sortFG with the f/g-composition
 * predicates compareAdv / indicatorAdv, single ciphertext (src/mehp24/mehp24_sort.cpp:219-249) and split -> multi-ciphertext
 * (:285-443) -> combine (sortLargeArrayFG :607-623); `sort` with the Chebyshev-interpolated compare / indicator predicates
 * of degree degree_c / degree_i on [left_bound_c, right_bound_c], single ciphertext (:23-47: the sorted values land in
 * column 0 of the matrix, NOT moved back to row 0) and over nc blocks (:49-217, out[nc]). */
int sortfhe_mehp24_sort_fg_adv(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int dg_c, int df_c, int dg_i, int df_i, ckks_ct_t** out);
int sortfhe_mehp24_sort_large_adv(ckks_ctx_t* ctx, const ckks_ct_t* input, int total_len, int sub_len, int dg_c, int df_c, int dg_i,
                                  int df_i, ckks_ct_t** out);
int sortfhe_mehp24_sort_cheb(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, double left_bound_c, double right_bound_c, int degree_c,
                             int degree_i, ckks_ct_t** out);
int sortfhe_mehp24_sort_cheb_multi(ckks_ctx_t* ctx, const ckks_ct_t* const* parts, int nc, int sub_len, double left_bound_c,
                                   double right_bound_c, int degree_c, int degree_i, ckks_ct_t** out);
/* Multi-ciphertext sortFG in shardable stages (the nc (nc + 1) / 2 block comparisons of
 * src/mehp24/mehp24_sort.cpp:484-522 and the nc target blocks of :554-603 are independent; the reference runs them under
 * OpenMP with `omp critical` accumulation).  split / combine = splitCiphertext / combineCiphertext (mehp24_utils.cpp:266-302;
 * `out` of split holds total_len / sub_len handles).  block_compare_partial runs comparison units [u_lo, u_hi) of the upper
 * triangle (j <= k, row-major) and returns the nc column-block sums cv_out[j] and the row-block sums ch_out[k] (k >= 1;
 * ch_out[0] = NULL); accumulators the range never touched are exact zeros of the same level and scale, so the partial sums
 * of all shards are combined limb-wise (ckks_ct_device_ptr + all-reduce + ckks_ct_mod_reduce).  block_ranks turns the
 * combined sums into the nc rank matrices (:524-552), block_sorted computes target block j from them (:554-603).  With the
 * full unit range the four calls are exactly sortfhe_mehp24_sort_large. */
int sortfhe_mehp24_split(ckks_ctx_t* ctx, const ckks_ct_t* input, int total_len, int sub_len, ckks_ct_t** out);
int sortfhe_mehp24_combine(ckks_ctx_t* ctx, const ckks_ct_t* const* parts, int nc, int sub_len, ckks_ct_t** out);
int sortfhe_mehp24_block_compare_partial(ckks_ctx_t* ctx, const ckks_ct_t* const* parts, int nc, int sub_len, int sign_n, int dg, int df,
                                         int u_lo, int u_hi, ckks_ct_t** cv_out, ckks_ct_t** ch_out);
int sortfhe_mehp24_block_ranks(ckks_ctx_t* ctx, const ckks_ct_t* const* cv, const ckks_ct_t* const* ch, int nc, int sub_len,
                               ckks_ct_t** s_out);
int sortfhe_mehp24_block_sorted(ckks_ctx_t* ctx, const ckks_ct_t* const* parts, const ckks_ct_t* const* s, int nc, int sub_len, int j,
                                int dg_i, int df_i, ckks_ct_t** out);
int sortfhe_mehp24_matrix_op(ckks_ctx_t* ctx, const ckks_ct_t* input, int n, int op, int mask_output, ckks_ct_t** out);
int sortfhe_mehp24_indicator_adv(ckks_ctx_t* ctx, const ckks_ct_t* x, double b, int dg, int df, ckks_ct_t** out);

/* k-way sorting network of HKC+21 (reference src/k-way/, src/kway_adapter.h): KWayAdapter<N>::getSizeParameters
 * (kway_adapter.h:41-64); Masking sortType / getRotateDistance / genIndices (Masking.cpp:26-167); Sorter::sorter
 * (Sorter.cpp:284-404) over stages [first_stage, last_stage) (last_stage < 0: all), k in {2, 3}, on a copy of the
 * input as KWayAdapter::sort does.  mult_depth is SignConfig::multDepth, the budget checkLevelAndBoot tests; the
 * call fails where the reference would bootstrap (EvalBootstrap is not implemented). */
int sortfhe_kway_params(int n, int* mult_depth, int* scaling_mod_size, int* first_mod_size, int32_t* rotations, int cap, int* count);
int sortfhe_kway_sort_type(int k, int M, int stage, int* m, int* log_dist, int* slope, int* stage_count, int64_t* rotate_distance);
int sortfhe_kway_gen_indices(int num_slots, int k, int M, int m, int log_dist, int slope, int32_t* out);
int sortfhe_kway_sort(ckks_ctx_t* ctx, const ckks_ct_t* input, int num_slots, int k, int M, int sign_n, int dg, int df,
                      int mult_depth, int first_stage, int last_stage, ckks_ct_t** out);

#ifdef __cplusplus
}
#endif
#endif
