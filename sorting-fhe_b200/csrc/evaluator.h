// evaluator.h -- internal C++ interface of the evaluator layer (see evaluator.cu, cheby.cu)
#pragma once
#include "engine.h"
#include <vector>

void context_generate(Context* c, int log_n, int mult_depth, int scaling_bits, int first_bits, int num_large_digits, int ext,
                      uint64_t seed, int dry, const uint64_t* given_moduli = nullptr, const uint64_t* given_roots = nullptr,
                      int given_count = 0);

Ct* ct_new(Context* c, int parts, int nl, int level, int deg, double sf, int slots, int lrep = 0);
Ct* ct_full(const Ct* a);   // full-length copy of a (possibly compact) plaintext / ciphertext
Ct* ct_clone(const Ct* a);
void ct_free(Ct* a);
void ct_drop_caches(const Ct* a);   // frees the cached twins and the ModUp digits of a handle whose limbs are about to change
void ct_wait(const Ct* a);   // order the main stream after the (side-stream) producer of a

void mult_core_const(Ct* a, double x);
void rescale(Ct* a);
void level_reduce(Ct* a, int levels);
void adjust_levels_and_depth(Ct* a, Ct* b);
u64* key_switch_modup(Context* c, const u64* src, int nl);
void key_switch(Context* c, const u64* src, int nl, const EvalKey& key, u64* out, const u64* add0, const u64* add1,
                uint32_t auto_k, const u64* modup = nullptr);

Ct* eval_add(const Ct* a, const Ct* b, int sign);
Ct* eval_add_const(const Ct* a, double x);
Ct* eval_const_sub(double x, const Ct* a);
Ct* eval_double_add_const(const Ct* a, double x);
Ct* eval_negate(const Ct* a);
Ct* eval_mult(const Ct* a, const Ct* b);
Ct* eval_mult_const(const Ct* a, double x);
Ct* eval_wsum_aligned(const std::vector<Ct*>& cts, const std::vector<double>& w);   // nullptr if operands are not aligned
Ct* eval_rotate(const Ct* a, int r);
Ct* fast_rotation_precompute(const Ct* a);
Ct* eval_fast_rotate(const Ct* a, int r, const Ct* pre);
uint32_t auto_index(const Context* c, int r);

void keygen(Context* c);
void relin_keygen(Context* c);
void rot_keygen(Context* c, const int* indices, int count);
Ct* encode(Context* c, const double* values, int nvals, int scale_deg, int level, int slots, int is_complex = 0);
void conj_keygen(Context* c);                                   // key of the conjugation automorphism X -> X^(2n-1)
Ct* eval_automorphism(const Ct* a, uint32_t k, const char* what);   // key switch + index permutation for a keyed automorphism
Ct* eval_conjugate(const Ct* a);
Ct* mod_raise(const Ct* a, int level);                          // bootstrapping: q0 residues re-read in every limb of `level`
Ct* eval_add_many(const std::vector<const Ct*>& cts);
// bootstrap.cu
void boot_setup(Context* c, int budget_c2s, int budget_s2c, int slots);
void boot_keygen(Context* c, int slots);
Ct* boot_eval(const Ct* ct);
void boot_release(Context* c);
Ct* encrypt(Context* c, const Ct* pt);
void decrypt(const Ct* ct, double* out, int out_len);

// cheby.cu
Ct* eval_linear_wsum_mutable(std::vector<Ct*>& cts, const std::vector<double>& weights);
Ct* eval_chebyshev_series_ps(const Ct* x, const std::vector<double>& coeffs, double a, double b);
Ct* eval_chebyshev_series_linear(const Ct* x, const std::vector<double>& coeffs, double a, double b);
Ct* eval_chebyshev_series(const Ct* x, const std::vector<double>& coeffs, double a, double b);
bool eval_wsum_aligned_multi(const std::vector<Ct*>& cts, const std::vector<std::vector<double>>& w, Ct** outs);
Ct* eval_inner_product_plain(const Ct* init, const Ct* const* cts, const Ct* const* pts, int cnt);
void eval_inner_product_plain_multi(const Ct* init, const Ct* const* cts, const Ct* const* pts, int cnt, int groups, Ct** outs);
Ct* eval_poly_linear(const Ct* x, const std::vector<double>& coeffs);
void long_division_chebyshev(const std::vector<double>& f, const std::vector<double>& g, std::vector<double>& q,
                             std::vector<double>& r);
void compute_degrees_ps(int n, int& k, int& m);
