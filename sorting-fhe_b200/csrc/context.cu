// context.cu -- CKKS parameter generation and device tables.
//
// Restates OpenFHE 1.1.4 ParameterGenerationCKKSRNS::ParamsGenCKKSRNS (moduli chain for
// FLEXIBLEAUTO / FLEXIBLEAUTOEXT, special primes for HYBRID key switching) and
// CryptoParametersCKKSRNS::PrecomputeCRTTables (per-level scaling factors, digit tables), i.e.
// what `GenCryptoContext(parameters)` does for the reference's fixtures
// (tests/DirectSortTest.cpp:29-44, src/sort_algo.h:87-201, src/kway_adapter.h:41-64).
// Host-side setup code only; nothing here is on the timed path.
#include "engine.h"
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <algorithm>
#include <atomic>
#include <sys/random.h>

typedef unsigned __int128 u128;

static u64 mulmod(u64 a, u64 b, u64 q) { return (u64)((u128)a * b % q); }
static u64 powmod(u64 a, u64 e, u64 q) { u64 r = 1 % q; a %= q; while (e) { if (e & 1) r = mulmod(r, a, q); a = mulmod(a, a, q); e >>= 1; } return r; }
static u64 invmod(u64 a, u64 q) { return powmod(a, q - 2, q); }
static bool is_prime(u64 n) {
    if (n < 2) return false;
    static const u64 sp[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    for (u64 p : sp) if (n % p == 0) return n == p;
    u64 d = n - 1; int s = 0; while (!(d & 1)) { d >>= 1; s++; }
    for (u64 a : sp) {
        u64 x = powmod(a, d, n); if (x == 1 || x == n - 1) continue;
        bool comp = true;
        for (int r = 1; r < s; r++) { x = mulmod(x, x, n); if (x == n - 1) { comp = false; break; } }
        if (comp) return false;
    }
    return true;
}
static u64 first_prime(int bits, u64 m) { u64 q = (u64)1 << bits, r = q % m, c = q + 1 - r; if (r) c += m; while (!is_prime(c)) c += m; return c; }
static u64 prev_prime(u64 q, u64 m) { u64 p = q - m; while (!is_prime(p)) p -= m; return p; }
static u64 next_prime(u64 q, u64 m) { u64 p = q + m; while (!is_prime(p)) p += m; return p; }
static u64 min_root(u64 m, u64 q) {
    u64 g = 0;
    for (u64 c = 2;; c++) { u64 r = powmod(c, (q - 1) / m, q); if (powmod(r, m / 2, q) == q - 1) { g = r; break; } }
    u64 g2 = mulmod(g, g, q), x = g, mn = g;
    for (u64 i = 1; i < m / 2; i++) { x = mulmod(x, g2, q); if (x < mn) mn = x; }
    return mn;
}
static uint32_t brev(uint32_t x, int bits) { uint32_t r = 0; for (int i = 0; i < bits; i++) { r = (r << 1) | (x & 1); x >>= 1; } return r; }
static u64 shoup(u64 w, u64 q) { return (u64)(((u128)w << 64) / q); }
static int bitlen(u64 x) { int b = 0; while (x) { b++; x >>= 1; } return b; }

u64* dev_alloc(Context* c, size_t words) {
    if (c->dry) return nullptr;
    const size_t bytes = words * sizeof(u64);
    if (c->arena_cap) {
        auto it = c->arena_bins.find({c->stream, bytes});
        if (it != c->arena_bins.end() && !it->second.empty()) {
            void* p = it->second.back(); it->second.pop_back();
            c->arena_cached -= bytes; c->arena_hits++;
            return (u64*)p;
        }
    }
    void* p = nullptr;
    HostTimer ht(c, 1);
    cudaError_t e = cudaMallocAsync(&p, bytes, c->stream);
    if (e != cudaSuccess && c->arena_cached) {       // the bins may hold what the pool lacks: hand everything back and retry once
        cudaGetLastError();
        arena_trim(c);
        e = cudaMallocAsync(&p, bytes, c->stream);
    }
    if (e != cudaSuccess) throw std::runtime_error(std::string("cudaMallocAsync: ") + cudaGetErrorString(e));
    if (c->arena_cap) { c->arena_sizes[p] = bytes; c->arena_misses++; }
    return (u64*)p;
}
void dev_free(Context* c, void* p) {
    if (!p || c->dry) return;
    if (c->arena_cap) {
        auto it = c->arena_sizes.find(p);
        if (it != c->arena_sizes.end()) {
            const size_t bytes = it->second;
            if (c->arena_cached + bytes <= c->arena_cap) {
                c->arena_bins[{c->stream, bytes}].push_back(p);   // every use of p is ordered before this point of c->stream
                c->arena_cached += bytes;
                return;
            }
            c->arena_sizes.erase(it);
        }
    }
    HostTimer ht(c, 2);
    cudaFreeAsync(p, c->stream);
}
// hand every cached buffer back to the pool (context teardown, allocation failure, ckks_ctx_set_option(ctx, 2, 0))
void arena_trim(Context* c) {
    for (auto& kv : c->arena_bins)
        for (void* p : kv.second) { c->arena_sizes.erase(p); cudaFreeAsync(p, kv.first.first); }
    c->arena_bins.clear();
    c->arena_cached = 0;
}

template <class T> static T* upload(Context* c, const std::vector<T>& h) {
    if (c->dry) return nullptr;
    T* d = nullptr;
    if (cudaMalloc(&d, h.size() * sizeof(T)) != cudaSuccess) throw std::runtime_error("cudaMalloc failed (tables)");
    cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
    return d;
}

static bool contains(const std::vector<u64>& v, size_t lo, size_t hi, u64 x) { for (size_t i = lo; i < hi; i++) if (v[i] == x) return true; return false; }

void context_generate(Context* c, int log_n, int mult_depth, int scaling_bits, int first_bits, int num_large_digits, int ext,
                      uint64_t seed, int dry, const uint64_t* given_moduli, const uint64_t* given_roots, int given_count) {
    c->log_n = log_n; c->n = 1 << log_n; c->mult_depth = mult_depth; c->scaling_bits = scaling_bits;
    c->first_bits = first_bits; c->ext = ext; c->seed = seed; c->dry = dry != 0;
    if (seed == 0) {   // no seed given: key material and encryption randomness are keyed from OS entropy (INTEGRATION.md "Randomness")
        uint64_t s = 0;
        if (getrandom(&s, sizeof s, 0) != (ssize_t)sizeof s || s == 0) throw std::runtime_error("getrandom failed: cannot seed the sampler");
        c->seed = s;
        if (getrandom(c->rng_key, sizeof c->rng_key, 0) != (ssize_t)sizeof c->rng_key) throw std::runtime_error("getrandom failed: cannot key the sampler");
        c->rng_secure = true;
    }
    { static std::atomic<uint64_t> next_uid{1}; c->uid = next_uid++; }
    if (log_n < 8 || log_n > 17) throw std::runtime_error("ring dimension must be 2^8..2^17");
    const u64 m = 2ull * c->n;
    const int num_primes = mult_depth + 1, size_q = num_primes + (ext ? 1 : 0);
    std::vector<u64> q(size_q, 0);
    q[num_primes - 1] = first_prime(scaling_bits, m);
    if (num_primes > 1) {
        double sf = (double)q[num_primes - 1];
        int cnt = 0;
        for (int i = num_primes - 2; i >= 1; i--) {
            sf = (sf * sf) / (double)q[i + 1];
            u64 sf_int = (u64)std::llround(sf), sf_rem = sf_int % m, cand;
            if (cnt % 2 == 0) { cand = sf_int - m - sf_rem + 1; do { cand = prev_prime(cand, m); } while (contains(q, i + 1, num_primes, cand)); }
            else { cand = sf_int + m - sf_rem + 1; do { cand = next_prime(cand, m); } while (contains(q, i + 1, num_primes, cand)); }
            q[i] = cand; cnt++;
        }
    }
    if (first_bits == scaling_bits) {
        u64 mn = q[num_primes > 1 ? 1 : 0];
        for (int i = 1; i < num_primes; i++) mn = std::min(mn, q[i]);
        q[0] = prev_prime(mn, m);
    } else q[0] = prev_prime(first_prime(first_bits, m), m);
    if (ext) {
        const int ext_bits = 20;
        if (ext_bits == scaling_bits || ext_bits == first_bits) q[num_primes] = prev_prime(q[0], m);
        else q[num_primes] = prev_prime(first_prime(ext_bits, m), m);
    }
    c->size_q = size_q;
    int dnum = num_large_digits;
    if (dnum == 0) dnum = mult_depth > 3 ? 3 : (mult_depth > 0 ? 2 : 1);
    c->alpha = (size_q + dnum - 1) / dnum;
    if (dnum < 1) throw std::runtime_error("numLargeDigits must be non-negative");
    c->num_part_q = (size_q + c->alpha - 1) / c->alpha;
    if (c->num_part_q > MAX_DIGITS)
        throw std::runtime_error("numLargeDigits = " + std::to_string(c->num_part_q) + ": the hybrid key switch supports at most " +
                                 std::to_string(MAX_DIGITS) + " digits");
    int max_bits = 0;
    for (int j = 0; j < c->num_part_q; j++) {
        // bit length of the digit modulus product, via long double log2 with an exact fallback margin
        std::vector<uint32_t> big(1, 1);  // little-endian base 2^32 big integer
        for (int k = j * c->alpha; k < std::min((j + 1) * c->alpha, size_q); k++) {
            u128 carry = 0;
            for (size_t w = 0; w < big.size(); w++) { u128 t = (u128)big[w] * q[k] + carry; big[w] = (uint32_t)t; carry = t >> 32; }
            while (carry) { big.push_back((uint32_t)carry); carry >>= 32; }
        }
        int bits = (int)(big.size() - 1) * 32 + bitlen(big.back());
        max_bits = std::max(max_bits, bits);
    }
    c->size_p = (max_bits + 59) / 60;
    std::vector<u64> p;
    u64 prev = first_prime(60, m);
    for (int i = 0; i < c->size_p; i++) {
        do { prev = prev_prime(prev, m); } while (contains(q, 0, size_q, prev));
        p.push_back(prev);
    }
    c->moduli = q; c->moduli.insert(c->moduli.end(), p.begin(), p.end());
    if (c->M() > MAX_LIMBS) throw std::runtime_error("too many RNS limbs");
    if (given_moduli) {
        // imported context (OpenFHE serialisation, host/serial.h): the file's moduli and roots of unity are used as they
        // are, never re-derived (SURVEY 8c'-2); the shape (limb counts) must be the one the parameters imply
        if (given_count != c->M())
            throw std::runtime_error("imported context: " + std::to_string(given_count) + " moduli given, the parameters imply " +
                                     std::to_string(c->M()) + " (Q then P)");
        for (int i = 0; i < given_count; i++) {
            const u64 qi = given_moduli[i];
            if (qi >> 61 || qi % m != 1 || !is_prime(qi)) throw std::runtime_error("imported context: modulus " + std::to_string(i) + " is not an NTT prime below 2^61");
            if (given_roots && (powmod(given_roots[i], m / 2, qi) != qi - 1)) throw std::runtime_error("imported context: root " + std::to_string(i) + " is not a primitive 2n-th root of unity");
            c->moduli[i] = qi;
        }
        for (int i = 0; i < size_q; i++) q[i] = c->moduli[i];
    }
    // scaling factors
    c->sf_real.assign(size_q, 0.0);
    c->sf_real[0] = (double)q[size_q - 1];
    if (ext && size_q > 1) c->sf_real[1] = (double)q[size_q - 2];
    for (int k = ext ? 2 : 1; k < size_q; k++) c->sf_real[k] = c->sf_real[k - 1] * c->sf_real[k - 1] / (double)q[size_q - k];
    c->sf_big.assign(size_q > 1 ? size_q - 1 : 0, 0.0);
    if (!c->sf_big.empty()) {
        c->sf_big[0] = ext ? c->sf_real[0] * c->sf_real[1] : c->sf_real[0] * c->sf_real[0];
        for (int k = 1; k < size_q - 1; k++) c->sf_big[k] = c->sf_real[k] * c->sf_real[k];
    }
    // per-modulus info + twiddles
    const int M = c->M(), n = c->n;
    c->h_mod.resize(M); c->psi.resize(M);
    std::vector<u64> tw((size_t)2 * M * n), itw((size_t)2 * M * n), ninvp((size_t)2 * M);
    for (int i = 0; i < M; i++) {
        u64 qi = c->moduli[i]; ModInfo& mi = c->h_mod[i];
        mi.q = qi; mi.bits = bitlen(qi); mi.pad = 0;
        mi.qrecip = (u64)((((u128)1) << 64) / qi);
        mi.mu = (u64)((((u128)1) << (2 * mi.bits)) / qi);
        u64 inv = 1; for (int k = 0; k < 6; k++) inv *= 2 - qi * inv;   // Newton: q^-1 mod 2^64
        mi.qinv = inv;
        u64 r = (u64)((((u128)1) << 64) % qi); mi.r2 = mulmod(r, r, qi);
        mi.ninv = invmod(n % qi, qi); mi.ninv_sh = shoup(mi.ninv, qi);
        ninvp[2 * i] = mi.ninv; ninvp[2 * i + 1] = mi.ninv_sh;
        u64 psi = (given_moduli && given_roots) ? given_roots[i] : min_root(m, qi), ipsi = invmod(psi, qi); c->psi[i] = psi;
        u64 x = 1, y = 1;
        for (int k = 0; k < n; k++) {
            uint32_t b = brev(k, log_n);
            const size_t o = 2 * ((size_t)i * n + b);
            tw[o] = x; tw[o + 1] = shoup(x, qi);
            itw[o] = y; itw[o + 1] = shoup(y, qi);
            x = mulmod(x, psi, qi); y = mulmod(y, ipsi, qi);
        }
    }
    if (getenv("CKKS_B200_NO_PDL")) c->pdl = false;
    if (getenv("CKKS_B200_HOST_PROF")) c->host_prof = true;
    if (getenv("CKKS_B200_NO_ENC_OFFLOAD")) c->enc_offload = false;
    if (!c->dry) {
        if (!c->stream) { cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking); c->own_stream = true; }
        if (!getenv("CKKS_B200_NO_ENCODE_STREAM")) {
            int lo = 0, hi = 0; cudaDeviceGetStreamPriorityRange(&lo, &hi);
            cudaStreamCreateWithPriority(&c->enc_stream, cudaStreamNonBlocking, hi);
        }
        cudaMemPool_t pool; int dev = 0; cudaGetDevice(&dev);
        if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) { uint64_t thr = UINT64_MAX; cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr); }
        {   // arena capacity: CKKS_B200_ARENA_MB (0 = off), default an eighth of the free device memory, at most 16 GiB
            size_t fr = 0, tot = 0; cudaMemGetInfo(&fr, &tot);
            const char* e = getenv("CKKS_B200_ARENA_MB");
            c->arena_cap = e ? (size_t)atoll(e) << 20 : std::min<size_t>(fr / 8, (size_t)16 << 30);
        }
    }
    c->d_mod = upload(c, c->h_mod);
    c->d_tw = upload(c, tw); c->d_itw = upload(c, itw); c->d_ninv_pairs = upload(c, ninvp);
    {   // double copies of the twiddles for the FP64 butterfly path (ntt.cu); rows of wider moduli stay zero
        std::vector<double> twd((size_t)M * n, 0.0), itwd((size_t)M * n, 0.0);
        bool any = false;
        for (int i = 0; i < M; i++) {
            if (c->h_mod[i].bits > 44) continue;
            any = true;
            for (int k = 0; k < n; k++) { const size_t o = (size_t)i * n + k; twd[o] = (double)tw[2 * o]; itwd[o] = (double)itw[2 * o]; }
        }
        if (any && !getenv("CKKS_B200_NO_FP64_NTT")) {
            c->d_twd = upload(c, twd); c->d_itwd = upload(c, itwd);
            if (n >= 64 && !getenv("CKKS_B200_NO_OT_TWIDDLES")) {   // every 4th twiddle of the last stage, packed
                std::vector<double> b((size_t)M * (n / 8)), ib((size_t)M * (n / 8));
                for (int i = 0; i < M; i++)
                    for (int e = 0; e < n / 8; e++) {
                        b[(size_t)i * (n / 8) + e] = twd[(size_t)i * n + n / 2 + 4 * e];
                        ib[(size_t)i * (n / 8) + e] = itwd[(size_t)i * n + n / 2 + 4 * e];
                    }
                c->d_twb = upload(c, b); c->d_itwb = upload(c, ib);
            }
        }
    }
    // special FFT tables (same expressions as oracle Context.fft_tables)
    const int Mc = 2 * n;
    std::vector<double> ksi(2 * (size_t)(Mc + 1));
    for (int k = 0; k <= Mc; k++) { double ang = 2.0 * M_PI * (double)k / (double)Mc; ksi[2 * k] = std::cos(ang); ksi[2 * k + 1] = std::sin(ang); }
    std::vector<uint32_t> rot(n / 2); u64 f = 1;
    for (int j = 0; j < n / 2; j++) { rot[j] = (uint32_t)f; f = f * 5 % Mc; }
    c->d_ksi = upload(c, ksi); c->d_rot = upload(c, rot);
}

static void fill_conv(Context* c, ConvTables& t, const std::vector<int>& in_ids, const std::vector<int>& out_ids) {
    t.nin = (int)in_ids.size(); t.nout = (int)out_ids.size();
    std::vector<u64> hatinv(t.nin), hatinv_sh(t.nin), hatmod((size_t)t.nin * t.nout), hatplain((size_t)t.nin * t.nout);
    for (int i = 0; i < t.nin; i++) {
        t.in_ids[i] = (short)in_ids[i];
        u64 qi = c->moduli[in_ids[i]], h = 1;
        for (int k = 0; k < t.nin; k++) if (k != i) h = mulmod(h, c->moduli[in_ids[k]] % qi, qi);
        hatinv[i] = invmod(h, qi); hatinv_sh[i] = shoup(hatinv[i], qi);
        for (int j = 0; j < t.nout; j++) {
            u64 tj = c->moduli[out_ids[j]], hm = 1;
            for (int k = 0; k < t.nin; k++) if (k != i) hm = mulmod(hm, c->moduli[in_ids[k]] % tj, tj);
            hatmod[(size_t)i * t.nout + j] = (u64)(((u128)hm << 64) % tj);   // Montgomery form
            hatplain[(size_t)i * t.nout + j] = hm;
        }
    }
    {   // k_conv_split: digit (< 2^20) x half (< 2^32, < 2^28) products, at most 12 per 64-bit column; all moduli < 2^60
        bool ok = t.nin <= 12;
        for (int id : in_ids) ok = ok && (c->moduli[id] >> 60) == 0;
        for (int id : out_ids) ok = ok && (c->moduli[id] >> 60) == 0;
        t.split_ok = ok;
    }
    t.wide = 0;
    for (int i = 0; i < t.nin && i < 32; i++) if (c->moduli[in_ids[i]] >> 42) t.wide |= 1u << i;
    for (int j = 0; j < t.nout; j++) t.out_ids[j] = (short)out_ids[j];
    t.d_hatinv = upload(c, hatinv); t.d_hatinv_sh = upload(c, hatinv_sh); t.d_hatmod = upload(c, hatmod); t.d_hatplain = upload(c, hatplain);
}

const KsTables& get_ks_tables(Context* c, int nl) {
    auto it = c->ks.find(nl);
    if (it != c->ks.end()) return it->second;
    KsTables t; t.nl = nl;
    t.nparts = std::min((nl + c->alpha - 1) / c->alpha, c->num_part_q);
    t.modup.resize(t.nparts);
    for (int j = 0; j < t.nparts; j++) {
        int lo = c->alpha * j, hi = std::min(c->alpha * (j + 1), nl);
        t.dig_lo.push_back(lo); t.dig_hi.push_back(hi);
        std::vector<int> in, out;
        for (int i = lo; i < hi; i++) in.push_back(i);
        for (int i = 0; i < nl; i++) if (i < lo || i >= hi) out.push_back(i);
        for (int k = 0; k < c->size_p; k++) out.push_back(c->size_q + k);
        fill_conv(c, t.modup[j], in, out);
    }
    std::vector<int> pin, qout;
    for (int k = 0; k < c->size_p; k++) pin.push_back(c->size_q + k);
    for (int i = 0; i < nl; i++) qout.push_back(i);
    fill_conv(c, t.moddown, pin, qout);
    std::vector<u64> pinv(nl), pinv_sh(nl);
    for (int i = 0; i < nl; i++) {
        u64 qi = c->moduli[i], pm = 1;
        for (int k = 0; k < c->size_p; k++) pm = mulmod(pm, c->moduli[c->size_q + k] % qi, qi);
        pinv[i] = invmod(pm, qi); pinv_sh[i] = shoup(pinv[i], qi);
    }
    t.d_pinv = upload(c, pinv); t.d_pinv_sh = upload(c, pinv_sh);
    {
        std::vector<u64> pairs(2 * (size_t)nl);
        for (int i = 0; i < nl; i++) { pairs[2 * i] = pinv[i]; pairs[2 * i + 1] = pinv_sh[i]; }
        t.d_pinv_pairs = upload(c, pairs);
    }
    // fused final factors of the inverse NTTs that feed the basis conversions
    std::vector<u64> ms(2 * (size_t)c->M(), 0), ds(2 * (size_t)c->M(), 0);
    for (int j = 0; j < t.nparts; j++) {
        for (int i = t.dig_lo[j]; i < t.dig_hi[j]; i++) {
            u64 qi = c->moduli[i], h = 1;
            for (int k = t.dig_lo[j]; k < t.dig_hi[j]; k++) if (k != i) h = mulmod(h, c->moduli[k] % qi, qi);
            u64 f = mulmod(c->h_mod[i].ninv, invmod(h, qi), qi);
            ms[2 * i] = f; ms[2 * i + 1] = shoup(f, qi);
        }
    }
    for (int k = 0; k < c->size_p; k++) {
        int id = c->size_q + k; u64 pk = c->moduli[id], h = 1;
        for (int k2 = 0; k2 < c->size_p; k2++) if (k2 != k) h = mulmod(h, c->moduli[c->size_q + k2] % pk, pk);
        u64 f = mulmod(c->h_mod[id].ninv, invmod(h, pk), pk);
        ds[2 * id] = f; ds[2 * id + 1] = shoup(f, pk);
    }
    t.d_modup_scale = upload(c, ms); t.d_moddown_scale = upload(c, ds);
    return c->ks.emplace(nl, std::move(t)).first->second;
}

// ---- per-class kernel timing ----
static cudaEvent_t prof_event(Context* c) {
    if (!c->prof.pool.empty()) { cudaEvent_t e = c->prof.pool.back(); c->prof.pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
}
ProfScope::ProfScope(Context* c_, int cls, double alg_bytes, int nlaunch) : c(c_) {
    if (c->dry || !(c->prof.mask & (1u << cls))) return;
    c->prof.bytes[cls] += alg_bytes; c->prof.launches[cls] += nlaunch;
    ProfRec r{cls, prof_event(c), prof_event(c)};
    cudaEventRecord(r.e0, c->stream);
    idx = (int)c->prof.pending.size();
    c->prof.pending.push_back(r);
}
ProfScope::~ProfScope() { if (idx >= 0) cudaEventRecord(c->prof.pending[idx].e1, c->stream); }
void prof_collect(Context* c) {
    if (c->dry) return;
    if (c->enc_stream) cudaStreamSynchronize(c->enc_stream);
    cudaStreamSynchronize(c->stream);
    for (auto& r : c->prof.pending) {
        float ms = 0; cudaEventElapsedTime(&ms, r.e0, r.e1);
        c->prof.ms[r.cls] += ms;
        c->prof.pool.push_back(r.e0); c->prof.pool.push_back(r.e1);
    }
    c->prof.pending.clear();
}
