// Device context: immutable device copies of everything SEALContext::ContextData owns for the hot
// path (per-modulus constants, NTT tables), uploaded switching keys, Galois permutation tables and
// lazily built per-level key-switch / rescale constants.
#pragma once
#include "../../include/gseal_b200.h"
#include "host_tables.h"
#include "ntt_kernels.cuh"
#include "gsb_bfv_kernels.cuh"
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace gsb
{
    struct ApiError
    {
        int code;
        std::string msg;
    };

    void set_error(const std::string &msg);
    unsigned long long &launch_counter();

#define GSB_CUDA_CHECK(expr)                                                                       \
    do                                                                                             \
    {                                                                                              \
        cudaError_t e__ = (expr);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            throw ::gsb::ApiError{ GSB_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__) }; \
    } while (0)

    inline void require(bool ok, int code, const char *msg)
    {
        if (!ok)
            throw ApiError{ code, msg };
    }

    // Constants of the hybrid key switch at one ciphertext level L (multi_special_primes.cpp).
    struct KsLevelConsts
    {
        // mod-up, bundle j with |S_j| > 1 (multi_special_primes.cpp:113-126):
        //   a[j][i]      = (prod_{i' in S_j \ i} q_i')^-1 mod q_i   (+ Shoup)
        //   b[j][i][kap] = prod_{i' in S_j \ i} q_i' mod q_kap
        uint64_t *d_modup_a = nullptr;  // [beta][k][2]
        uint64_t *d_modup_b = nullptr;  // [beta][k][L+k]
        // mod-down (multi_special_primes.cpp:186-234, 292-299):
        uint64_t *d_inv_hat = nullptr;  // [k][2]  (prod_{j'!=j} p_j')^-1 mod p_j (+ Shoup)
        uint64_t *d_neg_hat = nullptr;  // [L][k]  -(prod_{j'!=j} p_j') mod q_i
        uint64_t *d_inv_p = nullptr;    // [L][2]  (prod_j p_j)^-1 mod q_i (+ Shoup)
        uint64_t *d_neg_p = nullptr;    // [L][2]  -(prod_j p_j) mod q_i (+ Shoup): fused mul+relin+rescale tail
        // rescale (util/rns.cpp:719-728, 824): q_last^-1 mod q_i ; q_i - (half mod q_i)
        uint64_t *d_inv_qlast = nullptr; // [L-1][2] (+ Shoup)
        uint64_t *d_neg_half = nullptr;  // [L-1]
    };

    // Device side of the RNSTool of one (level L, plain modulus t) pair: NTT tables over the extended
    // base q u Bsk and the BEHZ conversion constants (util/rns.cpp:540-728).
    struct BfvTool
    {
        size_t L = 0, nB = 0, nBsk = 0;
        std::vector<uint64_t> bsk;
        Twiddle *d_WW = nullptr, *d_VV = nullptr;
        LimbC *d_limb = nullptr;
        uint64_t *d_consts = nullptr; // one allocation, carved up by `dev`
        BehzDev dev{};
        DecryptRoundArgs dec{}; // constants of decrypt_scale_and_round (dot / out / total filled per call)
    };

    struct SwitchKey
    {
        uint64_t *d = nullptr; // [beta0][2][K][N]
        size_t bundles = 0;
    };
} // namespace gsb

struct gsb_ws
{
    gsb_ctx *ctx = nullptr;
    void *buf = nullptr;
    size_t cap = 0;
    void *pinned = nullptr;
    size_t pinned_cap = 0;
    void *reserve(size_t bytes);
    void *reserve_pinned(size_t bytes);
};

struct gsb_ctx
{
    int device = 0;
    int logn = 0;
    size_t N = 0, K = 0, k = 0, L0 = 0;
    // no-wrap bounds over q_max = max modulus, n = log2 N, beta0 = ceil(L0 / k) (see KsDev, fwd_round):
    //   fast_ntt:   q_max (5n + 2) < 2^63             (a wide-range forward NTT grows by < 5q per stage
    //                                                  from an input < 2q; its multiply needs x < 2^63)
    //   lazy_modup: = fast_ntt                        (single-limb bundle enters the NTT unreduced)
    //   lazy_mac:   beta0 (5n + 2) q_max^2 < 2^128    (sum of beta0 products operand x key)
    bool fast_ntt = false;
    //   fast_fwd_public: q_max (5n + 4) < 2^63        (the public strict forward NTT may be handed lazy words < 4q)
    bool fast_fwd_public = false;
    //   wide_inv:   10 q_max < 2^63                   (wide-range inverse butterflies keep words in [0,5q);
    //                                                  no growth, so it also holds for 59-bit moduli)
    bool wide_inv = false;
    bool lazy_modup = false, lazy_mac = false;
    // testing aid (GSB_SPLIT_TAIL=1 at context creation): gsb_mul_relin_rescale_* runs mod-down and rescale as
    // the separate passes of the stand-alone entry points instead of the fused tail (identical outputs)
    bool split_tail = false;
    std::vector<gsb::host::LimbTables> tables; // host copies (introspection + constants)
    gsb::Twiddle *d_WW = nullptr, *d_VV = nullptr;
    gsb::LimbC *d_limb = nullptr;
    std::map<size_t, gsb::KsLevelConsts> level_consts;
    std::map<uint32_t, gsb::SwitchKey> relin; // RelinKeys::get_index(key_power) -> key
    std::map<uint32_t, gsb::SwitchKey> galois;
    std::map<uint32_t, uint32_t *> galois_tables; // device permutation tables per element
    std::map<std::pair<size_t, uint64_t>, gsb::BfvTool> bfv_tools; // (L, plain modulus) -> BEHZ tool
    // Decryptor::secret_key_array_ (decryptor.cpp:46-57, 151-216): NTT-form powers s^1.. at key level,
    // [sk_powers][K][N] on the device; extended on demand by dyadic products, never shrunk
    uint64_t *d_sk = nullptr;
    size_t sk_powers = 0;
    std::vector<std::pair<uint64_t *, size_t>> sk_retired; // superseded power arrays + bytes (kernels in flight may still read them)
    std::mutex mu;      // guards the lazily built maps below
    std::mutex host_mu; // serialises the *_host entry points (they share host_io / default_ws)
    gsb_ws default_ws;  // compute scratch for callers that pass ws = NULL
    gsb_ws host_io;     // device staging of the *_host entry points
    bool pipe_ready = false; // copy-in / run / copy-out streams and events of the *_host pipeline
    cudaStream_t pipe_stream[3] = {};
    cudaEvent_t pipe_event[6] = {};

    gsb::NttTables ntt_tables() const
    {
        return gsb::NttTables{ d_WW, d_VV, d_limb, static_cast<uint32_t>(N), logn };
    }
    const gsb::KsLevelConsts &consts_for(size_t L);
    const uint32_t *galois_table(uint32_t elt);
    const gsb::BfvTool &bfv_tool(size_t L, uint64_t plain_modulus);
};
