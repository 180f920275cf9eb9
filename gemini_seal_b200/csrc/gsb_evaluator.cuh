// CKKS evaluator hot path on the device: tensor product, hybrid key switching (k >= 1 special
// primes), mod-down, rescale and Galois rotation, built from the NTT rounds of ntt_core.cuh with
// fused prologues / epilogues.  Reference: native/src/seal/evaluator.cpp:447-527 (ckks_multiply),
// :772-827 (relinearize_internal), :829-892 (mod_switch_scale_to_next), :1841-1943
// (apply_galois_inplace), :2259-2368 (switch_key_inplace); multi_special_primes.cpp:80-304;
// util/rns.cpp:777-851; util/galois.cpp:18-47,188-214.  Included at the end of gsb_api.cu.
#pragma once
#include "gsb_context.cuh"
#include "gsb_ks_kernels.cuh"
#include <algorithm>

namespace gsb
{
    // ---- per-level constants (host, once per level; multi_special_primes.cpp / util/rns.cpp) ------
    static uint64_t reduce_mod(uint64_t x, const host::ModConst &m)
    {
        return x % m.q;
    }

    static uint64_t *upload_words(const std::vector<uint64_t> &h)
    {
        uint64_t *d = nullptr;
        const size_t n = h.empty() ? 1 : h.size();
        GSB_CUDA_CHECK(cudaMalloc(reinterpret_cast<void **>(&d), n * sizeof(uint64_t)));
        if (!h.empty())
            GSB_CUDA_CHECK(cudaMemcpy(d, h.data(), h.size() * sizeof(uint64_t), cudaMemcpyHostToDevice));
        return d;
    }
} // namespace gsb

const gsb::KsLevelConsts &gsb_ctx::consts_for(size_t L)
{
    using namespace gsb;
    std::lock_guard<std::mutex> lock(mu);
    auto it = level_consts.find(L);
    if (it != level_consts.end())
        return it->second;
    require(L >= 1 && L <= L0, GSB_E_INVALID_ARG, "no ciphertext level with that many limbs");
    const size_t E = L + k, beta = (L + k - 1) / k;
    auto modc = [&](size_t idx) -> const host::ModConst & { return tables[idx].mod; };
    auto ext_mod = [&](size_t kap) { return kap < L ? kap : L0 + (kap - L); };
    std::vector<uint64_t> a(beta * k * 2, 0), b(beta * k * E, 0);
    // mod-up constants (multi_special_primes.cpp:113-126)
    for (size_t j = 0; j < beta; j++)
    {
        const size_t s0 = j * k, s1 = std::min(s0 + k, L);
        if (s1 - s0 < 2)
            continue; // single-limb bundle: copy / barrett_reduce_63 path (:99-108)
        for (size_t i = s0; i < s1; i++)
        {
            uint64_t inv_prod = 1;
            for (size_t i2 = s0; i2 < s1; i2++)
                if (i2 != i)
                    inv_prod = host::mul_mod(inv_prod, reduce_mod(modc(i2).q, modc(i)), modc(i));
            uint64_t inv = 0;
            require(host::inv_mod(inv_prod, modc(i).q, inv), GSB_E_LOGIC, "modup: inv_mod fail");
            a[(j * k + (i - s0)) * 2] = inv;
            a[(j * k + (i - s0)) * 2 + 1] = host::shoup(inv, modc(i).q);
            for (size_t kap = 0; kap < E; kap++)
            {
                const host::ModConst &md = modc(ext_mod(kap));
                uint64_t prod = 1;
                for (size_t i2 = s0; i2 < s1; i2++)
                    if (i2 != i)
                        prod = host::mul_mod(prod, reduce_mod(modc(i2).q, md), md);
                b[(j * k + (i - s0)) * E + kap] = prod;
            }
        }
    }
    // mod-down constants (multi_special_primes.cpp:186-234, 292-299)
    std::vector<uint64_t> inv_hat(k * 2, 0), neg_hat(L * k, 0), inv_p(L * 2, 0), neg_p(L * 2, 0);
    for (size_t j = 0; j < k; j++)
    {
        const host::ModConst &pj = modc(L0 + j);
        uint64_t prod = 1;
        for (size_t j2 = 0; j2 < k; j2++)
            if (j2 != j)
                prod = host::mul_mod(prod, reduce_mod(modc(L0 + j2).q, pj), pj);
        uint64_t inv = 0;
        require(host::inv_mod(prod, pj.q, inv), GSB_E_LOGIC, "moddown: inv_mod fail");
        inv_hat[2 * j] = inv;
        inv_hat[2 * j + 1] = host::shoup(inv, pj.q);
    }
    for (size_t i = 0; i < L; i++)
    {
        const host::ModConst &qi = modc(i);
        uint64_t P = 1;
        for (size_t j = 0; j < k; j++)
        {
            uint64_t prod = 1;
            for (size_t j2 = 0; j2 < k; j2++)
                if (j2 != j)
                    prod = host::mul_mod(prod, reduce_mod(modc(L0 + j2).q, qi), qi);
            neg_hat[i * k + j] = prod ? qi.q - prod : 0;
            P = host::mul_mod(P, reduce_mod(modc(L0 + j).q, qi), qi);
        }
        uint64_t inv = 0;
        require(host::inv_mod(P, qi.q, inv), GSB_E_LOGIC, "moddown: inv_mod fail");
        inv_p[2 * i] = inv;
        inv_p[2 * i + 1] = host::shoup(inv, qi.q);
        neg_p[2 * i] = P ? qi.q - P : 0;
        neg_p[2 * i + 1] = host::shoup(neg_p[2 * i], qi.q);
    }
    // rescale constants (util/rns.cpp:719-728, 824)
    std::vector<uint64_t> inv_qlast, neg_half;
    if (L >= 2)
    {
        const uint64_t qL = modc(L - 1).q, half = qL >> 1;
        for (size_t i = 0; i + 1 < L; i++)
        {
            uint64_t inv = 0;
            require(host::inv_mod(qL, modc(i).q, inv), GSB_E_LOGIC, "invalid rns bases");
            inv_qlast.push_back(inv);
            inv_qlast.push_back(host::shoup(inv, modc(i).q));
            neg_half.push_back(modc(i).q - host::barrett63(half, modc(i)));
        }
    }
    GSB_CUDA_CHECK(cudaSetDevice(device));
    KsLevelConsts c;
    c.d_modup_a = upload_words(a);
    c.d_modup_b = upload_words(b);
    c.d_inv_hat = upload_words(inv_hat);
    c.d_neg_hat = upload_words(neg_hat);
    c.d_inv_p = upload_words(inv_p);
    c.d_neg_p = upload_words(neg_p);
    c.d_inv_qlast = upload_words(inv_qlast);
    c.d_neg_half = upload_words(neg_half);
    return level_consts.emplace(L, c).first->second;
}

const gsb::BfvTool &gsb_ctx::bfv_tool(size_t L, uint64_t plain_modulus)
{
    using namespace gsb;
    std::lock_guard<std::mutex> lock(mu);
    const auto key = std::make_pair(L, plain_modulus);
    auto it = bfv_tools.find(key);
    if (it != bfv_tools.end())
        return it->second;
    require(L >= 1 && L <= L0, GSB_E_INVALID_ARG, "no ciphertext level with that many limbs");
    // context.cpp:178-186: plain_modulus must be smaller than every coeff modulus prime, at least 2
    require(plain_modulus >= 2 && (plain_modulus >> 59) == 0, GSB_E_INVALID_ARG, "plain_modulus out of range");
    std::vector<uint64_t> q(L);
    for (size_t i = 0; i < L; i++)
        q[i] = tables[i].mod.q;
    host::BehzConsts bc;
    require(host::build_behz(logn, q.data(), L, plain_modulus, bc), GSB_E_LOGIC, "invalid rns bases");
    const size_t M = L + bc.nBsk;
    require(M <= BEHZ_MAX, GSB_E_LOGIC, "invalid rns bases");
    std::vector<Twiddle> ww(M * N), vv(M * N);
    std::vector<LimbC> limb(M);
    for (size_t i = 0; i < M; i++)
    {
        host::LimbTables aux;
        if (i >= L)
            require(host::build_limb_tables(logn, bc.bsk[i - L], aux), GSB_E_LOGIC, "invalid rns bases");
        const host::LimbTables &t = i < L ? tables[i] : aux;
        for (size_t d = 0; d < N; d++)
        {
            const uint32_t n32 = static_cast<uint32_t>(N), d32 = static_cast<uint32_t>(d);
            ww[i * N + fwd_tw_pos(d32, n32)] = make_ulonglong2(t.w[d], t.ws[d]);
            vv[i * N + inv_tw_pos(d32, n32)] = make_ulonglong2(t.v[d], t.vs[d]);
        }
        LimbC c;
        c.q = t.mod.q;
        c.two_q = t.mod.q << 1;
        c.rdp = t.rdp;
        c.cr0 = t.mod.cr0;
        c.cr1 = t.mod.cr1;
        c.inv_n = t.inv_n;
        c.inv_n_s = t.inv_n_s;
        c.pad = 0;
        limb_mu(c.q, c.mu, c.sh);
        limb[i] = c;
    }
    BfvTool tool;
    tool.L = L;
    tool.nB = bc.nB;
    tool.nBsk = bc.nBsk;
    tool.bsk = bc.bsk;
    GSB_CUDA_CHECK(cudaMalloc(reinterpret_cast<void **>(&tool.d_WW), ww.size() * sizeof(Twiddle)));
    GSB_CUDA_CHECK(cudaMalloc(reinterpret_cast<void **>(&tool.d_VV), vv.size() * sizeof(Twiddle)));
    GSB_CUDA_CHECK(cudaMalloc(reinterpret_cast<void **>(&tool.d_limb), limb.size() * sizeof(LimbC)));
    GSB_CUDA_CHECK(cudaMemcpy(tool.d_WW, ww.data(), ww.size() * sizeof(Twiddle), cudaMemcpyHostToDevice));
    GSB_CUDA_CHECK(cudaMemcpy(tool.d_VV, vv.data(), vv.size() * sizeof(Twiddle), cudaMemcpyHostToDevice));
    GSB_CUDA_CHECK(cudaMemcpy(tool.d_limb, limb.data(), limb.size() * sizeof(LimbC), cudaMemcpyHostToDevice));
    std::vector<uint64_t> all;
    auto put = [&](const std::vector<uint64_t> &v) {
        const size_t off = all.size();
        all.insert(all.end(), v.begin(), v.end());
        return off;
    };
    const size_t o1 = put(bc.inv_punct_q), o2 = put(bc.mat_q), o3 = put(bc.inv_punct_b), o4 = put(bc.mat_b),
                 o5 = put(bc.prod_q_mod_bsk), o6 = put(bc.inv_prod_q_mod_bsk), o7 = put(bc.inv_mtilde_mod_bsk), o8 = put(bc.prod_b_mod_q),
                 o9 = put(bc.prod_tg_mod_q), o10 = put(bc.mat_tg);
    tool.d_consts = upload_words(all);
    {
        auto as_limb = [](uint64_t m) {
            const host::ModConst mc = host::make_mod(m);
            LimbC c{};
            c.q = mc.q;
            c.two_q = mc.q << 1;
            c.cr0 = mc.cr0;
            c.cr1 = mc.cr1;
            limb_mu(c.q, c.mu, c.sh);
            return c;
        };
        tool.dec = DecryptRoundArgs{ nullptr, nullptr, tool.d_limb, tool.d_consts + o9, tool.d_consts + o1, tool.d_consts + o10,
                                     as_limb(plain_modulus), as_limb(bc.gamma), bc.neg_inv_q_mod_tg[0], bc.neg_inv_q_mod_tg[1],
                                     bc.inv_gamma_mod_t, static_cast<uint32_t>(N), static_cast<uint32_t>(L), 0 };
    }
    tool.dev = BehzDev{ tool.d_limb, tool.d_consts + o1, tool.d_consts + o2, tool.d_consts + o3, tool.d_consts + o4,
                        tool.d_consts + o5, tool.d_consts + o6, tool.d_consts + o7, tool.d_consts + o8,
                        bc.inv_prod_q_mod_mtilde, bc.inv_prod_b_mod_msk, plain_modulus,
                        static_cast<uint32_t>(L), static_cast<uint32_t>(bc.nB), static_cast<uint32_t>(bc.nBsk), static_cast<uint32_t>(N) };
    return bfv_tools.emplace(key, std::move(tool)).first->second;
}

const uint32_t *gsb_ctx::galois_table(uint32_t elt)
{
    using namespace gsb;
    std::lock_guard<std::mutex> lock(mu);
    auto it = galois_tables.find(elt);
    if (it != galois_tables.end())
        return it->second;
    // GaloisTool::generate_table_ntt argument checks (util/galois.cpp:20-24, galois.h:139-148)
    require((elt & 1u) && elt < 2u * N, GSB_E_INVALID_ARG, "Galois element is not valid");
    std::vector<uint32_t> t;
    host::galois_table_ntt(logn, elt, t);
    uint32_t *d = nullptr;
    GSB_CUDA_CHECK(cudaSetDevice(device));
    GSB_CUDA_CHECK(cudaMalloc(reinterpret_cast<void **>(&d), N * sizeof(uint32_t)));
    GSB_CUDA_CHECK(cudaMemcpy(d, t.data(), N * sizeof(uint32_t), cudaMemcpyHostToDevice));
    galois_tables[elt] = d;
    return d;
}

namespace gsb
{
    // ---- workspace carving ---------------------------------------------------------------------
    struct Arena
    {
        char *base;
        size_t off = 0;
        explicit Arena(void *p) : base(static_cast<char *>(p)) {}
        uint64_t *words(size_t n)
        {
            uint64_t *r = reinterpret_cast<uint64_t *>(base + off);
            off += (n * sizeof(uint64_t) + 255) / 256 * 256;
            return r;
        }
    };
    static size_t arena_words(size_t n)
    {
        return ((n * sizeof(uint64_t) + 255) / 256 * 256) / sizeof(uint64_t);
    }

    static KsDev make_ksdev(gsb_ctx *ctx, size_t L)
    {
        // SEALContext::using_keyswitching() is false when every prime is special (context.cpp:461-464):
        // the reference then refuses to key-switch (evaluator.cpp:2271-2274)
        require(ctx->K > ctx->k, GSB_E_LOGIC, "keyswitching is not supported by the context");
        const KsLevelConsts &lc = ctx->consts_for(L);
        KsDev ks;
        ks.limb = ctx->d_limb;
        ks.modup_a = lc.d_modup_a;
        ks.modup_b = lc.d_modup_b;
        ks.inv_hat = lc.d_inv_hat;
        ks.neg_hat = lc.d_neg_hat;
        ks.inv_p = lc.d_inv_p;
        ks.L = static_cast<uint32_t>(L);
        ks.L0 = static_cast<uint32_t>(ctx->L0);
        ks.k = static_cast<uint32_t>(ctx->k);
        ks.E = static_cast<uint32_t>(L + ctx->k);
        ks.beta = static_cast<uint32_t>((L + ctx->k - 1) / ctx->k);
        ks.beta_magic = ks.beta > 1 ? static_cast<uint32_t>(((uint64_t(1) << 32) + ks.beta - 1) / ks.beta) : 0u;
        ks.N = static_cast<uint32_t>(ctx->N);
        ks.lazy_modup = ctx->lazy_modup && ctx->k == 1;
        ks.lazy_mac = ctx->lazy_mac && ctx->k == 1;
        return ks;
    }

    // Words of scratch one ciphertext needs inside switch_key_core.
    static size_t ks_words_per_ct(const gsb_ctx *ctx, size_t L)
    {
        const size_t N = ctx->N, E = L + ctx->k, beta = (L + ctx->k - 1) / ctx->k;
        const bool two = ctx->logn >= GSB_TWO_PASS_MIN;
        return L * N + (two ? beta * E * N : 0) + 2 * E * N + (two ? 2 * L * N : 0);
    }

    struct KsOutput // where the key-switch result is added and stored (see MdStoreArgs)
    {
        const uint64_t *add_src;
        size_t add_bs, add_ls;
        uint64_t *dst;
        size_t dst_bs, dst_ls;
        const uint32_t *table;
        bool zero_l1 = false; // component 1 gets no addend (set with a Galois table, or for BFV Galois)
        bool bfv = false;     // BFV branch: coefficient-form target and result (is_ckks == false)
    };

    // switch_key_inplace for `B` ciphertexts (evaluator.cpp:2259-2368).  scratch must hold
    // B * ks_words_per_ct words (+ alignment slack).  front_only: stop after step 3 + the Barrett part of
    // step 4 (red = the reduced inner product of every ext slot, NTT form) and return `red`; the fused
    // mul+relin+rescale tail continues from there.
    template <int LOGN, bool K1>
    static uint64_t *switch_key_core_t(
        gsb_ctx *ctx, void *scratch, const uint64_t *target, size_t target_stride, const SwitchKey &key, const KsOutput &o,
        size_t B, size_t L, cudaStream_t st, bool front_only = false)
    {
        constexpr bool two = Split<LOGN>::two_pass;
        const size_t N = ctx->N;
        const KsDev ks = make_ksdev(ctx, L);
        const size_t E = ks.E, beta = ks.beta, k = ks.k;
        require(key.d != nullptr && key.bundles >= beta, GSB_E_INVALID_ARG, "kswitch_keys is not valid for encryption parameters");
        require(B * beta * beta < (uint64_t(1) << 32), GSB_E_RANGE, "key switching: batch chunk too large for KsDev::beta_magic");
        const NttTables tab = ctx->ntt_tables();
        Arena ar(scratch);
        uint64_t *coef = ar.words(B * L * N);
        uint64_t *ext1 = two ? ar.words(B * beta * E * N) : nullptr;
        uint64_t *red = ar.words(B * 2 * E * N);
        uint64_t *tmp = two ? ar.words(B * 2 * L * N) : nullptr;
        const int iB = static_cast<int>(B), iL = static_cast<int>(L), iE = static_cast<int>(E);

        // Step 1 (:2302-2307): strict inverse NTT of every target limb (all bundles at once); a BFV
        // target is in coefficient form already and is only gathered into the batch-dense layout.
        if (o.bfv)
            GSB_CUDA_CHECK(cudaMemcpy2DAsync(
                coef, L * N * sizeof(uint64_t), target, target_stride * sizeof(uint64_t), L * N * sizeof(uint64_t), B,
                cudaMemcpyDeviceToDevice, st));
        else
            ntt_inverse_plain<LOGN>(tab, ModMap{ 1 << 30, 0, 0 }, target, target_stride, coef, L * N, iL, iB, false, st, "ks_target", ctx->wide_inv);

        // Step 2 + first NTT pass of step 3 (:2310, :2322): mod-up fused into the column pass.
        if constexpr (two)
        {
            const ModMap mup{ iL, static_cast<int>(ctx->L0), 0 };
            if (K1 && ks.lazy_modup)
            {
                IoPair<ModupLoadT<K1, K1>, PlainStore<OUT_RAW>> io{ ModupArgs{ ks, coef }, PlainStoreArgs{ ext1, E * N } };
                launch_fwd_cols<LOGN, true>(io, tab, mup, iE, static_cast<int>(B * beta), st, "modup");
            }
            else
            {
                IoPair<ModupLoadT<K1, false>, PlainStore<OUT_RAW>> io{ ModupArgs{ ks, coef }, PlainStoreArgs{ ext1, E * N } };
                launch_fwd_cols<LOGN>(io, tab, mup, iE, static_cast<int>(B * beta), st, "modup");
            }
        }
        // Step 3 + Barrett part of step 4: last NTT pass fused with the key inner product.
        {
            typedef MacPlan<LOGN> MP;
            KsMacArgs a{ ks, target, target_stride, coef, ext1, key.d, red, static_cast<uint32_t>(ctx->K) };
            dim3 grid(static_cast<unsigned>(B), 1u << (LOGN - MP::T - 3), static_cast<unsigned>(E));
            auto launch = [&](auto kfn) {
                allow_smem(kfn, MP::smem_bytes);
                GSB_LAUNCH("ks_mac", "", kfn, grid, dim3(1 << MP::T), MP::smem_bytes, st, a, tab);
            };
            if (K1 && ks.lazy_mac)
                launch(ks_mac_kernel<LOGN, K1, K1>); // LAZY only exists for the k = 1 kernels
            else
                launch(ks_mac_kernel<LOGN, K1, false>);
        }
        if (front_only)
            return red;
        // Step 4 (:2353-2357): lazy inverse NTT of the special slots.
        ntt_inverse_plain<LOGN>(
            tab, ModMap{ 0, static_cast<int>(ctx->L0), 0 }, red + L * N, E * N, red + L * N, E * N, static_cast<int>(k),
            static_cast<int>(2 * B), true, st, "ks_special", ctx->wide_inv);
        // Step 5 (mod-down) + final add (:2360-2366).
        const MdLoadArgs la{ ks, red };
        const MdStoreArgs sa{ ks, red, o.add_src, o.add_bs, o.add_ls, o.dst, o.dst_bs, o.dst_ls, o.table, o.zero_l1 || o.table != nullptr };
        const ModMap mm{ 1 << 30, 0, 0 };
        if (o.bfv)
        {
            // multi_special_primes.cpp:287-289: the accumulated c_i goes back to coefficient form
            // (lazy inverse NTT) and the correction is added there.
            const MdStoreBfvArgs ba{ la, sa };
            if constexpr (two)
            {
                launch_inv_contig<LOGN>(IoPlain<OUT_RAW>{ red, tmp, E * N, L * N }, tab, mm, iL, static_cast<int>(2 * B), st, "moddown_bfv");
                IoPair<PlainLoad, MdStoreBfv> io2{ PlainArgs{ tmp, L * N }, ba };
                launch_inv_cols<LOGN>(io2, tab, mm, iL, static_cast<int>(2 * B), st, "moddown_bfv");
            }
            else
            {
                IoPair<PlainLoad, MdStoreBfv> io{ PlainArgs{ red, E * N }, ba };
                launch_inv_contig<LOGN>(io, tab, mm, iL, static_cast<int>(2 * B), st, "moddown_bfv");
            }
        }
        else if constexpr (two)
        {
            IoPair<MdLoadT<K1>, PlainStore<OUT_RAW>> io1{ la, PlainStoreArgs{ tmp, L * N } };
            IoPair<PlainLoad, MdStore> io2{ PlainArgs{ tmp, L * N }, sa };
            if (ctx->fast_ntt)
            {
                IoPair<MdLoadT<K1, true>, PlainStore<OUT_RAW>> io1f{ la, PlainStoreArgs{ tmp, L * N } };
                launch_fwd_cols<LOGN, true>(io1f, tab, mm, iL, static_cast<int>(2 * B), st, "moddown");
                launch_fwd_contig<LOGN, true>(io2, tab, mm, iL, static_cast<int>(2 * B), st, "moddown");
            }
            else
            {
                launch_fwd_cols<LOGN>(io1, tab, mm, iL, static_cast<int>(2 * B), st, "moddown");
                launch_fwd_contig<LOGN>(io2, tab, mm, iL, static_cast<int>(2 * B), st, "moddown");
            }
        }
        else
        {
            IoPair<MdLoadT<K1>, MdStore> io{ la, sa };
            if (ctx->fast_ntt)
                launch_fwd_contig<LOGN, true>(IoPair<MdLoadT<K1, true>, MdStore>{ la, sa }, tab, mm, iL, static_cast<int>(2 * B), st, "moddown");
            else
                launch_fwd_contig<LOGN>(io, tab, mm, iL, static_cast<int>(2 * B), st, "moddown");
        }
        return red;
    }

    template <int LOGN>
    static void switch_key_core(
        gsb_ctx *ctx, void *scratch, const uint64_t *target, size_t target_stride, const SwitchKey &key, const KsOutput &o,
        size_t B, size_t L, cudaStream_t st)
    {
        if (ctx->k == 1)
            switch_key_core_t<LOGN, true>(ctx, scratch, target, target_stride, key, o, B, L, st);
        else
            switch_key_core_t<LOGN, false>(ctx, scratch, target, target_stride, key, o, B, L, st);
    }

    // Fused tail of multiply + relinearize + rescale (see TailInvLoad in gsb_ks_kernels.cuh): from the
    // reduced inner product `red` and the tensor product `c3` = [B][3][L][N] straight to the rescaled
    // ciphertext out = [B][2][L-1][N].  Only for k = 1 contexts under gsb_ctx::fast_ntt / wide_inv, L >= 2.
    // buf = B*2*2*N words of scratch.
    template <int LOGN>
    static void mrr_tail_core(gsb_ctx *ctx, uint64_t *buf, const uint64_t *red, const uint64_t *c3, uint64_t *out, size_t B, size_t L, cudaStream_t st)
    {
        constexpr bool two = Split<LOGN>::two_pass;
        const size_t N = ctx->N, LN = L * N;
        const KsDev ks = make_ksdev(ctx, L);
        const KsLevelConsts &lc = ctx->consts_for(L);
        const NttTables tab = ctx->ntt_tables();
        const int nz = static_cast<int>(2 * B), iL = static_cast<int>(L);
        // inverse transforms of (u_{L-1} P^-1 + c3_{L-1}) and of the special slot
        const ModMap mlast{ iL, static_cast<int>(ctx->L0), iL - 1 };
        const TailInvArgs ia{ ks, red, c3, 3 * LN, LN };
        IoPair<TailInvLoad, PlainStore<OUT_RAW>> io_inv{ ia, PlainStoreArgs{ buf, 2 * N } };
        launch_inv_contig<LOGN, true>(io_inv, tab, mlast, 2, nz, st, "tail_last");
        if constexpr (two)
            launch_inv_cols<LOGN, true>(IoPlain<OUT_RAW>{ buf, buf, 2 * N, 2 * N }, tab, mlast, 2, nz, st, "tail_last");
        // t and t1 per coefficient
        {
            const host::LimbTables &tl = ctx->tables[L - 1], &tp = ctx->tables[ctx->L0];
            TailMidArgs a{};
            a.buf = buf;
            a.last.q = tl.mod.q;
            a.last.two_q = tl.mod.q << 1;
            a.last.cr1 = tl.mod.cr1;
            a.sp.q = tp.mod.q;
            a.sp.two_q = tp.mod.q << 1;
            uint64_t P = tp.mod.q % tl.mod.q, inv = 0;
            require(host::inv_mod(P, tl.mod.q, inv), GSB_E_LOGIC, "moddown: inv_mod fail");
            a.ip = inv;
            a.ips = host::shoup(inv, tl.mod.q);
            a.N = static_cast<uint32_t>(N);
            a.total = 2 * B * N;
            GSB_LAUNCH("tail_mid", "", tail_mid_kernel, dim3(grid_for(a.total, 256)), dim3(256), 0, st, a);
        }
        // one forward transform per remaining limb; `out` doubles as the buffer between the two passes
        const ModMap mm{ 1 << 30, 0, 0 };
        const int ny = iL - 1;
        const TailLoadArgs la{ buf, lc.d_neg_half, lc.d_neg_p };
        const TailStoreArgs sa{ ks, red, c3, 3 * LN, LN, out, lc.d_inv_qlast };
        if constexpr (two)
        {
            IoPair<TailLoad, PlainStore<OUT_RAW>> io1{ la, PlainStoreArgs{ out, (L - 1) * N } };
            IoPair<PlainLoad, TailStore> io2{ PlainArgs{ out, (L - 1) * N }, sa };
            launch_fwd_cols<LOGN, true>(io1, tab, mm, ny, nz, st, "tail");
            launch_fwd_contig<LOGN, true>(io2, tab, mm, ny, nz, st, "tail");
        }
        else
            launch_fwd_contig<LOGN, true>(IoPair<TailLoad, TailStore>{ la, sa }, tab, mm, ny, nz, st, "tail");
    }

    // rescale for n_polys polynomials of L limbs (util/rns.cpp:777-851). t = n_polys*N words scratch.
    template <int LOGN>
    static void rescale_core(gsb_ctx *ctx, uint64_t *t, const uint64_t *in, uint64_t *out, size_t n_polys, size_t L, cudaStream_t st)
    {
        constexpr bool two = Split<LOGN>::two_pass;
        const size_t N = ctx->N;
        const KsLevelConsts &lc = ctx->consts_for(L);
        const NttTables tab = ctx->ntt_tables();
        const int nz = static_cast<int>(n_polys);
        const ModMap mlast{ 1 << 30, 0, static_cast<int>(L - 1) };
        // stage 1: t = barrett63(strict_iNTT(last limb) + half)
        if constexpr (two)
        {
            launch_inv_contig<LOGN>(IoPlain<OUT_RAW>{ in + (L - 1) * N, t, L * N, N }, tab, mlast, 1, nz, st, "rescale_last");
            IoPair<PlainLoad, RsLastStore> io{ PlainArgs{ t, N }, PlainStoreArgs{ t, N } };
            launch_inv_cols<LOGN>(io, tab, mlast, 1, nz, st, "rescale_last");
        }
        else
        {
            IoPair<PlainLoad, RsLastStore> io{ PlainArgs{ in + (L - 1) * N, L * N }, PlainStoreArgs{ t, N } };
            launch_inv_contig<LOGN>(io, tab, mlast, 1, nz, st, "rescale_last");
        }
        // stage 2: per remaining limb
        const RsLoadArgs la{ t, lc.d_neg_half, ctx->tables[L - 1].mod.q };
        const RsStoreArgs sa{ in, out, lc.d_inv_qlast, static_cast<uint32_t>(L) };
        const ModMap mm{ 1 << 30, 0, 0 };
        const int ny = static_cast<int>(L - 1);
        if constexpr (two)
        {
            // `out` doubles as the buffer between the two passes (same element, same thread)
            IoPair<RsLoad, PlainStore<OUT_RAW>> io1{ la, PlainStoreArgs{ out, (L - 1) * N } };
            IoPair<PlainLoad, RsStore> io2{ PlainArgs{ out, (L - 1) * N }, sa };
            if (ctx->fast_ntt)
            {
                IoPair<RsLoadT<true>, PlainStore<OUT_RAW>> io1f{ la, PlainStoreArgs{ out, (L - 1) * N } };
                launch_fwd_cols<LOGN, true>(io1f, tab, mm, ny, nz, st, "rescale");
                launch_fwd_contig<LOGN, true>(io2, tab, mm, ny, nz, st, "rescale");
            }
            else
            {
                launch_fwd_cols<LOGN>(io1, tab, mm, ny, nz, st, "rescale");
                launch_fwd_contig<LOGN>(io2, tab, mm, ny, nz, st, "rescale");
            }
        }
        else
        {
            IoPair<RsLoad, RsStore> io{ la, sa };
            if (ctx->fast_ntt)
                launch_fwd_contig<LOGN, true>(IoPair<RsLoadT<true>, RsStore>{ la, sa }, tab, mm, ny, nz, st, "rescale");
            else
                launch_fwd_contig<LOGN>(io, tab, mm, ny, nz, st, "rescale");
        }
    }

    static void ckks_multiply_core(gsb_ctx *ctx, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t B, size_t L, cudaStream_t st)
    {
        CkksMulArgs p{ a, b, out, ctx->d_limb, static_cast<uint32_t>(ctx->N), static_cast<uint32_t>(L), B * L * ctx->N / 2 };
        GSB_LAUNCH("ckks_multiply", "", ckks_multiply_kernel<true>, dim3(grid_for(p.pairs, 256)), dim3(256), 0, st, p);
    }

    static void gather_core(
        gsb_ctx *ctx, const uint64_t *in, size_t in_zs, uint64_t *out, const uint32_t *table, size_t n, size_t L, cudaStream_t st)
    {
        GatherArgs p{ in, in_zs, out, table, static_cast<uint32_t>(ctx->N), static_cast<uint32_t>(L), n * L * ctx->N };
        GSB_LAUNCH("galois_gather", "", galois_gather_kernel, dim3(grid_for(p.total, 256)), dim3(256), 0, st, p);
    }

#ifndef GSB_SCRATCH_GIB
#define GSB_SCRATCH_GIB 16
#endif
    // How many ciphertexts to push through one pass of the pipeline: bounded scratch (<= GSB_SCRATCH_GIB, 16 GiB: 64 ciphertexts per pass at N = 2^15, K = 16).
    static size_t chunk_for(size_t words_per_ct, size_t B)
    {
        const size_t budget = (size_t(GSB_SCRATCH_GIB) << 30) / sizeof(uint64_t);
        size_t c = budget / (words_per_ct + 64);
        c = c < 1 ? 1 : c;
        if (c >= B)
            return B;
        const size_t n_chunks = (B + c - 1) / c; // balance the chunks
        return (B + n_chunks - 1) / n_chunks;
    }

    static const SwitchKey &find_key(gsb_ctx *ctx, int kind, uint32_t elt)
    {
        std::lock_guard<std::mutex> lock(ctx->mu);
        if (kind == GSB_KEY_RELIN)
        {
            auto it = ctx->relin.find(elt);
            require(it != ctx->relin.end(), GSB_E_INVALID_ARG, "not enough relinearization keys");
            return it->second;
        }
        require(kind == GSB_KEY_GALOIS, GSB_E_INVALID_ARG, "key kind");
        auto it = ctx->galois.find(elt);
        require(it != ctx->galois.end(), GSB_E_INVALID_ARG, "Galois key not present");
        return it->second;
    }

    static gsb_ws *pick_ws(gsb_ctx *ctx, gsb_ws *ws)
    {
        require(ws == nullptr || ws->ctx == ctx, GSB_E_INVALID_ARG, "workspace belongs to another context");
        return ws ? ws : &ctx->default_ws;
    }

    // Decryptor::compute_secret_key_array (decryptor.cpp:151-216): make sure s^1 .. s^n_pow exist on the
    // device; new powers are dyadic products of the previous power with s over all K key-level limbs.
    static const uint64_t *secret_key_powers(gsb_ctx *ctx, size_t n_pow, cudaStream_t st)
    {
        std::lock_guard<std::mutex> lock(ctx->mu);
        require(ctx->d_sk != nullptr, GSB_E_LOGIC, "secret key is not set");
        if (ctx->sk_powers >= n_pow)
            return ctx->d_sk;
        const size_t KN = ctx->K * ctx->N;
        uint64_t *d = nullptr;
        GSB_CUDA_CHECK(cudaMalloc(reinterpret_cast<void **>(&d), n_pow * KN * sizeof(uint64_t)));
        GSB_CUDA_CHECK(cudaMemcpyAsync(d, ctx->d_sk, ctx->sk_powers * KN * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
        for (size_t p = ctx->sk_powers; p < n_pow; p++)
        {
            PolyOpArgs a{};
            a.a = d + (p - 1) * KN;
            a.b = d;
            a.out = d + p * KN;
            a.limb = ctx->d_limb;
            a.N = static_cast<uint32_t>(ctx->N);
            a.n_limbs = static_cast<uint32_t>(ctx->K);
            a.limb_offset = 0;
            a.total = KN;
            launch_poly_op(GSB_OP_DYADIC, a, st);
        }
        GSB_CUDA_CHECK(cudaStreamSynchronize(st)); // published to every stream from here on
        ctx->sk_retired.emplace_back(ctx->d_sk, ctx->sk_powers * KN * sizeof(uint64_t));
        ctx->d_sk = d;
        ctx->sk_powers = n_pow;
        return d;
    }

    // dot_product_ct_sk_array for B ciphertexts; scratch (coefficient form only) = B*(size-1)*L*N words.
    template <int LOGN>
    static void dot_sk_core(
        gsb_ctx *ctx, uint64_t *scratch, const uint64_t *in, uint64_t *out, size_t B, size_t size, size_t L, bool ntt_form,
        cudaStream_t st)
    {
        const size_t N = ctx->N, LN = L * N;
        const uint64_t *skp = secret_key_powers(ctx, size - 1, st);
        DotSkArgs p{};
        p.skp = skp;
        p.out = out;
        p.limb = ctx->d_limb;
        p.N = static_cast<uint32_t>(N);
        p.L = static_cast<uint32_t>(L);
        p.K = static_cast<uint32_t>(ctx->K);
        p.n_pow = static_cast<uint32_t>(size - 1);
        p.pairs = B * LN / 2;
        const NttTables tab = ctx->ntt_tables();
        const ModMap mm{ 1 << 30, 0, 0 };
        if (ntt_form)
        {
            p.polys = in + LN;
            p.ct_stride = size * LN;
            p.c0 = in;
            p.c0_stride = size * LN;
            GSB_LAUNCH("dot_sk", "ntt", dot_sk_kernel, dim3(grid_for(p.pairs, 256)), dim3(256), 0, st, p);
            return;
        }
        // coefficient form (:241-262): lazy forward NTT of c_1.., dot product, strict inverse NTT, + c_0
        for (size_t pw = 0; pw + 1 < size; pw++)
            ntt_forward_plain<LOGN>(tab, mm, in + (pw + 1) * LN, size * LN, scratch + pw * LN, (size - 1) * LN, static_cast<int>(L),
                                    static_cast<int>(B), true, st, "dot_sk");
        p.polys = scratch;
        p.ct_stride = (size - 1) * LN;
        p.c0 = nullptr;
        GSB_LAUNCH("dot_sk", "coeff", dot_sk_kernel, dim3(grid_for(p.pairs, 256)), dim3(256), 0, st, p);
        ntt_inverse_plain<LOGN>(tab, mm, out, LN, out, LN, static_cast<int>(L), static_cast<int>(B), false, st, "dot_sk", ctx->wide_inv);
        // add_poly_coeffmod(destination, c_0) (:265): c_0 of ciphertext b sits at in + b*size*LN
        for (size_t b = 0; b < B; b++)
        {
            PolyOpArgs a{};
            a.a = out + b * LN;
            a.b = in + b * size * LN;
            a.out = out + b * LN;
            a.limb = ctx->d_limb;
            a.N = static_cast<uint32_t>(N);
            a.n_limbs = static_cast<uint32_t>(L);
            a.total = LN;
            launch_poly_op(GSB_OP_ADD, a, st);
        }
    }

    // Every *_dev entry point may be called from a thread whose current device is not the context's
    // (e.g. a worker thread of a multi-GPU host): allocations, attribute calls and launches must go there.
    static void check_batch(const gsb_ctx *ctx, size_t batch, size_t L, size_t min_L)
    {
        require(ctx != nullptr, GSB_E_INVALID_ARG, "ctx is null");
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        require(batch >= 1 && batch <= (size_t(1) << 20), GSB_E_INVALID_ARG, "batch out of range");
        require(L >= min_L && L <= ctx->L0, GSB_E_INVALID_ARG, "no ciphertext level with that many limbs");
    }
} // namespace gsb

using namespace gsb;

extern "C" {

int gsb_poly_op_host(
    gsb_ctx *ctx, int op, const uint64_t *a, const uint64_t *b, const uint64_t *scalars, uint64_t *out, size_t n_polys,
    size_t n_limbs, size_t limb_offset)
{
    int rc = GSB_OK;
    int g = guarded([&] {
        check_limbs(ctx, n_polys, n_limbs, limb_offset);
        require(a != nullptr && out != nullptr, GSB_E_INVALID_ARG, "null operand");
        const bool binary = op == GSB_OP_DYADIC || op == GSB_OP_ADD || op == GSB_OP_SUB;
        require(!binary || b != nullptr, GSB_E_INVALID_ARG, "operand2 is null");
        std::lock_guard<std::mutex> lock(ctx->host_mu);
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        const size_t words = n_polys * n_limbs * ctx->N, bytes = words * sizeof(uint64_t);
        uint64_t *d = static_cast<uint64_t *>(ctx->host_io.reserve(3 * bytes));
        GSB_CUDA_CHECK(cudaMemcpyAsync(d, a, bytes, cudaMemcpyHostToDevice, nullptr));
        if (binary)
            GSB_CUDA_CHECK(cudaMemcpyAsync(d + words, b, bytes, cudaMemcpyHostToDevice, nullptr));
        rc = gsb_poly_op_dev(ctx, op, d, binary ? d + words : nullptr, scalars, d + 2 * words, n_polys, n_limbs, limb_offset, nullptr);
        if (rc == GSB_OK)
        {
            GSB_CUDA_CHECK(cudaMemcpyAsync(out, d + 2 * words, bytes, cudaMemcpyDeviceToHost, nullptr));
            GSB_CUDA_CHECK(cudaStreamSynchronize(nullptr));
        }
    });
    return g != GSB_OK ? g : rc;
}

// Key upload: the bundles go to the device as they are (one plain H2D copy each); a kernel then checks every
// word against its limb's modulus (is_buffer_valid, valcheck.cpp) and permutes the 16-byte units inside every
// 1 KiB block into the bank-conflict-free device layout (key_word_pos, ntt_core.cuh) in place.
int gsb_key_upload_bundles(gsb_ctx *ctx, int kind, uint32_t galois_elt, const uint64_t *const *bundles, size_t n_bundles)
{
    return guarded([&] {
        require(ctx != nullptr && bundles != nullptr, GSB_E_INVALID_ARG, "null argument");
        require(kind == GSB_KEY_RELIN || kind == GSB_KEY_GALOIS, GSB_E_INVALID_ARG, "key kind");
        require(ctx->K > ctx->k, GSB_E_LOGIC, "keyswitching is not supported by the context");
        // keygenerator.cpp:333-343: one PublicKey per bundle of the first-level modulus
        const size_t beta0 = (ctx->L0 + ctx->k - 1) / ctx->k;
        require(n_bundles == beta0, GSB_E_INVALID_ARG, "kswitch_keys is not valid for encryption parameters");
        for (size_t j = 0; j < n_bundles; j++)
            require(bundles[j] != nullptr, GSB_E_INVALID_ARG, "null argument");
        if (kind == GSB_KEY_GALOIS)
            (void)ctx->galois_table(galois_elt); // validates the element and builds its permutation
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        const size_t per = 2 * ctx->K * ctx->N, words = n_bundles * per;
        uint64_t *d = nullptr;
        int *d_bad = nullptr;
        GSB_CUDA_CHECK(cudaMalloc(reinterpret_cast<void **>(&d), words * sizeof(uint64_t) + 256));
        d_bad = reinterpret_cast<int *>(d + words);
        struct Guard // frees the allocation unless the upload succeeded
        {
            uint64_t *p;
            ~Guard() { cudaFree(p); }
        } guard{ d };
        GSB_CUDA_CHECK(cudaMemsetAsync(d_bad, 0, sizeof(int), nullptr));
        for (size_t j = 0; j < n_bundles; j++)
            GSB_CUDA_CHECK(cudaMemcpyAsync(d + j * per, bundles[j], per * sizeof(uint64_t), cudaMemcpyHostToDevice, nullptr));
        KeySwizzleArgs a{ d, ctx->d_limb, static_cast<uint32_t>(ctx->N), static_cast<uint32_t>(ctx->K), words / 128, d_bad };
        GSB_LAUNCH("key_swizzle", "", key_swizzle_kernel, dim3(grid_for(a.blocks * 64, 256)), dim3(256), 0, nullptr, a);
        int bad = 0;
        GSB_CUDA_CHECK(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, nullptr));
        GSB_CUDA_CHECK(cudaStreamSynchronize(nullptr));
        require(bad == 0, GSB_E_INVALID_ARG, "kswitch_keys is not valid for encryption parameters");
        std::lock_guard<std::mutex> lock(ctx->mu);
        SwitchKey &slot = kind == GSB_KEY_RELIN ? ctx->relin[galois_elt] : ctx->galois[galois_elt];
        if (slot.d)
            cudaFree(slot.d);
        slot.d = d;
        slot.bundles = n_bundles;
        guard.p = nullptr;
    });
}

int gsb_key_upload(gsb_ctx *ctx, int kind, uint32_t galois_elt, const uint64_t *key_host, size_t n_bundles)
{
    if (ctx == nullptr || key_host == nullptr || n_bundles == 0 || n_bundles > 64)
        return guarded([&] { require(false, GSB_E_INVALID_ARG, ctx && key_host ? "kswitch_keys is not valid for encryption parameters" : "null argument"); });
    const uint64_t *ptrs[64];
    for (size_t j = 0; j < n_bundles; j++)
        ptrs[j] = key_host + j * 2 * ctx->K * ctx->N;
    return gsb_key_upload_bundles(ctx, kind, galois_elt, ptrs, n_bundles);
}

int gsb_key_drop(gsb_ctx *ctx, int kind, uint32_t galois_elt)
{
    return guarded([&] {
        require(ctx != nullptr, GSB_E_INVALID_ARG, "ctx is null");
        std::lock_guard<std::mutex> lock(ctx->mu);
        if (kind == GSB_KEY_RELIN)
        {
            auto it = ctx->relin.find(galois_elt);
            require(it != ctx->relin.end(), GSB_E_INVALID_ARG, "relinearization key not present");
            cudaFree(it->second.d);
            ctx->relin.erase(it);
        }
        else
        {
            auto it = ctx->galois.find(galois_elt);
            require(it != ctx->galois.end(), GSB_E_INVALID_ARG, "Galois key not present");
            cudaFree(it->second.d);
            ctx->galois.erase(it);
        }
    });
}

int gsb_galois_elt_from_step(const gsb_ctx *ctx, int step, uint32_t *elt)
{
    return guarded([&] {
        require(ctx != nullptr && elt != nullptr, GSB_E_INVALID_ARG, "null argument");
        const uint32_t e = host::galois_elt_from_step(ctx->logn, step);
        require(e != 0, GSB_E_INVALID_ARG, "step count too large"); // util/galois.cpp:62-65
        *elt = e;
    });
}

int gsb_galois_table(const gsb_ctx *ctx, uint32_t galois_elt, uint32_t *table_out)
{
    return guarded([&] {
        require(ctx != nullptr && table_out != nullptr, GSB_E_INVALID_ARG, "null argument");
        require((galois_elt & 1u) && galois_elt < 2u * ctx->N, GSB_E_INVALID_ARG, "Galois element is not valid");
        std::vector<uint32_t> t;
        host::galois_table_ntt(ctx->logn, galois_elt, t);
        memcpy(table_out, t.data(), t.size() * sizeof(uint32_t));
    });
}

int gsb_ckks_multiply_dev(gsb_ctx *ctx, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t batch, size_t L, void *stream)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 1);
        require(a && b && out, GSB_E_INVALID_ARG, "null operand");
        ckks_multiply_core(ctx, a, b, out, batch, L, static_cast<cudaStream_t>(stream));
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

int gsb_switch_key_dev(
    gsb_ctx *ctx, gsb_ws *ws, uint64_t *enc, size_t enc_size, const uint64_t *target, size_t target_stride, int key_kind,
    uint32_t galois_elt, size_t batch, size_t L, void *stream)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 1);
        require(enc && target && enc_size >= 2, GSB_E_INVALID_ARG, "invalid ciphertext");
        const SwitchKey &key = find_key(ctx, key_kind, galois_elt);
        gsb_ws *w = pick_ws(ctx, ws);
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const size_t N = ctx->N, per = ks_words_per_ct(ctx, L), chunk = chunk_for(per, batch);
        void *scratch = w->reserve((chunk * per + 1024) * sizeof(uint64_t));
        for (size_t b0 = 0; b0 < batch; b0 += chunk)
        {
            const size_t B = std::min(chunk, batch - b0);
            uint64_t *e = enc + b0 * enc_size * L * N;
            const KsOutput o{ e, enc_size * L * N, L * N, e, enc_size * L * N, L * N, nullptr };
            GSB_DISPATCH_LOGN(ctx->logn, switch_key_core<LOGN>(ctx, scratch, target + b0 * target_stride, target_stride, key, o, B, L, st));
        }
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

static int relinearize_impl(gsb_ctx *ctx, gsb_ws *ws, const uint64_t *in, uint64_t *out, size_t batch, size_t L, void *stream, bool bfv)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 1);
        require(in && out, GSB_E_INVALID_ARG, "null operand");
        const SwitchKey &key = find_key(ctx, GSB_KEY_RELIN, 0);
        gsb_ws *w = pick_ws(ctx, ws);
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const size_t N = ctx->N, LN = L * N, per = ks_words_per_ct(ctx, L), chunk = chunk_for(per, batch);
        void *scratch = w->reserve((chunk * per + 1024) * sizeof(uint64_t));
        for (size_t b0 = 0; b0 < batch; b0 += chunk)
        {
            const size_t B = std::min(chunk, batch - b0);
            const uint64_t *i3 = in + b0 * 3 * LN;
            const KsOutput o{ i3, 3 * LN, LN, out + b0 * 2 * LN, 2 * LN, LN, nullptr, false, bfv };
            GSB_DISPATCH_LOGN(ctx->logn, switch_key_core<LOGN>(ctx, scratch, i3 + 2 * LN, 3 * LN, key, o, B, L, st));
        }
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

int gsb_relinearize_dev(gsb_ctx *ctx, gsb_ws *ws, const uint64_t *in, uint64_t *out, size_t batch, size_t L, void *stream)
{
    return relinearize_impl(ctx, ws, in, out, batch, L, stream, false);
}

int gsb_bfv_relinearize_dev(gsb_ctx *ctx, gsb_ws *ws, const uint64_t *in, uint64_t *out, size_t batch, size_t L, void *stream)
{
    return relinearize_impl(ctx, ws, in, out, batch, L, stream, true);
}

int gsb_bfv_multiply_dev(
    gsb_ctx *ctx, gsb_ws *ws, uint64_t plain_modulus, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t batch, size_t L, void *stream)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 1);
        require(a && b && out, GSB_E_INVALID_ARG, "null operand");
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        const BfvTool &tool = ctx->bfv_tool(L, plain_modulus);
        gsb_ws *w = pick_ws(ctx, ws);
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const size_t N = ctx->N, M = L + tool.nBsk, LN = L * N, MN = M * N;
        const size_t per = 4 * MN + 3 * MN, chunk = chunk_for(per, batch);
        void *scratch = w->reserve((chunk * per + 1024) * sizeof(uint64_t));
        const NttTables tab{ tool.d_WW, tool.d_VV, tool.d_limb, static_cast<uint32_t>(N), ctx->logn };
        const ModMap mm{ 1 << 30, 0, 0 };
        for (size_t b0 = 0; b0 < batch; b0 += chunk)
        {
            const size_t B = std::min(chunk, batch - b0);
            Arena ar(scratch);
            uint64_t *ext = ar.words(4 * B * MN); // [2 operands][B][2][M][N]
            uint64_t *dest = ar.words(3 * B * MN); // [B][3][M][N]
            // (1)-(2): lift both operands to q u Bsk; (3): lazy NTT of every limb of every polynomial
            const size_t lift = 2 * B * N;
            GSB_LAUNCH("behz_lift", "", behz_lift_kernel, dim3(grid_for(lift, 256)), dim3(256), 0, st, tool.dev, a + b0 * 2 * LN, ext, lift);
            GSB_LAUNCH("behz_lift", "", behz_lift_kernel, dim3(grid_for(lift, 256)), dim3(256), 0, st, tool.dev, b + b0 * 2 * LN, ext + 2 * B * MN, lift);
            GSB_DISPATCH_LOGN(ctx->logn, ntt_forward_plain<LOGN>(tab, mm, ext, MN, ext, MN, static_cast<int>(M), static_cast<int>(4 * B), true, st, "bfv"));
            // (4): the size-2 x size-2 tensor product over all M limbs (same dyadic + add arithmetic as CKKS)
            CkksMulArgs p{ ext, ext + 2 * B * MN, dest, tool.d_limb, static_cast<uint32_t>(N), static_cast<uint32_t>(M), B * MN / 2 };
            GSB_LAUNCH("bfv_tensor", "", ckks_multiply_kernel<false>, dim3(grid_for(p.pairs, 256)), dim3(256), 0, st, p);
            // (5): strict inverse NTT
            GSB_DISPATCH_LOGN(ctx->logn, ntt_inverse_plain<LOGN>(tab, mm, dest, MN, dest, MN, static_cast<int>(M), static_cast<int>(3 * B), false, st, "bfv"));
            // (6)-(8): times t, divide by q and floor into Bsk, Shenoy-Kumaresan back to q
            const size_t fl = 3 * B * N;
            GSB_LAUNCH("behz_floor", "", behz_floor_kernel, dim3(grid_for(fl, 256)), dim3(256), 0, st, tool.dev, dest, out + b0 * 3 * LN, fl);
        }
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

int gsb_bfv_mod_switch_dev(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L, void *stream)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 2); // evaluator.cpp:1006-1009: needs a next level
        require(in && out && size >= 1 && size <= 16, GSB_E_INVALID_ARG, "invalid ciphertext");
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        const KsLevelConsts &lc = ctx->consts_for(L);
        BfvModSwitchArgs p{ in, out, ctx->d_limb, lc.d_inv_qlast, static_cast<uint32_t>(ctx->N), static_cast<uint32_t>(L),
                            batch * size * (L - 1) * ctx->N };
        GSB_LAUNCH("bfv_mod_switch", "", bfv_mod_switch_kernel, dim3(grid_for(p.total, 256)), dim3(256), 0,
                   static_cast<cudaStream_t>(stream), p);
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

int gsb_bfv_apply_galois_dev(gsb_ctx *ctx, gsb_ws *ws, uint32_t galois_elt, const uint64_t *in, uint64_t *out, size_t batch, size_t L, void *stream)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 1);
        require(in && out, GSB_E_INVALID_ARG, "null operand");
        require((galois_elt & 1u) && galois_elt < 2u * ctx->N, GSB_E_INVALID_ARG, "Galois element is not valid");
        const SwitchKey &key = find_key(ctx, GSB_KEY_GALOIS, galois_elt);
        gsb_ws *w = pick_ws(ctx, ws);
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const size_t N = ctx->N, LN = L * N, per = ks_words_per_ct(ctx, L) + 2 * LN, chunk = chunk_for(per, batch);
        char *scratch = static_cast<char *>(w->reserve((chunk * per + 1024) * sizeof(uint64_t)));
        for (size_t b0 = 0; b0 < batch; b0 += chunk)
        {
            const size_t B = std::min(chunk, batch - b0);
            // evaluator.cpp:1893-1904: both components through the coefficient-form automorphism; the
            // permuted c1 is the key-switch target, the permuted c0 the addend of component 0.
            uint64_t *perm = reinterpret_cast<uint64_t *>(scratch);
            GaloisCoeffArgs p{ in + b0 * 2 * LN, LN, perm, ctx->d_limb, galois_elt, static_cast<uint32_t>(N), static_cast<uint32_t>(L), ctx->logn, B * 2 * LN };
            GSB_LAUNCH("galois_coeff", "", galois_coeff_kernel, dim3(grid_for(p.total, 256)), dim3(256), 0, st, p);
            const KsOutput o{ perm, 2 * LN, LN, out + b0 * 2 * LN, 2 * LN, LN, nullptr, true, true };
            GSB_DISPATCH_LOGN(ctx->logn, switch_key_core<LOGN>(ctx, scratch + B * 2 * LN * sizeof(uint64_t), perm + LN, 2 * LN, key, o, B, L, st));
        }
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

int gsb_rescale_dev(gsb_ctx *ctx, gsb_ws *ws, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L, void *stream)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 2); // evaluator.cpp:845-848: needs a next level
        require(in && out && size >= 1 && size <= 16, GSB_E_INVALID_ARG, "invalid ciphertext");
        gsb_ws *w = pick_ws(ctx, ws);
        const size_t n_polys = batch * size;
        uint64_t *t = static_cast<uint64_t *>(w->reserve(n_polys * ctx->N * sizeof(uint64_t)));
        GSB_DISPATCH_LOGN(ctx->logn, rescale_core<LOGN>(ctx, t, in, out, n_polys, L, static_cast<cudaStream_t>(stream)));
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

int gsb_apply_galois_dev(gsb_ctx *ctx, gsb_ws *ws, uint32_t galois_elt, const uint64_t *in, uint64_t *out, size_t batch, size_t L, void *stream)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 1);
        require(in && out, GSB_E_INVALID_ARG, "null operand");
        const uint32_t *table = ctx->galois_table(galois_elt);
        const SwitchKey &key = find_key(ctx, GSB_KEY_GALOIS, galois_elt);
        gsb_ws *w = pick_ws(ctx, ws);
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const size_t N = ctx->N, LN = L * N, per = ks_words_per_ct(ctx, L) + LN, chunk = chunk_for(per, batch);
        char *scratch = static_cast<char *>(w->reserve((chunk * per + 1024) * sizeof(uint64_t)));
        for (size_t b0 = 0; b0 < batch; b0 += chunk)
        {
            const size_t B = std::min(chunk, batch - b0);
            const uint64_t *i2 = in + b0 * 2 * LN;
            // temp = apply_galois_ntt(c1) (evaluator.cpp:1920); c0's permutation (:1913-1917) and the
            // zeroed c1 (:1928) are folded into the mod-down epilogue.
            uint64_t *tperm = reinterpret_cast<uint64_t *>(scratch);
            gather_core(ctx, i2 + LN, 2 * LN, tperm, table, B, L, st);
            const KsOutput o{ i2, 2 * LN, LN, out + b0 * 2 * LN, 2 * LN, LN, table };
            GSB_DISPATCH_LOGN(ctx->logn, switch_key_core<LOGN>(ctx, scratch + B * LN * sizeof(uint64_t), tperm, LN, key, o, B, L, st));
        }
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

int gsb_mul_relin_rescale_dev(
    gsb_ctx *ctx, gsb_ws *ws, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t batch, size_t L, void *stream)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 2);
        require(a && b && out, GSB_E_INVALID_ARG, "null operand");
        const SwitchKey &key = find_key(ctx, GSB_KEY_RELIN, 0);
        gsb_ws *w = pick_ws(ctx, ws);
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const size_t N = ctx->N, LN = L * N;
        // one forward transform per limb for mod-down + rescale together (identical words, see mrr_tail_core)
        const bool fused_tail = ctx->k == 1 && ctx->fast_ntt && ctx->wide_inv && !ctx->split_tail;
        const size_t per = ks_words_per_ct(ctx, L) + 3 * LN + (fused_tail ? 4 * N : 2 * LN + 2 * N), chunk = chunk_for(per, batch);
        void *scratch = w->reserve((chunk * per + 2048) * sizeof(uint64_t));
        for (size_t b0 = 0; b0 < batch; b0 += chunk)
        {
            const size_t B = std::min(chunk, batch - b0);
            Arena ar(scratch);
            uint64_t *c3 = ar.words(B * 3 * LN);
            ckks_multiply_core(ctx, a + b0 * 2 * LN, b + b0 * 2 * LN, c3, B, L, st);
            if (fused_tail)
            {
                uint64_t *buf = ar.words(B * 4 * N);
                const KsOutput o{ c3, 3 * LN, LN, nullptr, 0, 0, nullptr };
                uint64_t *red = nullptr;
                GSB_DISPATCH_LOGN(ctx->logn, red = (switch_key_core_t<LOGN, true>(ctx, ar.base + ar.off, c3 + 2 * LN, 3 * LN, key, o, B, L, st, true)));
                GSB_DISPATCH_LOGN(ctx->logn, mrr_tail_core<LOGN>(ctx, buf, red, c3, out + b0 * 2 * (L - 1) * N, B, L, st));
                continue;
            }
            uint64_t *c2 = ar.words(B * 2 * LN), *t = ar.words(B * 2 * N);
            const KsOutput o{ c3, 3 * LN, LN, c2, 2 * LN, LN, nullptr };
            GSB_DISPATCH_LOGN(ctx->logn, switch_key_core<LOGN>(ctx, ar.base + ar.off, c3 + 2 * LN, 3 * LN, key, o, B, L, st));
            GSB_DISPATCH_LOGN(ctx->logn, rescale_core<LOGN>(ctx, t, c2, out + b0 * 2 * (L - 1) * N, B * 2, L, st));
        }
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

// ---- host-buffer forms -----------------------------------------------------------------------------
// Copy in, run, copy out -- pipelined: the batch is cut into chunks of a few ciphertexts and chunk
// i+1's host->device copy, chunk i's kernels and chunk i-1's device->host copy run concurrently on
// three streams over two staging sets (PCIe is full duplex; with pinned host buffers the copies
// are real DMA).  w0/w1/wo = words per ciphertext of the two inputs and the output.
static int host_call(
    gsb_ctx *ctx, const uint64_t *in0, size_t w0, const uint64_t *in1, size_t w1, uint64_t *out, size_t wo, size_t batch,
    const std::function<int(gsb_ws *, uint64_t *, uint64_t *, uint64_t *, size_t, void *)> &run)
{
    int rc = GSB_OK;
    int g = guarded([&] {
        require(ctx != nullptr, GSB_E_INVALID_ARG, "ctx is null");
        require(in0 != nullptr && out != nullptr && (in1 != nullptr || w1 == 0), GSB_E_INVALID_ARG, "null operand");
        require(batch >= 1 && batch <= (size_t(1) << 20), GSB_E_INVALID_ARG, "batch out of range");
        require(w0 + w1 > 0 && wo > 0, GSB_E_INVALID_ARG, "no ciphertext level with that many limbs");
        std::lock_guard<std::mutex> lock(ctx->host_mu);
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        if (!ctx->pipe_ready)
        {
            for (int i = 0; i < 3; i++)
                GSB_CUDA_CHECK(cudaStreamCreateWithFlags(&ctx->pipe_stream[i], cudaStreamNonBlocking));
            for (int i = 0; i < 6; i++)
                GSB_CUDA_CHECK(cudaEventCreateWithFlags(&ctx->pipe_event[i], cudaEventDisableTiming));
            ctx->pipe_ready = true;
        }
        cudaStream_t s_in = ctx->pipe_stream[0], s_run = ctx->pipe_stream[1], s_out = ctx->pipe_stream[2];
        // chunk: ~128 MiB of input per chunk, at least 1 ciphertext, at most the batch
        size_t chunk = (size_t(128) << 20) / ((w0 + w1) * sizeof(uint64_t));
        chunk = chunk < 1 ? 1 : chunk > batch ? batch : chunk;
        const size_t per_set = chunk * (w0 + w1 + wo);
        uint64_t *stage = static_cast<uint64_t *>(ctx->host_io.reserve(2 * per_set * sizeof(uint64_t)));
        cudaEvent_t *in_done = ctx->pipe_event, *run_done = ctx->pipe_event + 2, *out_done = ctx->pipe_event + 4;
        // Chunk schedule: full chunks, then a halving tail.  The host->device copies run back to back
        // and bound the call, so what matters is how little work is left once the last input byte has
        // arrived: the last chunk's kernels and its copy-out.
        std::vector<size_t> sizes;
        size_t rem = batch;
        for (; rem > chunk; rem -= chunk)
            sizes.push_back(chunk);
        for (; rem > 1; rem -= (rem + 1) / 2)
            sizes.push_back((rem + 1) / 2);
        if (rem)
            sizes.push_back(rem);
        size_t b0 = 0;
        for (size_t n = 0; n < sizes.size() && rc == GSB_OK; b0 += sizes[n], n++)
        {
            const size_t B = sizes[n];
            const int set = static_cast<int>(n & 1);
            uint64_t *d0 = stage + set * per_set, *d1 = d0 + chunk * w0, *dout = d1 + chunk * w1;
            if (n >= 2) // the set's inputs are free once its previous kernels ran, its output once copied out
            {
                GSB_CUDA_CHECK(cudaStreamWaitEvent(s_in, run_done[set], 0));
                GSB_CUDA_CHECK(cudaStreamWaitEvent(s_run, out_done[set], 0));
            }
            GSB_CUDA_CHECK(cudaMemcpyAsync(d0, in0 + b0 * w0, B * w0 * sizeof(uint64_t), cudaMemcpyHostToDevice, s_in));
            if (w1)
                GSB_CUDA_CHECK(cudaMemcpyAsync(d1, in1 + b0 * w1, B * w1 * sizeof(uint64_t), cudaMemcpyHostToDevice, s_in));
            GSB_CUDA_CHECK(cudaEventRecord(in_done[set], s_in));
            GSB_CUDA_CHECK(cudaStreamWaitEvent(s_run, in_done[set], 0));
            rc = run(&ctx->default_ws, d0, d1, dout, B, s_run);
            GSB_CUDA_CHECK(cudaEventRecord(run_done[set], s_run));
            GSB_CUDA_CHECK(cudaStreamWaitEvent(s_out, run_done[set], 0));
            GSB_CUDA_CHECK(cudaMemcpyAsync(out + b0 * wo, dout, B * wo * sizeof(uint64_t), cudaMemcpyDeviceToHost, s_out));
            GSB_CUDA_CHECK(cudaEventRecord(out_done[set], s_out));
        }
        for (int i = 0; i < 3; i++)
            GSB_CUDA_CHECK(cudaStreamSynchronize(ctx->pipe_stream[i]));
    });
    return g != GSB_OK ? g : rc;
}

int gsb_ckks_multiply_host(gsb_ctx *ctx, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t batch, size_t L)
{
    const size_t w = ctx ? L * ctx->N : 0;
    return host_call(ctx, a, 2 * w, b, 2 * w, out, 3 * w, batch, [&](gsb_ws *, uint64_t *da, uint64_t *db, uint64_t *dout, size_t B, void *st) {
        return gsb_ckks_multiply_dev(ctx, da, db, dout, B, L, st);
    });
}

int gsb_relinearize_host(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t L)
{
    const size_t w = ctx ? L * ctx->N : 0;
    return host_call(ctx, in, 3 * w, nullptr, 0, out, 2 * w, batch, [&](gsb_ws *ws, uint64_t *di, uint64_t *, uint64_t *dout, size_t B, void *st) {
        return gsb_relinearize_dev(ctx, ws, di, dout, B, L, st);
    });
}

int gsb_rescale_host(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L)
{
    const size_t w = ctx ? size * ctx->N : 0;
    return host_call(ctx, in, w * L, nullptr, 0, out, L ? w * (L - 1) : 0, batch, [&](gsb_ws *ws, uint64_t *di, uint64_t *, uint64_t *dout, size_t B, void *st) {
        return gsb_rescale_dev(ctx, ws, di, dout, B, size, L, st);
    });
}

int gsb_apply_galois_host(gsb_ctx *ctx, uint32_t galois_elt, const uint64_t *in, uint64_t *out, size_t batch, size_t L)
{
    const size_t w = ctx ? L * ctx->N : 0;
    return host_call(ctx, in, 2 * w, nullptr, 0, out, 2 * w, batch, [&](gsb_ws *ws, uint64_t *di, uint64_t *, uint64_t *dout, size_t B, void *st) {
        return gsb_apply_galois_dev(ctx, ws, galois_elt, di, dout, B, L, st);
    });
}

int gsb_bfv_multiply_host(gsb_ctx *ctx, uint64_t plain_modulus, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t batch, size_t L)
{
    const size_t w = ctx ? L * ctx->N : 0;
    return host_call(ctx, a, 2 * w, b, 2 * w, out, 3 * w, batch, [&](gsb_ws *ws, uint64_t *da, uint64_t *db, uint64_t *dout, size_t B, void *st) {
        return gsb_bfv_multiply_dev(ctx, ws, plain_modulus, da, db, dout, B, L, st);
    });
}

int gsb_bfv_aux_base(gsb_ctx *ctx, size_t L, uint64_t plain_modulus, uint64_t *moduli_out, size_t cap, size_t *n_out)
{
    return guarded([&] {
        require(ctx != nullptr && moduli_out != nullptr && n_out != nullptr, GSB_E_INVALID_ARG, "null argument");
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        const BfvTool &tool = ctx->bfv_tool(L, plain_modulus);
        require(cap >= tool.bsk.size(), GSB_E_RANGE, "output too small");
        for (size_t i = 0; i < tool.bsk.size(); i++)
            moduli_out[i] = tool.bsk[i];
        *n_out = tool.bsk.size();
    });
}

int gsb_bfv_mod_switch_host(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L)
{
    const size_t w = ctx ? size * ctx->N : 0;
    return host_call(ctx, in, w * L, nullptr, 0, out, L ? w * (L - 1) : 0, batch, [&](gsb_ws *, uint64_t *di, uint64_t *, uint64_t *dout, size_t B, void *st) {
        return gsb_bfv_mod_switch_dev(ctx, di, dout, B, size, L, st);
    });
}

int gsb_bfv_relinearize_host(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t L)
{
    const size_t w = ctx ? L * ctx->N : 0;
    return host_call(ctx, in, 3 * w, nullptr, 0, out, 2 * w, batch, [&](gsb_ws *ws, uint64_t *di, uint64_t *, uint64_t *dout, size_t B, void *st) {
        return gsb_bfv_relinearize_dev(ctx, ws, di, dout, B, L, st);
    });
}

int gsb_bfv_apply_galois_host(gsb_ctx *ctx, uint32_t galois_elt, const uint64_t *in, uint64_t *out, size_t batch, size_t L)
{
    const size_t w = ctx ? L * ctx->N : 0;
    return host_call(ctx, in, 2 * w, nullptr, 0, out, 2 * w, batch, [&](gsb_ws *ws, uint64_t *di, uint64_t *, uint64_t *dout, size_t B, void *st) {
        return gsb_bfv_apply_galois_dev(ctx, ws, galois_elt, di, dout, B, L, st);
    });
}

// ---- Decryptor ---------------------------------------------------------------
int gsb_secret_key_set(gsb_ctx *ctx, const uint64_t *sk_host)
{
    return guarded([&] {
        require(ctx != nullptr && sk_host != nullptr, GSB_E_INVALID_ARG, "null argument");
        // Decryptor::Decryptor (decryptor.cpp:34-37) -> is_valid_for(SecretKey): canonical residues at key level
        for (size_t i = 0; i < ctx->K; i++)
        {
            const uint64_t q = ctx->tables[i].mod.q;
            for (size_t d = 0; d < ctx->N; d++)
                require(sk_host[i * ctx->N + d] < q, GSB_E_INVALID_ARG, "secret key is not valid for encryption parameters");
        }
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        uint64_t *d = nullptr;
        const size_t bytes = ctx->K * ctx->N * sizeof(uint64_t);
        GSB_CUDA_CHECK(cudaMalloc(reinterpret_cast<void **>(&d), bytes));
        GSB_CUDA_CHECK(cudaMemcpy(d, sk_host, bytes, cudaMemcpyHostToDevice));
        std::lock_guard<std::mutex> lock(ctx->mu);
        if (ctx->d_sk)
            ctx->sk_retired.emplace_back(ctx->d_sk, ctx->sk_powers * bytes);
        ctx->d_sk = d;
        ctx->sk_powers = 1;
    });
}

int gsb_dot_product_ct_sk_dev(
    gsb_ctx *ctx, gsb_ws *ws, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L, int ntt_form, void *stream)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 1);
        require(in && out && size >= 2 && size <= 16, GSB_E_INVALID_ARG, "invalid ciphertext");
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        gsb_ws *w = pick_ws(ctx, ws);
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const size_t LN = L * ctx->N, per = ntt_form ? 0 : (size - 1) * LN;
        const size_t chunk = ntt_form ? batch : chunk_for(per, batch);
        uint64_t *scratch = per ? static_cast<uint64_t *>(w->reserve((chunk * per + 1024) * sizeof(uint64_t))) : nullptr;
        for (size_t b0 = 0; b0 < batch; b0 += chunk)
        {
            const size_t B = std::min(chunk, batch - b0);
            GSB_DISPATCH_LOGN(ctx->logn, dot_sk_core<LOGN>(ctx, scratch, in + b0 * size * LN, out + b0 * LN, B, size, L, ntt_form != 0, st));
        }
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

int gsb_bfv_decrypt_dev(
    gsb_ctx *ctx, gsb_ws *ws, uint64_t plain_modulus, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L, void *stream)
{
    return guarded([&] {
        check_batch(ctx, batch, L, 1);
        require(in && out && size >= 2 && size <= 16, GSB_E_INVALID_ARG, "invalid ciphertext");
        require(plain_modulus >= 2 && (plain_modulus >> 60) == 0, GSB_E_INVALID_ARG, "plain_modulus out of range");
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        const BfvTool &tool = ctx->bfv_tool(L, plain_modulus);
        gsb_ws *w = pick_ws(ctx, ws);
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        const size_t N = ctx->N, LN = L * N, per = (size - 1) * LN + LN, chunk = chunk_for(per, batch);
        uint64_t *scratch = static_cast<uint64_t *>(w->reserve((chunk * per + 1024) * sizeof(uint64_t)));
        for (size_t b0 = 0; b0 < batch; b0 += chunk)
        {
            const size_t B = std::min(chunk, batch - b0);
            uint64_t *dot = scratch + B * (size - 1) * LN;
            GSB_DISPATCH_LOGN(ctx->logn, dot_sk_core<LOGN>(ctx, scratch, in + b0 * size * LN, dot, B, size, L, false, st));
            DecryptRoundArgs p = tool.dec;
            p.dot = dot;
            p.out = out + b0 * N;
            p.total = B * N;
            GSB_LAUNCH("bfv_decrypt", "", decrypt_round_kernel, dim3(grid_for(p.total, 256)), dim3(256), 0, st, p);
        }
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

int gsb_dot_product_ct_sk_host(gsb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L, int ntt_form)
{
    const size_t w = ctx ? L * ctx->N : 0;
    return host_call(ctx, in, size * w, nullptr, 0, out, w, batch, [&](gsb_ws *ws, uint64_t *di, uint64_t *, uint64_t *dout, size_t B, void *st) {
        return gsb_dot_product_ct_sk_dev(ctx, ws, di, dout, B, size, L, ntt_form, st);
    });
}

int gsb_bfv_decrypt_host(gsb_ctx *ctx, uint64_t plain_modulus, const uint64_t *in, uint64_t *out, size_t batch, size_t size, size_t L)
{
    const size_t w = ctx ? L * ctx->N : 0;
    return host_call(ctx, in, size * w, nullptr, 0, out, ctx ? ctx->N : 0, batch, [&](gsb_ws *ws, uint64_t *di, uint64_t *, uint64_t *dout, size_t B, void *st) {
        return gsb_bfv_decrypt_dev(ctx, ws, plain_modulus, di, dout, B, size, L, st);
    });
}

int gsb_mul_relin_rescale_host(gsb_ctx *ctx, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t batch, size_t L)
{
    const size_t w = ctx ? ctx->N : 0;
    return host_call(ctx, a, 2 * w * L, b, 2 * w * L, out, L ? 2 * w * (L - 1) : 0, batch,
                     [&](gsb_ws *ws, uint64_t *da, uint64_t *db, uint64_t *dout, size_t B, void *st) {
                         return gsb_mul_relin_rescale_dev(ctx, ws, da, db, dout, B, L, st);
                     });
}

} // extern "C"
