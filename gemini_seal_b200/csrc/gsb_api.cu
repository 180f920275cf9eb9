// C-ABI implementation (include/gseal_b200.h): context, workspaces, NTT and elementwise entry
// points.  Key switching / multiply / rescale / Galois live in gsb_evaluator.cuh, included below.
#include "gsb_context.cuh"
#include "elementwise.cuh"
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <memory>

namespace gsb
{
    static thread_local std::string g_last_error;
    void set_error(const std::string &msg)
    {
        g_last_error = msg;
    }
    unsigned long long &launch_counter()
    {
        static unsigned long long n = 0;
        return n;
    }

    // ---- optional per-launch CUDA-event timing (gsb_prof_*) -------------------------------------
    struct ProfRec
    {
        std::string label;
        cudaEvent_t e0, e1;
    };
    static bool g_prof = false;
    static std::mutex g_prof_mu;
    static std::vector<ProfRec> g_prof_recs;
    bool prof_on()
    {
        return g_prof;
    }
    void prof_begin(const char *pass, const char *tag, cudaStream_t st)
    {
        ProfRec r;
        r.label = std::string(pass) + (tag && *tag ? std::string("/") + tag : std::string());
        cudaEventCreate(&r.e0);
        cudaEventCreate(&r.e1);
        cudaEventRecord(r.e0, st);
        std::lock_guard<std::mutex> lock(g_prof_mu);
        g_prof_recs.push_back(r);
    }
    void prof_end(cudaStream_t st)
    {
        std::lock_guard<std::mutex> lock(g_prof_mu);
        cudaEventRecord(g_prof_recs.back().e1, st);
    }
    static void prof_clear()
    {
        std::lock_guard<std::mutex> lock(g_prof_mu);
        for (auto &r : g_prof_recs)
        {
            cudaEventDestroy(r.e0);
            cudaEventDestroy(r.e1);
        }
        g_prof_recs.clear();
    }

    template <class F>
    static int guarded(F &&f)
    {
        try
        {
            f();
            return GSB_OK;
        }
        catch (const ApiError &e)
        {
            set_error(e.msg);
            return e.code;
        }
        catch (const std::bad_alloc &)
        {
            set_error("out of host memory");
            return GSB_E_NOMEM;
        }
        catch (const std::exception &e)
        {
            set_error(e.what());
            return GSB_E_LOGIC;
        }
    }

    template <class T>
    static T *upload(const std::vector<T> &h)
    {
        T *d = nullptr;
        GSB_CUDA_CHECK(cudaMalloc(reinterpret_cast<void **>(&d), h.size() * sizeof(T)));
        GSB_CUDA_CHECK(cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
        return d;
    }

    static void check_limbs(const gsb_ctx *ctx, size_t n_polys, size_t n_limbs, size_t limb_offset)
    {
        require(ctx != nullptr, GSB_E_INVALID_ARG, "ctx is null");
        require(n_limbs >= 1 && limb_offset + n_limbs <= ctx->K, GSB_E_INVALID_ARG, "limb range exceeds the key-level modulus");
        require(n_polys >= 1 && n_polys <= 65535 * 64, GSB_E_INVALID_ARG, "n_polys out of range");
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device)); // the calling thread's current device may be another GPU
    }
} // namespace gsb

using namespace gsb;

void *gsb_ws::reserve(size_t bytes)
{
    if (bytes > cap)
    {
        if (buf)
            GSB_CUDA_CHECK(cudaFree(buf));
        buf = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 8;
        GSB_CUDA_CHECK(cudaMalloc(&buf, want));
        cap = want;
    }
    return buf;
}

void *gsb_ws::reserve_pinned(size_t bytes)
{
    if (bytes > pinned_cap)
    {
        if (pinned)
            GSB_CUDA_CHECK(cudaFreeHost(pinned));
        pinned = nullptr;
        pinned_cap = 0;
        GSB_CUDA_CHECK(cudaMallocHost(&pinned, bytes));
        pinned_cap = bytes;
    }
    return pinned;
}

extern "C" {

const char *gsb_last_error(void)
{
    return g_last_error.c_str();
}

int gsb_version(void)
{
    return 100;
}

unsigned long long gsb_launch_count(void)
{
    return launch_counter();
}

int gsb_prof_enable(int on)
{
    prof_clear();
    g_prof = on != 0;
    return GSB_OK;
}

long gsb_prof_report(char *buf, size_t cap)
{
    long need = -1;
    guarded([&] {
        GSB_CUDA_CHECK(cudaDeviceSynchronize());
        std::map<std::string, std::pair<unsigned long long, double>> agg;
        std::vector<std::string> order;
        {
            std::lock_guard<std::mutex> lock(g_prof_mu);
            for (auto &r : g_prof_recs)
            {
                float ms = 0.f;
                GSB_CUDA_CHECK(cudaEventElapsedTime(&ms, r.e0, r.e1));
                if (!agg.count(r.label))
                    order.push_back(r.label);
                agg[r.label].first++;
                agg[r.label].second += ms;
            }
        }
        std::string js = "{";
        for (size_t i = 0; i < order.size(); i++)
        {
            char tmp[256];
            snprintf(tmp, sizeof(tmp), "%s\"%s\": {\"launches\": %llu, \"ms\": %.6f}", i ? ", " : "", order[i].c_str(),
                     agg[order[i]].first, agg[order[i]].second);
            js += tmp;
        }
        js += "}";
        need = static_cast<long>(js.size()) + 1;
        if (buf && cap >= js.size() + 1)
            memcpy(buf, js.c_str(), js.size() + 1);
    });
    return need;
}

int gsb_ctx_create(int device, int logn, const uint64_t *moduli, size_t K, size_t k_special, gsb_ctx **out)
{
    return guarded([&] {
        require(out != nullptr && moduli != nullptr, GSB_E_INVALID_ARG, "null argument");
        require(logn >= 10 && logn <= 16, GSB_E_INVALID_ARG, "poly_modulus_degree must be 2^10..2^16 on the device path");
        require(K >= 1 && K <= 64, GSB_E_INVALID_ARG, "coeff_modulus size out of range (1..64)");
        // encryptionparams.h:205-214 / context.cpp:524-530
        require(k_special >= 1 && (K == 1 || k_special < K), GSB_E_LOGIC, "n_special_primes must be in [1, #moduli)");
        int prev_device = 0;
        GSB_CUDA_CHECK(cudaGetDevice(&prev_device));
        GSB_CUDA_CHECK(cudaSetDevice(device));
        struct Restore // the creating thread keeps its current device
        {
            int d;
            ~Restore() { cudaSetDevice(d); }
        } restore{ prev_device };
        std::unique_ptr<gsb_ctx> ctx(new gsb_ctx());
        ctx->device = device;
        ctx->logn = logn;
        ctx->N = size_t(1) << logn;
        ctx->K = K;
        ctx->k = k_special;
        ctx->L0 = K > k_special ? K - k_special : K;
        ctx->tables.resize(K);
        std::vector<Twiddle> ww(K * ctx->N), vv(K * ctx->N);
        std::vector<LimbC> limb(K);
        for (size_t i = 0; i < K; i++)
        {
            for (size_t j = 0; j < i; j++)
                require(moduli[i] != moduli[j], GSB_E_INVALID_ARG, "coeff_modulus primes must be distinct");
            // SEAL_MOD_BIT_COUNT_MAX = 61 (util/defines.h:33)
            require(moduli[i] > 1 && (moduli[i] >> 61) == 0, GSB_E_INVALID_ARG, "modulus out of range");
            require(host::build_limb_tables(logn, moduli[i], ctx->tables[i]), GSB_E_INVALID_ARG,
                    "invalid modulus: no primitive 2N-th root (NTTTables, util/ntt.cpp:56-59)");
            const host::LimbTables &t = ctx->tables[i];
            for (size_t d = 0; d < ctx->N; d++)
            {
                // device layout: small-gap stages swizzled against bank conflicts (ntt_core.cuh)
                const uint32_t n32 = static_cast<uint32_t>(ctx->N), d32 = static_cast<uint32_t>(d);
                ww[i * ctx->N + fwd_tw_pos(d32, n32)] = make_ulonglong2(t.w[d], t.ws[d]);
                vv[i * ctx->N + inv_tw_pos(d32, n32)] = make_ulonglong2(t.v[d], t.vs[d]);
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
        {
            typedef unsigned __int128 u128;
            uint64_t qmax = 0;
            for (size_t i = 0; i < K; i++)
                qmax = std::max(qmax, moduli[i]);
            // wide-range butterflies grow a word by < 5q per stage from an input < 2q (see fwd_round)
            const u128 growth = 5 * static_cast<u128>(logn) + 2;
            const u128 beta0 = (ctx->L0 + ctx->k - 1) / ctx->k;
            ctx->fast_ntt = static_cast<u128>(qmax) * growth < (static_cast<u128>(1) << 63);
            ctx->lazy_modup = ctx->fast_ntt;
            ctx->fast_fwd_public = static_cast<u128>(qmax) * (growth + 2) < (static_cast<u128>(1) << 63);
            ctx->lazy_mac = ctx->fast_ntt && static_cast<u128>(qmax) * qmax <= ~static_cast<u128>(0) / (beta0 * growth);
            // GSB_LITERAL=1 (read at context creation; a testing aid): every congruent-word shortcut off, the
            // kernels carry the reference's own intermediate words.  tests/test_gpu_parity.py runs the full
            // size both ways and requires identical outputs.
            ctx->wide_inv = static_cast<u128>(qmax) * 10 < (static_cast<u128>(1) << 63);
            const char *split = std::getenv("GSB_SPLIT_TAIL");
            ctx->split_tail = split && split[0] == '1';
            const char *lit = std::getenv("GSB_LITERAL");
            if (lit && lit[0] == '1')
            {
                ctx->fast_ntt = ctx->fast_fwd_public = ctx->lazy_modup = ctx->lazy_mac = ctx->wide_inv = false;
                for (auto &c : limb)
                    c.mu = 0;
            }
        }
        ctx->d_WW = upload(ww);
        ctx->d_VV = upload(vv);
        ctx->d_limb = upload(limb);
        ctx->default_ws.ctx = ctx.get();
        ctx->host_io.ctx = ctx.get();
        *out = ctx.release();
    });
}

void gsb_ctx_destroy(gsb_ctx *ctx)
{
    if (!ctx)
        return;
    cudaSetDevice(ctx->device);
    cudaFree(ctx->d_WW);
    cudaFree(ctx->d_VV);
    cudaFree(ctx->d_limb);
    for (auto &kv : ctx->level_consts)
    {
        cudaFree(kv.second.d_modup_a);
        cudaFree(kv.second.d_modup_b);
        cudaFree(kv.second.d_inv_hat);
        cudaFree(kv.second.d_neg_hat);
        cudaFree(kv.second.d_inv_p);
        cudaFree(kv.second.d_neg_p);
        cudaFree(kv.second.d_inv_qlast);
        cudaFree(kv.second.d_neg_half);
    }
    for (auto &kv : ctx->relin)
        cudaFree(kv.second.d);
    for (auto &kv : ctx->galois)
        cudaFree(kv.second.d);
    for (auto &kv : ctx->galois_tables)
        cudaFree(kv.second);
    for (auto &kv : ctx->bfv_tools)
    {
        cudaFree(kv.second.d_WW);
        cudaFree(kv.second.d_VV);
        cudaFree(kv.second.d_limb);
        cudaFree(kv.second.d_consts);
    }
    // secret-key material is wiped before its memory goes back to the allocator (the reference keeps
    // SecretKey in a clear-on-destruction pool, secretkey.h:48-57); retired arrays hold at most as many
    // powers as the current one
    if (ctx->d_sk)
        cudaMemset(ctx->d_sk, 0, ctx->sk_powers * ctx->K * ctx->N * sizeof(uint64_t));
    cudaFree(ctx->d_sk);
    for (auto &r : ctx->sk_retired)
    {
        cudaMemset(r.first, 0, r.second);
        cudaFree(r.first);
    }
    cudaFree(ctx->default_ws.buf);
    cudaFree(ctx->host_io.buf);
    if (ctx->pipe_ready)
    {
        for (int i = 0; i < 3; i++)
            cudaStreamDestroy(ctx->pipe_stream[i]);
        for (int i = 0; i < 6; i++)
            cudaEventDestroy(ctx->pipe_event[i]);
    }
    if (ctx->default_ws.pinned)
        cudaFreeHost(ctx->default_ws.pinned);
    delete ctx;
}

int gsb_ctx_get_tables(
    const gsb_ctx *ctx, size_t mod_index, uint64_t *w, uint64_t *ws, uint64_t *v, uint64_t *vs, uint64_t *scalars)
{
    return guarded([&] {
        require(ctx != nullptr, GSB_E_INVALID_ARG, "ctx is null");
        require(mod_index < ctx->K, GSB_E_RANGE, "mod_index out of range");
        const host::LimbTables &t = ctx->tables[mod_index];
        const size_t bytes = ctx->N * sizeof(uint64_t);
        if (w)
            memcpy(w, t.w.data(), bytes);
        if (ws)
            memcpy(ws, t.ws.data(), bytes);
        if (v)
            memcpy(v, t.v.data(), bytes);
        if (vs)
            memcpy(vs, t.vs.data(), bytes);
        if (scalars)
        {
            scalars[0] = t.mod.q;
            scalars[1] = t.mod.cr0;
            scalars[2] = t.mod.cr1;
            scalars[3] = t.root;
            scalars[4] = t.inv_n;
            scalars[5] = t.inv_n_s;
            scalars[6] = t.rdp;
        }
    });
}

int gsb_ctx_info(const gsb_ctx *ctx, int *logn, size_t *K, size_t *k_special)
{
    return guarded([&] {
        require(ctx != nullptr, GSB_E_INVALID_ARG, "ctx is null");
        if (logn)
            *logn = ctx->logn;
        if (K)
            *K = ctx->K;
        if (k_special)
            *k_special = ctx->k;
    });
}

int gsb_ws_create(gsb_ctx *ctx, gsb_ws **out)
{
    return guarded([&] {
        require(ctx != nullptr && out != nullptr, GSB_E_INVALID_ARG, "null argument");
        gsb_ws *ws = new gsb_ws();
        ws->ctx = ctx;
        *out = ws;
    });
}

void gsb_ws_destroy(gsb_ws *ws)
{
    if (!ws)
        return;
    cudaFree(ws->buf);
    if (ws->pinned)
        cudaFreeHost(ws->pinned);
    delete ws;
}

int gsb_malloc(gsb_ctx *ctx, size_t bytes, void **out)
{
    return guarded([&] {
        require(ctx != nullptr && out != nullptr, GSB_E_INVALID_ARG, "null argument");
        GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
        GSB_CUDA_CHECK(cudaMalloc(out, bytes));
    });
}

int gsb_free(gsb_ctx *ctx, void *ptr)
{
    return guarded([&] {
        require(ctx != nullptr, GSB_E_INVALID_ARG, "ctx is null");
        GSB_CUDA_CHECK(cudaFree(ptr));
    });
}

int gsb_malloc_host(size_t bytes, void **out)
{
    return guarded([&] {
        require(out != nullptr, GSB_E_INVALID_ARG, "null argument");
        GSB_CUDA_CHECK(cudaMallocHost(out, bytes));
    });
}

int gsb_free_host(void *ptr)
{
    return guarded([&] { GSB_CUDA_CHECK(cudaFreeHost(ptr)); });
}

int gsb_memcpy_h2d(gsb_ctx *ctx, void *dst, const void *src, size_t bytes, void *stream)
{
    return guarded([&] {
        require(ctx != nullptr, GSB_E_INVALID_ARG, "ctx is null");
        GSB_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream)));
    });
}

int gsb_memcpy_d2h(gsb_ctx *ctx, void *dst, const void *src, size_t bytes, void *stream)
{
    return guarded([&] {
        require(ctx != nullptr, GSB_E_INVALID_ARG, "ctx is null");
        GSB_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream)));
    });
}

int gsb_stream_sync(gsb_ctx *ctx, void *stream)
{
    return guarded([&] {
        require(ctx != nullptr, GSB_E_INVALID_ARG, "ctx is null");
        GSB_CUDA_CHECK(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    });
}

// ---- NTT -------------------------------------------------------------------------------------
static void ntt_dev(
    gsb_ctx *ctx, uint64_t *data, size_t n_polys, size_t n_limbs, size_t limb_offset, int mode, bool inverse,
    cudaStream_t st)
{
    check_limbs(ctx, n_polys, n_limbs, limb_offset);
    require(data != nullptr, GSB_E_INVALID_ARG, "data is null");
    require(mode == GSB_NTT_STRICT || mode == GSB_NTT_LAZY, GSB_E_INVALID_ARG, "mode");
    const NttTables tab = ctx->ntt_tables();
    const ModMap mm{ 1 << 30, 0, static_cast<int>(limb_offset) };
    const bool lazy = mode == GSB_NTT_LAZY;
    // gridDim.z is limited to 65535: split very large batches
    const size_t per = ctx->N * n_limbs;
    for (size_t p0 = 0; p0 < n_polys; p0 += 32768)
    {
        const int nz = static_cast<int>(n_polys - p0 < 32768 ? n_polys - p0 : 32768);
        uint64_t *p = data + p0 * per;
        if (inverse)
        {
            GSB_DISPATCH_LOGN(ctx->logn, ntt_inverse_plain<LOGN>(tab, mm, p, per, p, per, static_cast<int>(n_limbs), nz, lazy, st, "plain", !lazy && ctx->wide_inv));
        }
        else
        {
            // strict output is canonical either way: wide-range stages when no word can reach 2^63 from an input < 4q
            GSB_DISPATCH_LOGN(ctx->logn, ntt_forward_plain<LOGN>(tab, mm, p, per, p, per, static_cast<int>(n_limbs), nz, lazy, st, "plain", !lazy && ctx->fast_fwd_public));
        }
    }
    GSB_CUDA_CHECK(cudaGetLastError());
}

int gsb_ntt_forward_dev(gsb_ctx *ctx, uint64_t *data, size_t n_polys, size_t n_limbs, size_t limb_offset, int mode, void *stream)
{
    return guarded([&] { ntt_dev(ctx, data, n_polys, n_limbs, limb_offset, mode, false, static_cast<cudaStream_t>(stream)); });
}

int gsb_ntt_inverse_dev(gsb_ctx *ctx, uint64_t *data, size_t n_polys, size_t n_limbs, size_t limb_offset, int mode, void *stream)
{
    return guarded([&] { ntt_dev(ctx, data, n_polys, n_limbs, limb_offset, mode, true, static_cast<cudaStream_t>(stream)); });
}

static void ntt_host(gsb_ctx *ctx, uint64_t *data, size_t n_polys, size_t n_limbs, size_t limb_offset, int mode, bool inverse)
{
    check_limbs(ctx, n_polys, n_limbs, limb_offset);
    require(data != nullptr, GSB_E_INVALID_ARG, "data is null");
    std::lock_guard<std::mutex> lock(ctx->host_mu);
    GSB_CUDA_CHECK(cudaSetDevice(ctx->device));
    const size_t bytes = n_polys * n_limbs * ctx->N * sizeof(uint64_t);
    uint64_t *d = static_cast<uint64_t *>(ctx->host_io.reserve(bytes));
    GSB_CUDA_CHECK(cudaMemcpyAsync(d, data, bytes, cudaMemcpyHostToDevice, nullptr));
    ntt_dev(ctx, d, n_polys, n_limbs, limb_offset, mode, inverse, nullptr);
    GSB_CUDA_CHECK(cudaMemcpyAsync(data, d, bytes, cudaMemcpyDeviceToHost, nullptr));
    GSB_CUDA_CHECK(cudaStreamSynchronize(nullptr));
}

int gsb_ntt_forward_host(gsb_ctx *ctx, uint64_t *data, size_t n_polys, size_t n_limbs, size_t limb_offset, int mode)
{
    return guarded([&] { ntt_host(ctx, data, n_polys, n_limbs, limb_offset, mode, false); });
}

int gsb_ntt_inverse_host(gsb_ctx *ctx, uint64_t *data, size_t n_polys, size_t n_limbs, size_t limb_offset, int mode)
{
    return guarded([&] { ntt_host(ctx, data, n_polys, n_limbs, limb_offset, mode, true); });
}

// ---- elementwise -------------------------------------------------------------------------------
int gsb_poly_op_dev(
    gsb_ctx *ctx, int op, const uint64_t *a, const uint64_t *b, const uint64_t *scalars, uint64_t *out, size_t n_polys,
    size_t n_limbs, size_t limb_offset, void *stream)
{
    return guarded([&] {
        check_limbs(ctx, n_polys, n_limbs, limb_offset);
        require(a != nullptr && out != nullptr, GSB_E_INVALID_ARG, "null operand");
        require(op >= GSB_OP_DYADIC && op <= GSB_OP_MUL_SCALAR, GSB_E_INVALID_ARG, "op");
        require(op == GSB_OP_NEGATE || op == GSB_OP_MUL_SCALAR || b != nullptr, GSB_E_INVALID_ARG, "operand2 is null");
        require(op != GSB_OP_MUL_SCALAR || scalars != nullptr, GSB_E_INVALID_ARG, "scalars is null");
        PolyOpArgs args{};
        args.a = a;
        args.b = b;
        args.out = out;
        args.limb = ctx->d_limb;
        args.N = static_cast<uint32_t>(ctx->N);
        args.n_limbs = static_cast<uint32_t>(n_limbs);
        args.limb_offset = static_cast<uint32_t>(limb_offset);
        args.total = n_polys * n_limbs * ctx->N;
        if (op == GSB_OP_MUL_SCALAR)
        {
            require(n_limbs <= 64, GSB_E_INVALID_ARG, "n_limbs");
            for (size_t i = 0; i < n_limbs; i++)
                args.scalar[i] = scalars[i];
        }
        launch_poly_op(op, args, static_cast<cudaStream_t>(stream));
        GSB_CUDA_CHECK(cudaGetLastError());
    });
}

int gsb_ctx_fast_ntt(const gsb_ctx *ctx)
{
    return ctx != nullptr && ctx->fast_ntt ? 1 : 0;
}

int gsb_ctx_shortcuts(const gsb_ctx *ctx)
{
    if (ctx == nullptr)
        return 0;
    return (ctx->fast_ntt ? 1 : 0) | (ctx->lazy_modup ? 2 : 0) | (ctx->lazy_mac ? 4 : 0) | (ctx->wide_inv ? 8 : 0) |
           (ctx->fast_fwd_public ? 16 : 0);
}

} // extern "C"

#include "gsb_evaluator.cuh"
#include "gsb_general.cuh"
#include "gsb_keygen.cuh"
#include "gsb_noalloc.cuh"
