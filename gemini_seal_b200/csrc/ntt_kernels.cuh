// __global__ wrappers around ntt_core.cuh and the per-N launch plans.
//
//   N <= 2^11 : one pass, the whole limb is one tile (8T = N words).
//   N >= 2^12 : two passes (GSB_TWO_PASS_MIN) through L2/HBM.  Forward = column pass (stages with gaps >= N >> A on
//               2^A x 2^C strided tiles) then contiguous pass (remaining stages on (8 << GSB_CONTIG_T)-word
//               tiles, 1024 words by default); inverse = contiguous pass then column pass.
// grid = (tiles per limb, limb slots, polys); the IO policy turns (y, z, coefficient) into memory.
#pragma once
#include "ntt_core.cuh"

#ifndef GSB_PREFETCH
#define GSB_PREFETCH 1
#endif
#ifndef GSB_TILE_MIN_CTAS
#define GSB_TILE_MIN_CTAS 2
#endif

namespace gsb
{
    enum : int
    {
        OUT_RAW = 0,        // lazy result: [0,4q) forward, [0,2q) inverse
        OUT_FWD_STRICT = 1, // util/ntt.h:232-245
        OUT_INV_STRICT = 2  // util/ntt.h:328-333
    };

    template <int MODE>
    GSB_DEV uint64_t finish(uint64_t v, const LimbC &c)
    {
        if (MODE == OUT_FWD_STRICT)
        {
            v -= (v >= c.two_q) ? c.two_q : 0;
            v -= (v >= c.q) ? c.q : 0;
        }
        else if (MODE == OUT_INV_STRICT)
        {
            v -= (v >= c.q) ? c.q : 0;
        }
        return v;
    }

    // Plain limb-major IO: element (z, y, d) at base[z*zs + y*N + d] (zs = word stride between
    // polynomials; separate strides for source and destination allow out-of-place transforms).
    struct PlainPro
    {
        const uint64_t *p;
        GSB_DEV PlainPro(const uint64_t *base, size_t zs, int y, int z, uint32_t N)
        {
            p = base + static_cast<size_t>(z) * zs + static_cast<size_t>(y) * N;
            GSB_PIN_GLOBAL(p);
        }
        // raw / raw2: the memory loads only; xform: the prologue arithmetic on a loaded word
        GSB_DEV uint64_t raw(uint32_t d) const { return p[d]; }
        GSB_DEV void raw2(uint32_t d, uint64_t &a, uint64_t &b) const
        {
            ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(p + d);
            a = v.x;
            b = v.y;
        }
        GSB_DEV uint64_t xform(uint64_t v) const { return v; }
    };

    template <int MODE>
    struct PlainEpi
    {
        uint64_t *p;
        const LimbC &c;
        GSB_DEV PlainEpi(uint64_t *base, size_t zs, int y, int z, const LimbC &c_, uint32_t N)
            : c(c_)
        {
            p = base + static_cast<size_t>(z) * zs + static_cast<size_t>(y) * N;
            GSB_PIN_GLOBAL(p);
        }
        GSB_DEV void store(uint32_t d, uint64_t v, int) const { p[d] = finish<MODE>(v, c); }
        GSB_DEV void store2(uint32_t d, uint64_t a, uint64_t b, int) const
        {
            *reinterpret_cast<ulonglong2 *>(p + d) = make_ulonglong2(finish<MODE>(a, c), finish<MODE>(b, c));
        }
    };

    template <int MODE>
    struct IoPlain
    {
        const uint64_t *src;
        uint64_t *dst;
        size_t src_zs, dst_zs;
        GSB_DEV bool skip(int, int) const { return false; }
        struct Pro : PlainPro
        {
            GSB_DEV Pro(const IoPlain &io, int y, int z, int, const LimbC &, uint32_t N)
                : PlainPro(io.src, io.src_zs, y, z, N)
            {}
        };
        struct Epi : PlainEpi<MODE>
        {
            GSB_DEV Epi(const IoPlain &io, int y, int z, int, const LimbC &c_, uint32_t N)
                : PlainEpi<MODE>(io.dst, io.dst_zs, y, z, c_, N)
            {}
        };
    };

    // ---- the tile kernel ----------------------------------------------------------------------
    // One CTA owns tile blockIdx.x of limb slot y = blockIdx.y and walks `zb` consecutive polynomials
    // z (same modulus, same twiddles):
    //   * the tile's twiddles are staged ONCE into shared memory by TMA bulk copies (one per stage
    //     for a contiguous tile, one for a column tile) and reused for every z;
    //   * the global loads of polynomial z+1 are issued before the butterflies of polynomial z run,
    //     so HBM/L2 latency hides behind integer work.
    // A = 0: contiguous tile (ContigMap);  A > 0: column tile with 2^A rows (ColMap).
    template <int T_LOG, int G_HI, int A>
    struct TilePlan
    {
        static constexpr int TILE_LOG = T_LOG + 3;
        static constexpr bool COLS = A > 0;
        typedef TwContigSmem<TILE_LOG, G_HI> TwC; // every stage of the tile staged
        typedef TwColsSmem<(A > 0 ? A : 1), TILE_LOG - (A > 0 ? A : 1)> TwR;
        static constexpr uint32_t PAIRS = COLS ? TwR::PAIRS : TwC::PAIRS;
        static constexpr size_t exch_words = size_t(2) << TILE_LOG;
        static constexpr size_t smem_bytes = exch_words * sizeof(uint64_t) + PAIRS * sizeof(Twiddle) + 16;
    };

    // FAST: wide-range butterflies, see fwd_round / inv_round.
#ifndef GSB_TILE_FWD_FORM
#define GSB_TILE_FWD_FORM 7 // which sums of the wide-range forward butterfly are forced onto the ALU pipe (fwd_round)
#endif
    // Per-IO tuning: resident CTAs per SM the register allocation aims at (per 256 threads), whether polynomial z+1 is
    // loaded into registers before the butterflies of z (16 more live registers), and the butterfly form.  An IO type
    // may specialise it; none does at present (with 256-thread CTAs the mod-up column pass was faster at 3 CTAs of 80
    // registers without prefetch, 1.76 vs 1.82 ms; with 128-thread CTAs the default is faster again, 1.73 ms).
    // min_ctas is quoted for 256-thread CTAs; smaller CTAs keep the same number of resident threads
    __host__ __device__ constexpr int tile_min_ctas(int t_log, int per256)
    {
        return t_log > 8 ? 1 : t_log == 8 ? per256 : t_log == 7 ? 2 * per256 : 4 * per256;
    }
    template <class IO>
    struct TileTuning
    {
        static constexpr int min_ctas = GSB_TILE_MIN_CTAS;
        static constexpr bool prefetch = GSB_PREFETCH != 0;
        static constexpr int fwd_form = GSB_TILE_FWD_FORM; // fwd_round's FAST form of the wide-range forward transforms
    };

    template <bool INVERSE, int T_LOG, int G_HI, int G_LO, bool LAST, int A, class IO, bool FAST = false>
    __global__ void __launch_bounds__(1 << T_LOG, tile_min_ctas(T_LOG, TileTuning<IO>::min_ctas)) ntt_tile_kernel(IO io, NttTables tab, ModMap mm, int nz, int zb)
    {
        typedef TilePlan<T_LOG, G_HI, A> TP;
        const int y = blockIdx.y, mod = mm.mod(y);
        const uint32_t tid = threadIdx.x, tile = blockIdx.x, N = tab.N;
        const int z0 = blockIdx.z * zb, z1 = z0 + zb < nz ? z0 + zb : nz;
        uint64_t *smem = dyn_smem();
        Twiddle *tws = reinterpret_cast<Twiddle *>(smem + TP::exch_words);
        uint64_t *bar = reinterpret_cast<uint64_t *>(tws + TP::PAIRS);
        const Twiddle *table = (INVERSE ? tab.VV : tab.WW) + static_cast<size_t>(mod) * N;
        const LimbC c = tab.limb[mod];
        if (tid == 0)
            mbar_init(bar);
        __syncthreads();
        uint64_t xn[8];
        int z = z0;
        while (z < z1 && io.skip(y, z))
            z++;
        auto body = [&](const auto &map, const auto &tw) {
            if (z < z1)
            {
                const typename IO::Pro pro(io, y, z, mod, c, N);
                xf_load_raw<T_LOG, G_HI, G_LO, INVERSE>(xn, tid, tile, map, pro);
            }
            mbar_wait(bar, 0); // twiddles have landed (every thread waits, also before an early exit)
            if constexpr (FAST && !INVERSE)
            {
                // wide-range stages want w' >> 1 (mulmod_wide_add): shift the staged Shoup halves once per
                // CTA.  The literal last stage of a contiguous transform (LAST: gap 2^0, segment 0 = the
                // first TILE/2 pairs) keeps its exact w'.
                constexpr uint32_t first = (LAST && !TP::COLS && G_LO == 0) ? (1u << (TP::TILE_LOG - 1)) : 0u;
                for (uint32_t i = first + tid; i < TP::PAIRS; i += (1u << T_LOG))
                    tws[i].y >>= 1;
                __syncthreads();
            }
            int parity = 0;
            // One polynomial: `xc` holds its raw words, `xo` receives the next polynomial's.  The loop below calls it
            // with the two buffers swapped every other time, so the prefetched words are used where they landed
            // (a single buffer costs 16 register moves per polynomial on the multiplier pipe).
            auto one = [&](uint64_t (&xc)[8], uint64_t (&xo)[8]) {
                uint64_t x[8];
                {
                    const typename IO::Pro pro(io, y, z, mod, c, N);
#pragma unroll
                    for (int i = 0; i < 8; i++)
                        x[i] = pro.xform(xc[i]);
                }
                int zn = z + 1;
                while (zn < z1 && io.skip(y, zn))
                    zn++;
                if constexpr (TileTuning<IO>::prefetch)
                {
                    if (zn < z1)
                    {
                        const typename IO::Pro pro(io, y, zn, mod, c, N);
                        xf_load_raw<T_LOG, G_HI, G_LO, INVERSE>(xo, tid, tile, map, pro);
                    }
                }
                const typename IO::Epi epi(io, y, z, mod, c, N);
                if constexpr (has_prefetch<typename IO::Epi>::value)
                    xf_prefetch<T_LOG, G_HI, G_LO, INVERSE>(tid, tile, map, epi);
                if constexpr (INVERSE)
                    inv_compute<T_LOG, G_HI, G_LO, LAST, 0, FAST>(x, smem, tid, tw, c, parity);
                else if constexpr (FAST)
                    fwd_compute<T_LOG, G_HI, G_LO, LAST, 0, TileTuning<IO>::fwd_form>(x, smem, tid, TwPreshifted<std::decay_t<decltype(tw)>>{ tw }, c, parity);
                else
                    fwd_compute<T_LOG, G_HI, G_LO, LAST, 0, false>(x, smem, tid, tw, c, parity);
                xf_store<T_LOG, G_HI, G_LO, INVERSE>(x, tid, tile, map, epi);
                if constexpr (!TileTuning<IO>::prefetch)
                {
                    if (zn < z1)
                    {
                        const typename IO::Pro pro(io, y, zn, mod, c, N);
                        xf_load_raw<T_LOG, G_HI, G_LO, INVERSE>(xo, tid, tile, map, pro);
                    }
                }
                z = zn;
            };
            uint64_t xm[8];
            while (z < z1)
            {
                one(xn, xm);
                if (z >= z1)
                    break;
                one(xm, xn);
            }
        };
        if constexpr (TP::COLS)
        {
            if (tid == 0)
                TP::TwR::stage(tws, table, N, INVERSE, bar);
            const ColMap<T_LOG, A> map{ N, tab.logn };
            const typename TP::TwR tw{ tws };
            body(map, tw);
        }
        else
        {
            const ContigMap<T_LOG> map{ N };
            if (tid == 0)
                TP::TwC::stage(tws, table, N, map.off(tile), INVERSE, bar);
            const typename TP::TwC tw{ tws, TwGlobal{ table, N, map.off(tile), N } };
            body(map, tw);
        }
    }

    // ---- launch plans ------------------------------------------------------------------------
    template <class K>
    inline void allow_smem(K kfn, size_t bytes)
    {
        if (bytes > 48 * 1024)
            cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes));
    }

    // Two-pass split for a given log2(N): A = log2(rows) of the column pass, TC = T_LOG of the
    // column pass.  The contiguous pass uses (8 << GSB_CONTIG_T)-word tiles.  A is chosen so the
    // two passes need as few register rounds (<= 3 stages each) as possible.
    // log2(threads) of the CTAs of a two-pass transform (tile = 8 << T words).  128-thread CTAs (twice as many resident)
    // measured 3-8 % faster than 256-thread ones in every pass: a barrier or an mbarrier wait holds 4 warps, not 8
    // (step 6.53 -> 6.44 ms; 64-thread CTAs gain nothing more).
#ifndef GSB_COLS_T
#define GSB_COLS_T 7
#endif
#ifndef GSB_TWO_PASS_MIN
#define GSB_TWO_PASS_MIN 12 // smallest log2 N that runs as two passes (smaller limbs are one tile of N words); 12: +5 % strict NTT, +9 % BFV multiply at N = 4096 over the single 512-thread CTA per limb
#endif
#ifndef GSB_CONTIG_T
#define GSB_CONTIG_T 7
#endif
    template <int LOGN>
    struct Split
    {
        static constexpr bool two_pass = LOGN >= GSB_TWO_PASS_MIN;
        static constexpr int A = LOGN <= 12 ? 3 : LOGN == 13 ? 6 : LOGN == 14 ? 6 : LOGN == 15 ? 6 : 8;
        static constexpr int TC = LOGN == 16 ? 9 : GSB_COLS_T;
        static constexpr int TK = GSB_CONTIG_T; // T_LOG of the contiguous pass of a two-pass transform
        static constexpr int T1 = LOGN - 3; // single-pass T_LOG
    };

    // SMs of the current device (the launching thread's current device is the context's: every entry point sets it)
    inline size_t sm_count()
    {
#ifdef GSB_CUSIM
        return 148;
#else
        static thread_local int cached_dev = -1;
        static thread_local int cached_sms = 148;
        int dev = 0;
        if (cudaGetDevice(&dev) == cudaSuccess && dev != cached_dev)
        {
            int n = 0;
            if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
                cached_sms = n;
            cached_dev = dev;
        }
        return static_cast<size_t>(cached_sms);
#endif
    }

#ifndef GSB_ZB_MAX
#define GSB_ZB_MAX 8
#endif
    // polynomials per CTA: enough CTAs left for ~8 waves of (SMs x 3 resident CTAs)
    inline int pick_zb(size_t ctas_per_z, int nz)
    {
        const size_t want = sm_count() * 3 * 8;
        size_t zb = (ctas_per_z * static_cast<size_t>(nz)) / want;
        zb = zb < 1 ? 1 : zb > GSB_ZB_MAX ? GSB_ZB_MAX : zb;
        return static_cast<int>(zb);
    }

    template <bool INVERSE, int T_LOG, int G_HI, int G_LO, bool LAST, int A, bool FAST = false, class IO>
    inline void launch_tile(
        const char *pass, const char *tag, const IO &io, const NttTables &tab, const ModMap &mm, unsigned tiles, int ny, int nz,
        cudaStream_t st)
    {
        typedef TilePlan<T_LOG, G_HI, A> TP;
        auto kfn = ntt_tile_kernel<INVERSE, T_LOG, G_HI, G_LO, LAST, A, IO, FAST>;
        allow_smem(kfn, TP::smem_bytes);
        const int zb = pick_zb(static_cast<size_t>(tiles) * ny, nz);
        dim3 grid(tiles, ny, (nz + zb - 1) / zb);
        GSB_LAUNCH(pass, tag, kfn, grid, dim3(1 << T_LOG), TP::smem_bytes, st, io, tab, mm, nz, zb);
    }

    // Forward pass 1 (columns) with arbitrary IO.  tiles = N / 2^C column groups.
    template <int LOGN, bool FAST = false, class IO>
    inline void launch_fwd_cols(
        const IO &io, const NttTables &tab, const ModMap &mm, int ny, int nz, cudaStream_t st, const char *tag = "plain")
    {
        using SP = Split<LOGN>;
        constexpr int A = SP::A, TC = SP::TC, C = TC + 3 - A;
        launch_tile<false, TC, A + C - 1, C, false, A, FAST>("fwd_cols", tag, io, tab, mm, (1u << (LOGN - A)) >> C, ny, nz, st);
    }

    // Forward pass 2 (contiguous blocks of N >> A) -- or the whole transform when !two_pass.
    template <int LOGN, bool FAST = false, class IO>
    inline void launch_fwd_contig(
        const IO &io, const NttTables &tab, const ModMap &mm, int ny, int nz, cudaStream_t st, const char *tag = "plain")
    {
        using SP = Split<LOGN>;
        if constexpr (SP::two_pass)
            launch_tile<false, SP::TK, LOGN - SP::A - 1, 0, true, 0, FAST>("fwd_contig", tag, io, tab, mm, 1u << (LOGN - SP::TK - 3), ny, nz, st);
        else
            launch_tile<false, SP::T1, LOGN - 1, 0, true, 0, FAST>("fwd_contig", tag, io, tab, mm, 1u, ny, nz, st);
    }

    // Inverse pass 1 (contiguous) -- or the whole transform when !two_pass.
    template <int LOGN, bool FAST = false, class IO>
    inline void launch_inv_contig(
        const IO &io, const NttTables &tab, const ModMap &mm, int ny, int nz, cudaStream_t st, const char *tag = "plain")
    {
        using SP = Split<LOGN>;
        if constexpr (SP::two_pass)
            launch_tile<true, SP::TK, LOGN - SP::A - 1, 0, false, 0, FAST>("inv_contig", tag, io, tab, mm, 1u << (LOGN - SP::TK - 3), ny, nz, st);
        else
            launch_tile<true, SP::T1, LOGN - 1, 0, true, 0, FAST>("inv_contig", tag, io, tab, mm, 1u, ny, nz, st);
    }

    // Inverse pass 2 (columns; carries the global last stage with n^-1).
    template <int LOGN, bool FAST = false, class IO>
    inline void launch_inv_cols(
        const IO &io, const NttTables &tab, const ModMap &mm, int ny, int nz, cudaStream_t st, const char *tag = "plain")
    {
        using SP = Split<LOGN>;
        constexpr int A = SP::A, TC = SP::TC, C = TC + 3 - A;
        launch_tile<true, TC, A + C - 1, C, true, A, FAST>("inv_cols", tag, io, tab, mm, (1u << (LOGN - A)) >> C, ny, nz, st);
    }

    // Plain batched transforms: element (z, y, d) of src/dst at base + z*zs + y*N + d; in place
    // allowed (src == dst with equal strides).
    // wide: run the wide-range forward butterflies (fwd_round FAST) in every stage but the global last one.
    // Callers pass it only under gsb_ctx::fast_ntt, for canonical inputs (< q) and only when the output is
    // strict (canonicalised by finish<OUT_FWD_STRICT>, hence the same words as the literal transform).
    template <int LOGN, bool FAST>
    inline void ntt_forward_plain_t(
        const NttTables &tab, const ModMap &mm, const uint64_t *src, size_t src_zs, uint64_t *dst, size_t dst_zs, int ny,
        int nz, bool lazy, cudaStream_t st, const char *tag)
    {
        if constexpr (Split<LOGN>::two_pass)
        {
            launch_fwd_cols<LOGN, FAST>(IoPlain<OUT_RAW>{ src, dst, src_zs, dst_zs }, tab, mm, ny, nz, st, tag);
            src = dst;
            src_zs = dst_zs;
        }
        if (lazy)
            launch_fwd_contig<LOGN, FAST>(IoPlain<OUT_RAW>{ src, dst, src_zs, dst_zs }, tab, mm, ny, nz, st, tag);
        else
            launch_fwd_contig<LOGN, FAST>(IoPlain<OUT_FWD_STRICT>{ src, dst, src_zs, dst_zs }, tab, mm, ny, nz, st, tag);
    }

    template <int LOGN>
    inline void ntt_forward_plain(
        const NttTables &tab, const ModMap &mm, const uint64_t *src, size_t src_zs, uint64_t *dst, size_t dst_zs, int ny,
        int nz, bool lazy, cudaStream_t st, const char *tag = "plain", bool wide = false)
    {
        if (wide && !lazy)
            ntt_forward_plain_t<LOGN, true>(tab, mm, src, src_zs, dst, dst_zs, ny, nz, lazy, st, tag);
        else
            ntt_forward_plain_t<LOGN, false>(tab, mm, src, src_zs, dst, dst_zs, ny, nz, lazy, st, tag);
    }

    // wide: run the wide-range inverse butterflies (inv_round FAST).  Callers pass it only under
    // gsb_ctx::wide_inv and only when the output is strict (canonical) or stays inside the library.
    template <int LOGN, bool FAST>
    inline void ntt_inverse_plain_t(
        const NttTables &tab, const ModMap &mm, const uint64_t *src, size_t src_zs, uint64_t *dst, size_t dst_zs, int ny,
        int nz, bool lazy, cudaStream_t st, const char *tag)
    {
        if constexpr (Split<LOGN>::two_pass)
        {
            launch_inv_contig<LOGN, FAST>(IoPlain<OUT_RAW>{ src, dst, src_zs, dst_zs }, tab, mm, ny, nz, st, tag);
            if (lazy)
                launch_inv_cols<LOGN, FAST>(IoPlain<OUT_RAW>{ dst, dst, dst_zs, dst_zs }, tab, mm, ny, nz, st, tag);
            else
                launch_inv_cols<LOGN, FAST>(IoPlain<OUT_INV_STRICT>{ dst, dst, dst_zs, dst_zs }, tab, mm, ny, nz, st, tag);
        }
        else
        {
            if (lazy)
                launch_inv_contig<LOGN, FAST>(IoPlain<OUT_RAW>{ src, dst, src_zs, dst_zs }, tab, mm, ny, nz, st, tag);
            else
                launch_inv_contig<LOGN, FAST>(IoPlain<OUT_INV_STRICT>{ src, dst, src_zs, dst_zs }, tab, mm, ny, nz, st, tag);
        }
    }

    template <int LOGN>
    inline void ntt_inverse_plain(
        const NttTables &tab, const ModMap &mm, const uint64_t *src, size_t src_zs, uint64_t *dst, size_t dst_zs, int ny,
        int nz, bool lazy, cudaStream_t st, const char *tag = "plain", bool wide = false)
    {
        if (wide)
            ntt_inverse_plain_t<LOGN, true>(tab, mm, src, src_zs, dst, dst_zs, ny, nz, lazy, st, tag);
        else
            ntt_inverse_plain_t<LOGN, false>(tab, mm, src, src_zs, dst, dst_zs, ny, nz, lazy, st, tag);
    }

// Dispatch a templated call on the runtime log2(N) (10..16 supported on the device path).
#define GSB_DISPATCH_LOGN(logn, CALL)                       \
    switch (logn)                                           \
    {                                                       \
    case 10: { constexpr int LOGN = 10; CALL; } break;      \
    case 11: { constexpr int LOGN = 11; CALL; } break;      \
    case 12: { constexpr int LOGN = 12; CALL; } break;      \
    case 13: { constexpr int LOGN = 13; CALL; } break;      \
    case 14: { constexpr int LOGN = 14; CALL; } break;      \
    case 15: { constexpr int LOGN = 15; CALL; } break;      \
    case 16: { constexpr int LOGN = 16; CALL; } break;      \
    default: break;                                         \
    }
} // namespace gsb
