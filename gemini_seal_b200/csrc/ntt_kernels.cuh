// __global__ wrappers around ntt_core.cuh and the per-N launch plans.
//
//   N <= 2^12 : one pass, the whole limb is one tile (8T = N words, <= 32 KiB).
//   N >= 2^13 : two passes through L2/HBM.  Forward = column pass (stages with gaps >= N >> A on
//               2^A x 2^C strided tiles) then contiguous pass (remaining stages on 2048-word
//               tiles); inverse = contiguous pass then column pass.
// grid = (tiles per limb, limb slots, polys); the IO policy turns (y, z, coefficient) into memory.
#pragma once
#include "ntt_core.cuh"

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
            : p(base + static_cast<size_t>(z) * zs + static_cast<size_t>(y) * N)
        {}
        GSB_DEV uint64_t load(uint32_t d) const { return p[d]; }
        GSB_DEV void load2(uint32_t d, uint64_t &a, uint64_t &b) const
        {
            ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(p + d);
            a = v.x;
            b = v.y;
        }
    };

    template <int MODE>
    struct PlainEpi
    {
        uint64_t *p;
        const LimbC &c;
        GSB_DEV PlainEpi(uint64_t *base, size_t zs, int y, int z, const LimbC &c_, uint32_t N)
            : p(base + static_cast<size_t>(z) * zs + static_cast<size_t>(y) * N), c(c_)
        {}
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

    template <int T_LOG, int G_HI, int G_LO, bool LAST, class Map, class IO>
    __global__ void __launch_bounds__(1 << T_LOG) ntt_fwd_kernel(IO io, NttTables tab, ModMap mm, Map map)
    {
        const int y = blockIdx.y, z = blockIdx.z, mod = mm.mod(y);
        if (io.skip(y, z))
            return;
        const LimbC c = tab.limb[mod];
        typename IO::Pro pro(io, y, z, mod, c, tab.N);
        typename IO::Epi epi(io, y, z, mod, c, tab.N);
        uint64_t x[8];
        int parity = 0;
        fwd_rounds<T_LOG, G_HI, G_LO, LAST, 0>(
            x, dyn_smem(), threadIdx.x, blockIdx.x, map, pro, epi, tab.WW + static_cast<size_t>(mod) * tab.N, c, parity);
    }

    template <int T_LOG, int G_HI, int G_LO, bool LAST, class Map, class IO>
    __global__ void __launch_bounds__(1 << T_LOG) ntt_inv_kernel(IO io, NttTables tab, ModMap mm, Map map)
    {
        const int y = blockIdx.y, z = blockIdx.z, mod = mm.mod(y);
        if (io.skip(y, z))
            return;
        const LimbC c = tab.limb[mod];
        typename IO::Pro pro(io, y, z, mod, c, tab.N);
        typename IO::Epi epi(io, y, z, mod, c, tab.N);
        uint64_t x[8];
        int parity = 0;
        inv_rounds<T_LOG, G_HI, G_LO, LAST, 0>(
            x, dyn_smem(), threadIdx.x, blockIdx.x, map, pro, epi, tab.VV + static_cast<size_t>(mod) * tab.N, tab.N, c,
            parity);
    }

    // ---- launch plans ------------------------------------------------------------------------
    template <int T_LOG>
    constexpr size_t tile_smem_bytes()
    {
        return size_t(2) * (size_t(8) << T_LOG) * sizeof(uint64_t);
    }

    template <class K>
    inline void allow_smem(K kfn, size_t bytes)
    {
        if (bytes > 48 * 1024)
            cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes));
    }

    // Two-pass split for a given log2(N): A = log2(rows) of the column pass, TC = T_LOG of the
    // column pass.  The contiguous pass always uses 2048-word tiles (T_LOG = 8).
    template <int LOGN>
    struct Split
    {
        static constexpr bool two_pass = LOGN >= 13;
        static constexpr int A = LOGN == 13 ? 6 : LOGN == 14 ? 7 : LOGN == 15 ? 7 : 8;
        static constexpr int TC = LOGN == 16 ? 9 : 8;
        static constexpr int T1 = LOGN - 3; // single-pass T_LOG
    };

    // Forward pass 1 (columns) with arbitrary IO.  grid.x = N / 2^C column groups.
    template <int LOGN, class IO>
    inline void launch_fwd_cols(
        const IO &io, const NttTables &tab, const ModMap &mm, int ny, int nz, cudaStream_t st, const char *tag = "plain")
    {
        using SP = Split<LOGN>;
        constexpr int A = SP::A, TC = SP::TC, C = TC + 3 - A;
        ColMap<TC, A> map{ tab.N, tab.logn };
        auto kfn = ntt_fwd_kernel<TC, A + C - 1, C, false, ColMap<TC, A>, IO>;
        allow_smem(kfn, tile_smem_bytes<TC>());
        dim3 grid((1u << (LOGN - A)) >> C, ny, nz);
        GSB_LAUNCH("fwd_cols", tag, kfn, grid, dim3(1 << TC), tile_smem_bytes<TC>(), st, io, tab, mm, map);
    }

    // Forward pass 2 (contiguous blocks of N >> A) -- or the whole transform when !two_pass.
    template <int LOGN, class IO>
    inline void launch_fwd_contig(
        const IO &io, const NttTables &tab, const ModMap &mm, int ny, int nz, cudaStream_t st, const char *tag = "plain")
    {
        using SP = Split<LOGN>;
        if constexpr (SP::two_pass)
        {
            constexpr int T = 8;
            ContigMap<T> map{ tab.N };
            auto kfn = ntt_fwd_kernel<T, LOGN - SP::A - 1, 0, true, ContigMap<T>, IO>;
            allow_smem(kfn, tile_smem_bytes<T>());
            dim3 grid(1u << (LOGN - T - 3), ny, nz);
            GSB_LAUNCH("fwd_contig", tag, kfn, grid, dim3(1 << T), tile_smem_bytes<T>(), st, io, tab, mm, map);
        }
        else
        {
            constexpr int T = SP::T1;
            ContigMap<T> map{ tab.N };
            auto kfn = ntt_fwd_kernel<T, LOGN - 1, 0, true, ContigMap<T>, IO>;
            allow_smem(kfn, tile_smem_bytes<T>());
            GSB_LAUNCH("fwd_contig", tag, kfn, dim3(1, ny, nz), dim3(1 << T), tile_smem_bytes<T>(), st, io, tab, mm, map);
        }
    }

    // Inverse pass 1 (contiguous) -- or the whole transform when !two_pass.
    template <int LOGN, class IO>
    inline void launch_inv_contig(
        const IO &io, const NttTables &tab, const ModMap &mm, int ny, int nz, cudaStream_t st, const char *tag = "plain")
    {
        using SP = Split<LOGN>;
        if constexpr (SP::two_pass)
        {
            constexpr int T = 8;
            ContigMap<T> map{ tab.N };
            auto kfn = ntt_inv_kernel<T, LOGN - SP::A - 1, 0, false, ContigMap<T>, IO>;
            allow_smem(kfn, tile_smem_bytes<T>());
            dim3 grid(1u << (LOGN - T - 3), ny, nz);
            GSB_LAUNCH("inv_contig", tag, kfn, grid, dim3(1 << T), tile_smem_bytes<T>(), st, io, tab, mm, map);
        }
        else
        {
            constexpr int T = SP::T1;
            ContigMap<T> map{ tab.N };
            auto kfn = ntt_inv_kernel<T, LOGN - 1, 0, true, ContigMap<T>, IO>;
            allow_smem(kfn, tile_smem_bytes<T>());
            GSB_LAUNCH("inv_contig", tag, kfn, dim3(1, ny, nz), dim3(1 << T), tile_smem_bytes<T>(), st, io, tab, mm, map);
        }
    }

    // Inverse pass 2 (columns; carries the global last stage with n^-1).
    template <int LOGN, class IO>
    inline void launch_inv_cols(
        const IO &io, const NttTables &tab, const ModMap &mm, int ny, int nz, cudaStream_t st, const char *tag = "plain")
    {
        using SP = Split<LOGN>;
        constexpr int A = SP::A, TC = SP::TC, C = TC + 3 - A;
        ColMap<TC, A> map{ tab.N, tab.logn };
        auto kfn = ntt_inv_kernel<TC, A + C - 1, C, true, ColMap<TC, A>, IO>;
        allow_smem(kfn, tile_smem_bytes<TC>());
        dim3 grid((1u << (LOGN - A)) >> C, ny, nz);
        GSB_LAUNCH("inv_cols", tag, kfn, grid, dim3(1 << TC), tile_smem_bytes<TC>(), st, io, tab, mm, map);
    }

    // Plain batched transforms: element (z, y, d) of src/dst at base + z*zs + y*N + d; in place
    // allowed (src == dst with equal strides).
    template <int LOGN>
    inline void ntt_forward_plain(
        const NttTables &tab, const ModMap &mm, const uint64_t *src, size_t src_zs, uint64_t *dst, size_t dst_zs, int ny,
        int nz, bool lazy, cudaStream_t st, const char *tag = "plain")
    {
        if constexpr (Split<LOGN>::two_pass)
        {
            launch_fwd_cols<LOGN>(IoPlain<OUT_RAW>{ src, dst, src_zs, dst_zs }, tab, mm, ny, nz, st, tag);
            src = dst;
            src_zs = dst_zs;
        }
        if (lazy)
            launch_fwd_contig<LOGN>(IoPlain<OUT_RAW>{ src, dst, src_zs, dst_zs }, tab, mm, ny, nz, st, tag);
        else
            launch_fwd_contig<LOGN>(IoPlain<OUT_FWD_STRICT>{ src, dst, src_zs, dst_zs }, tab, mm, ny, nz, st, tag);
    }

    template <int LOGN>
    inline void ntt_inverse_plain(
        const NttTables &tab, const ModMap &mm, const uint64_t *src, size_t src_zs, uint64_t *dst, size_t dst_zs, int ny,
        int nz, bool lazy, cudaStream_t st, const char *tag = "plain")
    {
        if constexpr (Split<LOGN>::two_pass)
        {
            launch_inv_contig<LOGN>(IoPlain<OUT_RAW>{ src, dst, src_zs, dst_zs }, tab, mm, ny, nz, st, tag);
            if (lazy)
                launch_inv_cols<LOGN>(IoPlain<OUT_RAW>{ dst, dst, dst_zs, dst_zs }, tab, mm, ny, nz, st, tag);
            else
                launch_inv_cols<LOGN>(IoPlain<OUT_INV_STRICT>{ dst, dst, dst_zs, dst_zs }, tab, mm, ny, nz, st, tag);
        }
        else
        {
            if (lazy)
                launch_inv_contig<LOGN>(IoPlain<OUT_RAW>{ src, dst, src_zs, dst_zs }, tab, mm, ny, nz, st, tag);
            else
                launch_inv_contig<LOGN>(IoPlain<OUT_INV_STRICT>{ src, dst, src_zs, dst_zs }, tab, mm, ny, nz, st, tag);
        }
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
