// Register-blocked radix-8 negacyclic NTT building blocks (forward = Cooley-Tukey, inverse =
// Gentleman-Sande) that reproduce the reference's per-butterfly u64 arithmetic exactly
// (SlothfulNTT, util/ntt.cpp:213-282) while regrouping its radix-2 stages into rounds of up to
// three stages held in registers.  Regrouping does not change any butterfly's inputs, so every
// intermediate word -- lazy ranges and mod-2^64 wraps included -- equals the CPU's.
//
// Geometry.  A CTA of T = 2^T_LOG threads owns a tile of 8T words; each thread keeps 8 words in
// registers per round.  A "round" (TOP, R) runs the R stages whose butterfly gaps, measured in
// tile indices, are 2^TOP ... 2^(TOP-R+1).  The 8 words of a thread are 8>>R groups of 2^R words
// spaced 2^(TOP-R+1) apart, so lanes of a warp always touch consecutive tile indices (coalesced
// global access, conflict-free 64-bit shared access); the stride-1 round uses 128-bit accesses.
// Between rounds the tile is exchanged through XOR-swizzled, ping-ponged shared memory.
//
// Twiddle indexing (util/ntt.cpp:292-342, 345-404).  With S = size of the index space a stage's
// gap lives in (N for a contiguous tile at offset `off`, 2^(a+c) for a 2^a x 2^c column tile) the
// stage with gap 2^g has m = S >> (g+1) groups and the group of tile index i is (off+i) >> (g+1):
//   forward : W[m + group]                      (root_powers, bit-reversed order)
//   inverse : V[N - 2m + 1 + group]             (inv_root_powers, re-ordered at ntt.cpp:85-96)
#pragma once
#include "modarith.cuh"

namespace gsb
{
    // {w, floor(w*2^64/q)} pairs so one 128-bit load fetches both halves of a Shoup twiddle.
    typedef ulonglong2 Twiddle;

    GSB_DEV uint32_t swz(uint32_t idx)
    {
        return idx ^ (((idx >> 4) & 7u) << 1);
    }

    // ---- compile-time round plan ---------------------------------------------------------------
    // Split `ns` stages into rounds of <= 3 so that the round touching gap 2^0 has 3 stages.
    __host__ __device__ constexpr int plan_next(int left)
    {
        return left <= 3 ? left : left == 4 ? 1 : (left == 5 || left == 7) ? 2 : 3;
    }
    __host__ __device__ constexpr int plan_count(int ns)
    {
        int n = 0;
        while (ns > 0)
        {
            ns -= plan_next(ns);
            n++;
        }
        return n;
    }
    // size of round `idx` (0 = the round with the largest gaps)
    __host__ __device__ constexpr int plan_size(int ns, int idx)
    {
        int r = plan_next(ns);
        while (idx > 0)
        {
            ns -= r;
            r = plan_next(ns);
            idx--;
        }
        return r;
    }
    // number of stages in rounds [0, idx)
    __host__ __device__ constexpr int plan_before(int ns, int idx)
    {
        int done = 0;
        for (int i = 0; i < idx; i++)
            done += plan_size(ns, i);
        return done;
    }

    // tile index of word 0 of group `sub` of thread `tid` in round (TOP, R)
    template <int TOP, int R, int T_LOG>
    GSB_DEV uint32_t group_base(uint32_t tid, int sub)
    {
        constexpr int low = TOP - R + 1;
        uint32_t v = (static_cast<uint32_t>(sub) << T_LOG) + tid;
        uint32_t lo = v & ((1u << low) - 1u), hi = v >> low;
        return (hi << (TOP + 1)) | lo;
    }

    // ---- shared-memory exchange ----------------------------------------------------------------
    template <int TOP, int R, int T_LOG>
    GSB_DEV void tile_store(uint64_t *buf, const uint64_t (&x)[8], uint32_t tid)
    {
        constexpr int low = TOP - R + 1;
        if (low == 0 && R == 3)
        {
#pragma unroll
            for (int i = 0; i < 8; i += 2)
                *reinterpret_cast<ulonglong2 *>(buf + swz(tid * 8 + i)) = make_ulonglong2(x[i], x[i + 1]);
        }
        else
        {
#pragma unroll
            for (int sub = 0; sub < (8 >> R); sub++)
            {
                uint32_t base = group_base<TOP, R, T_LOG>(tid, sub);
#pragma unroll
                for (int i = 0; i < (1 << R); i++)
                    buf[swz(base + (static_cast<uint32_t>(i) << low))] = x[sub * (1 << R) + i];
            }
        }
    }

    template <int TOP, int R, int T_LOG>
    GSB_DEV void tile_load(const uint64_t *buf, uint64_t (&x)[8], uint32_t tid)
    {
        constexpr int low = TOP - R + 1;
        if (low == 0 && R == 3)
        {
#pragma unroll
            for (int i = 0; i < 8; i += 2)
            {
                ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(buf + swz(tid * 8 + i));
                x[i] = v.x;
                x[i + 1] = v.y;
            }
        }
        else
        {
#pragma unroll
            for (int sub = 0; sub < (8 >> R); sub++)
            {
                uint32_t base = group_base<TOP, R, T_LOG>(tid, sub);
#pragma unroll
                for (int i = 0; i < (1 << R); i++)
                    x[sub * (1 << R) + i] = buf[swz(base + (static_cast<uint32_t>(i) << low))];
            }
        }
    }

    // ---- forward round: ForwardLazy / ForwardLazyLast (util/ntt.cpp:245-261) -------------------
    // LAST: this transform's gap-2^0 stage is the global last stage (h = 1), where x0 is first
    // reduced with reduceBarrettLazy.
    template <int TOP, int R, int T_LOG, bool LAST>
    GSB_DEV void fwd_round(
        uint64_t (&x)[8], uint32_t tid, uint32_t off, uint32_t S, const Twiddle *__restrict__ WW, const LimbC &c)
    {
#pragma unroll
        for (int sub = 0; sub < (8 >> R); sub++)
        {
            const uint32_t base = off + group_base<TOP, R, T_LOG>(tid, sub);
#pragma unroll
            for (int s = 0; s < R; s++)
            {
                const int g = TOP - s;
                const int half = 1 << (R - 1 - s);
                const uint32_t tw0 = (S >> (g + 1)) + (base >> (g + 1));
#pragma unroll
                for (int grp = 0; grp < (1 << s); grp++)
                {
                    const Twiddle w = __ldg(WW + tw0 + grp);
#pragma unroll
                    for (int i = 0; i < half; i++)
                    {
                        const int a = sub * (1 << R) + grp * 2 * half + i, b = a + half;
                        uint64_t u = x[a];
                        if (LAST && g == 0)
                            u = reduce_barrett_lazy(u, c.rdp, c.q);
                        const uint64_t v = mulmod_lazy(x[b], w.x, w.y, c.q);
                        x[a] = u + v;
                        x[b] = u - v + c.two_q;
                    }
                }
            }
        }
    }

    // ---- inverse round: BackwardLazy / BackwardLazyLast (util/ntt.cpp:265-281) -----------------
    // Round covers gaps 2^BOT ... 2^(BOT+R-1), smallest first.  LAST: the gap-2^(BOT+R-1) stage is
    // the global last stage (m = 1), where n^-1 is folded in (util/ntt.cpp:391-402).
    template <int BOT, int R, int T_LOG, bool LAST>
    GSB_DEV void inv_round(
        uint64_t (&x)[8], uint32_t tid, uint32_t off, uint32_t S, uint32_t N, const Twiddle *__restrict__ VV,
        const LimbC &c)
    {
        constexpr int TOP = BOT + R - 1;
#pragma unroll
        for (int sub = 0; sub < (8 >> R); sub++)
        {
            const uint32_t base = off + group_base<TOP, R, T_LOG>(tid, sub);
#pragma unroll
            for (int s = 0; s < R; s++)
            {
                const int g = BOT + s;
                const uint32_t m = S >> (g + 1);
                const uint32_t tw0 = N - 2u * m + 1u + (base >> (g + 1));
#pragma unroll
                for (int grp = 0; grp < (1 << (R - 1 - s)); grp++)
                {
                    const Twiddle w = __ldg(VV + tw0 + grp);
#pragma unroll
                    for (int i = 0; i < (1 << s); i++)
                    {
                        const int a = sub * (1 << R) + (grp << (s + 1)) + i, b = a + (1 << s);
                        const uint64_t u = x[a], v = x[b];
                        uint64_t t = u + v;
                        t -= (t >= c.two_q) ? c.two_q : 0;
                        if (LAST && s == R - 1)
                            t = mulmod_lazy(t, c.inv_n, c.inv_n_s, c.q);
                        x[a] = t;
                        x[b] = mulmod_lazy(u - v + c.two_q, w.x, w.y, c.q);
                    }
                }
            }
        }
    }

    // ---- tile geometry: maps a tile index to a coefficient index of the limb --------------------
    // Contiguous tile: 8T consecutive coefficients starting at tile*8T; gaps live in the limb (S = N).
    template <int T_LOG>
    struct ContigMap
    {
        uint32_t N;
        GSB_DEV uint32_t S() const { return N; }
        GSB_DEV uint32_t off(uint32_t tile) const { return tile << (T_LOG + 3); }
        GSB_DEV uint32_t coeff(uint32_t tile, uint32_t idx) const { return (tile << (T_LOG + 3)) + idx; }
    };
    // Column tile: 2^A rows (stride N >> A) x 2^C adjacent columns, 2^(A+C) = 8T; row-major tile
    // index, so only gaps >= 2^C are butterfly gaps and S = 2^(A+C).
    template <int T_LOG, int A>
    struct ColMap
    {
        static constexpr int C = T_LOG + 3 - A;
        uint32_t N;
        int logn;
        GSB_DEV uint32_t S() const { return 1u << (A + C); }
        GSB_DEV uint32_t off(uint32_t) const { return 0; }
        GSB_DEV uint32_t coeff(uint32_t tile, uint32_t idx) const
        {
            return ((idx >> C) << (logn - A)) + (tile << C) + (idx & ((1u << C) - 1u));
        }
    };

    // mod(y): which key-level modulus an extended limb slot uses (evaluator.cpp:2317,2343):
    // slots [0,L) are ciphertext limbs, slots >= L are the special primes at key index L0 + (y-L).
    struct ModMap
    {
        int L, L0, offset;
        GSB_DEV int mod(int y) const
        {
            int s = y + offset;
            return s < L ? s : L0 + (s - L);
        }
    };

    struct NttTables
    {
        const Twiddle *WW; // [K][N] forward {w, w'}
        const Twiddle *VV; // [K][N] inverse {v, v'} (re-ordered, n^-1 folded into the last entry)
        const LimbC *limb; // [K]
        uint32_t N;
        int logn;
    };

    // ---- generic forward / inverse tile transforms ----------------------------------------------
    // Runs the stages with tile gaps 2^G_HI ... 2^G_LO.  First round loads through `pro`, last
    // round stores through `epi`; both see coefficient indices, never tile indices.
    template <int T_LOG, int G_HI, int G_LO, bool LAST, int IDX, class Map, class Pro, class Epi>
    GSB_DEV void fwd_rounds(
        uint64_t (&x)[8], uint64_t *smem, uint32_t tid, uint32_t tile, const Map &map, const Pro &pro, const Epi &epi,
        const Twiddle *__restrict__ WW, const LimbC &c, int &parity)
    {
        constexpr int NS = G_HI - G_LO + 1;
        constexpr int NR = plan_count(NS);
        if constexpr (IDX < NR)
        {
            constexpr int R = plan_size(NS, IDX);
            constexpr int TOP = G_HI - plan_before(NS, IDX);
            constexpr int low = TOP - R + 1;
            if constexpr (IDX == 0)
            {
#pragma unroll
                for (int sub = 0; sub < (8 >> R); sub++)
                {
                    uint32_t base = group_base<TOP, R, T_LOG>(tid, sub);
                    if (low == 0 && R == 3)
                    {
#pragma unroll
                        for (int i = 0; i < 8; i += 2)
                            pro.load2(map.coeff(tile, base + i), x[i], x[i + 1]);
                    }
                    else
                    {
#pragma unroll
                        for (int i = 0; i < (1 << R); i++)
                            x[sub * (1 << R) + i] = pro.load(map.coeff(tile, base + (static_cast<uint32_t>(i) << low)));
                    }
                }
            }
            else
            {
                uint64_t *buf = smem + (static_cast<size_t>(parity) << (T_LOG + 3));
                parity ^= 1;
                __syncthreads();
                tile_load<TOP, R, T_LOG>(buf, x, tid);
            }
            fwd_round<TOP, R, T_LOG, LAST>(x, tid, map.off(tile), map.S(), WW, c);
            if constexpr (IDX == NR - 1)
            {
#pragma unroll
                for (int sub = 0; sub < (8 >> R); sub++)
                {
                    uint32_t base = group_base<TOP, R, T_LOG>(tid, sub);
                    if (low == 0 && R == 3)
                    {
#pragma unroll
                        for (int i = 0; i < 8; i += 2)
                            epi.store2(map.coeff(tile, base + i), x[i], x[i + 1], i);
                    }
                    else
                    {
#pragma unroll
                        for (int i = 0; i < (1 << R); i++)
                            epi.store(map.coeff(tile, base + (static_cast<uint32_t>(i) << low)), x[sub * (1 << R) + i], sub * (1 << R) + i);
                    }
                }
            }
            else
            {
                uint64_t *buf = smem + (static_cast<size_t>(parity) << (T_LOG + 3));
                tile_store<TOP, R, T_LOG>(buf, x, tid);
                fwd_rounds<T_LOG, G_HI, G_LO, LAST, IDX + 1>(x, smem, tid, tile, map, pro, epi, WW, c, parity);
            }
        }
    }

    // Inverse: rounds run in the reverse order of the forward plan (smallest gaps first).
    template <int T_LOG, int G_HI, int G_LO, bool LAST, int IDX, class Map, class Pro, class Epi>
    GSB_DEV void inv_rounds(
        uint64_t (&x)[8], uint64_t *smem, uint32_t tid, uint32_t tile, const Map &map, const Pro &pro, const Epi &epi,
        const Twiddle *__restrict__ VV, uint32_t N, const LimbC &c, int &parity)
    {
        constexpr int NS = G_HI - G_LO + 1;
        constexpr int NR = plan_count(NS);
        if constexpr (IDX < NR)
        {
            constexpr int PI = NR - 1 - IDX; // forward-plan round executed at this step
            constexpr int R = plan_size(NS, PI);
            constexpr int TOP = G_HI - plan_before(NS, PI);
            constexpr int BOT = TOP - R + 1;
            if constexpr (IDX == 0)
            {
#pragma unroll
                for (int sub = 0; sub < (8 >> R); sub++)
                {
                    uint32_t base = group_base<TOP, R, T_LOG>(tid, sub);
                    if (BOT == 0 && R == 3)
                    {
#pragma unroll
                        for (int i = 0; i < 8; i += 2)
                            pro.load2(map.coeff(tile, base + i), x[i], x[i + 1]);
                    }
                    else
                    {
#pragma unroll
                        for (int i = 0; i < (1 << R); i++)
                            x[sub * (1 << R) + i] = pro.load(map.coeff(tile, base + (static_cast<uint32_t>(i) << BOT)));
                    }
                }
            }
            else
            {
                uint64_t *buf = smem + (static_cast<size_t>(parity) << (T_LOG + 3));
                parity ^= 1;
                __syncthreads();
                tile_load<TOP, R, T_LOG>(buf, x, tid);
            }
            inv_round<BOT, R, T_LOG, (LAST && IDX == NR - 1)>(x, tid, map.off(tile), map.S(), N, VV, c);
            if constexpr (IDX == NR - 1)
            {
#pragma unroll
                for (int sub = 0; sub < (8 >> R); sub++)
                {
                    uint32_t base = group_base<TOP, R, T_LOG>(tid, sub);
                    if (BOT == 0 && R == 3)
                    {
#pragma unroll
                        for (int i = 0; i < 8; i += 2)
                            epi.store2(map.coeff(tile, base + i), x[i], x[i + 1], i);
                    }
                    else
                    {
#pragma unroll
                        for (int i = 0; i < (1 << R); i++)
                            epi.store(map.coeff(tile, base + (static_cast<uint32_t>(i) << BOT)), x[sub * (1 << R) + i], sub * (1 << R) + i);
                    }
                }
            }
            else
            {
                uint64_t *buf = smem + (static_cast<size_t>(parity) << (T_LOG + 3));
                tile_store<TOP, R, T_LOG>(buf, x, tid);
                inv_rounds<T_LOG, G_HI, G_LO, LAST, IDX + 1>(x, smem, tid, tile, map, pro, epi, VV, N, c, parity);
            }
        }
    }
} // namespace gsb
