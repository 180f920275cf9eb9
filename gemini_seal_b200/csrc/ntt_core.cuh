// Register-blocked radix-8 negacyclic NTT building blocks (forward = Cooley-Tukey, inverse =
// Gentleman-Sande) that reproduce the reference's per-butterfly u64 arithmetic exactly
// (SlothfulNTT, util/ntt.cpp:213-282) while regrouping its radix-2 stages into rounds of up to
// three stages held in registers.  Regrouping does not change any butterfly's inputs, so every
// intermediate word -- lazy ranges and mod-2^64 wraps included -- equals the CPU's.
// That literal form serves the public util::ntt_* entry points, every inverse transform and every
// context whose moduli fail gsb_ctx::fast_ntt.  Forward transforms that stay inside the library and
// end in a canonical reduction may instead run the wide-range butterfly (`FAST`, see fwd_round):
// congruent words, one wide multiply less, identical outputs.
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
#include <type_traits>
#include <utility>
#include "modarith.cuh"

namespace gsb
{
    // {w, floor(w*2^64/q)} pairs so one 128-bit load fetches both halves of a Shoup twiddle.
    typedef ulonglong2 Twiddle;

    GSB_DEV uint32_t swz(uint32_t idx)
    {
        return idx ^ (((idx >> 4) & 7u) << 1);
    }

    // ---- bank-conflict-free layout of the small-gap twiddles and of the switching keys -----------
    // In the stride-1 round a thread owns 8 adjacent coefficients, so lane l of a warp reads the
    // twiddles of the gap-1 / gap-2 stages (and its key words) at a 64 / 32 byte stride: 4- / 2-way
    // shared-memory bank conflicts on 128-bit loads.  The DEVICE copies of those arrays are therefore
    // stored with their 16-byte units XOR-swizzled inside every block of 64 units (unit u of the block
    // lives at u ^ ((u >> 3) & 7)): a linear bulk copy of any 64-unit-aligned range carries the
    // permutation into shared memory, where 8 lanes at a 64- or 32-byte stride then hit 8 different
    // 16-byte bank groups.  Only the device layout changes; every word read is the same word.
    __host__ __device__ inline uint32_t swz8(uint32_t u)
    {
        return u ^ ((u >> 3) & 7u);
    }
    // Device position of forward twiddle `idx` (root_powers order): the entries of the three
    // smallest-gap stages, idx >= N/8, are swizzled.
    __host__ __device__ inline uint32_t fwd_tw_pos(uint32_t idx, uint32_t N)
    {
        return idx >= (N >> 3) ? swz8(idx) : idx;
    }
    // Device position of inverse twiddle `idx` (re-ordered table, entries 1 .. N-1): the three
    // smallest-gap stages are idx - 1 < 7N/8; their blocks start at idx - 1 = multiple of 64.
    __host__ __device__ inline uint32_t inv_tw_pos(uint32_t idx, uint32_t N)
    {
        return (idx >= 1 && idx - 1 < N - (N >> 3)) ? 1u + swz8(idx - 1) : idx;
    }
    // Device position of word d of a switching-key limb (16-byte unit = 2 words).
    __host__ __device__ inline uint32_t key_word_pos(uint32_t d)
    {
        return (swz8(d >> 1) << 1) | (d & 1u);
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

    // ---- twiddle sources ------------------------------------------------------------------------
    // fwd(g, local, grp) / inv(g, local, grp): the Shoup twiddle of the stage with tile gap 2^g for the
    // butterfly group that contains tile-local index `local`, plus `grp` groups.  (Indexing rules at
    // the top of this file.)
    //
    // TwGlobal reads the tables straight from global memory (L2 / L1).
    struct TwGlobal
    {
        const Twiddle *tab; // WW (forward) or VV (inverse) of the limb's modulus
        uint32_t S, off, N;
        GSB_DEV Twiddle fwd(int g, uint32_t local, int grp) const
        {
            return __ldg(tab + fwd_tw_pos((S >> (g + 1)) + ((off + local) >> (g + 1)) + grp, N));
        }
        GSB_DEV Twiddle inv(int g, uint32_t local, int grp) const
        {
            return __ldg(tab + inv_tw_pos(N - 2u * (S >> (g + 1)) + 1u + ((off + local) >> (g + 1)) + grp, N));
        }
    };

    // TwContigSmem: contiguous tile of 2^TILE_LOG words at coefficient offset `off`.  The stage with
    // gap 2^g uses the 2^TILE_LOG >> (g+1) consecutive table entries starting at
    //   forward: W[(N >> (g+1)) + (off >> (g+1))]      inverse: V[N - 2(N >> (g+1)) + 1 + (off >> (g+1))]
    // which a single thread stages into shared memory with one TMA bulk copy per stage
    // (segment g at pair offset TILE - (TILE >> g)); stages g > G_SM stay on the global path.
    template <int TILE_LOG, int G_SM>
    struct TwContigSmem
    {
        static constexpr uint32_t TILE = 1u << TILE_LOG;
        static constexpr uint32_t PAIRS = TILE - (TILE >> (G_SM + 1)); // pairs staged for g = 0..G_SM
        const Twiddle *sm;
        TwGlobal gl;
        GSB_DEV static uint32_t seg(int g) { return TILE - (TILE >> g); }
        // stages g <= 2 arrive swizzled (fwd_tw_pos / inv_tw_pos): every staged range starts at a
        // multiple of 64 entries, so the position inside the segment is swizzled the same way
        GSB_DEV static uint32_t pos(int g, uint32_t rel) { return g <= 2 ? swz8(rel) : rel; }
        GSB_DEV Twiddle fwd(int g, uint32_t local, int grp) const
        {
            return g <= G_SM ? sm[seg(g) + pos(g, (local >> (g + 1)) + grp)] : gl.fwd(g, local, grp);
        }
        GSB_DEV Twiddle inv(int g, uint32_t local, int grp) const
        {
            return g <= G_SM ? sm[seg(g) + pos(g, (local >> (g + 1)) + grp)] : gl.inv(g, local, grp);
        }
        // issued by ONE thread after mbar_init + a CTA barrier
        GSB_DEV static void stage(Twiddle *sm, const Twiddle *tab, uint32_t N, uint32_t off, bool inverse, uint64_t *bar)
        {
            mbar_expect(bar, PAIRS * static_cast<uint32_t>(sizeof(Twiddle)));
#pragma unroll
            for (int g = 0; g <= G_SM; g++)
            {
                const uint32_t m = N >> (g + 1);
                const Twiddle *src = tab + (inverse ? N - 2u * m + 1u : m) + (off >> (g + 1));
                bulk_g2s(sm + seg(g), src, (TILE >> (g + 1)) * static_cast<uint32_t>(sizeof(Twiddle)), bar);
            }
        }
    };

    // TwColsSmem: column tile (2^A rows): every stage indexes the first 2^A forward entries, or the
    // last 2^A inverse entries, of the table -- 2^A pairs staged with one bulk copy.
    template <int A, int C>
    struct TwColsSmem
    {
        static constexpr uint32_t PAIRS = 1u << A;
        const Twiddle *sm;
        GSB_DEV Twiddle fwd(int g, uint32_t local, int grp) const
        {
            return sm[((1u << (A + C)) >> (g + 1)) + (local >> (g + 1)) + grp];
        }
        GSB_DEV Twiddle inv(int g, uint32_t local, int grp) const
        {
            return sm[PAIRS - 2u * ((1u << (A + C)) >> (g + 1)) + 1u + (local >> (g + 1)) + grp];
        }
        GSB_DEV static void stage(Twiddle *sm, const Twiddle *tab, uint32_t N, bool inverse, uint64_t *bar)
        {
            mbar_expect(bar, PAIRS * static_cast<uint32_t>(sizeof(Twiddle)));
            bulk_g2s(sm, inverse ? tab + N - PAIRS : tab, PAIRS * static_cast<uint32_t>(sizeof(Twiddle)), bar);
        }
    };

    // A twiddle source whose Shoup halves were already shifted right by one in shared memory (what
    // mulmod_wide_add wants): saves the two funnel shifts per twiddle in kernels that reuse a staged
    // table many times.  Only valid when every stage of the transform is a wide-range stage.
    template <class TW>
    struct TwPreshifted
    {
        TW tw;
        static constexpr bool kPreshifted = true;
        GSB_DEV Twiddle fwd(int g, uint32_t local, int grp) const { return tw.fwd(g, local, grp); }
    };
    template <class T, class = void>
    struct tw_preshifted : std::false_type
    {};
    template <class T>
    struct tw_preshifted<T, std::void_t<decltype(T::kPreshifted)>> : std::true_type
    {};

    // ---- forward round: ForwardLazy / ForwardLazyLast (util/ntt.cpp:245-261) -------------------
    // LAST: this transform's gap-2^0 stage is the global last stage (h = 1), where x0 is first
    // reduced with reduceBarrettLazy.
    //
    // FAST (wide-range butterflies): only for transforms whose words stay inside the library and end
    // in a canonical reduction, and only when gsb_ctx::fast_ntt proves no word can reach 2^63.  The
    // twiddle product comes from mulmod_wide_add (one IMAD.WIDE and four carry adds fewer), which
    // returns a word CONGRUENT to the reference's mulmodLazy in [0,5q) instead of [0,2q); the
    // butterfly therefore uses u - v + 5q.  A word grows by < 5q per stage, so after s stages it is
    // < in + 5qs.  The global last stage (LAST, g = 0) stays the literal ForwardLazyLast, whose
    // reduceBarrettLazy / exact Shoup product accept any 64-bit word: its outputs are again in
    // [0,4q), the range every epilogue of the reference expects.
    // FAST is an int: 0 = literal; otherwise 1 + a bit mask of how the wide-range butterfly is written:
    //   1 (H3)    the quotient's carry word as a three-input sum (mulmod_wide_add): that carry on the ALU pipe;
    //   2 (D3)    the difference as two three-input sums (twice_plus_minus) instead of the C++ expression, which NVVM
    //             regroups into two-input sums whose carry words ptxas may turn into IMAD.X on the multiplier pipe;
    //   4 (EARLY) the h-independent narrow products ahead of the wide chain (mulmod_wide_add).
    // `true` (= 1) is the plain form.  Which one is fastest depends on which of the two pipes the rest of the kernel
    // loads more and on its register budget ("two-pipe model"; scripts/sass_pipes.py): the tile
    // kernels use D3 + EARLY (7), ks_mac D3 (3).
    template <int TOP, int R, int T_LOG, bool LAST, int FAST, class TW>
    GSB_DEV void fwd_round(uint64_t (&x)[8], uint32_t tid, const TW &tw, const LimbC &c)
    {
        const uint64_t five_q = c.two_q + c.two_q + c.q, neg_two_q = 0 - c.two_q;
#pragma unroll
        for (int sub = 0; sub < (8 >> R); sub++)
        {
            const uint32_t base = group_base<TOP, R, T_LOG>(tid, sub);
#pragma unroll
            for (int s = 0; s < R; s++)
            {
                const int g = TOP - s;
                const int half = 1 << (R - 1 - s);
#pragma unroll
                for (int grp = 0; grp < (1 << s); grp++)
                {
                    const Twiddle w = tw.fwd(g, base, grp);
                    const bool wide = FAST && !(LAST && g == 0);
                    const uint64_t wsh = tw_preshifted<TW>::value ? w.y : (w.y >> 1);
#pragma unroll
                    for (int i = 0; i < half; i++)
                    {
                        const int a = sub * (1 << R) + grp * 2 * half + i, b = a + half;
                        uint64_t u = x[a];
                        if (LAST && g == 0)
                            u = reduce_barrett_lazy(u, c.rdp, c.q);
                        // x[a] = u + v with u folded into v's multiply chain; x[b] = u - v + 2q
                        //      = (2u + 2q) - x[a]  (all mod 2^64, exactly the reference's words)
                        // (c.pad is a run-time zero: it keeps both sums three-input IADD3s on the ALU pipe
                        // instead of IMAD.X on the FMA-heavy pipe that bounds these kernels)
                        if (wide)
                        {
                            const uint64_t s = mulmod_wide_add<(((FAST - 1) & 1) != 0), (((FAST - 1) & 4) != 0)>(x[b], w.x, wsh, neg_two_q, u, c.pad);
                            if constexpr (((FAST - 1) & 2) != 0)
                                x[b] = twice_plus_minus(u, five_q, s, c.pad);
                            else
                                x[b] = (u + u + five_q) + (c.pad - s);
                            x[a] = s;
                            GSB_SIM_ASSERT((x[a] >> 63) == 0 && (x[b] >> 63) == 0);
                        }
                        else
                        {
                            const uint64_t s = mulmod_lazy_add(x[b], w.x, w.y, c.q, u);
                            x[b] = twice_plus_minus(u, c.two_q, s, c.pad);
                            x[a] = s;
                        }
                    }
                }
            }
        }
    }

    // ---- inverse round: BackwardLazy / BackwardLazyLast (util/ntt.cpp:265-281) -----------------
    // Round covers gaps 2^BOT ... 2^(BOT+R-1), smallest first.  LAST: the gap-2^(BOT+R-1) stage is
    // the global last stage (m = 1), where n^-1 is folded in (util/ntt.cpp:391-402).
    // FAST (wide-range inverse butterflies, gsb_ctx::wide_inv: 10 q_max < 2^63): words live in [0,5q)
    // instead of [0,2q) -- the sum takes a conditional subtraction of 5q, the difference u - v + 5q < 10q
    // goes through mulmod_wide_add -- with no growth from stage to stage.  The global last stage stays
    // the literal BackwardLazyLast (exact Shoup products accept any 64-bit word), so the transform's
    // outputs are in [0,2q) as in the reference.  Only for transforms whose output is canonicalised
    // (strict) or consumed inside the library.
    template <int BOT, int R, int T_LOG, bool LAST, bool FAST = false, class TW>
    GSB_DEV void inv_round(uint64_t (&x)[8], uint32_t tid, const TW &tw, const LimbC &c)
    {
        constexpr int TOP = BOT + R - 1;
        const uint64_t five_q = c.two_q + c.two_q + c.q, neg_two_q = 0 - c.two_q;
#pragma unroll
        for (int sub = 0; sub < (8 >> R); sub++)
        {
            const uint32_t base = group_base<TOP, R, T_LOG>(tid, sub);
#pragma unroll
            for (int s = 0; s < R; s++)
            {
                const int g = BOT + s;
#pragma unroll
                for (int grp = 0; grp < (1 << (R - 1 - s)); grp++)
                {
                    const Twiddle w = tw.inv(g, base, grp);
                    const bool last = LAST && s == R - 1;
                    const uint64_t wsh = w.y >> 1;
#pragma unroll
                    for (int i = 0; i < (1 << s); i++)
                    {
                        const int a = sub * (1 << R) + (grp << (s + 1)) + i, b = a + (1 << s);
                        const uint64_t u = x[a], v = x[b];
                        uint64_t t = u + v + c.pad; // c.pad == 0 at run time, see fwd_round
                        if (FAST)
                        {
                            GSB_SIM_ASSERT(u < five_q && v < five_q);
                            if (last)
                            {
                                x[a] = mulmod_lazy(t, c.inv_n, c.inv_n_s, c.q);
                                x[b] = mulmod_lazy(u - v + five_q, w.x, w.y, c.q);
                            }
                            else
                            {
                                t -= (t >= five_q) ? five_q : 0;
                                x[a] = t;
                                x[b] = mulmod_wide_add<false, false>(u - v + five_q, w.x, wsh, neg_two_q, 0, c.pad); // the inverse rounds are ALU-bound: carries left to ptxas
                            }
                        }
                        else
                        {
                            t -= (t >= c.two_q) ? c.two_q : 0;
                            if (last)
                                t = mulmod_lazy(t, c.inv_n, c.inv_n_s, c.q);
                            x[a] = t;
                            x[b] = mulmod_lazy(u - v + c.two_q, w.x, w.y, c.q);
                        }
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
    // A transform over tile gaps 2^G_HI ... 2^G_LO is: tile_load (raw global loads in the register
    // layout of the first round) -> tile_xform (the prologue's arithmetic) -> *_compute (all rounds,
    // exchanging through shared memory) -> tile_store (epilogue, register layout of the last round).
    // Loads are split from the prologue arithmetic so a kernel can issue the NEXT tile's loads before
    // it computes on the current one.  Pro/Epi see coefficient indices, never tile indices.
    template <int NS, bool INVERSE>
    struct RoundPlan
    {
        static constexpr int NR = plan_count(NS);
        // plan index of the first / last round EXECUTED (the inverse walks the plan backwards)
        static constexpr int first = INVERSE ? NR - 1 : 0;
        static constexpr int last = INVERSE ? 0 : NR - 1;
    };

    template <int TOP, int R, int T_LOG, class Map, class Pro>
    GSB_DEV void tile_load_raw(uint64_t (&x)[8], uint32_t tid, uint32_t tile, const Map &map, const Pro &pro)
    {
        constexpr int low = TOP - R + 1;
#pragma unroll
        for (int sub = 0; sub < (8 >> R); sub++)
        {
            const uint32_t base = group_base<TOP, R, T_LOG>(tid, sub);
            if (low == 0 && R == 3)
            {
#pragma unroll
                for (int i = 0; i < 8; i += 2)
                    pro.raw2(map.coeff(tile, base + i), x[i], x[i + 1]);
            }
            else
            {
#pragma unroll
                for (int i = 0; i < (1 << R); i++)
                    x[sub * (1 << R) + i] = pro.raw(map.coeff(tile, base + (static_cast<uint32_t>(i) << low)));
            }
        }
    }

    template <class Pro>
    GSB_DEV void tile_xform(uint64_t (&x)[8], const Pro &pro)
    {
#pragma unroll
        for (int i = 0; i < 8; i++)
            x[i] = pro.xform(x[i]);
    }

    template <int TOP, int R, int T_LOG, class Map, class Epi>
    GSB_DEV void tile_store_global(const uint64_t (&x)[8], uint32_t tid, uint32_t tile, const Map &map, const Epi &epi)
    {
        constexpr int low = TOP - R + 1;
#pragma unroll
        for (int sub = 0; sub < (8 >> R); sub++)
        {
            const uint32_t base = group_base<TOP, R, T_LOG>(tid, sub);
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

    // Epilogues that read extra operands may offer prefetch(d): called for the elements the thread will
    // store, before its butterflies, so those operands are on their way while the integer pipe works.
    template <class E, class = void>
    struct has_prefetch : std::false_type
    {};
    template <class E>
    struct has_prefetch<E, std::void_t<decltype(std::declval<const E &>().prefetch(0u))>> : std::true_type
    {};

    template <int TOP, int R, int T_LOG, class Map, class Epi>
    GSB_DEV void tile_prefetch(uint32_t tid, uint32_t tile, const Map &map, const Epi &epi)
    {
        constexpr int low = TOP - R + 1;
#pragma unroll
        for (int sub = 0; sub < (8 >> R); sub++)
        {
            const uint32_t base = group_base<TOP, R, T_LOG>(tid, sub);
            if (low == 0 && R == 3)
                epi.prefetch(map.coeff(tile, base)); // 8 adjacent words: one 64-byte segment
            else
            {
#pragma unroll
                for (int i = 0; i < (1 << R); i++)
                    epi.prefetch(map.coeff(tile, base + (static_cast<uint32_t>(i) << low)));
            }
        }
    }

    // Loads / stores of a whole transform (first / last executed round's layout).
    template <int T_LOG, int G_HI, int G_LO, bool INVERSE, class Map, class Pro>
    GSB_DEV void xf_load_raw(uint64_t (&x)[8], uint32_t tid, uint32_t tile, const Map &map, const Pro &pro)
    {
        constexpr int NS = G_HI - G_LO + 1;
        constexpr int PI = RoundPlan<NS, INVERSE>::first;
        tile_load_raw<G_HI - plan_before(NS, PI), plan_size(NS, PI), T_LOG>(x, tid, tile, map, pro);
    }
    template <int T_LOG, int G_HI, int G_LO, bool INVERSE, class Map, class Epi>
    GSB_DEV void xf_prefetch(uint32_t tid, uint32_t tile, const Map &map, const Epi &epi)
    {
        constexpr int NS = G_HI - G_LO + 1;
        constexpr int PI = RoundPlan<NS, INVERSE>::last;
        tile_prefetch<G_HI - plan_before(NS, PI), plan_size(NS, PI), T_LOG>(tid, tile, map, epi);
    }
    template <int T_LOG, int G_HI, int G_LO, bool INVERSE, class Map, class Epi>
    GSB_DEV void xf_store(const uint64_t (&x)[8], uint32_t tid, uint32_t tile, const Map &map, const Epi &epi)
    {
        constexpr int NS = G_HI - G_LO + 1;
        constexpr int PI = RoundPlan<NS, INVERSE>::last;
        tile_store_global<G_HI - plan_before(NS, PI), plan_size(NS, PI), T_LOG>(x, tid, tile, map, epi);
    }

    // All rounds of a forward transform; x enters in the first round's layout, leaves in the last's.
    template <int T_LOG, int G_HI, int G_LO, bool LAST, int IDX, int FAST = 0, class TW>
    GSB_DEV void fwd_compute(uint64_t (&x)[8], uint64_t *smem, uint32_t tid, const TW &tw, const LimbC &c, int &parity)
    {
        constexpr int NS = G_HI - G_LO + 1;
        constexpr int NR = plan_count(NS);
        if constexpr (IDX < NR)
        {
            constexpr int R = plan_size(NS, IDX);
            constexpr int TOP = G_HI - plan_before(NS, IDX);
            if constexpr (IDX > 0)
            {
                uint64_t *buf = smem + (static_cast<size_t>(parity) << (T_LOG + 3));
                parity ^= 1;
                __syncthreads();
                tile_load<TOP, R, T_LOG>(buf, x, tid);
            }
            fwd_round<TOP, R, T_LOG, LAST, FAST>(x, tid, tw, c);
            if constexpr (IDX < NR - 1)
            {
                uint64_t *buf = smem + (static_cast<size_t>(parity) << (T_LOG + 3));
                tile_store<TOP, R, T_LOG>(buf, x, tid);
                fwd_compute<T_LOG, G_HI, G_LO, LAST, IDX + 1, FAST>(x, smem, tid, tw, c, parity);
            }
        }
    }

    // Inverse: rounds run in the reverse order of the forward plan (smallest gaps first).
    template <int T_LOG, int G_HI, int G_LO, bool LAST, int IDX, bool FAST = false, class TW>
    GSB_DEV void inv_compute(uint64_t (&x)[8], uint64_t *smem, uint32_t tid, const TW &tw, const LimbC &c, int &parity)
    {
        constexpr int NS = G_HI - G_LO + 1;
        constexpr int NR = plan_count(NS);
        if constexpr (IDX < NR)
        {
            constexpr int PI = NR - 1 - IDX; // forward-plan round executed at this step
            constexpr int R = plan_size(NS, PI);
            constexpr int TOP = G_HI - plan_before(NS, PI);
            constexpr int BOT = TOP - R + 1;
            if constexpr (IDX > 0)
            {
                uint64_t *buf = smem + (static_cast<size_t>(parity) << (T_LOG + 3));
                parity ^= 1;
                __syncthreads();
                tile_load<TOP, R, T_LOG>(buf, x, tid);
            }
            inv_round<BOT, R, T_LOG, (LAST && IDX == NR - 1), FAST>(x, tid, tw, c);
            if constexpr (IDX < NR - 1)
            {
                uint64_t *buf = smem + (static_cast<size_t>(parity) << (T_LOG + 3));
                tile_store<TOP, R, T_LOG>(buf, x, tid);
                inv_compute<T_LOG, G_HI, G_LO, LAST, IDX + 1, FAST>(x, smem, tid, tw, c, parity);
            }
        }
    }
} // namespace gsb
