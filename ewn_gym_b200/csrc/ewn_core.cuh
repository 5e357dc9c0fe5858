// ewn_core.cuh -- EinStein wuerfelt nicht rules, leaf evaluation and Philox draw schedule on a
// bit-packed piece list.  Every function here is the per-game body of a CUDA kernel in
// ewn_kernels.cu; the functions are marked __host__ __device__ only so that tests/hostsim can
// single-step the very same code under gdb / gcc sanitizers.  The shipped library has no host path.
//
// Reference semantics (jchen8tw/ewn-gym, cited as file:line relative to the reference root):
//   envs/ewn.py:94-107   initial layout            envs/ewn.py:131-142  check_win
//   envs/ewn.py:144-215  nearest cube / cube to move   envs/ewn.py:217-263  execute_move (captures ANY occupant)
//   envs/ewn.py:325-375  legal directions / ordered legal list
//   envs/ewn.py:436-486  step                       envs/training_ewn.py:40-99 shaped step
//   envs/minimax_ewn.py:29-213 leaf heuristics      classical_policies/minimax.py:19-73 expectiminimax
//   classical_policies/mcts.py:21-45 random playout
//
// State layout (DESIGN.md "Packed state"):
//   P5 (board_size <= 5): one u64 per game. low u32 = TOP_LEFT side, high u32 = BOTTOM_RIGHT side.
//        side word bits [5k,5k+5) = cell of cube k+1, cell = x*5+y, 31 = captured / not on this board.
//        low[31:30] = (dice-1)&3, high[30] = (dice-1)>>2, high[31] = side to move (0 TL, 1 BR).
//   P8 (board_size <= 9): two u64 per game (16-byte vector). word0 = TL side, word1 = BR side.
//        side word bits [8k,8k+8) = cell of cube k+1, cell = (x<<4)|y, 0xFF = captured.
//        word0[50:48] = dice-1, word1[48] = side to move.
//   PG (cube_layer >= 4, i.e. 10..28 cubes per side; envs/ewn.py:47,51 allow any cube_layer < board_size - 1):
//        2 * cw + 1 u64 per game, cw = ceil(cube_num / 8): cw TL words, cw BR words (byte k of a side = cell of cube k+1 as in P8,
//        0xFF = captured or no such cube), then one meta word: bits [7:0] = dice, bit 8 = side to move.
//        P5 / P8 stay the specialised six-cube fast paths; PG runs the same rule code on byte arrays.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define EWN_HD __host__ __device__ __forceinline__
#else
#define EWN_HD inline
#endif

namespace ewn {

typedef uint32_t u32;
typedef uint64_t u64;

// ------------------------------------------------------------------------------------------------
// bit intrinsics (device instructions when compiled for the GPU); ctz_nz / msb_nz return -1 for x == 0 on both sides
// ------------------------------------------------------------------------------------------------
EWN_HD int ctz_nz(u32 x) {
#if defined(__CUDA_ARCH__)
    return __ffs((int)x) - 1;
#else
    return x ? __builtin_ctz(x) : -1;
#endif
}
EWN_HD int ctz_nz(u64 x) {
#if defined(__CUDA_ARCH__)
    return __ffsll((long long)x) - 1;
#else
    return x ? __builtin_ctzll(x) : -1;
#endif
}
EWN_HD int msb_nz(u32 x) {
#if defined(__CUDA_ARCH__)
    int r; asm("bfind.u32 %0, %1;" : "=r"(r) : "r"(x)); return r;        // one FLO; -1 for x == 0
#else
    return x ? 31 - __builtin_clz(x) : -1;
#endif
}
EWN_HD int msb_nz(u64 x) {
#if defined(__CUDA_ARCH__)
    int r; asm("bfind.u64 %0, %1;" : "=r"(r) : "l"(x)); return r;
#else
    return x ? 63 - __builtin_clzll(x) : -1;
#endif
}
EWN_HD int popc(u32 x) {
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}
EWN_HD int popc(u64 x) {
#if defined(__CUDA_ARCH__)
    return __popcll(x);
#else
    return __builtin_popcountll(x);
#endif
}
EWN_HD u32 mulhi(u32 a, u32 b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (u32)(((u64)a * (u64)b) >> 32);
#endif
}
// IEEE-754 binary64 with round-to-nearest and NO fused contraction: the reference computes in CPython doubles.
EWN_HD double f64_mul(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dmul_rn(a, b);
#else
    volatile double r = a * b; return r;
#endif
}
EWN_HD double f64_add(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dadd_rn(a, b);
#else
    volatile double r = a + b; return r;
#endif
}
EWN_HD double f64_sub(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dsub_rn(a, b);
#else
    volatile double r = a - b; return r;
#endif
}
EWN_HD double f64_div(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __ddiv_rn(a, b);
#else
    volatile double r = a / b; return r;
#endif
}

// ------------------------------------------------------------------------------------------------
// geometry of one (board_size, cube_layer) configuration; filled on the host (ewn_kernels.cu make_geo)
// ------------------------------------------------------------------------------------------------
struct Geo {
    int32_t N, L, cube_num;
    int32_t corner;                          // packed cell of (N-1, N-1)
    u32 tl_col, tl_row, br_col, br_row;      // P5 only: cells from which dir 0 / dir 1 leaves the board
    u64 tl0, br0;                            // side words of the initial layout (envs/ewn.py:94-107); PG: word 0 of g_tl0 / g_br0
    int32_t cw, sw;                          // PG: u64 words per side / per game (P5: 0 / 1, P8: 0 / 2)
    u64 g_tl0[4], g_br0[4];                  // PG: the initial layout, unused bytes 0xFF
};

// Geo plus the shared-memory address of a per-block table  dirs[side][cell]  (k_vstep, k_playouts): the legal-direction set of
// a cube becomes one LDS instead of ~9 integer instructions.  Functions below are generic in the geometry type G.
struct GeoLut : Geo {
    u32 lut;              // shared address of dirs[2][C::LUT]; the table min(x, y)[C::LUT] (dead cube: -1) follows it
    const double* vt;     // optional global table of 'hybrid' leaf values (k_emm_table), nullptr = compute
};

enum { SIDE_TL = 0, SIDE_BR = 1 };
enum { H_HYBRID = 0, H_MIN_DIST = 1, H_TWO_MIN_DIST = 2, H_ATTK = 3, H_SIM_WINRATE = 4 };
enum { CODE_ONGOING = 0, CODE_AGENT_ILLEGAL = 1, CODE_AGENT_WON = 2, CODE_OPP_ILLEGAL = 3, CODE_AGENT_LOST = 4,
       CODE_ILLEGAL_TOLERATED = 5 };

// ------------------------------------------------------------------------------------------------
// codecs
// ------------------------------------------------------------------------------------------------
struct P5 {
    typedef u32 W;
    static constexpr int FB = 5, S = 5, WORDS = 1, MAXN = 5;
    static constexpr W FM = 31u, ONES = 0x02108421u, ALL = 0x3FFFFFFFu;
    static constexpr W LOWS = 0x02108421u * 15u, HIGHS = 0x02108421u << 4;
    static EWN_HD int cell(int x, int y) { return x * 5 + y; }
    static EWN_HD int x_of(int c) { return (c * 13) >> 6; }          // c / 5 for c < 32
    static EWN_HD int y_of(int c) { return c - 5 * x_of(c); }
    static EWN_HD W dead(W s) { W t = s & (s >> 1); W u = t & (t >> 2); return u & (s >> 4) & ONES; }
    template <int BR> static EWN_HD int dirs(const Geo& g, int c) {
        u32 e0 = BR ? g.br_col : g.tl_col, e1 = BR ? g.br_row : g.tl_row;
        u32 ok0 = ~(e0 >> c) & 1u, ok1 = ~(e1 >> c) & 1u;
        return (int)(ok0 | (ok1 << 1) | ((ok0 & ok1) << 2));
    }
    static constexpr bool GENERIC = false;
    static constexpr int MAXC = 6;
    static EWN_HD W all() { return ALL; }
    static EWN_HD int get(W s, int k) { return (int)((s >> (FB * k)) & FM); }
    static EWN_HD void set(W& s, int k, int c) { s = (s & ~(FM << (FB * k))) | ((W)c << (FB * k)); }
    static EWN_HD int ncubes(const Geo&) { return 6; }
    static constexpr int LUT = 32;                                   // table entries per side (every 5-bit field value)
    template <int BR> static EWN_HD int dirs(const GeoLut& g, int c) {
#if defined(__CUDA_ARCH__)
        u32 v; asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(g.lut + (u32)(BR * LUT) + (u32)c)); return (int)v;
#else
        return dirs<BR>(static_cast<const Geo&>(g), c);
#endif
    }
};

struct P8 {
    typedef u64 W;
    static constexpr int FB = 8, S = 16, WORDS = 2, MAXN = 9;
    static constexpr W FM = 0xFFull, ONES = 0x0000010101010101ull, ALL = 0x0000FFFFFFFFFFFFull;
    static constexpr W LOWS = 0x0000010101010101ull * 0x7Full, HIGHS = 0x0000010101010101ull << 7;
    static EWN_HD int cell(int x, int y) { return (x << 4) | y; }
    static EWN_HD int x_of(int c) { return c >> 4; }
    static EWN_HD int y_of(int c) { return c & 15; }
    static EWN_HD W dead(W s) { W t = s & (s >> 4); t &= t >> 2; t &= t >> 1; return t & ONES; }
    template <int BR> static EWN_HD int dirs(const Geo& g, int c) {
        int x = c >> 4, y = c & 15;
        u32 ok0 = BR ? (y > 0) : (y < g.N - 1), ok1 = BR ? (x > 0) : (x < g.N - 1);
        return (int)(ok0 | (ok1 << 1) | ((ok0 & ok1) << 2));
    }
    static constexpr bool GENERIC = false;
    static constexpr int MAXC = 6;
    static EWN_HD W all() { return ALL; }
    static EWN_HD int get(W s, int k) { return (int)((s >> (FB * k)) & FM); }
    static EWN_HD void set(W& s, int k, int c) { s = (s & ~(FM << (FB * k))) | ((W)c << (FB * k)); }
    static EWN_HD int ncubes(const Geo&) { return 6; }
    static constexpr int LUT = 256;
    template <int BR> static EWN_HD int dirs(const GeoLut& g, int c) {
#if defined(__CUDA_ARCH__)
        u32 v; asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(g.lut + (u32)(BR * LUT) + (u32)c)); return (int)v;
#else
        return dirs<BR>(static_cast<const Geo&>(g), c);
#endif
    }
};

// Generic codec: up to 32 cubes per side, one byte per cube ((x << 4) | y, 0xFF = captured / no such cube), four u64 words per
// side in registers (words are picked with selects, never with a run-time index, so nothing spills to local memory).
struct PG {
    struct W { u64 w[4]; };
    static constexpr bool GENERIC = true;
    static constexpr int FB = 8, S = 16, WORDS = 9, MAXN = 9, MAXC = 32, LUT = 256;
    static constexpr u64 FM = 0xFFull, ONES = 0x0101010101010101ull, LOWS = ONES * 0x7Full, HIGHS = ONES << 7;
    static EWN_HD int cell(int x, int y) { return (x << 4) | y; }
    static EWN_HD int x_of(int c) { return c >> 4; }
    static EWN_HD int y_of(int c) { return c & 15; }
    template <int BR, class G> static EWN_HD int dirs(const G& g, int c) {
        int x = c >> 4, y = c & 15;
        u32 ok0 = BR ? (y > 0) : (y < g.N - 1), ok1 = BR ? (x > 0) : (x < g.N - 1);
        return (int)(ok0 | (ok1 << 1) | ((ok0 & ok1) << 2));
    }
    static EWN_HD W all() { W s; s.w[0] = s.w[1] = s.w[2] = s.w[3] = ~0ull; return s; }
    static EWN_HD int ncubes(const Geo& g) { return g.cube_num; }
    static EWN_HD u64 word(const W& s, int j) { return j == 0 ? s.w[0] : j == 1 ? s.w[1] : j == 2 ? s.w[2] : s.w[3]; }
    static EWN_HD int get(const W& s, int k) { return (int)((word(s, k >> 3) >> (8 * (k & 7))) & 0xFFull); }
    static EWN_HD void set(W& s, int k, int c) {
        const int j = k >> 3, sh = 8 * (k & 7);
        const u64 clr = ~(0xFFull << sh), ins = (u64)c << sh;
#pragma unroll
        for (int q = 0; q < 4; q++) if (q == j) s.w[q] = (s.w[q] & clr) | ins;
    }
    // bit k set <=> cube k+1 stands on the board
    static EWN_HD u32 alive(const W& s) {
        u32 dead = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            u64 t = s.w[q] & (s.w[q] >> 4); t &= t >> 2; t &= t >> 1; t &= ONES;       // bit 8i <=> byte i == 0xFF
            dead |= (u32)((t * 0x0102040810204080ull) >> 56) << (8 * q);               // gather the eight flags into one byte
        }
        return ~dead;
    }
    static EWN_HD u64 match1(u64 w, u64 rep) { u64 x = w ^ rep; return ((((x & LOWS) + LOWS) | x) & HIGHS) ^ HIGHS; }
    static EWN_HD bool any_at(const W& s, int c) {
        const u64 rep = (u64)c * ONES;
        return (match1(s.w[0], rep) | match1(s.w[1], rep) | match1(s.w[2], rep) | match1(s.w[3], rep)) != 0;
    }
    static EWN_HD void capture(W& s, int c) {                        // envs/ewn.py:254-258
        const u64 rep = (u64)c * ONES;
#pragma unroll
        for (int q = 0; q < 4; q++) { const u64 z = match1(s.w[q], rep); s.w[q] |= (z << 1) - (z >> 7); }
    }
    static EWN_HD bool all_dead(const W& s) { return (s.w[0] & s.w[1] & s.w[2] & s.w[3]) == ~0ull; }
};

template <class C> struct Game {
    typename C::W tl, br;
    int dice;      // 1..6
    int player;    // SIDE_TL / SIDE_BR: side to move
};

inline int codec_of(int N, int L) { return L > 3 ? 2 : (N <= 5 ? 0 : 1); }      // 0 = P5, 1 = P8, 2 = PG
EWN_HD void unpack_regs(Game<P5>& g, u64 v, u64) {     // from words already in registers (prefetching kernels)
    u32 lo = (u32)v, hi = (u32)(v >> 32);
    g.tl = lo & P5::ALL; g.br = hi & P5::ALL;
    g.dice = 1 + (int)((lo >> 30) | ((hi >> 28) & 4u));
    g.player = (int)(hi >> 31);
}
EWN_HD void unpack(Game<P5>& g, const u64* p, long i) { unpack_regs(g, p[i], 0); }
EWN_HD void unpack(Game<P5>& g, const u64* p, long i, const Geo&) { unpack_regs(g, p[i], 0); }
EWN_HD u64 pack_word(const Game<P5>& g) {
    u32 d = (u32)(g.dice - 1);
    u32 lo = g.tl | ((d & 3u) << 30);
    u32 hi = g.br | ((d >> 2) << 30) | ((u32)g.player << 31);
    return (u64)lo | ((u64)hi << 32);
}
EWN_HD void pack(const Game<P5>& g, u64* p, long i) { p[i] = pack_word(g); }
EWN_HD void pack(const Game<P5>& g, u64* p, long i, const Geo&) { p[i] = pack_word(g); }

EWN_HD void unpack_words(Game<P8>& g, u64 w0, u64 w1) {
    g.tl = w0 & P8::ALL; g.br = w1 & P8::ALL;
    g.dice = 1 + (int)((w0 >> 48) & 7u);
    g.player = (int)((w1 >> 48) & 1u);
}
EWN_HD void pack_words(const Game<P8>& g, u64& w0, u64& w1) {
    w0 = g.tl | ((u64)(g.dice - 1) << 48);
    w1 = g.br | ((u64)g.player << 48);
}
EWN_HD void unpack_regs(Game<P8>& g, u64 w0, u64 w1) { unpack_words(g, w0, w1); }
EWN_HD void unpack(Game<P8>& g, const u64* p, long i) {
#if defined(__CUDA_ARCH__)
    ulonglong2 v = reinterpret_cast<const ulonglong2*>(p)[i];
    unpack_words(g, v.x, v.y);
#else
    unpack_words(g, p[2 * i], p[2 * i + 1]);
#endif
}
EWN_HD void pack(const Game<P8>& g, u64* p, long i) {
    u64 w0, w1; pack_words(g, w0, w1);
#if defined(__CUDA_ARCH__)
    reinterpret_cast<ulonglong2*>(p)[i] = make_ulonglong2(w0, w1);
#else
    p[2 * i] = w0; p[2 * i + 1] = w1;
#endif
}

EWN_HD void unpack(Game<P8>& g, const u64* p, long i, const Geo&) { unpack(g, p, i); }
EWN_HD void pack(const Game<P8>& g, u64* p, long i, const Geo&) { pack(g, p, i); }

EWN_HD void unpack(Game<PG>& g, const u64* p, long i, const Geo& geo) {
    const u64* q = p + i * geo.sw;
    g.tl = PG::all(); g.br = PG::all();
#pragma unroll
    for (int j = 0; j < 4; j++) if (j < geo.cw) { g.tl.w[j] = q[j]; g.br.w[j] = q[geo.cw + j]; }
    const u64 meta = q[2 * geo.cw];
    g.dice = (int)(meta & 0xFFull); g.player = (int)((meta >> 8) & 1ull);
}
EWN_HD void pack(const Game<PG>& g, u64* p, long i, const Geo& geo) {
    u64* q = p + i * geo.sw;
#pragma unroll
    for (int j = 0; j < 4; j++) if (j < geo.cw) { q[j] = g.tl.w[j]; q[geo.cw + j] = g.br.w[j]; }
    q[2 * geo.cw] = (u64)(g.dice & 0xFF) | ((u64)(g.player & 1) << 8);
}

template <class C> EWN_HD void setup(Game<C>& g, const Geo& geo, int dice) {   // envs/ewn.py:94-107
    if constexpr (C::GENERIC) {
#pragma unroll
        for (int j = 0; j < 4; j++) { g.tl.w[j] = geo.g_tl0[j]; g.br.w[j] = geo.g_br0[j]; }
    } else { g.tl = (typename C::W)geo.tl0; g.br = (typename C::W)geo.br0; }
    g.dice = dice; g.player = SIDE_TL;
}

inline bool use_p5(int N) { return N <= 5; }
// host side: geometry + initial layout (envs/ewn.py:94-107)
inline Geo make_geo(int N, int L) {
    Geo g;
    g = Geo();
    g.N = N; g.L = L; g.cube_num = L * (L + 1) / 2;
    const bool p5 = use_p5(N);
    auto cell = [&](int x, int y) { return p5 ? P5::cell(x, y) : P8::cell(x, y); };
    g.corner = cell(N - 1, N - 1);
    if (p5) {
        for (int k = 0; k < N; k++) {
            g.tl_col |= 1u << cell(k, N - 1);   // y == N-1: TL cannot go dir 0
            g.tl_row |= 1u << cell(N - 1, k);   // x == N-1: TL cannot go dir 1
            g.br_col |= 1u << cell(k, 0);       // y == 0  : BR cannot go dir 0
            g.br_row |= 1u << cell(0, k);       // x == 0  : BR cannot go dir 1
        }
    }
    if (codec_of(N, L) == 2) {
        g.cw = (g.cube_num + 7) / 8; g.sw = 2 * g.cw + 1;
        PG::W tl = PG::all(), br = PG::all();
        int cnt = 0;
        for (int i = 1; i <= L; i++)
            for (int j = 0; j < i; j++) {
                PG::set(tl, cnt, PG::cell(j, i - j - 1));
                PG::set(br, cnt, PG::cell(N - 1 - j, N - i + j));
                cnt++;
            }
        for (int j = 0; j < 4; j++) { g.g_tl0[j] = tl.w[j]; g.g_br0[j] = br.w[j]; }
        g.tl0 = tl.w[0]; g.br0 = br.w[0];
        return g;
    }
    g.cw = 0; g.sw = p5 ? 1 : 2;
    const int fb = p5 ? P5::FB : P8::FB;
    u64 tl = p5 ? (u64)P5::ALL : P8::ALL, br = tl;
    const u64 fm = p5 ? (u64)P5::FM : P8::FM;
    int cnt = 1;
    for (int i = 1; i <= L; i++)
        for (int j = 0; j < i; j++) {
            int sh = fb * (cnt - 1);
            tl = (tl & ~(fm << sh)) | ((u64)cell(j, i - j - 1) << sh);
            br = (br & ~(fm << sh)) | ((u64)cell(N - 1 - j, N - i + j) << sh);
            cnt++;
        }
    g.tl0 = tl; g.br0 = br;
    return g;
}

// ------------------------------------------------------------------------------------------------
// rules
// ------------------------------------------------------------------------------------------------
// flag per field (at the field's top bit) where the field of `w` equals the replicated cell `rep`; exact, carry-free
template <class C> EWN_HD typename C::W match(typename C::W w, typename C::W rep) {
    typename C::W x = w ^ rep;
    typename C::W nz = (((x & C::LOWS) + C::LOWS) | x) & C::HIGHS;
    return nz ^ C::HIGHS;
}
// envs/ewn.py:254-258: whatever stands on the target cell is captured (own cubes too)
template <class C> EWN_HD typename C::W capture(typename C::W w, typename C::W rep) {
    typename C::W z = match<C>(w, rep);
    return w | ((z << 1) - (z >> (C::FB - 1)));
}

// envs/ewn.py:131-142 check_win on an arbitrary state
// is one of the side's cubes standing on `cell`? / has the side no cube left?
template <class C> EWN_HD bool any_at(const typename C::W& side, int cell) {
    if constexpr (C::GENERIC) return C::any_at(side, cell);
    else return match<C>(side, (typename C::W)cell * C::ONES) != 0;
}
template <class C> EWN_HD bool all_dead(const typename C::W& side) {
    if constexpr (C::GENERIC) return C::all_dead(side);
    else return side == C::ALL;
}
template <class C> EWN_HD bool full_win(const Game<C>& g, const Geo& geo) {
    if constexpr (C::GENERIC) return any_at<C>(g.tl, geo.corner) || any_at<C>(g.br, 0) || all_dead<C>(g.tl) || all_dead<C>(g.br);
    else {
        typedef typename C::W W;
        W at_corner = match<C>(g.tl, (W)geo.corner * C::ONES);     // board[-1,-1] > 0
        W at_origin = match<C>(g.br, (W)0);                         // board[0,0] < 0
        return (at_corner | at_origin) != 0 || g.tl == C::ALL || g.br == C::ALL;
    }
}
// winner test used after a terminal state is known (mcts.py:39-41, minimax_ewn.py:42-47): did TOP_LEFT win?
template <class C> EWN_HD bool tl_won(const Game<C>& g, const Geo& geo) {
    return any_at<C>(g.tl, geo.corner) || all_dead<C>(g.br);
}

// Shifts (bit offsets of the cube's field) of the cubes the side may move with the current die.
// envs/ewn.py:144-176 find_near_cube, :202-215 find_cube_to_move.
struct Movers { int d, hi, lo; };   // d >= 0: the dice cube itself is alive; else hi / lo = nearest larger / smaller (or -1)
template <class C> EWN_HD Movers movers(const typename C::W& side, int dice) {      // branch-free: lanes of a warp do not diverge here
    if constexpr (C::GENERIC) {                           // one alive bit per cube; shifts are byte offsets (8 * cube index)
        const u32 alive = C::alive(side);
        const int k = dice - 1;
        const bool has = (alive >> k) & 1u;
        const u32 above = k >= 31 ? 0u : alive & (~0u << (k + 1));
        const u32 below = alive & ((1u << k) - 1u);
        Movers m;
        m.d = has ? 8 * k : -1;
        m.hi = (has || !above) ? -1 : 8 * ctz_nz(above);
        m.lo = (has || !below) ? -1 : 8 * msb_nz(below);
        return m;
    } else {
    typedef typename C::W W;
    W alive = C::dead(side) ^ C::ONES;
    int sh = C::FB * (dice - 1);
    bool has = (alive >> sh) & 1;
    W above = alive & (~(W)0 << (sh + 1));
    W below = alive & (((W)1 << sh) - 1);
    Movers m;
    m.d = has ? sh : -1;
    m.hi = has ? -1 : ctz_nz(above);
    m.lo = has ? -1 : msb_nz(below);
    return m;
    }
}
// envs/ewn.py:178-215: the flag is only consulted when the dice cube is gone, and falls back to the other side
EWN_HD int cube_to_move(const Movers& m, bool larger) {
    if (m.d >= 0) return m.d;
    if (larger) return m.hi >= 0 ? m.hi : m.lo;
    return m.lo >= 0 ? m.lo : m.hi;
}

template <class C> EWN_HD int cell_at(const typename C::W& side, int sh) {
    if constexpr (C::GENERIC) return C::get(side, (sh >> 3) & 31);
    else return (int)((side >> sh) & C::FM);
}

// Ordered legal list of envs/ewn.py:338-375 as two 3-bit direction sets: `first` is listed before `second`.
//   dice cube alive : first = dirs(dice cube) with flag 0, second = empty
//   otherwise       : first = dirs(larger neighbour) with flag 1, second = dirs(smaller neighbour) with flag 0
struct Legal { int first, second, first_flag, sh_first, sh_second; };
template <class C, int BR, class G> EWN_HD Legal legal(const Game<C>& g, const G& geo, const Movers& m) {
    const typename C::W& side = BR ? g.br : g.tl;
    Legal l;
    const bool has = m.d >= 0;
    l.first_flag = has ? 0 : 1;
    l.sh_first = has ? m.d : m.hi;
    l.sh_second = m.lo;                                       // -1 when the dice cube is alive (movers)
    const int fa = C::template dirs<BR>(geo, cell_at<C>(side, l.sh_first & (int)(8 * sizeof(typename C::W) - 1)));
    const int fb = C::template dirs<BR>(geo, cell_at<C>(side, l.sh_second & (int)(8 * sizeof(typename C::W) - 1)));
    l.first = l.sh_first >= 0 ? fa : 0;
    l.second = l.sh_second >= 0 ? fb : 0;
    return l;
}
// 6-bit mask, bit = larger*3 + dir (the action id convention of alpha_zero/ewn/EWNLogic.py:328)
EWN_HD int legal_mask(const Legal& l) { return l.first_flag ? ((l.first << 3) | l.second) : l.first; }
EWN_HD int legal_count(const Legal& l) { return popc((u32)(l.first | (l.second << 3))); }
EWN_HD int nth_dir(int m, int j) {           // j-th (0-based) set bit of a 3-bit set
    if (j >= 1) m &= m - 1;
    if (j >= 2) m &= m - 1;
    return ctz_nz((u32)(m | 8));
}
// idx-th entry of the ordered list -> (larger flag, dir, shift of the cube that moves)
EWN_HD void legal_pick(const Legal& l, int idx, int& larger, int& dir, int& sh) {
    int nf = popc((u32)l.first);
    if (idx < nf) { larger = l.first_flag; dir = nth_dir(l.first, idx); sh = l.sh_first; }
    else          { larger = 0;            dir = nth_dir(l.second, idx - nf); sh = l.sh_second; }
}

// envs/ewn.py:217-263 execute_move for a move already known to stay on the board. Returns the target cell.
// dir 3 is the reference's "no branch matches" case (:229-247): (x, y) stays put, the cube is lifted and put back.
template <class C, int BR> EWN_HD int apply_move(Game<C>& g, int sh, int dir) {
    if constexpr (C::GENERIC) {
        const int k = (sh >> 3) & 31;
        const int c = C::get(BR ? g.br : g.tl, k);
        int delta = (dir != 1) + C::S * (dir != 0);        // envs/ewn.py:301-319 update_position
        delta = dir > 2 ? 0 : delta;
        const int t = BR ? c - delta : c + delta;
        C::capture(g.tl, t); C::capture(g.br, t);          // :254-258 (the mover itself too when dir 3 keeps it in place ...)
        if (BR) C::set(g.br, k, t); else C::set(g.tl, k, t);   // ... and is put back here, :259-261
        return t;
    } else {
    typedef typename C::W W;
    W own = BR ? g.br : g.tl;
    int c = (int)((own >> sh) & C::FM);
    int delta = (dir != 1) + C::S * (dir != 0);            // envs/ewn.py:301-319 update_position
    delta = dir > 2 ? 0 : delta;
    int t = BR ? c - delta : c + delta;
    W rep = (W)t * C::ONES;
    g.tl = capture<C>(g.tl, rep);
    g.br = capture<C>(g.br, rep);
    if (BR) g.br = (g.br & ~(C::FM << sh)) | ((W)t << sh);
    else    g.tl = (g.tl & ~(C::FM << sh)) | ((W)t << sh);
    return t;
    }
}
// check_win right after `BR ? BOTTOM_RIGHT : TOP_LEFT` moved to cell t from a non-terminal position
template <class C, int BR> EWN_HD bool win_after(const Game<C>& g, const Geo& geo, int t) {
    return BR ? (t == 0 || all_dead<C>(g.tl)) : (t == geo.corner || all_dead<C>(g.br));
}

// One ply with an externally chosen action (find_cube_to_move + execute_move, envs/ewn.py:439-442).
// Returns false (state untouched, :262-263) when the move leaves the board or the side has no cube.
// A direction outside {0,1,2} matches no branch of :229-247: (x, y) stays put, the bounds test passes and the cube is
// put back where it stood -- a VALID no-op move (the action space never produces it; parity keeps it).  It is folded
// into the data path as pseudo-direction 3 (always "on the board", displacement 0) so that the ply stays branch-free.
template <class C, int BR, class G> EWN_HD bool ply_action(Game<C>& g, const G& geo, int larger, int dir, int& t_out, int& sh_out) {
    const typename C::W side = BR ? g.br : g.tl;
    Movers m = movers<C>(side, g.dice);
    int sh = cube_to_move(m, larger == 1);
    sh_out = sh;
    const int d = (unsigned)dir > 2u ? 3 : dir;
    const int ok = C::template dirs<BR>(geo, cell_at<C>(side, sh & (int)(8 * sizeof(typename C::W) - 1))) | 8;
    if (sh < 0 || !((ok >> d) & 1)) return false;
    t_out = apply_move<C, BR>(g, sh, d);
    return true;
}

// The board as the other side sees it: np.rot90(-board, 2) (envs/ewn.py:289-296 opponent_action).  Cube k of one side at
// (x, y) becomes cube k of the other side at (N-1-x, N-1-y); the die is kept, the side to move is mirrored.
template <class C> EWN_HD typename C::W flip_side(const typename C::W& side, const Geo& geo) {
    typedef typename C::W W;
    if constexpr (C::GENERIC) {
        W out = C::all();
        for (int k = 0; k < geo.cube_num; k++) {
            const int c = C::get(side, k);
            if (c != (int)C::FM) C::set(out, k, geo.corner - c);
        }
        return out;
    } else {
    W out = 0;
#pragma unroll
    for (int k = 0; k < 6; k++) {
        int c = (int)((side >> (C::FB * k)) & C::FM);
        int f = c == (int)C::FM ? (int)C::FM : geo.corner - c;      // every codec: cell(N-1-x, N-1-y) = corner - cell(x, y)
        out |= (W)f << (C::FB * k);
    }
    return out;
    }
}
template <class C> EWN_HD void flip_game(Game<C>& g, const Geo& geo) {
    typename C::W tl = flip_side<C>(g.br, geo), br = flip_side<C>(g.tl, geo);
    g.tl = tl; g.br = br; g.player ^= 1;
}

// ------------------------------------------------------------------------------------------------
// leaf evaluation, envs/minimax_ewn.py:29-213 (agent is always TOP_LEFT: minimax_ewn.py:17-19)
// ------------------------------------------------------------------------------------------------
struct SideStats { int cnt, best, second; };   // best/second = two largest values of min(x,y) over alive cubes (-1 if none)
template <class C> EWN_HD SideStats side_stats(const typename C::W& side) {
    SideStats s; s.cnt = 0; s.best = -1; s.second = -1;
#pragma unroll
    for (int k = 0; k < C::MAXC; k++) {
        int c = C::get(side, k);
        if (c != (int)C::FM) {
            int x = C::x_of(c), y = C::y_of(c);
            int mn = x < y ? x : y;
            s.cnt++;
            if (mn > s.best) { s.second = s.best; s.best = mn; }
            else if (mn > s.second) s.second = mn;
        }
    }
    return s;
}
EWN_HD double inv_count(int c) {   // 1.0 / c, correctly rounded (compile-time IEEE division up to six cubes, f64_div above)
    switch (c) {
        case 1: return 1.0; case 2: return 1.0 / 2.0; case 3: return 1.0 / 3.0;
        case 4: return 1.0 / 4.0; case 5: return 1.0 / 5.0; case 6: case 0: return 1.0 / 6.0;
        default: return f64_div(1.0, (double)c);
    }
}
template <class C> EWN_HD double evaluate(const Game<C>& g, const Geo& geo, int heuristic) {
    // 'sim_winrate' is dispatched before the terminal tests (minimax_ewn.py:36-37): on a finished game simulate() plays nothing and
    // returns win_count / num_simulations = 1.0 or 0.0 (:231-233).  Live positions are evaluated by sim_unit_value, not here.
    if (heuristic == H_SIM_WINRATE) return tl_won<C>(g, geo) ? 1.0 : 0.0;
    if (tl_won<C>(g, geo)) return 10.0;                                   // minimax_ewn.py:42-44
    if (any_at<C>(g.br, 0) || all_dead<C>(g.tl)) return -10.0;            // :45-47
    SideStats p = side_stats<C>(g.tl), n = side_stats<C>(g.br);
    // Both colours measure max(N-1-x, N-1-y), the distance to the BOTTOM-RIGHT corner (minimax_ewn.py:67-76),
    // so N - min_dist = 1 + max over cubes of min(x, y).
    int N = geo.N;
    if (heuristic == H_HYBRID) {                                          // :79-86
        double a = f64_mul((double)(1 + p.best), inv_count(p.cnt));
        double b = f64_mul((double)(1 + n.best), inv_count(n.cnt));
        return f64_sub(a, b);
    }
    if (heuristic == H_MIN_DIST) return (double)((1 + p.best) - (1 + n.best));     // :126-131
    if (heuristic == H_TWO_MIN_DIST) {                                    // :168-178 sum of the (up to) two smallest distances
        int sp = (N - 1 - p.best) + (p.cnt >= 2 ? (N - 1 - p.second) : 0);
        int sn = (N - 1 - n.best) + (n.cnt >= 2 ? (N - 1 - n.second) : 0);
        return (double)(sn - sp);
    }
    return (double)(-(p.cnt + n.cnt));                                     // :212-213
}
// classical_policies/minimax.py:65-73: expectation over six identical children, e = 0; e += v/6 six times
EWN_HD double six_fold(double v) {
    double q = f64_div(v, 6.0), e = 0.0;
#pragma unroll
    for (int i = 0; i < 6; i++) e = f64_add(e, q);
    return e;
}

// 'hybrid' leaf values as a table: the heuristic depends on (1 + max min(x,y), #cubes) of either side only (minimax_ewn.py:79-86),
// i.e. on at most N*6*N*6 combinations.  vt[fold][pb][pc-1][nb][nc-1], fold = the value after minimax.py:65-73 over six identical
// leaves.  The entries are produced on the device by the very same f64_mul / f64_sub / six_fold calls, so they are bit-identical.
EWN_HD long emm_table_index(int N, int fold, int pb, int pc, int nb, int nc) {
    return ((((long)fold * N + pb) * 6 + (pc - 1)) * N + nb) * 6 + (nc - 1);
}
EWN_HD long emm_table_size(int N) { return 2L * N * 6 * N * 6; }
EWN_HD double emm_table_entry(int N, long idx) {
    int nc = 1 + (int)(idx % 6); idx /= 6;
    int nb = (int)(idx % N); idx /= N;
    int pc = 1 + (int)(idx % 6); idx /= 6;
    int pb = (int)(idx % N); idx /= N;
    double a = f64_mul((double)(1 + pb), inv_count(pc)), b = f64_mul((double)(1 + nb), inv_count(nc));
    double v = f64_sub(a, b);
    return idx ? six_fold(v) : v;
}
// value stored for a position reached by a move of side `mover_br` (over = that move won), six-folded when `fold`
template <class C> EWN_HD double leaf_value(const Game<C>& g, const Geo& geo, int heuristic, bool over, int mover_br, bool fold) {
    (void)over; (void)mover_br;
    double e = evaluate<C>(g, geo, heuristic);
    return fold ? six_fold(e) : e;
}
#if defined(__CUDA_ARCH__)
// index of a LIVE position's 'hybrid' value in vt (min(x, y) per cell from the shared-memory table behind geo.lut)
template <class C> __device__ __forceinline__ long leaf_table_index(const Game<C>& g, const GeoLut& geo, bool fold) {
    int pb = -1, nb = -1;
    const u32 mn = geo.lut + 2u * (u32)C::LUT;
#pragma unroll
    for (int k = 0; k < 6; k++) {
        int a, b;
        asm("ld.shared.s8 %0, [%1];" : "=r"(a) : "r"(mn + (u32)((g.tl >> (C::FB * k)) & C::FM)));
        asm("ld.shared.s8 %0, [%1];" : "=r"(b) : "r"(mn + (u32)((g.br >> (C::FB * k)) & C::FM)));
        pb = a > pb ? a : pb; nb = b > nb ? b : nb;
    }
    const int pc = popc(C::dead(g.tl) ^ C::ONES), nc = popc(C::dead(g.br) ^ C::ONES);
    return emm_table_index(geo.N, fold ? 1 : 0, pb, pc, nb, nc);
}
#endif
template <class C> EWN_HD double leaf_value(const Game<C>& g, const GeoLut& geo, int heuristic, bool over, int mover_br, bool fold) {
#if defined(__CUDA_ARCH__)
    if (heuristic == H_HYBRID && geo.vt) {
        // a move that does not win from a live position leaves a live position: evaluate()'s terminal tests are `over`
        if (over) return mover_br ? -10.0 : 10.0;                          // minimax_ewn.py:42-47 (never six-folded: fold implies !over)
        return __ldg(geo.vt + leaf_table_index<C>(g, geo, fold));
    }
#endif
    return leaf_value<C>(g, static_cast<const Geo&>(geo), heuristic, over, mover_br, fold);
}

// ------------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11) and the draw schedule (DESIGN.md "RNG schedule")
// ------------------------------------------------------------------------------------------------
EWN_HD void philox4x32_10(u32 c0, u32 c1, u32 c2, u32 c3, u32 k0, u32 k1, u32 out[4]) {
#pragma unroll
    for (int r = 0; r < 10; r++) {
        u32 h0 = mulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
        u32 h1 = mulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
        u32 n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
        c0 = n0; c1 = l1; c2 = n2; c3 = l0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
// same function with the ten round keys precomputed on the host (they depend on the seed only): the key schedule
// leaves the per-thread instruction stream and the keys become constant-bank operands
struct RoundKeys { u32 k[20]; };
inline RoundKeys make_round_keys(u64 seed) {
    RoundKeys r; u32 k0 = (u32)seed, k1 = (u32)(seed >> 32);
    for (int i = 0; i < 10; i++) { r.k[2 * i] = k0; r.k[2 * i + 1] = k1; k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
    return r;
}
EWN_HD void philox4x32_10_rk(u32 c0, u32 c1, u32 c2, u32 c3, const RoundKeys& rk, u32 out[4]) {
#pragma unroll
    for (int r = 0; r < 10; r++) {
        u32 h0 = mulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
        u32 h1 = mulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
        u32 n0 = h1 ^ c1 ^ rk.k[2 * r], n2 = h0 ^ c3 ^ rk.k[2 * r + 1];
        c0 = n0; c1 = l1; c2 = n2; c3 = l0;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
EWN_HD u32 bounded(u32 r, u32 k) { return mulhi(r, k); }   // uniform on [0,k) up to a bias < k / 2^32

enum { TAG_STEP = 0, TAG_RESET = 1, TAG_PLAYOUT = 2 };
// per (env, step): [0] die for the opponent's ply (envs/ewn.py:458), [1] opponent's index into its legal list
// (random_policy.py:12-14), [2] die for the next agent turn (:483) -- or, if this step ended the episode and the
// env auto-resets, the first die of the new episode (:108) --, [3] the agent's index when the agent is random too.
EWN_HD void env_block(u64 seed, u64 env_id, u64 step, u32 tag, u32 out[4]) {
    philox4x32_10((u32)env_id, (u32)(env_id >> 32), (u32)step, ((u32)(step >> 32) & 0x0FFFFFFFu) | (tag << 28),
                  (u32)seed, (u32)(seed >> 32), out);
}
EWN_HD void env_block_rk(const RoundKeys& rk, u64 env_id, u64 step, u32 tag, u32 out[4]) {
    philox4x32_10_rk((u32)env_id, (u32)(env_id >> 32), (u32)step, ((u32)(step >> 32) & 0x0FFFFFFFu) | (tag << 28), rk, out);
}
EWN_HD void playout_block_rk(const RoundKeys& rk, u32 root_id, u32 child, u32 playout, u32 j, u32 out[4]) {
    philox4x32_10_rk(playout, root_id, j, ((u32)TAG_PLAYOUT << 28) | child, rk, out);
}
// per (root, child, playout, j): plies 2j and 2j+1 of that playout; [0],[2] die, [1],[3] index into the legal list
EWN_HD void playout_block(u64 seed, u32 root_id, u32 child, u32 playout, u32 j, u32 out[4]) {
    philox4x32_10(playout, root_id, j, ((u32)TAG_PLAYOUT << 28) | child, (u32)seed, (u32)(seed >> 32), out);
}

// ------------------------------------------------------------------------------------------------
// one uniformly random ply (random_policy.py:11-15 / mcts.py:29-35). Returns true if the mover has won.
// ------------------------------------------------------------------------------------------------
template <class C, int BR, class G> EWN_HD bool random_ply(Game<C>& g, const G& geo, u32 r_index, int* larger_out, int* dir_out) {
    Movers m = movers<C>(BR ? g.br : g.tl, g.dice);
    Legal l = legal<C, BR>(g, geo, m);
    int idx = (int)bounded(r_index, (u32)legal_count(l));
    int larger, dir, sh;
    legal_pick(l, idx, larger, dir, sh);
    if (larger_out) { *larger_out = larger; *dir_out = dir; }
    int t = apply_move<C, BR>(g, sh, dir);
    return win_after<C, BR>(g, geo, t);
}

// classical_policies/mcts.py:24-41: BOTTOM_RIGHT moves first, 6-face die, uniform over the legal list.
// Returns 1 if TOP_LEFT won.
template <class C> EWN_HD int playout(Game<C> g, const Geo& geo, u64 seed, u32 root_id, u32 child, u32 index, int* plies_out) {
    int plies = 0;
    bool over = full_win<C>(g, geo);
    u32 j = 0;
    while (!over) {
        u32 blk[4];
        playout_block(seed, root_id, child, index, j++, blk);
        g.dice = 1 + (int)bounded(blk[0], 6u);
        over = random_ply<C, SIDE_BR>(g, geo, blk[1], nullptr, nullptr);
        plies++;
        if (over) break;
        g.dice = 1 + (int)bounded(blk[2], 6u);
        over = random_ply<C, SIDE_TL>(g, geo, blk[3], nullptr, nullptr);
        plies++;
    }
    if (plies_out) *plies_out = plies;
    return tl_won<C>(g, geo) ? 1 : 0;
}

// ------------------------------------------------------------------------------------------------
// MinimaxEnv.simulate (envs/minimax_ewn.py:215-238), the 'sim_winrate' leaf: num_simulations random playouts, fraction TOP_LEFT wins.
// The loop is switch_player -> roll -> random legal move (:220-229), and NOTHING restores current_player afterwards: a playout
// ends on the winner's move, so the next playout of the same env is opened by the side that just LOST.  The side that opens is
// therefore a two-state chain over the playouts of a position (opened by BOTTOM_RIGHT -> TOP_LEFT wins -> BOTTOM_RIGHT opens again
// ...), and the returned fraction is the mixture that chain produces -- not the win rate under either fixed opener.  Reproduced here:
// both openers are played for every playout index (independent Philox streams), and the chain picks, for playout j, the result under
// the opener the previous result dictates.  The first playout of a position is opened by BOTTOM_RIGHT (a fresh MinimaxEnv has
// current_player = TOP_LEFT, envs/ewn.py:71; the carry-over from the previous position's last playout is not modelled: 1 of n).
// ------------------------------------------------------------------------------------------------
// one playout from g with the given opener (1 = BOTTOM_RIGHT, 0 = TOP_LEFT); Philox stream (unit, opener, index).  Returns 1 if TOP_LEFT won.
template <class C, class G> EWN_HD int sim_playout(Game<C> g, const G& geo, const RoundKeys& rk, u32 unit, int opener_br, u32 index) {
    bool over = full_win<C>(g, geo);
    u32 j = 0;
    while (!over) {
        u32 blk[4];
        playout_block_rk(rk, unit, (u32)opener_br, index, j++, blk);
        g.dice = 1 + (int)bounded(blk[0], 6u);                            // random.randint(1, 6): six faces whatever the layout (:223)
        over = opener_br ? random_ply<C, SIDE_BR>(g, geo, blk[1], nullptr, nullptr) : random_ply<C, SIDE_TL>(g, geo, blk[1], nullptr, nullptr);
        if (over) break;
        g.dice = 1 + (int)bounded(blk[2], 6u);
        over = opener_br ? random_ply<C, SIDE_TL>(g, geo, blk[3], nullptr, nullptr) : random_ply<C, SIDE_BR>(g, geo, blk[3], nullptr, nullptr);
    }
    return tl_won<C>(g, geo) ? 1 : 0;
}
// the chain over n playouts: bits_br / bits_tl = TOP_LEFT-win bit of playout j under either opener (32 per word). -> win count
EWN_HD int sim_chain(const u32* bits_br, const u32* bits_tl, int n) {
    int wins = 0, opener_br = 1;
    for (int j = 0; j < n; j++) {
        const u32 w = ((opener_br ? bits_br : bits_tl)[j >> 5] >> (j & 31)) & 1u;
        wins += (int)w;
        opener_br = (int)w;                                               // TOP_LEFT won -> BOTTOM_RIGHT lost -> BOTTOM_RIGHT opens the next one
    }
    return wins;
}
EWN_HD int sim_words(int n) { return (n + 31) >> 5; }
// evaluate('sim_winrate') of one LIVE position, everything in one worker (host simulator; the CUDA path spreads the playouts over threads)
template <class C, class G> EWN_HD double sim_unit_value(const Game<C>& g, const G& geo, const RoundKeys& rk, u32 unit, int n) {
    u32 bb[8] = {0, 0, 0, 0, 0, 0, 0, 0}, bt[8] = {0, 0, 0, 0, 0, 0, 0, 0};          // n <= 256
    for (int j = 0; j < n; j++) {
        if (sim_playout<C>(g, geo, rk, unit, 1, (u32)j)) bb[j >> 5] |= 1u << (j & 31);
        if (sim_playout<C>(g, geo, rk, unit, 0, (u32)j)) bt[j >> 5] |= 1u << (j & 31);
    }
    return f64_div((double)sim_chain(bb, bt, n), (double)n);
}
static constexpr int SIM_MAX = 256;                  // num_simulations supported (the reference hard-codes 100, minimax_ewn.py:21)

// ------------------------------------------------------------------------------------------------
// The reference's OWN random stream: numpy's legacy global RandomState, i.e. MT19937 seeded by `np.random.seed(seed)`
// (init_genrand) and `np.random.randint(low, high)` = masked rejection sampling on 32-bit draws (numpy
// random/src/distributions: buffered_bounded_masked_uint32; a range of one value consumes no draw).  With it the whole
// eval_random.py game of a seed is reproduced from the seed alone -- dice (envs/ewn.py:90-92), agent and opponent picks
// (random_policy.py:12-14) in the consumption order of SURVEY 3.5-20 -- without any recorded log.
// ------------------------------------------------------------------------------------------------
struct MT19937 {
    u32 mt[624];
    int idx;
    EWN_HD void seed(u32 s) {
        mt[0] = s;
        for (int i = 1; i < 624; i++) mt[i] = 1812433253u * (mt[i - 1] ^ (mt[i - 1] >> 30)) + (u32)i;
        idx = 624;
    }
    EWN_HD void twist() {
        for (int k = 0; k < 624; k++) {
            const u32 y = (mt[k] & 0x80000000u) | (mt[k == 623 ? 0 : k + 1] & 0x7fffffffu);
            mt[k] = mt[k < 227 ? k + 397 : k - 227] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        idx = 0;
    }
    EWN_HD u32 next() {
        if (idx >= 624) twist();
        u32 y = mt[idx++];
        y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
        return y;
    }
    EWN_HD int randint(int low, int high) {                    // np.random.randint(low, high): uniform on [low, high)
        const u32 rng = (u32)(high - 1 - low);
        if (rng == 0) return low;
        u32 mask = rng;
        mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16;
        for (;;) { const u32 v = next() & mask; if (v <= rng) return low + (int)v; }
    }
};

// The same stream consumed across kernel launches (tournament driver with the reference's RNG): the first K tempered outputs of a
// seed are stored once (ewn_numpy_stream) and a per-game cursor walks them; randint is the same masked rejection.
struct RawStream {
    const u32* raw; int K; int cur;
    EWN_HD int randint(int low, int high) {
        const u32 rng = (u32)(high - 1 - low);
        if (rng == 0) return low;
        u32 mask = rng;
        mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16;
        while (cur < K) { const u32 v = raw[cur++] & mask; if (v <= rng) return low + (int)v; }
        return low;                                            // stream exhausted (never for a live game with K >= 512)
    }
};

// eval_random.py:17-51 for ONE seed: env.reset(seed) (np.random.seed + first die, envs/ewn.py:488-494), then RandomAgent vs the
// env's RandomAgent opponent until the game ends.  Returns +1 (agent = TOP_LEFT won) or -1; g = final position.
// trace: FNV-1a over the per-step log (dice at the agent's turn, agent action, die :458, opponent action, die :483) so that
// the whole trajectory can be compared with the recorded one, not only its result.
EWN_HD void trace_mix(u64& h, int v) { h = (h ^ (u64)(u32)(v & 0xff)) * 0x100000001b3ull; }
template <class C> EWN_HD int numpy_random_game(Game<C>& g, const Geo& geo, u32 seed, MT19937& rng, int& agent_steps, u64& trace) {
    rng.seed(seed);
    setup<C>(g, geo, rng.randint(1, geo.cube_num + 1));
    agent_steps = 0;
    trace = 0xcbf29ce484222325ull;
    for (;;) {
        int la, di, sh, t;
        {
            Legal l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
            legal_pick(l, rng.randint(0, legal_count(l)), la, di, sh);
            trace_mix(trace, g.dice); trace_mix(trace, la); trace_mix(trace, di);
            t = apply_move<C, SIDE_TL>(g, sh, di);
            agent_steps++;
            if (win_after<C, SIDE_TL>(g, geo, t)) return 1;                       // envs/ewn.py:451-454
        }
        g.player = SIDE_BR;
        g.dice = rng.randint(1, geo.cube_num + 1);                               // :458
        {
            Legal l = legal<C, SIDE_BR>(g, geo, movers<C>(g.br, g.dice));
            legal_pick(l, rng.randint(0, legal_count(l)), la, di, sh);
            trace_mix(trace, g.dice); trace_mix(trace, la); trace_mix(trace, di);
            t = apply_move<C, SIDE_BR>(g, sh, di);
            if (win_after<C, SIDE_BR>(g, geo, t)) return -1;                      // :475-478
        }
        g.player = SIDE_TL;
        g.dice = rng.randint(1, geo.cube_num + 1);                               // :483
        trace_mix(trace, g.dice);
    }
}

// ------------------------------------------------------------------------------------------------
// env step, envs/ewn.py:436-486 (+ envs/training_ewn.py:40-99 when shaped). Agent = TOP_LEFT.
// ------------------------------------------------------------------------------------------------
struct StepParams {
    Geo geo;
    double reward;                 // goal reward (envs/ewn.py:38)
    double illegal_move_reward;    // training_ewn.py:25
    u64 seed, step, env_id0;
    RoundKeys rk;                  // make_round_keys(seed)
    int32_t agent_random;          // agent = uniform over its legal list (blk[3])
    int32_t autoreset;             // SB3 VecEnv contract: re-initialise finished games in the same call
    int32_t shaped;                // MiniMaxHeuristicEnv semantics
    int32_t replay;                // dice / opponent answers injected by the caller
};
struct StepInject { int dice_opp, o_larger, o_dir, dice_next; };
struct StepOut {
    double reward; int term, trunc, code;
    int a_larger, a_dir;           // the action that was played (useful when agent_random)
};

// g: in = state at the agent's turn; out = next state (after auto-reset when the episode ended).
// term_state: the terminal state (valid when out.term).
// `geo` and the three mode flags are passed beside `p` so that a kernel can specialise them at compile time while `p`
// itself stays a read-only kernel parameter (constant bank): nothing of it is copied into registers or local memory.
template <class C, class G> EWN_HD void step_one_cfg(Game<C>& g, Game<C>& term_state, const StepParams& p, const G& geo, const int replay,
                                            const int shaped, const int agent_random, const u64 step, long i, int a_larger, int a_dir,
                                            const StepInject& inj, int* tolerance, double* prev_score, StepOut& out) {
    u32 blk[4] = {0, 0, 0, 0};
    if (!replay || agent_random) env_block_rk(p.rk, p.env_id0 + (u64)i, step, TAG_STEP, blk);
    bool pre = full_win<C>(g, geo);        // stepping an already finished game keeps reporting a win (check_win is sticky)
    out.reward = 0.0; out.term = 0; out.trunc = 0; out.code = CODE_ONGOING;
    bool valid; int t = 0, sh = -1;
    if (agent_random) {
        Movers m = movers<C>(g.tl, g.dice);
        Legal l = legal<C, SIDE_TL>(g, geo, m);
        int cnt = legal_count(l);
        valid = cnt > 0;
        if (valid) {
            legal_pick(l, (int)bounded(blk[3], (u32)cnt), a_larger, a_dir, sh);
            t = apply_move<C, SIDE_TL>(g, sh, a_dir);
        }
    } else {
        valid = ply_action<C, SIDE_TL>(g, geo, a_larger, a_dir, t, sh);
    }
    out.a_larger = a_larger; out.a_dir = a_dir;
    if (!valid) {
        if (shaped) {                                                   // training_ewn.py:48-56
            int tol = *tolerance - 1; *tolerance = tol;
            if (tol <= 0) { out.reward = -p.reward; out.term = 1; out.trunc = 1; out.code = CODE_AGENT_ILLEGAL; }
            else { out.reward = p.illegal_move_reward; out.code = CODE_ILLEGAL_TOLERATED; }
        } else { out.reward = -p.reward; out.term = 1; out.trunc = 1; out.code = CODE_AGENT_ILLEGAL; }   // ewn.py:445-448
    } else if (pre || win_after<C, SIDE_TL>(g, geo, t)) {                 // :451-454
        out.reward = p.reward; out.term = 1; out.code = CODE_AGENT_WON;
    } else {
        g.player = SIDE_BR;                                               // :457-458
        g.dice = replay ? inj.dice_opp : 1 + (int)bounded(blk[0], (u32)geo.cube_num);
        bool ovalid; int ot = 0, osh;
        if (replay) ovalid = ply_action<C, SIDE_BR>(g, geo, inj.o_larger, inj.o_dir, ot, osh);   // :461-467
        else {
            Movers m = movers<C>(g.br, g.dice);
            Legal l = legal<C, SIDE_BR>(g, geo, m);
            int cnt = legal_count(l);
            ovalid = cnt > 0;
            if (ovalid) {
                int ol, od;
                legal_pick(l, (int)bounded(blk[1], (u32)cnt), ol, od, osh);
                ot = apply_move<C, SIDE_BR>(g, osh, od);
            }
        }
        if (!ovalid) { out.reward = 0.0; out.term = 1; out.trunc = 1; out.code = CODE_OPP_ILLEGAL; }       // :470-473
        else if (win_after<C, SIDE_BR>(g, geo, ot)) { out.reward = -p.reward; out.term = 1; out.code = CODE_AGENT_LOST; }   // :475-478
        else {
            g.player = SIDE_TL;                                           // :481-483
            g.dice = replay ? inj.dice_next : 1 + (int)bounded(blk[2], (u32)geo.cube_num);
            if (shaped) {                                               // training_ewn.py:94-96
                double s = evaluate<C>(g, geo, H_HYBRID);
                out.reward = f64_sub(s, *prev_score); *prev_score = s;
            }
        }
    }
    if (out.term) {
        term_state = g;
        if (p.autoreset) setup<C>(g, geo, replay ? inj.dice_next : 1 + (int)bounded(blk[2], (u32)geo.cube_num));
    }
}

template <class C> EWN_HD void step_one(Game<C>& g, Game<C>& term_state, const StepParams& p, long i, int a_larger, int a_dir,
                                        const StepInject& inj, int* tolerance, double* prev_score, StepOut& out) {
    step_one_cfg<C>(g, term_state, p, p.geo, p.replay, p.shaped, p.agent_random, p.step, i, a_larger, a_dir, inj, tolerance, prev_score, out);
}

// ------------------------------------------------------------------------------------------------
// expectiminimax tree bookkeeping (classical_policies/minimax.py:19-73), see ewn_kernels.cu
// ------------------------------------------------------------------------------------------------
// Move level k (1-based) holds the positions reached after k moves: 6 * 36^(k-1) slots per root,
// slot index = ((..(a1)*6 + die2)*6 + a2)*6 + die3)*6 + a3 ...   (a = position in the ordered legal list, die 0..5)
EWN_HD long emm_level_size(int k) { long s = 6; for (int i = 1; i < k; i++) s *= 36; return s; }
EWN_HD long emm_level_offset(int k) { long o = 0; for (int i = 1; i < k; i++) o += emm_level_size(i); return o; }
EWN_HD int emm_levels(int depth) { return (depth + 1) / 2; }
// node records are doubles; two NaN payloads mark "no such child" and "non-terminal interior node"
static constexpr u64 EMM_ABSENT = 0xFFFFFFFFFFFFFFFFull;
static constexpr u64 EMM_INTERIOR = 0xFFF8000000000001ull;

EWN_HD double bits_to_f64(u64 b) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)b);
#else
    double d; __builtin_memcpy(&d, &b, 8); return d;
#endif
}
EWN_HD u64 f64_to_bits(double d) {
#if defined(__CUDA_ARCH__)
    return (u64)__double_as_longlong(d);
#else
    u64 b; __builtin_memcpy(&b, &d, 8); return b;
#endif
}
EWN_HD double f64_inf() { return bits_to_f64(0x7FF0000000000000ull); }

// Record codecs.  double: the value itself (any heuristic).  uint16_t ('hybrid' with the value table vt, device only): the
// table index of a live position, or one of four marks -- a quarter of the bytes, and sixteen records per 32-byte sector.
static constexpr uint16_t EMM16_ABSENT = 0xFFFF, EMM16_INTERIOR = 0xFFFE, EMM16_WIN = 0xFFFD, EMM16_LOSS = 0xFFFC;
static constexpr u64 EMM_PENDING = 0xFFF8000000000002ull;      // 'sim_winrate' leaf whose playouts have not been played yet (| 4: six-fold leaf)
EWN_HD void emm_mark(double* p, bool absent) { *p = bits_to_f64(absent ? EMM_ABSENT : EMM_INTERIOR); }
EWN_HD void emm_mark(uint16_t* p, bool absent) { *p = absent ? EMM16_ABSENT : EMM16_INTERIOR; }
EWN_HD bool emm_is_absent(double r) { return f64_to_bits(r) == EMM_ABSENT; }
EWN_HD bool emm_is_absent(uint16_t r) { return r == EMM16_ABSENT; }
EWN_HD bool emm_is_interior(double r) { return f64_to_bits(r) == EMM_INTERIOR; }
EWN_HD bool emm_is_interior(uint16_t r) { return r == EMM16_INTERIOR; }
EWN_HD double emm_value(double r, const double*) { return r; }
EWN_HD double emm_value(uint16_t r, const double* vt) {
    if (r >= EMM16_LOSS) return r == EMM16_WIN ? 10.0 : -10.0;
#if defined(__CUDA_ARCH__)
    return __ldg(vt + r);
#else
    return vt[r];
#endif
}
template <class C, class G> EWN_HD void emm_leaf(double* p, const Game<C>& g, const G& geo, int heuristic, bool over, int mover_br, bool fold) {
    *p = leaf_value<C>(g, geo, heuristic, over, mover_br, fold);
}
// 'sim_winrate' search: geometry + where the leaf POSITIONS go (a u64[WORDS] slot per record, parallel to the record array); the
// leaf's value is filled in after its playouts have been played (ewn_expectiminimax_sim)
template <class Base> struct GeoSimT : Base { u64* sim_states; const double* sim_base; };
template <class C, class Base> EWN_HD void emm_leaf(double* p, const Game<C>& g, const GeoSimT<Base>& geo, int, bool over, int mover_br, bool fold) {
    if (over) { *p = mover_br ? 0.0 : 1.0; return; }                      // simulate() on a finished game: the mover has just won
    *p = bits_to_f64(EMM_PENDING | (fold ? 4ull : 0ull));
    Game<C> t = g; t.dice = 1; t.player = SIDE_TL;
    pack(t, geo.sim_states, (long)(p - geo.sim_base), static_cast<const Geo&>(geo));
}
#if defined(__CUDA_ARCH__)
template <class C> __device__ __forceinline__ void emm_leaf(uint16_t* p, const Game<C>& g, const GeoLut& geo, int, bool over, int mover_br, bool fold) {
    *p = over ? (mover_br ? EMM16_LOSS : EMM16_WIN) : (uint16_t)leaf_table_index<C>(g, geo, fold);
}
#endif

// Phase 1, one call per PARENT tp of the deepest move level of one root (g = non-terminal root, TOP_LEFT to move; the level
// has 6 * 36^(m-1) slots, slot = tp * 6 + a): walk the path root -> parent once, write the record of every earlier position this
// path owns, build the parent's legal list once and write the records of its (up to six) children.
//   last level      : value of the chance node on that position (evaluate, six-folded when depth is even and the
//                     game is not over: minimax.py:22-23 + :64-73 over six identical depth-0 leaves)
//   earlier levels  : EMM_INTERIOR, or evaluate() if the game ended there (minimax.py:22-23)
//   a >= #legal     : EMM_ABSENT
template <class C, int SIDE, class G, class R> EWN_HD void emm_expand_children(const Game<C>& g, const G& geo, int depth, int heuristic, R* out) {
    const Legal l = legal<C, SIDE>(g, geo, movers<C>(SIDE == SIDE_TL ? g.tl : g.br, g.dice));
    const int cnt = legal_count(l);
    for (int a = 0; a < 6; a++) {
        if (a >= cnt) { emm_mark(out + a, true); continue; }
        Game<C> c = g;
        int la, di, sh; legal_pick(l, a, la, di, sh);
        const int tc = apply_move<C, SIDE>(c, sh, di);
        const bool over = win_after<C, SIDE>(c, geo, tc);
        emm_leaf<C>(out + a, c, geo, heuristic, over, SIDE == SIDE_TL ? 0 : 1, !(depth & 1) && !over);
    }
}

template <class C, class G, class R> EWN_HD void emm_expand_parent(Game<C> g, const G& geo, int depth, int heuristic, u32 tp, R* v) {
    const int m = emm_levels(depth);
    const u32 t = tp * 6u;                         // first slot of this parent at the last level
    u32 pow36 = (u32)emm_level_size(m) / 6;        // 36^(m-k) at level k; t % pow36 == 0 <=> this path owns its level-k node
    long off = 0, lvl = 6;
    for (int k = 1; k < m; k++) {
        const u32 idx = t / pow36;                 // node index at level k
        const bool owner = (t - idx * pow36) == 0;
        const int a = (int)(idx % 6);
        if (k > 1) g.dice = 1 + (int)((idx / 6) % 6);
        bool over = false; int cnt;
        if (k & 1) {
            Legal l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
            cnt = legal_count(l);
            if (a < cnt) { int la, di, sh; legal_pick(l, a, la, di, sh); int tc = apply_move<C, SIDE_TL>(g, sh, di); over = win_after<C, SIDE_TL>(g, geo, tc); }
        } else {
            Legal l = legal<C, SIDE_BR>(g, geo, movers<C>(g.br, g.dice));
            cnt = legal_count(l);
            if (a < cnt) { int la, di, sh; legal_pick(l, a, la, di, sh); int tc = apply_move<C, SIDE_BR>(g, sh, di); over = win_after<C, SIDE_BR>(g, geo, tc); }
        }
        if (a >= cnt) { if (owner) emm_mark(v + off + idx, true); return; }
        if (over) { if (owner) emm_leaf<C>(v + off + idx, g, geo, heuristic, true, (k & 1) ? 0 : 1, false); return; }
        if (owner) emm_mark(v + off + idx, false);
        off += lvl; lvl *= 36; pow36 /= 36;
    }
    if (m > 1) g.dice = 1 + (int)(tp % 6);         // (slot / 6) % 6 of every child
    if (m & 1) emm_expand_children<C, SIDE_TL, G, R>(g, geo, depth, heuristic, v + off + t);
    else       emm_expand_children<C, SIDE_BR, G, R>(g, geo, depth, heuristic, v + off + t);
}

// Phase 2: classical_policies/minimax.py:19-73 replayed over the stored records of one root.
template <int K, int M> struct Emm {
    static constexpr long OFF = (K == 1) ? 0 : (K == 2) ? 6 : (K == 3) ? 222 : 7998;
    // value of the chance node sitting on position `idx` of level K
    template <class R> static EWN_HD double chance(const R* v, const double* vt, long idx, double alpha, double beta, int pruned) {
        const R rec = v[OFF + idx];
        if (K == M) return emm_value(rec, vt);
        if (!emm_is_interior(rec)) return emm_value(rec, vt);         // finished game: minimax.py:22-23
        double e = 0.0;
        for (int die = 0; die < 6; die++) {                           // minimax.py:64-73
            double val = Emm<K + 1, M>::move(v, vt, idx * 6 + die, alpha, beta, pruned, nullptr);
            e = f64_add(e, f64_div(val, 6.0));
        }
        return e;
    }
    // value of the move node whose children are positions base*6 + a of level K (K odd: TOP_LEFT maximises)
    template <class R> static EWN_HD double move(const R* v, const double* vt, long base, double alpha, double beta, int pruned, int* best_a) {
        if (K & 1) {                                                  // minimax.py:26-43
            double best = -f64_inf();
            for (int a = 0; a < 6; a++) {
                if (emm_is_absent(v[OFF + base * 6 + a])) break;
                double val = chance(v, vt, base * 6 + a, alpha, beta, pruned);
                if (val > best) { best = val; if (best_a) *best_a = a; }
                alpha = alpha > best ? alpha : best;
                if (pruned && beta <= alpha) break;
            }
            return best;
        } else {                                                      // minimax.py:45-62
            double worst = f64_inf();
            for (int a = 0; a < 6; a++) {
                if (emm_is_absent(v[OFF + base * 6 + a])) break;
                double val = chance(v, vt, base * 6 + a, alpha, beta, pruned);
                if (val < worst) worst = val;
                beta = beta < worst ? beta : worst;
                if (pruned && beta <= alpha) break;
            }
            return worst;
        }
    }
};
template <int M> struct Emm<5, M> {
    template <class R> static EWN_HD double chance(const R*, const double*, long, double, double, int) { return 0.0; }
    template <class R> static EWN_HD double move(const R*, const double*, long, double, double, int, int*) { return 0.0; }
};

// ExpectiMinimaxAgent.predict (minimax.py:89-93) for one root; (la, di) = (-1,-1) when the root is finished (action None)
template <class C, int M, class R> EWN_HD double emm_root(const Game<C>& g, const Geo& geo, int heuristic, int pruned, const R* v,
                                                           int& la, int& di, const double* vt = nullptr) {
    la = -1; di = -1;
    if (full_win<C>(g, geo)) return evaluate<C>(g, geo, heuristic);   // minimax.py:22-23
    int best_a = 0;
    double val = Emm<1, M>::move(v, vt, 0, -f64_inf(), f64_inf(), pruned, &best_a);
    Legal l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
    int sh; legal_pick(l, best_a, la, di, sh);
    return val;
}

// ------------------------------------------------------------------------------------------------
// Fused depth-3 / depth-4 expectiminimax (the closed-form prefix rule).  Tree of minimax.py:19-73 at these depths:
//   root MAX (TOP_LEFT, <= 6 moves a1)  ->  CHANCE (6 faces d2)  ->  MIN (BOTTOM_RIGHT, <= 6 moves a2)  ->  leaf
//   leaf = evaluate (depth 3), or the chance node over six identical depth-0 evaluations (depth 4: six_fold), or +-10 when the
//   move ended the game.  Leaf values do not depend on the (alpha, beta) window, beta is +inf everywhere (the root never lowers
//   it), so the only cut is the MIN node's `beta <= alpha` against the ROOT's running alpha: a MIN node returns its first
//   prefix-minimum <= alpha, else its full minimum (minimax.py:53-58).  alpha only changes between root children, so the a1 loop
//   is the one sequential dimension; the six dice of one root child are independent given alpha.
// Leaves are not materialised: a MIN-level move changes one cube of the mover and captures at most one cube, and every heuristic
// of envs/minimax_ewn.py depends only on the per-side multiset of min(x, y) over alive cubes, kept as a histogram word.
// ------------------------------------------------------------------------------------------------
// histogram word of one side: 3-bit field v = number of alive cubes with min(x, y) == v (v <= 8, count <= 6)
EWN_HD u32 hist_unit(int mn) { return 1u << (3 * mn); }
EWN_HD int hist_best(u32 h) { return (msb_nz(h | 1u) * 11) >> 5; }                          // largest v with a non-zero field (0 if none)
EWN_HD int hist_second(u32 h) { return hist_best(h - hist_unit(hist_best(h))); }             // second largest, multiplicity counted

// per-block tables of the fused kernel (shared memory on the device, plain arrays on the host):
//   inc[cell] = hist_unit(min(x, y)), 0 for the captured sentinel and for cells off the board
//   v / q     = 'hybrid' leaf values of the fold the depth needs and their quotients v / 6 (what minimax.py:72 adds up; looked up,
//               so the hot loop has no fp64 division), index [pb][pc-1][nb][nc-1]; both pointers are EMM_PAD(N) entries into
//               their arrays so that the summaries of slots that are not legal (pc - 1 = -1 ...) still index allocated memory;
//               entry -1 holds the value of a lost game (-10).  v == nullptr: values are computed in place.
struct EmmTabs { const u32* inc; const double* v; const double* q; };
EWN_HD int emm_pad(int N) { return 12 * N + 8; }
EWN_HD int emm_tab_entries(int N) { return N * 6 * N * 6; }
template <class C> EWN_HD u32 emm_inc_entry(int N, int c) {
    const int x = C::x_of(c), y = C::y_of(c);
    if (c == (int)C::FM || x >= N || y >= N) return 0u;
    return hist_unit(x < y ? x : y);
}
template <class C> EWN_HD u32 side_hist(typename C::W side, const EmmTabs& tb) {
    u32 h = 0;
#pragma unroll
    for (int k = 0; k < 6; k++) h += tb.inc[(int)((side >> (C::FB * k)) & C::FM)];
    return h;
}
// occupancy bit board of one side (P5 only: cell < 32; bit 31 collects the captured cubes and is never a target)
EWN_HD u32 side_occ(u32 side) {
    u32 o = 0;
#pragma unroll
    for (int k = 0; k < 6; k++) o |= 1u << ((side >> (5 * k)) & 31u);
    return o;
}
EWN_HD u32 side_occ(u64) { return 0u; }
// envs/minimax_ewn.py:29-213 on the summary of a LIVE position (N - min_dist = 1 + best), computed in place;
// fold = minimax.py:64-73 over six equal leaves
EWN_HD double emm_summary_value(int heuristic, int N, u32 htl, int pc, u32 hbr, int nc, bool fold) {
    const int pb = hist_best(htl), nb = hist_best(hbr);
    double v;
    if (heuristic == H_HYBRID) v = f64_sub(f64_mul((double)(1 + pb), inv_count(pc)), f64_mul((double)(1 + nb), inv_count(nc)));
    else if (heuristic == H_MIN_DIST) v = (double)(pb - nb);
    else if (heuristic == H_TWO_MIN_DIST) {
        const int sp = (N - 1 - pb) + (pc >= 2 ? (N - 1 - hist_second(htl)) : 0);
        const int sn = (N - 1 - nb) + (nc >= 2 ? (N - 1 - hist_second(hbr)) : 0);
        v = (double)(sn - sp);
    } else v = (double)(-(pc + nc));
    return fold ? six_fold(v) : v;
}

// what the six (root child, die) workers need to know about one root child: the position after TOP_LEFT's move and its summaries.
// meta: bits 0-15 = 1024 + (pc - 1) * 6N + (nc - 1) (the count part of the value-table index), 16-18 = pc, 19-21 = nc,
//       30 = the move ended the game (TOP_LEFT won), 31 = the child exists (a1 < #legal)
template <class C> struct EmmChild {
    typename C::W tl, br;
    u32 htl, hbr, otl, obr, meta;
};
static constexpr u32 EMMC_EXISTS = 0x80000000u, EMMC_OVER = 0x40000000u;

template <class C, class G> EWN_HD EmmChild<C> emm2_child(const Game<C>& root, const G& geo, const EmmTabs& tb, int k) {
    EmmChild<C> ch;
    Game<C> g = root;
    const Legal l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
    u32 flags = 0;
    if (k < legal_count(l)) {
        int la, di, sh; legal_pick(l, k, la, di, sh);
        const int t = apply_move<C, SIDE_TL>(g, sh, di);
        flags = EMMC_EXISTS | (win_after<C, SIDE_TL>(g, geo, t) ? EMMC_OVER : 0u);
    }
    ch.tl = g.tl; ch.br = g.br;
    ch.htl = side_hist<C>(g.tl, tb); ch.hbr = side_hist<C>(g.br, tb);
    ch.otl = side_occ(g.tl); ch.obr = side_occ(g.br);
    const int pc = popc(C::dead(g.tl) ^ C::ONES), nc = popc(C::dead(g.br) ^ C::ONES);
    ch.meta = (u32)(1024 + (pc - 1) * 6 * geo.N + (nc - 1)) | ((u32)pc << 16) | ((u32)nc << 19) | flags;
    return ch;
}

// MIN node under root child `ch` with die `die` (1..6): minimax.py:45-62 with beta = +inf on entry.  `alpha` = the root's running
// alpha (-inf for the unpruned search).  Six fixed slots = (larger neighbour | dice cube, smaller neighbour) x direction, in the
// order of the legal list (envs/ewn.py:338-375); slots that are not legal never update the minimum.  Returns the node's value
// and, through q, value / 6.
template <bool TAB, class C, class G> EWN_HD double emm2_min(const EmmChild<C>& ch, const G& geo, const EmmTabs& tb, int die, int heuristic, bool fold,
                                                              double alpha, double& q) {
    typedef typename C::W W;
    Game<C> g; g.tl = ch.tl; g.br = ch.br; g.dice = die; g.player = SIDE_BR;
    const Legal l = legal<C, SIDE_BR>(g, geo, movers<C>(g.br, die));
    const int N = geo.N;
    const int kbase = (int)(ch.meta & 0xFFFFu) - 1024, pc = (int)((ch.meta >> 16) & 7u), nc = (int)((ch.meta >> 19) & 7u);
    double worst = f64_inf();
    int worst_i = -1;
    bool cut = false;
#pragma unroll
    for (int s = 0; s < 2; s++) {
        const int set = s ? l.second : l.first, sh = s ? l.sh_second : l.sh_first;
#if defined(__CUDA_ARCH__)
        if (!__any_sync(0xffffffffu, set != 0)) continue;                     // every lane of the warp calls this function together (k_emm_fused)
#endif
        const int c = cell_at<C>(g.br, sh & (int)(8 * sizeof(W) - 1));
        const u32 inc_c = sh >= 0 ? tb.inc[c] : 0u;                            // no such mover: keep the histogram arithmetic borrow-free
#pragma unroll
        for (int dir = 0; dir < 3; dir++) {
            const int delta = (dir != 1) + C::S * (dir != 0);                 // envs/ewn.py:301-319, BOTTOM_RIGHT moves towards (0, 0)
            const int t = (c - delta) & (C::LUT - 1);                          // in range for every legal slot; wrapped garbage otherwise
            bool cap_t, cap_b;                                                 // envs/ewn.py:254-258: whatever stands on t is captured
            if (C::WORDS == 1) { cap_t = (ch.otl >> t) & 1u; cap_b = (ch.obr >> t) & 1u; }
            else { const W rep = (W)t * C::ONES; cap_t = match<C>(g.tl, rep) != 0; cap_b = match<C>(g.br, rep) != 0; }
            const u32 inc_t = tb.inc[t];
            const u32 h_tl = ch.htl - (cap_t ? inc_t : 0u);
            const u32 h_br = ch.hbr - inc_c + (cap_b ? 0u : inc_t);
            const bool over = t == 0 || (cap_t && pc == 1);                    // check_win after BOTTOM_RIGHT's move
            double v; int idx = -1;
            if (TAB) {
                idx = hist_best(h_tl) * (36 * N) + hist_best(h_br) * 6 + kbase - (cap_t ? 6 * N : 0) - (cap_b ? 1 : 0);
                idx = over ? -1 : idx;                                         // entry -1: minimax_ewn.py:45-47, -10
#if defined(EWN_CHECKS) && defined(__CUDA_ARCH__)
                if (idx < -emm_pad(N) || idx >= emm_tab_entries(N)) { printf("EWN_CHECK failed: value-table index %d\n", idx); __trap(); }
#endif
                v = tb.v[idx];
            } else {
                v = over ? -10.0 : emm_summary_value(heuristic, N, h_tl, pc - (cap_t ? 1 : 0), h_br, nc - (cap_b ? 1 : 0), fold);
            }
            const bool take = ((set >> dir) & 1) && !cut;
            if (take && v < worst) { worst = v; worst_i = idx; }
            cut = cut || (take && worst <= alpha);
        }
    }
    q = TAB ? tb.q[worst_i] : f64_div(worst, 6.0);
    return worst;
}

// ExpectiMinimaxAgent.predict at depth 3 or 4 for one root, one worker doing everything (host simulator; the CUDA kernel spreads the
// same calls over six lanes per root, see k_emm_fused)
template <class C, class G> EWN_HD double emm2_root_serial(const Game<C>& root, const G& geo, const EmmTabs& tb, int depth, int heuristic, int pruned,
                                                            int& la, int& di) {
    la = -1; di = -1;
    if (full_win<C>(root, geo)) return evaluate<C>(root, geo, heuristic);       // minimax.py:22-23
    const bool fold = !(depth & 1);
    double best = -f64_inf(), alpha = -f64_inf();
    int best_a = 0;
    for (int a1 = 0; a1 < 6; a1++) {
        const EmmChild<C> ch = emm2_child<C>(root, geo, tb, a1);
        if (!(ch.meta & EMMC_EXISTS)) break;
        double e = 0.0;
        if (ch.meta & EMMC_OVER) e = 10.0;                                        // minimax.py:22-23 at the chance node: TOP_LEFT has won
        else
            for (int die = 1; die <= 6; die++) {                                  // minimax.py:64-73
                double q;
                if (tb.v) emm2_min<true, C>(ch, geo, tb, die, heuristic, fold, pruned ? alpha : -f64_inf(), q);
                else      emm2_min<false, C>(ch, geo, tb, die, heuristic, fold, pruned ? alpha : -f64_inf(), q);
                e = f64_add(e, q);
            }
        if (e > best) { best = e; best_a = a1; }
        alpha = alpha > best ? alpha : best;
    }
    const Legal l = legal<C, SIDE_TL>(root, geo, movers<C>(root.tl, root.dice));
    int sh; legal_pick(l, best_a, la, di, sh);
    return best;
}

// ------------------------------------------------------------------------------------------------
// UCT at the root (the intent of the reference's unfinished ucth_temp.py:6-89: expand the root's legal moves, pick the child with the best
// UCB1 score, simulate from it, back the result up; final move = best win rate).  Batched: every round each root picks ONE child and
// plays `batch` random playouts from it (mcts.py:24-41 playouts, BOTTOM_RIGHT to move), wins / visits are kept per legal-list position.
//   score = wins / visits + c * sqrt(ln(total visits) / visits)   (ucth_temp.py:37-40, c = 1.41), unvisited children first, ties -> first.
// ln(total) is passed in by the host (one libm, the same for the CUDA path and the host build); division and sqrt are IEEE-exact.
// ------------------------------------------------------------------------------------------------
EWN_HD double f64_sqrt(double a) {
#if defined(__CUDA_ARCH__)
    return __dsqrt_rn(a);
#else
    volatile double r = __builtin_sqrt(a); return r;
#endif
}
// -> list position of the child to simulate next (n_children >= 1)
EWN_HD int uct_select(const u32* wins, const u32* visits, int n_children, double c, double ln_total) {
    int best = 0; double best_score = -f64_inf();
    for (int j = 0; j < n_children; j++) {
        double score;
        if (visits[j] == 0) score = f64_inf();                                   // ucth_temp.py:35-36
        else {
            const double v = (double)visits[j];
            score = f64_add(f64_div((double)wins[j], v), f64_mul(c, f64_sqrt(f64_div(ln_total, v))));
        }
        if (score > best_score) { best_score = score; best = j; }
    }
    return best;
}
// final move: the child with the best win rate (ucth_temp.py:76), first maximum; children never visited count as 0
EWN_HD int uct_best(const u32* wins, const u32* visits, int n_children) {
    int best = 0; double bv = -1.0;
    for (int j = 0; j < n_children; j++) {
        const double r = visits[j] ? f64_div((double)wins[j], (double)visits[j]) : 0.0;
        if (r > bv) { bv = r; best = j; }
    }
    return best;
}
// One round for one root: fold the previous round's result in, pick the next child, hand its position out.  Returns the chosen list
// position (or -1 for a finished root).  g = root (TOP_LEFT to move), child = position after the chosen move (BOTTOM_RIGHT to move).
template <class C, class G> EWN_HD int uct_round(const Game<C>& g, const G& geo, u32* wins, u32* visits, int prev, u32 prev_wins, u32 batch,
                                                  double c, double ln_total, Game<C>& child, int& n_children) {
    n_children = 0;
    if (full_win<C>(g, geo)) return -1;
    const Legal l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
    n_children = legal_count(l);
    if (prev >= 0) { wins[prev] += prev_wins; visits[prev] += batch; }
    const int j = uct_select(wins, visits, n_children, c, ln_total);
    int la, di, sh; legal_pick(l, j, la, di, sh);
    child = g;
    apply_move<C, SIDE_TL>(child, sh, di);
    child.player = SIDE_BR;
    return j;
}

}  // namespace ewn
