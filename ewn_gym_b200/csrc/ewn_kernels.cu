// ewn_kernels.cu -- sm_100a kernels and the C ABI (include/ewn_b200.h) of the EinStein-wuerfelt-nicht engine.
//
//   K0  k_reset / k_encode / k_decode     packed-state codec                        (envs/ewn.py:94-108, minimax.py:75-87)
//   K1  k_step                            fused env step, one thread per game       (envs/ewn.py:436-486, training_ewn.py:40-99)
//   K2  k_legal / k_ply / k_random_action legal list, single ply, random agent      (envs/ewn.py:338-375, random_policy.py:11-15)
//   K3  k_mc_expand / k_playouts          flat Monte-Carlo playouts, lane per game  (classical_policies/mcts.py:21-84)
//   K4  k_emm_expand / k_emm_replay       expectiminimax, level-parallel expansion  (classical_policies/minimax.py:19-93)
//
// All of it is integer / bit manipulation on the packed piece list of ewn_core.cuh; nothing here is a contraction,
// so there is no tensor-core path.  K1 streams 8 (5x5) or 16 (<=9x9) bytes of state per game with fully coalesced
// vector accesses and stages observation rows through shared memory so that global stores are full 16-byte vectors.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <mutex>
#include <new>

#include "../../include/ewn_b200.h"
#include "ewn_core.cuh"

using namespace ewn;

namespace {

constexpr int TPB = 256;               // threads per block for the per-game kernels
constexpr int WARPS = TPB / 32;

thread_local char g_err[256] = "";
std::atomic<long long> g_launches{0};

// -DEWN_CHECKS: bounds assertions of our own on the hand-rolled indexing (compute-sanitizer is closed on the GPU pool: profiles/
// r2_sanitize_unavailable.txt).  A violated check traps, the launch fails with an error and every test that follows fails loudly.
// tools/checked_build.sh builds build/checked/libewn_b200.so; EWN_B200_LIB=<that file> makes the test-suite load it.
#if defined(EWN_CHECKS)
#define EWN_CHECK(cond) do { if (!(cond)) { printf("EWN_CHECK failed: %s (%s:%d) block %d thread %d\n", #cond, __FILE__, __LINE__, (int)blockIdx.x, (int)threadIdx.x); __trap(); } } while (0)
#else
#define EWN_CHECK(cond) do { } while (0)
#endif

inline int cuda_fail(cudaError_t e, const char* what) {
    snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
    return EWN_ECUDA;
}
#define CU(call)                                            \
    do {                                                    \
        cudaError_t e__ = (call);                           \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call); \
    } while (0)
#define LAUNCHED(name)                                      \
    do {                                                    \
        g_launches.fetch_add(1, std::memory_order_relaxed); \
        cudaError_t e__ = cudaGetLastError();               \
        if (e__ != cudaSuccess) return cuda_fail(e__, name); \
    } while (0)

// envs/ewn.py:47: any cube_layer < board_size - 1 (up to 28 cubes per side on 9x9)
inline bool config_ok(int N, int L) { return N >= 3 && N <= EWN_MAX_BOARD && L >= 1 && L < N - 1; }
inline unsigned blocks_for(int64_t n) { return (unsigned)((n + TPB - 1) / TPB); }

// ------------------------------------------------------------------------------------------------
// warp-cooperative observation writer: lanes hold one game each, rows leave as coalesced 16-byte vectors
// ------------------------------------------------------------------------------------------------
// NT = board size known at compile time (0: read it from geo) -- the hot configurations (5x5, 7x7) get fully unrolled
// staging loops and, for 5x5, a packed cell index that already is the row-major board index.
template <class C, int NT> __device__ __forceinline__ int lin_cell(const Geo& geo, int c) {
    const int N = NT ? NT : geo.N;
    return (C::S == N) ? c : C::x_of(c) * N + C::y_of(c);
}

template <class C, int NT> __device__ __forceinline__ void stage_rows(const Game<C>& g, bool mine, const Geo& geo, int16_t* sm, int lane) {
    const int NN = NT ? NT * NT : geo.N * geo.N;
    uint4* v = reinterpret_cast<uint4*>(sm);
    const int nvec = 4 * NN;                     // 32 rows * NN * 2 bytes / 16
    if (NT) {
#pragma unroll
        for (int k = 0; k < (4 * NT * NT + 31) / 32; k++) if (lane + 32 * k < nvec) v[lane + 32 * k] = make_uint4(0, 0, 0, 0);
    } else {
        for (int k = lane; k < nvec; k += 32) v[k] = make_uint4(0, 0, 0, 0);
    }
    __syncwarp();
    if (mine) {
        int16_t* row = sm + lane * NN;
        const int nc = C::ncubes(geo);
#pragma unroll
        for (int k = 0; k < C::MAXC; k++) {
            if (k >= nc) break;
            int c = C::get(g.tl, k);
            if (c != (int)C::FM) row[lin_cell<C, NT>(geo, c)] = (int16_t)(k + 1);
            int d = C::get(g.br, k);
            if (d != (int)C::FM) row[lin_cell<C, NT>(geo, d)] = (int16_t)(-(k + 1));
        }
    }
    __syncwarp();
}
// dense: rows of all lanes with env index < n.  warp_env0 = env index of lane 0 (a multiple of 32).
template <class C, int NT> __device__ __forceinline__ void write_rows_dense(int16_t* out, int64_t warp_env0, int64_t n, const Geo& geo,
                                                                            const int16_t* sm, int lane) {
    const int NN = NT ? NT * NT : geo.N * geo.N;
    int64_t rows = n - warp_env0; if (rows > 32) rows = 32;
    int16_t* dst = out + warp_env0 * NN;
    if (rows == 32 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
        const uint4* v = reinterpret_cast<const uint4*>(sm);
        uint4* d = reinterpret_cast<uint4*>(dst);
        if (NT) {
#pragma unroll
            for (int k = 0; k < (4 * NT * NT + 31) / 32; k++) if (lane + 32 * k < 4 * NN) __stcs(d + lane + 32 * k, v[lane + 32 * k]);
        } else {
            for (int k = lane; k < 4 * NN; k += 32) __stcs(d + k, v[k]);
        }
    } else {
        const int total = (int)rows * NN;
        for (int k = lane; k < total; k += 32) dst[k] = sm[k];
    }
    __syncwarp();
}
// sparse: only rows whose bit is set in `rows`; the k-th such row goes to row dst_row0 + (compact ? k : its lane)
template <class C, int NT> __device__ __forceinline__ void write_rows_sparse(int16_t* out, int64_t dst_row0, bool compact, unsigned rows,
                                                                             const Geo& geo, const int16_t* sm, int lane) {
    const int NN = NT ? NT * NT : geo.N * geo.N;
    int k = 0;
    while (rows) {
        int r = __ffs((int)rows) - 1; rows &= rows - 1;
        int16_t* dst = out + (dst_row0 + (compact ? k : r)) * NN;
        for (int e = lane; e < NN; e += 32) dst[e] = sm[r * NN + e];
        k++;
    }
    __syncwarp();
}

// ------------------------------------------------------------------------------------------------
// K0 codec
// ------------------------------------------------------------------------------------------------
template <class C> __global__ void __launch_bounds__(TPB) k_reset(u64* state, const int32_t* dice_in, int64_t n, Geo geo, u64 seed,
                                                                  u64 step, u64 env_id0) {
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    int dice;
    if (dice_in) { dice = dice_in[i]; dice = dice < 1 ? 1 : (dice > geo.cube_num ? geo.cube_num : dice); }   // a die outside 1..cube_num would spill into the flag bits
    else { u32 blk[4]; env_block(seed, env_id0 + (u64)i, step, TAG_RESET, blk); dice = 1 + (int)bounded(blk[0], (u32)geo.cube_num); }
    Game<C> g; setup<C>(g, geo, dice);
    pack(g, state, i, geo);
}

template <class C> __global__ void __launch_bounds__(TPB) k_encode(const int16_t* board, const int32_t* dice, const int32_t* player,
                                                                   u64* state, int64_t n, Geo geo) {
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    const int N = geo.N, NN = N * N, nc = geo.cube_num;
    Game<C> g; g.tl = C::all(); g.br = C::all();
    const int16_t* b = board + i * NN;
    for (int x = 0; x < N; x++)
        for (int y = 0; y < N; y++) {
            int v = b[x * N + y];
            if (v > 0 && v <= nc) C::set(g.tl, v - 1, C::cell(x, y));
            else if (v < 0 && v >= -nc) C::set(g.br, -v - 1, C::cell(x, y));
        }
    int d = dice ? dice[i] : 1;
    g.dice = d < 1 ? 1 : (d > nc ? nc : d);
    g.player = (player && player[i] == EWN_BOTTOM_RIGHT) ? SIDE_BR : SIDE_TL;
    pack(g, state, i, geo);
}

template <class C> __global__ void __launch_bounds__(TPB) k_decode(const u64* state, int16_t* board, int64_t* dice, int32_t* player,
                                                                   int64_t n, Geo geo) {
    extern __shared__ uint4 smem_raw[];
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool mine = i < n;
    Game<C> g; g.tl = C::all(); g.br = C::all(); g.dice = 1; g.player = 0;
    if (mine) {
        unpack(g, state, i, geo);
        if (dice) dice[i] = g.dice;
        if (player) player[i] = g.player == SIDE_BR ? EWN_BOTTOM_RIGHT : EWN_TOP_LEFT;
    }
    if (board) {
        const int NN = geo.N * geo.N;
        int16_t* sm = reinterpret_cast<int16_t*>(smem_raw) + warp * 32 * NN;
        const int64_t warp_env0 = i - lane;
        if (warp_env0 < n) {
            stage_rows<C, 0>(g, mine, geo, sm, lane);
            write_rows_dense<C, 0>(board, warp_env0, n, geo, sm, lane);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K2 rules
// ------------------------------------------------------------------------------------------------
template <class C> __global__ void __launch_bounds__(TPB) k_legal(const u64* state, int player, uint8_t* mask, int8_t* list,
                                                                  int32_t* count, int64_t n, Geo geo) {
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    Game<C> g; unpack(g, state, i, geo);
    int side = player == 0 ? g.player : (player == EWN_BOTTOM_RIGHT ? SIDE_BR : SIDE_TL);
    Legal l;
    if (side == SIDE_BR) l = legal<C, SIDE_BR>(g, geo, movers<C>(g.br, g.dice));
    else                 l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
    int c = legal_count(l);
    if (mask) mask[i] = (uint8_t)legal_mask(l);
    if (count) count[i] = c;
    if (list) {
        int8_t* o = list + i * 12;
        for (int j = 0; j < 6; j++) {
            int la = -1, di = -1, sh;
            if (j < c) legal_pick(l, j, la, di, sh);
            o[2 * j] = (int8_t)la; o[2 * j + 1] = (int8_t)di;
        }
    }
}

template <class C> __global__ void __launch_bounds__(TPB) k_check_win(const u64* state, uint8_t* win, int64_t n, Geo geo) {
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    Game<C> g; unpack(g, state, i, geo);
    win[i] = full_win<C>(g, geo) ? 1 : 0;
}

template <class C> __global__ void __launch_bounds__(TPB) k_ply(u64* state, int player, const int32_t* action, uint8_t* valid,
                                                                uint8_t* win, int32_t* moved, int64_t n, Geo geo) {
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    Game<C> g; unpack(g, state, i, geo);
    int side = player == 0 ? g.player : (player == EWN_BOTTOM_RIGHT ? SIDE_BR : SIDE_TL);
    int t = 0, sh = -1;
    bool ok;
    if (side == SIDE_BR) ok = ply_action<C, SIDE_BR>(g, geo, action[2 * i], action[2 * i + 1], t, sh);
    else                 ok = ply_action<C, SIDE_TL>(g, geo, action[2 * i], action[2 * i + 1], t, sh);
    if (ok) g.player = side ^ 1;
    pack(g, state, i, geo);
    if (valid) valid[i] = ok ? 1 : 0;
    if (win) win[i] = full_win<C>(g, geo) ? 1 : 0;
    if (moved) moved[i] = sh < 0 ? 0 : sh / C::FB + 1;
}

template <class C> __global__ void __launch_bounds__(TPB) k_set_dice(u64* state, const int32_t* dice, const int32_t* player, int64_t n, Geo geo) {
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    Game<C> g; unpack(g, state, i, geo);
    if (dice) { int d = dice[i]; g.dice = d < 1 ? 1 : (d > geo.cube_num ? geo.cube_num : d); }
    if (player) g.player = player[i] == EWN_BOTTOM_RIGHT ? SIDE_BR : SIDE_TL;
    pack(g, state, i, geo);
}

// NN-facing observation (classical_policies/alpha_zero/alpha_zero.py:39-58 ewn_obs_one_hot): board[x][y][k] = +1 where TOP_LEFT's
// cube k+1 stands, -1 for BOTTOM_RIGHT's, dice one-hot.  One warp per 32 games: the block's tile is zero-filled and written with
// coalesced float4 stores, then every lane drops its <= 12 non-zeros (and the die) into its own game's row.
template <class C> __global__ void __launch_bounds__(TPB) k_decode_onehot(const u64* state, float* board, float* dice, int64_t n, Geo geo) {
    const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    const int cn = geo.cube_num, NN = geo.N * geo.N, row = NN * cn;
    const int64_t blk0 = (int64_t)blockIdx.x * TPB;
    const int64_t rows = (n - blk0 < TPB) ? (n - blk0) : TPB;
    if (rows <= 0) return;
    {   // zero fill of this block's contiguous slab
        float* b0 = board + blk0 * row;
        const int64_t tot = rows * row;
        if ((reinterpret_cast<uintptr_t>(b0) & 15) == 0) {
            float4* v = reinterpret_cast<float4*>(b0);
            for (int64_t k = threadIdx.x; k < (tot >> 2); k += TPB) v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int64_t k = (tot & ~(int64_t)3) + threadIdx.x; k < tot; k += TPB) b0[k] = 0.f;
        } else {
            for (int64_t k = threadIdx.x; k < tot; k += TPB) b0[k] = 0.f;
        }
        if (dice) { float* d0 = dice + blk0 * cn; for (int64_t k = threadIdx.x; k < rows * cn; k += TPB) d0[k] = 0.f; }
    }
    __syncthreads();
    if (i >= n) return;
    Game<C> g; unpack(g, state, i, geo);
    float* b = board + i * row;
#pragma unroll
    for (int k = 0; k < C::MAXC; k++) {
        if (k >= cn) break;
        const int c = C::get(g.tl, k), d = C::get(g.br, k);
        if (c != (int)C::FM) b[(C::x_of(c) * geo.N + C::y_of(c)) * cn + k] = 1.f;
        if (d != (int)C::FM) b[(C::x_of(d) * geo.N + C::y_of(d)) * cn + k] = -1.f;
    }
    if (dice && g.dice >= 1 && g.dice <= cn) dice[i * cn + g.dice - 1] = 1.f;
}

// eval_random.py from seeds alone: one thread per game, the reference's MT19937 stream in local memory (2.5 KB per thread)
template <class C> __global__ void __launch_bounds__(128) k_numpy_random_games(const u32* seeds, int64_t n, Geo geo, int8_t* result,
                                                                               int32_t* agent_steps, u64* final_state, u64* trace) {
    const int64_t i = (int64_t)blockIdx.x * 128 + threadIdx.x;
    if (i >= n) return;
    MT19937 rng;
    Game<C> g;
    int steps; u64 h;
    const int r = numpy_random_game<C>(g, geo, seeds[i], rng, steps, h);
    if (result) result[i] = (int8_t)r;
    if (agent_steps) agent_steps[i] = steps;
    if (final_state) pack(g, final_state, i, geo);
    if (trace) trace[i] = h;
}

__global__ void __launch_bounds__(128) k_numpy_stream(const u32* seeds, int64_t n, int K, u32* raw) {
    const int64_t i = (int64_t)blockIdx.x * 128 + threadIdx.x;
    if (i >= n) return;
    MT19937 rng; rng.seed(seeds[i]);
    for (int k = 0; k < K; k++) raw[i * K + k] = rng.next();
}
// roll_dice (envs/ewn.py:90-92) and / or RandomAgent.predict (random_policy.py:12-14) for every game from its stored stream
template <class C> __global__ void __launch_bounds__(TPB) k_numpy_draw(u64* state, const u32* raw, int32_t* cursor, int K, int64_t n, Geo geo,
                                                                       int roll, int pick_player, int32_t* action) {
    const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    Game<C> g; unpack(g, state, i, geo);
    RawStream rs; rs.raw = raw + i * K; rs.K = K; rs.cur = cursor[i];
    if (roll) { g.dice = rs.randint(1, geo.cube_num + 1); pack(g, state, i, geo); }
    if (pick_player) {
        Legal l;
        if (pick_player == EWN_BOTTOM_RIGHT) l = legal<C, SIDE_BR>(g, geo, movers<C>(g.br, g.dice));
        else                                 l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
        const int cnt = legal_count(l);
        int la = 0, di = 0, sh;
        if (cnt > 0) legal_pick(l, rs.randint(0, cnt), la, di, sh);
        action[2 * i] = la; action[2 * i + 1] = di;
    }
    cursor[i] = rs.cur;
}

template <class C> __global__ void __launch_bounds__(TPB) k_flip(const u64* in, u64* out, int64_t n, Geo geo) {
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    Game<C> g; unpack(g, in, i, geo);
    flip_game<C>(g, geo);
    pack(g, out, i, geo);
}

template <class C> __global__ void __launch_bounds__(TPB) k_random_action(const u64* state, int64_t* action, int64_t n, Geo geo,
                                                                          u64 seed, u64 step, u64 env_id0) {
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    Game<C> g; unpack(g, state, i, geo);
    u32 blk[4]; env_block(seed, env_id0 + (u64)i, step, TAG_STEP, blk);
    Legal l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
    int c = legal_count(l);
    int la = 0, di = 0, sh;
    if (c > 0) legal_pick(l, (int)bounded(blk[3], (u32)c), la, di, sh);
    reinterpret_cast<longlong2*>(action)[i] = make_longlong2(la, di);
}

// ------------------------------------------------------------------------------------------------
// K1 env step
// ------------------------------------------------------------------------------------------------
struct StepKArgs {
    ewn_step_args a;
    StepParams p;
};

// VEC = the VecEnv surface and nothing else (int64 actions in; f32 reward, terminated, truncated, int16 board and int64
// dice out; Philox dice, random opponent): every optional-pointer test below folds away at compile time.
//
// Finished games are handled at BLOCK level: warps append the packed terminal states of their finished lanes to a shared
// list (one shared-memory atomic per warp, which also carries the won / lost tallies), the block reserves F consecutive
// rows of the compacted terminal-observation arrays with ONE global atomic, and then all threads decode the F boards
// together, one (game, cube) pair per thread -- no lane sits idle while a few lanes of its warp write boards.
constexpr int STPB = 128;                // threads per block of k_step: small blocks keep the tail of the last wave short
constexpr int SWARPS = STPB / 32;

template <class C, int NT, bool REPLAY, bool SHAPED, bool VEC> __global__ void __launch_bounds__(STPB)
k_step(const __grid_constant__ StepKArgs ka) {
    extern __shared__ uint4 smem_raw[];
    __shared__ unsigned s_tot, s_row0;                          // s_tot = finished | won << 10 | lost << 20 of this block
    __shared__ __align__(16) u64 s_term[STPB * C::WORDS];       // packed terminal states of the finished games (PG: geo.sw <= 9 words each)
    __shared__ unsigned char s_who[STPB];                       // ... and the thread that owns each of them
    __shared__ unsigned char s_code[STPB];                      // ... and how each of them ended (EWN_CODE_*)
    const ewn_step_args& a = ka.a;
    const int64_t n = a.n;
    const unsigned tid = threadIdx.x;
    const int64_t i = (int64_t)blockIdx.x * STPB + tid;
    const int lane = tid & 31, warp = tid >> 5;
    const bool mine = i < n;
    const bool compact = a.term_count != nullptr;
    // programmatic dependent launch (as k_vstep): the next kernel of the stream / graph may be scheduled while this grid runs, and nothing
    // here touches global memory before the previous grid has completed.  A rollout of small batches is a chain of launch gaps: this
    // shortens every one of them.  No-ops without the launch attribute.
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (tid == 0) {
        s_tot = 0;
        if (a.term_count_clear && blockIdx.x == 0) *a.term_count_clear = 0u;
    }

    Game<C> g, term;
    g.tl = C::all(); g.br = C::all(); g.dice = 1; g.player = 0; term = g;
    StepOut out; out.term = 0; out.code = CODE_ONGOING;
    if (mine) {
        StepParams p = ka.p;
        p.replay = REPLAY; p.shaped = SHAPED;
        if (NT) { p.geo.N = NT; p.geo.corner = C::cell(NT - 1, NT - 1); }
        if (VEC) p.agent_random = 0;
        unpack(g, a.state, i, ka.p.geo);
        int al = 0, ad = 0;
        if (!p.agent_random) {
            if (VEC || a.action_i64) { longlong2 v = __ldcs(reinterpret_cast<const longlong2*>(a.action_i64) + i); al = (int)v.x; ad = (v.y < 0 || v.y > 2) ? 3 : (int)v.y; }
            else if (a.action_i32) { int2 v = reinterpret_cast<const int2*>(a.action_i32)[i]; al = v.x; ad = v.y; }
            else { int v = a.action_u8[i]; al = v >= 3; ad = v - 3 * al; }
        }
        StepInject inj; inj.dice_opp = 1; inj.o_larger = 0; inj.o_dir = 0; inj.dice_next = 1;
        if (REPLAY) {
            const int cn = ka.p.geo.cube_num;                        // a die outside 1..cube_num would spill into the packed state's flag bits
            inj.dice_opp = min(max(a.inject_dice_opp[i], 1), cn); inj.dice_next = min(max(a.inject_dice_next[i], 1), cn);
            int2 o = reinterpret_cast<const int2*>(a.inject_opp_action)[i]; inj.o_larger = o.x; inj.o_dir = o.y;
        }
        int tol = 0; double prev = 0.0;
        if (SHAPED) { tol = a.tolerance[i]; prev = a.prev_score[i]; }
        if (a.step_dev) p.step += __ldg(a.step_dev);
        step_one<C>(g, term, p, i, al, ad, inj, &tol, &prev, out);
        if (SHAPED) { a.tolerance[i] = tol; a.prev_score[i] = prev; }
        pack(g, a.state, i, ka.p.geo);
        if (VEC) {
            __stcs(a.reward_f32 + i, (float)out.reward);
            __stcs(a.terminated + i, (uint8_t)out.term);
            __stcs(a.truncated + i, (uint8_t)out.trunc);
            __stcs(reinterpret_cast<long long*>(a.obs_dice) + i, (long long)g.dice);
        } else {
            if (a.reward_f32) a.reward_f32[i] = (float)out.reward;
            if (a.reward_f64) a.reward_f64[i] = out.reward;
            if (a.terminated) a.terminated[i] = (uint8_t)out.term;
            if (a.truncated) a.truncated[i] = (uint8_t)out.trunc;
            if (a.code) a.code[i] = out.code;
            if (a.action_out) a.action_out[i] = (uint8_t)(out.a_larger * 3 + out.a_dir);
            if (a.obs_dice) a.obs_dice[i] = g.dice;
        }
    }
    Geo geo = ka.p.geo;
    if (NT) geo.N = NT;
    const int NN = geo.N * geo.N;
    int16_t* const smb = reinterpret_cast<int16_t*>(smem_raw);
    const int64_t warp_env0 = i - lane;
    const bool fin = mine && out.term;
    if ((VEC || a.obs_board) && warp_env0 < n) {                 // observation after the step: warp-level staging
        int16_t* sm = smb + warp * 32 * NN;
        stage_rows<C, NT>(g, mine, geo, sm, lane);
        write_rows_dense<C, NT>(a.obs_board, warp_env0, n, geo, sm, lane);
    }
    if (!compact && (a.term_board || a.term_dice || a.term_code)) {   // dense terminal rows (row = env index), warp-level
        const unsigned done = __ballot_sync(0xffffffffu, fin);
        if (done) {
            if (fin && a.term_dice) a.term_dice[i] = term.dice;
            if (fin && a.term_code) a.term_code[i] = (uint8_t)out.code;
            if (a.term_board) {
                int16_t* sm = smb + warp * 32 * NN;
                stage_rows<C, NT>(term, fin, geo, sm, lane);
                write_rows_sparse<C, NT>(a.term_board, warp_env0, false, done, geo, sm, lane);
            }
        }
    }
    if (!(a.stats || compact)) return;                            // uniform over the grid

    __syncthreads();                                              // s_tot is zero; the staging area is free again
    {
        const unsigned done = __ballot_sync(0xffffffffu, fin);
        if (done) {
            const unsigned won = __ballot_sync(0xffffffffu, fin && out.code == CODE_AGENT_WON);
            const unsigned lost = __ballot_sync(0xffffffffu, fin && out.code == CODE_AGENT_LOST);
            unsigned base = 0;
            if (lane == 0) base = atomicAdd(&s_tot, (unsigned)__popc(done) | ((unsigned)__popc(won) << 10) | ((unsigned)__popc(lost) << 20));
            if (compact) {
                base = __shfl_sync(0xffffffffu, base, 0) & 1023u;
                if (fin) {
                    const unsigned k = base + (unsigned)__popc(done & ((1u << lane) - 1u));
                    pack(term, s_term, (long)k, geo);
                    s_who[k] = (unsigned char)tid;
                    s_code[k] = (unsigned char)out.code;
                }
            }
        }
    }
    __syncthreads();
    const unsigned tot = s_tot;
    const int F = (int)(tot & 1023u);
    if (F == 0) return;                                           // uniform over the block
    if (a.stats && tid < 4) {
        const unsigned w = (tot >> 10) & 1023u, l = tot >> 20;
        const unsigned v = tid == 0 ? w : tid == 1 ? l : tid == 2 ? (unsigned)F - w - l : (unsigned)F;
        if (v) atomicAdd(a.stats + tid, (unsigned long long)v);
    }
    if (!compact) return;
    if (tid == 0) s_row0 = atomicAdd(a.term_count, (unsigned)F);
    const bool boards = a.term_board != nullptr;
    if (boards) {                                                 // rows are staged at the 16-byte phase of their destination
        uint4* v = reinterpret_cast<uint4*>(smb);
        const int nvec = (F * NN + 14) >> 3;
        for (int k = tid; k < nvec; k += STPB) v[k] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();
    const int64_t row0 = (int64_t)s_row0;
    EWN_CHECK(F <= STPB && row0 >= 0 && row0 + F <= a.term_index_base + n);      // the caller's terminal arrays hold one row per game
    const int head = (int)(((unsigned)row0 * (unsigned)NN) & 7u);
    if ((int)tid < F) {
        Game<C> t; unpack(t, s_term, (long)tid, geo);
        if (a.term_dice) __stcs(reinterpret_cast<long long*>(a.term_dice) + row0 + tid, (long long)t.dice);
        __stcs(a.term_index + row0 + tid, (int32_t)(a.term_index_base + (int64_t)blockIdx.x * STPB + s_who[tid]));
        if (a.term_code) a.term_code[row0 + tid] = s_code[tid];
        if constexpr (C::GENERIC) {                               // generic codec: one finished game per thread, all its cubes
            if (boards)
                for (int k = 0; k < geo.cube_num; k++) {
                    const int c = C::get(t.tl, k), d = C::get(t.br, k);
                    if (c != (int)C::FM) smb[head + tid * NN + lin_cell<C, NT>(geo, c)] = (int16_t)(k + 1);
                    if (d != (int)C::FM) smb[head + tid * NN + lin_cell<C, NT>(geo, d)] = (int16_t)(-(k + 1));
                }
        }
    }
    if (!boards) return;
    if constexpr (!C::GENERIC) {
        for (int p = tid; p < F * 12; p += STPB) {                // one (finished game, cube) pair per thread
            const int r = (int)__umulhi((unsigned)p, 0x15555556u);    // p / 12
            const int k = p - 12 * r;
            const int br = k >= 6, kk = br ? k - 6 : k;
            const typename C::W side = (typename C::W)(C::WORDS == 1 ? (br ? s_term[r] >> 32 : s_term[r]) : s_term[2 * r + br]);
            const int c = (int)((side >> (C::FB * kk)) & C::FM);
            if (c != (int)C::FM) smb[head + r * NN + lin_cell<C, NT>(geo, c)] = (int16_t)(br ? -(kk + 1) : kk + 1);
        }
    }
    __syncthreads();
    int16_t* const dst = a.term_board + row0 * NN - head;         // element `e` of the staging area goes to dst[e]
    const int lo = head, hi = head + F * NN;
    if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
        const int v0 = (lo + 7) >> 3, v1 = hi >> 3;               // full 16-byte vectors [v0, v1)
        const uint4* sv = reinterpret_cast<const uint4*>(smb);
        uint4* dv = reinterpret_cast<uint4*>(dst);
        for (int k = v0 + tid; k < v1; k += STPB) __stcs(dv + k, sv[k]);
        const int h1 = min(v0 << 3, hi), t0 = max(v1 << 3, h1);
        if ((int)tid < 8) { const int e = lo + tid; if (e < h1) dst[e] = smb[e]; }
        else if (tid < 16) { const int e = t0 + (tid - 8); if (e < hi) dst[e] = smb[e]; }
    } else {
        for (int e = lo + tid; e < hi; e += STPB) dst[e] = smb[e];
    }
}

// ------------------------------------------------------------------------------------------------
// K1v  persistent VecEnv step.  Same arithmetic as k_step<.., VEC = true>, different schedule:
//   * the grid is sized to the machine (sm_count x VMINB resident blocks); a block walks units of VTPB games with a
//     grid stride and loads the NEXT unit's state and action before it computes the current one, so DRAM latency and
//     the block-launch ramp are paid once per block instead of once per game;
//   * observation rows are staged per warp (__syncwarp only) -- there is no block barrier inside the loop;
//   * finished games are only APPENDED (packed state + env index) to a block-local shared list inside the loop, one
//     shared atomic per warp; after the loop the block reserves its rows of the compacted terminal-observation arrays
//     with ONE global atomic and all threads decode the boards together, one (game, cube) pair per thread.
// ------------------------------------------------------------------------------------------------
// 128-thread blocks x 10 per SM = 40 resident warps at 48 registers: measured best on B200 (profiles/r2_k1_sweeps.txt: 26.0 us per
// 1 Mi games against 26.6 for 32 x 32 = 32 warps at 60 registers, 26.1 for 64 x 20, 26.2 for 256 x 5; 48 warps need 40 registers,
// which spills inside the loop: 27.6).  An SM holds at most 32 blocks, so one-warp blocks cannot go beyond 32 warps.
#ifndef EWN_VTPB
#define EWN_VTPB 128
#endif
#ifndef EWN_VMINB
#define EWN_VMINB 10
#endif
#ifndef EWN_VITER_MAX
#define EWN_VITER_MAX 8
#endif
constexpr int VTPB = EWN_VTPB, VMINB = EWN_VMINB, VITER_MAX = EWN_VITER_MAX;

// fills the per-block table dirs[side][cell] and returns the geometry that reads it (call by all threads, then __syncthreads)
template <class C, int NTHREADS> __device__ __forceinline__ GeoLut make_geo_lut(const Geo& geo, unsigned char* tab) {
    for (int c = threadIdx.x; c < C::LUT; c += NTHREADS) {
        tab[c] = (unsigned char)C::template dirs<SIDE_TL>(geo, c);
        tab[C::LUT + c] = (unsigned char)C::template dirs<SIDE_BR>(geo, c);
        const int x = C::x_of(c), y = C::y_of(c);
        tab[2 * C::LUT + c] = (unsigned char)(signed char)(c == (int)C::FM ? -1 : (x < y ? x : y));   // min(x, y), dead cube: -1
    }
    GeoLut gl;
    static_cast<Geo&>(gl) = geo;
    gl.lut = (u32)__cvta_generic_to_shared(tab);
    gl.vt = nullptr;
    return gl;
}

// lane-private scatter of the twelve cubes into the lane's row of a zeroed [32][NN] int16 tile
template <class C, int NT> __device__ __forceinline__ void scatter_row(const Game<C>& g, const Geo& geo, int16_t* sm, int lane) {
    const int NN = NT ? NT * NT : geo.N * geo.N;
    char* row = reinterpret_cast<char*>(sm) + lane * (2 * NN);
#pragma unroll
    for (int k = 0; k < 6; k++) {
        if (C::S == (NT ? NT : -1)) {                 // packed cell == row-major index (5x5): byte offset straight from the field
            const u32 a = (u32)(k == 0 ? (g.tl << 1) : (g.tl >> (C::FB * k - 1))) & (2u * (u32)C::FM);
            const u32 b = (u32)(k == 0 ? (g.br << 1) : (g.br >> (C::FB * k - 1))) & (2u * (u32)C::FM);
            if (a != 2u * (u32)C::FM) *reinterpret_cast<int16_t*>(row + a) = (int16_t)(k + 1);
            if (b != 2u * (u32)C::FM) *reinterpret_cast<int16_t*>(row + b) = (int16_t)(-(k + 1));
        } else {
            const int c = (int)((g.tl >> (C::FB * k)) & C::FM), d = (int)((g.br >> (C::FB * k)) & C::FM);
            if (c != (int)C::FM) *reinterpret_cast<int16_t*>(row + 2 * lin_cell<C, NT>(geo, c)) = (int16_t)(k + 1);
            if (d != (int)C::FM) *reinterpret_cast<int16_t*>(row + 2 * lin_cell<C, NT>(geo, d)) = (int16_t)(-(k + 1));
        }
    }
}

#ifndef EWN_VFLUSH
#define EWN_VFLUSH 1
#endif
#ifndef EWN_VPRE
#define EWN_VPRE 1
#endif
// EWN_VSTAG = 1: every block (one warp: EWN_VTPB == 32) also flushes the terminal rows it has collected so far after unit
// EWN_VSTAG_LO + blockIdx % (iters - 1 - EWN_VSTAG_LO) of its walk, so that most rows leave while other warps of the SM are
// computing and the flush at the end of the grid -- where every warp of the machine is in its latency chain at the same moment --
// has only the rows of the last units left.
#ifndef EWN_VSTAG
#define EWN_VSTAG 0
#endif
#ifndef EWN_VSTAG_LO
#define EWN_VSTAG_LO 2
#endif
// Terminal rows of one block, after its loop: thread r owns finished game r of the pass (its dice, its env index and the
// scatter of its twelve cubes into the staging tile), then the tile leaves as 16-byte vectors.  Kept out of line so that its
// registers and schedule do not disturb the step loop.
template <class C, int NT> __device__ __noinline__ void vstep_flush(int16_t* term_board, int64_t* term_dice, int32_t* term_index, uint8_t* term_code, int64_t index_base,
                                                                    Geo geo, int16_t* smb, const u64* s_term, const u32* s_who,
                                                                    const unsigned* s_wcnt, unsigned seg, int F, int64_t row0) {
    const int tid = (int)threadIdx.x;
    const int NN = NT ? NT * NT : geo.N * geo.N;
    unsigned pre[VTPB / 32 + 1];
    pre[0] = 0;
#pragma unroll
    for (int w = 0; w < VTPB / 32; w++) pre[w + 1] = pre[w] + s_wcnt[w];
    auto src = [&](int r) -> int {                                 // row r of this block's output <-> entry of the segmented list
        int w = 0;
#pragma unroll
        for (int k = 1; k < VTPB / 32; k++) w += (unsigned)r >= pre[k];
        unsigned base = 0;
#pragma unroll
        for (int k = 1; k < VTPB / 32; k++) base = w == k ? pre[k] : base;
        return w * (int)seg + (r - (int)base);
    };
    for (int r0 = 0; r0 < F; r0 += VTPB) {                         // VTPB rows per pass through the staging tile
        const int Fc = min(VTPB, F - r0);
        const int64_t rowc = row0 + r0;
        const int head = (int)(((unsigned)rowc * (unsigned)NN) & 7u);   // 16-byte phase of the destination
        uint4* v = reinterpret_cast<uint4*>(smb);
        const int nvec = (head + Fc * NN + 7) >> 3;
        EWN_CHECK(nvec * 8 <= VTPB * NN + 16 && rowc >= 0);            // the staging tile: VTPB rows + 32 bytes of slack
        if (term_board) for (int k = tid; k < nvec; k += VTPB) v[k] = make_uint4(0, 0, 0, 0);
        __syncthreads();
        if (tid < Fc) {
            const int e = src(r0 + tid);
            EWN_CHECK(e >= 0 && (unsigned)e < seg * (VTPB / 32));
            Game<C> t; unpack(t, s_term, (long)e);
            if (term_dice) __stcs(reinterpret_cast<long long*>(term_dice) + rowc + tid, (long long)t.dice);
            __stcs(term_index + rowc + tid, (int32_t)(index_base + (int64_t)(s_who[e] & 0x3FFFFFFFu)));
            if (term_code) term_code[rowc + tid] = (uint8_t)(1u + (s_who[e] >> 30));        // EWN_CODE_* 1..4
            if (term_board) scatter_row<C, NT>(t, geo, smb + head, tid);
        }
        if (!term_board) continue;
        __syncthreads();
        int16_t* const dst = term_board + rowc * NN - head;        // element e of the staging tile goes to dst[e]
        const int lo = head, hi = head + Fc * NN;
        if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
            const int v0 = (lo + 7) >> 3, v1 = hi >> 3;            // full 16-byte vectors [v0, v1)
            const uint4* sv = reinterpret_cast<const uint4*>(smb);
            uint4* dv = reinterpret_cast<uint4*>(dst);
            for (int k = v0 + tid; k < v1; k += VTPB) __stcs(dv + k, sv[k]);
            const int h1 = min(v0 << 3, hi), t0 = max(v1 << 3, h1);
            if (tid < 8) { const int e = lo + tid; if (e < h1) dst[e] = smb[e]; }
            else if (tid < 16) { const int e = t0 + (tid - 8); if (e < hi) dst[e] = smb[e]; }
        } else {
            for (int e = lo + tid; e < hi; e += VTPB) dst[e] = smb[e];
        }
        if (r0 + VTPB < F) __syncthreads();
    }
}

// asynchronous copy of one game's packed state and int64 action pair into the thread's 32-byte shared slot
template <class C, bool AR> __device__ __forceinline__ void prefetch_game(unsigned slot, const u64* state, const int64_t* action, int i) {
    if (C::WORDS == 1) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(slot), "l"(state + i) : "memory");
    else               asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(slot), "l"(state + 2 * (size_t)i) : "memory");
    if (!AR) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(slot + 16u), "l"(action + 2 * (size_t)i) : "memory");
}

#if defined(EWN_EXP_TIMELINE)    /* experiment build: per-warp clock stamps of the last k_vstep launch (tools/k1_sweep.py EWN_K1_TIMELINE=..., tools/k1_timeline.py) */
__device__ unsigned long long g_tl[148 * 64 * 16];
#define EWN_TL_SLOT ((((size_t)blockIdx.x * (VTPB / 32) + (threadIdx.x >> 5)) % (148 * 64)) * 16)
#define EWN_TL(k) do { if ((threadIdx.x & 31) == 0) g_tl[EWN_TL_SLOT + (k)] = (unsigned long long)clock64(); } while (0)
#else
#define EWN_TL(k) do { } while (0)
#endif
// AR: the agent is the in-kernel uniform-over-the-legal-list player (random_policy.py:11-15, Philox word 3): no action array is read
template <class C, int NT, bool AR> __global__ void __launch_bounds__(VTPB, VMINB)
k_vstep(const __grid_constant__ StepKArgs ka, const int cap) {
    extern __shared__ uint4 smem_raw[];
    __shared__ unsigned s_won, s_lost, s_row0, s_wcnt[VTPB / 32];
    __shared__ ulonglong2 s_pf[2 * VTPB];
    __shared__ __align__(16) unsigned char s_dirs[3 * C::LUT];     // legal-direction set per (side, cell): one LDS per cube
    const ewn_step_args& a = ka.a;
    const int n = (int)a.n;                                        // the host keeps n below 2^31 on this path: 32-bit indices
    const unsigned tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const bool compact = a.term_count != nullptr;
    const bool tally = compact || a.stats != nullptr;
    Geo geo = ka.p.geo;
    if (NT) { geo.N = NT; geo.corner = C::cell(NT - 1, NT - 1); }
    const int NN = geo.N * geo.N;
    int16_t* const smb = reinterpret_cast<int16_t*>(smem_raw);                       // [VTPB][NN] staging (+ 32 B slack)
    u64* const s_term = reinterpret_cast<u64*>(reinterpret_cast<char*>(smem_raw) + (((size_t)VTPB * NN * 2 + 32 + 15) & ~(size_t)15));
    u32* const s_who = reinterpret_cast<u32*>(s_term + (size_t)cap * C::WORDS);
    // Programmatic dependent launch: let the NEXT kernel of the stream become resident while this grid drains (its blocks
    // take the SM slots our blocks free and park at their own griddepcontrol.wait), and do not touch global memory
    // before the PREVIOUS grid has completed and flushed.  Both are no-ops for launches without the PDL attribute.
    asm volatile("griddepcontrol.launch_dependents;");
    if (tid == 0) { s_won = 0; s_lost = 0; }
#if defined(EWN_EXP_TIMELINE)
    if ((tid & 31) == 0) {
        unsigned long long gt; unsigned smid;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        g_tl[EWN_TL_SLOT + 15] = gt; g_tl[EWN_TL_SLOT + 14] = smid;
        for (int k = 1; k < 14; k++) g_tl[EWN_TL_SLOT + k] = 0;    // stamps of earlier launches
    }
    EWN_TL(0);
    int it_ = 0;
#endif
#if EWN_VPRE
    // Pull the first unit's lines into L2 while the previous grid is still draining.  A prefetch has no architectural effect
    // and L2 is the coherence point, so this is safe although the previous grid may still be writing these very addresses;
    // the loads proper stay behind griddepcontrol.wait.
    {
        const int i0 = (int)blockIdx.x * VTPB + (int)tid;
        if (i0 < n) {                                  // the first unit only: all seven units measured slower (27.3 vs 26.5 us)
            if ((tid & 15u) == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.state + (size_t)i0 * C::WORDS));
            if (!AR && (tid & 7u) == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.action_i64 + 2 * (size_t)i0));
        }
    }
#endif
    const GeoLut geol = make_geo_lut<C, VTPB>(geo, s_dirs);
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (tid == 0 && a.term_count_clear && blockIdx.x == 0) *a.term_count_clear = 0u;
    EWN_TL(1);
    __syncthreads();

    const u64 step = ka.p.step + (a.step_dev ? __ldg(a.step_dev) : 0ull);   // device-resident step counter: CUDA-graph replays
    const int n_units = (n + VTPB - 1) / VTPB;
    int u = blockIdx.x;
    int i = u * VTPB + tid;
    bool mine = i < n;
    // prefetch slot of this thread: state (8 / 16 B) + action (16 B), filled by cp.async while the previous unit is computed
    const unsigned pf = (unsigned)__cvta_generic_to_shared(s_pf) + tid * 32u;
    if (mine) prefetch_game<C, AR>(pf, a.state, a.action_i64, i);
    const bool obs_vec = (reinterpret_cast<uintptr_t>(a.obs_board) & 15) == 0;
    unsigned tally_c = 0;                                          // wins | losses << 16 of this thread
    unsigned wcnt = 0;                                             // finished games of this warp so far (warp-uniform)
    const unsigned seg = (unsigned)cap / (VTPB / 32);              // list entries per warp: one per game the warp can meet
    int16_t* const sm = smb + warp * 32 * NN;
#if EWN_VSTAG
    static_assert(VTPB == 32, "the staggered flush is written for one-warp blocks");
    const int iters_b = (n_units + (int)gridDim.x - 1) / (int)gridDim.x;
    const int stag_at = iters_b - 1 > EWN_VSTAG_LO ? EWN_VSTAG_LO + (int)(blockIdx.x % (unsigned)(iters_b - 1 - EWN_VSTAG_LO)) : -1;
    unsigned flushed = 0;
    // two passes over the same loop body: units up to the block's flush point, the flush (a call: kept OUT of the unit loop so that
    // nothing of the loop's register allocation has to survive it), then the rest
    const int u_split = stag_at < 0 ? n_units : min(n_units, (int)blockIdx.x + (stag_at + 1) * (int)gridDim.x);
    for (int phase = 0; phase < 2; phase++) {
    const int u_end = phase == 0 ? u_split : n_units;
#else
    const int u_end = n_units;
#endif
    while (u < u_end) {                                            // block-uniform
#if defined(EWN_EXP_VSYNC)       /* experiment build: the block's warps walk in lock step (the warp scheduler is not fair: profiles/r2_k1_timeline.txt) */
        __syncthreads();
#endif
        asm volatile("cp.async.wait_all;" ::: "memory");
#if defined(EWN_EXP_TIMELINE)
        if (it_ < 8) EWN_TL(2 + it_);
        it_++;
#endif
        const ulonglong2 sw = s_pf[2 * tid];                        // own slot: no cross-thread hazard
        longlong2 act = make_longlong2(0, 0);
        if (!AR) act = reinterpret_cast<const longlong2*>(s_pf)[2 * tid + 1];
        const int un = u + (int)gridDim.x, in = un * VTPB + (int)tid;
        const bool mine_n = in < n;
        if (mine_n) prefetch_game<C, AR>(pf, a.state, a.action_i64, in);
        Game<C> g, term;
        g.tl = C::ALL; g.br = C::ALL; g.dice = 1; g.player = 0; term = g;
        StepOut out; out.term = 0; out.code = CODE_ONGOING;
        if (mine) {
            unpack_regs(g, sw.x, sw.y);
            const int al = (int)act.x, ad = (act.y < 0 || act.y > 2) ? 3 : (int)act.y;
            StepInject inj; inj.dice_opp = 1; inj.o_larger = 0; inj.o_dir = 0; inj.dice_next = 1;
            int tol = 0; double prev = 0.0;
#if defined(EWN_EXP_NOSTEP)      /* experiment build: the I/O skeleton without the rules (tools/step_sweep.py) */
            out.reward = 0.0; out.trunc = 0; g.dice = 1 + ((g.dice + al + ad) % 6);
#elif defined(EWN_EXP_PHILOX)    /* experiment build: I/O skeleton + the Philox block only */
            { u32 blk[4]; env_block_rk(ka.p.rk, ka.p.env_id0 + (u64)i, step, TAG_STEP, blk);
              out.reward = 0.0; out.trunc = 0; g.dice = 1 + (int)bounded(blk[0] ^ blk[1] ^ blk[2] ^ blk[3] ^ (u32)(al + ad), 6u); }
#else
            step_one_cfg<C>(g, term, ka.p, geol, 0, 0, AR ? 1 : 0, step, i, al, ad, inj, &tol, &prev, out);
#endif
            pack(g, a.state, i);
            __stcs(a.reward_f32 + i, (float)out.reward);
            __stcs(a.terminated + i, (uint8_t)out.term);
            __stcs(a.truncated + i, (uint8_t)out.trunc);
            __stcs(reinterpret_cast<long long*>(a.obs_dice) + i, (long long)g.dice);
        }
        if (tally) {                                               // finished games: append to this WARP's segment of the list
            const bool fin = mine && out.term;
            const unsigned done = __ballot_sync(0xffffffffu, fin);
            if (done) {                                            // warp-uniform; no atomics, the counter lives in a register
                if (compact && fin) {
                    const unsigned k = (unsigned)warp * seg + wcnt + (unsigned)__popc(done & ((1u << lane) - 1u));
                    EWN_CHECK(k < ((unsigned)warp + 1u) * seg && k < (unsigned)cap);       // the warp's own segment of the list
                    pack(term, s_term, (long)k);
                    s_who[k] = (u32)i | ((u32)(out.code - 1) << 30);        // i < 2^30; the result code (1..4) rides in the top bits
                }
                wcnt += (unsigned)__popc(done);
                if (fin) tally_c += (out.code == CODE_AGENT_WON ? 1u : 0u) + (out.code == CODE_AGENT_LOST ? 0x10000u : 0u);
            }
        }
        const int warp_env0 = i - lane;
        if (warp_env0 < n) {                                       // warp-uniform: observation rows of this warp's 32 games
            uint4* v = reinterpret_cast<uint4*>(sm);
            const int nvec = 4 * NN;
            if (NT) {
#pragma unroll
                for (int k = 0; k < (4 * NT * NT + 31) / 32; k++) if (lane + 32 * k < nvec) v[lane + 32 * k] = make_uint4(0, 0, 0, 0);
            } else {
                for (int k = lane; k < nvec; k += 32) v[k] = make_uint4(0, 0, 0, 0);
            }
            __syncwarp();
            if (mine) scatter_row<C, NT>(g, geo, sm, lane);
            __syncwarp();
            if (obs_vec && warp_env0 + 32 <= n) {                  // 32 full rows, 16-byte aligned: the warp's tile as vectors
                uint4* d = reinterpret_cast<uint4*>(a.obs_board) + (size_t)(warp_env0 >> 5) * nvec;
                if (NT) {
#pragma unroll
                    for (int k = 0; k < (4 * NT * NT + 31) / 32; k++) if (lane + 32 * k < nvec) __stcs(d + lane + 32 * k, v[lane + 32 * k]);
                } else {
                    for (int k = lane; k < nvec; k += 32) __stcs(d + k, v[k]);
                }
                __syncwarp();
            } else {
                write_rows_dense<C, NT>(a.obs_board, warp_env0, n, geo, sm, lane);
            }
        }
        u = un; i = in; mine = mine_n;
    }
#if EWN_VSTAG
        if (phase == 0 && compact && wcnt) {                       // block-uniform (one warp): rows of the units walked so far
            if (tid == 0) { s_wcnt[0] = wcnt; s_row0 = atomicAdd(a.term_count, wcnt); }
            __syncthreads();
            vstep_flush<C, NT>(a.term_board, a.term_dice, a.term_index, a.term_code, a.term_index_base, geo, smb, s_term, s_who, s_wcnt, seg, (int)wcnt,
                               (int64_t)s_row0);
            __syncthreads();
            flushed += wcnt; wcnt = 0;
        }
    }
#endif
    EWN_TL(10);
#if defined(EWN_EXP_TIMELINE)
    if (lane == 0) g_tl[EWN_TL_SLOT + 13] = (unsigned long long)it_;
#endif
    if (!tally) return;
    {
        const unsigned wl = __reduce_add_sync(0xffffffffu, tally_c);          // <= 32 * VITER_MAX per half: no carry
        if (lane == 0) { s_wcnt[warp] = wcnt; if (wl) { atomicAdd(&s_won, wl & 0xffffu); atomicAdd(&s_lost, wl >> 16); } }
    }
    __syncthreads();
    unsigned pre[VTPB / 32 + 1];                                   // exclusive prefix of the warps' counts
    pre[0] = 0;
#pragma unroll
    for (int w = 0; w < VTPB / 32; w++) pre[w + 1] = pre[w] + s_wcnt[w];
    const int F = (int)pre[VTPB / 32];                             // rows still to flush
#if EWN_VSTAG
    const int Ftot = F + (int)flushed;
#else
    const int Ftot = F;
#endif
    if (Ftot == 0) return;                                         // uniform over the block
    if (a.stats && tid < 4) {
        const unsigned w = s_won, l = s_lost;
        const unsigned v = tid == 0 ? w : tid == 1 ? l : tid == 2 ? (unsigned)Ftot - w - l : (unsigned)Ftot;
        if (v) atomicAdd(a.stats + tid, (unsigned long long)v);
    }
    if (!compact || F == 0) return;
#if defined(EWN_EXP_NOFLUSH)         /* diagnostic build: in-loop append kept, no reservation, no rows written (tools/k1_sweep.py) */
    return;
#elif defined(EWN_EXP_NOATOMIC)      /* diagnostic build: rows at a fixed per-block offset instead of the atomic reservation */
    if (tid == 0) s_row0 = blockIdx.x * 96u;                     /* timing only: rows are not compact (and F <= 96 is assumed) */
#elif defined(EWN_EXP_ATOMIC_ONLY)   /* diagnostic build: the reservation alone, no rows written */
    if (tid == 0) s_row0 = atomicAdd(a.term_count, (unsigned)F);
    __syncthreads();
    if (s_row0 == 0xFFFFFFFFu) a.term_index[0] = 0;                /* keep the result live */
    return;
#else
    if (tid == 0) s_row0 = atomicAdd(a.term_count, (unsigned)F);
#endif
#if EWN_VFLUSH
    __syncthreads();
    EWN_TL(11);
    vstep_flush<C, NT>(a.term_board, a.term_dice, a.term_index, a.term_code, a.term_index_base, geo, smb, s_term, s_who, s_wcnt, seg, F, (int64_t)s_row0);
    EWN_TL(12);
}
#else
    // row r of this block's output <-> entry src(r) of the segmented list
    auto src = [&](int r) -> int {
        int w = 0;
#pragma unroll
        for (int k = 1; k < VTPB / 32; k++) w += (unsigned)r >= pre[k];
        unsigned base = 0;
#pragma unroll
        for (int k = 1; k < VTPB / 32; k++) base = w == k ? pre[k] : base;
        return w * (int)seg + (r - (int)base);
    };
    __syncthreads();
    const int64_t row0 = (int64_t)s_row0;
    for (int r = tid; r < F; r += VTPB) {
        const int e = src(r);
        Game<C> t; unpack(t, s_term, (long)e);
        if (a.term_dice) __stcs(reinterpret_cast<long long*>(a.term_dice) + row0 + r, (long long)t.dice);
        __stcs(a.term_index + row0 + r, (int32_t)(a.term_index_base + (int64_t)(s_who[e] & 0x3FFFFFFFu)));
        if (a.term_code) a.term_code[row0 + r] = (uint8_t)(1u + (s_who[e] >> 30));
    }
    if (!a.term_board) return;
    for (int r0 = 0; r0 < F; r0 += VTPB) {                         // VTPB rows per pass through the staging tile
        const int Fc = min(VTPB, F - r0);
        const int64_t rowc = row0 + r0;
        const int head = (int)(((unsigned)rowc * (unsigned)NN) & 7u);   // 16-byte phase of the destination
        uint4* v = reinterpret_cast<uint4*>(smb);
        const int nvec = (head + Fc * NN + 7) >> 3;
        for (int k = tid; k < nvec; k += VTPB) v[k] = make_uint4(0, 0, 0, 0);
        __syncthreads();
        for (int q = tid; q < Fc * 12; q += VTPB) {                // one (finished game, cube) pair per thread
            const int r = (int)__umulhi((unsigned)q, 0x15555556u);  // q / 12
            const int k = q - 12 * r;
            const int br = k >= 6, kk = br ? k - 6 : k;
            const int e = src(r0 + r);
            const u64 wsel = C::WORDS == 1 ? (br ? s_term[e] >> 32 : s_term[e]) : s_term[2 * e + br];
            const int c = (int)(((typename C::W)wsel >> (C::FB * kk)) & C::FM);
            if (c != (int)C::FM) smb[head + r * NN + lin_cell<C, NT>(geo, c)] = (int16_t)(br ? -(kk + 1) : kk + 1);
        }
        __syncthreads();
        int16_t* const dst = a.term_board + rowc * NN - head;      // element e of the staging tile goes to dst[e]
        const int lo = head, hi = head + Fc * NN;
        if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
            const int v0 = (lo + 7) >> 3, v1 = hi >> 3;            // full 16-byte vectors [v0, v1)
            const uint4* sv = reinterpret_cast<const uint4*>(smb);
            uint4* dv = reinterpret_cast<uint4*>(dst);
            for (int k = v0 + tid; k < v1; k += VTPB) __stcs(dv + k, sv[k]);
            const int h1 = min(v0 << 3, hi), t0 = max(v1 << 3, h1);
            if ((int)tid < 8) { const int e = lo + tid; if (e < h1) dst[e] = smb[e]; }
            else if (tid < 16) { const int e = t0 + (tid - 8); if (e < hi) dst[e] = smb[e]; }
        } else {
            for (int e = lo + tid; e < hi; e += VTPB) dst[e] = smb[e];
        }
        if (r0 + VTPB < F) __syncthreads();
    }
}
#endif

// ------------------------------------------------------------------------------------------------
// evaluate
// ------------------------------------------------------------------------------------------------
template <class C> __global__ void __launch_bounds__(TPB) k_evaluate(const u64* state, double* value, int64_t n, Geo geo, int heuristic) {
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    Game<C> g; unpack(g, state, i, geo);
    value[i] = evaluate<C>(g, geo, heuristic);
}

// ------------------------------------------------------------------------------------------------
// K4 expectiminimax.  Phase 1 (k_emm_expand): one thread per parent of the deepest move level walks its path from the
// root (<= 3 plies), writes the value of every position it owns, then expands its up-to-six children from ONE legal list
// (a thread per child would redo the whole path and the list six times, and 40 % of them would find no such child);
// no inter-thread communication.  Phase 2 (k_emm_replay): one thread per root replays minimax.py:19-73 over the
// stored values, so the reference's window hand-through at chance nodes is reproduced bit for bit.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB) k_emm_table(double* vt, int N) {
    const long i = (long)blockIdx.x * TPB + threadIdx.x;
    if (i < emm_table_size(N)) vt[i] = emm_table_entry(N, i);
}

// sim_states != nullptr ('sim_winrate' search, R = double): leaves are stored as POSITIONS (a u64[WORDS] slot per record) and marked
// pending; their values are filled in by k_sim_playouts + k_sim_values before the replay
template <class C, class R> __global__ void __launch_bounds__(TPB) k_emm_expand(const u64* state, R* vals, int64_t n_roots, Geo geo,
                                                                                int depth, int heuristic, const double* vt, u64* sim_states) {
    __shared__ __align__(16) unsigned char s_dirs[3 * C::LUT];
    GeoLut geol = make_geo_lut<C, TPB>(geo, s_dirs);
    geol.vt = vt;
    __syncthreads();
    const int m = emm_levels(depth);
    const u32 parents = (u32)emm_level_size(m) / 6;
    const u32 gid = blockIdx.x * (u32)TPB + threadIdx.x;          // host keeps n_roots * slots below 2^31
    if (gid >= (u32)n_roots * parents) return;
    const u32 root = gid / parents, tp = gid - root * parents;
    Game<C> g; unpack(g, state, (long)root);
    g.player = SIDE_TL;
    if (full_win<C>(g, geo)) return;               // finished root: phase 2 evaluates it directly
    const int64_t per_root = emm_level_offset(m + 1);
    if constexpr (sizeof(R) == sizeof(double)) {
        if (sim_states) {
            GeoSimT<GeoLut> gs;
            static_cast<GeoLut&>(gs) = geol;
            gs.sim_base = reinterpret_cast<const double*>(vals) + (int64_t)root * per_root;
            gs.sim_states = sim_states + (int64_t)root * per_root * C::WORDS;
            emm_expand_parent<C>(g, gs, depth, heuristic, tp, reinterpret_cast<double*>(vals) + (int64_t)root * per_root);
            return;
        }
    }
    emm_expand_parent<C>(g, geol, depth, heuristic, tp, vals + (int64_t)root * per_root);
}

// ------------------------------------------------------------------------------------------------
// 'sim_winrate' leaves (envs/minimax_ewn.py:215-238).  recs: one double per position, EMM_PENDING (| 4: evaluated six times, minimax.py:
// 64-73 over six depth-0 calls) where playouts are wanted; states: the positions (u64[WORDS] each); bits: per (position, fold, opener)
// the TOP_LEFT-win bit of every playout index.  One thread per (position, fold, opener, playout): a slow, simple kernel for a leaf
// evaluation that costs 200 to 1,200 random games per position by definition.
// ------------------------------------------------------------------------------------------------
template <class C> __global__ void __launch_bounds__(TPB) k_sim_playouts(const double* recs, const u64* states, int64_t n_slots, u32 unit0, int nsim,
                                                                         Geo geo, const RoundKeys rk, u32* bits) {
    __shared__ __align__(16) unsigned char s_dirs[3 * C::LUT];
    const GeoLut geol = make_geo_lut<C, TPB>(geo, s_dirs);
    __syncthreads();
    const int64_t gid = (int64_t)blockIdx.x * TPB + threadIdx.x;
    const int64_t per_slot = 12LL * nsim;                           // 6 folds x 2 openers x playouts
    if (gid >= n_slots * per_slot) return;
    const int64_t slot = gid / per_slot;
    const int r = (int)(gid - slot * per_slot);
    const int fo = r / nsim, j = r - fo * nsim, f = fo >> 1, opener_br = 1 - (fo & 1);
    const u64 rec = f64_to_bits(recs[slot]);
    if ((rec & ~4ull) != EMM_PENDING || f >= ((rec & 4ull) ? 6 : 1)) return;
    Game<C> g; unpack(g, states, (long)slot, geo);
    if (sim_playout<C>(g, geol, rk, (unit0 + (u32)slot) * 6u + (u32)f, opener_br, (u32)j))
        atomicOr(bits + ((slot * 6 + f) * 2 + (opener_br ? 0 : 1)) * sim_words(nsim) + (j >> 5), 1u << (j & 31));
}

__global__ void __launch_bounds__(TPB) k_sim_values(double* recs, int64_t n_slots, int nsim, const u32* bits) {
    const int64_t slot = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (slot >= n_slots) return;
    const u64 rec = f64_to_bits(recs[slot]);
    if ((rec & ~4ull) != EMM_PENDING) return;
    const int W = sim_words(nsim), folds = (rec & 4ull) ? 6 : 1;
    double e = 0.0, v = 0.0;
    for (int f = 0; f < folds; f++) {
        const u32* b = bits + ((slot * 6 + f) * 2) * W;
        v = f64_div((double)sim_chain(b, b + W, nsim), (double)nsim);       // win_count / self.num_simulations (:238)
        e = f64_add(e, f64_div(v, 6.0));                                     // minimax.py:64-73 over six depth-0 evaluations
    }
    recs[slot] = folds == 1 ? v : e;
}

// evaluate('sim_winrate') for a batch of positions: marks finished ones 1.0 / 0.0, the others pending
template <class C> __global__ void __launch_bounds__(TPB) k_sim_mark(const u64* state, double* recs, int64_t n, Geo geo) {
    const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    Game<C> g; unpack(g, state, i, geo);
    recs[i] = full_win<C>(g, geo) ? (tl_won<C>(g, geo) ? 1.0 : 0.0) : bits_to_f64(EMM_PENDING);
}

template <class C, int M, class R> __global__ void __launch_bounds__(128) k_emm_replay(const u64* state, const R* vals, const double* vt, int64_t n_roots,
                                                                                        int64_t root0, Geo geo, int heuristic, int pruned,
                                                                                        double* value, int8_t* action) {
    int64_t r = (int64_t)blockIdx.x * 128 + threadIdx.x;
    if (r >= n_roots) return;
    Game<C> g; unpack(g, state, root0 + r);
    g.player = SIDE_TL;
    int la, di;
    value[root0 + r] = emm_root<C, M>(g, geo, heuristic, pruned, vals + r * emm_level_offset(M + 1), la, di, vt);
    if (action) { action[2 * (root0 + r)] = (int8_t)la; action[2 * (root0 + r) + 1] = (int8_t)di; }
}

// K4f  depth 3 / 4 in ONE kernel, nothing but the roots and the results touches HBM (18 B per root).  Six lanes per root, lane d
// owns die face d + 1; five roots per warp (lanes 30, 31 idle).  Lane d first plays root child d (one TOP_LEFT ply + the two
// histogram words); then the a1 loop runs in lock step: the child of the round reaches all six lanes by shuffle, every lane
// evaluates ITS MIN node against the running alpha (emm2_min: six move slots, leaf values from the shared 'hybrid' table),
// and the expectation is the six v / 6 added in face order (minimax.py:64-73) from shuffles -- every lane of the group ends the
// round with the same best / alpha, so the cut rule of minimax.py:53-58 sees exactly the reference's window.
constexpr int FTPB = 256, FROOTS = (FTPB / 32) * 5;       // threads and roots per block

// `tabs` (global, built once per board size by k_emm_fused_table): [fold][2][pad + N*6*N*6] doubles = values, then quotients
template <class C, int H> __global__ void __launch_bounds__(FTPB) k_emm_fused(const u64* state, int64_t n_roots, Geo geo, int depth, int pruned,
                                                                               const double* tabs, double* value, int8_t* action) {
    extern __shared__ double s_tab[];                                 // 'hybrid' only: values [pad + entries], quotients [pad + entries]
    __shared__ u32 s_inc[C::LUT];
    __shared__ __align__(16) unsigned char s_dirs[3 * C::LUT];
    const GeoLut geol = make_geo_lut<C, FTPB>(geo, s_dirs);
    for (int c = threadIdx.x; c < C::LUT; c += FTPB) s_inc[c] = emm_inc_entry<C>(geo.N, c);
    const bool fold = !(depth & 1);
    const int span = emm_pad(geo.N) + emm_tab_entries(geo.N);
    if (H == H_HYBRID) {
        const double2* src = reinterpret_cast<const double2*>(tabs + (fold ? 2 * span : 0));      // span is even
        double2* dst = reinterpret_cast<double2*>(s_tab);
        for (int k = threadIdx.x; k < span; k += FTPB) dst[k] = __ldg(src + k);
    }
    __syncthreads();
    EmmTabs tb; tb.inc = s_inc;
    tb.v = H == H_HYBRID ? s_tab + emm_pad(geo.N) : nullptr;
    tb.q = H == H_HYBRID ? s_tab + span + emm_pad(geo.N) : nullptr;
    // The a1 loop of a warp runs as long as ANY of its five roots has another child: order the block's roots by fan-out (largest
    // first) so that the roots sharing a warp have similar loop lengths.  Results are written by root index, so the order is invisible.
    __shared__ unsigned char s_cnt[FROOTS], s_perm[FROOTS];
    const int64_t blk0 = (int64_t)blockIdx.x * FROOTS;
    if (threadIdx.x < FROOTS) {
        int c = 0;
        if (blk0 + threadIdx.x < n_roots) {
            Game<C> g; unpack(g, state, (long)(blk0 + threadIdx.x)); g.player = SIDE_TL;
            if (!full_win<C>(g, geo)) c = legal_count(legal<C, SIDE_TL>(g, geol, movers<C>(g.tl, g.dice)));
        }
        s_cnt[threadIdx.x] = (unsigned char)c;
    }
    __syncthreads();
    if (threadIdx.x < FROOTS) {
        const int c = s_cnt[threadIdx.x];
        int rank = 0;
        for (int j = 0; j < FROOTS; j++) { const int cj = s_cnt[j]; rank += (cj > c) || (cj == c && j < (int)threadIdx.x); }
        s_perm[rank] = (unsigned char)threadIdx.x;
    }
    __syncthreads();
    const unsigned lane = threadIdx.x & 31u, grp = lane / 6u, d = lane - 6u * grp;
    const unsigned slot = (threadIdx.x >> 5) * 5u + (grp < 5u ? grp : 4u);
    const int64_t r = blk0 + s_perm[slot];
    const bool active = lane < 30u && r < n_roots;
    Game<C> root; setup<C>(root, geo, 1);                             // idle lanes search the initial position and drop the result
    if (active) { unpack(root, state, (long)r); root.player = SIDE_TL; }
    const bool fin = full_win<C>(root, geo);
    Game<C> work = root;
    if (fin) setup<C>(work, geo, 1);                                  // finished roots too: every summary the tables see is of a live position
    EmmChild<C> mine = emm2_child<C>(work, geol, tb, (int)d);
    if (fin || !active) mine.meta &= ~EMMC_EXISTS;
    const double ninf = -f64_inf();
    double best = ninf, alpha = ninf;
    int best_a = 0;
    const unsigned g0 = grp * 6u;
    for (unsigned a1 = 0; a1 < 6u; a1++) {
        EmmChild<C> ch;
        const unsigned src = g0 + a1;
        ch.br = __shfl_sync(0xffffffffu, mine.br, src);
        ch.htl = __shfl_sync(0xffffffffu, mine.htl, src); ch.hbr = __shfl_sync(0xffffffffu, mine.hbr, src);
        ch.meta = __shfl_sync(0xffffffffu, mine.meta, src);
        if (C::WORDS == 1) { ch.otl = __shfl_sync(0xffffffffu, mine.otl, src); ch.obr = __shfl_sync(0xffffffffu, mine.obr, src); ch.tl = 0; }
        else { ch.tl = __shfl_sync(0xffffffffu, mine.tl, src); ch.otl = 0; ch.obr = 0; }
        const bool exists = lane < 30u && (ch.meta & EMMC_EXISTS) != 0;          // lanes 30, 31: their shuffle source wraps to root 0's lanes
        if (!__any_sync(0xffffffffu, exists)) break;                  // children exist for a1 < #legal: no root of this warp has more
        double q;
        emm2_min<H == H_HYBRID, C>(ch, geol, tb, (int)d + 1, H, fold, pruned ? alpha : ninf, q);
        double e = 0.0;
#pragma unroll
        for (unsigned f = 0; f < 6u; f++) e = f64_add(e, __shfl_sync(0xffffffffu, q, g0 + f));      // minimax.py:64-73, face order
        if (ch.meta & EMMC_OVER) e = 10.0;                            // minimax.py:22-23 at the chance node: TOP_LEFT has just won
        if (exists) {
            if (e > best) { best = e; best_a = (int)a1; }
            alpha = alpha > best ? alpha : best;
        }
    }
    if (active && d == 0) {
        int la = -1, di = -1;
        if (fin) best = evaluate<C>(root, geo, H);                    // minimax.py:22-23 at the root: action None
        else { const Legal l = legal<C, SIDE_TL>(root, geol, movers<C>(root.tl, root.dice)); int sh; legal_pick(l, best_a, la, di, sh); }
        value[r] = best;
        if (action) reinterpret_cast<char2*>(action)[r] = make_char2((signed char)la, (signed char)di);
    }
}

// builds `tabs` for k_emm_fused from the value table vt[fold][pb][pc-1][nb][nc-1] (k_emm_table): per fold the padded values, then
// the padded quotients v / 6 (IEEE division, once per entry per process instead of once per MIN node); entry pad-1 = a lost game
__global__ void __launch_bounds__(TPB) k_emm_fused_table(const double* vt, double* tabs, int N) {
    const int pad = emm_pad(N), ent = emm_tab_entries(N), span = pad + ent;
    const int i = (int)(blockIdx.x * TPB + threadIdx.x);
    if (i >= 2 * span) return;
    const int fold = i / span, k = i - fold * span;
    double v = 0.0;
    if (k >= pad) v = vt[(long)fold * ent + (k - pad)];
    else if (k == pad - 1) v = -10.0;
    tabs[(long)fold * 2 * span + k] = v;
    tabs[(long)fold * 2 * span + span + k] = f64_div(v, 6.0);
}

// ------------------------------------------------------------------------------------------------
// K3 flat Monte-Carlo.  k_mc_expand turns roots into (root, child) pairs; k_playouts plays `count` random games
// per pair with one lane per game.  Lanes that finish pull the next game of their warp's contiguous range, so a
// warp stays full until its range is exhausted; results depend only on (seed, root id, child, playout index).
// ------------------------------------------------------------------------------------------------
// One (root, child) pair = one AoS record, so that a lane picks up its next game with ONE vector load:
//   P5: 16 B {packed state, root id | (slot * 8 + child) << 32};   P8: 32 B {word 0, word 1, the same meta word, unused}
struct PairBuf {
    void* rec;           // [pairs] records
    int* n_pairs;        // device counter
};
template <class C> struct PairRec;
template <> struct PairRec<P5> {
    static __device__ __forceinline__ void store(void* rec, u32 s, const Game<P5>& g, u32 rid, u32 sc) {
        reinterpret_cast<ulonglong2*>(rec)[s] = make_ulonglong2(pack_word(g), (u64)rid | ((u64)sc << 32));
    }
    static __device__ __forceinline__ void load(const void* rec, u32 s, Game<P5>& g, u32& rid, u32& sc) {
        const ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2*>(rec) + s);
        unpack_regs(g, v.x, 0); rid = (u32)v.y; sc = (u32)(v.y >> 32);
    }
};
template <> struct PairRec<P8> {
    static __device__ __forceinline__ void store(void* rec, u32 s, const Game<P8>& g, u32 rid, u32 sc) {
        u64 w0, w1; pack_words(g, w0, w1);
        ulonglong2* p = reinterpret_cast<ulonglong2*>(rec) + 2 * (size_t)s;
        p[0] = make_ulonglong2(w0, w1); p[1] = make_ulonglong2((u64)rid | ((u64)sc << 32), 0);
    }
    static __device__ __forceinline__ void load(const void* rec, u32 s, Game<P8>& g, u32& rid, u32& sc) {
        const ulonglong2* p = reinterpret_cast<const ulonglong2*>(rec) + 2 * (size_t)s;
        const ulonglong2 a = __ldg(p), b = __ldg(p + 1);
        unpack_words(g, a.x, a.y); rid = (u32)b.x; sc = (u32)(b.x >> 32);
    }
};

// `count`: playouts this launch sequence will play per pair.  A child that already ends the game (the agent's move wins,
// mcts.py:28: the playout loop body never runs) is settled here -- all its playouts have the same result -- and gets no record.
// zero_wins: the root's six counters are cleared here first (ewn_mc_predict: no memset launch before the decision).
// The three kernels of a decision (expand -> playouts -> arg-max) are launched with programmatic dependent launch: each lets its
// successor become resident early (griddepcontrol.launch_dependents) and touches global memory only after its predecessor has
// completed (griddepcontrol.wait); both instructions are no-ops in a launch without the attribute.
template <class C> __global__ void __launch_bounds__(TPB) k_mc_expand(const u64* state, const u32* root_ids, int64_t n, Geo geo,
                                                                      PairBuf pb, int32_t* n_children, u32 count, u32* wins, int zero_wins) {
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    if (zero_wins) {
        uint2* w2 = reinterpret_cast<uint2*>(wins + i * 6);
        w2[0] = make_uint2(0, 0); w2[1] = make_uint2(0, 0); w2[2] = make_uint2(0, 0);
    }
    Game<C> g; unpack(g, state, i, geo);
    g.player = SIDE_TL;
    int cnt = 0;
    if (!full_win<C>(g, geo)) {
        Legal l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
        cnt = legal_count(l);
        const u32 rid = root_ids ? root_ids[i] : (u32)i;
        for (int j = 0; j < cnt; j++) {
            Game<C> c = g;
            int la, di, sh; legal_pick(l, j, la, di, sh);
            apply_move<C, SIDE_TL>(c, sh, di);
            c.player = SIDE_BR;
            const u32 slot = (u32)(i * 6 + j);
            if (full_win<C>(c, geo)) {
                if (count && tl_won<C>(c, geo)) atomicAdd(&wins[slot], count);
                continue;
            }
            const int s = atomicAdd(pb.n_pairs, 1);
            PairRec<C>::store(pb.rec, (u32)s, c, rid, slot * 8u + (u32)j);
        }
    }
    if (n_children) n_children[i] = cnt;
}

// q = n / d for n < 2^31 as one wide multiply and a shift: m = floor(2^sh / d) + 1, sh = 31 + ceil(log2 d)
// (the error m * d - 2^sh is in (0, d] <= 2^(sh-31), so n * error < 2^sh for every n < 2^31)
struct Div31 { u32 m, sh; };
inline Div31 make_div31(u32 d) {
    int L = 0;
    while ((1ull << L) < d) L++;
    Div31 r; r.sh = 31u + (u32)L; r.m = (u32)(((1ull << r.sh) / d) + 1ull);
    return r;
}

// DIRECT: `pairs` is the caller's packed state array (ewn_playouts): Philox root id from root_ids (or the index), child tag 7,
// result slot = index, and a state that is already finished is settled when it is picked up.
// otherwise: `pairs` are the records k_mc_expand wrote (never finished states).
// n_pairs * count < 2^31 (the host slices larger jobs), so all game indexing is 32-bit.
// reset_hdr != nullptr: {pair counter, ticket} of the workspace header -- the last block to finish puts both back to zero, so the next
// decision's k_mc_expand finds a clean counter without a memset launch (every block has read the counter before it can finish).
// DYN (needs reset_hdr): chunks beyond a warp's first one are handed out by an atomic counter in the workspace header (word 2) instead
// of round-robin -- a warp that draws short games takes more chunks, so all warps finish within about one game of each other; the next
// chunk's number is fetched one chunk ahead (the atomic's result is first read when the current chunk runs out).  Which warp plays a
// game does not matter: results depend only on (root id, child, playout index).
#ifndef EWN_MC_TAIL
#define EWN_MC_TAIL 1
#endif
#ifndef EWN_MC_TAIL_R
#define EWN_MC_TAIL_R 2
#endif
constexpr int MC_TAIL_R = EWN_MC_TAIL_R;
struct TailRec { u64 tl, br; u32 rid, sc, pidx, j; };           // a live game on its way to another lane
template <class C, bool DIRECT, bool DYN> __global__ void __launch_bounds__(TPB) k_playouts(const void* pairs, const u32* root_ids, const int* n_pairs_dev,
                                                                                  u32 n_pairs_host, u32 first, u32 count, Div31 dv, u32* wins,
                                                                                  unsigned long long* plies, Geo geo, const RoundKeys rk, int* reset_hdr) {
    __shared__ __align__(16) unsigned char s_dirs[3 * C::LUT];
#if EWN_MC_TAIL
    __shared__ unsigned s_tailw, s_cnt[2][TPB / 32];
    __shared__ TailRec s_rec[TPB];
    if (threadIdx.x == 0) s_tailw = 0;
#endif
    asm volatile("griddepcontrol.launch_dependents;");
    const GeoLut geol = make_geo_lut<C, TPB>(geo, s_dirs);       // legal directions of a cube: one LDS (two cubes per ply)
    asm volatile("griddepcontrol.wait;" ::: "memory");
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const u32 warp_id = (blockIdx.x * TPB + threadIdx.x) >> 5;
    const u32 n_warps = (gridDim.x * TPB) >> 5;
    const u32 n_pairs = n_pairs_dev ? (u32)*n_pairs_dev : n_pairs_host;
    auto finish = [&]() {                                        // one ticket per warp; every warp has read the counter above
        if (reset_hdr && lane == 0) {
            __threadfence();
            if (atomicAdd(reset_hdr + 1, 1) == (int)n_warps - 1) { reset_hdr[0] = 0; reset_hdr[1] = 0; reset_hdr[2] = 0; }
        }
    };
    const u32 total = n_pairs * count;
    // Games are dealt to warps in chunks of CH consecutive games, round-robin (chunk c belongs to warp c % n_warps): playout
    // length depends on the root position, so a warp that owned ONE contiguous range of ~1,500 games of one or two roots
    // would finish much earlier or later than its neighbours; many small chunks average over the roots.  One warp-wide
    // pick-up per chunk is the finest useful grain and measured best at every job size (profiles/r1_playout_chunk_sweep.txt):
    // a decision split over 8 GPUs leaves only ~190 games per warp, and 128-game chunks then give half the warps twice the work.
    constexpr u32 CH = 32;
    const u32 n_chunks = (total + CH - 1) / CH;
    const u32 my_chunks = n_chunks > warp_id ? (n_chunks - warp_id + n_warps - 1) / n_warps : 0;
    u32 tnext = 0;                                             // warp-uniform cursor over the warp's local game indices
    const u32 tend = my_chunks * CH;
#if !EWN_MC_TAIL
    if (tend == 0) { finish(); return; }
#endif
    constexpr u32 NONE = 0xFFFFFFFFu;
    u32 base0 = warp_id * CH, used = 0, pending = 0;           // DYN: current chunk (first: chunk warp_id), games dealt from it, next chunk's ticket
    auto fetch = [&]() { if (lane == 0) pending = atomicAdd(reinterpret_cast<unsigned*>(reset_hdr) + 2, 1u); };
    if (DYN && tend != 0) fetch();

    Game<C> g; g.tl = C::ALL; g.br = C::ALL; g.dice = 1; g.player = SIDE_BR;
    bool busy = false;
    u32 rid = 0, child = 0, slot = 0, pidx = 0, j = 0;
    unsigned long long my_plies = 0;
    const unsigned lt = (1u << lane) - 1u;
    auto play = [&]() {                                            // one iteration (two plies) of this lane's game
        u32 blk[4];
        playout_block_rk(rk, rid, child, pidx, j++, blk);
        g.dice = 1 + (int)bounded(blk[0], 6u);
        bool over = random_ply<C, SIDE_BR>(g, geol, blk[1], nullptr, nullptr);
        my_plies++;
        if (!over) {
            g.dice = 1 + (int)bounded(blk[2], 6u);
            over = random_ply<C, SIDE_TL>(g, geol, blk[3], nullptr, nullptr);
            my_plies++;
        }
        if (over) {
            if (tl_won<C>(g, geo)) atomicAdd(&wins[slot], 1u);
            busy = false;
        }
    };
    for (; tend != 0;) {
        const unsigned need = __ballot_sync(0xffffffffu, !busy);
        if (need) {
            u32 gi = NONE;                                         // global game index this lane picks up
            if (DYN) {
                const u32 k = (u32)__popc(need);
                const bool roll = base0 != NONE && used + k >= CH; // this deal uses the current chunk up (warp-uniform)
                u32 nb = NONE;
                if (roll) { const u32 c = n_warps + __shfl_sync(0xffffffffu, pending, 0); nb = c < n_chunks ? c * CH : NONE; }
                if (!busy && base0 != NONE) {
                    const u32 r = used + (u32)__popc(need & lt);
                    if (r < CH) gi = base0 + r; else if (nb != NONE) gi = nb + (r - CH);
                }
                if (roll) { used = used + k - CH; base0 = nb; if (nb != NONE) fetch(); }
                else if (base0 != NONE) used += k;
            } else {
                const u32 t = tnext + (u32)__popc(need & lt);
                if (!busy && t < tend) gi = ((t / CH) * n_warps + warp_id) * CH + (t % CH);
            }
            if (gi != NONE) {
                if (gi < total) {
                    const u32 pair = (u32)(((u64)gi * dv.m) >> dv.sh);
                    pidx = first + (gi - pair * count);
                    j = 0;
                    busy = true;
                    if (DIRECT) {
                        unpack(g, reinterpret_cast<const u64*>(pairs), pair);
                        rid = root_ids ? root_ids[pair] : pair;
                        child = 7u; slot = pair;
                        if (full_win<C>(g, geo)) {               // mcts.py:28: the loop body never runs
                            if (tl_won<C>(g, geo)) atomicAdd(&wins[slot], 1u);
                            busy = false;
                        }
                    } else {
                        u32 sc;
                        PairRec<C>::load(pairs, pair, g, rid, sc);
                        child = sc & 7u; slot = sc >> 3;
                    }
                }
            }
            tnext += (u32)__popc(need);
#if EWN_MC_TAIL
            if (DYN ? base0 == NONE : tnext >= tend) break;        // no fresh games left for this warp: its live games go on in the tail
#else
            if (!__ballot_sync(0xffffffffu, busy) && (DYN ? base0 == NONE : tnext >= tend)) break;
#endif
        }
        if (busy) play();
    }
#if EWN_MC_TAIL
    // Tail.  Once every warp of the block has run out of fresh games, the block's live games are packed into as few warps as hold them
    // every MC_TAIL_R iterations (through shared memory: a game is its two side words, its Philox coordinates and its slot), and the
    // warps left without a game retire: at the 8-GPU share of a decision (eight games per lane) a third of the kernel used to be warps
    // iterating for one or two long games.  Which lane plays a game changes nothing: counters are atomics, streams are keyed by game.
    {
        constexpr int NW = TPB / 32;
        if (lane == 0) atomicAdd(&s_tailw, 1u);
        while (*reinterpret_cast<volatile unsigned*>(&s_tailw) < (unsigned)NW) {       // the others still deal fresh games: play on
            if (__ballot_sync(0xffffffffu, busy)) { if (busy) play(); } else __nanosleep(200);
        }
        const int warp = threadIdx.x >> 5;
        int nw = NW;                                               // warps still taking part (warps 0 .. nw-1)
        for (unsigned round = 0;; round++) {
            const unsigned live = __ballot_sync(0xffffffffu, busy);
            unsigned* cnt = s_cnt[round & 1];
            if (lane == 0) cnt[warp] = (unsigned)__popc(live);
            asm volatile("bar.sync 1, %0;" ::"r"(32 * nw) : "memory");
            unsigned before = 0, total = 0;
            for (int w = 0; w < nw; w++) { const unsigned c = cnt[w]; before += w < warp ? c : 0u; total += c; }
            if (total == 0) break;
            const int nw_new = (int)((total + 31u) >> 5);
            if (nw_new < nw) {                                     // packing frees at least one warp
                if (busy) {
                    TailRec& r = s_rec[before + (unsigned)__popc(live & lt)];
                    r.tl = (u64)g.tl; r.br = (u64)g.br; r.rid = rid; r.sc = (slot << 3) | child; r.pidx = pidx; r.j = j;
                }
                asm volatile("bar.sync 1, %0;" ::"r"(32 * nw) : "memory");
                const unsigned t = (unsigned)(warp * 32 + lane);
                busy = t < total;
                if (busy) {
                    const TailRec& r = s_rec[t];
                    g.tl = (typename C::W)r.tl; g.br = (typename C::W)r.br; rid = r.rid; child = r.sc & 7u; slot = r.sc >> 3; pidx = r.pidx; j = r.j;
                }
                if (warp >= nw_new) break;
                nw = nw_new;
            }
            for (int it = 0; it < MC_TAIL_R; it++) {
                if (!__ballot_sync(0xffffffffu, busy)) break;
                if (busy) play();
            }
        }
    }
#endif
    if (plies) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) my_plies += __shfl_xor_sync(0xffffffffu, my_plies, o);
        if (lane == 0) atomicAdd(plies, my_plies);
    }
    finish();
}

// UCT at the root, one thread per root and round (ewn_uct_search).  res = TOP_LEFT wins of the playouts of the previous round's child.
template <class C> __global__ void __launch_bounds__(TPB) k_uct_round(const u64* state, const u32* root_ids, int64_t n, Geo geo, u32* wins, u32* visits,
                                                                      int32_t* chosen, u32* res, u64* child, u32* child_rid, u32 batch, double c,
                                                                      double ln_total, int first_round) {
    const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    Game<C> g, ch; unpack(g, state, i, geo);
    g.player = SIDE_TL;
    u32 w[6], v[6];
#pragma unroll
    for (int j = 0; j < 6; j++) { w[j] = first_round ? 0u : wins[i * 6 + j]; v[j] = first_round ? 0u : visits[i * 6 + j]; }
    int nch;
    setup<C>(ch, geo, 1);
    const int j = uct_round<C>(g, geo, w, v, first_round ? -1 : chosen[i], res[i], batch, c, ln_total, ch, nch);
#pragma unroll
    for (int k = 0; k < 6; k++) { wins[i * 6 + k] = w[k]; visits[i * 6 + k] = v[k]; }
    chosen[i] = j;
    res[i] = 0u;
    pack(ch, child, i, geo);                                  // a finished root hands out the initial position; its result is dropped
    child_rid[i] = (root_ids ? root_ids[i] : (u32)i) * 6u + (u32)(j < 0 ? 0 : j);
}
template <class C> __global__ void __launch_bounds__(TPB) k_uct_final(const u64* state, int64_t n, Geo geo, u32* wins, u32* visits, const int32_t* chosen,
                                                                      const u32* res, u32 batch, int8_t* action) {
    const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    Game<C> g; unpack(g, state, i, geo);
    int la = -1, di = -1;
    if (!full_win<C>(g, geo)) {
        u32 w[6], v[6];
#pragma unroll
        for (int j = 0; j < 6; j++) { w[j] = wins[i * 6 + j]; v[j] = visits[i * 6 + j]; }
        const int p = chosen[i];
#pragma unroll
        for (int j = 0; j < 6; j++) if (j == p) { w[j] += res[i]; v[j] += batch; }
#pragma unroll
        for (int j = 0; j < 6; j++) { wins[i * 6 + j] = w[j]; visits[i * 6 + j] = v[j]; }
        const Legal l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
        int sh; legal_pick(l, uct_best(w, v, legal_count(l)), la, di, sh);
    }
    if (action) { action[2 * i] = (int8_t)la; action[2 * i + 1] = (int8_t)di; }
}

template <class C> __global__ void __launch_bounds__(TPB) k_mc_best(const u64* state, const u32* wins, int8_t* action, int64_t n, Geo geo) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    Game<C> g; unpack(g, state, i, geo);
    int la = -1, di = -1;
    if (!full_win<C>(g, geo)) {
        Legal l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
        int cnt = legal_count(l), best = 0;
        for (int j = 1; j < cnt; j++) if (wins[i * 6 + j] > wins[i * 6 + best]) best = j;      // np.argmax: first maximum
        int sh; legal_pick(l, best, la, di, sh);
    }
    action[2 * i] = (int8_t)la; action[2 * i + 1] = (int8_t)di;
}

// root-statistics merge fused with the arg-max: one-shot all-reduce of the win counters over NVLink peer loads
struct PeerSet { const u32* wins[EWN_MAX_PEERS]; u32* flags[EWN_MAX_PEERS]; };

__device__ __forceinline__ void st_release_sys(u32* p, u32 v) { asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ u32 ld_acquire_sys(const u32* p) { u32 v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ unsigned long long global_ns() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

template <class C> __global__ void __launch_bounds__(TPB) k_mc_best_peers(const u64* state, PeerSet ps, int rank, int world, u32 epoch,
                                                                           unsigned long long timeout_ns, u32* total, int8_t* action, int64_t n, Geo geo) {
    // (selects instead of a run-time index: kernel parameters indexed dynamically would be copied to local memory)
    u32* mine = ps.flags[0]; u32* theirs = ps.flags[0];
#pragma unroll
    for (int r = 1; r < EWN_MAX_PEERS; r++) {
        if (r == rank) mine = ps.flags[r];
        if (r == (int)threadIdx.x) theirs = ps.flags[r];
    }
    // publish: the playout kernel that filled my counters precedes this kernel on the stream, so they are complete
    if (blockIdx.x == 0 && threadIdx.x < world) { __threadfence_system(); st_release_sys(theirs + rank, epoch); }
    // wait for the same epoch of every rank; the flag words live in MY memory, the spin never crosses NVLink
    int timed_out = 0;
    if (threadIdx.x < world) {
        const u32* f = mine + threadIdx.x;
        unsigned long long t0 = global_ns();
        while ((int)(ld_acquire_sys(f) - epoch) < 0) {
            if (global_ns() - t0 > timeout_ns) { atomicExch(mine + EWN_PEER_ERR_WORD, epoch); timed_out = 1; break; }
            __nanosleep(64);
        }
    }
    const int dead = __syncthreads_or(timed_out);
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    if (dead) {                                                  // a peer never published this epoch: no move, and the caller can see it
        action[2 * i] = (int8_t)-1; action[2 * i + 1] = (int8_t)-1;
        return;
    }
    uint2 w[EWN_MAX_PEERS][3];                                   // 24-byte rows: three 8-byte peer loads per rank, all in flight together
#pragma unroll
    for (int r = 0; r < EWN_MAX_PEERS; r++)
        if (r < world) {
            const uint2* src = reinterpret_cast<const uint2*>(ps.wins[r]) + i * 3;
#pragma unroll
            for (int k = 0; k < 3; k++) w[r][k] = __ldcg(src + k);
        }
    u32 sum[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int r = 0; r < EWN_MAX_PEERS; r++)
        if (r < world) {
#pragma unroll
            for (int k = 0; k < 3; k++) { sum[2 * k] += w[r][k].x; sum[2 * k + 1] += w[r][k].y; }
        }
    if (total) {
        uint2* dst = reinterpret_cast<uint2*>(total) + i * 3;
#pragma unroll
        for (int k = 0; k < 3; k++) dst[k] = make_uint2(sum[2 * k], sum[2 * k + 1]);
    }
    Game<C> g; unpack(g, state, i, geo);
    int la = -1, di = -1;
    if (!full_win<C>(g, geo)) {
        Legal l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
        int cnt = legal_count(l), best = 0;
        u32 bv = sum[0];
#pragma unroll
        for (int j = 1; j < 6; j++) if (j < cnt && sum[j] > bv) { best = j; bv = sum[j]; }       // np.argmax: first maximum
        int sh; legal_pick(l, best, la, di, sh);
    }
    action[2 * i] = (int8_t)la; action[2 * i + 1] = (int8_t)di;
}

// ------------------------------------------------------------------------------------------------
// K5  rollout collection with a policy in the loop (train.py:35-63 "MultiInputPolicy", :144-155 model.learn -> collect_rollouts).
// k_policy_forward is SB3's ActorCriticPolicy.forward for MultiInputPolicy on this env's observation -- CombinedExtractor (Flatten
// of the int16 board, one-hot of the Discrete die), two Tanh towers of two layers each (pi / vf), action_net -> the logits of
// MultiDiscrete([2, 3]), value_net -> V(s) -- and the draw from the two categoricals, in ONE launch, reading the torch
// parameters where they lie (Linear.weight is [out][in], fp32).  With the step kernel it makes a rollout step two launches, and a
// rollout one CUDA graph.
// A block serves PE = 32 envs with 256 threads.  The 2 x 64 hidden units of the two towers are one 128-wide output row; a thread
// owns FOUR adjacent units (lane -> unit group, so a warp spans the whole row) for FOUR envs (warp -> env group): 16 accumulators,
// and per k step one 16-byte shared load of weights (distinct per lane, conflict-free) and one of activations (warp-uniform) feed
// 16 FMAs.  At 4096 envs the launch is 128 blocks on 148 SMs, so it is latency-, not throughput-bound (profiles/r2_k5_policy.txt: one
// block alone takes as long as 128); 4 x 8 tiles at 128 threads and 4 x 2 tiles at 512 threads both measured slower.  Both layers'
// weights are requested up front with cp.async (4-byte copies, transposed on the way: wt[k][tower * 64 + unit]) -- one round trip
// in all, not one per element.  fp32 FMA in k order; tanh on the fast exponential.
// ------------------------------------------------------------------------------------------------
#ifndef EWN_PT
#define EWN_PT 256
#endif
#ifndef EWN_PE
#define EWN_PE 32
#endif
constexpr int PE = EWN_PE, PH = 64, PO = 2 * PH, PWS = PO + 4, PT = EWN_PT, PEV = PE / (PT / 32);   // PWS: row stride of a transposed weight
                                                                   // tile (16-byte aligned rows); PEV = 4 (2) envs per thread at 256 (512) threads
constexpr int PHS = PE + 4;                                        // row stride of the activation tiles h1 / h2 (see pol_row)
// Output row (tower * 64 + unit) number j of lane l: the four rows of a thread are 16 apart inside ONE tower and adjacent lanes own
// adjacent rows, so that the 16-byte activation stores of a quarter-warp fall 36 words apart -- eight different bank groups -- instead of
// 128 words apart (rows 4 l + j: a 32-way conflict on every store, 1 us of the 8.9 us step of k_rollout).  The weight tiles hold the
// column of row pol_row(l, j) at position 4 l + j, so a thread's four weights of a k step are still one 16-byte load.
__host__ __device__ __forceinline__ constexpr int pol_row(int l, int j) { return (l >> 4) * PH + (l & 15) + 16 * j; }
__host__ __device__ __forceinline__ constexpr int pol_col(int r) { return 4 * ((r / PH) * 16 + (r % PH) % 16) + (r % PH) / 16; }
struct PolicyNet {
    const float *w1[2], *b1[2], *w2[2], *b2[2];      // [0] = pi tower, [1] = vf tower: Linear(F, 64), Linear(64, 64)
    const float *wa, *ba, *wv, *bv;                  // action_net Linear(64, 5), value_net Linear(64, 1)
};
enum { TAG_POLICY = 3 };
// shared memory of one block: the two layers' weights of both towers (transposed), features, activations, head outputs, small parameters
struct PolSmem { float *wt1, *wt2, *x, *h1, *h2, *outv, *small; };
constexpr int POL_SMALL = 10 * PH + 8;               // b1[2][64], b2[2][64], action_net w[5][64], value_net w[64], their six biases
inline size_t policy_smem(int F) { return ((size_t)F * PE + 2 * (size_t)PO * PHS + 6 * PE + POL_SMALL + (size_t)(F + PH) * PWS) * sizeof(float); }
__device__ __forceinline__ PolSmem pol_carve(float* base, int F) {
    PolSmem m;
    m.wt1 = base;                                    // [F][PWS]   layer 1, both towers, transposed
    m.wt2 = m.wt1 + (size_t)F * PWS;                 // [PH][PWS]  layer 2
    m.x = m.wt2 + (size_t)PH * PWS;                  // [F][PE]    features, feature-major
    m.h1 = m.x + (size_t)F * PE;                     // [PO][PHS]  row = tower * 64 + unit
    m.h2 = m.h1 + (size_t)PO * PHS;                  // [PO][PHS]
    m.outv = m.h2 + (size_t)PO * PHS;                // [6][PE]: five logits and the value
    m.small = m.outv + 6 * PE;
    return m;
}
// request all parameters (no wait: cp.async.wait_all + __syncthreads by the caller)
__device__ __forceinline__ void pol_stage(const PolSmem& m, const PolicyNet& net, int F) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // W[tower][unit][k] -> dst[k][tower * 64 + unit].  Rows of a multiple of four floats (7x7: 56, 5x5: 32, layer 2: 64): a thread takes
    // 16-byte pieces of ONE output row (lanes = adjacent rows: 32 sectors per load, twice the minimum, but conflict-free transposed
    // stores), all loads of a layer in flight before the first store.  Other widths: 4-byte cp.async, a warp per row.
    auto stage = [&](float* dst, const float* wa_, const float* wb_, int K) {
        if ((K & 3) == 0 && K <= 64) {
            constexpr int TPR = PT / PO, NQ = 16 / TPR;                     // threads per row, 16-byte pieces per thread
            const int r = tid & (PO - 1), c0 = tid / PO, nc = K >> 2;
            const float4* src = reinterpret_cast<const float4*>(r < PH ? wa_ + (size_t)r * K : wb_ + (size_t)(r - PH) * K);
            float4 v[NQ];
#pragma unroll
            for (int q = 0; q < NQ; q++) { const int c = c0 + TPR * q; if (c < nc) v[q] = __ldg(src + c); }
#pragma unroll
            for (int q = 0; q < NQ; q++) {
                const int c = c0 + TPR * q;
                if (c < nc) { float* d = dst + (size_t)(4 * c) * PWS + pol_col(r); d[0] = v[q].x; d[PWS] = v[q].y; d[2 * PWS] = v[q].z; d[3 * PWS] = v[q].w; }
            }
        } else {
            for (int r = warp; r < PO; r += PT / 32) {
                const float* src = (r < PH ? wa_ + (size_t)r * K : wb_ + (size_t)(r - PH) * K);
                const unsigned d = (unsigned)__cvta_generic_to_shared(dst + pol_col(r));
                for (int k = lane; k < K; k += 32) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + (unsigned)(k * PWS * 4)), "l"(src + k) : "memory");
            }
        }
    };
    // biases of the four hidden layers and the two heads' weights and biases: small, but a load from L2 in the middle of a phase is a
    // round trip nobody hides at one or two warps per scheduler
    for (int idx = tid; idx < 4 * PH + 6 * PH + 6; idx += PT) {
        const float* src = idx < PH ? net.b1[0] + idx : idx < 2 * PH ? net.b1[1] + (idx - PH) : idx < 3 * PH ? net.b2[0] + (idx - 2 * PH)
                         : idx < 4 * PH ? net.b2[1] + (idx - 3 * PH) : idx < 9 * PH ? net.wa + (idx - 4 * PH) : idx < 10 * PH ? net.wv + (idx - 9 * PH)
                         : idx < 10 * PH + 5 ? net.ba + (idx - 10 * PH) : net.bv;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(m.small + idx)), "l"(src) : "memory");
    }
#if !defined(EWN_EXP_POL_NOSTAGE)
    stage(m.wt1, net.w1[0], net.w1[1], F);
    stage(m.wt2, net.w2[0], net.w2[1], PH);
#endif
}
// x (features of the block's PE envs) and the parameters are in shared memory and visible: two hidden layers of both towers and the
// heads -> m.outv[6][PE].  Ends with a block barrier.
__device__ __forceinline__ void pol_layers(const PolSmem& m, int F) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tower = lane >> 4;                                   // this thread's four units: rows pol_row(lane, 0..3), all in this tower
    static_assert(PEV == 2 || PEV == 4, "one float2 / float4 of activations per k step");
    float acc[4][PEV];
    auto layer = [&](const float* wl, const float* bias, const float* in, int K, int stride) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const float b = bias[pol_row(lane, j)];
#pragma unroll
            for (int e = 0; e < PEV; e++) acc[j][e] = b;
        }
        const float* w = wl + 4 * lane;
        in += PEV * warp;
#if defined(EWN_EXP_POL_NOLAYER)
        K = 0;
#endif
#ifndef EWN_POL_UNROLL
#define EWN_POL_UNROLL 8
#endif
        constexpr int PUN = EWN_POL_UNROLL;
#pragma unroll PUN
        for (int k = 0; k < K; k++) {
            const float4 w4 = *reinterpret_cast<const float4*>(w + (size_t)k * PWS);
            float av[PEV];
            if (PEV == 4) { const float4 a0 = *reinterpret_cast<const float4*>(in + (size_t)k * stride); av[0] = a0.x; av[1] = a0.y; av[PEV - 2] = a0.z; av[PEV - 1] = a0.w; }
            else { const float2 a0 = *reinterpret_cast<const float2*>(in + (size_t)k * stride); av[0] = a0.x; av[1] = a0.y; }
            const float wv[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
            for (int j = 0; j < 4; j++)
#pragma unroll
                for (int e = 0; e < PEV; e++) acc[j][e] = fmaf(wv[j], av[e], acc[j][e]);
        }
    };
    // tanh(v) = 1 - 2 / (1 + e^(2v)) on the fast exponential and reciprocal: absolute error ~1e-7 (the accurate tanhf is ~35 instructions
    // and was a quarter of the kernel); saturates correctly (e^(2v) = inf -> 1, = 0 -> -1)
    auto th = [](float v) { return 1.f - __fdividef(2.f, __expf(2.f * v) + 1.f); };
    auto squash = [&](float* h) {                                  // tanh of the sums -> h[tower * 64 + unit][env]
#pragma unroll
        for (int j = 0; j < 4; j++)
            if (PEV == 4) *reinterpret_cast<float4*>(h + (size_t)pol_row(lane, j) * PHS + PEV * warp) = make_float4(th(acc[j][0]), th(acc[j][1]), th(acc[j][PEV - 2]), th(acc[j][PEV - 1]));
            else *reinterpret_cast<float2*>(h + (size_t)pol_row(lane, j) * PHS + PEV * warp) = make_float2(th(acc[j][0]), th(acc[j][1]));
    };
    layer(m.wt1, m.small, m.x, F, PE);
    squash(m.h1);
    __syncthreads();
    layer(m.wt2, m.small + 2 * PH, m.h1 + (size_t)tower * PH * PHS, PH, PHS);
    squash(m.h2);
    __syncthreads();
    for (int idx = tid; idx < 6 * PE; idx += PT) {               // heads: (output o, env e)
        const int o = idx / PE, e = idx - o * PE;
        const float* w = m.small + 4 * PH + o * PH;               // action_net rows 0..4, then value_net
        const float* in = m.h2 + (size_t)(o < 5 ? 0 : PH) * PHS + e;
        float a = m.small[10 * PH + o];
#pragma unroll 8
        for (int k = 0; k < PH; k++) a = fmaf(w[k], in[(size_t)k * PHS], a);
        m.outv[o * PE + e] = a;
    }
    __syncthreads();
}
// the draw for env e of the block (global index i) from m.outv: Categorical(logits) x 2 (normalised logits = logits - logsumexp, probs =
// softmax: torch.distributions), Philox block of (seed; env id, step), tag POLICY.  -> action pair, log-probability
__device__ __forceinline__ void pol_draw(const float* outv, int e, u64 env, u64 seed, u64 st, int deterministic, int& a0, int& a1, float& lp) {
    const float l0 = outv[e], l1 = outv[PE + e], m0 = outv[2 * PE + e], m1 = outv[3 * PE + e], m2 = outv[4 * PE + e];
    const float mx = fmaxf(l0, l1), z = mx + logf(expf(l0 - mx) + expf(l1 - mx));
    const float my = fmaxf(m0, fmaxf(m1, m2)), zy = my + logf(expf(m0 - my) + expf(m1 - my) + expf(m2 - my));
    if (deterministic) {                                          // probs.argmax: the first maximum
        a0 = l1 > l0 ? 1 : 0;
        a1 = (m1 > m0 ? (m2 > m1 ? 2 : 1) : (m2 > m0 ? 2 : 0));
    } else {
        u32 blk[4];
        philox4x32_10((u32)env, (u32)(env >> 32), (u32)st, ((u32)(st >> 32) & 0x0FFFFFFFu) | ((u32)TAG_POLICY << 28), (u32)seed, (u32)(seed >> 32), blk);
        const float r0 = (float)(blk[0] >> 8) * (1.0f / 16777216.0f), r1 = (float)(blk[1] >> 8) * (1.0f / 16777216.0f);   // [0, 1)
        a0 = r0 < expf(l0 - z) ? 0 : 1;
        const float q0 = expf(m0 - zy), q1 = expf(m1 - zy);
        a1 = r1 < q0 ? 0 : (r1 < q0 + q1 ? 1 : 2);
    }
    lp = ((a0 ? l1 : l0) - z) + ((a1 == 0 ? m0 : a1 == 1 ? m1 : m2) - zy);
}

__global__ void __launch_bounds__(PT) k_policy_forward(const int16_t* board, const int64_t* dice, int64_t n, int NN, int D, PolicyNet net,
                                                       u64 seed, u64 step, const unsigned long long* step_dev, u64 env_id0, int deterministic,
                                                       int64_t* actions, float* values, float* logp, float* logits_out) {
    extern __shared__ float4 pol_smem[];
    const int F = NN + D;
    const PolSmem m = pol_carve(reinterpret_cast<float*>(pol_smem), F);
    const int tid = threadIdx.x;
    pol_stage(m, net, F);
    // a block walks env groups with a grid stride (the host launches at most two blocks per SM): the parameters are staged ONCE per block,
    // not once per 32 envs -- at 4,096 envs every block has one group; at 1 Mi envs staging 61 KB per group was two thirds of the time
    const int64_t groups = (n + PE - 1) / PE;
    for (int64_t grp = blockIdx.x; grp < groups; grp += gridDim.x) {
        const int64_t e0 = grp * PE;
        const int ne = (int)min((int64_t)PE, n - e0);
        for (int idx = tid; idx < PE * NN; idx += PT) {          // Flatten(board) as float: SB3 does not rescale a non-image Box
            const int e = idx / NN, k = idx - e * NN;
            m.x[k * PE + e] = e < ne ? (float)board[(e0 + e) * NN + k] : 0.f;
        }
        for (int idx = tid; idx < PE * D; idx += PT) {           // one-hot(Discrete(cube_num + 1, start = 1)): index dice - 1
            const int e = idx / D, d = idx - e * D;
            int v = e < ne ? (int)dice[e0 + e] - 1 : -1;
            v = v < 0 ? 0 : (v > D - 1 ? D - 1 : v);
            m.x[(NN + d) * PE + e] = (e < ne && d == v) ? 1.f : 0.f;
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncthreads();
        pol_layers(m, F);
        if (tid < ne) {
            const int e = tid;
            const int64_t i = e0 + e;
            if (values) values[i] = m.outv[5 * PE + e];
            if (logits_out) { float* q = logits_out + 5 * i; for (int o = 0; o < 5; o++) q[o] = m.outv[o * PE + e]; }
            if (actions) {
                int a0, a1; float lp;
                pol_draw(m.outv, e, env_id0 + (u64)i, seed, step + (step_dev ? *step_dev : 0ull), deterministic, a0, a1, lp);
                actions[2 * i] = a0; actions[2 * i + 1] = a1;
                if (logp) logp[i] = lp;
            }
        }
        __syncthreads();                                          // m.outv is read above, m.x is rewritten by the next group
    }
}

// K5p: the whole rollout of a block's 32 envs in ONE launch.  The games of one block never need another block's data, so policy forward
// and env step alternate on chip for all T steps: parameters staged once, no launch gaps (two dependent launches per step cost ~7 us of
// gaps alone at this size), games / tolerance / previous score in the registers of warp 0 (lane = env) from the first step to the last.
// Per step: warp 0 stages its 32 boards (the observation of step t) and writes them to the rollout buffer; all threads turn the tile
// into features; layers + heads (pol_layers); warp 0 draws the actions, plays the env step (step_one_cfg: the same function, Philox
// blocks and arithmetic as k_step) and writes actions / values / log-probs / rewards / dones of step t.  After the last step: the
// observation and V of slot T, the games back to HBM, GAE over the block's own columns.  The buffers are bit-identical to the
// two-launches-per-step collector's (tests/test_policy_kernel.py).
struct RolloutArgs {
    u64* state; int64_t n; int T;
    int16_t* board; int64_t* dice;                   // [T + 1][n][N][N], [T + 1][n]
    int64_t* actions; float* rewards; uint8_t* dones; float* values; float* logp;    // [T][n][2], [T][n] ...
    float* last_values; float* adv; float* ret;      // [n], [T][n], [T][n]
    int32_t* tolerance; double* prev_score; unsigned long long* stats;
    const unsigned long long* step_dev;
    u64 policy_seed; float gamma, gamma_lambda;
};
template <class C, int NT, bool SHAPED> __global__ void __launch_bounds__(PT)
k_rollout(const __grid_constant__ RolloutArgs ra, const __grid_constant__ StepParams sp, const __grid_constant__ PolicyNet net, const int D) {
    extern __shared__ float4 pol_smem[];
    Geo geo = sp.geo;
    if (NT) { geo.N = NT; geo.corner = C::cell(NT - 1, NT - 1); }
    const int NN = geo.N * geo.N, F = NN + D;
    const PolSmem m = pol_carve(reinterpret_cast<float*>(pol_smem), F);
    int16_t* const tile = reinterpret_cast<int16_t*>(m.small + POL_SMALL);            // [32][NN] int16 (16-byte aligned: every part before it is)
    int* const s_dice = reinterpret_cast<int*>(reinterpret_cast<char*>(tile) + (((size_t)32 * NN * 2 + 15) & ~(size_t)15));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t n = ra.n, e0 = (int64_t)blockIdx.x * PE, i = e0 + lane;
    const bool mine = warp == 0 && lane < PE && i < n;
    const u64 step0 = sp.step + (ra.step_dev ? *ra.step_dev : 0ull);
    constexpr bool has_net = true;                              // (without a network ewn_rollout runs k_rollout_env)
    if (has_net) pol_stage(m, net, F);
    Game<C> g; g.tl = C::all(); g.br = C::all(); g.dice = 1; g.player = 0;
    int tol = 0; double prev = 0.0;
    unsigned won = 0, lost = 0, fin_n = 0;
    if (mine) {
        unpack(g, ra.state, i, sp.geo);
        if (SHAPED) { tol = ra.tolerance[i]; prev = ra.prev_score[i]; }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    for (int t = 0; t <= ra.T; t++) {
        if (warp == 0) {                                          // observation of slot t: the block's 32 rows are contiguous in the buffer
            stage_rows<C, NT>(g, mine, geo, tile, lane);
            write_rows_dense<C, NT>(ra.board + (size_t)t * n * NN, e0, min(n, e0 + PE), geo, tile, lane);
            if (mine) ra.dice[(size_t)t * n + i] = g.dice;
            s_dice[lane] = mine ? g.dice : 0;
        }
        __syncthreads();
        for (int idx = tid; idx < PE * NN; idx += PT) {          // x[k][e] = board[e][k] (rows of envs beyond n are zero)
            const int k = idx / PE, e = idx % PE;
            m.x[idx] = (float)tile[e * NN + k];
        }
        for (int idx = tid; idx < PE * D; idx += PT) {
            const int d = idx / PE, e = idx % PE;
            int v = s_dice[e] - 1;
            const bool live = v >= 0;
            v = v < 0 ? 0 : (v > D - 1 ? D - 1 : v);
            m.x[(NN + d) * PE + e] = (live && d == v) ? 1.f : 0.f;
        }
        __syncthreads();
        pol_layers(m, F);
        if (t == ra.T) {                                          // V of the observation after the last step
            if (mine) ra.last_values[i] = m.outv[5 * PE + lane];
            break;
        }
        if (mine) {
            const size_t k = (size_t)t * n + i;
            int a0, a1; float lp;
            pol_draw(m.outv, lane, sp.env_id0 + (u64)i, ra.policy_seed, step0 + (u64)t, 0, a0, a1, lp);
            ra.actions[2 * k] = a0; ra.actions[2 * k + 1] = a1;
            ra.values[k] = m.outv[5 * PE + lane]; ra.logp[k] = lp;
            Game<C> term = g;
            StepOut out; out.term = 0; out.code = CODE_ONGOING;
            StepInject inj; inj.dice_opp = 1; inj.o_larger = 0; inj.o_dir = 0; inj.dice_next = 1;
            step_one_cfg<C>(g, term, sp, geo, 0, SHAPED ? 1 : 0, 0, step0 + (u64)t, (long)i, a0, a1, inj, &tol, &prev, out);
            ra.rewards[k] = (float)out.reward;
            ra.dones[k] = (uint8_t)out.term;
            if (out.term) { fin_n++; won += out.code == CODE_AGENT_WON; lost += out.code == CODE_AGENT_LOST; }
        }
        // (the next iteration's first barrier orders this step's reads of m.outv before the next pol_layers writes it)
    }
    if (warp != 0) return;
    if (mine) {
        pack(g, ra.state, i, sp.geo);
        if (SHAPED) { ra.tolerance[i] = tol; ra.prev_score[i] = prev; }
        // RolloutBuffer.compute_returns_and_advantage over this env's column (k_gae's arithmetic; the thread reads its own stores)
        float gae = 0.f, next_v = has_net ? m.outv[5 * PE + lane] : 0.f;
        for (int t = has_net ? ra.T - 1 : -1; t >= 0; t--) {
            const size_t k = (size_t)t * n + i;
            const float nt = __fsub_rn(1.0f, ra.dones[k] ? 1.0f : 0.0f), v = ra.values[k];
            const float delta = __fsub_rn(__fadd_rn(ra.rewards[k], __fmul_rn(__fmul_rn(ra.gamma, next_v), nt)), v);
            gae = __fadd_rn(delta, __fmul_rn(__fmul_rn(ra.gamma_lambda, nt), gae));
            ra.adv[k] = gae;
            ra.ret[k] = __fadd_rn(gae, v);
            next_v = v;
        }
    }
    if (ra.stats) {
        const unsigned f = __reduce_add_sync(0xffffffffu, fin_n), w = __reduce_add_sync(0xffffffffu, won), l = __reduce_add_sync(0xffffffffu, lost);
        if (lane < 4 && f) {
            const unsigned v = lane == 0 ? w : lane == 1 ? l : lane == 2 ? f - w - l : f;
            if (v) atomicAdd(ra.stats + lane, (unsigned long long)v);
        }
    }
}
// ewn_rollout without a network at any batch size: a thread owns a game for all T steps (actions read from the buffer), a warp stages and
// writes its 32 observation rows per step exactly as k_step does -- T env steps per launch, the games in registers, no per-step launch
// ramp / tail and no state round trip through HBM (79 B per env step on 5x5 instead of 96).
#ifndef EWN_RTPB
#define EWN_RTPB 128
#endif
constexpr int RTPB = EWN_RTPB;
template <class C, int NT, bool SHAPED> __global__ void __launch_bounds__(RTPB)
k_rollout_env(const __grid_constant__ RolloutArgs ra, const __grid_constant__ StepParams sp) {
    extern __shared__ uint4 smem_raw[];
    Geo geo = sp.geo;
    if (NT) { geo.N = NT; geo.corner = C::cell(NT - 1, NT - 1); }
    const int NN = geo.N * geo.N;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int16_t* const tile = reinterpret_cast<int16_t*>(smem_raw) + (size_t)warp * 32 * NN;
    const int64_t n = ra.n, i = (int64_t)blockIdx.x * RTPB + tid, warp_env0 = i - lane;
    if (warp_env0 >= n) return;                                   // warp-uniform
    const bool mine = i < n;
    const u64 step0 = sp.step + (ra.step_dev ? *ra.step_dev : 0ull);
    Game<C> g; g.tl = C::all(); g.br = C::all(); g.dice = 1; g.player = 0;
    int tol = 0; double prev = 0.0;
    unsigned won = 0, lost = 0, fin_n = 0;
    if (mine) {
        unpack(g, ra.state, i, sp.geo);
        if (SHAPED) { tol = ra.tolerance[i]; prev = ra.prev_score[i]; }
    }
    const int ar = sp.agent_random;                               // the agent plays uniformly over its legal list inside the kernel: `actions` is written
    longlong2 act = make_longlong2(0, 0);
    if (mine && !ar) act = __ldcs(reinterpret_cast<const longlong2*>(ra.actions) + i);
    for (int t = 0; t <= ra.T; t++) {
        stage_rows<C, NT>(g, mine, geo, tile, lane);
        write_rows_dense<C, NT>(ra.board + (size_t)t * n * NN, warp_env0, n, geo, tile, lane);
        if (mine) __stcs(reinterpret_cast<long long*>(ra.dice) + (size_t)t * n + i, (long long)g.dice);
        if (t == ra.T) break;
        if (mine) {
            const size_t k = (size_t)t * n + i;
            const int a0 = (int)act.x, a1 = (act.y < 0 || act.y > 2) ? 3 : (int)act.y;              // as k_step reads an int64 action pair
            if (!ar && t + 1 < ra.T) act = __ldcs(reinterpret_cast<const longlong2*>(ra.actions) + k + n);   // the next step's action: in flight during this step
            Game<C> term = g;
            StepOut out; out.term = 0; out.code = CODE_ONGOING;
            StepInject inj; inj.dice_opp = 1; inj.o_larger = 0; inj.o_dir = 0; inj.dice_next = 1;
            step_one_cfg<C>(g, term, sp, geo, 0, SHAPED ? 1 : 0, ar, step0 + (u64)t, (long)i, a0, a1, inj, &tol, &prev, out);
            if (ar && ra.actions) __stcs(reinterpret_cast<longlong2*>(ra.actions) + k, make_longlong2(out.a_larger, out.a_dir));
            __stcs(ra.rewards + k, (float)out.reward);
            __stcs(ra.dones + k, (uint8_t)out.term);
            if (out.term) { fin_n++; won += out.code == CODE_AGENT_WON; lost += out.code == CODE_AGENT_LOST; }
        }
    }
    if (mine) {
        pack(g, ra.state, i, sp.geo);
        if (SHAPED) { ra.tolerance[i] = tol; ra.prev_score[i] = prev; }
    }
    if (ra.stats) {
        const unsigned f = __reduce_add_sync(0xffffffffu, fin_n), w = __reduce_add_sync(0xffffffffu, won), l = __reduce_add_sync(0xffffffffu, lost);
        if (lane < 4 && f) {
            const unsigned v = lane == 0 ? w : lane == 1 ? l : lane == 2 ? f - w - l : f;
            if (v) atomicAdd(ra.stats + lane, (unsigned long long)v);
        }
    }
}
template <class C, int NT, bool SHAPED> int launch_rollout_env(const RolloutArgs& ra, const StepParams& sp, cudaStream_t st) {
    const int NN = sp.geo.N * sp.geo.N;
    const size_t sm = (size_t)RTPB * NN * sizeof(int16_t);
    k_rollout_env<C, NT, SHAPED><<<(unsigned)((ra.n + RTPB - 1) / RTPB), RTPB, sm, st>>>(ra, sp);
    LAUNCHED("k_rollout_env");
    return EWN_OK;
}
inline size_t rollout_smem(int NN, int D) { return policy_smem(NN + D) + (((size_t)32 * NN * 2 + 15) & ~(size_t)15) + 32 * sizeof(int); }
template <class C, int NT, bool SHAPED> int launch_rollout(const RolloutArgs& ra, const StepParams& sp, const PolicyNet& pn, int D, cudaStream_t st) {
    const int NN = sp.geo.N * sp.geo.N;
    const size_t sm = rollout_smem(NN, D);
    {   // more than the 48 KB default: opt in once per device and instantiation
        static std::mutex mu;
        static bool done[64] = {false};
        int dev = 0;
        CU(cudaGetDevice(&dev));
        std::lock_guard<std::mutex> lock(mu);
        if (dev >= 0 && dev < 64 && !done[dev]) {
            CU(cudaFuncSetAttribute(k_rollout<C, NT, SHAPED>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rollout_smem(EWN_MAX_BOARD * EWN_MAX_BOARD, 7)));
            done[dev] = true;
        }
    }
    k_rollout<C, NT, SHAPED><<<(unsigned)((ra.n + PE - 1) / PE), PT, sm, st>>>(ra, sp, pn, D);
    LAUNCHED("k_rollout");
    return EWN_OK;
}

// RolloutBuffer.compute_returns_and_advantage (GAE; A2C: lambda = 1) for rewards / values / dones laid out [T][n]; dones[t] = the
// episode ended at step t (= episode_starts[t + 1]).  One thread per env walks t downwards; fp32, rounded operation by operation
// in torch's order (no contraction), so the result equals the tensor loop of OnPolicyAlgorithm bit for bit.
__global__ void __launch_bounds__(TPB) k_gae(const float* __restrict__ rewards, const float* __restrict__ values, const uint8_t* __restrict__ dones,
                                             const float* __restrict__ last_values, int T, int64_t n, float gamma, float gamma_lambda,
                                             float* __restrict__ adv, float* __restrict__ returns) {
    const int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    float gae = 0.f, next_v = last_values[i];
    constexpr int CHK = 8;                                         // the loads of eight steps are issued together: the walk is a chain of
    for (int t1 = T; t1 > 0; t1 -= CHK) {                          // dependent adds, not of dependent DRAM round trips
        float r[CHK], v[CHK]; uint8_t d[CHK];
#pragma unroll
        for (int q = 0; q < CHK; q++) {
            const int t = t1 - 1 - q;
            const size_t k = (size_t)(t < 0 ? 0 : t) * n + i;
            r[q] = rewards[k]; v[q] = values[k]; d[q] = dones[k];
        }
#pragma unroll
        for (int q = 0; q < CHK; q++) {
            const int t = t1 - 1 - q;
            if (t < 0) break;
            const size_t k = (size_t)t * n + i;
            const float nt = __fsub_rn(1.0f, d[q] ? 1.0f : 0.0f);
            const float delta = __fsub_rn(__fadd_rn(r[q], __fmul_rn(__fmul_rn(gamma, next_v), nt)), v[q]);
            gae = __fadd_rn(delta, __fmul_rn(__fmul_rn(gamma_lambda, nt), gae));
            adv[k] = gae;
            if (returns) returns[k] = __fadd_rn(gae, v[q]);
            next_v = v[q];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// library-owned scratch (one per device), grown on demand; used when the caller passes no workspace
// ------------------------------------------------------------------------------------------------
// One allocation per (device, stream): calls on different streams never share scratch, so ewn_expectiminimax / ewn_mc_root_wins
// without a caller workspace are safe from several host threads as long as each thread uses its own stream (calls on ONE stream
// are ordered by the stream).  The table itself is guarded by a mutex; growing an entry waits for that stream only.
struct Scratch { int dev = -1; cudaStream_t st = nullptr; int kind = 0; void* p = nullptr; size_t bytes = 0; };
constexpr int MAX_SCRATCH = 256;
Scratch g_scratch[MAX_SCRATCH];
int g_scratch_n = 0;
std::mutex g_scratch_mutex;

// kind 1 = the Monte-Carlo workspace: its 256-byte header {pair counter, ticket} is zero whenever no decision is in flight (cleared
// when allocated, put back by the last warp of k_playouts), so ewn_mc_predict / ewn_mc_root_wins launch no memset for it
int scratch_get(size_t bytes, void** out, cudaStream_t st, int kind = 0) {
    int dev = 0;
    CU(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    Scratch* s = nullptr;
    for (int i = 0; i < g_scratch_n; i++) if (g_scratch[i].dev == dev && g_scratch[i].st == st && g_scratch[i].kind == kind) { s = &g_scratch[i]; break; }
    if (!s) {
        if (g_scratch_n == MAX_SCRATCH) return EWN_ESPACE;            // pass a workspace instead
        s = &g_scratch[g_scratch_n++]; s->dev = dev; s->st = st; s->kind = kind;
    }
    if (s->bytes < bytes) {
        if (s->p) { CU(cudaStreamSynchronize(st)); CU(cudaFree(s->p)); s->p = nullptr; s->bytes = 0; }
        CU(cudaMalloc(&s->p, bytes));
        s->bytes = bytes;
        if (kind == 1) CU(cudaMemsetAsync(s->p, 0, 256, st));
    }
    *out = s->p;
    return EWN_OK;
}

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (!cached[dev]) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        cached[dev] = v;
    }
    return cached[dev];
}

template <class C> size_t obs_smem(const Geo& geo) { return (size_t)WARPS * 32 * geo.N * geo.N * sizeof(int16_t); }
// k_step: one staging row per thread + slack for the 16-byte phase of the compacted terminal rows
inline size_t step_smem(const Geo& geo) { return (size_t)STPB * geo.N * geo.N * sizeof(int16_t) + 32; }

// grid of k_playouts: 8 blocks of 256 threads per SM.  Only 6 are resident at once (36 registers), so the last third of
// the blocks runs as a short second wave at low occupancy; measured on B200 this is FASTER than exactly one resident wave
// (1.66 ms vs 1.82 ms for 14.5 M playouts): the kernel is bound by the alu pipe, and the second wave's warps run with less
// contention while the equal-share ranges stay shorter (profiles/r1_k_playouts_grid_experiment.txt).
template <class C> int playout_grid(int64_t total_games) {
    int64_t warps_needed = (total_games + 31) / 32;
    int64_t blocks = (warps_needed + WARPS - 1) / WARPS;
    int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

// k_playouts indexes games with 32 bits: a job of 2^31 games or more runs as several launches over slices of the playout
// index range (results depend only on the global playout index, so the slicing is invisible)
inline bool dynamic_chunks() {            // EWN_MC_STATIC=1 in the environment selects the round-robin hand-out (A/B measurements)
    static const bool on = [] { const char* e = getenv("EWN_MC_STATIC"); return !(e && e[0] == '1'); }();
    return on;
}

// kernel launch with the programmatic-dependent-launch attribute (pdl) or a plain one
template <class... KArgs, class... Args>
static cudaError_t launch_k(void (*kernel)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t st, bool pdl, Args... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

template <class C, bool DIRECT>
static int launch_playouts(const void* pairs, const uint32_t* root_ids, const int* n_pairs_dev, int64_t max_pairs, uint32_t first, uint32_t count,
                           uint32_t* wins, unsigned long long* plies, const Geo& geo, uint64_t seed, cudaStream_t st, int* reset_hdr = nullptr,
                           bool pdl = false) {
    const RoundKeys rk = make_round_keys(seed);
    const int64_t lim = (((int64_t)1 << 31) - 256) / (max_pairs > 0 ? max_pairs : 1);
    if (lim < 1) return EWN_ENOSUP;                       // more than 2^31 pairs in one call
    for (uint64_t done = 0; done < count;) {
        const uint32_t c = (uint32_t)((count - done) < (uint64_t)lim ? (count - done) : (uint64_t)lim);
        const int grid = playout_grid<C>(max_pairs * (int64_t)c);
        const bool last = done + c >= count;
        // dynamic chunk hand-out needs the header's counter clean at launch: single-slice jobs on a workspace whose header resets itself
        if (reset_hdr && done == 0 && last && !DIRECT && dynamic_chunks())
            CU(launch_k(k_playouts<C, DIRECT, true>, (unsigned)grid, TPB, 0, st, pdl, pairs, root_ids, n_pairs_dev, (u32)max_pairs, first, c,
                        make_div31(c), wins, plies, geo, rk, reset_hdr));
        else
            CU(launch_k(k_playouts<C, DIRECT, false>, (unsigned)grid, TPB, 0, st, pdl && done == 0, pairs, root_ids, n_pairs_dev, (u32)max_pairs, first + (uint32_t)done, c,
                        make_div31(c), wins, plies, geo, rk, last ? reset_hdr : (int*)nullptr));
        LAUNCHED("k_playouts");
        done += c;
    }
    return EWN_OK;
}

// run `...` with C = the codec of (N, L): P5 / P8 for the six-cube configurations, PG for cube_layer >= 4
#define FOR_CODEC(N, L, ...)                                             \
    do {                                                                 \
        switch (codec_of(N, L)) {                                        \
            case 0: { typedef P5 C; __VA_ARGS__; } break;                \
            case 1: { typedef P8 C; __VA_ARGS__; } break;                \
            default: { typedef PG C; __VA_ARGS__; } break;               \
        }                                                                \
    } while (0)

// pending records -> values: zero the bit planes, play, reduce.  `bits` holds n_slots * 12 * sim_words(nsim) words.
template <class C> static int sim_fill(double* recs, const u64* states, int64_t n_slots, uint32_t unit0, int nsim, const Geo& geo, uint64_t seed,
                                       uint32_t* bits, cudaStream_t st) {
    const int64_t threads = n_slots * 12LL * nsim;
    if (threads >= ((int64_t)1 << 31) * TPB) return EWN_ENOSUP;
    CU(cudaMemsetAsync(bits, 0, (size_t)n_slots * 12 * sim_words(nsim) * sizeof(uint32_t), st));
    k_sim_playouts<C><<<(unsigned)((threads + TPB - 1) / TPB), TPB, 0, st>>>(recs, states, n_slots, unit0, nsim, geo, make_round_keys(seed), bits);
    LAUNCHED("k_sim_playouts");
    k_sim_values<<<blocks_for(n_slots), TPB, 0, st>>>(recs, n_slots, nsim, bits);
    LAUNCHED("k_sim_values");
    return EWN_OK;
}

// k_vstep launch; shared memory above the 48 KB default (large EWN_VTPB experiment builds, 9x9 boards) needs the opt-in attribute
template <class C, int NT, bool AR> cudaError_t launch_vstep(const cudaLaunchConfig_t& cfg, const StepKArgs& ka, int cap) {
    if (cfg.dynamicSmemBytes > 40 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_vstep<C, NT, AR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg.dynamicSmemBytes);
        if (e != cudaSuccess) return e;
    }
    return cudaLaunchKernelEx(&cfg, k_vstep<C, NT, AR>, ka, cap);
}

}  // namespace

// ================================================================================================
// C ABI
// ================================================================================================
extern "C" {
#if defined(EWN_EXP_TIMELINE)
__attribute__((visibility("default"))) int ewn_debug_timeline(unsigned long long* host_out) {
    return (int)cudaMemcpyFromSymbol(host_out, g_tl, sizeof(g_tl));
}
#endif

int ewn_version(void) { return 100; }

const char* ewn_strerror(int code) {
    switch (code) {
        case EWN_OK: return "ok";
        case EWN_EARG: return "bad argument";
        case EWN_ECUDA: return "CUDA error";
        case EWN_ENOSUP: return "not supported";
        case EWN_ESPACE: return "workspace too small";
        default: return "unknown error";
    }
}
const char* ewn_last_cuda_error(void) { return g_err; }

int ewn_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount");
    return n;
}
int ewn_device_pci_bus_id(int device, char* out, int len) {
    if (!out || len < 13) return EWN_EARG;
    CU(cudaDeviceGetPCIBusId(out, len, device));
    return EWN_OK;
}
int ewn_state_words(int N) { return (N >= 3 && N <= EWN_MAX_BOARD) ? (use_p5(N) ? 1 : 2) : EWN_EARG; }
int ewn_state_words_cfg(int N, int L) { return config_ok(N, L) ? make_geo(N, L).sw : EWN_EARG; }
int64_t ewn_launch_count(void) { return g_launches.load(); }

int ewn_reset(uint64_t* state, int64_t n, int N, int L, uint64_t seed, uint64_t step, uint64_t env_id0, void* stream) {
    if (!state || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    FOR_CODEC(N, L, k_reset<C><<<blocks_for(n), TPB, 0, st>>>(state, nullptr, n, geo, seed, step, env_id0));
    LAUNCHED("k_reset");
    return EWN_OK;
}

int ewn_reset_dice(uint64_t* state, const int32_t* dice, int64_t n, int N, int L, void* stream) {
    if (!state || !dice || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    FOR_CODEC(N, L, k_reset<C><<<blocks_for(n), TPB, 0, st>>>(state, dice, n, geo, 0, 0, 0));
    LAUNCHED("k_reset");
    return EWN_OK;
}

int ewn_encode(const int16_t* board, const int32_t* dice, const int32_t* player, uint64_t* state, int64_t n, int N, int L, void* stream) {
    if (!board || !state || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    FOR_CODEC(N, L, k_encode<C><<<blocks_for(n), TPB, 0, st>>>(board, dice, player, state, n, geo));
    LAUNCHED("k_encode");
    return EWN_OK;
}

int ewn_decode(const uint64_t* state, int16_t* board, int64_t* dice, int32_t* player, int64_t n, int N, int L, void* stream) {
    if (!state || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    FOR_CODEC(N, L, k_decode<C><<<blocks_for(n), TPB, obs_smem<C>(geo), st>>>(state, board, dice, player, n, geo));
    LAUNCHED("k_decode");
    return EWN_OK;
}

int ewn_legal(const uint64_t* state, int player, uint8_t* mask, int8_t* list, int32_t* count, int64_t n, int N, int L, void* stream) {
    if (!state || n < 0 || !config_ok(N, L) || player < 0 || player > 2) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    FOR_CODEC(N, L, k_legal<C><<<blocks_for(n), TPB, 0, st>>>(state, player, mask, list, count, n, geo));
    LAUNCHED("k_legal");
    return EWN_OK;
}

int ewn_check_win(const uint64_t* state, uint8_t* win, int64_t n, int N, int L, void* stream) {
    if (!state || !win || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    FOR_CODEC(N, L, k_check_win<C><<<blocks_for(n), TPB, 0, st>>>(state, win, n, geo));
    LAUNCHED("k_check_win");
    return EWN_OK;
}

int ewn_ply(uint64_t* state, int player, const int32_t* action, uint8_t* valid, uint8_t* win, int32_t* moved_cube, int64_t n, int N, int L,
            void* stream) {
    if (!state || !action || n < 0 || !config_ok(N, L) || player < 0 || player > 2) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    FOR_CODEC(N, L, k_ply<C><<<blocks_for(n), TPB, 0, st>>>(state, player, action, valid, win, moved_cube, n, geo));
    LAUNCHED("k_ply");
    return EWN_OK;
}

int ewn_set_dice(uint64_t* state, const int32_t* dice, const int32_t* player, int64_t n, int N, int L, void* stream) {
    if (!state || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    FOR_CODEC(N, L, k_set_dice<C><<<blocks_for(n), TPB, 0, st>>>(state, dice, player, n, geo));
    LAUNCHED("k_set_dice");
    return EWN_OK;
}

int ewn_decode_onehot(const uint64_t* state, float* board, float* dice, int64_t n, int N, int L, void* stream) {
    if (!state || !board || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    FOR_CODEC(N, L, k_decode_onehot<C><<<blocks_for(n), TPB, 0, st>>>(state, board, dice, n, geo));
    LAUNCHED("k_decode_onehot");
    return EWN_OK;
}

int ewn_numpy_random_games(const uint32_t* seeds, int64_t n, int N, int L, int8_t* result, int32_t* agent_steps, uint64_t* final_state,
                           uint64_t* trace, void* stream) {
    if (!seeds || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = (unsigned)((n + 127) / 128);
    FOR_CODEC(N, L, k_numpy_random_games<C><<<grid, 128, 0, st>>>(seeds, n, geo, result, agent_steps, final_state, trace));
    LAUNCHED("k_numpy_random_games");
    return EWN_OK;
}

int ewn_numpy_stream(const uint32_t* seeds, int64_t n, int K, uint32_t* raw, void* stream) {
    if (!seeds || !raw || n < 0 || K < 1) return EWN_EARG;
    if (n == 0) return EWN_OK;
    k_numpy_stream<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(seeds, n, K, raw);
    LAUNCHED("k_numpy_stream");
    return EWN_OK;
}

int ewn_numpy_draw(uint64_t* state, const uint32_t* raw, int32_t* cursor, int K, int64_t n, int N, int L, int roll_dice, int pick_player,
                   int32_t* action, void* stream) {
    if (!state || !raw || !cursor || n < 0 || K < 1 || !config_ok(N, L) || pick_player < 0 || pick_player > 2 || (pick_player && !action)) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    FOR_CODEC(N, L, k_numpy_draw<C><<<blocks_for(n), TPB, 0, st>>>(state, raw, cursor, K, n, geo, roll_dice, pick_player, action));
    LAUNCHED("k_numpy_draw");
    return EWN_OK;
}

int ewn_flip(const uint64_t* state, uint64_t* out, int64_t n, int N, int L, void* stream) {
    if (!state || !out || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    FOR_CODEC(N, L, k_flip<C><<<blocks_for(n), TPB, 0, st>>>(state, out, n, geo));
    LAUNCHED("k_flip");
    return EWN_OK;
}

int ewn_random_action(const uint64_t* state, int64_t* action, int64_t n, int N, int L, uint64_t seed, uint64_t step, uint64_t env_id0,
                      void* stream) {
    if (!state || !action || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    FOR_CODEC(N, L, k_random_action<C><<<blocks_for(n), TPB, 0, st>>>(state, action, n, geo, seed, step, env_id0));
    LAUNCHED("k_random_action");
    return EWN_OK;
}

int ewn_step(const ewn_step_args* a, void* stream) {
    if (!a || !a->state || a->n < 0 || !config_ok(a->board_size, a->cube_layer)) return EWN_EARG;
    if (a->n == 0) return EWN_OK;
    if (!a->agent_random && !a->action_i64 && !a->action_i32 && !a->action_u8) return EWN_EARG;
    const bool replay = a->inject_dice_opp || a->inject_opp_action || a->inject_dice_next;
    if (replay && !(a->inject_dice_opp && a->inject_opp_action && a->inject_dice_next)) return EWN_EARG;
    if (a->shaped && !(a->tolerance && a->prev_score)) return EWN_EARG;
    StepKArgs ka;
    ka.a = *a;
    memset(&ka.p, 0, sizeof(ka.p));
    ka.p.geo = make_geo(a->board_size, a->cube_layer);
    ka.p.reward = a->reward; ka.p.illegal_move_reward = a->illegal_move_reward;
    ka.p.seed = a->seed; ka.p.step = a->step; ka.p.env_id0 = a->env_id0; ka.p.rk = make_round_keys(a->seed);
    ka.p.agent_random = a->agent_random != 0; ka.p.autoreset = a->autoreset != 0;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned grid = (unsigned)((a->n + STPB - 1) / STPB);
    const bool need_smem = a->obs_board || a->term_board;
    if (a->term_count && !a->term_index) return EWN_EARG;
    const bool vec_out = a->reward_f32 && a->terminated && a->truncated && a->obs_board && a->obs_dice && !a->reward_f64 && !a->code &&
                         !a->action_out && !replay && !a->shaped;
    const bool vec = vec_out && a->action_i64 && !a->agent_random;
    const bool vec_ar = vec_out && a->agent_random;               // the same surface with the in-kernel random agent
#define LAUNCH_STEP(C, NT)                                                                           \
    do {                                                                                             \
        size_t sm = need_smem ? step_smem(ka.p.geo) : 0;                                             \
        const bool pdl_ = !(a->flags & EWN_STEP_NO_PDL);                                             \
        cudaError_t le_;                                                                             \
        if (vec) le_ = launch_k(k_step<C, NT, false, false, true>, grid, STPB, sm, st, pdl_, ka);     \
        else if (replay) { if (a->shaped) le_ = launch_k(k_step<C, NT, true, true, false>, grid, STPB, sm, st, pdl_, ka); else le_ = launch_k(k_step<C, NT, true, false, false>, grid, STPB, sm, st, pdl_, ka); } \
        else        { if (a->shaped) le_ = launch_k(k_step<C, NT, false, true, false>, grid, STPB, sm, st, pdl_, ka); else le_ = launch_k(k_step<C, NT, false, false, false>, grid, STPB, sm, st, pdl_, ka); } \
        if (le_ != cudaSuccess) return cuda_fail(le_, "k_step");                                     \
    } while (0)
    const int N = a->board_size;
    // the VecEnv surface with compacted (or no) terminal observations runs on the persistent kernel
    const int codec = codec_of(N, a->cube_layer);
    if (codec != 2 && (vec || vec_ar) && !(a->flags & EWN_STEP_BLOCK_KERNEL) && (a->term_count || (!a->term_board && !a->term_dice && !a->term_code)) && a->n < ((int64_t)1 << 30)) {
        const int64_t n_units = (a->n + VTPB - 1) / VTPB;
        int64_t blocks = (int64_t)sm_count() * VMINB;
        if (blocks > n_units) blocks = n_units;
        if (blocks * VITER_MAX < n_units) blocks = (n_units + VITER_MAX - 1) / VITER_MAX;
        const int iters = (int)((n_units + blocks - 1) / blocks);
        const int cap = a->term_count ? iters * VTPB : 0;
        const int words = use_p5(N) ? 1 : 2;
        const size_t sm = (((size_t)VTPB * N * N * 2 + 32 + 15) & ~(size_t)15) + (size_t)cap * (8 * words + 4);
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3((unsigned)blocks); cfg.blockDim = dim3(VTPB); cfg.dynamicSmemBytes = sm; cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr; cfg.numAttrs = (a->flags & EWN_STEP_NO_PDL) ? 0 : 1;
        cudaError_t le;
        if (vec_ar) {
            if (N == 5) le = launch_vstep<P5, 5, true>(cfg, ka, cap);
            else if (use_p5(N)) le = launch_vstep<P5, 0, true>(cfg, ka, cap);
            else if (N == 7) le = launch_vstep<P8, 7, true>(cfg, ka, cap);
            else le = launch_vstep<P8, 0, true>(cfg, ka, cap);
        } else if (N == 5) le = launch_vstep<P5, 5, false>(cfg, ka, cap);
        else if (use_p5(N)) le = launch_vstep<P5, 0, false>(cfg, ka, cap);
        else if (N == 7) le = launch_vstep<P8, 7, false>(cfg, ka, cap);
        else le = launch_vstep<P8, 0, false>(cfg, ka, cap);
        if (le != cudaSuccess) return cuda_fail(le, "k_vstep");
        LAUNCHED("k_vstep");
        return EWN_OK;
    }
    if (codec == 2) LAUNCH_STEP(PG, 0);          // cube_layer >= 4: generic codec on the block kernel
    else if (N == 5) LAUNCH_STEP(P5, 5); else if (use_p5(N)) LAUNCH_STEP(P5, 0); else if (N == 7) LAUNCH_STEP(P8, 7); else LAUNCH_STEP(P8, 0);
#undef LAUNCH_STEP
    LAUNCHED("k_step");
    return EWN_OK;
}

// the two dice a step draws (STEP block words 0 and 2), for callers that run the opponent's policy themselves between the two plies
__global__ void __launch_bounds__(TPB) k_step_dice(int32_t* dice_opp, int32_t* dice_next, int64_t n, int cube_num, u64 seed, u64 step, u64 env_id0,
                                                   const unsigned long long* step_dev) {
    int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x;
    if (i >= n) return;
    u32 blk[4]; env_block(seed, env_id0 + (u64)i, step + (step_dev ? *step_dev : 0ull), TAG_STEP, blk);
    if (dice_opp) dice_opp[i] = 1 + (int)bounded(blk[0], (u32)cube_num);
    if (dice_next) dice_next[i] = 1 + (int)bounded(blk[2], (u32)cube_num);
}

__global__ void k_counter_add(unsigned long long* c, unsigned long long k) { if (threadIdx.x == 0 && blockIdx.x == 0) *c += k; }

int ewn_step_dice(int32_t* dice_opp, int32_t* dice_next, int64_t n, int N, int L, uint64_t seed, uint64_t step, uint64_t env_id0,
                  const uint64_t* step_dev, void* stream) {
    if ((!dice_opp && !dice_next) || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (n == 0) return EWN_OK;
    k_step_dice<<<blocks_for(n), TPB, 0, (cudaStream_t)stream>>>(dice_opp, dice_next, n, L * (L + 1) / 2, seed, step, env_id0,
                                                                 reinterpret_cast<const unsigned long long*>(step_dev));
    LAUNCHED("k_step_dice");
    return EWN_OK;
}

int ewn_counter_add(uint64_t* counter, uint64_t k, void* stream) {
    if (!counter) return EWN_EARG;
    k_counter_add<<<1, 32, 0, (cudaStream_t)stream>>>(reinterpret_cast<unsigned long long*>(counter), (unsigned long long)k);
    LAUNCHED("k_counter_add");
    return EWN_OK;
}

int ewn_policy_forward(const ewn_policy_net* net, const int16_t* board, const int64_t* dice, int64_t n, int N, int L, uint64_t seed, uint64_t step,
                       const uint64_t* step_dev, uint64_t env_id0, int deterministic, int64_t* actions, float* values, float* logp, float* logits,
                       void* stream) {
    if (!net || !board || !dice || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (net->hidden != PH) return EWN_ENOSUP;
    if (!net->pi_w1 || !net->pi_b1 || !net->pi_w2 || !net->pi_b2 || !net->vf_w1 || !net->vf_b1 || !net->vf_w2 || !net->vf_b2 || !net->action_w ||
        !net->action_b || !net->value_w || !net->value_b) return EWN_EARG;
    if (logp && !actions) return EWN_EARG;
    if (n == 0) return EWN_OK;
    PolicyNet pn;
    pn.w1[0] = net->pi_w1; pn.b1[0] = net->pi_b1; pn.w2[0] = net->pi_w2; pn.b2[0] = net->pi_b2;
    pn.w1[1] = net->vf_w1; pn.b1[1] = net->vf_b1; pn.w2[1] = net->vf_w2; pn.b2[1] = net->vf_b2;
    pn.wa = net->action_w; pn.ba = net->action_b; pn.wv = net->value_w; pn.bv = net->value_b;
    const int NN = N * N, D = L * (L + 1) / 2 + 1;
    const size_t sm = policy_smem(NN + D);
    unsigned grid = (unsigned)std::min<int64_t>((n + PE - 1) / PE, (int64_t)2 * sm_count());       // two blocks fit an SM (shared memory): one wave
    {   // more than the 48 KB default: opt in once per device
        static std::mutex mu;
        static bool done[64] = {false};
        int dev = 0;
        CU(cudaGetDevice(&dev));
        std::lock_guard<std::mutex> lock(mu);
        if (dev >= 0 && dev < 64 && !done[dev]) {
            CU(cudaFuncSetAttribute(k_policy_forward, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)policy_smem(EWN_MAX_BOARD * EWN_MAX_BOARD + 32)));
            done[dev] = true;
        }
    }
    k_policy_forward<<<grid, PT, sm, (cudaStream_t)stream>>>(board, dice, n, NN, D, pn, seed, step, reinterpret_cast<const unsigned long long*>(step_dev),
                                                                 env_id0, deterministic, actions, values, logp, logits);
    LAUNCHED("k_policy_forward");
    return EWN_OK;
}

int ewn_rollout(const ewn_rollout_args* a, void* stream) {
    if (!a || !a->state || a->n < 0 || a->T < 1 || !config_ok(a->board_size, a->cube_layer)) return EWN_EARG;
    if (!a->board || !a->dice || !a->rewards || !a->dones) return EWN_EARG;
    if (a->agent_random ? a->net != nullptr : !a->actions) return EWN_EARG;       // in-kernel random agent: no network, actions optional (written)
    if (a->net && (!a->values || !a->logp || !a->last_values || !a->advantages || !a->returns)) return EWN_EARG;
    if (a->shaped && !(a->tolerance && a->prev_score)) return EWN_EARG;
    const ewn_policy_net* net = a->net;
    if ((net && net->hidden != PH) || codec_of(a->board_size, a->cube_layer) == 2) return EWN_ENOSUP;
    if (net && (!net->pi_w1 || !net->pi_b1 || !net->pi_w2 || !net->pi_b2 || !net->vf_w1 || !net->vf_b1 || !net->vf_w2 || !net->vf_b2 || !net->action_w ||
                !net->action_b || !net->value_w || !net->value_b)) return EWN_EARG;
    if (a->n == 0) return EWN_OK;
    PolicyNet pn;
    memset(&pn, 0, sizeof(pn));
    if (net) {
        pn.w1[0] = net->pi_w1; pn.b1[0] = net->pi_b1; pn.w2[0] = net->pi_w2; pn.b2[0] = net->pi_b2;
        pn.w1[1] = net->vf_w1; pn.b1[1] = net->vf_b1; pn.w2[1] = net->vf_w2; pn.b2[1] = net->vf_b2;
        pn.wa = net->action_w; pn.ba = net->action_b; pn.wv = net->value_w; pn.bv = net->value_b;
    }
    StepParams sp;
    memset(&sp, 0, sizeof(sp));
    sp.geo = make_geo(a->board_size, a->cube_layer);
    sp.reward = a->reward; sp.illegal_move_reward = a->illegal_move_reward;
    sp.seed = a->seed; sp.step = a->step; sp.env_id0 = a->env_id0; sp.rk = make_round_keys(a->seed);
    sp.agent_random = a->agent_random != 0; sp.autoreset = 1; sp.shaped = a->shaped != 0; sp.replay = 0;
    RolloutArgs ra;
    ra.state = a->state; ra.n = a->n; ra.T = a->T;
    ra.board = a->board; ra.dice = a->dice; ra.actions = a->actions; ra.rewards = a->rewards; ra.dones = a->dones; ra.values = a->values;
    ra.logp = a->logp; ra.last_values = a->last_values; ra.adv = a->advantages; ra.ret = a->returns;
    ra.tolerance = a->tolerance; ra.prev_score = a->prev_score; ra.stats = reinterpret_cast<unsigned long long*>(a->stats);
    ra.step_dev = reinterpret_cast<const unsigned long long*>(a->step_dev);
    ra.policy_seed = a->policy_seed; ra.gamma = (float)a->gamma; ra.gamma_lambda = (float)(a->gamma * a->gae_lambda);
    const int N = a->board_size, D = a->cube_layer * (a->cube_layer + 1) / 2 + 1;
    cudaStream_t st = (cudaStream_t)stream;
#define ROLL(C, NT) (!net ? (a->shaped ? launch_rollout_env<C, NT, true>(ra, sp, st) : launch_rollout_env<C, NT, false>(ra, sp, st)) \
                          : (a->shaped ? launch_rollout<C, NT, true>(ra, sp, pn, D, st) : launch_rollout<C, NT, false>(ra, sp, pn, D, st)))
    if (N == 5) return ROLL(P5, 5);
    if (use_p5(N)) return ROLL(P5, 0);
    if (N == 7) return ROLL(P8, 7);
    return ROLL(P8, 0);
#undef ROLL
}

int ewn_gae(const float* rewards, const float* values, const uint8_t* dones, const float* last_values, int T, int64_t n, double gamma,
            double gae_lambda, float* advantages, float* returns, void* stream) {
    if (!rewards || !values || !dones || !last_values || !advantages || T < 0 || n < 0) return EWN_EARG;
    if (n == 0 || T == 0) return EWN_OK;
    k_gae<<<blocks_for(n), TPB, 0, (cudaStream_t)stream>>>(rewards, values, dones, last_values, T, n, (float)gamma, (float)(gamma * gae_lambda),
                                                          advantages, returns);
    LAUNCHED("k_gae");
    return EWN_OK;
}

int ewn_evaluate(const uint64_t* state, double* value, int64_t n, int N, int L, int heuristic, void* stream) {
    if (!state || !value || n < 0 || !config_ok(N, L) || heuristic < 0 || heuristic > 3) return EWN_EARG;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    FOR_CODEC(N, L, k_evaluate<C><<<blocks_for(n), TPB, 0, st>>>(state, value, n, geo, heuristic));
    LAUNCHED("k_evaluate");
    return EWN_OK;
}

size_t ewn_expectiminimax_workspace(int64_t n, int depth) {
    if (n <= 0 || depth < 1 || depth > EWN_MAX_DEPTH) return 0;
    return (size_t)n * (size_t)emm_level_offset(emm_levels(depth) + 1) * sizeof(double);
}

// 'hybrid' leaf-value table of one board size, built once per device on first use (k_emm_table) and kept for the process
// (the fused kernel's padded value / quotient tables follow the value table in the same allocation: *fused_out)
static std::mutex g_table_mutex;
static const double* emm_value_table(int N, cudaStream_t st, const double** fused_out = nullptr) {
    static double* tab[64][EWN_MAX_BOARD + 1] = {{nullptr}};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    const long cnt = emm_table_size(N);
    std::lock_guard<std::mutex> lock(g_table_mutex);
    if (!tab[dev][N]) {
        double* p = nullptr;
        const long fused = 4L * (emm_pad(N) + emm_tab_entries(N));
        if (cudaMalloc(&p, (size_t)(cnt + fused) * sizeof(double)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        k_emm_table<<<(unsigned)((cnt + TPB - 1) / TPB), TPB, 0, st>>>(p, N);
        k_emm_fused_table<<<(unsigned)((fused / 2 + TPB - 1) / TPB), TPB, 0, st>>>(p, p + cnt, N);
        g_launches.fetch_add(2, std::memory_order_relaxed);
        if (cudaStreamSynchronize(st) != cudaSuccess) { cudaGetLastError(); cudaFree(p); return nullptr; }
        tab[dev][N] = p;
    }
    if (fused_out) *fused_out = tab[dev][N] + cnt;
    return tab[dev][N];
}

int ewn_expectiminimax(const uint64_t* state, int64_t n, int N, int L, int depth, int heuristic, int pruned, double* value, int8_t* action,
                       void* workspace, size_t workspace_bytes, void* stream) {
    if (!state || !value || n < 0 || !config_ok(N, L) || heuristic < 0 || heuristic > 3) return EWN_EARG;
    if (depth < 1 || depth > EWN_MAX_DEPTH) return EWN_ENOSUP;
    if (L != 3) return EWN_ENOSUP;      // the reference's search hard-codes a six-face die (minimax.py:68)
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    const int m = emm_levels(depth);
    const double* tabs = nullptr;
    const double* vt = heuristic == H_HYBRID ? emm_value_table(N, st, &tabs) : nullptr;      // nullptr: leaves are computed in place
    if (m == 2 && !(pruned & EWN_EMM_REPLAY_PATH) && (heuristic != H_HYBRID || vt)) {  // depth 3 / 4: fused, on-chip (k_emm_fused)
        const unsigned grid = (unsigned)((n + FROOTS - 1) / FROOTS);
        const size_t sm = heuristic == H_HYBRID ? (size_t)2 * (emm_pad(N) + emm_tab_entries(N)) * sizeof(double) : 0;
        const int pr = pruned & 1;
#define LAUNCH_FUSED(C)                                                                                                     \
    if (sm > 40 * 1024) CU(cudaFuncSetAttribute(k_emm_fused<C, H_HYBRID>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); /* 9x9: static + dynamic > 48 KB */ \
    switch (heuristic) {                                                                                                    \
        case H_HYBRID: k_emm_fused<C, H_HYBRID><<<grid, FTPB, sm, st>>>(state, n, geo, depth, pr, tabs, value, action); break; \
        case H_MIN_DIST: k_emm_fused<C, H_MIN_DIST><<<grid, FTPB, sm, st>>>(state, n, geo, depth, pr, tabs, value, action); break; \
        case H_TWO_MIN_DIST: k_emm_fused<C, H_TWO_MIN_DIST><<<grid, FTPB, sm, st>>>(state, n, geo, depth, pr, tabs, value, action); break; \
        default: k_emm_fused<C, H_ATTK><<<grid, FTPB, sm, st>>>(state, n, geo, depth, pr, tabs, value, action); break;          \
    }
        if (use_p5(N)) { LAUNCH_FUSED(P5) } else { LAUNCH_FUSED(P8) }
#undef LAUNCH_FUSED
        LAUNCHED("k_emm_fused");
        return EWN_OK;
    }
    pruned &= 1;
    // 'hybrid' with the value table: 16-bit records (table index or mark) instead of doubles -- a quarter of the record traffic
    const bool compact = vt != nullptr && emm_table_size(N) < (long)EMM16_LOSS;
    const size_t per_root = (size_t)emm_level_offset(m + 1) * (compact ? sizeof(uint16_t) : sizeof(double));
    if (!workspace || workspace_bytes == 0) {
        size_t want = per_root * (size_t)n;
        const size_t cap = (size_t)1 << 30;
        if (want > cap) want = cap > per_root ? cap : per_root;
        int rc = scratch_get(want, &workspace, st);
        if (rc) return rc;
        workspace_bytes = want;
    }
    int64_t chunk = (int64_t)(workspace_bytes / per_root);
    if (chunk < 1) return EWN_ESPACE;
    const int64_t slots = emm_level_size(m);
    if (chunk > ((int64_t)1 << 31) / slots - 1) chunk = ((int64_t)1 << 31) / slots - 1;   // 32-bit thread ids in k_emm_expand
    for (int64_t r0 = 0; r0 < n; r0 += chunk) {
        const int64_t nr = (n - r0 < chunk) ? (n - r0) : chunk;
        const uint64_t* s0 = state + r0 * (use_p5(N) ? 1 : 2);
        const unsigned grid1 = blocks_for(nr * (slots / 6));
        const unsigned grid2 = (unsigned)((nr + 127) / 128);
#define LAUNCH_EMM(C, R)                                                                                           \
    do {                                                                                                           \
        R* vals = (R*)workspace;                                                                                   \
        k_emm_expand<C, R><<<grid1, TPB, 0, st>>>(s0, vals, nr, geo, depth, heuristic, vt, nullptr);               \
        LAUNCHED("k_emm_expand");                                                                                  \
        switch (m) {                                                                                               \
            case 1: k_emm_replay<C, 1, R><<<grid2, 128, 0, st>>>(state, vals, vt, nr, r0, geo, heuristic, pruned, value, action); break; \
            case 2: k_emm_replay<C, 2, R><<<grid2, 128, 0, st>>>(state, vals, vt, nr, r0, geo, heuristic, pruned, value, action); break; \
            case 3: k_emm_replay<C, 3, R><<<grid2, 128, 0, st>>>(state, vals, vt, nr, r0, geo, heuristic, pruned, value, action); break; \
            default: k_emm_replay<C, 4, R><<<grid2, 128, 0, st>>>(state, vals, vt, nr, r0, geo, heuristic, pruned, value, action); break; \
        }                                                                                                          \
        LAUNCHED("k_emm_replay");                                                                                  \
    } while (0)
        if (use_p5(N)) { if (compact) LAUNCH_EMM(P5, uint16_t); else LAUNCH_EMM(P5, double); }
        else           { if (compact) LAUNCH_EMM(P8, uint16_t); else LAUNCH_EMM(P8, double); }
#undef LAUNCH_EMM
    }
    return EWN_OK;
}

int ewn_sim_winrate(const uint64_t* state, int64_t n, int N, int L, int num_simulations, uint64_t seed, uint32_t unit0, double* value, void* stream) {
    if (!state || !value || n < 0 || !config_ok(N, L) || num_simulations < 1 || num_simulations > SIM_MAX) return EWN_EARG;
    if (L != 3) return EWN_ENOSUP;      // six-face die hard-coded in simulate() (minimax_ewn.py:223)
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    void* ws = nullptr;
    int rc = scratch_get((size_t)n * 12 * sim_words(num_simulations) * sizeof(uint32_t), &ws, st);
    if (rc) return rc;
    if (use_p5(N)) { k_sim_mark<P5><<<blocks_for(n), TPB, 0, st>>>(state, value, n, geo); LAUNCHED("k_sim_mark"); return sim_fill<P5>(value, state, n, unit0, num_simulations, geo, seed, (uint32_t*)ws, st); }
    k_sim_mark<P8><<<blocks_for(n), TPB, 0, st>>>(state, value, n, geo); LAUNCHED("k_sim_mark");
    return sim_fill<P8>(value, state, n, unit0, num_simulations, geo, seed, (uint32_t*)ws, st);
}

int ewn_expectiminimax_sim(const uint64_t* state, int64_t n, int N, int L, int depth, int pruned, int num_simulations, uint64_t seed,
                           double* value, int8_t* action, void* stream) {
    if (!state || !value || n < 0 || !config_ok(N, L) || num_simulations < 1 || num_simulations > SIM_MAX) return EWN_EARG;
    if (depth < 1 || depth > 4) return EWN_ENOSUP;       // 216 leaf positions x up to 1,200 playouts per root at depth 3 / 4 is the practical limit
    if (L != 3) return EWN_ENOSUP;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    const int m = emm_levels(depth), words = use_p5(N) ? 1 : 2, W = sim_words(num_simulations);
    const int64_t per_root = emm_level_offset(m + 1), slots = emm_level_size(m);
    const size_t bytes_root = (size_t)per_root * (8 + 8 * words + 12 * W * 4);
    int64_t chunk = (int64_t)(((size_t)512 << 20) / bytes_root);
    if (chunk < 1) chunk = 1;
    if (chunk > n) chunk = n;
    void* ws = nullptr;
    int rc = scratch_get(bytes_root * (size_t)chunk, &ws, st);
    if (rc) return rc;
    double* vals = (double*)ws;
    uint64_t* sst = (uint64_t*)(vals + chunk * per_root);
    uint32_t* bits = (uint32_t*)(sst + chunk * per_root * words);
    for (int64_t r0 = 0; r0 < n; r0 += chunk) {
        const int64_t nr = (n - r0 < chunk) ? (n - r0) : chunk;
        if ((uint64_t)(r0 + nr) * (uint64_t)per_root * 6u >= ((uint64_t)1 << 32)) return EWN_ENOSUP;     // Philox unit ids are 32-bit
        const uint64_t* s0 = state + r0 * words;
        const unsigned grid1 = blocks_for(nr * (slots / 6)), grid2 = (unsigned)((nr + 127) / 128);
        const uint32_t unit0 = (uint32_t)(r0 * per_root);
#define SIM_SEARCH(C)                                                                                                              \
    do {                                                                                                                           \
        k_emm_expand<C, double><<<grid1, TPB, 0, st>>>(s0, vals, nr, geo, depth, H_SIM_WINRATE, nullptr, sst);                     \
        LAUNCHED("k_emm_expand");                                                                                                  \
        rc = sim_fill<C>(vals, sst, nr * per_root, unit0, num_simulations, geo, seed, bits, st);                                   \
        if (rc) return rc;                                                                                                         \
        if (m == 1) k_emm_replay<C, 1, double><<<grid2, 128, 0, st>>>(state, vals, nullptr, nr, r0, geo, H_SIM_WINRATE, pruned & 1, value, action); \
        else        k_emm_replay<C, 2, double><<<grid2, 128, 0, st>>>(state, vals, nullptr, nr, r0, geo, H_SIM_WINRATE, pruned & 1, value, action); \
        LAUNCHED("k_emm_replay");                                                                                                  \
    } while (0)
        if (use_p5(N)) SIM_SEARCH(P5); else SIM_SEARCH(P8);
#undef SIM_SEARCH
    }
    return EWN_OK;
}

size_t ewn_mc_workspace(int64_t n) {
    if (n <= 0) return 0;
    return 256 + (size_t)n * 6 * 32;      // counter + one record (16 B for board_size <= 5, 32 B above) per (root, child)
}

int ewn_playouts(const uint64_t* state, const uint32_t* root_ids, int64_t n, int N, int L, uint64_t seed, uint32_t first, uint32_t count,
                 uint32_t* wins, unsigned long long* plies, void* stream) {
    if (!state || !wins || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (L != 3) return EWN_ENOSUP;      // six-face die hard-coded in the reference's playouts (mcts.py:29)
    if (n == 0 || count == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    return use_p5(N) ? launch_playouts<P5, true>(state, root_ids, nullptr, n, first, count, wins, plies, geo, seed, st)
                     : launch_playouts<P8, true>(state, root_ids, nullptr, n, first, count, wins, plies, geo, seed, st);
}

// expand -> playouts (-> arg-max): the launches of one Monte-Carlo decision.  predict: wins are cleared by the expansion, the arg-max
// follows when `action` is given, and the three launches are chained with programmatic dependent launch.
static int mc_run(const uint64_t* state, const uint32_t* root_ids, int64_t n, int N, int L, uint64_t seed, uint32_t first, uint32_t count,
                  uint32_t* wins, int32_t* n_children, unsigned long long* plies, int8_t* action, void* workspace, size_t workspace_bytes,
                  cudaStream_t st, bool predict) {
    if (!state || !wins || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (L != 3) return EWN_ENOSUP;
    if (n == 0) return EWN_OK;
    if (n > ((int64_t)1 << 28)) return EWN_ENOSUP;        // slot * 8 + child is a 32-bit word
    const size_t need = ewn_mc_workspace(n);
    bool own = false;
    if (!workspace || workspace_bytes == 0) {
        int rc = scratch_get(need, &workspace, st, 1);
        if (rc) return rc;
        own = true;
    } else if (workspace_bytes < need) return EWN_ESPACE;
    Geo geo = make_geo(N, L);
    char* w = (char*)workspace;
    PairBuf pb;
    pb.n_pairs = (int*)w;
    pb.rec = w + 256;
    // the library's own workspace keeps its header clean between decisions; a caller's buffer is cleared here
#if defined(EWN_MC_NO_TICKET)      /* experiment build: memset per decision instead of the self-resetting counter */
    const bool hdr_clean = false;
#else
    const bool hdr_clean = own && count != 0;
#endif
    if (!hdr_clean) CU(cudaMemsetAsync(pb.n_pairs, 0, 2 * sizeof(int), st));
#if defined(EWN_MC_NO_PDL)         /* experiment build: plain stream order between the three kernels */
    const bool pdl = false;
#else
    const bool pdl = predict;
#endif
    if (use_p5(N)) CU(launch_k(k_mc_expand<P5>, blocks_for(n), TPB, 0, st, false, state, root_ids, n, geo, pb, n_children, count, wins, predict ? 1 : 0));
    else           CU(launch_k(k_mc_expand<P8>, blocks_for(n), TPB, 0, st, false, state, root_ids, n, geo, pb, n_children, count, wins, predict ? 1 : 0));
    LAUNCHED("k_mc_expand");
    if (count != 0) {
        int* hdr = hdr_clean ? pb.n_pairs : nullptr;
        int rc = use_p5(N) ? launch_playouts<P5, false>(pb.rec, nullptr, pb.n_pairs, n * 6, first, count, wins, plies, geo, seed, st, hdr, pdl)
                           : launch_playouts<P8, false>(pb.rec, nullptr, pb.n_pairs, n * 6, first, count, wins, plies, geo, seed, st, hdr, pdl);
        if (rc) return rc;
    }
    if (predict && action) {
        if (use_p5(N)) CU(launch_k(k_mc_best<P5>, blocks_for(n), TPB, 0, st, pdl, state, (const u32*)wins, action, n, geo));
        else           CU(launch_k(k_mc_best<P8>, blocks_for(n), TPB, 0, st, pdl, state, (const u32*)wins, action, n, geo));
        LAUNCHED("k_mc_best");
    }
    return EWN_OK;
}

int ewn_mc_root_wins(const uint64_t* state, const uint32_t* root_ids, int64_t n, int N, int L, uint64_t seed, uint32_t first, uint32_t count,
                     uint32_t* wins, int32_t* n_children, unsigned long long* plies, void* workspace, size_t workspace_bytes, void* stream) {
    return mc_run(state, root_ids, n, N, L, seed, first, count, wins, n_children, plies, nullptr, workspace, workspace_bytes, (cudaStream_t)stream, false);
}

int ewn_mc_predict(const uint64_t* state, const uint32_t* root_ids, int64_t n, int N, int L, uint64_t seed, uint32_t first, uint32_t count,
                   uint32_t* wins, int8_t* action, int32_t* n_children, unsigned long long* plies, void* workspace, size_t workspace_bytes,
                   void* stream) {
    return mc_run(state, root_ids, n, N, L, seed, first, count, wins, n_children, plies, action, workspace, workspace_bytes, (cudaStream_t)stream, true);
}

// ---- ewn_mc_predict without any allocation (no counterpart in the reference) ----------
// The Monte-Carlo workspace {256-byte header, one record per (root, child)} lives in a file-scope device array instead of the
// library's cached cudaMalloc: callers under CUDA graph capture, or in loops / harnesses that must not allocate, use this entry point.
// Every GPU that loads the library holds the array (3 MiB) until the process exits.  All calls share it: calls that overlap on the GPU
// (two streams, two host threads) race on it; calls on one stream are ordered by the stream.  Nothing in it needs to survive a call:
// the header is cleared by a memset node at the head of every call, the records are rewritten by the expansion kernel.
static __device__ __align__(256) unsigned char g_mc_static_ws[256 + (size_t)EWN_MC_STATIC_MAX_ROOTS * 6 * 32];

// device address of g_mc_static_ws on the CURRENT device (in host code the array's name is only a handle for the runtime, not a pointer
// a kernel could use): looked up once per device, with that device current
static int mc_static_ptr(void** out) {
    static void* cached[64] = {nullptr};
    static std::mutex mu;
    int dev = 0;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { snprintf(g_err, sizeof(g_err), "ewn_mc_predict_static: device ordinal %d not supported", dev); return EWN_ENOSUP; }
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[dev]) {
        void* p = nullptr;
        CU(cudaGetSymbolAddress(&p, g_mc_static_ws));
        if (!p) { snprintf(g_err, sizeof(g_err), "ewn_mc_predict_static: cudaGetSymbolAddress returned a null address"); return EWN_ECUDA; }
        cached[dev] = p;
    }
    *out = cached[dev];
    return EWN_OK;
}

int64_t ewn_mc_static_max_roots(void) { return EWN_MC_STATIC_MAX_ROOTS; }

int ewn_mc_predict_static(const uint64_t* state, const uint32_t* root_ids, int64_t n, int N, int L, uint64_t seed, uint32_t first, uint32_t count,
                          uint32_t* wins, int8_t* action, int32_t* n_children, unsigned long long* plies, void* stream) {
    if (n > EWN_MC_STATIC_MAX_ROOTS) {
        snprintf(g_err, sizeof(g_err), "ewn_mc_predict_static: %lld roots exceed the static workspace (EWN_MC_STATIC_MAX_ROOTS = %d); "
                 "split the batch or use ewn_mc_predict", (long long)n, EWN_MC_STATIC_MAX_ROOTS);
        return EWN_ESPACE;
    }
    if (!state || !wins || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (n == 0) return EWN_OK;
    void* ws = nullptr;
    int rc = mc_static_ptr(&ws);
    if (rc) return rc;
    // a caller-style workspace: mc_run clears its header (cudaMemsetAsync, 8 bytes) and launches expand -> playouts -> arg-max; no
    // cudaMalloc / cudaFree on this path
    return mc_run(state, root_ids, n, N, L, seed, first, count, wins, n_children, plies, action, ws, sizeof(g_mc_static_ws), (cudaStream_t)stream, true);
}

size_t ewn_uct_workspace(int64_t n) { return n <= 0 ? 0 : (size_t)n * (16 + 4 + 4 + 4) + 256; }

int ewn_uct_search(const uint64_t* state, const uint32_t* root_ids, int64_t n, int N, int L, uint64_t seed, int rounds, uint32_t batch, double c,
                   uint32_t* wins, uint32_t* visits, int8_t* action, void* workspace, size_t workspace_bytes, void* stream) {
    if (!state || !wins || !visits || n < 0 || !config_ok(N, L) || rounds < 1 || batch < 1) return EWN_EARG;
    if (L != 3) return EWN_ENOSUP;      // six-face die in the playouts (mcts.py:29)
    if (n == 0) return EWN_OK;
    if (n > ((int64_t)1 << 29)) return EWN_ENOSUP;        // root id * 6 + child is a 32-bit Philox stream id
    cudaStream_t st = (cudaStream_t)stream;
    const size_t need = ewn_uct_workspace(n);
    if (!workspace || workspace_bytes == 0) { int rc = scratch_get(need, &workspace, st); if (rc) return rc; }
    else if (workspace_bytes < need) return EWN_ESPACE;
    Geo geo = make_geo(N, L);
    char* w = (char*)workspace;
    uint64_t* child = (uint64_t*)w;                       // n x 2 words
    uint32_t* rid = (uint32_t*)(w + (size_t)n * 16);
    uint32_t* res = rid + n;
    int32_t* chosen = (int32_t*)(res + n);
    CU(cudaMemsetAsync(res, 0, (size_t)n * 4, st));
    for (int r = 0; r < rounds; r++) {
        const double ln_total = r ? log((double)r * (double)batch) : 0.0;       // visits of the root so far; host libm on every path
        if (use_p5(N)) k_uct_round<P5><<<blocks_for(n), TPB, 0, st>>>(state, root_ids, n, geo, wins, visits, chosen, res, child, rid, batch, c, ln_total, r == 0);
        else           k_uct_round<P8><<<blocks_for(n), TPB, 0, st>>>(state, root_ids, n, geo, wins, visits, chosen, res, child, rid, batch, c, ln_total, r == 0);
        LAUNCHED("k_uct_round");
        int rc = use_p5(N) ? launch_playouts<P5, true>(child, rid, nullptr, n, (uint32_t)r * batch, batch, res, nullptr, geo, seed, st)
                           : launch_playouts<P8, true>(child, rid, nullptr, n, (uint32_t)r * batch, batch, res, nullptr, geo, seed, st);
        if (rc) return rc;
    }
    if (use_p5(N)) k_uct_final<P5><<<blocks_for(n), TPB, 0, st>>>(state, n, geo, wins, visits, chosen, res, batch, action);
    else           k_uct_final<P8><<<blocks_for(n), TPB, 0, st>>>(state, n, geo, wins, visits, chosen, res, batch, action);
    LAUNCHED("k_uct_final");
    return EWN_OK;
}

int ewn_mc_best_action(const uint64_t* state, const uint32_t* wins, int8_t* action, int64_t n, int N, int L, void* stream) {
    if (!state || !wins || !action || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (L != 3) return EWN_ENOSUP;
    if (n == 0) return EWN_OK;
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    if (use_p5(N)) k_mc_best<P5><<<blocks_for(n), TPB, 0, st>>>(state, wins, action, n, geo);
    else           k_mc_best<P8><<<blocks_for(n), TPB, 0, st>>>(state, wins, action, n, geo);
    LAUNCHED("k_mc_best");
    return EWN_OK;
}

int ewn_peer_alloc(size_t bytes, void** ptr, ewn_peer_handle* handle) {
    if (!ptr || !handle || bytes == 0) return EWN_EARG;
    static_assert(sizeof(cudaIpcMemHandle_t) <= sizeof(ewn_peer_handle), "handle size");
    void* p = nullptr;
    CU(cudaMalloc(&p, bytes));
    cudaError_t e = cudaMemset(p, 0, bytes);
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); CU(e); }
    memset(handle, 0, sizeof(*handle));
    memcpy(handle->bytes, &h, sizeof(h));
    *ptr = p;
    return EWN_OK;
}

int ewn_peer_open(const ewn_peer_handle* handle, void** ptr) {
    if (!handle || !ptr) return EWN_EARG;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle->bytes, sizeof(h));
    CU(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return EWN_OK;
}

int ewn_peer_close(void* ptr) {
    if (!ptr) return EWN_EARG;
    CU(cudaIpcCloseMemHandle(ptr));
    return EWN_OK;
}

int ewn_peer_free(void* ptr) {
    if (!ptr) return EWN_EARG;
    CU(cudaFree(ptr));
    return EWN_OK;
}

int ewn_mc_best_action_peers(const uint64_t* state, const uint32_t* const* wins_peers, uint32_t* const* flag_peers, int rank, int world,
                             uint32_t epoch, uint32_t timeout_ms, uint32_t* wins_total, int8_t* action, int64_t n, int N, int L, void* stream) {
    if (!state || !wins_peers || !flag_peers || !action || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (world < 1 || world > EWN_MAX_PEERS || rank < 0 || rank >= world || epoch == 0) return EWN_EARG;
    if (L != 3) return EWN_ENOSUP;
    PeerSet ps;
    for (int r = 0; r < EWN_MAX_PEERS; r++) {
        ps.wins[r] = r < world ? wins_peers[r] : nullptr;
        ps.flags[r] = r < world ? flag_peers[r] : nullptr;
        if (r < world && (!ps.wins[r] || !ps.flags[r])) return EWN_EARG;
    }
    Geo geo = make_geo(N, L);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long ns = (unsigned long long)(timeout_ms ? timeout_ms : 2000u) * 1000000ull;
    int64_t blocks = n ? blocks_for(n) : 1;                                                        // n == 0 still takes part in the epoch
    if (use_p5(N)) k_mc_best_peers<P5><<<blocks, TPB, 0, st>>>(state, ps, rank, world, epoch, ns, wins_total, action, n, geo);
    else           k_mc_best_peers<P8><<<blocks, TPB, 0, st>>>(state, ps, rank, world, epoch, ns, wins_total, action, n, geo);
    LAUNCHED("k_mc_best_peers");
    return EWN_OK;
}

// ------------------------------------------------------------------------------------------------
// host-buffer vector environment
// ------------------------------------------------------------------------------------------------
struct ewn_vec {
    ewn_vec_config cfg;
    int words;
    int n_streams;
    cudaStream_t streams[4];
    cudaStream_t h2d;              // action uploads run ahead of the compute / download streams (PCIe is full duplex)
    cudaEvent_t ev[64];            // ev[k]: chunk k's actions are on the device
    uint64_t step;
    uint64_t* state;
    int64_t* action;
    int16_t* obs_board; int64_t* obs_dice; float* reward; uint8_t* done; uint8_t* trunc;
    int16_t* term_board; int64_t* term_dice; int32_t* term_index; uint32_t* term_count; uint8_t* term_code;
    uint32_t* count_host;          // page-locked landing slot for *term_count
    int32_t* tolerance; double* prev_score;
};

static void vec_free(ewn_vec* v) {
    if (!v) return;
    cudaSetDevice(v->cfg.device);
    for (int i = 0; i < v->n_streams; i++) if (v->streams[i]) cudaStreamDestroy(v->streams[i]);
    if (v->h2d) cudaStreamDestroy(v->h2d);
    for (int i = 0; i < 64; i++) if (v->ev[i]) cudaEventDestroy(v->ev[i]);
    cudaFree(v->state); cudaFree(v->action); cudaFree(v->obs_board); cudaFree(v->obs_dice); cudaFree(v->reward);
    cudaFree(v->done); cudaFree(v->trunc); cudaFree(v->term_board); cudaFree(v->term_dice); cudaFree(v->term_index); cudaFree(v->term_count); cudaFree(v->term_code);
    if (v->count_host) cudaFreeHost(v->count_host);
    cudaFree(v->tolerance); cudaFree(v->prev_score);
    delete v;
}

int ewn_vec_create(const ewn_vec_config* cfg, ewn_vec** out) {
    if (!cfg || !out || cfg->n <= 0 || !config_ok(cfg->board_size, cfg->cube_layer)) return EWN_EARG;
    ewn_vec* v = new (std::nothrow) ewn_vec();
    if (!v) return EWN_EARG;
    memset(v, 0, sizeof(*v));
    v->cfg = *cfg;
    if (v->cfg.chunks <= 0) v->cfg.chunks = 8;
    if (v->cfg.chunks > 64) v->cfg.chunks = 64;
    v->words = make_geo(cfg->board_size, cfg->cube_layer).sw;
    v->n_streams = 3;
    const int64_t n = cfg->n;
    const int NN = cfg->board_size * cfg->board_size;
    cudaError_t e = cudaSetDevice(cfg->device);
    if (e != cudaSuccess) { delete v; return cuda_fail(e, "cudaSetDevice"); }
#define VEC_TRY(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { int rc = cuda_fail(e__, #call); vec_free(v); return rc; } } while (0)
    for (int i = 0; i < v->n_streams; i++) VEC_TRY(cudaStreamCreateWithFlags(&v->streams[i], cudaStreamNonBlocking));
    VEC_TRY(cudaStreamCreateWithFlags(&v->h2d, cudaStreamNonBlocking));
    for (int i = 0; i < v->cfg.chunks; i++) VEC_TRY(cudaEventCreateWithFlags(&v->ev[i], cudaEventDisableTiming));
    VEC_TRY(cudaMalloc(&v->state, (size_t)n * v->words * 8));
    VEC_TRY(cudaMalloc(&v->action, (size_t)n * 16));
    VEC_TRY(cudaMalloc(&v->obs_board, (size_t)n * NN * 2));
    VEC_TRY(cudaMalloc(&v->obs_dice, (size_t)n * 8));
    VEC_TRY(cudaMalloc(&v->reward, (size_t)n * 4));
    VEC_TRY(cudaMalloc(&v->done, (size_t)n));
    VEC_TRY(cudaMalloc(&v->trunc, (size_t)n));
    VEC_TRY(cudaMalloc(&v->term_board, (size_t)n * NN * 2));
    VEC_TRY(cudaMalloc(&v->term_dice, (size_t)n * 8));
    VEC_TRY(cudaMalloc(&v->term_index, (size_t)n * 4));
    VEC_TRY(cudaMalloc(&v->term_code, (size_t)n));
    VEC_TRY(cudaMalloc(&v->term_count, 256));
    VEC_TRY(cudaMemset(v->term_count, 0, 256));
    VEC_TRY(cudaHostAlloc((void**)&v->count_host, 64, cudaHostAllocDefault));
    if (cfg->shaped) {
        VEC_TRY(cudaMalloc(&v->tolerance, (size_t)n * 4));
        VEC_TRY(cudaMalloc(&v->prev_score, (size_t)n * 8));
    }
#undef VEC_TRY
    *out = v;
    return EWN_OK;
}

void ewn_vec_destroy(ewn_vec* v) { vec_free(v); }
uint64_t* ewn_vec_state(ewn_vec* v) { return v ? v->state : nullptr; }
uint64_t ewn_vec_step_index(const ewn_vec* v) { return v ? v->step : 0; }

// fills an int32 / f64 array with a constant (shaped-env bookkeeping at creation time)
__global__ void k_fill_i32(int32_t* p, int32_t v, int64_t n) { int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x; if (i < n) p[i] = v; }

int ewn_vec_reset(ewn_vec* v, int16_t* obs_board, int64_t* obs_dice) {
    if (!v) return EWN_EARG;
    CU(cudaSetDevice(v->cfg.device));
    const ewn_vec_config& c = v->cfg;
    cudaStream_t st = v->streams[0];
    const int NN = c.board_size * c.board_size;
    int rc = ewn_reset(v->state, c.n, c.board_size, c.cube_layer, c.seed, v->step, c.env_id0, st);
    if (rc) return rc;
    if (c.shaped && v->step == 0) {
        // training_ewn.py:35-38: tolerance budget and prev_score are set once, at construction
        k_fill_i32<<<blocks_for(c.n), TPB, 0, st>>>(v->tolerance, c.illegal_move_tolerance, c.n);
        LAUNCHED("k_fill_i32");
        rc = ewn_evaluate(v->state, v->prev_score, c.n, c.board_size, c.cube_layer, EWN_H_HYBRID, st);
        if (rc) return rc;
    }
    // a reset shifts the parity of the step index: neither append counter may carry a count over (the step that would have
    // cleared it never ran)
    CU(cudaMemsetAsync(v->term_count, 0, 2 * sizeof(uint32_t), st));
    rc = ewn_decode(v->state, v->obs_board, v->obs_dice, nullptr, c.n, c.board_size, c.cube_layer, st);
    if (rc) return rc;
    if (obs_board) CU(cudaMemcpyAsync(obs_board, v->obs_board, (size_t)c.n * NN * 2, cudaMemcpyDeviceToHost, st));
    if (obs_dice) CU(cudaMemcpyAsync(obs_dice, v->obs_dice, (size_t)c.n * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    v->step++;
    return EWN_OK;
}

// device-visible alias of a page-locked host buffer (unified addressing maps every cudaHostAlloc / cudaHostRegister
// allocation into the device address space), or nullptr if `p` is pageable
static void* pinned_alias(const void* p) {
    if (!p) return nullptr;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return at.type == cudaMemoryTypeHost ? at.devicePointer : nullptr;
}

int ewn_vec_set_seed(ewn_vec* v, uint64_t seed) {
    if (!v) return EWN_EARG;
    v->cfg.seed = seed;
    return EWN_OK;
}

int ewn_vec_step(ewn_vec* v, const int64_t* action, int16_t* obs_board, int64_t* obs_dice, float* reward, uint8_t* done, uint8_t* truncated,
                 int16_t* term_board, int64_t* term_dice, int32_t* term_index, uint32_t* term_count) {
    return ewn_vec_step2(v, action, obs_board, obs_dice, reward, done, truncated, term_board, term_dice, term_index, nullptr, term_count);
}

int ewn_vec_step2(ewn_vec* v, const int64_t* action, int16_t* obs_board, int64_t* obs_dice, float* reward, uint8_t* done, uint8_t* truncated,
                  int16_t* term_board, int64_t* term_dice, int32_t* term_index, uint8_t* term_code, uint32_t* term_count) {
    if (!v || !action) return EWN_EARG;
    if (term_count && !term_index) return EWN_EARG;
    CU(cudaSetDevice(v->cfg.device));
    const ewn_vec_config& c = v->cfg;
    const int NN = c.board_size * c.board_size;
    // DIRECT mode: when every output buffer the caller handed in is page-locked, the kernel stores its results straight into
    // host memory over PCIe (coalesced 16-byte vectors) -- no device staging copy, no D2H memcpy, no pipeline ramp: the step is
    // one pass that runs at the PCIe write rate.  Pageable buffers take the staged path (device arrays + cudaMemcpyAsync).
    int16_t* d_board = (int16_t*)pinned_alias(obs_board); int64_t* d_dice = (int64_t*)pinned_alias(obs_dice);
    float* d_rew = (float*)pinned_alias(reward); uint8_t* d_done = (uint8_t*)pinned_alias(done); uint8_t* d_trunc = (uint8_t*)pinned_alias(truncated);
    int16_t* d_tb = (int16_t*)pinned_alias(term_board); int64_t* d_td = (int64_t*)pinned_alias(term_dice); int32_t* d_ti = (int32_t*)pinned_alias(term_index);
    uint8_t* d_tc = (uint8_t*)pinned_alias(term_code);
    const bool direct = !c.shaped && !(c.flags & EWN_VEC_STAGED) && d_board && d_dice && d_rew && d_done && d_trunc &&
                        (!term_count || ((!term_board || d_tb) && (!term_dice || d_td) && d_ti && (!term_code || d_tc)));
    // chunk boundaries are multiples of 1024 games so that every chunk starts 16-byte aligned in every array
    int64_t per = (c.n + c.chunks - 1) / c.chunks;
    per = (per + 1023) & ~(int64_t)1023;
    int k = 0;
    for (int64_t lo = 0; lo < c.n; lo += per, k++) {       // all uploads first, on their own stream: they never queue behind a download
        const int64_t m = (c.n - lo < per) ? (c.n - lo) : per;
        CU(cudaMemcpyAsync(v->action + 2 * lo, action + 2 * lo, (size_t)m * 16, cudaMemcpyHostToDevice, v->h2d));
        CU(cudaEventRecord(v->ev[k], v->h2d));
    }
    k = 0;
    for (int64_t lo = 0; lo < c.n; lo += per, k++) {
        const int64_t m = (c.n - lo < per) ? (c.n - lo) : per;
        cudaStream_t st = v->streams[k % v->n_streams];
        CU(cudaStreamWaitEvent(st, v->ev[k], 0));
        ewn_step_args a;
        memset(&a, 0, sizeof(a));
        a.state = v->state + lo * v->words; a.n = m; a.board_size = c.board_size; a.cube_layer = c.cube_layer;
        a.action_i64 = v->action + 2 * lo;
        if (direct) {
            a.reward_f32 = d_rew + lo; a.terminated = d_done + lo; a.truncated = d_trunc + lo;
            a.obs_board = d_board + lo * NN; a.obs_dice = d_dice + lo;
        } else {
            a.reward_f32 = v->reward + lo; a.terminated = v->done + lo; a.truncated = v->trunc + lo;
            a.obs_board = v->obs_board + lo * NN; a.obs_dice = v->obs_dice + lo;
        }
        if (term_count) {               // compacted terminal observations, one list for all chunks
            if (direct) { a.term_board = d_tb; a.term_dice = d_td; a.term_index = d_ti; a.term_code = d_tc; }
            else { a.term_board = term_board ? v->term_board : nullptr; a.term_dice = term_dice ? v->term_dice : nullptr; a.term_index = v->term_index;
                   a.term_code = term_code ? v->term_code : nullptr; }
            a.term_count = v->term_count + (v->step & 1); a.term_index_base = lo;
        }
        // every step clears the counter the NEXT one appends to, whether or not this call asked for terminal observations: a caller
        // that skips them for one step must not find the count of two steps ago (rows would be appended past it)
        if (lo == 0) a.term_count_clear = v->term_count + ((v->step + 1) & 1);
        if (c.shaped) { a.tolerance = v->tolerance + lo; a.prev_score = v->prev_score + lo; a.shaped = 1; }
        a.reward = c.reward; a.illegal_move_reward = c.illegal_move_reward;
        a.seed = c.seed; a.step = v->step; a.env_id0 = c.env_id0 + (uint64_t)lo;
        a.autoreset = 1;
        a.flags = EWN_STEP_NO_PDL;      // chunks run on different streams: nothing to chain
        int rc = ewn_step(&a, st);
        if (rc) return rc;
        if (direct) continue;
        if (obs_board) CU(cudaMemcpyAsync(obs_board + lo * NN, v->obs_board + lo * NN, (size_t)m * NN * 2, cudaMemcpyDeviceToHost, st));
        if (obs_dice) CU(cudaMemcpyAsync(obs_dice + lo, v->obs_dice + lo, (size_t)m * 8, cudaMemcpyDeviceToHost, st));
        if (reward) CU(cudaMemcpyAsync(reward + lo, v->reward + lo, (size_t)m * 4, cudaMemcpyDeviceToHost, st));
        if (done) CU(cudaMemcpyAsync(done + lo, v->done + lo, (size_t)m, cudaMemcpyDeviceToHost, st));
        if (truncated) CU(cudaMemcpyAsync(truncated + lo, v->trunc + lo, (size_t)m, cudaMemcpyDeviceToHost, st));
    }
    // every kernel has run once the streams drain; the append counter is final after all of them
    for (int i = 1; i < v->n_streams; i++) CU(cudaStreamSynchronize(v->streams[i]));
    cudaStream_t s0 = v->streams[0];
    if (term_count) {
        CU(cudaMemcpyAsync(v->count_host, v->term_count + (v->step & 1), 4, cudaMemcpyDeviceToHost, s0));
        CU(cudaStreamSynchronize(s0));
        const uint32_t cnt = *v->count_host;
        *term_count = cnt;
        if (cnt && !direct) {
            if (term_board) CU(cudaMemcpyAsync(term_board, v->term_board, (size_t)cnt * NN * 2, cudaMemcpyDeviceToHost, s0));
            if (term_dice) CU(cudaMemcpyAsync(term_dice, v->term_dice, (size_t)cnt * 8, cudaMemcpyDeviceToHost, s0));
            CU(cudaMemcpyAsync(term_index, v->term_index, (size_t)cnt * 4, cudaMemcpyDeviceToHost, s0));
            if (term_code) CU(cudaMemcpyAsync(term_code, v->term_code, (size_t)cnt, cudaMemcpyDeviceToHost, s0));
        }
    }
    CU(cudaStreamSynchronize(s0));
    v->step++;
    return EWN_OK;
}

int ewn_host_alloc(void** ptr, size_t bytes) {
    if (!ptr) return EWN_EARG;
    CU(cudaHostAlloc(ptr, bytes ? bytes : 1, cudaHostAllocDefault));
    return EWN_OK;
}
int ewn_host_free(void* ptr) {
    if (ptr) CU(cudaFreeHost(ptr));
    return EWN_OK;
}

}  // extern "C"
