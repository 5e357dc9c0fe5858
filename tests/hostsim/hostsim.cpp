// hostsim.cpp -- TEST INFRASTRUCTURE ONLY.  Compiles ewn_gym_b200/csrc/ewn_core.cuh (the per-game bodies of the
// CUDA kernels) with g++ and drives them with plain loops behind the SAME C signatures as include/ewn_b200.h, so
// the CPU-only test tier can check the packed-state logic against the oracle before GPU time is spent, and so the
// kernel bodies can be stepped in gdb / run under sanitizers.  Never loaded by the ewn_gym_b200 package; the
// product has no CPU path (libewn_b200.so returns EWN_ECUDA without a device).
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include "../../include/ewn_b200.h"
#include "../../ewn_gym_b200/csrc/ewn_core.cuh"

using namespace ewn;

namespace {
inline bool config_ok(int N, int L) { return N >= 3 && N <= EWN_MAX_BOARD && L >= 1 && L < N - 1; }
// f(codec tag) with the codec of (N, L): P5 / P8 for the six-cube layouts, PG for cube_layer >= 4
template <class F> auto with_codec(int N, int L, F&& f) {
    switch (codec_of(N, L)) { case 0: return f(P5{}); case 1: return f(P8{}); default: return f(PG{}); }
}
#define WITH_CODEC(N, L, expr) with_codec(N, L, [&](auto c_) { typedef decltype(c_) C; return (expr); })

template <class C> void decode_one(const Game<C>& g, const Geo& geo, int16_t* row) {
    const int N = geo.N;
    memset(row, 0, sizeof(int16_t) * N * N);
    for (int k = 0; k < C::ncubes(geo); k++) {
        int c = C::get(g.tl, k);
        if (c != (int)C::FM) row[C::x_of(c) * N + C::y_of(c)] = (int16_t)(k + 1);
        int d = C::get(g.br, k);
        if (d != (int)C::FM) row[C::x_of(d) * N + C::y_of(d)] = (int16_t)(-(k + 1));
    }
}

template <class C> int t_reset(uint64_t* state, const int32_t* dice_in, int64_t n, const Geo& geo, u64 seed, u64 step, u64 env_id0) {
    for (int64_t i = 0; i < n; i++) {
        int dice;
        if (dice_in) { dice = dice_in[i]; dice = dice < 1 ? 1 : (dice > geo.cube_num ? geo.cube_num : dice); }
        else { u32 blk[4]; env_block(seed, env_id0 + (u64)i, step, TAG_RESET, blk); dice = 1 + (int)bounded(blk[0], (u32)geo.cube_num); }
        Game<C> g; setup<C>(g, geo, dice);
        pack(g, state, i, geo);
    }
    return EWN_OK;
}

template <class C> int t_encode(const int16_t* board, const int32_t* dice, const int32_t* player, uint64_t* state, int64_t n, const Geo& geo) {
    const int N = geo.N, NN = N * N, nc = geo.cube_num;
    for (int64_t i = 0; i < n; i++) {
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
    return EWN_OK;
}

template <class C> int t_decode(const uint64_t* state, int16_t* board, int64_t* dice, int32_t* player, int64_t n, const Geo& geo) {
    const int NN = geo.N * geo.N;
    for (int64_t i = 0; i < n; i++) {
        Game<C> g; unpack(g, state, i, geo);
        if (dice) dice[i] = g.dice;
        if (player) player[i] = g.player == SIDE_BR ? EWN_BOTTOM_RIGHT : EWN_TOP_LEFT;
        if (board) decode_one<C>(g, geo, board + i * NN);
    }
    return EWN_OK;
}

template <class C> int t_legal(const uint64_t* state, int player, uint8_t* mask, int8_t* list, int32_t* count, int64_t n, const Geo& geo) {
    for (int64_t i = 0; i < n; i++) {
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
    return EWN_OK;
}

template <class C> int t_ply(uint64_t* state, int player, const int32_t* action, uint8_t* valid, uint8_t* win, int32_t* moved, int64_t n,
                             const Geo& geo) {
    for (int64_t i = 0; i < n; i++) {
        Game<C> g; unpack(g, state, i, geo);
        int side = player == 0 ? g.player : (player == EWN_BOTTOM_RIGHT ? SIDE_BR : SIDE_TL);
        int t = 0, sh = -1; bool ok;
        if (side == SIDE_BR) ok = ply_action<C, SIDE_BR>(g, geo, action[2 * i], action[2 * i + 1], t, sh);
        else                 ok = ply_action<C, SIDE_TL>(g, geo, action[2 * i], action[2 * i + 1], t, sh);
        if (ok) g.player = side ^ 1;
        pack(g, state, i, geo);
        if (valid) valid[i] = ok ? 1 : 0;
        if (win) win[i] = full_win<C>(g, geo) ? 1 : 0;
        if (moved) moved[i] = sh < 0 ? 0 : sh / C::FB + 1;
    }
    return EWN_OK;
}

template <class C> int t_step(const ewn_step_args& a, const Geo& geo) {
    const int NN = geo.N * geo.N;
    const bool replay = a.inject_dice_opp != nullptr;
    StepParams p; memset(&p, 0, sizeof(p));
    p.geo = geo; p.reward = a.reward; p.illegal_move_reward = a.illegal_move_reward;
    p.seed = a.seed; p.step = a.step + (a.step_dev ? *a.step_dev : 0); p.env_id0 = a.env_id0; p.rk = make_round_keys(a.seed);
    p.agent_random = a.agent_random != 0; p.autoreset = a.autoreset != 0; p.shaped = a.shaped != 0; p.replay = replay;
    if (a.term_count_clear) *a.term_count_clear = 0u;
    for (int64_t i = 0; i < a.n; i++) {
        Game<C> g, term; unpack(g, a.state, i, geo); term = g;
        int al = 0, ad = 0;
        if (!p.agent_random) {
            if (a.action_i64) { al = (int)a.action_i64[2 * i]; int64_t y = a.action_i64[2 * i + 1]; ad = (y < 0 || y > 2) ? 3 : (int)y; }
            else if (a.action_i32) { al = a.action_i32[2 * i]; ad = a.action_i32[2 * i + 1]; }
            else { int v = a.action_u8[i]; al = v >= 3; ad = v - 3 * al; }
        }
        StepInject inj = {1, 0, 0, 1};
        if (replay) { const int cn = geo.cube_num; auto cl = [&](int d) { return d < 1 ? 1 : (d > cn ? cn : d); };
            inj.dice_opp = cl(a.inject_dice_opp[i]); inj.dice_next = cl(a.inject_dice_next[i]); inj.o_larger = a.inject_opp_action[2 * i]; inj.o_dir = a.inject_opp_action[2 * i + 1]; }
        int tol = 0; double prev = 0.0;
        if (p.shaped) { tol = a.tolerance[i]; prev = a.prev_score[i]; }
        StepOut out;
        step_one<C>(g, term, p, i, al, ad, inj, &tol, &prev, out);
        if (p.shaped) { a.tolerance[i] = tol; a.prev_score[i] = prev; }
        pack(g, a.state, i, geo);
        if (a.reward_f32) a.reward_f32[i] = (float)out.reward;
        if (a.reward_f64) a.reward_f64[i] = out.reward;
        if (a.terminated) a.terminated[i] = (uint8_t)out.term;
        if (a.truncated) a.truncated[i] = (uint8_t)out.trunc;
        if (a.code) a.code[i] = out.code;
        if (a.action_out) a.action_out[i] = (uint8_t)(out.a_larger * 3 + out.a_dir);
        if (a.obs_dice) a.obs_dice[i] = g.dice;
        if (a.obs_board) decode_one<C>(g, geo, a.obs_board + i * NN);
        if (out.term && a.stats) {
            if (out.code == CODE_AGENT_WON) a.stats[0]++; else if (out.code == CODE_AGENT_LOST) a.stats[1]++; else a.stats[2]++;
            a.stats[3]++;
        }
        if (out.term) {
            int64_t row = i;
            if (a.term_count) { row = (*a.term_count)++; a.term_index[row] = (int32_t)(a.term_index_base + i); }
            if (a.term_dice) a.term_dice[row] = term.dice;
            if (a.term_code) a.term_code[row] = (uint8_t)out.code;
            if (a.term_board) decode_one<C>(term, geo, a.term_board + row * NN);
        }
    }
    return EWN_OK;
}

template <class C, int M> void emm_replay_all(const uint64_t* state, const double* vals, int64_t n, const Geo& geo, int heuristic, int pruned,
                                              double* value, int8_t* action) {
    for (int64_t r = 0; r < n; r++) {
        Game<C> g; unpack(g, state, r); g.player = SIDE_TL;
        int la, di;
        value[r] = emm_root<C, M>(g, geo, heuristic, pruned, vals + r * emm_level_offset(M + 1), la, di);
        if (action) { action[2 * r] = (int8_t)la; action[2 * r + 1] = (int8_t)di; }
    }
}

// depth 3 / 4: the closed-form prefix rule (emm2_*), one worker per root; every second root looks its 'hybrid' values up in a table
// built like the device's (k_emm_table), the others compute them in place -- both paths of emm_summary_value are exercised
template <class C> int t_emm_fused(const uint64_t* state, int64_t n, const Geo& geo, int depth, int heuristic, int pruned, double* value, int8_t* action) {
    std::vector<u32> inc(C::LUT);
    for (int c = 0; c < C::LUT; c++) inc[c] = emm_inc_entry<C>(geo.N, c);
    const int half = emm_tab_entries(geo.N), pad = emm_pad(geo.N);
    std::vector<double> tv((size_t)(pad + half), 0.0), tq((size_t)(pad + half), 0.0);
    const bool fold = !(depth & 1);
    for (int i = 0; i < half; i++) { tv[pad + i] = emm_table_entry(geo.N, (fold ? half : 0) + i); tq[pad + i] = f64_div(tv[pad + i], 6.0); }
    tv[pad - 1] = -10.0; tq[pad - 1] = f64_div(-10.0, 6.0);
    for (int64_t r = 0; r < n; r++) {
        Game<C> g; unpack(g, state, r); g.player = SIDE_TL;
        const bool tab = (r & 1) && heuristic == H_HYBRID;
        EmmTabs tb; tb.inc = inc.data(); tb.v = tab ? tv.data() + pad : nullptr; tb.q = tab ? tq.data() + pad : nullptr;
        int la, di;
        value[r] = emm2_root_serial<C>(g, geo, tb, depth, heuristic, pruned, la, di);
        if (action) { action[2 * r] = (int8_t)la; action[2 * r + 1] = (int8_t)di; }
    }
    return EWN_OK;
}

template <class C> int t_emm(const uint64_t* state, int64_t n, const Geo& geo, int depth, int heuristic, int pruned, double* value, int8_t* action) {
    const int m = emm_levels(depth);
    if (m == 2 && !(pruned & 2)) return t_emm_fused<C>(state, n, geo, depth, heuristic, pruned & 1, value, action);
    pruned &= 1;
    const long slots = emm_level_size(m), per_root = emm_level_offset(m + 1);
    const int64_t chunk = 256;
    std::vector<double> vals((size_t)chunk * per_root);
    for (int64_t r0 = 0; r0 < n; r0 += chunk) {
        const int64_t nr = n - r0 < chunk ? n - r0 : chunk;
        for (int64_t r = 0; r < nr; r++) {
            Game<C> g; unpack(g, state, r0 + r); g.player = SIDE_TL;
            if (full_win<C>(g, geo)) continue;
            for (long tp = 0; tp < slots / 6; tp++) emm_expand_parent<C>(g, geo, depth, heuristic, (u32)tp, vals.data() + r * per_root);
        }
        const uint64_t* s0 = state + r0 * C::WORDS;
        double* v0 = value + r0; int8_t* a0 = action ? action + 2 * r0 : nullptr;
        switch (m) {
            case 1: emm_replay_all<C, 1>(s0, vals.data(), nr, geo, heuristic, pruned, v0, a0); break;
            case 2: emm_replay_all<C, 2>(s0, vals.data(), nr, geo, heuristic, pruned, v0, a0); break;
            case 3: emm_replay_all<C, 3>(s0, vals.data(), nr, geo, heuristic, pruned, v0, a0); break;
            default: emm_replay_all<C, 4>(s0, vals.data(), nr, geo, heuristic, pruned, v0, a0); break;
        }
    }
    return EWN_OK;
}

template <class C> int t_playouts(const uint64_t* state, const uint32_t* root_ids, int64_t n, const Geo& geo, u64 seed, u32 first, u32 count,
                                  uint32_t* wins, unsigned long long* plies) {
    for (int64_t i = 0; i < n; i++) {
        Game<C> g; unpack(g, state, i, geo);
        for (u32 p = first; p < first + count; p++) {
            int pl = 0;
            wins[i] += (u32)playout<C>(g, geo, seed, root_ids ? root_ids[i] : (u32)i, 7u, p, &pl);
            if (plies) *plies += (unsigned long long)pl;
        }
    }
    return EWN_OK;
}

template <class C> int t_uct(const uint64_t* state, const uint32_t* root_ids, int64_t n, const Geo& geo, u64 seed, int rounds, u32 batch, double c,
                             uint32_t* wins, uint32_t* visits, int8_t* action) {
    for (int64_t i = 0; i < n; i++) {
        Game<C> g, ch; unpack(g, state, i, geo); g.player = SIDE_TL;
        u32 w[6] = {0, 0, 0, 0, 0, 0}, v[6] = {0, 0, 0, 0, 0, 0}, res = 0;
        int prev = -1, nch = 0;
        const u32 rid = root_ids ? root_ids[i] : (u32)i;
        for (int r = 0; r < rounds; r++) {
            const double ln_total = r ? log((double)r * (double)batch) : 0.0;
            const int j = uct_round<C>(g, geo, w, v, prev, res, batch, c, ln_total, ch, nch);
            if (j < 0) break;
            res = 0;
            for (u32 p = (u32)r * batch; p < (u32)(r + 1) * batch; p++) res += (u32)playout<C>(ch, geo, seed, rid * 6u + (u32)j, 7u, p, nullptr);
            prev = j;
        }
        int la = -1, di = -1;
        if (prev >= 0) {
            w[prev] += res; v[prev] += batch;
            const Legal l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
            int sh; legal_pick(l, uct_best(w, v, legal_count(l)), la, di, sh);
        }
        for (int j = 0; j < 6; j++) { wins[i * 6 + j] = w[j]; visits[i * 6 + j] = v[j]; }
        if (action) { action[2 * i] = (int8_t)la; action[2 * i + 1] = (int8_t)di; }
    }
    return EWN_OK;
}

template <class C> int t_mc(const uint64_t* state, const uint32_t* root_ids, int64_t n, const Geo& geo, u64 seed, u32 first, u32 count,
                            uint32_t* wins, int32_t* n_children, unsigned long long* plies) {
    for (int64_t i = 0; i < n; i++) {
        Game<C> g; unpack(g, state, i, geo); g.player = SIDE_TL;
        int cnt = 0;
        if (!full_win<C>(g, geo)) {
            Legal l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
            cnt = legal_count(l);
            for (int j = 0; j < cnt; j++) {
                Game<C> c = g; int la, di, sh; legal_pick(l, j, la, di, sh);
                apply_move<C, SIDE_TL>(c, sh, di);
                for (u32 p = first; p < first + count; p++) {
                    int pl = 0;
                    wins[i * 6 + j] += (u32)playout<C>(c, geo, seed, root_ids ? root_ids[i] : (u32)i, (u32)j, p, &pl);
                    if (plies) *plies += (unsigned long long)pl;
                }
            }
        }
        if (n_children) n_children[i] = cnt;
    }
    return EWN_OK;
}

template <class C> int t_mc_best(const uint64_t* state, const uint32_t* wins, int8_t* action, int64_t n, const Geo& geo) {
    for (int64_t i = 0; i < n; i++) {
        Game<C> g; unpack(g, state, i, geo);
        int la = -1, di = -1;
        if (!full_win<C>(g, geo)) {
            Legal l = legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
            int cnt = legal_count(l), best = 0;
            for (int j = 1; j < cnt; j++) if (wins[i * 6 + j] > wins[i * 6 + best]) best = j;
            int sh; legal_pick(l, best, la, di, sh);
        }
        action[2 * i] = (int8_t)la; action[2 * i + 1] = (int8_t)di;
    }
    return EWN_OK;
}
}  // namespace

template <class C> static int t_flip(const uint64_t* in, uint64_t* out, int64_t n, const Geo& geo) {
    for (int64_t i = 0; i < n; i++) { Game<C> g; unpack(g, in, i, geo); flip_game<C>(g, geo); pack(g, out, i, geo); }
    return EWN_OK;
}
template <class C> int t_sim_winrate(const uint64_t* state, int64_t n, const Geo& geo, int nsim, u64 seed, u32 unit0, double* value) {
    const RoundKeys rk = make_round_keys(seed);
    for (int64_t i = 0; i < n; i++) {
        Game<C> g; unpack(g, state, i, geo);
        value[i] = full_win<C>(g, geo) ? (tl_won<C>(g, geo) ? 1.0 : 0.0) : sim_unit_value<C>(g, geo, rk, (unit0 + (u32)i) * 6u, nsim);
    }
    return EWN_OK;
}
// the 'sim_winrate' search as the CUDA path does it: expand to leaf positions, evaluate the pending ones, replay
template <class C> int t_emm_sim(const uint64_t* state, int64_t n, const Geo& geo, int depth, int pruned, int nsim, u64 seed, double* value, int8_t* action) {
    const int m = emm_levels(depth);
    const long slots = emm_level_size(m), per_root = emm_level_offset(m + 1);
    const RoundKeys rk = make_round_keys(seed);
    std::vector<double> vals((size_t)per_root);
    std::vector<u64> sst((size_t)per_root * C::WORDS);
    for (int64_t r = 0; r < n; r++) {
        Game<C> g; unpack(g, state, r); g.player = SIDE_TL;
        if (!full_win<C>(g, geo)) {
            GeoSimT<Geo> gs; static_cast<Geo&>(gs) = geo; gs.sim_base = vals.data(); gs.sim_states = sst.data();
            for (long tp = 0; tp < slots / 6; tp++) emm_expand_parent<C>(g, gs, depth, H_SIM_WINRATE, (u32)tp, vals.data());
            for (long s = 0; s < per_root; s++) {
                const u64 rec = f64_to_bits(vals[s]);
                if ((rec & ~4ull) != EMM_PENDING) continue;
                Game<C> leaf; unpack(leaf, sst.data(), s, geo);
                const int folds = (rec & 4ull) ? 6 : 1;
                double e = 0.0, v = 0.0;
                for (int f = 0; f < folds; f++) {
                    v = sim_unit_value<C>(leaf, geo, rk, (u32)((r * per_root + s) * 6 + f), nsim);
                    e = f64_add(e, f64_div(v, 6.0));
                }
                vals[s] = folds == 1 ? v : e;
            }
        }
        int la, di;
        value[r] = m == 1 ? emm_root<C, 1>(g, geo, H_SIM_WINRATE, pruned & 1, vals.data(), la, di) : emm_root<C, 2>(g, geo, H_SIM_WINRATE, pruned & 1, vals.data(), la, di);
        if (action) { action[2 * r] = (int8_t)la; action[2 * r + 1] = (int8_t)di; }
    }
    return EWN_OK;
}
#define BY_CODEC(N, call_p5, call_p8) (use_p5(N) ? (call_p5) : (call_p8))

extern "C" {
int ewn_version(void) { return -100; }   // negative: this is the host simulator, not the product library
const char* ewn_strerror(int) { return "hostsim"; }
const char* ewn_last_cuda_error(void) { return ""; }
int ewn_device_count(void) { return 0; }
int ewn_state_words(int N) { return (N >= 3 && N <= EWN_MAX_BOARD) ? (use_p5(N) ? 1 : 2) : EWN_EARG; }
int ewn_state_words_cfg(int N, int L) { return config_ok(N, L) ? make_geo(N, L).sw : EWN_EARG; }
int64_t ewn_launch_count(void) { return 0; }

int ewn_reset(uint64_t* state, int64_t n, int N, int L, uint64_t seed, uint64_t step, uint64_t env_id0, void*) {
    if (!state || n < 0 || !config_ok(N, L)) return EWN_EARG;
    Geo geo = make_geo(N, L);
    return WITH_CODEC(N, L, t_reset<C>(state, nullptr, n, geo, seed, step, env_id0));
}
int ewn_reset_dice(uint64_t* state, const int32_t* dice, int64_t n, int N, int L, void*) {
    if (!state || !dice || n < 0 || !config_ok(N, L)) return EWN_EARG;
    Geo geo = make_geo(N, L);
    return WITH_CODEC(N, L, t_reset<C>(state, dice, n, geo, 0, 0, 0));
}
int ewn_encode(const int16_t* board, const int32_t* dice, const int32_t* player, uint64_t* state, int64_t n, int N, int L, void*) {
    if (!board || !state || n < 0 || !config_ok(N, L)) return EWN_EARG;
    Geo geo = make_geo(N, L);
    return WITH_CODEC(N, L, t_encode<C>(board, dice, player, state, n, geo));
}
int ewn_decode(const uint64_t* state, int16_t* board, int64_t* dice, int32_t* player, int64_t n, int N, int L, void*) {
    if (!state || n < 0 || !config_ok(N, L)) return EWN_EARG;
    Geo geo = make_geo(N, L);
    return WITH_CODEC(N, L, t_decode<C>(state, board, dice, player, n, geo));
}
int ewn_legal(const uint64_t* state, int player, uint8_t* mask, int8_t* list, int32_t* count, int64_t n, int N, int L, void*) {
    if (!state || n < 0 || !config_ok(N, L) || player < 0 || player > 2) return EWN_EARG;
    Geo geo = make_geo(N, L);
    return WITH_CODEC(N, L, t_legal<C>(state, player, mask, list, count, n, geo));
}
int ewn_check_win(const uint64_t* state, uint8_t* win, int64_t n, int N, int L, void*) {
    if (!state || !win || n < 0 || !config_ok(N, L)) return EWN_EARG;
    Geo geo = make_geo(N, L);
    return WITH_CODEC(N, L, ([&] { for (int64_t i = 0; i < n; i++) { Game<C> g; unpack(g, state, i, geo); win[i] = full_win<C>(g, geo); } return (int)EWN_OK; })());
}
int ewn_ply(uint64_t* state, int player, const int32_t* action, uint8_t* valid, uint8_t* win, int32_t* moved, int64_t n, int N, int L, void*) {
    if (!state || !action || n < 0 || !config_ok(N, L) || player < 0 || player > 2) return EWN_EARG;
    Geo geo = make_geo(N, L);
    return WITH_CODEC(N, L, t_ply<C>(state, player, action, valid, win, moved, n, geo));
}
int ewn_set_dice(uint64_t* state, const int32_t* dice, const int32_t* player, int64_t n, int N, int L, void*) {
    if (!state || n < 0 || !config_ok(N, L)) return EWN_EARG;
    Geo geo = make_geo(N, L);
    return WITH_CODEC(N, L, ([&] {
        for (int64_t i = 0; i < n; i++) {
            Game<C> g; unpack(g, state, i, geo);
            if (dice) { int d = dice[i]; g.dice = d < 1 ? 1 : (d > geo.cube_num ? geo.cube_num : d); }
            if (player) g.player = player[i] == EWN_BOTTOM_RIGHT;
            pack(g, state, i, geo);
        }
        return (int)EWN_OK; })());
}
int ewn_random_action(const uint64_t* state, int64_t* action, int64_t n, int N, int L, uint64_t seed, uint64_t step, uint64_t env_id0, void*) {
    if (!state || !action || n < 0 || !config_ok(N, L)) return EWN_EARG;
    Geo geo = make_geo(N, L);
    for (int64_t i = 0; i < n; i++) {
        u32 blk[4]; env_block(seed, env_id0 + (u64)i, step, TAG_STEP, blk);
        Legal l = WITH_CODEC(N, L, ([&] { Game<C> g; unpack(g, state, i, geo); return legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice)); })());
        int c = legal_count(l), la = 0, di = 0, sh;
        if (c > 0) legal_pick(l, (int)bounded(blk[3], (u32)c), la, di, sh);
        action[2 * i] = la; action[2 * i + 1] = di;
    }
    return EWN_OK;
}
int ewn_step(const ewn_step_args* a, void*) {
    if (!a || !a->state || a->n < 0 || !config_ok(a->board_size, a->cube_layer)) return EWN_EARG;
    if (!a->agent_random && !a->action_i64 && !a->action_i32 && !a->action_u8) return EWN_EARG;
    const bool replay = a->inject_dice_opp || a->inject_opp_action || a->inject_dice_next;
    if (replay && !(a->inject_dice_opp && a->inject_opp_action && a->inject_dice_next)) return EWN_EARG;
    if (a->shaped && !(a->tolerance && a->prev_score)) return EWN_EARG;
    Geo geo = make_geo(a->board_size, a->cube_layer);
    return WITH_CODEC(a->board_size, a->cube_layer, t_step<C>(*a, geo));
}
int ewn_step_dice(int32_t* dice_opp, int32_t* dice_next, int64_t n, int N, int L, uint64_t seed, uint64_t step, uint64_t env_id0,
                  const uint64_t* step_dev, void*) {
    if ((!dice_opp && !dice_next) || n < 0 || !config_ok(N, L)) return EWN_EARG;
    const int cn = L * (L + 1) / 2;
    for (int64_t i = 0; i < n; i++) {
        u32 blk[4]; env_block(seed, env_id0 + (u64)i, step + (step_dev ? *step_dev : 0), TAG_STEP, blk);
        if (dice_opp) dice_opp[i] = 1 + (int)bounded(blk[0], (u32)cn);
        if (dice_next) dice_next[i] = 1 + (int)bounded(blk[2], (u32)cn);
    }
    return EWN_OK;
}
int ewn_numpy_random_games(const uint32_t* seeds, int64_t n, int N, int L, int8_t* result, int32_t* agent_steps, uint64_t* final_state,
                           uint64_t* trace, void*) {
    if (!seeds || n < 0 || !config_ok(N, L)) return EWN_EARG;
    Geo geo = make_geo(N, L);
    MT19937 rng;
    for (int64_t i = 0; i < n; i++) {
        int steps; u64 h; int r;
        r = WITH_CODEC(N, L, ([&] { Game<C> g; int q = numpy_random_game<C>(g, geo, seeds[i], rng, steps, h); if (final_state) pack(g, final_state, i, geo); return q; })());
        if (result) result[i] = (int8_t)r;
        if (agent_steps) agent_steps[i] = steps;
        if (trace) trace[i] = h;
    }
    return EWN_OK;
}
int ewn_numpy_stream(const uint32_t* seeds, int64_t n, int K, uint32_t* raw, void*) {
    if (!seeds || !raw || n < 0 || K < 1) return EWN_EARG;
    MT19937 rng;
    for (int64_t i = 0; i < n; i++) { rng.seed(seeds[i]); for (int k = 0; k < K; k++) raw[i * K + k] = rng.next(); }
    return EWN_OK;
}
int ewn_numpy_draw(uint64_t* state, const uint32_t* raw, int32_t* cursor, int K, int64_t n, int N, int L, int roll, int pick_player,
                   int32_t* action, void*) {
    if (!state || !raw || !cursor || n < 0 || K < 1 || !config_ok(N, L) || pick_player < 0 || pick_player > 2 || (pick_player && !action)) return EWN_EARG;
    Geo geo = make_geo(N, L);
    for (int64_t i = 0; i < n; i++) {
        RawStream rs; rs.raw = raw + i * K; rs.K = K; rs.cur = cursor[i];
        int la = 0, di = 0, sh;
        WITH_CODEC(N, L, ([&] { Game<C> g; unpack(g, state, i, geo); if (roll) { g.dice = rs.randint(1, geo.cube_num + 1); pack(g, state, i, geo); }
            if (pick_player) { Legal l = pick_player == EWN_BOTTOM_RIGHT ? legal<C, SIDE_BR>(g, geo, movers<C>(g.br, g.dice)) : legal<C, SIDE_TL>(g, geo, movers<C>(g.tl, g.dice));
                int cnt = legal_count(l); if (cnt > 0) legal_pick(l, rs.randint(0, cnt), la, di, sh); } return 0; })());
        if (pick_player) { action[2 * i] = la; action[2 * i + 1] = di; }
        cursor[i] = rs.cur;
    }
    return EWN_OK;
}
int ewn_decode_onehot(const uint64_t* state, float* board, float* dice, int64_t n, int N, int L, void*) {
    if (!state || !board || n < 0 || !config_ok(N, L)) return EWN_EARG;
    Geo geo = make_geo(N, L);
    const int cn = geo.cube_num, row = N * N * cn;
    for (int64_t i = 0; i < n; i++) {
        float* b = board + i * row;
        for (int k = 0; k < row; k++) b[k] = 0.f;
        int dice_v = WITH_CODEC(N, L, ([&] { Game<C> g; unpack(g, state, i, geo);
            for (int k = 0; k < cn; k++) { int c = C::get(g.tl, k), d = C::get(g.br, k);
                if (c != (int)C::FM) b[(C::x_of(c) * N + C::y_of(c)) * cn + k] = 1.f; if (d != (int)C::FM) b[(C::x_of(d) * N + C::y_of(d)) * cn + k] = -1.f; }
            return g.dice; })());
        if (dice) { for (int k = 0; k < cn; k++) dice[i * cn + k] = 0.f; if (dice_v >= 1 && dice_v <= cn) dice[i * cn + dice_v - 1] = 1.f; }
    }
    return EWN_OK;
}
int ewn_flip(const uint64_t* state, uint64_t* out, int64_t n, int N, int L, void*) {
    if (!state || !out || n < 0 || !config_ok(N, L)) return EWN_EARG;
    Geo geo = make_geo(N, L);
    return WITH_CODEC(N, L, t_flip<C>(state, out, n, geo));
}
int ewn_counter_add(uint64_t* counter, uint64_t k, void*) { if (!counter) return EWN_EARG; *counter += k; return EWN_OK; }
int ewn_evaluate(const uint64_t* state, double* value, int64_t n, int N, int L, int heuristic, void*) {
    if (!state || !value || n < 0 || !config_ok(N, L) || heuristic < 0 || heuristic > 3) return EWN_EARG;
    Geo geo = make_geo(N, L);
    for (int64_t i = 0; i < n; i++) {
        value[i] = WITH_CODEC(N, L, ([&] { Game<C> g; unpack(g, state, i, geo); return evaluate<C>(g, geo, heuristic); })());
    }
    return EWN_OK;
}
size_t ewn_expectiminimax_workspace(int64_t, int) { return 0; }
int ewn_expectiminimax(const uint64_t* state, int64_t n, int N, int L, int depth, int heuristic, int pruned, double* value, int8_t* action,
                       void*, size_t, void*) {
    if (!state || !value || n < 0 || !config_ok(N, L) || heuristic < 0 || heuristic > 3) return EWN_EARG;
    if (depth < 1 || depth > EWN_MAX_DEPTH || L != 3) return EWN_ENOSUP;
    Geo geo = make_geo(N, L);
    return BY_CODEC(N, t_emm<P5>(state, n, geo, depth, heuristic, pruned, value, action), t_emm<P8>(state, n, geo, depth, heuristic, pruned, value, action));
}
int ewn_sim_winrate(const uint64_t* state, int64_t n, int N, int L, int nsim, uint64_t seed, uint32_t unit0, double* value, void*) {
    if (!state || !value || n < 0 || !config_ok(N, L) || nsim < 1 || nsim > SIM_MAX) return EWN_EARG;
    if (L != 3) return EWN_ENOSUP;
    Geo geo = make_geo(N, L);
    return BY_CODEC(N, t_sim_winrate<P5>(state, n, geo, nsim, seed, unit0, value), t_sim_winrate<P8>(state, n, geo, nsim, seed, unit0, value));
}
int ewn_expectiminimax_sim(const uint64_t* state, int64_t n, int N, int L, int depth, int pruned, int nsim, uint64_t seed, double* value,
                           int8_t* action, void*) {
    if (!state || !value || n < 0 || !config_ok(N, L) || nsim < 1 || nsim > SIM_MAX) return EWN_EARG;
    if (depth < 1 || depth > 4 || L != 3) return EWN_ENOSUP;
    Geo geo = make_geo(N, L);
    return BY_CODEC(N, t_emm_sim<P5>(state, n, geo, depth, pruned, nsim, seed, value, action), t_emm_sim<P8>(state, n, geo, depth, pruned, nsim, seed, value, action));
}
int ewn_playouts(const uint64_t* state, const uint32_t* root_ids, int64_t n, int N, int L, uint64_t seed, uint32_t first, uint32_t count,
                 uint32_t* wins, unsigned long long* plies, void*) {
    if (!state || !wins || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (L != 3) return EWN_ENOSUP;
    Geo geo = make_geo(N, L);
    return BY_CODEC(N, t_playouts<P5>(state, root_ids, n, geo, seed, first, count, wins, plies), t_playouts<P8>(state, root_ids, n, geo, seed, first, count, wins, plies));
}
size_t ewn_mc_workspace(int64_t) { return 0; }
int ewn_mc_root_wins(const uint64_t* state, const uint32_t* root_ids, int64_t n, int N, int L, uint64_t seed, uint32_t first, uint32_t count,
                     uint32_t* wins, int32_t* n_children, unsigned long long* plies, void*, size_t, void*) {
    if (!state || !wins || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (L != 3) return EWN_ENOSUP;
    Geo geo = make_geo(N, L);
    return BY_CODEC(N, t_mc<P5>(state, root_ids, n, geo, seed, first, count, wins, n_children, plies), t_mc<P8>(state, root_ids, n, geo, seed, first, count, wins, n_children, plies));
}
int ewn_mc_best_action(const uint64_t* state, const uint32_t* wins, int8_t* action, int64_t n, int N, int L, void*);
int ewn_mc_predict(const uint64_t* state, const uint32_t* root_ids, int64_t n, int N, int L, uint64_t seed, uint32_t first, uint32_t count,
                   uint32_t* wins, int8_t* action, int32_t* n_children, unsigned long long* plies, void* ws, size_t wsb, void* st) {
    if (!state || !wins || n < 0 || !config_ok(N, L)) return EWN_EARG;
    memset(wins, 0, sizeof(uint32_t) * 6 * (size_t)n);
    int rc = ewn_mc_root_wins(state, root_ids, n, N, L, seed, first, count, wins, n_children, plies, ws, wsb, st);
    if (rc || !action) return rc;
    return ewn_mc_best_action(state, wins, action, n, N, L, st);
}
size_t ewn_uct_workspace(int64_t) { return 0; }
int ewn_uct_search(const uint64_t* state, const uint32_t* root_ids, int64_t n, int N, int L, uint64_t seed, int rounds, uint32_t batch, double c,
                   uint32_t* wins, uint32_t* visits, int8_t* action, void*, size_t, void*) {
    if (!state || !wins || !visits || n < 0 || !config_ok(N, L) || rounds < 1 || batch < 1) return EWN_EARG;
    if (L != 3) return EWN_ENOSUP;
    Geo geo = make_geo(N, L);
    return BY_CODEC(N, t_uct<P5>(state, root_ids, n, geo, seed, rounds, batch, c, wins, visits, action),
                    t_uct<P8>(state, root_ids, n, geo, seed, rounds, batch, c, wins, visits, action));
}
int ewn_mc_best_action(const uint64_t* state, const uint32_t* wins, int8_t* action, int64_t n, int N, int L, void*) {
    if (!state || !wins || !action || n < 0 || !config_ok(N, L)) return EWN_EARG;
    Geo geo = make_geo(N, L);
    return BY_CODEC(N, t_mc_best<P5>(state, wins, action, n, geo), t_mc_best<P8>(state, wins, action, n, geo));
}
// peer exchange on the host: "handles" carry the pointer itself (one process plays all ranks)
int ewn_peer_alloc(size_t bytes, void** ptr, ewn_peer_handle* handle) {
    if (!ptr || !handle || bytes == 0) return EWN_EARG;
    void* p = calloc(1, bytes);
    if (!p) return EWN_ECUDA;
    memset(handle, 0, sizeof(*handle));
    memcpy(handle->bytes, &p, sizeof(p));
    *ptr = p;
    return EWN_OK;
}
int ewn_peer_open(const ewn_peer_handle* handle, void** ptr) {
    if (!handle || !ptr) return EWN_EARG;
    memcpy(ptr, handle->bytes, sizeof(*ptr));
    return EWN_OK;
}
int ewn_peer_close(void* ptr) { return ptr ? EWN_OK : EWN_EARG; }
int ewn_peer_free(void* ptr) { if (!ptr) return EWN_EARG; free(ptr); return EWN_OK; }
int ewn_mc_best_action_peers(const uint64_t* state, const uint32_t* const* wins_peers, uint32_t* const* flag_peers, int rank, int world,
                             uint32_t epoch, uint32_t, uint32_t* wins_total, int8_t* action, int64_t n, int N, int L, void*) {
    if (!state || !wins_peers || !flag_peers || !action || n < 0 || !config_ok(N, L)) return EWN_EARG;
    if (world < 1 || world > EWN_MAX_PEERS || rank < 0 || rank >= world || epoch == 0) return EWN_EARG;
    for (int r = 0; r < world; r++) if (!wins_peers[r] || !flag_peers[r]) return EWN_EARG;
    for (int r = 0; r < world; r++) flag_peers[r][rank] = epoch;
    for (int r = 0; r < world; r++)                                   // a sequential host cannot wait: an unpublished peer is the time-out case
        if ((int)(flag_peers[rank][r] - epoch) < 0) { flag_peers[rank][EWN_PEER_ERR_WORD] = epoch; break; }
    uint32_t* sum = (uint32_t*)calloc((size_t)(n ? n : 1) * 6, sizeof(uint32_t));
    for (int r = 0; r < world; r++) for (int64_t k = 0; k < n * 6; k++) sum[k] += wins_peers[r][k];
    if (wins_total) memcpy(wins_total, sum, (size_t)n * 6 * sizeof(uint32_t));
    Geo geo = make_geo(N, L);
    int rc = BY_CODEC(N, t_mc_best<P5>(state, sum, action, n, geo), t_mc_best<P8>(state, sum, action, n, geo));
    free(sum);
    return rc;
}
}  // extern "C"
