/*
 * ewn_oracle.c -- TEST INFRASTRUCTURE ONLY (parity oracle + timed CPU baseline).
 *
 * A plain-C, scalar, board-centric restatement of the EinStein-wuerfelt-nicht rules and search of the
 * reference jchen8tw/ewn-gym (pure Python).  Every function cites the reference file:line it follows.
 * It deliberately keeps the reference's own data model (an N x N signed board plus a per-cube piece
 * list with a "masked" flag) so that it is an independent check of the GPU engine, which uses a
 * bit-packed piece list and SWAR arithmetic instead.
 *
 * Parity pinning: this file is validated against golden vectors produced by running the UNMODIFIED
 * reference in the build container (tests/golden/make_golden.py -> tests/golden/ *.npz) and, when
 * /root/reference is present, against the live reference (tests/test_oracle_vs_reference.py: marker `reference`).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library.  The product package (ewn_gym_b200) never does.
 *
 * Random numbers: the reference draws from numpy's global MT19937 (envs/ewn.py:91,
 * classical_policies/random_policy.py:13) and CPython's `random` (classical_policies/mcts.py:29-32).
 * The engine uses counter-based Philox4x32-10 instead; the draw schedule restated here (orc_*_philox)
 * is the one documented in DESIGN.md "RNG schedule", so that engine and oracle agree bit-for-bit.
 * Replay entry points (orc_step_replay, orc_ply) take dice and actions from the caller so that the
 * reference's own trajectories can be reproduced exactly.
 */
#include <stdint.h>
#include <string.h>
#include <math.h>
#include <stdlib.h>

#define ORC_MAXN 9
#define ORC_MAXC 28 /* cube_num = L(L+1)/2, cube_layer < board_size - 1 <= 8 (envs/ewn.py:47,51): up to 28 cubes per side on 9x9 */

#define TOP_LEFT 1     /* constants/player.py:5 */
#define BOTTOM_RIGHT 2 /* constants/player.py:6 */
#define CHANCE 3       /* constants/player.py:8 */

typedef struct {
    int N, L, cube_num;
    int dice;                              /* envs/ewn.py:59 */
    int player;                            /* envs/ewn.py:71 (current_player) */
    int16_t board[ORC_MAXN * ORC_MAXN];    /* envs/ewn.py:49-50, row-major [x][y] */
    int8_t px[2][ORC_MAXC], py[2][ORC_MAXC]; /* envs/ewn.py:57-58; [0]=TL cube k at k-1, [1]=BR cube k; */
    int8_t alive[2][ORC_MAXC];             /* !mask */
} orc_game;

/* an empty game of this configuration: only what the configuration uses is cleared (the piece-list arrays are sized for 28 cubes,
 * the common six-cube games should not pay for that) */
static void g_clear(orc_game *g, int N, int L) {
    g->N = N; g->L = L; g->cube_num = L * (L + 1) / 2;
    memset(g->board, 0, sizeof(int16_t) * (size_t)(N * N));
    for (int s = 0; s < 2; s++) { memset(g->alive[s], 0, (size_t)g->cube_num); memset(g->px[s], 0, (size_t)g->cube_num); memset(g->py[s], 0, (size_t)g->cube_num); }
}

static inline int side_of(int player) { return player == TOP_LEFT ? 0 : 1; }
static inline int opponent_of(int player) { return player == TOP_LEFT ? BOTTOM_RIGHT : TOP_LEFT; } /* constants/player.py:10-17 */

/* ------------------------------------------------------------------------------------------------
 * setup / restore
 * ---------------------------------------------------------------------------------------------- */

/* envs/ewn.py:94-107 setup_game (without the trailing roll_dice) */
static void g_setup(orc_game *g, int N, int L) {
    g_clear(g, N, L);
    g->dice = 1; g->player = TOP_LEFT;
    int cnt = 1;
    for (int i = 1; i <= L; i++) {
        for (int j = 0; j < i; j++) {
            int x = j, y = i - j - 1;
            g->board[x * N + y] = (int16_t)cnt;
            g->px[0][cnt - 1] = (int8_t)x; g->py[0][cnt - 1] = (int8_t)y; g->alive[0][cnt - 1] = 1;
            int bx = N - 1 - j, by = N - i + j;
            g->board[bx * N + by] = (int16_t)(-cnt);
            g->px[1][cnt - 1] = (int8_t)bx; g->py[1][cnt - 1] = (int8_t)by; g->alive[1][cnt - 1] = 1;
            cnt++;
        }
    }
}

/* classical_policies/minimax.py:75-87 restore_env_with_obs: piece list rebuilt by scanning the board */
static void g_from_board(orc_game *g, int N, int L, const int16_t *board, int dice, int player) {
    g_clear(g, N, L);
    g->dice = dice; g->player = player;
    for (int i = 0; i < N; i++)
        for (int j = 0; j < N; j++) {
            int c = board[i * N + j];
            g->board[i * N + j] = (int16_t)c;
            if (c > 0 && c <= ORC_MAXC) { g->px[0][c - 1] = (int8_t)i; g->py[0][c - 1] = (int8_t)j; g->alive[0][c - 1] = 1; }
            else if (c < 0 && -c <= ORC_MAXC) { g->px[1][-c - 1] = (int8_t)i; g->py[1][-c - 1] = (int8_t)j; g->alive[1][-c - 1] = 1; }
        }
}

static void g_to_board(const orc_game *g, int16_t *board) {
    memcpy(board, g->board, sizeof(int16_t) * g->N * g->N);
}

/* ------------------------------------------------------------------------------------------------
 * rules
 * ---------------------------------------------------------------------------------------------- */

/* envs/ewn.py:131-142 check_win */
static int g_check_win(const orc_game *g) {
    int N = g->N;
    if (g->board[0] < 0 || g->board[N * N - 1] > 0) return 1;
    int any_pos = 0, any_neg = 0;
    for (int i = 0; i < N * N; i++) { any_pos |= g->board[i] > 0; any_neg |= g->board[i] < 0; }
    if (!any_pos) return 1;
    if (!any_neg) return 1;
    return 0;
}

/* envs/ewn.py:144-176 find_near_cube; k is the 1-based cube number, returns 1-based number or 0 (None).
 * "larger" walks k+1..cube_num, "smaller" walks k-1..1, for both colours (the BR loops :158-162,:171-174
 * walk negative indices, which is the same order in |k|). */
static int g_find_near(const orc_game *g, int k, int larger, int player) {
    int s = side_of(player);
    if (larger) { for (int i = k + 1; i <= g->cube_num; i++) if (g->alive[s][i - 1]) return i; }
    else        { for (int i = k - 1; i >= 1; i--)           if (g->alive[s][i - 1]) return i; }
    return 0;
}

/* envs/ewn.py:178-215 find_cube_to_move */
static int g_find_cube_to_move(const orc_game *g, int larger, int player) {
    int s = side_of(player), k = g->dice;
    if (k >= 1 && k <= g->cube_num && g->alive[s][k - 1]) return k;          /* :202-203 */
    int near = g_find_near(g, k, larger, player);                            /* :206-209 */
    if (near) return near;
    near = g_find_near(g, k, !larger, player);                               /* :210-215 flag ignored */
    return near; /* 0 only if the side has no cube at all (reference asserts :214) */
}

/* envs/ewn.py:301-319 update_position */
static void g_update_pos(int *x, int *y, int dir, int cube) {
    if (cube > 0) { if (dir == 0) (*y)++; else if (dir == 1) (*x)++; else if (dir == 2) { (*x)++; (*y)++; } }
    else          { if (dir == 0) (*y)--; else if (dir == 1) (*x)--; else if (dir == 2) { (*x)--; (*y)--; } }
}

/* envs/ewn.py:217-263 execute_move (dry_run omitted); any occupant of the target is captured (:254-258) */
static int g_execute_move(orc_game *g, int player, int k, int dir) {
    int s = side_of(player), N = g->N;
    int ox = g->px[s][k - 1], oy = g->py[s][k - 1];
    int cube = g->board[ox * N + oy];
    int x = ox, y = oy;
    g_update_pos(&x, &y, dir, cube);
    if (0 <= x && x < N && 0 <= y && y < N) {
        g->board[ox * N + oy] = 0;
        int t = g->board[x * N + y];
        if (t != 0) { if (t > 0) g->alive[0][t - 1] = 0; else g->alive[1][-t - 1] = 0; }
        g->board[x * N + y] = (int16_t)cube;
        g->px[s][k - 1] = (int8_t)x; g->py[s][k - 1] = (int8_t)y;
        return 1;
    }
    return 0; /* :262-263 state untouched */
}

/* envs/ewn.py:325-336 get_cube_legal_directions; returns bit mask of dirs */
static int g_cube_dirs(const orc_game *g, int player, int k) {
    int s = side_of(player), N = g->N, m = 0;
    int cube = (s == 0) ? k : -k;
    for (int d = 0; d < 3; d++) {
        int x = g->px[s][k - 1], y = g->py[s][k - 1];
        g_update_pos(&x, &y, d, cube);
        if (0 <= x && x < N && 0 <= y && y < N) m |= 1 << d;
    }
    return m;
}

/* envs/ewn.py:338-375 get_legal_actions; list[i] = {larger, dir}; returns length */
static int g_legal_actions(const orc_game *g, int player, int list[6][2]) {
    int s = side_of(player), k = g->dice, n = 0;
    if (k >= 1 && k <= g->cube_num && g->alive[s][k - 1]) {              /* :347-353 */
        int m = g_cube_dirs(g, player, k);
        for (int d = 0; d < 3; d++) if (m >> d & 1) { list[n][0] = 0; list[n][1] = d; n++; }
        return n;
    }
    int near = g_find_near(g, k, 1, player);                             /* :356-363 */
    if (near) { int m = g_cube_dirs(g, player, near); for (int d = 0; d < 3; d++) if (m >> d & 1) { list[n][0] = 1; list[n][1] = d; n++; } }
    near = g_find_near(g, k, 0, player);                                 /* :366-373 */
    if (near) { int m = g_cube_dirs(g, player, near); for (int d = 0; d < 3; d++) if (m >> d & 1) { list[n][0] = 0; list[n][1] = d; n++; } }
    return n;
}

/* envs/ewn.py:377-412 make_simulated_action (no history: callers copy the struct instead of undoing) */
static void g_make_action(orc_game *g, int player, int larger, int dir) {
    int k = g_find_cube_to_move(g, larger, player);
    if (!k) return;
    g_execute_move(g, player, k, dir);
}

/* ------------------------------------------------------------------------------------------------
 * leaf evaluation  (envs/minimax_ewn.py:29-213; agent_player is always TOP_LEFT, minimax_ewn.py:17-19)
 * ---------------------------------------------------------------------------------------------- */
enum { H_HYBRID = 0, H_MIN_DIST = 1, H_TWO_MIN_DIST = 2, H_ATTK = 3 };

static double g_evaluate(const orc_game *g, int heuristic) {
    int N = g->N;
    int any_pos = 0, any_neg = 0;
    for (int i = 0; i < N * N; i++) { any_pos |= g->board[i] > 0; any_neg |= g->board[i] < 0; }
    /* minimax_ewn.py:41-47 (and the identical preambles :91-97, :136-142, :183-189) */
    if (g->board[N * N - 1] > 0 || !any_neg) return 10.0;
    if (g->board[0] < 0 || !any_pos) return -10.0;

    int md_p = 1 << 20, md_n = 1 << 20, c_p = 0, c_n = 0;
    int dp[ORC_MAXC * 2], dn[ORC_MAXC * 2], np_ = 0, nn_ = 0;
    for (int i = 0; i < N; i++)
        for (int j = 0; j < N; j++) {
            int v = g->board[i * N + j];
            int a = N - 1 - i, b = N - 1 - j;
            int dist = a > b ? a : b; /* both colours measure to the bottom-right corner: minimax_ewn.py:67-76 */
            if (v > 0) { if (dist < md_p) md_p = dist; c_p++; dp[np_++] = dist; }
            else if (v < 0) { if (dist < md_n) md_n = dist; c_n++; dn[nn_++] = dist; }
        }
    if (heuristic == H_HYBRID) { /* minimax_ewn.py:79-86 */
        double score = 0;
        score += (double)(N - md_p) * (1.0 / (double)c_p);
        score -= (double)(N - md_n) * (1.0 / (double)c_n);
        return score;
    }
    if (heuristic == H_MIN_DIST) { /* minimax_ewn.py:126-131 */
        return (double)((N - md_p) - (N - md_n));
    }
    if (heuristic == H_TWO_MIN_DIST) { /* minimax_ewn.py:168-178 */
        for (int a = 0; a < np_; a++) for (int b = a + 1; b < np_; b++) if (dp[b] < dp[a]) { int t = dp[a]; dp[a] = dp[b]; dp[b] = t; }
        for (int a = 0; a < nn_; a++) for (int b = a + 1; b < nn_; b++) if (dn[b] < dn[a]) { int t = dn[a]; dn[a] = dn[b]; dn[b] = t; }
        int sp = 0, sn = 0;
        for (int a = 0; a < np_ && a < 2; a++) sp += dp[a];
        for (int a = 0; a < nn_ && a < 2; a++) sn += dn[a];
        return (double)(sn - sp);
    }
    /* H_ATTK minimax_ewn.py:212-213 */
    return (double)(-(c_p + c_n));
}

/* ------------------------------------------------------------------------------------------------
 * expectiminimax  (classical_policies/minimax.py:19-73), including the alpha/beta hand-through at
 * chance nodes (:34-35,:70-71) and the /6 sequential expectation (:65-73).
 * pruned == 0 disables the two `break`s (:42,:61) -- textbook expectiminimax, not in the reference.
 * ---------------------------------------------------------------------------------------------- */
typedef struct { long leaves; long nodes; } orc_stats;

static double g_emm(const orc_game *g, int depth, int player, int parent, double alpha, double beta,
                    int heuristic, int pruned, int *best_larger, int *best_dir, orc_stats *st) {
    if (st) st->nodes++;
    if (g_check_win(g) || depth == 0) { /* :22-23 */
        if (st) st->leaves++;
        if (best_larger) { *best_larger = -1; *best_dir = -1; }
        return g_evaluate(g, heuristic);
    }
    if (player == TOP_LEFT) { /* :26-43 maximizing (agent_player == TOP_LEFT always) */
        double best_val = -INFINITY; int bl = -1, bd = -1;
        int list[6][2]; int n = g_legal_actions(g, player, list);
        for (int i = 0; i < n; i++) {
            orc_game c = *g;                         /* set_dice_roll(curr) + make + ... + undo */
            g_make_action(&c, player, list[i][0], list[i][1]);
            double val = g_emm(&c, depth - 1, CHANCE, player, alpha, beta, heuristic, pruned, 0, 0, st);
            if (val > best_val) { best_val = val; bl = list[i][0]; bd = list[i][1]; }
            alpha = alpha > best_val ? alpha : best_val;   /* max(alpha, best_val) */
            if (pruned && beta <= alpha) break;
        }
        if (best_larger) { *best_larger = bl; *best_dir = bd; }
        return best_val;
    } else if (player == BOTTOM_RIGHT) { /* :45-62 minimizing */
        double worst_val = INFINITY;
        int list[6][2]; int n = g_legal_actions(g, player, list);
        for (int i = 0; i < n; i++) {
            orc_game c = *g;
            g_make_action(&c, player, list[i][0], list[i][1]);
            double val = g_emm(&c, depth - 1, CHANCE, player, alpha, beta, heuristic, pruned, 0, 0, st);
            if (val < worst_val) worst_val = val;
            beta = beta < worst_val ? beta : worst_val;    /* min(beta, worst_val) */
            if (pruned && beta <= alpha) break;
        }
        if (best_larger) { *best_larger = -1; *best_dir = -1; } /* :47,:58 worst_action stays None */
        return worst_val;
    } else { /* :64-73 chance */
        double expected = 0.0;
        int next_player = (parent == BOTTOM_RIGHT) ? TOP_LEFT : BOTTOM_RIGHT;
        for (int d = 1; d <= 6; d++) {
            orc_game c = *g; c.dice = d;
            double val = g_emm(&c, depth - 1, next_player, CHANCE, alpha, beta, heuristic, pruned, 0, 0, st);
            expected += val / 6;
        }
        if (best_larger) { *best_larger = -1; *best_dir = -1; }
        return expected;
    }
}

/* ------------------------------------------------------------------------------------------------
 * Philox4x32-10 (Salmon et al., SC'11) and the engine's draw schedule (DESIGN.md "RNG schedule")
 * ---------------------------------------------------------------------------------------------- */
static void philox4x32_10(const uint32_t ctr_in[4], const uint32_t key_in[2], uint32_t out[4]) {
    uint32_t c0 = ctr_in[0], c1 = ctr_in[1], c2 = ctr_in[2], c3 = ctr_in[3];
    uint32_t k0 = key_in[0], k1 = key_in[1];
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static inline uint32_t bounded(uint32_t r, uint32_t k) { return (uint32_t)(((uint64_t)r * (uint64_t)k) >> 32); }

#define TAG_STEP 0u
#define TAG_RESET 1u
#define TAG_PLAYOUT 2u

static void env_block(uint64_t seed, uint64_t env_id, uint64_t step, uint32_t tag, uint32_t out[4]) {
    uint32_t ctr[4] = { (uint32_t)env_id, (uint32_t)(env_id >> 32), (uint32_t)step,
                        ((uint32_t)(step >> 32) & 0x0FFFFFFFu) | (tag << 28) };
    uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
    philox4x32_10(ctr, key, out);
}

static void playout_block(uint64_t seed, uint32_t root_id, uint32_t child, uint32_t playout, uint32_t j, uint32_t out[4]) {
    uint32_t ctr[4] = { playout, root_id, j, (TAG_PLAYOUT << 28) | child };
    uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
    philox4x32_10(ctr, key, out);
}

/* One random playout, classical_policies/mcts.py:24-41: BOTTOM_RIGHT moves first (:26), 6-face die (:29),
 * uniform over the legal *list* (:31-33), TL win test (:39-41).  Returns 1 if TOP_LEFT won. */
static int g_playout(orc_game *g, uint64_t seed, uint32_t root_id, uint32_t child, uint32_t playout, int *plies_out) {
    int player = BOTTOM_RIGHT; uint32_t p = 0; uint32_t blk[4] = {0, 0, 0, 0};
    while (!g_check_win(g)) {
        if ((p & 1u) == 0) playout_block(seed, root_id, child, playout, p >> 1, blk);
        uint32_t rd = blk[(p & 1u) * 2], ra = blk[(p & 1u) * 2 + 1];
        g->dice = 1 + (int)bounded(rd, 6);
        int list[6][2]; int n = g_legal_actions(g, player, list);
        int idx = (int)bounded(ra, (uint32_t)n);
        g_make_action(g, player, list[idx][0], list[idx][1]);
        player = opponent_of(player);
        p++;
    }
    if (plies_out) *plies_out = (int)p;
    int N = g->N, any_neg = 0;
    for (int i = 0; i < N * N; i++) any_neg |= g->board[i] < 0;
    return (g->board[N * N - 1] > 0 || !any_neg) ? 1 : 0;
}

/* ================================================================================================
 * exported batch API (SoA numpy arrays via ctypes).  board: int16[n][N*N]
 * ============================================================================================== */
#define API __attribute__((visibility("default")))

API int orc_version(void) { return 1; }

API void orc_philox(const uint32_t *ctr, const uint32_t *key, uint32_t *out) { philox4x32_10(ctr, key, out); }

API void orc_init_boards(int16_t *board, long n, int N, int L) {
    orc_game g; g_setup(&g, N, L);
    for (long i = 0; i < n; i++) g_to_board(&g, board + i * N * N);
}

API void orc_check_win(const int16_t *board, long n, int N, int L, uint8_t *win) {
    for (long i = 0; i < n; i++) { orc_game g; g_from_board(&g, N, L, board + i * N * N, 1, TOP_LEFT); win[i] = (uint8_t)g_check_win(&g); }
}

/* mask bit = larger*3+dir (same convention as the reference's EWNLogic.py:328 action id);
 * list: the ordered list of envs/ewn.py:338-375, padded with -1. */
API void orc_legal(const int16_t *board, const int32_t *dice, const int32_t *player, long n, int N, int L,
                   uint8_t *mask, int8_t *list_out, int32_t *count) {
    for (long i = 0; i < n; i++) {
        orc_game g; g_from_board(&g, N, L, board + i * N * N, dice[i], player[i]);
        int list[6][2]; int c = g_legal_actions(&g, player[i], list);
        uint8_t m = 0;
        for (int j = 0; j < 6; j++) {
            if (j < c) { m |= (uint8_t)(1u << (list[j][0] * 3 + list[j][1])); if (list_out) { list_out[(i * 6 + j) * 2] = (int8_t)list[j][0]; list_out[(i * 6 + j) * 2 + 1] = (int8_t)list[j][1]; } }
            else if (list_out) { list_out[(i * 6 + j) * 2] = -1; list_out[(i * 6 + j) * 2 + 1] = -1; }
        }
        if (mask) mask[i] = m;
        if (count) count[i] = c;
    }
}

/* One ply for `player` with the current dice: find_cube_to_move + execute_move + check_win
 * (envs/ewn.py:439-451).  Does not switch player or roll. */
API void orc_ply(int16_t *board, const int32_t *dice, const int32_t *player, const int32_t *larger, const int32_t *dir,
                 long n, int N, int L, uint8_t *valid, uint8_t *win, int32_t *moved_cube) {
    for (long i = 0; i < n; i++) {
        orc_game g; g_from_board(&g, N, L, board + i * N * N, dice[i], player[i]);
        int k = g_find_cube_to_move(&g, larger[i] == 1, player[i]);
        int v = k ? g_execute_move(&g, player[i], k, dir[i]) : 0;
        g_to_board(&g, board + i * N * N);
        if (valid) valid[i] = (uint8_t)v;
        if (win) win[i] = (uint8_t)g_check_win(&g);
        if (moved_cube) moved_cube[i] = k;
    }
}

/* Result codes shared with the engine (include/ewn_b200.h EWN_CODE_*). */
enum { CODE_ONGOING = 0, CODE_AGENT_ILLEGAL = 1, CODE_AGENT_WON = 2, CODE_OPP_ILLEGAL = 3, CODE_AGENT_LOST = 4, CODE_ILLEGAL_TOLERATED = 5 };

/* envs/ewn.py:436-486 step(), agent = TOP_LEFT, with the die values and the opponent's answer injected.
 * dice_opp = the roll at :458, dice_next = the roll at :483.  The dice value left in the state follows the
 * reference: a terminal step keeps the die that was current when the game ended. */
static int g_step_replay(orc_game *g, int a_larger, int a_dir, int dice_opp, int o_larger, int o_dir, int dice_next,
                         double R, double *reward, int *term, int *trunc) {
    int k = g_find_cube_to_move(g, a_larger == 1, g->player);
    int valid = g_execute_move(g, g->player, k, a_dir);
    if (!valid) { *reward = -R; *term = 1; *trunc = 1; return CODE_AGENT_ILLEGAL; }   /* :445-448 */
    if (g_check_win(g)) { *reward = R; *term = 1; *trunc = 0; return CODE_AGENT_WON; }   /* :451-454 */
    g->player = opponent_of(g->player); g->dice = dice_opp;                            /* :457-458 */
    k = g_find_cube_to_move(g, o_larger == 1, g->player);                              /* :464 */
    valid = g_execute_move(g, g->player, k, o_dir);                                     /* :467 */
    if (!valid) { *reward = 0; *term = 1; *trunc = 1; return CODE_OPP_ILLEGAL; }        /* :470-473 */
    if (g_check_win(g)) { *reward = -R; *term = 1; *trunc = 0; return CODE_AGENT_LOST; } /* :475-478 */
    g->player = opponent_of(g->player); g->dice = dice_next;                           /* :481-483 */
    *reward = 0; *term = 0; *trunc = 0; return CODE_ONGOING;                            /* :485-486 */
}

API void orc_step_replay(int16_t *board, int32_t *dice, int32_t *player, const int32_t *action /*[n,2]*/,
                         const int32_t *dice_opp, const int32_t *opp_action /*[n,2]*/, const int32_t *dice_next,
                         long n, int N, int L, double R, double *reward, uint8_t *term, uint8_t *trunc, int32_t *code) {
    for (long i = 0; i < n; i++) {
        orc_game g; g_from_board(&g, N, L, board + i * N * N, dice[i], player[i]);
        double r; int te, tr;
        int c = g_step_replay(&g, action[2 * i], action[2 * i + 1], dice_opp[i], opp_action[2 * i], opp_action[2 * i + 1], dice_next[i], R, &r, &te, &tr);
        g_to_board(&g, board + i * N * N); dice[i] = g.dice; player[i] = g.player;
        reward[i] = r; term[i] = (uint8_t)te; trunc[i] = (uint8_t)tr; if (code) code[i] = c;
    }
}

API void orc_eval(const int16_t *board, long n, int N, int L, int heuristic, double *out) {
    for (long i = 0; i < n; i++) { orc_game g; g_from_board(&g, N, L, board + i * N * N, 1, TOP_LEFT); out[i] = g_evaluate(&g, heuristic); }
}

/* classical_policies/minimax.py:89-93 predict(): root = TOP_LEFT to move with obs dice. action = -1,-1 for None. */
API void orc_expectiminimax(const int16_t *board, const int32_t *dice, long n, int N, int L, int depth, int heuristic, int pruned,
                            double *value, int8_t *action /*[n,2]*/, int64_t *leaves) {
    for (long i = 0; i < n; i++) {
        orc_game g; g_from_board(&g, N, L, board + i * N * N, dice[i], TOP_LEFT);
        int bl, bd; orc_stats st = {0, 0};
        value[i] = g_emm(&g, depth, TOP_LEFT, 0, -INFINITY, INFINITY, heuristic, pruned, &bl, &bd, &st);
        action[2 * i] = (int8_t)bl; action[2 * i + 1] = (int8_t)bd;
        if (leaves) leaves[i] = st.leaves;
    }
}

/* Flat Monte-Carlo over root children, classical_policies/mcts.py:47-69 with the engine's Philox schedule.
 * For every root i and every legal root action (list position c): wins[i][c] = number of playouts with index in
 * [first, first+count) that TOP_LEFT wins, after TL plays the action.  n_children[i] = len(legal list). */
API void orc_mc_root_wins(const int16_t *board, const int32_t *dice, const uint32_t *root_ids, long n, int N, int L,
                          uint64_t seed, uint32_t first, uint32_t count, uint32_t *wins /*[n,6]*/, int32_t *n_children,
                          int64_t *total_plies) {
    int64_t plies_sum = 0;
    for (long i = 0; i < n; i++) {
        orc_game g; g_from_board(&g, N, L, board + i * N * N, dice[i], TOP_LEFT);
        int list[6][2]; int c = g_legal_actions(&g, TOP_LEFT, list);
        n_children[i] = c;
        for (int j = 0; j < 6; j++) wins[i * 6 + j] = 0;
        for (int j = 0; j < c; j++) {
            orc_game child = g; g_make_action(&child, TOP_LEFT, list[j][0], list[j][1]);
            uint32_t w = 0;
            for (uint32_t p = first; p < first + count; p++) {
                orc_game sim = child; int plies;
                w += (uint32_t)g_playout(&sim, seed, root_ids[i], (uint32_t)j, p, &plies);
                plies_sum += plies;
            }
            wins[i * 6 + j] = w;
        }
    }
    if (total_plies) *total_plies = plies_sum;
}

/* Playouts from given states (no root move): state i, playout indices [first, first+count), BR to move first.
 * Philox child field = 7 (never a list position). */
API void orc_playouts(const int16_t *board, const uint32_t *root_ids, long n, int N, int L, uint64_t seed,
                      uint32_t first, uint32_t count, uint32_t *wins, int64_t *total_plies) {
    int64_t plies_sum = 0;
    for (long i = 0; i < n; i++) {
        orc_game g; g_from_board(&g, N, L, board + i * N * N, 1, BOTTOM_RIGHT);
        uint32_t w = 0;
        for (uint32_t p = first; p < first + count; p++) { orc_game sim = g; int plies; w += (uint32_t)g_playout(&sim, seed, root_ids[i], 7u, p, &plies); plies_sum += plies; }
        wins[i] = w;
    }
    if (total_plies) *total_plies = plies_sum;
}

/* ------------------------------------------------------------------------------------------------
 * Free-running batched env step with the engine's Philox schedule (the CPU twin of kernel K1).
 *   blk = philox(seed; env_id, step, TAG_STEP):  blk[0] die for the opponent (:458), blk[1] opponent's index
 *   into its legal list (random_policy.py:12-14), blk[2] die for the next agent turn (:483), blk[3] the agent's
 *   index into its legal list when agent_random != 0.
 *   explicit reset (orc_reset_philox): die = 1 + bounded(philox(seed; env_id, step, TAG_RESET)[0], cube_num)  (:108)
 *   auto-reset inside a step: die = 1 + bounded(blk[2], cube_num) (blk[2] is unused by a step that ends the game)
 * shaped != 0 follows envs/training_ewn.py:40-99 (tolerance and prev_score persist across resets).
 * autoreset != 0 follows the SB3 VecEnv contract: terminal board/dice are saved, the env is re-initialised.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    int32_t N, L;
    double reward;             /* goal reward (envs/ewn.py:38 / training_ewn.py:24) */
    int32_t agent_random;      /* 1: agent = uniform over legal list from blk[3] */
    int32_t autoreset;
    int32_t shaped;            /* MiniMaxHeuristicEnv */
    double illegal_move_reward;
    uint64_t seed;
    uint64_t step;             /* global step index */
    uint64_t env_id0;          /* global id of env 0 of this batch */
} orc_step_cfg;

API void orc_reset_philox(int16_t *board, int32_t *dice, long n, int N, int L, uint64_t seed, uint64_t step, uint64_t env_id0) {
    orc_game g0; g_setup(&g0, N, L);
    for (long i = 0; i < n; i++) {
        uint32_t blk[4]; env_block(seed, env_id0 + (uint64_t)i, step, TAG_RESET, blk);
        g_to_board(&g0, board + i * N * N);
        dice[i] = 1 + (int)bounded(blk[0], (uint32_t)g0.cube_num);
    }
}

API void orc_step_philox(int16_t *board, int32_t *dice, const int32_t *action /*[n,2] or NULL*/, long n, const orc_step_cfg *cfg,
                         double *reward, uint8_t *term, uint8_t *trunc, int32_t *code,
                         int16_t *term_board, int32_t *term_dice, int32_t *tolerance, double *prev_score,
                         int8_t *agent_action_out) {
    int N = cfg->N, L = cfg->L;
    for (long i = 0; i < n; i++) {
        orc_game g; g_from_board(&g, N, L, board + i * N * N, dice[i], TOP_LEFT);
        uint64_t env_id = cfg->env_id0 + (uint64_t)i;
        uint32_t blk[4]; env_block(cfg->seed, env_id, cfg->step, TAG_STEP, blk);
        int al, ad;
        if (cfg->agent_random) {
            int list[6][2]; int c = g_legal_actions(&g, TOP_LEFT, list);
            int idx = (int)bounded(blk[3], (uint32_t)c); al = list[idx][0]; ad = list[idx][1];
        } else { al = action[2 * i]; ad = action[2 * i + 1]; }
        if (agent_action_out) { agent_action_out[2 * i] = (int8_t)al; agent_action_out[2 * i + 1] = (int8_t)ad; }
        double r = 0; int te = 0, tr = 0, c = CODE_ONGOING;
        int k = g_find_cube_to_move(&g, al == 1, TOP_LEFT);
        int valid = g_execute_move(&g, TOP_LEFT, k, ad);
        if (!valid) {
            if (cfg->shaped) { /* training_ewn.py:48-56 */
                tolerance[i] -= 1;
                if (tolerance[i] <= 0) { r = -cfg->reward; te = 1; tr = 1; c = CODE_AGENT_ILLEGAL; }
                else { r = cfg->illegal_move_reward; te = 0; tr = 0; c = CODE_ILLEGAL_TOLERATED; }
            } else { r = -cfg->reward; te = 1; tr = 1; c = CODE_AGENT_ILLEGAL; }
        } else if (g_check_win(&g)) { r = cfg->reward; te = 1; tr = 0; c = CODE_AGENT_WON; }
        else {
            g.player = BOTTOM_RIGHT; g.dice = 1 + (int)bounded(blk[0], (uint32_t)g.cube_num);
            int list[6][2]; int cn = g_legal_actions(&g, BOTTOM_RIGHT, list);
            int idx = (int)bounded(blk[1], (uint32_t)cn);
            k = g_find_cube_to_move(&g, list[idx][0] == 1, BOTTOM_RIGHT);
            valid = g_execute_move(&g, BOTTOM_RIGHT, k, list[idx][1]);
            if (!valid) { r = 0; te = 1; tr = 1; c = CODE_OPP_ILLEGAL; }
            else if (g_check_win(&g)) { r = -cfg->reward; te = 1; tr = 0; c = CODE_AGENT_LOST; }
            else {
                g.player = TOP_LEFT; g.dice = 1 + (int)bounded(blk[2], (uint32_t)g.cube_num);
                if (cfg->shaped) { /* training_ewn.py:94-96 */
                    double s = g_evaluate(&g, H_HYBRID); r = s - prev_score[i]; prev_score[i] = s;
                }
            }
        }
        reward[i] = r; term[i] = (uint8_t)te; trunc[i] = (uint8_t)tr; if (code) code[i] = c;
        if (te && cfg->autoreset) {
            if (term_board) g_to_board(&g, term_board + i * N * N);
            if (term_dice) term_dice[i] = g.dice;
            /* the episode ended, so blk[2] (the "next agent turn" die) was not consumed: it becomes the first die of
             * the new episode (envs/ewn.py:108) -- one Philox block per env step, never two */
            g_setup(&g, N, L); g.dice = 1 + (int)bounded(blk[2], (uint32_t)g.cube_num);
        }
        g_to_board(&g, board + i * N * N); dice[i] = g.dice;
    }
}

/* ------------------------------------------------------------------------------------------------
 * CPU baseline driver: the eval_random.py:17-51 evaluation() loop (random agent vs random opponent), one
 * game per seed, Philox dice in place of numpy's MT19937.  Returns wins; counts agent steps and plies.
 * Game g: env_id = g, steps t = 0,1,2,... ; first die from the TAG_RESET block at step 0.
 * ---------------------------------------------------------------------------------------------- */
API long orc_random_games_range(long g_lo, long g_hi, int N, int L, uint64_t seed, int64_t *steps_out, int64_t *plies_out) {
    long wins = 0; int64_t steps = 0, plies = 0;
    for (long gi = g_lo; gi < g_hi; gi++) {
        orc_game g; g_setup(&g, N, L);
        uint32_t blk[4]; env_block(seed, (uint64_t)gi, 0, TAG_RESET, blk);
        g.dice = 1 + (int)bounded(blk[0], (uint32_t)g.cube_num);
        for (uint64_t t = 0;; t++) {
            env_block(seed, (uint64_t)gi, t, TAG_STEP, blk);
            int list[6][2]; int c = g_legal_actions(&g, TOP_LEFT, list);
            int idx = (int)bounded(blk[3], (uint32_t)c);
            int k = g_find_cube_to_move(&g, list[idx][0] == 1, TOP_LEFT);
            g_execute_move(&g, TOP_LEFT, k, list[idx][1]);
            steps++; plies++;
            if (g_check_win(&g)) { wins++; break; }
            g.player = BOTTOM_RIGHT; g.dice = 1 + (int)bounded(blk[0], (uint32_t)g.cube_num);
            c = g_legal_actions(&g, BOTTOM_RIGHT, list);
            idx = (int)bounded(blk[1], (uint32_t)c);
            k = g_find_cube_to_move(&g, list[idx][0] == 1, BOTTOM_RIGHT);
            g_execute_move(&g, BOTTOM_RIGHT, k, list[idx][1]);
            plies++;
            if (g_check_win(&g)) break;
            g.player = TOP_LEFT; g.dice = 1 + (int)bounded(blk[2], (uint32_t)g.cube_num);
        }
    }
    if (steps_out) *steps_out = steps;
    if (plies_out) *plies_out = plies;
    return wins;
}
