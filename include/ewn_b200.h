/* ewn_b200.h -- C ABI of libewn_b200.so: the B200 (sm_100a) engine behind the ewn-gym surface.
 *
 * The reference (jchen8tw/ewn-gym) is pure Python and has no FFI of its own; the boundary it exposes for this
 * path is three Python interfaces.  Each entry point below states which reference code it
 * replaces (file:line relative to the reference root).  INTEGRATION.md shows the ctypes binding a maintainer of
 * the reference would add.
 *
 * Conventions
 *   - plain C, no torch types: device pointers + sizes + an explicit cudaStream_t (passed as void*; NULL = the
 *     legacy default stream).  Entry points only enqueue work on `stream`; they never synchronise, except the
 *     ewn_vec_* host-buffer calls, which return after the outputs are in the caller's host buffers.
 *   - every function returns 0 on success or a negative EWN_E* code; nothing throws.  The library keeps two process-wide caches,
 *     both mutex-guarded: the 'hybrid' value tables of ewn_expectiminimax (one per device and board size, read-only once built)
 *     and the scratch used when a search call passes no workspace (one per device and stream -- calls on one stream are ordered
 *     by the stream, calls on different streams never share scratch).
 *   - `state` is the packed game array: ewn_state_words_cfg(N, L) u64 per game (1 for N<=5, 2 for N<=9 with cube_layer <= 3;
 *     2*ceil(cube_num/8)+1 for cube_layer >= 4), game-major.
 *     Layout is documented in DESIGN.md "Packed state" and ewn_gym_b200/csrc/ewn_core.cuh.
 *   - boards are int16 [n][N][N] exactly as the reference's observation (envs/ewn.py:63-69): +k TOP_LEFT cube k,
 *     -k BOTTOM_RIGHT cube k, 0 empty.  Players: 1 = TOP_LEFT, 2 = BOTTOM_RIGHT (constants/player.py:5-6).
 *   - actions are [larger, direction] pairs (envs/ewn.py:61-62).
 *   - legal masks use bit (larger*3 + direction).
 *   - there is no CPU fallback: without a CUDA device every compute entry point returns EWN_ECUDA.
 */
#ifndef EWN_B200_H
#define EWN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define EWN_API __attribute__((visibility("default")))
#else
#define EWN_API
#endif

#define EWN_OK 0
#define EWN_EARG (-1)      /* bad argument (NULL pointer, size, board_size / cube_layer out of range) */
#define EWN_ECUDA (-2)     /* CUDA runtime error; ewn_last_cuda_error() has the text */
#define EWN_ENOSUP (-3)    /* combination not supported (e.g. search with cube_layer != 3, depth > EWN_MAX_DEPTH) */
#define EWN_ESPACE (-4)    /* workspace too small */

#define EWN_MAX_BOARD 9
#define EWN_MAX_DEPTH 8

#define EWN_TOP_LEFT 1
#define EWN_BOTTOM_RIGHT 2

/* step result codes (info["message"] of envs/ewn.py:448,454,473,478 and training_ewn.py:56) */
#define EWN_CODE_ONGOING 0
#define EWN_CODE_AGENT_ILLEGAL 1      /* "Invalid move for player! End the game."    reward -R, terminated, truncated */
#define EWN_CODE_AGENT_WON 2          /* "You won!"                 reward +R, terminated            */
#define EWN_CODE_OPP_ILLEGAL 3        /* "Invalid move for opponent! End the game."  reward 0,  terminated, truncated */
#define EWN_CODE_AGENT_LOST 4         /* "You lost!"                reward -R, terminated            */
#define EWN_CODE_ILLEGAL_TOLERATED 5  /* training_ewn.py:48-56      reward illegal_move_reward, state unchanged */

#define EWN_H_HYBRID 0        /* envs/minimax_ewn.py:39-86   */
#define EWN_H_MIN_DIST 1      /* envs/minimax_ewn.py:88-131  */
#define EWN_H_TWO_MIN_DIST 2  /* envs/minimax_ewn.py:133-178 */
#define EWN_H_ATTK 3          /* envs/minimax_ewn.py:180-213 */

EWN_API int ewn_version(void);
EWN_API const char *ewn_strerror(int code);
EWN_API const char *ewn_last_cuda_error(void);
/* number of CUDA devices visible, or EWN_ECUDA */
EWN_API int ewn_device_count(void);
/* "dddd:bb:dd.f" of a CUDA device (len >= 13): lets the host layer find the NUMA node the GPU hangs off
 * (/sys/bus/pci/devices/<id>/numa_node) and keep its page-locked buffers and stepping thread there (vec_env.py) */
EWN_API int ewn_device_pci_bus_id(int device, char *out, int len);
/* u64 words per game for this board size with cube_layer <= 3 (1 or 2), or EWN_EARG */
EWN_API int ewn_state_words(int board_size);
/* u64 words per game for (board_size, cube_layer): 1 / 2 for the six-cube layouts, 2 * ceil(cube_num / 8) + 1 for cube_layer >= 4
 * (envs/ewn.py:47,51: any cube_layer < board_size - 1, cube_num = L (L + 1) / 2), or EWN_EARG */
EWN_API int ewn_state_words_cfg(int board_size, int cube_layer);
/* total kernels launched by this library in this process (bench.py reports it as gpu_launches) */
EWN_API int64_t ewn_launch_count(void);

/* ---- state codec ----------------------------------------------------------------------------------------- */

/* setup_game + first roll_dice (envs/ewn.py:94-108, :488-494).  Die = Philox(seed; env_id0+i, step, RESET). */
EWN_API int ewn_reset(uint64_t *state, int64_t n, int board_size, int cube_layer, uint64_t seed, uint64_t step, uint64_t env_id0,
              void *stream);
/* Same layout, dice supplied by the caller (int32[n], values 1..cube_num). */
EWN_API int ewn_reset_dice(uint64_t *state, const int32_t *dice, int64_t n, int board_size, int cube_layer, void *stream);

/* observation -> packed state: the piece-list rebuild of classical_policies/minimax.py:75-87 restore_env_with_obs.
 * dice: int32[n] or NULL (=1); player: int32[n] (1/2) or NULL (= TOP_LEFT). */
EWN_API int ewn_encode(const int16_t *board, const int32_t *dice, const int32_t *player, uint64_t *state, int64_t n, int board_size,
               int cube_layer, void *stream);
/* packed state -> observation (envs/ewn.py:485: {"board", "dice_roll"}); any output pointer may be NULL. */
EWN_API int ewn_decode(const uint64_t *state, int16_t *board, int64_t *dice, int32_t *player, int64_t n, int board_size, int cube_layer,
               void *stream);

/* packed state -> the NN-facing observation of classical_policies/alpha_zero/alpha_zero.py:39-58 (ewn_obs_one_hot, also
 * minimax.py:118-139): board f32[n][N][N][cube_num] with +1 where TOP_LEFT's cube k+1 stands and -1 for BOTTOM_RIGHT's,
 * dice f32[n][cube_num] one-hot (or NULL).  Combine with ewn_flip for the canonical form of EWNGame.py:85-92. */
EWN_API int ewn_decode_onehot(const uint64_t *state, float *board, float *dice, int64_t n, int board_size, int cube_layer, void *stream);

/* ---- rules ----------------------------------------------------------------------------------------------- */

/* envs/ewn.py:338-375 get_legal_actions(player).  player: 0 = the state's side to move, 1 / 2 = forced.
 * mask: u8[n]; list: int8[n][6][2] in the reference's list order, padded with -1 (or NULL); count: int32[n] (or NULL). */
EWN_API int ewn_legal(const uint64_t *state, int player, uint8_t *mask, int8_t *list, int32_t *count, int64_t n, int board_size,
              int cube_layer, void *stream);
/* envs/ewn.py:131-142 check_win -> u8[n] */
EWN_API int ewn_check_win(const uint64_t *state, uint8_t *win, int64_t n, int board_size, int cube_layer, void *stream);
/* One ply for `player` (0 = side to move): find_cube_to_move + execute_move + check_win (envs/ewn.py:439-451, :377-412
 * make_simulated_action).  action: int32[n][2].  Does not roll the die; flips the side to move when the move was valid.
 * valid / win: u8[n] (or NULL); moved_cube: int32[n] 1-based cube number, 0 if none (or NULL). */
EWN_API int ewn_ply(uint64_t *state, int player, const int32_t *action, uint8_t *valid, uint8_t *win, int32_t *moved_cube, int64_t n,
            int board_size, int cube_layer, void *stream);
/* MinimaxEnv.set_dice_roll (envs/minimax_ewn.py:26-27): overwrite the die (int32[n]) and optionally the side to move. */
EWN_API int ewn_set_dice(uint64_t *state, const int32_t *dice, const int32_t *player, int64_t n, int board_size, int cube_layer,
                 void *stream);
/* envs/ewn.py:289-296 opponent_action: the position as the other side sees it, np.rot90(-board, 2) -- cube k at (x, y) becomes
 * the other colour's cube k at (N-1-x, N-1-y); die kept, side to move mirrored.  `out` may alias `state`.  Policies reason as
 * TOP_LEFT (minimax.py:92, mcts.py:49), so a BOTTOM_RIGHT player is served by flip -> policy -> play the same [larger, dir]. */
EWN_API int ewn_flip(const uint64_t *state, uint64_t *out, int64_t n, int board_size, int cube_layer, void *stream);
/* classical_policies/random_policy.py:11-15 RandomAgent.predict for the TOP_LEFT side of every game:
 * action int64[n][2] = uniform over the ordered legal list, index = bounded(Philox(seed; env_id0+i, step, STEP)[3]). */
EWN_API int ewn_random_action(const uint64_t *state, int64_t *action, int64_t n, int board_size, int cube_layer, uint64_t seed,
                      uint64_t step, uint64_t env_id0, void *stream);

/* ---- env step (envs/ewn.py:436-486; envs/training_ewn.py:40-99 when shaped) ------------------------------ */

typedef struct ewn_step_args {
    uint64_t *state;              /* in/out, n games, agent (TOP_LEFT) to move                                  */
    int64_t n;
    int32_t board_size, cube_layer;
    /* agent action: exactly one of the three, or none when agent_random != 0                                   */
    const int64_t *action_i64;    /* [n][2]  (SB3 VecEnv.step_async dtype)                                      */
    const int32_t *action_i32;    /* [n][2]                                                                     */
    const uint8_t *action_u8;     /* [n]     packed larger*3+dir (engine-native)                                */
    /* outputs (any may be NULL)                                                                                */
    float *reward_f32;            /* [n]                                                                        */
    double *reward_f64;           /* [n]  bit-exact shaped rewards                                              */
    uint8_t *terminated;          /* [n]  "done"                                                                */
    uint8_t *truncated;           /* [n]                                                                        */
    int32_t *code;                /* [n]  EWN_CODE_*                                                            */
    uint8_t *action_out;          /* [n]  larger*3+dir actually played (useful with agent_random)               */
    int16_t *obs_board;           /* [n][N][N] observation after the step (after auto-reset for finished games) */
    int64_t *obs_dice;            /* [n]                                                                        */
    int16_t *term_board;          /* [n][N][N] rows written only where terminated: info["terminal_observation"] */
    int64_t *term_dice;           /* [n]       likewise                                                         */
    /* compacted terminal observations: when term_count != NULL, finished games append their terminal observation  */
    /* at row k = atomicAdd(term_count) of term_board / term_dice and record their env index in term_index[k]     */
    /* (k order is arbitrary; the caller zeroes *term_count).  With term_count == NULL the rows are dense.        */
    uint32_t *term_count;         /* [1] device counter                                                         */
    int32_t *term_index;          /* [n] term_index[k] = term_index_base + position of the game in this call    */
    int64_t term_index_base;
    uint32_t *term_count_clear;   /* [1] optional: set to 0 by this launch -- the counter the NEXT step will append to,   */
                                  /*     so that a rollout loop alternates two counters and never launches a memset      */
    unsigned long long *stats;    /* [4] += {agent wins, agent losses, illegal-move endings, finished episodes}: the  */
                                  /*     counters that are all-reduced across GPUs (eval_random.py:101-108 win rate)   */
    /* MiniMaxHeuristicEnv state (required when shaped != 0), persists across resets: training_ewn.py:35-38     */
    int32_t *tolerance;           /* [n] in/out                                                                 */
    double *prev_score;           /* [n] in/out                                                                 */
    /* replay mode: dice and opponent answers recorded from the reference instead of Philox (all int32)         */
    const int32_t *inject_dice_opp;   /* [n]     roll at envs/ewn.py:458                                        */
    const int32_t *inject_opp_action; /* [n][2]  opponent_action() at :461                                      */
    const int32_t *inject_dice_next;  /* [n]     roll at :483 (or the first roll after an auto-reset)           */
    double reward;                /* goal reward, envs/ewn.py:38                                                */
    double illegal_move_reward;   /* training_ewn.py:25                                                         */
    uint64_t seed, step, env_id0; /* Philox key / counter: env i uses (env_id0+i, step)                         */
    int32_t agent_random;         /* TOP_LEFT plays uniformly over its legal list inside the kernel             */
    int32_t autoreset;            /* SB3 VecEnv contract                                                        */
    int32_t shaped;               /* MiniMaxHeuristicEnv.step                                                   */
    int32_t flags;                /* EWN_STEP_* bits, 0 = defaults                                              */
    const uint64_t *step_dev;     /* optional device counter added to `step` when the kernel runs: a CUDA graph of K   */
                                  /* captured steps (step = 1..K baked in) is replayed with fresh Philox counters by   */
                                  /* ewn_counter_add(step_dev, K) at the end of the graph                              */
    uint8_t *term_code;           /* optional, rows as term_dice (compacted with term_count, else dense): EWN_CODE_* of the   */
                                  /* finished game -- info["message"] of envs/ewn.py:448,454,473,478 without guessing it     */
                                  /* from the reward's sign                                                                  */
} ewn_step_args;

/* flags: run the VecEnv surface on the one-block-per-128-games kernel instead of the persistent one (A/B measurements) */
#define EWN_STEP_BLOCK_KERNEL 1
/* flags: launch the persistent kernel without the programmatic-dependent-launch attribute (A/B measurements) */
#define EWN_STEP_NO_PDL 2

EWN_API int ewn_step(const ewn_step_args *args, void *stream);
/* The two dice a step draws for games i = 0..n-1 (Philox STEP block of (env_id0+i, step [+ *step_dev]): word 0 = the die rolled for
 * the opponent's ply at envs/ewn.py:458, word 2 = the die at :483 / the first die after an auto-reset), as int32[n] arrays (either
 * may be NULL).  For opponents that are not the in-kernel random player (envs/ewn.py:265-296 opponent_policy): the caller plays the
 * agent's ply on a copy (ewn_ply), sets dice_opp, flips (ewn_flip), asks its policy, and then runs the real step in replay mode with
 * inject_dice_opp = dice_opp, inject_opp_action = the policy's answers, inject_dice_next = dice_next -- the same dice the fused
 * random-opponent step would have drawn. */
EWN_API int ewn_step_dice(int32_t *dice_opp, int32_t *dice_next, int64_t n, int board_size, int cube_layer, uint64_t seed, uint64_t step,
                  uint64_t env_id0, const uint64_t *step_dev, void *stream);
/* *counter += k on `stream` (one tiny kernel; capturable).  See ewn_step_args.step_dev. */
EWN_API int ewn_counter_add(uint64_t *counter, uint64_t k, void *stream);

/* ---- rollout collection with a policy in the loop (BASELINE configs[4]) ----------------------------------- */

/* The parameters of SB3's "MultiInputPolicy" as train.py:35-63 builds it (policy_kwargs = dict(activation_fn = Tanh), default
 * net_arch pi = vf = [64, 64]), as device pointers to fp32 arrays in torch's Linear layout ([out][in] row-major; bias [out]):
 * features = Flatten(board int16 N*N as float) ++ one-hot(dice - 1, cube_num + 1) (CombinedExtractor on envs/ewn.py:63-69);
 * latent_pi = tanh(pi_w2 tanh(pi_w1 f + pi_b1) + pi_b2), latent_vf likewise; logits = action_w latent_pi + action_b (2 + 3 logits of
 * MultiDiscrete([2, 3]), envs/ewn.py:61-62); value = value_w latent_vf + value_b.  stable_baselines3 is a dependency of the
 * reference that is not vendored in it (requirement.txt); the architecture restated here is its published default. */
typedef struct ewn_policy_net {
    int32_t hidden;               /* width of the four hidden layers; 64 is supported (EWN_ENOSUP otherwise) */
    int32_t reserved;
    const float *pi_w1, *pi_b1, *pi_w2, *pi_b2;     /* [hidden][N*N + cube_num + 1], [hidden], [hidden][hidden], [hidden] */
    const float *vf_w1, *vf_b1, *vf_w2, *vf_b2;
    const float *action_w, *action_b;               /* [5][hidden], [5] */
    const float *value_w, *value_b;                 /* [1][hidden], [1] */
} ewn_policy_net;
/* ActorCriticPolicy.forward(obs, deterministic) of train.py:144-155's model.learn() -> collect_rollouts, for n observations in ONE
 * launch: actions int64[n][2] sampled from the two categoricals (Philox block of (seed; env_id0 + i, step [+ *step_dev]), tag POLICY;
 * deterministic != 0: the first arg-max), values f32[n], logp f32[n] = log-probability of the sampled action, logits f32[n][5]
 * (any output may be NULL; logp needs actions).  With actions == NULL this is predict_values(). */
EWN_API int ewn_policy_forward(const ewn_policy_net *net, const int16_t *board, const int64_t *dice, int64_t n, int board_size, int cube_layer,
                       uint64_t seed, uint64_t step, const uint64_t *step_dev, uint64_t env_id0, int deterministic, int64_t *actions,
                       float *values, float *logp, float *logits, void *stream);
/* The WHOLE rollout of train.py:144-155 (collect_rollouts: T x (policy forward, draw, env.step, buffer add), V of the last
 * observation, returns and advantages) in ONE launch: the games of a block of 32 envs never need another block's data, so policy
 * forward and env step alternate on chip with the parameters staged once.  Same Philox schedule as T calls of ewn_policy_forward
 * (policy_seed; env, step + t) and ewn_step (seed; env, step + t): the buffers are bit-identical to that sequence.  In-kernel random
 * opponent, auto-reset, MultiDiscrete actions from the policy; six-cube layouts and 64-wide towers (EWN_ENOSUP otherwise).
 * board int16 [T + 1][n][N][N] and dice int64 [T + 1][n]: slot t = the observation step t's action was computed from, slot T = the one
 * after the last step; actions int64 [T][n][2]; rewards / values / logp / advantages / returns f32 [T][n]; dones u8 [T][n]; last_values
 * f32 [n]; tolerance / prev_score as in ewn_step_args (required when shaped); stats as there (or NULL).
 * net == NULL: no policy -- the actions are READ from `actions` (a pooled / pre-sampled tensor) and only board / dice / rewards / dones are
 * written (the other outputs may be NULL): T env steps of every game in one launch, identical to T calls of ewn_step. */
typedef struct ewn_rollout_args {
    uint64_t *state;              /* in/out, n games, agent to move */
    int64_t n;
    int32_t board_size, cube_layer, T, shaped;
    const ewn_policy_net *net;
    int16_t *board; int64_t *dice; int64_t *actions;
    float *rewards; uint8_t *dones; float *values, *logp, *last_values, *advantages, *returns;
    int32_t *tolerance; double *prev_score; unsigned long long *stats;
    double reward, illegal_move_reward;
    uint64_t seed, step, env_id0; /* env Philox key / first step index / first env id */
    const uint64_t *step_dev;     /* optional device counter added to `step` (CUDA-graph replays) */
    uint64_t policy_seed;
    double gamma, gae_lambda;
    int32_t agent_random;         /* net == NULL only: TOP_LEFT plays uniformly over its legal list inside the kernel (random_policy.py:11-15,  */
    int32_t reserved;             /* as ewn_step_args.agent_random); `actions` (may be NULL) then RECEIVES the moves played                 */
} ewn_rollout_args;
EWN_API int ewn_rollout(const ewn_rollout_args *args, void *stream);
/* RolloutBuffer.compute_returns_and_advantage of the same loop (GAE(gamma, lambda); A2C's default lambda = 1 gives the n-step return):
 * rewards / values f32[T][n], dones u8[T][n] (the episode ended AT step t), last_values f32[n] = V of the observation after the last
 * step -> advantages f32[T][n], returns f32[T][n] (or NULL).  fp32, in the operation order of the reference's tensor loop. */
EWN_API int ewn_gae(const float *rewards, const float *values, const uint8_t *dones, const float *last_values, int T, int64_t n, double gamma,
            double gae_lambda, float *advantages, float *returns, void *stream);

/* eval_random.py:17-51 evaluation() for n seeds, from the seeds ALONE: every game draws its dice (envs/ewn.py:90-92) and both
 * random players' picks (random_policy.py:12-14) from the reference's own stream -- numpy's legacy global RandomState, MT19937
 * seeded by np.random.seed(seed) in env.reset(seed) (:488-494), np.random.randint by masked rejection -- in the reference's
 * consumption order.  result int8[n] (+1 agent = TOP_LEFT won, -1 lost), agent_steps int32[n] (env.step calls), final_state
 * (packed, ewn_decode gives the last board), trace u64[n] (FNV-1a over the per-step dice / action log).  Any output may be NULL. */
EWN_API int ewn_numpy_random_games(const uint32_t *seeds, int64_t n, int board_size, int cube_layer, int8_t *result, int32_t *agent_steps,
                           uint64_t *final_state, uint64_t *trace, void *stream);

/* The same reference stream across launches (tournament driver with a deterministic policy on one side): ewn_numpy_stream stores the
 * first K outputs of MT19937(seed) per game, raw u32[n][K]; ewn_numpy_draw consumes them through a per-game cursor int32[n]:
 * roll_dice != 0 -> die = randint(1, cube_num + 1) (envs/ewn.py:90-92), then pick_player (1 / 2, 0 = none) -> action int32[n][2] =
 * RandomAgent.predict for that side: legal list[randint(0, len)] (random_policy.py:12-14). */
EWN_API int ewn_numpy_stream(const uint32_t *seeds, int64_t n, int K, uint32_t *raw, void *stream);
EWN_API int ewn_numpy_draw(uint64_t *state, const uint32_t *raw, int32_t *cursor, int K, int64_t n, int board_size, int cube_layer,
                   int roll_dice, int pick_player, int32_t *action, void *stream);

/* ---- search ---------------------------------------------------------------------------------------------- */

/* envs/minimax_ewn.py:29-213 evaluate(heuristic) -> f64[n] */
EWN_API int ewn_evaluate(const uint64_t *state, double *value, int64_t n, int board_size, int cube_layer, int heuristic, void *stream);

/* classical_policies/minimax.py:19-93 ExpectiMinimaxAgent.predict on n roots (TOP_LEFT to move, die from the state).
 * pruned bit 0 set reproduces the reference's alpha/beta hand-through at chance nodes bit for bit; clear = textbook.
 * Depth 3 and 4 run in one kernel with the whole tree on chip (no workspace is touched); other depths store 2-byte / 8-byte node
 * records in the workspace and replay minimax.py over them.
 * value: f64[n]; action: int8[n][2] (-1,-1 = None for a finished root).
 * workspace: device scratch of at least ewn_expectiminimax_workspace(n, depth) bytes, or NULL / 0 to let the library
 * use its own cached allocation (one per device: such calls must be ordered on one stream; pass a workspace to run several
 * searches concurrently). */
/* `pruned` bit 1: run depth 3 / 4 on the record + replay kernels that serve depth >= 5 instead of the fused on-chip kernel (A/B) */
#define EWN_EMM_REPLAY_PATH 2
EWN_API size_t ewn_expectiminimax_workspace(int64_t n, int depth);
EWN_API int ewn_expectiminimax(const uint64_t *state, int64_t n, int board_size, int cube_layer, int depth, int heuristic, int pruned,
                       double *value, int8_t *action, void *workspace, size_t workspace_bytes, void *stream);

/* ---- the 'sim_winrate' heuristic (envs/minimax_ewn.py:36-37, :215-238 MinimaxEnv.simulate) ------------------------------------
 * evaluate('sim_winrate') of n positions -> f64[n]: num_simulations (<= 256; the reference hard-codes 100, :21) random playouts with a
 * six-face die, fraction that TOP_LEFT wins; 1.0 / 0.0 on a finished game.  simulate() never restores current_player, so the side
 * that opens a playout is the one that lost the previous playout of the same call sequence; the first playout of a position is
 * opened by BOTTOM_RIGHT here (a fresh env).  Philox streams: (seed; unit = (unit0 + i) * 6, opener, playout index). */
#define EWN_H_SIM_WINRATE 4
EWN_API int ewn_sim_winrate(const uint64_t *state, int64_t n, int board_size, int cube_layer, int num_simulations, uint64_t seed,
                    uint32_t unit0, double *value, void *stream);
/* ExpectiMinimaxAgent(heuristic='sim_winrate').predict (eval_minimax.py --heuristic sim_winrate) for depth 1..4: the tree is expanded
 * to leaf POSITIONS, every live leaf is evaluated by ewn_sim_winrate's playouts (six independent evaluations at even depth,
 * minimax.py:64-73), finished positions are worth 1.0 / 0.0, then minimax.py:19-73 is replayed over the values.  Results are
 * reproducible in (seed, root index) and statistically -- not bit for bit -- those of the reference (CPython's `random`). */
EWN_API int ewn_expectiminimax_sim(const uint64_t *state, int64_t n, int board_size, int cube_layer, int depth, int pruned, int num_simulations,
                           uint64_t seed, double *value, int8_t *action, void *stream);

/* classical_policies/mcts.py:21-45 simulate: random playouts from n states with BOTTOM_RIGHT to move.
 * wins[i] += number of playouts with index in [first, first+count) that TOP_LEFT wins (u32[n], caller zeroes).
 * root_ids: u32[n] Philox stream ids (NULL = 0..n-1). plies: optional u64[1] accumulator of plies played. */
EWN_API int ewn_playouts(const uint64_t *state, const uint32_t *root_ids, int64_t n, int board_size, int cube_layer, uint64_t seed,
                 uint32_t first, uint32_t count, uint32_t *wins, unsigned long long *plies, void *stream);
/* classical_policies/mcts.py:47-84 flat Monte-Carlo over the root's legal actions: for every root i (TOP_LEFT to move)
 * and list position c, wins[i][c] += TL wins of playouts [first, first+count) started after playing that action.
 * wins: u32[n][6] (caller zeroes); n_children: int32[n] (or NULL).
 * workspace: at least ewn_mc_workspace(n) bytes of device scratch, or NULL / 0 for the library's cached allocation. */
EWN_API size_t ewn_mc_workspace(int64_t n);
EWN_API int ewn_mc_root_wins(const uint64_t *state, const uint32_t *root_ids, int64_t n, int board_size, int cube_layer, uint64_t seed,
                     uint32_t first, uint32_t count, uint32_t *wins, int32_t *n_children, unsigned long long *plies,
                     void *workspace, size_t workspace_bytes, void *stream);
/* MctsAgent.predict (mcts.py:47-69) as ONE call: wins u32[n][6] is OVERWRITTEN (cleared by the expansion kernel, no memset launch), the
 * playouts [first, first+count) are played and, when action != NULL, the first arg-max is taken -- three kernels chained with
 * programmatic dependent launch; with the library's own workspace (workspace == NULL) nothing else is launched.  action == NULL leaves
 * the arg-max to the caller (ewn_mc_best_action_peers when the playout range is split over GPUs). */
EWN_API int ewn_mc_predict(const uint64_t *state, const uint32_t *root_ids, int64_t n, int board_size, int cube_layer, uint64_t seed,
                   uint32_t first, uint32_t count, uint32_t *wins, int8_t *action, int32_t *n_children, unsigned long long *plies,
                   void *workspace, size_t workspace_bytes, void *stream);
/* ewn_mc_predict for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * NO COUNTERPART IN THE REFERENCE.  Same arguments and results as ewn_mc_predict with
 * workspace == NULL, but the Monte-Carlo workspace is a file-scope device array of the library (3 MiB, held by every GPU that loads
 * the library until the process exits) and the call performs no cudaMalloc / cudaFree.  Limit: n <= EWN_MC_STATIC_MAX_ROOTS, above it
 * the call returns EWN_ESPACE and ewn_last_cuda_error() says so.  Every call on a device shares the one array: calls that overlap on
 * the GPU (different streams or host threads) race -- keep them on one stream.  Nothing in the array has to survive between calls. */
#define EWN_MC_STATIC_MAX_ROOTS 16384
EWN_API int64_t ewn_mc_static_max_roots(void);
EWN_API int ewn_mc_predict_static(const uint64_t *state, const uint32_t *root_ids, int64_t n, int board_size, int cube_layer, uint64_t seed,
                          uint32_t first, uint32_t count, uint32_t *wins, int8_t *action, int32_t *n_children,
                          unsigned long long *plies, void *stream);
/* first arg-max over the legal children (np.argmax, mcts.py:68) -> action int8[n][2] */
EWN_API int ewn_mc_best_action(const uint64_t *state, const uint32_t *wins, int8_t *action, int64_t n, int board_size, int cube_layer,
                       void *stream);

/* UCT at the root on top of the playout kernel (the intent of the reference's unfinished ucth_temp.py:6-89): `rounds` times,
 * every root picks the legal move with the best UCB1 score wins/visits + c * sqrt(ln(total)/visits) (unvisited first, ties -> list order)
 * and plays `batch` random playouts (mcts.py:24-41) from the position after it; the final move is the one with the best win rate.
 * wins / visits: u32[n][6] by legal-list position (out; overwritten); action: int8[n][2] (or NULL; -1,-1 for a finished root).
 * Playout streams are keyed (seed; root id * 6 + move, playout index), so results do not depend on how roots are batched or sharded.
 * workspace: at least ewn_uct_workspace(n) bytes, or NULL / 0 for the library's per-stream scratch. */
EWN_API size_t ewn_uct_workspace(int64_t n);
EWN_API int ewn_uct_search(const uint64_t *state, const uint32_t *root_ids, int64_t n, int board_size, int cube_layer, uint64_t seed, int rounds,
                   uint32_t batch, double c, uint32_t *wins, uint32_t *visits, int8_t *action, void *workspace, size_t workspace_bytes,
                   void *stream);

/* ---- root-statistics merge over NVLink peer memory (the one exchange step of the path) ----------- */
/* mcts.py:59-68 sums the wins of all env copies and takes np.argmax.  With the playout range of every (root, action) split
 * over the GPUs of one node, that sum is an all-reduce of u32[n][6] followed by the arg-max.  ewn_mc_best_action_peers does
 * both in ONE kernel: every rank publishes "my wins of epoch e are complete" into each peer's flag words, waits for the
 * peers' flags in its own memory, then reads all ranks' win counters with peer loads, sums and picks the first maximum.
 * No NCCL call, no intermediate buffer.
 *
 * Buffers come from ewn_peer_alloc (cudaMalloc, zero-filled) and travel between the processes of a node as 64-byte handles
 * (cudaIpcMemHandle_t) which the host layer exchanges (torch.distributed all_gather_object in ewn_gym_b200/peer.py).
 * flags: u32[EWN_PEER_FLAG_WORDS] per rank; word s = last epoch published by rank s, word EWN_PEER_ERR_WORD = the epoch at
 * which a wait timed out (0 = never).  Epochs start at 1 and increase; the caller alternates between two win buffers so
 * that a rank never rewrites counters a peer may still be reading (the wait of epoch e+1 proves every peer finished e). */
#define EWN_MAX_PEERS 8
#define EWN_PEER_FLAG_WORDS 32
#define EWN_PEER_ERR_WORD 16
typedef struct ewn_peer_handle { unsigned char bytes[64]; } ewn_peer_handle;
EWN_API int ewn_peer_alloc(size_t bytes, void **ptr, ewn_peer_handle *handle);
EWN_API int ewn_peer_open(const ewn_peer_handle *handle, void **ptr);     /* another process's buffer, mapped on the current device */
EWN_API int ewn_peer_close(void *ptr);                                    /* undo ewn_peer_open  */
EWN_API int ewn_peer_free(void *ptr);                                     /* undo ewn_peer_alloc */
/* wins_peers / flag_peers: HOST arrays of `world` device pointers, index = rank (entry `rank` is the caller's own buffer).
 * wins_total: optional u32[n][6] out (the summed counters); action: int8[n][2] as ewn_mc_best_action.
 * timeout_ms: bound on the wait for the peers (0 = 2000); on expiry the kernel records the epoch in its error word and
 * returns without hanging the device -- results of that call are undefined. */
EWN_API int ewn_mc_best_action_peers(const uint64_t *state, const uint32_t *const *wins_peers, uint32_t *const *flag_peers, int rank,
                             int world, uint32_t epoch, uint32_t timeout_ms, uint32_t *wins_total, int8_t *action, int64_t n,
                             int board_size, int cube_layer, void *stream);

/* ---- host-buffer vector environment (the call a numpy / SB3 user makes) ----------------------------------- */
/* Owns device state, pinned staging buffers and streams for n games on the current device.  step: actions come from
 * host memory, observations / rewards / dones land in host memory; copies and the kernel are pipelined in chunks. */
typedef struct ewn_vec ewn_vec;
typedef struct ewn_vec_config {
    int64_t n;
    int32_t board_size, cube_layer;
    double reward, illegal_move_reward;
    uint64_t seed, env_id0;
    int32_t shaped, illegal_move_tolerance;
    int32_t chunks;               /* pipeline depth (0 = default)                                               */
    int32_t device;               /* CUDA device ordinal                                                        */
    int32_t flags;                /* EWN_VEC_* bits, 0 = defaults                                               */
    int32_t reserved;
} ewn_vec_config;
/* flags: always stage outputs in device memory and copy them back (default: page-locked output buffers are written by the
 * kernel directly over PCIe) */
#define EWN_VEC_STAGED 1

EWN_API int ewn_vec_create(const ewn_vec_config *cfg, ewn_vec **out);
EWN_API void ewn_vec_destroy(ewn_vec *v);
/* VecEnv.reset(): obs_board int16[n][N][N], obs_dice int64[n] (host) */
EWN_API int ewn_vec_reset(ewn_vec *v, int16_t *obs_board, int64_t *obs_dice);
/* VecEnv.step(actions int64[n][2]) with auto-reset.  Terminal observations of the finished games come back
 * COMPACTED: *term_count rows in term_board / term_dice, row k belonging to env term_index[k] (all host pointers, each
 * sized for n rows; pass term_count == NULL to skip terminal observations altogether). */
EWN_API int ewn_vec_step(ewn_vec *v, const int64_t *action, int16_t *obs_board, int64_t *obs_dice, float *reward, uint8_t *done,
                 uint8_t *truncated, int16_t *term_board, int64_t *term_dice, int32_t *term_index, uint32_t *term_count);
/* the same with the result code (EWN_CODE_*) of every finished game, term_code u8[n] compacted like term_dice (or NULL) */
EWN_API int ewn_vec_step2(ewn_vec *v, const int64_t *action, int16_t *obs_board, int64_t *obs_dice, float *reward, uint8_t *done,
                  uint8_t *truncated, int16_t *term_board, int64_t *term_dice, int32_t *term_index, uint8_t *term_code,
                  uint32_t *term_count);
/* VecEnv.seed(seed) (train.py:135): the Philox key the NEXT ewn_vec_reset and every step after it use; game i keeps its own
 * stream (key = seed, counter = env_id0 + i), which is what SubprocVecEnv's seed + i per worker amounts to */
EWN_API int ewn_vec_set_seed(ewn_vec *v, uint64_t seed);
/* pinned host memory helpers so that callers can hand in page-locked numpy buffers */
EWN_API int ewn_host_alloc(void **ptr, size_t bytes);
EWN_API int ewn_host_free(void *ptr);
/* device pointer to the packed state owned by the vec env (for zero-copy use from torch) */
EWN_API uint64_t *ewn_vec_state(ewn_vec *v);
EWN_API uint64_t ewn_vec_step_index(const ewn_vec *v);

#ifdef __cplusplus
}
#endif
#endif /* EWN_B200_H */
