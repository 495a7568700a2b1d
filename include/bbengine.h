/* bbengine.h -- C-ABI of the B200-native BlackBird self-play engine (libbbengine.so).
 *
 * The reference (ZackAttack614/BlackBird) is pure Python with no FFI layer; its "operator API"
 * for the self-play path is the duck-typed Python surface of GameState / MCTS / Model
 * (SURVEY.md 8b).  Each entry point below names the reference interface it replaces
 * (file:line under /root/reference/src).  The Python mirror in blackbird_b200/ binds these
 * with ctypes (INTEGRATION.md shows the stub a reference maintainer would add).
 *
 * Conventions: plain C, opaque handle, int status (0 = OK, <0 = error, bb_last_error() gives
 * text).  Pointers named *_dev are CUDA device pointers (e.g. torch.Tensor.data_ptr()),
 * *_host are host pointers.  `stream` is a cudaStream_t passed as void* (NULL = default
 * stream).  One engine = one GPU; an engine is not thread-safe, distinct engines are.
 */
#ifndef BBENGINE_H
#define BBENGINE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { BB_OK = 0, BB_EINVAL = -1, BB_EILLEGAL = -2, BB_EASSERT = -3, BB_ENOMEM = -4,
       BB_ECUDA = -5, BB_ESTATE = -6 };
enum { BB_TICTACTOE = 0, BB_CONNECT4 = 1, BB_DRAGONCHESS = 2 };   /* TicTacToe.py / Connect4.py / DragonChess.py */
enum { BB_FIXED = 0, BB_DYNAMIC = 1 };                            /* FixedMCTS.py:21-34 / DynamicMCTS.py:14-34 */
enum { BB_EVAL_KAT = 0,        /* integer evaluator of SURVEY.md Appendix B.2 (parity vehicle) */
       BB_EVAL_NET = 1,        /* built-in policy-value network (NetworkFactory.py:22-183)     */
       BB_EVAL_EXTERNAL = 2,   /* caller serves requests (bb_eval_pending / bb_eval_submit)    */
       BB_EVAL_ROLLOUT = 3 };  /* the base class's own evaluator (MCTS.py:346-383): all-ones priors, a fresh random
                                  rollout on every leaf visit, exact values (bb_config.value_f64 = 1)  */
/* winner codes: None, python int 0 (draw), 1.0, 2.0, np.float64(0.0) */
enum { BB_WIN_NONE = -1, BB_WIN_DRAW = 0, BB_WIN_P1 = 1, BB_WIN_P2 = 2, BB_WIN_FLOAT0 = 3 };
/* per-game phases reported by bb_search_phases */
enum { BB_PH_IDLE = 0, BB_PH_RUN = 1, BB_PH_WAIT_EXPAND = 2, BB_PH_WAIT_LEAF = 3,
       BB_PH_MOVE_READY = 4, BB_PH_GAME_OVER = 5, BB_PH_ERROR = 6,
       BB_PH_COMPACT = 7 /* copying the new root's subtree after a move (busy) */ };

const char* bb_last_error(void);
int bb_abi_version(void);

/* ---- game description -------------------------------------------------------------------- */
int bb_game_state_bytes(int game);     /* packed state size: 4 / 16 / 40                      */
int bb_game_actions(int game);         /* BoardState.LegalMoves: 9 / 7 / 4032                 */
int bb_game_plane_shape(int game, int* h, int* w, int* c);   /* AsInputArray() is [1,h,w,c]   */

/* ---- batched rule kernels over n packed states (all pointers device) ---------------------- */
/* BoardState.__init__ (TicTacToe.py:18-21, Connect4.py:19-22, DragonChess.py:38-60) */
int bb_game_init(int game, void* states_dev, int n, void* stream);
/* BoardState.LegalActions -> float64[n][A] of 0/1 (TicTacToe.py:29-36, Connect4.py:30-36, DragonChess.py:78-106) */
int bb_game_legal(int game, const void* states_dev, int n, double* out_dev, void* stream);
/* BoardState.ApplyAction; ok_dev[i]=0 where the reference raises ValueError('Tried to make an
 * illegal move.') and the state is left untouched (TicTacToe.py:41-48, Connect4.py:41-53, DragonChess.py:127-214) */
int bb_game_apply(int game, void* states_dev, const int32_t* actions_dev, int n, uint8_t* ok_dev, void* stream);
/* BoardState.Winner(prevAction): prev_actions_dev NULL or entry <0 means None
 * (TicTacToe.py:57-105, Connect4.py:62-108, DragonChess.py:161-167) */
int bb_game_winner(int game, const void* states_dev, const int32_t* prev_actions_dev, int n, int32_t* out_dev, void* stream);
/* BoardState.AsInputArray -> int8[n][h*w*c] (TicTacToe.py:50-55, Connect4.py:55-60, DragonChess.py:111-125) */
int bb_game_planes(int game, const void* states_dev, int n, int8_t* out_dev, void* stream);
/* BoardState.__eq__ (TicTacToe.py:132-135, Connect4.py:129-132, DragonChess.py:226-237) */
int bb_game_equal(int game, const void* a_dev, const void* b_dev, int n, uint8_t* out_dev, void* stream);
/* Build packed states from the reference's array form: cells int8[n][h*w] (0/1/2 for the line
 * games; signed piece codes for DragonChess), player[n], prev[n] (0=None), castle[n] bit0..3 */
int bb_game_pack(int game, const int8_t* cells_dev, const uint8_t* player_dev, const uint8_t* prev_dev,
                 const uint8_t* castle_dev, int n, void* states_dev, void* stream);
/* Inverse of bb_game_pack (any output pointer may be NULL). */
int bb_game_unpack(int game, const void* states_dev, int n, int8_t* cells_dev, uint8_t* player_dev,
                   uint8_t* prev_dev, uint8_t* castle_dev, void* stream);

/* ---- engine ------------------------------------------------------------------------------ */
typedef struct bb_engine bb_engine;

typedef struct bb_config {
  int32_t game;              /* BB_TICTACTOE | BB_CONNECT4 | BB_DRAGONCHESS                      */
  int32_t search;            /* BB_FIXED | BB_DYNAMIC                                            */
  int32_t max_depth;         /* FixedMCTS maxDepth (FixedMCTS.py:9,15-16); ignored for dynamic   */
  int32_t evaluator;         /* BB_EVAL_*                                                        */
  int32_t n_games;           /* concurrent game slots on this GPU                                */
  int32_t node_cap;          /* tree nodes per half-arena of a game: one move's growth + the
                                subtree carried over MoveRoot (overflow -> BB_ENOMEM, never silent) */
  int32_t max_plies;         /* record slots per game; engine-side ply cap      */
  int32_t device;            /* CUDA device ordinal                                              */
  double exploration_rate;   /* MCTS.ExplorationRate (MCTS.py:119)                               */
  int32_t value_f64;         /* 0: Node.Value accumulates in f32 (Model path, A.9); 1: f64       */
  int32_t lanes;             /* independent sub-populations of slots (own stream + mailbox each) whose ticks
                                overlap on the GPU; 0 = choose (2 from 512 slots up).  Results do not depend on it. */
  int32_t contexts;          /* 0: exact mode -- one simulation in flight per game, the reference's order (default).
                                W >= 1: virtual-loss mode -- W simulations of a game in flight, visits counted on the way down,
                                values on the way up, contexts resumed in strict cyclic order (deterministic).  W = 1 gives the
                                exact mode's results; W > 1 is NOT the reference's search (north_star (2), SURVEY.md 7.3-2:
                                reported separately).  Needs BB_EVAL_NET with the streaming kernel pair (bf16, 16 filters). */
  int32_t reserved[5];
} bb_config;

int bb_engine_create(const bb_config* cfg, bb_engine** out);
int bb_engine_destroy(bb_engine* e);

/* ---- built-in network (NetworkFactory.py:22-183 forward, Network.py:48-64) ---------------- */
typedef struct bb_net_desc {
  int32_t blocks, filters, dense;   /* parameters_template.yaml:12-16.  filters: 16 (the template: shared-memory-resident kernels, streaming
                                       self-play); 32 / 64 / 128 / 256 with precision 1 (layer-by-layer tensor-core kernels, tick-mode
                                       searches); any width up to 1024 with precision 0 */
  int32_t cin;                      /* input planes (3 / 3 / 17)      */
  float epsilon;                    /* policy noise weight (A.17); 0 = deterministic */
  float alpha;
  uint64_t noise_seed;
  int32_t precision;                /* 0: fp32 SIMT kernel (1e-5 parity); 1: bf16 tcgen05 tensor-core kernel (2e-2) */
  int32_t reserved[3];
} bb_net_desc;
/* weights_host: flat fp32 in the order documented in DESIGN.md ("weight layout"). */
int bb_net_load(bb_engine* e, const bb_net_desc* d, const float* weights_host, size_t n_floats);
/* Network.getPolicy + getEvaluation for n packed states: policy f32[n][A] (softmax, noise,
 * renormalised), value f32[n] (tanh). */
int bb_net_forward(bb_engine* e, const void* states_dev, int n, float* policy_dev, float* value_dev, void* stream);
/* n draws of the policy-noise sampler the network kernels use: dist.sample([1, A])[0][:, 0] of Dirichlet([alpha,
 * 1 - alpha]) = Beta(alpha, 1 - alpha) per action (NetworkFactory.py:176-180).  TensorFlow's stream cannot
 * be matched, so the tests check the distribution itself (Kolmogorov-Smirnov against the Beta cdf). */
int bb_debug_beta_noise(float alpha, uint64_t seed, int n, float* out_dev, void* stream);
/* Appendix B.2 integer evaluator, same signature (parity vehicle). */
int bb_kat_forward(int game, const void* states_dev, int n, float* policy_dev, float* value_dev, void* stream);

/* ---- search (MCTS.py:101-399) ------------------------------------------------------------ */
/* Start games: every slot gets the initial position and no Root (Model.Game() + DropRoot,
 * Blackbird.py:241-244).  states_dev != NULL starts from those packed states instead. */
int bb_search_reset(bb_engine* e, const void* states_dev);
/* MCTS.DropRoot (MCTS.py:141-144) for all slots. */
int bb_search_drop_root(bb_engine* e);
/* FindMove's search half (MCTS.py:184-195): create Root if None, then run play_limit
 * simulations per slot (Root.Plays grows by exactly play_limit, MCTS.py:297-299).
 * Runs ticks (advance kernel + evaluator) until every slot is MOVE_READY.  With
 * BB_EVAL_EXTERNAL it returns after one advance with *pending = number of open requests. */
int bb_search_begin_move(bb_engine* e, int play_limit);
int bb_search_run(bb_engine* e, int* pending);
/* Root statistics after the search: Root.ChildPlays() as sparse (action,plays) lists padded to
 * kmax per slot, Root.Plays, Root.Value (as double), and child count.
 * (MCTS.py:43-98; used by FindMove's return value MCTS.py:198-199) */
int bb_search_root_stats(bb_engine* e, int kmax, int32_t* n_children_host, int32_t* actions_host,
                         uint32_t* plays_host, float* child_value_host, double* child_prior_host,
                         uint32_t* root_plays_host, double* root_value_host, int32_t* root_has_parent_host);
/* _selectAction(exploring=False) (MCTS.py:335-338) on the device for temp==1 (uniforms: one
 * double per slot = RandomState.random_sample()) or temp==0 (PUCT argmax, A.15).  actions out. */
int bb_search_choose(bb_engine* e, double temp, const double* uniforms_host, int32_t* actions_host);
/* state = nextState; MoveRoot(state) (Blackbird.py:249-251, MCTS.py:260-282), by action. */
int bb_search_move_root(bb_engine* e, const int32_t* actions_host);
/* BB_EVAL_ROLLOUT: the draws of MCTS.SampleValue's rollouts (np.random.choice(actions), MCTS.py:378-379) come from a
 * per-slot stream of raw 32-bit MT19937 outputs, words_host [n_games][words_per_slot] -- numpy's own global stream when
 * the reference's games are to be replayed bit for bit (legacy RandomState.randint: one masked-rejection draw per
 * rollout step, none when a single action is legal).  The call rewinds every slot's cursor; bb_search_rng_used reports
 * how many words each slot consumed so that the caller can advance its own generator by the same amount.  A slot that
 * runs out of words stops with BB_ESTATE. */
int bb_search_set_rng(bb_engine* e, const uint32_t* words_host, uint32_t words_per_slot);
int bb_search_rng_used(bb_engine* e, uint32_t* used_host);
/* current root states / phases / status codes per slot */
int bb_search_states(bb_engine* e, void* states_host);
int bb_search_phases(bb_engine* e, int32_t* phases_host, int32_t* status_host);
/* External evaluator (Model.GetPriors / SampleValue overrides, MCTS.py:346-383): list open
 * requests (slot ids + packed states, device pointers valid until submit), then answer them. */
int bb_eval_pending(bb_engine* e, int32_t* n_host, const int32_t** slots_dev, const void** states_dev);
/* Copy the open requests (packed states, slot ids) into caller-owned device buffers. */
int bb_eval_fetch(bb_engine* e, void* states_out_dev, int32_t* slots_out_dev);
int bb_eval_submit(bb_engine* e, const float* policy_dev, const float* value_dev);

/* ---- whole self-play (Blackbird.GenerateTrainingSamples, Blackbird.py:219-268) ----------- */
/* Plays every slot's game to the end on the device: play_limit simulations per move, move
 * choice at temp 1 from uniforms_host[n_games][max_plies] (one double per move, A.15), root
 * motion, Winner(None) after every move, records kept on the device. */
int bb_selfplay_run(bb_engine* e, int play_limit, double temp, const double* uniforms_host);
/* The same loop for nGames > slots (Blackbird.py:238 `for _ in range(nGames)`): n_rows >= n_games whole games on the
 * engine's n_games slots; a slot that ends a game starts the next unplayed one, so the GPU stays full until the last
 * games run out.  Game r draws its moves from uniforms_host[r][max_plies] and its records land in row r whichever
 * slot played it; bb_records_export / bb_records_device then cover bb_records_rows() rows. */
int bb_selfplay_games(bb_engine* e, int n_rows, int play_limit, double temp, const double* uniforms_host);
int bb_records_rows(bb_engine* e);   /* rows of the last bb_selfplay_run / bb_selfplay_games call (n_games after create) */
/* Steady-state stepping for throughput measurement and streaming generation: every slot plays
 * n_moves further moves (play_limit simulations each); a slot whose game ends starts the next
 * game in place (Blackbird.py:238-244 loop).  uniforms_host may be NULL to keep the table. */
int bb_selfplay_step(bb_engine* e, int n_moves, int play_limit, double temp, const double* uniforms_host);
/* Steady state without a per-move barrier (throughput measurement, streaming generation): slots
 * keep searching, moving and restarting finished games in place until the engine as a whole has
 * played target_moves further moves; slots are left mid-search and a later call carries on. */
int bb_selfplay_stream(bb_engine* e, long long target_moves, int play_limit, double temp, const double* uniforms_host);
/* The cudaStream_t all engine kernels are launched on (for CUDA-event timing by the caller). */
void* bb_engine_stream(bb_engine* e);
/* Per-tick CUDA-event timing of the advance kernel and the evaluator kernel on the engine's
 * stream (measurement only; up to max_ticks ticks are recorded after enabling). */
int bb_profile_enable(bb_engine* e, int enable, int max_ticks);
int bb_profile_read(bb_engine* e, double* advance_ms, double* eval_ms, int* ticks);
/* Fixed-stride record export (host buffers).  Per slot: n_moves, winner code (BB_WIN_*; NONE =
 * truncated by max_plies), and per record r in [0, n_moves] (the last one is the terminal
 * record): packed state, player to move, and sparse child plays (offsets into a per-slot pool).
 * Sizes: states [n_games][max_plies+1][state_bytes]; rec_off [n_games][max_plies+2];
 * pool_actions/pool_plays [n_games][pool_cap].  Use bb_records_pool_cap() to size. */
int bb_records_pool_cap(bb_engine* e);
int bb_records_export(bb_engine* e, int32_t* n_moves_host, int32_t* winner_host, void* states_host,
                      uint32_t* rec_off_host, uint16_t* pool_actions_host, uint32_t* pool_plays_host,
                      uint16_t* chosen_host);
/* Same buffers as device pointers (for the NCCL all-gather of records). */
int bb_records_device(bb_engine* e, void** n_moves_dev, void** winner_dev, void** states_dev,
                      void** rec_off_dev, void** pool_actions_dev, void** pool_plays_dev, void** chosen_dev);

/* ExampleState.SerializeState (Blackbird.py:68-81, proto/state.proto:3-9) with the outcome labels of
 * GenerateTrainingSamples (Blackbird.py:260-264), on the device: every record of the last self-play call as its final
 * protobuf bytes, game after game (row order), record after record, in ONE compact stream -- what Connection.PutGames
 * stores (DataManager.py:41-50) and what the ranks all-gather for the trainer.  All records of a game have the same
 * size (Connect4: 200 bytes with a label, 195 for draws), so record r of game g starts at
 * game_off_host[g] + r * rec_bytes_host[g]; game g has (game_off_host[g+1] - game_off_host[g]) / rec_bytes_host[g]
 * records.  out_dev == NULL only reports the sizes (n_records, n_bytes, game_off_host [rows+1], rec_bytes_host [rows];
 * any of them may be NULL). */
int bb_records_serialize(bb_engine* e, void* out_dev, uint64_t cap_bytes, uint64_t* n_records, uint64_t* n_bytes,
                         uint64_t* game_off_host, uint32_t* rec_bytes_host);

/* bb_records_serialize for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  This entry point has NO counterpart in the reference.
 * It produces the same byte stream, but the per-game offsets its pack kernel reads live in a file-scope
 * `static __device__` array (512 KB, held by every GPU that loads the library until the process exits) instead of a
 * cudaMalloc'ed buffer, and are computed on the device: the call makes no cudaMalloc / cudaFree and uses no allocating
 * library routine.  Limits: at most bb_records_static_max_rows() = 65536 games per call (more -> BB_EINVAL, use
 * bb_records_serialize); cap_bytes below the stream's size -> BB_EINVAL with *n_bytes set to what is needed.  The
 * caller passes every buffer: out_dev (device, cap_bytes), game_off_dev (device, uint64 [rows + 1], may be NULL),
 * rec_bytes_dev (device, uint32 [rows], may be NULL); n_records / n_bytes are host outputs (may be NULL).  All calls on
 * one GPU share the array, and it keeps the last call's contents: a call rewrites every word it reads, so nothing needs
 * clearing, but two calls that OVERLAP on the GPU (two engines on one device driven from two threads) race. */
int bb_records_serialize_static(bb_engine* e, void* out_dev, uint64_t cap_bytes, uint64_t* game_off_dev, uint32_t* rec_bytes_dev,
                                uint64_t* n_records, uint64_t* n_bytes);
int bb_records_static_max_rows(void);

/* ---- trainer step: Network.train (Network.py:66-84) over the loss / optimizers of
 *      NetworkFactory.py:185-245, on the GPU ------------------------------------------------------------ */
typedef struct bb_trainer bb_trainer;
typedef struct bb_train_desc {
  int32_t game;                     /* BB_TICTACTOE | BB_CONNECT4 | BB_DRAGONCHESS: input planes and action count    */
  int32_t blocks, filters, dense;   /* parameters_template.yaml:12-16                                                 */
  int32_t optimizer;                /* 0 adam, 1 momentum, 2 gradient descent (network.training.optimizer, :229-242) */
  float momentum;                   /* network.training.momentum (MomentumOptimizer only)                            */
  float epsilon, alpha;             /* policy noise of the TRAINING forward (the loss sees the noisy policy, :176-194) */
  uint64_t noise_seed;
  int32_t device;
  int32_t reserved[5];
} bb_train_desc;
/* weights_host: the flat fp32 vector of bb_net_load (DESIGN.md "weight layout"); BN moving statistics stay frozen. */
int bb_trainer_create(const bb_train_desc* d, const float* weights_host, size_t n_floats, bb_trainer** out);
int bb_trainer_destroy(bb_trainer* t);
/* One sess.run(trainingOp) (Network.py:78): forward, loss as written, gradients, optimizer update.  Device pointers:
 * planes int8 [B][H][W][C] (np.vstack of ExampleState.Board), eval f32 [B], policy f32 [B][A].  noise_host: the [A]
 * noise vector the batch shares (tests); NULL draws it on the device when epsilon > 0. */
int bb_trainer_step(bb_trainer* t, const int8_t* planes_dev, const float* eval_dev, const float* policy_dev, int B,
                    float learning_rate, const float* noise_host);
/* The same step with the batch assembled on the device from a compact record stream (bb_records_serialize output, or the
 * all-gathered streams of every rank): rec_off_host[i] = first byte of example i.  No sqlite round trip, no host parsing. */
int bb_trainer_step_stream(bb_trainer* t, const void* stream_dev, const uint64_t* rec_off_host, int B, float learning_rate,
                           const float* noise_host);
/* loss, lossEvaluation, lossPolicy, lossParam of the last step's forward (NetworkFactory.py:187-204; the summaries of :221-228) */
int bb_trainer_losses(bb_trainer* t, float out[4]);
int bb_trainer_weights(bb_trainer* t, float* weights_host, size_t n_floats);
int bb_trainer_set_weights(bb_trainer* t, const float* weights_host, size_t n_floats);
int bb_trainer_grads(bb_trainer* t, float* grads_host, size_t n_floats);     /* gradients of the last step (tests) */
uint64_t bb_trainer_steps(bb_trainer* t);

/* ---- counters (sims, evals, nodes, depth, expansions, launches) ------------- */
typedef struct bb_counters {
  uint64_t sims, evals, nodes, expansions, select_steps, ticks, kernel_launches, moves, games_done;
  /* streaming mode keeps the evaluator's answers in a state-keyed cache shared by all game slots -- the engine-wide form of
   * the reference's functools.lru_cache on Model.SampleValue / Model.GetPriors (Blackbird.py:350,372).  `evals` counts the
   * evaluations the network computed; these two count the evaluations asked of the cache and answered by it.
   * BB_EVAL_CACHE=0 in the environment switches it off (log2 of the entry count otherwise, default 22). */
  uint64_t cache_lookups, cache_hits;
} bb_counters;
int bb_counters_read(bb_engine* e, bb_counters* out, int reset);
int bb_engine_lanes(bb_engine* e);   /* lanes the engine runs with */
/* 1 when searches run in streaming mode: a persistent search kernel and a persistent network kernel side by side,
 * talking through a request ring (built-in tensor-core network only; BB_STREAM=0 forces tick mode).  Results do not
 * depend on the mode.  bb_profile_read then returns the two kernels' durations and the batches served. */
int bb_engine_streaming(bb_engine* e);
/* Streaming-mode busy accounting since engine creation: out[0] batches served by the network kernel, out[1..3]
 * SM cycles its pipelines spent waiting for tickets, in the convolution tower, and in heads + publish (summed over
 * pipelines); out[4] pipelines x SMs serving; out[5..7] search warps: cycles inside the per-slot advance routine, cycles
 * resident, advance calls.  Zeros in tick mode.  bench.py turns them into busy fractions. */
int bb_stream_stats(bb_engine* e, uint64_t out[8]);
/* Streaming mode keeps, per slot, the last position it asked the network about and the answer it got (the serve-mode
 * path of Network.getPolicy / getEvaluation, Network.py:48-64).  After a self-play call that ran to completion every
 * request has been answered, so (state[s], policy[s], value[s]) are evaluator input / output pairs produced through the
 * request ring: the parity tests compare them with the oracle forward.  Host buffers: states [n_games][state_bytes],
 * policy [n_games][A], value [n_games], seq [n_games] (ticket + 1 of the last answer; 0 = the slot never asked).
 * BB_ESTATE when the engine is not in streaming mode. */
int bb_stream_last_answers(bb_engine* e, void* states_host, float* policy_host, float* value_host, uint32_t* seq_host);
/* 1 when two kernels on two streams run concurrently on this device, 0 when launches are being serialised (a
 * profiler in kernel-replay mode): the persistent search / network pair of streaming mode needs real concurrency
 * (each side waits for the other), so under such a profiler a streaming call ends in the engine's watchdog
 * (BB_ESTATE) instead of a result.  Diagnostic only; the engine never changes its path on it. */
int bb_debug_kernels_concurrent(int device);

#ifdef __cplusplus
}
#endif
#endif
