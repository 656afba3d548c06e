/*
 * battlespace.h -- C ABI of libbattlespace.so, the B200 (sm_100a) batched implementation of the
 * Deep-RL-Battlespace environment step + instinct policy.
 *
 * Each entry point replaces one piece of the reference's Python surface for this path (reference
 * paths are relative to the reference root):
 *
 *   bs_create            parallel_env.__init__            envs/battle_env.py:73-184
 *   bs_reset[_host]      parallel_env.reset               envs/battle_env.py:246-279
 *                        (Base.reset / Plane.reset        envs/sprites.py:238-252, 74-91)
 *   bs_step[_host]       parallel_env.step                envs/battle_env.py:281-381
 *                        (process_action :383-424, Plane.forward/rotate sprites.py:93-141,
 *                         Bullet.__init__/update sprites.py:293-351, observe :202-244,
 *                         tie/win :469-496)
 *   bs_instinct_actions  instinct Team.choose_actions     instinct/team.py:11-14, instinct/agent.py:10-62
 *   bs_step_team         one game-loop iteration with the scripted team in-kernel
 *                                                         main.py:179-181, evaluate.py:69-70
 *   bs_step_team_replay  the same + ReplayBuffer.store_transition   main.py:205, maddpg/buffer.py:25-47
 *   bs_rollout_fused     the whole inner game loop incl. ReplayBuffer.store_transition
 *                                                         main.py:178-213, maddpg/buffer.py:25-47
 *   bs_rollout_fused_host[_begin/_end]   the same with the learner's actions in HOST memory
 *   bs_actor_forward     NetworkedAgent.choose_action for a whole team (row f1)
 *                                                         maddpg/networks.py:54-85, maddpg/agent.py:25-33,
 *                                                         utils/noise.py:17-21, maddpg/team.py:21-28
 *   bs_rollout_actor     the inner game loop with the learner team played by its actors
 *                                                         main.py:178-213
 *   bs_replay_sample     ReplayBuffer.sample_buffer (row f2)   maddpg/buffer.py:49-67
 *   bs_render            parallel_env.render into a device RGB buffer (row f4; no window)
 *                                                         envs/battle_env.py:498-560, envs/sprites.py:155-178,265-278,353-360
 *   bs_stats             env.team[c]['wins'], ties, total_games + report sums
 *                                                         envs/battle_env.py:102-103,169-170; main.py:218-253
 *   bs_episode_log       score_dict / scores.json history (row f3)     main.py:219-227
 *   bs_get_state/bs_set_state   (no reference equivalent: checkpoint + parity injection)
 *
 * Conventions
 *   - Plain C types only.  Unless a name ends in _host, every buffer is a CALLER-OWNED DEVICE
 *     pointer on the handle's device and the call is asynchronous on `stream` (a cudaStream_t
 *     passed as void*; NULL = the legacy default stream).  *_host entry points take host
 *     pointers, do their own staged copies and return after the results are in host memory.
 *   - Return value: 0 on success, negative BS_E_* otherwise; bs_last_error() gives the
 *     message (thread local).  A handle is bound to one device and is not thread-safe.
 *   - Agents: index i in [0,2N): red = [0,N), blue = [N,2N) (battle_env.py:106-108).
 *     obs row length L = 3N+2 (battle_env.py:132).  A = 3 (continuous) or 4 (discrete).
 *   - There is no CPU fallback: without a CUDA device every compute call fails with BS_E_CUDA.
 *
 * Random draws.  Either injected per call (`draws` arguments: the reference's own draws are
 * replayed for parity) or Philox4x32-10 with
 *     key     = (seed lo32, seed hi32)
 *     counter = (g lo32, g hi32, episode, tick<<16 | agent<<8 | purpose),  g = env_offset + env
 * so that env g sees the same stream however the envs are sharded over GPUs.
 *   purpose 0 (agent tick): words w0..w3 ->
 *       noise_u[k] = (w_k >> 8) / 2^24  (k=0..2)      instinct noise  instinct/agent.py:53
 *       shoot_u    = ((w0&255)<<16 | (w1&255)<<8 | (w2&255)) / 2^24    instinct/agent.py:44
 *       spread_u   = w3 / 2^32                        Bullet spread   envs/sprites.py:314
 *       random policy: continuous a_k = 2*noise_u[k]-1, discrete a = w0 >> 30 (= floor(4*noise_u[0]), which is
 *       what an injected noise_u[0] yields)
 *   purpose 1 (plane spawn, agent = plane): x, y, dir = lo + mulhi(w_k, hi-lo+1)
 *   purpose 2 (base spawn): red x,y from w0,w1; blue x,y from w2,w3
 *   `tick` is the env's step counter before the call, `episode` starts at 1.
 * Injected unit draws are laid out [.., 5] doubles per agent: shoot_u, noise_u0..2, spread_u.
 * Injected spawn draws are [E, 4+6N] int32 in the reference's call order (battle_env.py:257-268):
 * red base x,y, blue base x,y, then x,y,dir for each red plane, then each blue plane.
 */
#ifndef BATTLESPACE_H
#define BATTLESPACE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BS_VERSION 100
#define BS_MAX_AGENTS 8     /* planes per team */
#define BS_BULLET_SLOTS 12  /* per plane: a bullet is removed on its 12th update (sprites.py:334-337) */

enum { BS_OK = 0, BS_E_INVALID = -1, BS_E_CUDA = -2, BS_E_UNSUPPORTED = -3 };
enum { BS_ACT_I32 = 0, BS_ACT_F32 = 1, BS_ACT_F64 = 2, BS_ACT_NONE = 3 };
enum { BS_POL_EXTERNAL = 0, BS_POL_RANDOM = 1, BS_POL_INSTINCT = 2, BS_POL_INSTINCT_FF = 3 };
enum { BS_WINNER_NONE = 0, BS_WINNER_RED = 1, BS_WINNER_BLUE = 2, BS_WINNER_TIE = 3 };

typedef struct bs_handle bs_handle;

/* parallel_env(**kwargs) -- battle_env.py:73 (show / fps are host-side rendering options) */
typedef struct bs_config {
    int32_t n_agents;           /* planes per team, 1..BS_MAX_AGENTS */
    int32_t continuous_actions; /* 0: Discrete(4), 1: Box(-1,1,(3,)) */
    double hit_base_reward;
    double hit_plane_reward;
    double miss_punishment;
    double die_punishment;
    double lose_punishment;
} bs_config;

/* Replay ring in the row layout of maddpg/buffer.py:11-23 (float32; rows of the learner team).
 * The row of (tick t, env e) of a call is (cursor + t*n_envs + e) % capacity.  Rows of different envs must
 * never collide inside one call: capacity >= n_envs, and a call that wraps the ring (T*n_envs > capacity)
 * needs capacity % n_envs == 0 -- the ring then keeps the last capacity/n_envs ticks of that call.  Anything
 * else is answered with BS_E_INVALID. */
typedef struct bs_replay {
    float *state;       /* [capacity, N*L]   state_mem / actor_states[k] = columns k*L..(k+1)*L */
    float *new_state;   /* [capacity, N*L] */
    float *action;      /* [capacity, N, A]  action_mem[k] */
    float *reward;      /* [capacity, N] */
    uint8_t *done;      /* [capacity, N] */
    int64_t capacity;   /* rows; 0 = write nothing */
} bs_replay;

/* One NetworkedAgent.actor -- ActorNetwork, maddpg/networks.py:54-85: float32 DEVICE pointers in the
 * nn.Linear ([out,in] row-major) / nn.LayerNorm layouts of its state_dict. */
typedef struct bs_actor {
    const float *fc1_w, *fc1_b; /* [H,L], [H]   fc1 */
    const float *ln1_g, *ln1_b; /* [H], [H]     bn1 (a LayerNorm, networks.py:61) */
    const float *fc2_w, *fc2_b; /* [H,H], [H]   fc2 */
    const float *ln2_g, *ln2_b; /* [H], [H]     bn2 */
    const float *pi_w, *pi_b;   /* [A,H], [A]   pi */
} bs_actor;

/* OUNoise of every (env, team slot) -- utils/noise.py:4-21, added in choose_action (agent.py:30-32). */
typedef struct bs_ou_noise {
    float *state;              /* [E,N,A] device, updated in place */
    const float *normals;      /* [E,N,A] device randn draws to use, or NULL: Philox + Box-Muller keyed by
                                  (global env, step, team slot) under `seed` */
    const uint8_t *reset_mask; /* [E] device or NULL: != 0 -> OUNoise.reset() first (main.py:150) */
    float mu, theta, sigma, scale; /* reference defaults 0, 0.15, 0.2, 0.1 */
    uint64_t seed;
    uint32_t step;             /* draw index; bs_rollout_actor uses step, step+1, ... */
} bs_ou_noise;

/* Host-side view used by bs_get_state / bs_set_state (any pointer may be NULL = skip). */
typedef struct bs_state_view {
    int16_t *plane_x, *plane_y; /* [E,2N] rect centre */
    double *plane_dir;          /* [E,2N] degrees */
    int8_t *plane_hp;           /* [E,2N] */
    int16_t *base_pos;          /* [E,2,2] (team, xy) */
    int16_t *base_hp;           /* [E,2] may be negative */
    int16_t *bullet_pos;        /* [E,2N,12,2] x,y ; -1,-1 = empty ; slot = age-1 */
    int8_t *bullet_vel;         /* [E,2N,12,3] per-update integer step sx, sy and flags
                                   (bit0 valid, bit1 sx exact, bit2 sy exact) */
    int32_t *tick;              /* [E] */
    uint8_t *env_done;          /* [E] */
    int8_t *winner;             /* [E] BS_WINNER_* */
    uint32_t *episode;          /* [E] */
    double *team_return;        /* [E,2] running return of the current episode */
} bs_state_view;

int bs_version(void);
const char *bs_last_error(void);

int bs_create(const bs_config *cfg, int64_t n_envs, int64_t env_offset, int device, uint64_t seed,
              bs_handle **out);
int bs_destroy(bs_handle *h);

/* Resets the envs with env_mask[e] != 0 (NULL = all).  draws: NULL (Philox) or [E,4+6N] int32.
 * obs: NULL or [E,2N,L] float32, written for every env. */
int bs_reset(bs_handle *h, const uint8_t *env_mask, const int32_t *draws, float *obs, void *stream);

/* One tick for every env.  actions: BS_ACT_I32 [E,2N] | BS_ACT_F32 [E,2N,A] (continuous values,
 * or discrete logits -> first argmax, battle_env.py:327-328) | BS_ACT_F64 [E,2N,3].
 * BS_ACT_NONE (actions NULL) is the reference's "empty action dict": every running env ties
 * (battle_env.py:309-313).
 * draws: NULL or [E,2N,5].  Outputs (each may be NULL): obs [E,2N,L] f32, rew [E,2N] f32,
 * done [E,2N] u8, env_done [E] u8, winner [E] i8.  Envs that are already done return their
 * observation, zero rewards and their dones (battle_env.py:303-306); nothing auto-resets. */
int bs_step(bs_handle *h, const void *actions, int action_dtype, const double *draws, float *obs,
            float *rew, uint8_t *done, uint8_t *env_done, int8_t *winner, void *stream);

/* As bs_step, but the caller supplies only learner_team's actions ([E,N] / [E,N,A] / [E,N,3]);
 * the other team is played in-kernel by other_policy (BS_POL_INSTINCT, _INSTINCT_FF or _RANDOM). */
int bs_step_team(bs_handle *h, int learner_team, const void *team_actions, int action_dtype,
                 int other_policy, const double *draws, float *obs, float *rew, uint8_t *done,
                 uint8_t *env_done, int8_t *winner, void *stream);

/* bs_step_team + ReplayBuffer.store_transition (main.py:205, maddpg/buffer.py:25-47) in the same
 * launch: one row per env for learner_team at row (cursor + e) mod capacity -- state = the observation
 * before the tick, action = the actions as given (continuous: clipped; BS_ACT_I32: one-hot),
 * new_state / reward / done after it.  replay may be NULL.
 * auto_reset != 0: an env that finishes in this tick is reset (Philox spawns) before the call
 * returns, i.e. the game loop's next env.reset() (main.py:148) is folded in: rew/done/env_done/winner
 * and the stored row describe the finished game, obs is the first observation of the new one.
 * auto_reset == 0: the caller resets done envs between calls (bs_reset with the env_done mask). */
int bs_step_team_replay(bs_handle *h, int learner_team, const void *team_actions, int action_dtype,
                        int other_policy, const double *draws, const bs_replay *replay, int64_t cursor,
                        int auto_reset, float *obs, float *rew, uint8_t *done, uint8_t *env_done,
                        int8_t *winner, void *stream);

/* instinct Team.choose_actions for `team`: obs [E,2N,L] f32 in; actions_out [E,N,3] f64
 * (continuous) or [E,N] i32 (discrete).  draws: NULL or [E,2N,5]. */
int bs_instinct_actions(bs_handle *h, int team, const float *obs, const double *draws,
                        void *actions_out, void *stream);

/* T ticks in one launch: done envs reset themselves (Philox spawns), both teams act by policy
 * (learner_team may be BS_POL_EXTERNAL with ext_actions [T,E,N,A] f32), and one replay row per
 * env-tick is written at row (cursor + t*E + e) mod capacity.  replay may be NULL. */
int bs_rollout_fused(bs_handle *h, int T, int learner_team, int policy_red, int policy_blue,
                     const float *ext_actions, const bs_replay *replay, int64_t cursor, void *stream);

/* NetworkedAgent.choose_action for every plane of `team` in every env (maddpg/agent.py:25-33,
 * Team.choose_actions maddpg/team.py:21-28): actions_out[e,k,:] = tanh(pi(relu(ln(fc2(relu(ln(fc1(
 * obs[e, team*N+k, :])))))))) with actor k's parameters, float32 like the reference's network;
 * with `noise` != NULL the OU noise is advanced and added and the result clamped to [-1,1].
 * actors: HOST array of N structs of device pointers.  hidden must be 64 (fc1_dims == fc2_dims,
 * params.json).  obs [E,2N,L] f32 (the layout bs_step / bs_reset write); actions_out [E,N,A] f32 --
 * exactly what bs_step_team takes as BS_ACT_F32 (discrete: logits -> first argmax). */
int bs_actor_forward(bs_handle *h, int team, const bs_actor *actors, int hidden, const float *obs,
                     const bs_ou_noise *noise, float *actions_out, void *stream);

/* T iterations of the training game loop with learner_team played by its actors (main.py:178-213):
 * per tick  bs_actor_forward(obs) -> bs_step_team_replay(auto_reset = 1, row cursor + t*E + e),
 * 2 launches per tick and no host synchronisation.  obs [E,2N,L]: in = the current observation of
 * every env (bs_reset / the previous call), out = the observation after the last tick.
 * act_scratch [E,N,A] f32 and env_done [E] u8 are required work buffers (last tick's values on
 * return); rew/done/winner may be NULL.  With noise != NULL, envs that finished in the previous tick
 * get OUNoise.reset() (reset_mask of the FIRST tick is noise->reset_mask) and draw t uses step+t. */
int bs_rollout_actor(bs_handle *h, int T, int learner_team, const bs_actor *actors, int hidden,
                     int other_policy, const bs_ou_noise *noise, const bs_replay *replay, int64_t cursor,
                     float *obs, float *act_scratch, float *rew, uint8_t *done, uint8_t *env_done,
                     int8_t *winner, void *stream);

/* bs_rollout_actor for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and bit-identical results, but there is no act_scratch argument and env_done may be NULL: the
 * action rows [E,N,A] f32 handed from the actor kernel to the step kernel (and env_done [E] u8 when NULL) live in
 * a file-scope `static __device__` array of BS_STATIC_SCRATCH_BYTES inside the library, which every GPU that
 * loads the library holds until the process exits.  The call makes no cudaMalloc / cudaFree / allocating library
 * call.  Limit: round_up(n_envs*N*A*4, 256) + (env_done ? 0 : n_envs) <= BS_STATIC_SCRATCH_BYTES (8 MiB: about
 * 226,000 envs of 3v3 continuous); a larger handle is answered with BS_E_INVALID and a message naming both numbers.
 * The array is shared by every call on the device and keeps the last call's contents (nothing in it needs
 * clearing: each call writes what it reads): calls that OVERLAP on the GPU -- two streams or two handles on one
 * device -- race on it; serialise them.  The first call on a device resolves the array's device address
 * (cudaGetSymbolAddress); under stream capture, make one call on that device before the capture begins. */
#define BS_STATIC_SCRATCH_BYTES (8u << 20)
int bs_rollout_actor_static(bs_handle *h, int T, int learner_team, const bs_actor *actors, int hidden,
                            int other_policy, const bs_ou_noise *noise, const bs_replay *replay,
                            int64_t cursor, float *obs, float *rew, uint8_t *done, uint8_t *env_done,
                            int8_t *winner, void *stream);

/* ReplayBuffer.sample_buffer (maddpg/buffer.py:49-67) as one gather launch: idx [B] int64 device row
 * indices (the caller draws them: np.random.choice(max_mem, batch_size) with replacement, :51-52).
 * Outputs (device, float32 unless noted; any may be NULL): actor_states [N,B,L], states [B,N*L],
 * actions [N,B,A], rewards [B,N], actor_new_states [N,B,L], states_ [B,N*L], dones [B,N] u8.
 * Needs no env handle: the ring is the caller's; the launch goes to the device that owns `idx`. */
int bs_replay_sample(const bs_replay *replay, int n_agents, int obs_len, int n_actions,
                     const int64_t *idx, int64_t batch,
                     float *actor_states, float *states, float *actions, float *rewards,
                     float *actor_new_states, float *new_states, uint8_t *dones, void *stream);

/* Sprites for bs_render: RGBA8 DEVICE images in the sizes of the reference's assets (sprites.py:61,226:
 * plane 50 wide x 48 high, base 62 x 62), index 0 = red, 1 = blue. */
typedef struct bs_sprites {
    const uint8_t *plane[2]; /* [48][50][4] */
    const uint8_t *base[2];  /* [62][62][4] */
    /* Explosion animation (sprites.py:375-419): 9 frames scaled to 64 x 64, colour key black = alpha 0.
     * NULL or n_explosions == 0: none drawn.  explosions = DEVICE int32 [n][3] {x, y, frame}: the caller
     * keeps the list the reference keeps in self.explosions (one per killed plane, battle_env.py:356,
     * frame + 1 per drawn frame, sprites.py:405-419); entries with frame outside 0..8 are skipped. */
    const uint8_t *explosion_img;  /* [9][64][64][4] */
    const int32_t *explosions;
    int32_t n_explosions;          /* at most 64 are drawn */
} bs_sprites;

/* parallel_env.render (battle_env.py:498-560) for env `env` into rgb [800][1200][3] u8 (device), no
 * window: white background, bullets as 6x3 quads in the shooter's colour, both bases and the alive
 * planes (sprites rotated counter-clockwise by their direction about the rect centre, nearest
 * neighbour) with their hp bars; explosions (between bullets and bases, battle_env.py:534-536) from the
 * caller's list.  Left out: plane names / winner text (fonts), rounded bar corners.  Integer geometry (half-pixels, Q14 rotation): reproducible bit for bit. */
int bs_render(bs_handle *h, int64_t env, const bs_sprites *sprites, uint8_t *rgb, void *stream);

/* out[8] = {red wins, blue wins, ties, total_games, sum(red episode return)*2^20,
 * sum(blue episode return)*2^20, env ticks, agent steps}; synchronises `stream`. */
int bs_stats(bs_handle *h, int64_t out[8], void *stream);
/* Per-episode score log (row f3; score_dict of main.py:219-221, scores.json): attach a DEVICE ring
 * log [capacity][3] f32, capacity a power of two (NULL detaches).  Every game that ends in any kernel appends {red score,
 * blue score, code} at (count mod capacity); code 1 = red won, 2 = blue won, 3 = tie, 6 = both bases
 * fell in the same tick (winner 'blue', both wins counted, battle_env.py:363-372).  The order of
 * games that end in the same launch is unspecified.  bs_episode_log_count synchronises `stream`. */
int bs_episode_log(bs_handle *h, float *log, int64_t capacity);
int bs_episode_log_count(bs_handle *h, int64_t *count, void *stream);

/* Device address of the same 8 int64 counters (for an NCCL all-reduce without a host hop). */
int bs_stats_device(bs_handle *h, int64_t **dev_counters);

int bs_get_state(bs_handle *h, const bs_state_view *host_view);
int bs_set_state(bs_handle *h, const bs_state_view *host_view);

/* ---- host-buffer entry points (same meaning; pinned staging + copies inside) ------------- */
int bs_reset_host(bs_handle *h, const uint8_t *env_mask, const int32_t *draws, float *obs);
int bs_step_host(bs_handle *h, const void *actions, int action_dtype, const double *draws,
                 float *obs, float *rew, uint8_t *done, uint8_t *env_done, int8_t *winner);
int bs_step_team_host(bs_handle *h, int learner_team, const void *team_actions, int action_dtype,
                      int other_policy, const double *draws, float *obs, float *rew, uint8_t *done,
                      uint8_t *env_done, int8_t *winner);
int bs_instinct_actions_host(bs_handle *h, int team, const float *obs, const double *draws,
                             void *actions_out);
/* ext_actions is a HOST buffer [T,E,N,A] (or NULL); replay rows stay on the device; the eight
 * counters are copied back into stats_out (may be NULL). */
int bs_rollout_fused_host(bs_handle *h, int T, int learner_team, int policy_red, int policy_blue,
                          const float *ext_actions, const bs_replay *replay, int64_t cursor,
                          int64_t stats_out[8]);
/* The same call in two phases, for a caller that prepares the next action block while the device works
 * (main.py:178-213 with the actors on the host): _begin returns once ext_actions has been copied (the host
 * buffer may be refilled; the kernels may still run), _end waits for the OLDEST call in flight and returns the
 * counters as they stood after it.  At most two calls may be in flight (BS_E_INVALID otherwise); the action
 * copies of call k+1 then overlap the kernels of call k. */
int bs_rollout_fused_host_begin(bs_handle *h, int T, int learner_team, int policy_red, int policy_blue,
                                const float *ext_actions, const bs_replay *replay, int64_t cursor);
int bs_rollout_fused_host_end(bs_handle *h, int64_t stats_out[8]);

#ifdef __cplusplus
}
#endif
#endif /* BATTLESPACE_H */
