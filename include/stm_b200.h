/*
 * stm_b200.h — C ABI of libstm_b200.so: the B200-native tree-expansion hot path of SoftTreeMax.
 *
 * Boundary rules:
 *   - plain pointers and sizes only; no torch / C++ types cross this boundary;
 *   - unless a function says "host", every data pointer is a DEVICE pointer owned by the caller
 *     (PyTorch allocates, exactly as the reference passes `tensor.data_ptr()` integers into
 *     torchcule, policies/cule_bfs.py:154-167); the ctx owns only its cartridge image, decode
 *     table, intermediate-level pools and native-frame scratch;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, no internal
 *     streams, no host synchronisation unless stated;
 *   - every entry point returns 0 (STM_OK) or a negative stm_status and never throws;
 *     the message is available from stm_last_error();
 *   - one ctx per device, not re-entrant;
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails with
 *     STM_ERR_CUDA.
 */
#ifndef STM_B200_H
#define STM_B200_H

#include <stddef.h>
#include <stdint.h>
#include "stm_state.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct stm_ctx stm_ctx;

enum stm_status { STM_OK = 0, STM_ERR_ARG = -1, STM_ERR_CUDA = -2, STM_ERR_NOMEM = -3 };

/* Replaces torchcule.atari.Rom(env_name) + Env(num_envs, device, action_set, ...) construction:
 * policies/cule_bfs.py:29-39,55-62.  `rom` is a HOST pointer to desc->rom_size bytes. */
int stm_ctx_create(stm_ctx **out, const stm_game_desc *desc, const uint8_t *rom, int device);
int stm_ctx_destroy(stm_ctx *ctx);
/* last error of `ctx`, or of the last failed stm_ctx_create when ctx == NULL */
const char *stm_last_error(const stm_ctx *ctx);
/* size of one env state blob (include/stm_state.h): S in the roofline arithmetic */
size_t stm_state_bytes(const stm_ctx *ctx);
/* size of the minimal action set: torchcule Rom.minimal_actions(), cule_bfs.py:29-31 */
int stm_num_actions(const stm_ctx *ctx);

/* Colour-index -> grey table of the observation path (`generate_frames` in grey mode, cule_bfs.py:162-167): 128 bytes, entry
 * i = the 8-bit grey of NTSC colour index i = COLUxx >> 1.  `grey128` is a HOST pointer.  It is DATA of the ctx, not of the
 * build: the default (include/stm_palette.h) is computed from a YIQ formula because ALE's NTSC table — what CuLE's grey path
 * uses — is not available offline; a user who has that table installs it here and nothing is recompiled.  Synchronises the
 * device.  (The 160x210 -> 84x84 resize rule is still fixed at build time: DESIGN.md §2 lists what is unpinned.) */
int stm_set_palette(stm_ctx *ctx, const uint8_t *grey128);

/* Power-on state written to `states[0..n)`: replaces `super(AtariEnv, env).reset(0)`,
 * cule_bfs.py:57.  rewards/done/lives may be NULL. */
int stm_reset_states(stm_ctx *ctx, uint8_t *states, float *rewards, uint8_t *done, int32_t *lives,
                     int n, void *stream);

/* K1 — clone every env `fanout` times: child i <- parent i / fanout.
 * Replaces `depth_env.expand(num_envs)` (cule_bfs.py:143) and `new_expand` (:81-96).
 * src/dst must not overlap.  rewards/done/lives pairs may be NULL together. */
int stm_expand(stm_ctx *ctx, const uint8_t *src_states, const float *src_rewards,
               const uint8_t *src_done, const int32_t *src_lives, uint8_t *dst_states,
               float *dst_rewards, uint8_t *dst_done, int32_t *dst_lives, int n_src, int fanout,
               void *stream);

/* K2 — one agent step for n envs = `frameskip` emulated frames, with the per-frame
 * reward/done/lives update fused in.  Replaces the loop cule_bfs.py:152-158
 * (`super(AtariEnv, env).step(...)` + `get_data(episodic_life, gamma**depth, done, rewards, lives)`).
 *   actions  : n policy-action indices (0..A-1), or NULL for "env i takes action i % A"
 *              (cule_bfs.py:144);
 *   weight   : gamma ** depth, rewards[i] += weight * r_frame;
 *   render   : non-zero -> the last two frames are kept as native 160x210 frames in the ctx
 *              scratch for stm_frames (cule_bfs.py:162-167 `generate_frames` at skip index
 *              frameskip-2 and frameskip-1).
 * An env whose done flag is set on entry is frozen.  n must not exceed stm_max_step_envs(). */
int stm_step(stm_ctx *ctx, uint8_t *states, const uint8_t *actions, int n, int frameskip,
             int episodic_life, float weight, float *rewards, uint8_t *done, int32_t *lives,
             int render, void *stream);
/* how many envs one stm_step(render=1) + stm_frames pair can hold in the native-frame scratch;
 * stm_set_scratch_envs changes it (reallocates; not for use inside a timed loop). */
int stm_max_step_envs(const stm_ctx *ctx);
int stm_set_scratch_envs(stm_ctx *ctx, int n_envs);

/* K3 — observation kernel: grey + 84x84 bilinear resize of the two native frames left by the last
 * stm_step(render=1), per-pixel max, and frame stacking:
 *   child_stacks[i] = concat(parent_stacks[i / fanout][1:4], new_frame_i)       (u8, 4x84x84)
 * Replaces generate_frames' grey/rescale stage + cule_bfs.py:168-172
 * (`th.max`, `replicate_state`, `th.cat`).  fanout = 1 with parent = the env's previous stack gives the
 * plain frame-stack update of a root env (environments/cule_env.py:83,93).  parent_stacks and child_stacks
 * must NOT overlap (the kernel moves the three older planes from many thread blocks at once): an
 * overlapping pair is refused with STM_ERR_ARG.  The same holds for stm_step_frames_noalloc. */
int stm_frames(stm_ctx *ctx, const uint8_t *parent_stacks, uint8_t *child_stacks, int n, int fanout,
               void *stream);

/* One emulated FRAME (not a whole agent step) + that frame's reward/done/lives update, for hosts that drive the emulator
 * frame by frame through raw pointers as the reference does: `super(AtariEnv, env).step(fire_reset, False, actions_ptr,
 * 0, done_ptr, 0)` followed by `env.get_data(episodic_life, gamma**depth, done_ptr, rewards_ptr, lives_ptr)`
 * (cule_bfs.py:154-158).  Differences from stm_step(frameskip = 1): `action_codes` are ALE action codes (what
 * `action_set.repeat(...)` holds, cule_bfs.py:144), not indices into the minimal set; and `frozen` (n bytes, or NULL =
 * done[]) names the envs that sit the frame out — pass the done flags as they were when the AGENT step began and the
 * four calls of an agent step equal one stm_step(frameskip = 4) bit for bit (an env that finishes in frame 1 of 4 still
 * plays frames 2 and 3, as inside stm_step).  The frame is always kept for stm_frame_obs. */
int stm_step_frame(stm_ctx *ctx, uint8_t *states, const uint8_t *action_codes, int n, int episodic_life, float weight,
                   float *rewards, uint8_t *done, int32_t *lives, const uint8_t *frozen, void *stream);
/* Grey 84x84 observation of the frame the last stm_step_frame (or stm_step with render) emulated, n x 84 x 84 u8; an env
 * that sat the frame out gets zeros.  Replaces `env.generate_frames(rescale, False, num_channels, obs_ptr)`
 * (cule_bfs.py:162-167); the max over two such frames is the caller's (`th.max`, cule_bfs.py:168). */
int stm_frame_obs(stm_ctx *ctx, uint8_t *obs, int n, void *stream);

/* Whole breadth-first expansion of n_roots roots to `depth` levels: cule_bfs.py:98-209, batched
 * over roots (reference: one root per call).
 *   root_states  n_roots x 256                 (cule_bfs.py:110-112)
 *   root_stacks  n_roots x 4 x 84 x 84 u8      (cule_bfs.py:99)
 *   leaf_stacks  n_roots x A^depth x 4 x 84 x 84 u8
 *   leaf_rewards n_roots x A^depth f32, sign() applied when clip_reward (cule_bfs.py:204-206)
 *   leaf_done    n_roots x A^depth u8 ; leaf_lives i32 ; leaf_states x256 — each may be NULL.
 * Leaf j of root r sits at r*A^depth + j; its first action is j / A^(depth-1).
 * Everything is ordered on `stream`; inside the call a second stream of the ctx (forked from and joined back into
 * `stream` by events, so the call may be captured into a CUDA graph) copies the three older planes of every child
 * stack while the level is being emulated. */
int stm_bfs(stm_ctx *ctx, const uint8_t *root_states, const uint8_t *root_stacks, int n_roots,
            int depth, float gamma, int frameskip, int episodic_life, int clip_reward,
            uint8_t *leaf_stacks, float *leaf_rewards, uint8_t *leaf_done, int32_t *leaf_lives,
            uint8_t *leaf_states, void *stream);

/* stm_bfs for root stacks that are still on their way: `stacks_ready` is a CUDA event (cudaEvent_t) after which
 * root_stacks holds the roots' observation stacks — e.g. recorded behind their host-to-device copy on another stream.
 * Only the copy of the older planes into the first level waits for it; the emulation of the tree starts at once, as
 * stm_bfs_host does with its own upload.  NULL = stm_bfs.  No reference counterpart (the reference's roots live on the
 * device of the env, cule_bfs.py:99,110-112). */
int stm_bfs_after(stm_ctx *ctx, const uint8_t *root_states, const uint8_t *root_stacks, int n_roots,
                  int depth, float gamma, int frameskip, int episodic_life, int clip_reward,
                  uint8_t *leaf_stacks, float *leaf_rewards, uint8_t *leaf_done, int32_t *leaf_lives,
                  uint8_t *leaf_states, void *stream, void *stacks_ready);

/* Continue a tree from interior nodes: `n_nodes` nodes of level `level0` (their emulator states, observation
 * stacks, accumulated UNCLIPPED discounted rewards, done flags and lives, e.g. a slice of what stm_bfs returned
 * with clip_reward = 0) are expanded `depth_more` further levels exactly as stm_bfs would have expanded them:
 * child order, frozen done nodes, reward weight gamma^(level0 + k), sign() of the total at the end when clip_reward.
 * The reference has no counterpart (it always expands the whole tree on one device, cule_bfs.py:129-173); this is
 * the leaf-block partition of one tree over several GPUs: every rank expands the first levels, keeps its
 * contiguous block of level-`level0` nodes and resumes from there (softtreemax_b200/shard.py). */
int stm_bfs_resume(stm_ctx *ctx, const uint8_t *node_states, const uint8_t *node_stacks, const float *node_rewards,
                   const uint8_t *node_done, const int32_t *node_lives, int n_nodes, int level0, int depth_more,
                   float gamma, int frameskip, int episodic_life, int clip_reward, uint8_t *leaf_stacks,
                   float *leaf_rewards, uint8_t *leaf_done, int32_t *leaf_lives, uint8_t *leaf_states, void *stream);

/* Same, with HOST buffers (pinned or pageable) for the roots and for rewards/done/lives: copies
 * the roots host->device, runs stm_bfs, copies the results device->host and synchronises
 * `stream` before returning.  This is the call a non-CUDA host program makes, and it mirrors the
 * reference boundary: `bfs(state, depth)` takes the root observation from host memory
 * (environments/cule_env.py:93 hands SB3 numpy) and returns the leaves ON THE GPU
 * (cule_bfs.py:170,204,209).  leaf_stacks is a DEVICE pointer when leaf_stacks_on_device != 0
 * (the reference behaviour), else a HOST pointer that receives a full copy of the leaf stacks. */
int stm_bfs_host(stm_ctx *ctx, const uint8_t *root_states, const uint8_t *root_stacks, int n_roots,
                 int depth, float gamma, int frameskip, int episodic_life, int clip_reward,
                 uint8_t *leaf_stacks, int leaf_stacks_on_device, float *leaf_rewards,
                 uint8_t *leaf_done, int32_t *leaf_lives, void *stream);

/* K5 — SoftTreeMax logits for R roots of a full-width tree (L = A^d leaves each), fused:
 * m = val_coef*head + reward; per root action (L/A consecutive leaves x A head outputs) logsumexp(beta*m)
 * - log(count) [cumulative == 0] or beta*mean(m) [cumulative != 0]; NaN -> 0; alpha-mix with the depth-0
 * logits; eps-regularisation log((1-eps)*softmax + eps/A).
 * Replaces policies/actor_critic_ts.py:60-95 and :147-188 and utils.py:52-57.
 *   head R x L x A f32, rewards R x L f32, depth0 R x A f32 or NULL (alpha-mix term treated as 0),
 *   values R x L f32 or NULL; logits R x A (out), raw R x A or NULL (z before the mix),
 *   vmax R or NULL (max_l val_coef*values + reward: the use_leaves_v root value, :62-63);
 *   group R x L i32 or NULL: the root action of every leaf of a width-limited tree (the reference's
 *   first_action, actor_critic_ts.py:70-79,163-171) — then count = members*A, actions without a
 *   surviving leaf get -inf, and the reference's zero-filled scatter is reproduced (each non-member
 *   slot adds exp(0) to the E-mode sum).  NULL = full-width tree, leaf l belongs to action l/(L/A). */
int stm_logits_fwd(stm_ctx *ctx, const float *head, const float *rewards, const float *depth0,
                   const float *values, const int *group, int R, int L, int A, float val_coef, float beta,
                   float alpha, float eps, int cumulative, float *logits, float *raw, float *vmax,
                   void *stream);
/* Backward of stm_logits_fwd w.r.t. head, depth0, beta and alpha (what evaluate_actions needs from
 * autograd, actor_critic_ts.py:106-194).  dhead R x L x A, ddepth0 R x A or NULL, dbeta / dalpha: R
 * per-root partial sums (the caller adds them up). */
int stm_logits_bwd(stm_ctx *ctx, const float *head, const float *rewards, const float *depth0,
                   const int *group, int R, int L, int A, float val_coef, float beta, float alpha, float eps,
                   int cumulative, const float *dlogits, float *dhead, float *ddepth0, float *dbeta,
                   float *dalpha, void *stream);

/* K4 — NatureCNN + heads.  stm_cnn_set_weights takes DEVICE pointers to fp32 parameters in the
 * layout SB3 1.4.0's ActorCriticCnnPolicy holds them (features_extractor.cnn.{0,2,4}, .linear.0,
 * action_net, value_net: conv (out,in,kh,kw), linear (out,in)) and re-tiles them to bf16 for the
 * tensor cores (call again after every optimizer step; synchronises `stream`; allocates on the first call only).
 * stm_cnn_forward evaluates n observation stacks (n x 4 x 84 x 84 u8, as stm_bfs / stm_frames write
 * them; the /255 of SB3's preprocess_obs is applied inside) and writes heads n x (A+1) f32:
 * columns 0..A-1 = action_net(features), column A = value_net(features).
 * Replaces extract_features -> mlp_extractor -> action_net / value_net,
 * policies/actor_critic_ts.py:199-218 (cuDNN/cuBLAS fp32 there; bf16 x bf16 -> f32 tcgen05 here). */
int stm_cnn_set_weights(stm_ctx *ctx, const float *conv1_w, const float *conv1_b, const float *conv2_w,
                        const float *conv2_b, const float *conv3_w, const float *conv3_b,
                        const float *fc_w, const float *fc_b, const float *action_w,
                        const float *action_b, const float *value_w, const float *value_b,
                        int n_actions, void *stream);
int stm_cnn_forward(stm_ctx *ctx, const uint8_t *obs, long long n, float *heads, void *stream);

/* K4, update path.  policies/actor_critic_ts.py:139 runs `extract_features(th.cat(all_leaves_obs))` under autograd on
 * 32*L + 32 images per PPO minibatch (main.py:54) and SB3's PPO.train backpropagates through it (cuDNN / cuBLAS fp32 in the
 * reference).  stm_cnn_forward_train = stm_cnn_forward for one minibatch of n <= 32 768 images that also keeps the bf16
 * activations inside the ctx; stm_cnn_backward must follow it with the same obs and n and turns dL/dheads (n x (A+1) f32,
 * columns as stm_cnn_forward writes them) into dL/dparameter for the twelve tensors stm_cnn_set_weights took, in their torch
 * layouts (conv (out, in, kh, kw), linear (out, in)), OVERWRITING the twelve device buffers.  Data- and weight-gradient
 * contractions run on the tensor cores (tcgen05, bf16 operands, fp32 accumulation; split-K for the weight gradients).
 * One minibatch in flight per ctx; a rollout-time stm_cnn_forward in between does not disturb it. */
int stm_cnn_forward_train(stm_ctx *ctx, const uint8_t *obs, long long n, float *heads, void *stream);
int stm_cnn_backward(stm_ctx *ctx, const uint8_t *obs, long long n, const float *dheads, float *g_conv1_w, float *g_conv1_b,
                     float *g_conv2_w, float *g_conv2_b, float *g_conv3_w, float *g_conv3_b, float *g_fc_w, float *g_fc_b,
                     float *g_action_w, float *g_action_b, float *g_value_w, float *g_value_b, void *stream);

/* (No counterpart in the reference: the reference never manages device memory itself.)
 * stm_step(render=1) followed by stm_frames, for callers that must not cause a device allocation:
 * hosts whose own pool already owns all GPU memory, so a cudaMalloc inside the library could only
 * fail, and latency-critical loops that cannot take cudaMalloc's device-wide synchronisation.
 * The native frames between the two kernels live in a scratch array built into the library image
 * (about 8.6 MB on every GPU that loads the library, held until exit) instead of in allocated
 * memory, so n is limited to STM_NOALLOC_MAX_ENVS; a larger n is refused with STM_ERR_ARG.
 * The call itself never allocates or frees.  The first call on a device also looks up the device
 * address of that array, so make that first call before starting any stream capture.
 * All calls share the one array and it keeps the last call's contents (nothing in it needs
 * clearing: everything read was written earlier in the same call); calls that overlap on the GPU
 * race on it — issue them one after another.  Arguments as for stm_step and stm_frames. */
#define STM_NOALLOC_MAX_ENVS 128
int stm_step_frames_noalloc(stm_ctx *ctx, uint8_t *states, const uint8_t *actions, int n,
                            int frameskip, int episodic_life, float weight, float *rewards,
                            uint8_t *done, int32_t *lives, const uint8_t *parent_stacks,
                            uint8_t *child_stacks, int fanout, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* STM_B200_H */
