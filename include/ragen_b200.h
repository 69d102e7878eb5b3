/*
 * ragen_b200.h — C ABI of libragen_b200.so (hand-written sm_100a CUDA kernels for RAGEN's rollout
 * hot path).  This is the drop-in boundary: plain pointers and sizes, no torch / C++ types.
 *
 * Conventions
 *   - every function returns 0 on success, a negative RAGEN_E* code otherwise; the message is in
 *     ragen_last_error() (thread-local).  No exceptions cross the ABI.
 *   - every pointer not marked [host] is a DEVICE pointer owned by the caller (e.g. a torch tensor's
 *     data_ptr); the library never allocates user-visible memory.  The only internal allocations
 *     are scratch buffers owned by the opaque ragen_ctx (generator search workspace).
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream).  All work is
 *     asynchronous on that stream; nothing synchronises unless stated.
 *   - one ragen_ctx per host thread / stream.
 *
 * The reference has no FFI for this path (pure Python); each entry point cites the reference
 * function(s) it replaces.  The ctypes binding a maintainer would add is in INTEGRATION.md.
 */
#ifndef RAGEN_B200_H
#define RAGEN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RAGEN_OK 0
#define RAGEN_EINVAL (-1)   /* bad argument */
#define RAGEN_ECUDA (-2)    /* CUDA runtime error (message has cudaGetErrorString) */
#define RAGEN_ENOMEM (-3)   /* workspace allocation failed */

typedef struct ragen_ctx ragen_ctx;

/* ---- context ---------------------------------------------------------------------------- */
int ragen_create(int device, ragen_ctx** out);
int ragen_destroy(ragen_ctx* ctx);
const char* ragen_last_error(void);
int ragen_version(void);
/* number of kernels this ctx has launched so far (bench.py reports it as gpu_launches) */
uint64_t ragen_launch_count(const ragen_ctx* ctx);
int ragen_sm_count(const ragen_ctx* ctx);

/* ---- Sokoban ---------------------------------------------------------------------------- */
/* Cell codes (ragen/env/sokoban/utils.py:563-570): 0 wall 1 floor 2 target 3 box-on-target 4 box
 * 5 player.  Rooms are row-major uint8[n][H*W]; `fixed` holds codes 0..2 only.  Rooms must have a
 * wall border (every generated room does, utils.py:352-353).
 * Glyph rows: uint8[n][H*(W+1)], codes 0..5 as above, 6 = player on target, 7 = newline; the last
 * byte of each env (position H*(W+1)-1) is 7 as well and is not part of the reference string
 * (which has no trailing newline, sokoban/env.py:56). */

/* Per-env scalars live in ONE 8-byte record so a step moves them with a single 64-bit access:
 *   int16 meta[n][4]:  [0] player cell index r*W+c (int16)      [1] num_env_steps (uint16)
 *                      [2] low byte boxes_on_target (int8, as the reference stores it: 0 after reset),
 *                          high byte 1 if the player stands on a target (room_fixed[player]==2)
 *                      [3] low byte num_boxes - #uncovered targets of the current board (int8), high byte 0
 * room_fixed is read only by ragen_sokoban_adopt: it verifies that room_fixed is a function of room_state
 * (+ the one bit above), which holds for every room the reference generates, so step/render never
 * touch it again. */

/* SokobanEnv.step (ragen/env/sokoban/env.py:44-51) -> gym_sokoban SokobanEnv.step/_push/_move/
 * _calc_reward/_check_if_done, fused with SokobanEnv.render text branch (env.py:53-57), for up to
 * K actions per env in one launch (EnvStateManager._execute_actions, es_manager.py:116-128: stop
 * at the first `done`, rewards summed in order as doubles).
 *   actions    [n][K]  1..4 = Up/Down/Left/Right (sokoban/config.py:12), 0 = gym no-op step
 *   n_actions  [n] or NULL  how many of the K actions to attempt for env i (NULL = K for all)
 *   state      [n][H*W] in/out        meta [n][4] in/out (see above)
 *   acc_reward [n] out  sum of step rewards (0.0 when nothing executed)
 *   done       [n] out  1 if the last executed step returned done (0 when nothing executed)
 *   flags      [n] out  bit0 action_is_effective, bit1 success (of the last executed step),
 *                       bit2 at least one step executed
 *   n_executed [n] out
 *   obs        [n][H*(W+1)] out glyph rows after the last executed step (NULL = skip render)   */
int ragen_sokoban_step(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                       const uint8_t* actions, int K, const uint8_t* n_actions,
                       uint8_t* state, int16_t* meta,
                       double* acc_reward, uint8_t* done, uint8_t* flags, uint8_t* n_executed,
                       uint8_t* obs, void* stream);

/* The same step with the glyph rows written as two 4-bit codes per byte (low nibble = the earlier cell):
 * obs_packed [n][H*(W+1)/2], H*(W+1) must be even.  Half the bytes to write and to send to the host; a host turns a
 * byte into its two glyph strings with one 256-entry translate, exactly as it turns a code into one. */
int ragen_sokoban_step_packed(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                              const uint8_t* actions, int K, const uint8_t* n_actions,
                              uint8_t* state, int16_t* meta,
                              double* acc_reward, uint8_t* done, uint8_t* flags, uint8_t* n_executed,
                              uint8_t* obs_packed, void* stream);
int ragen_sokoban_render_packed(ragen_ctx* ctx, int64_t n, int H, int W, const uint8_t* state,
                                const int16_t* meta, uint8_t* obs_packed, void* stream);

/* The same step / render with THREE-BIT glyph rows: the H*W cell codes (0..6), row-major, no separators (the reader
 * knows W), as a little-endian bit stream (cell i in bits 3i..3i+2) padded with zero bits to an even number of bytes:
 * obs_packed3 [n][((3*H*W + 15) / 16) * 2] - 38 B for 10x10, 14 B for 6x6.  The bulk-transport format: where a step's
 * results go back to host memory the device->host link bounds the step, and this is 31 % fewer bytes than the nibble
 * rows.  Built for 6x6, 8x8 and 10x10 rooms (other shapes: RAGEN_EINVAL). */
int ragen_sokoban_step_packed3(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                               const uint8_t* actions, int K, const uint8_t* n_actions,
                               uint8_t* state, int16_t* meta,
                               double* acc_reward, uint8_t* done, uint8_t* flags, uint8_t* n_executed,
                               uint8_t* obs_packed3, void* stream);
/* The step with the rooms kept as NIBBLES in HBM: state4 [n][H*W/2], cell 2i in the low nibble of byte i (cells are
 * 0..5) - half the bytes of the dominant read-modify-write of a step; the tile is staged, stepped and rendered in that
 * form.  Glyph rows come out three-bit (obs_packed3, may be NULL).  Everything else as ragen_sokoban_step
 * (SokobanEnv.step, ragen/env/sokoban/env.py:44-60).  6x6, 8x8 and 10x10 rooms (other shapes: RAGEN_EINVAL).  The caller
 * converts between state and state4 when rooms are loaded / inspected (SokobanBatch(compact=True) does). */
int ragen_sokoban_step_compact(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                               const uint8_t* actions, int K, const uint8_t* n_actions,
                               uint8_t* state4, int16_t* meta,
                               double* acc_reward, uint8_t* done, uint8_t* flags, uint8_t* n_executed,
                               uint8_t* obs_packed3, void* stream);
/* The compact step with its results as DELTAS: a step changes at most three cells, so what a consumer that mirrors the rooms
 * needs is 4 bytes per executed action instead of reward + flags + a re-rendered room (49 B per 10x10 env) - the transport
 * format for callers that keep the results in host memory, where the device->host link bounds the step.
 *   delta [n][K][4]: byte 0 player cell before the action, byte 1 player cell after, byte 2 cell the pushed box went to
 *   (0xff: no push), byte 3 = reward code | done << 3 | action_is_effective << 4 | success << 5 | executed << 6, with reward code
 *   0: -0.1, 1: -0.1 + 1, 2: -0.1 - 1, 3: -0.1 + 1 + 10, 4: -0.1 + 10 (gym_sokoban _calc_reward; the caller adds the doubles in
 *   order, which reproduces acc_reward bit for bit).  Records of actions that were not executed are zero.
 * The vacated cell shows room_fixed again, the box lands as 3 on a target else 4 (the consumer has room_fixed from the reset).
 * Rooms as in ragen_sokoban_step_compact (6x6, 8x8, 10x10).  No glyph rows, no result arrays are written. */
int ragen_sokoban_step_delta(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                             const uint8_t* actions, int K, const uint8_t* n_actions,
                             uint8_t* state4, int16_t* meta, uint8_t* delta, void* stream);
/* The step on GLYPH ROOMS: state3 [n][((3*H*W + 15) / 16) * 2] holds every room as three-bit glyph codes (0..5 as room_state, 6 =
 * the player on a target; the layout of obs_packed3 above), so the room IS its rendered observation (SokobanEnv.render,
 * sokoban/env.py:53-57: where(state == 5 & fixed == 2, 6, state)): a step reads and writes 38 B per 10x10 env and nothing is
 * left to render.  room_state = the codes with 6 read as 5.  Results as ragen_sokoban_step, or - delta != NULL - as the 4-byte
 * records of ragen_sokoban_step_delta (the result arrays may then be NULL).  ragen_sokoban_adopt_slots writes state3.  6x6, 8x8, 10x10. */
int ragen_sokoban_step_glyph3(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                              const uint8_t* actions, int K, const uint8_t* n_actions,
                              uint8_t* state3, int16_t* meta,
                              double* acc_reward, uint8_t* done, uint8_t* flags, uint8_t* n_executed,
                              uint8_t* delta, void* stream);
/* The same step reporting ONE byte per action, delta1 [n][K]: byte 3 of the record above with bit 7 = "a box moved".  The consumer
 * sent the action and mirrors the player's cell, so "moved" (bit 4) and "a box moved" say the rest: the player went one cell in
 * the action's direction (1 up, 2 down, 3 left, 4 right: -W, +W, -1, +1) and the box, if any, one cell further.  Two bytes per
 * env-step cross the host link (one action up, one record down).  rooms_format 3: `rooms` is state4 as in
 * ragen_sokoban_step_delta; 4: state3 as in ragen_sokoban_step_glyph3.  Any room size the format covers (no cell index is sent). */
int ragen_sokoban_step_delta1(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                              const uint8_t* actions, int K, const uint8_t* n_actions, int rooms_format, uint8_t* rooms,
                              int16_t* meta, uint8_t* delta1, void* stream);
int ragen_sokoban_render_packed3(ragen_ctx* ctx, int64_t n, int H, int W, const uint8_t* state,
                                 const int16_t* meta, uint8_t* obs_packed3, void* stream);

/* SokobanEnv.render text branch only (env.py:53-57). */
int ragen_sokoban_render(ragen_ctx* ctx, int64_t n, int H, int W, const uint8_t* state,
                         const int16_t* meta, uint8_t* obs, void* stream);

/* SokobanEnv.render rgb_array branch (ragen/env/sokoban/env.py:58-59 -> gym_sokoban get_image -> room_to_rgb): every cell
 * receives the S x S x 3 sprite of its surface (room_state code; 6 = player on a target) in a uint8 image
 * out [m][H*S][W*S][3].  atlas uint8 [7][S][S][3] is caller data (upstream's 16 x 16 PNG sprites, which live in the
 * un-vendored gym_sokoban package).  rows int64 [n_rows] selects the envs to draw (image i = env rows[i], m = n_rows);
 * rows NULL draws all n envs.  S in [1, 64]; S = 16 with 16-byte aligned atlas/out takes the vector path. */
int ragen_sokoban_render_rgb(ragen_ctx* ctx, int64_t n, int H, int W, int S, const uint8_t* state,
                             const int16_t* meta, const uint8_t* atlas, const int64_t* rows, int64_t n_rows,
                             uint8_t* out, void* stream);

/* Post-reset bookkeeping of SokobanEnv.reset (env.py:37-38) for rooms already in device memory:
 * player = argwhere(state == 5)[0], num_env_steps = 0, boxes_on_target = 0 -> meta.
 * status[i] (may be NULL) = 0 ok, 1 no player cell, 2 border not wall, 3 room_fixed is not the
 * function of room_state described above (such rooms cannot come from the reference generator). */
int ragen_sokoban_adopt(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, const uint8_t* state,
                        const uint8_t* fixed, int16_t* meta, int32_t* status, void* stream);

/* The same for a SUBSET of a shared batch (BaseEnv.reset on one env of a manager-owned batch, ragen/env/base.py:16,
 * es_manager.py:93-95): env i is adopted only when only[i] != 0 (only NULL = all); the others keep their record
 * (num_env_steps, stored box count) and their status entry.  state4 (may be NULL) [n][H*W/2] receives the adopted rooms
 * as nibbles, the form ragen_sokoban_step_compact steps (H*W even); state3 (may be NULL) as three-bit glyph codes
 * (ragen_sokoban_step_glyph3). */
int ragen_sokoban_adopt_slots(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, const uint8_t* state,
                              const uint8_t* fixed, int16_t* meta, int32_t* status, const uint8_t* only,
                              uint8_t* state4, uint8_t* state3, void* stream);

/* SokobanEnv.reset(seed) -> generate_room (ragen/env/sokoban/env.py:28-42, ragen/env/sokoban/utils.py:221-278 with
 * room_topology_generation :281-355, place_boxes_and_player :358-397, reverse_playing / depth_first_search /
 * reverse_move :408-542, box_displacement_score :545-560, add_random_player_movement :152-213) for a batch of seeds.
 * Stream-exact: seeds[i] seeds CPython's `random` and NumPy's legacy `np.random` exactly as all_seed(seed)
 * (ragen/utils.py:7-18) does, and the search visits states in the reference's order, so fixed/state [n][H*W] are
 * the room_fixed / room_state the reference builds for that seed.  num_gen_steps is gym_sokoban's
 * int(1.7 * (H + W)).  status[i] = 0 ok, 1 "Not enough free spots" (RuntimeError in the reference), 2 "score == 0"
 * after 4 tries (RuntimeWarning), 3 seed >= 2^32 (np.random.seed raises); on non-zero status the room is undefined
 * and the caller re-seeds as SokobanEnv.reset does (env.py:40-42).  explored (may be NULL) [n] = states explored by
 * the last search.  H*W <= 256, num_boxes <= 4.  Uses ctx-owned workspace (grown on first use). */
int ragen_sokoban_generate(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int search_depth,
                           int num_gen_steps, const uint64_t* seeds, uint8_t* fixed, uint8_t* state,
                           int32_t* status, int32_t* explored, void* stream);

/* get_shortest_action_path(room_fixed, room_state, MAX_DEPTH) (ragen/env/sokoban/utils.py:11-101) for n rooms: the
 * reference's breadth-first search (moves tried up, down, left, right; duplicates dropped; a path longer than max_depth
 * ends the search) with its discovery order, so path [n][max_depth] (may be NULL) receives the reference's action list
 * (ids 1..4) in its first length[n] bytes.  status [n]: 0 solved (length 0 = already solved: the reference returns []
 * for that too), 1 no solution within max_depth (the reference's other []), 2 more than max_states distinct states
 * (0 = 131072, enough for every 10x10 / 2-box room), 3 no player / more than 4 boxes.  explored (may be NULL) [n] =
 * states popped.  Uses ctx-owned workspace (grown on first use). */
int ragen_sokoban_solve(ragen_ctx* ctx, int64_t n, int H, int W, const uint8_t* fixed, const uint8_t* state,
                        int max_depth, int64_t max_states, int32_t* status, int32_t* length, uint8_t* path,
                        int32_t* explored, void* stream);

/* ---- FrozenLake -------------------------------------------------------------------------- */
/* Maps are row-major uint8[n][size*size] with tile codes 0 'S' 1 'F' 2 'H' 3 'G' (size 2..8).  Glyph rows are
 * uint8[n][size*(size+1)]: 0 player, 1 frozen (S is drawn as F), 2 hole, 3 goal, 4 player in hole, 5 player on
 * goal (ragen/env/frozen_lake/config.py:17-20), 7 = row separator (also the last byte, not part of the string).
 * Every env owns NumPy's PCG64 generator: rng is uint64[2][n][2] = a plane of 128-bit states {hi, lo} followed
 * by a plane of 128-bit increments {hi, lo} (16-byte aligned). */

/* FrozenLakeEnv.reset(seed) (ragen/env/frozen_lake/env.py:28-37): generate_random_map(size, p, seed)
 * (frozen_lake/utils.py:25-47, with the is_valid reachability test :6-22) from Generator(PCG64(SeedSequence(seed))),
 * then gymnasium's reset: the env generator is re-seeded with the same seed and draws one uniform for the start
 * state.  status[i] (may be NULL) = 0 ok, 1 no valid map within 100000 tries.  obs may be NULL. */
int ragen_frozenlake_reset(ragen_ctx* ctx, int64_t n, int size, double p, const uint64_t* seeds,
                           uint8_t* tiles, uint8_t* pos, uint64_t* rng, int32_t* status, uint8_t* obs,
                           void* stream);

/* FrozenLakeEnv.step (env.py:39-45) -> gymnasium FrozenLakeEnv.step, fused with the text render (env.py:47-57),
 * up to K actions per env as in ragen_sokoban_step.  actions 1..4 = Left/Down/Right/Up (config.py:15,22); any
 * other value ends the env's action list (the reference raises KeyError).  One uniform is drawn per executed
 * step.  acc_reward/done/flags/n_executed/obs as in ragen_sokoban_step (flags bit1 = standing on the goal). */
int ragen_frozenlake_step(ragen_ctx* ctx, int64_t n, int size, int is_slippery, const uint8_t* actions, int K,
                          const uint8_t* n_actions, const uint8_t* tiles, uint8_t* pos, uint64_t* rng,
                          double* acc_reward, uint8_t* done, uint8_t* flags, uint8_t* n_executed,
                          uint8_t* obs, void* stream);

int ragen_frozenlake_render(ragen_ctx* ctx, int64_t n, int size, const uint8_t* tiles, const uint8_t* pos,
                            uint8_t* obs, void* stream);

/* ---- Bandit ------------------------------------------------------------------------------ */
/* BanditEnv.reset(seed) (ragen/env/bandit/env.py:50-60): lo_first[i] = 1 when action id `start` is the lo arm
 * (_randomize_arms :24-38).  rng layout as above. */
int ragen_bandit_reset(ragen_ctx* ctx, int64_t n, const uint64_t* seeds, uint64_t* rng, uint8_t* lo_first,
                       void* stream);
/* BanditEnv.step (env.py:62-77): reward[i] = lo_arm_score, or hi_arm_hiscore with probability
 * hi_arm_hiscore_prob else hi_arm_loscore (one uniform, hi arm only).  active (may be NULL) masks envs; an action
 * outside {start, start+1} executes nothing (the reference asserts).  flags: bit0 effective, bit1 success (hi arm
 * pulled), bit2 executed. */
int ragen_bandit_step(ragen_ctx* ctx, int64_t n, int action_space_start, double lo_arm_score,
                      double hi_arm_loscore, double hi_arm_hiscore, double hi_arm_hiscore_prob,
                      const uint8_t* actions, const uint8_t* active, uint64_t* rng, const uint8_t* lo_first,
                      double* reward, uint8_t* flags, void* stream);

/* ---- response parsing / action lookup / Countdown scoring ---------------------------------- */
/* ContextManager._parse_response (ragen/llm_agent/ctx_manager.py:148-173) + EnvStateManager.
 * _extract_map_valid_actions (ragen/llm_agent/es_manager.py:230-240) over a batch of responses.
 *   text, offsets   UTF-8 blob and int64 [n+1] byte offsets (response i = text[offsets[i], offsets[i+1]))
 *   action_sep      [host] NUL-terminated separator, 1..8 bytes ("||")
 *   lookup tables   table t = entries [table_first[t], table_first[t+1]); entry k has name
 *                   names[name_off[k], name_off[k+1]) and action id codes[k]; table_id[i] selects the table of
 *                   response i (-1 = env without action_lookup; NULL = table 0 for all).  Names match
 *                   case-insensitively over ASCII letters.
 *   scratch         blob of the same size as text: cleaned group contents that are not a sub-span of the input
 *   flags      [n]  bit0 regex matched, bit1 think span lives in scratch, bit2 answer span and action spans live
 *                   in scratch, bit3 more than max_actions actions were found (the answer text of llm_response is
 *                   " <sep> ".join(kept actions)), bit4 codes/n_valid are meaningful
 *   spans      [n][4]  think start, think length, answer start, answer length (bytes, relative to offsets[i])
 *   act_spans  [n][max_actions][2]  start, length of each kept action string
 *   n_parsed   [n]  kept actions (<= max_actions);  n_total [n] non-empty actions before the cut
 *   codes_out  [n][max_actions], n_valid [n]  ids of the kept actions present in the table, in order */
int ragen_parse_actions(ragen_ctx* ctx, int64_t n, const uint8_t* text, const int64_t* offsets,
                        int enable_think, const char* action_sep, int max_actions, const uint8_t* names,
                        const int32_t* name_off, const uint8_t* codes, const int32_t* table_first, int n_tables,
                        const int32_t* table_id, uint8_t* scratch, uint8_t* flags, int32_t* spans,
                        int32_t* act_spans, uint8_t* n_parsed, int32_t* n_total, uint8_t* codes_out,
                        uint8_t* n_valid, void* stream);

/* CountdownEnv.compute_reward (ragen/env/countdown/env.py:69-78): result[i] = 0 check_format failed, 1 format ok
 * but eval != target (or raised), 2 correct.  Answer i = text[start[i], start[i]+len[i]); nums int32 [n][8] with
 * n_nums[i] used.  unsupported[i] (may be NULL) = 1 when the expression leaves the evaluator's grammar (ints,
 * + - * /, unary + -, parentheses, blanks): such answers are scored as if eval() had raised. */
int ragen_countdown_score(ragen_ctx* ctx, int64_t n, const uint8_t* text, const int64_t* start,
                          const int32_t* len, const int32_t* nums, const int32_t* n_nums, const int64_t* target,
                          uint8_t* result, uint8_t* unsupported, void* stream);

/* ---- ContextManager tensor passes ---------------------------------------------------------- */
/* get_masks_and_scores (ragen/llm_agent/ctx_manager.py:35-70).  input_ids int64 [B][L]; outputs [B][L-1].
 * turn_scores double [B][max_turns] with n_turns[i] valid entries per row (the Python floats of
 * history[k]['reward']); use_turn_scores=0 puts sum(turn_scores[i]) on the last column, =1 puts score k on the
 * <|im_end|> of turn indicator 2k+3 (last column when the row has none), then rolls by one when roll_scores
 * (the Qwen branch :60-62). */
int ragen_masks_scores(ragen_ctx* ctx, int B, int L, const int64_t* input_ids, int64_t im_start_id,
                       int64_t im_end_id, int use_turn_scores, int enable_response_mask, int roll_scores,
                       const double* turn_scores, const int32_t* n_turns, int max_turns, uint8_t* loss_mask,
                       uint8_t* response_mask, float* score, void* stream);

/* get_lm_inputs' tensor assembly (ragen/llm_agent/ctx_manager.py:269-280: apply_chat_template -> tokenizer(texts, padding=True,
 * padding_side="left") -> position_ids = attention_mask.cumsum(-1)) as a gather of pre-tokenised segments.  pool int32
 * [pool_tokens] holds token ids; row b is the concatenation of segments e in [ent_off[b], ent_off[b+1]): ent_len[e] tokens from
 * pool[ent_src[e]] written at column ent_dst[e] of the row (the caller lays rows out left-padded: the first segment starts at
 * L - row_len[b]).  Outputs int64 [B][L]: input_ids (pad_id in the padding), attention_mask (0 / 1), position_ids (0 in the
 * padding, then 1, 2, ...).  A segment that leaves the pool or its row's real region is skipped. */
int ragen_traj_assemble(ragen_ctx* ctx, int B, int L, const int32_t* pool, int64_t pool_tokens, const int64_t* ent_src,
                        const int32_t* ent_dst, const int32_t* ent_len, const int32_t* ent_off, const int32_t* row_len,
                        int64_t pad_id, int64_t* input_ids, int64_t* attention_mask, int64_t* position_ids, void* stream);

/* Text render of a grid observation straight into TOKEN ids (SokobanEnv.render, ragen/env/sokoban/env.py:53-57; FrozenLakeEnv
 * .render, frozen_lake/env.py:47-57, followed by the tokenizer): codes [n][G] (or [n][G/2] when packed: two 4-bit codes per
 * byte, low nibble first), env i at codes + i * pitch bytes, are the glyph rows the env kernels write (code 7 = row separator, the last byte of an env included:
 * the "\n" that follows the state in the prompt, ctx_manager.py:259); [host] table int32 [8][4] / table_len int32 [8] give the
 * 0..4 token ids of each code under the caller's tokenizer.  tokens int32 [n][stride] receives env i's ids, n_tokens (may be
 * NULL) [n] their count.  Valid for tokenizers that split a grid per glyph (the caller verifies that against the tokenizer). */
int ragen_glyph_tokens(ragen_ctx* ctx, int64_t n, int G, int packed, const uint8_t* codes, int64_t pitch, const int32_t* table,
                       const int32_t* table_len, int stride, int32_t* tokens, int32_t* n_tokens, void* stream);

/* The same render into UTF-8 TEXT (SokobanEnv.render's string, sokoban/env.py:53-57): [host] table int32 [8][4] / table_len hold
 * the 0..4 UTF-8 bytes of each glyph ("√" is three), text uint8 [n][stride] receives env i's string NUL-padded to the stride
 * (a numpy 'S<stride>' view of it is an array of byte strings), n_bytes (may be NULL) its length.  format: 0 / 1 as `packed`
 * above, 3 = the three-bit rows of ragen_sokoban_step_packed3 / _compact (row-major cells, no separators; W = room width, pitch
 * >= the row's even byte count).  The last separator is rendered as "\n" too; the reference string has none (drop one byte). */
int ragen_glyph_utf8(ragen_ctx* ctx, int64_t n, int G, int format, int W, const uint8_t* codes, int64_t pitch, const int32_t* table,
                     const int32_t* table_len, int stride, uint8_t* text, int32_t* n_bytes, void* stream);

/* _normalize_score_tensor (ctx_manager.py:175-226) on the last score column: out = score_last + penalty, then,
 * when n_groups < B, per group (dense ids 0..n_groups-1): method 0 identity, 1 mean, 2 mean_std, 3 asym_clip. */
int ragen_group_norm(ragen_ctx* ctx, int B, const float* score_last, const float* penalty, const int32_t* group,
                     int n_groups, int method, float* out, void* stream);

/* The same in two halves, for rows sharded over ranks when a group spans ranks (grouping "inductive" = per tag, "batch" = all
 * rows; ctx_manager.py:186-189): ragen_group_norm_stats writes x = score_last + penalty and this rank's per-group {count, sum,
 * sum of squares} (double [n_groups][3], zeroed by the call; B = 0 is allowed), the caller all-reduces gstats (and the row count:
 * the reference normalises only when there are fewer groups than rows, :219), ragen_group_norm_apply normalises x in place. */
int ragen_group_norm_stats(ragen_ctx* ctx, int B, const float* score_last, const float* penalty, const int32_t* group,
                           int n_groups, float* x, double* gstats, void* stream);
int ragen_group_norm_apply(ragen_ctx* ctx, int B, float* x, const int32_t* group, int n_groups, const double* gstats,
                           int method, void* stream);

/* ragen_group_norm without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments, same results as ragen_group_norm; the per-group statistics live in a file-scope
 * `static __device__` array of RAGEN_NOALLOC_MAX_GROUPS * 3 doubles (96 KiB, held by every GPU that loads the library
 * until the process exits) instead of the ctx-owned scratch.  n_groups > RAGEN_NOALLOC_MAX_GROUPS -> RAGEN_EINVAL.
 * Calls no cudaMalloc / cudaFree / allocating routine; the first call on a device resolves the array's device address
 * (cudaGetSymbolAddress), so make that first call outside a graph capture.  Every call on a device SHARES the array:
 * a call clears what it uses on `stream` first, and calls that overlap on the GPU (unordered streams) race on it. */
#define RAGEN_NOALLOC_MAX_GROUPS 4096
int ragen_group_norm_noalloc(ragen_ctx* ctx, int B, const float* score_last, const float* penalty,
                             const int32_t* group, int n_groups, int method, float* out, void* stream);

/* ---- advantage pass ------------------------------------------------------------------------- */
/* compute_gae_advantage_return (verl/verl/trainer/ppo/core_algos.py:67-109) WITHOUT the final whitening:
 * adv/ret fp32 [B][T] (bit-identical to the reference's sequential fp32 loop) and, when moments != NULL, the
 * masked moments {sum m, sum adv*m, sum adv^2*m} (double[3], zeroed by the call) that masked_whiten needs -
 * the only quantity that has to be all-reduced when rows are sharded over GPUs. */
int ragen_gae(ragen_ctx* ctx, int B, int T, const float* rewards, const float* values, const uint8_t* mask,
              double gamma, double lam, float* adv, float* ret, double* moments, void* stream);

/* The same quantities from a PARALLEL scan of the recurrence (a CTA per row, 2304 words per step staged by bulk copies - 2048
 * tokens by 4-byte cp.async when a base pointer is off a 16-byte boundary or B * T % 16 != 0; per-thread runs of 8 tokens
 * combined by warp-shuffle suffix sums): adv / ret equal the
 * reference's sequential fp32 loop up to rounding (a few 1e-7 relative; north_star's bar is 1e-5) instead of bit for bit, at
 * HBM speed for any number of rows.  For callers that consume the whitened advantages; ragen_gae keeps the reference's bits. */
int ragen_gae_scan(ragen_ctx* ctx, int B, int T, const float* rewards, const float* values, const uint8_t* mask,
                   double gamma, double lam, float* adv, float* ret, double* moments, void* stream);

/* compute_bi_level_gae_advantage_return (ragen/trainer/core_algos.py:4-92) without the final whitening. */
int ragen_bilevel_gae(ragen_ctx* ctx, int B, int T, const float* rewards, const float* values,
                      const uint8_t* loss_mask, double gamma, double lam, double high_level_gamma, float* adv,
                      float* ret, double* moments, void* stream);

/* The same estimator as block scans (every piece of :44-88 is an associative composition of maps on a state that travels right to
 * left: next loss-mask value, next reward-position value, turn-level and token-level lastgaelam): equal to ragen_bilevel_gae up
 * to fp32 rounding (north_star's bar is 1e-5) instead of bit for bit, at HBM speed for any number of rows.  Tiles travel by bulk
 * copies, which need 16-byte-aligned base pointers and B * T % 16 == 0; other tensors are handed to ragen_bilevel_gae. */
int ragen_bilevel_gae_scan(ragen_ctx* ctx, int B, int T, const float* rewards, const float* values,
                           const uint8_t* loss_mask, double gamma, double lam, double high_level_gamma, float* adv,
                           float* ret, double* moments, void* stream);

/* masked moments of any fp32 tensor, and masked_whiten (verl/verl/utils/torch_functional.py:122-151) applied
 * in place from (possibly all-reduced) moments: x <- (x - mean) * rsqrt(var + 1e-8). */
int ragen_masked_moments(ragen_ctx* ctx, int64_t n, const float* x, const uint8_t* mask, double* moments,
                         void* stream);
int ragen_whiten(ragen_ctx* ctx, int64_t n, float* x, const double* moments, void* stream);

/* compute_grpo_outcome_advantage (verl core_algos.py:113-165): scores[b] = sum_t rewards[b][t]; per group
 * (dense ids; a group of one has mean 0 / std 1) (score - mean)/(std + epsilon) or score - mean, times mask. */
int ragen_grpo_outcome(ragen_ctx* ctx, int B, int T, const float* rewards, const uint8_t* mask,
                       const int32_t* group, int n_groups, double epsilon, int norm_by_std, float* scores,
                       float* adv, void* stream);

/* ---- the rest of compute_advantage's dispatch table and its neighbours ---------------- */
/* Outcome estimators over per-group statistics of the row scores (scores[b] = sum_t rewards[b][t]), broadcast
 * over the row and multiplied by the mask.  mode 0: score - mean; 1: GRPO (as ragen_grpo_outcome); 2: RLOO
 * (compute_rloo_outcome_advantage, verl core_algos.py:205-242: score*n/(n-1) - mean*n/(n-1) for groups of n > 1);
 * 3: REINFORCE++-baseline before its whitening (compute_reinforce_plus_plus_baseline_outcome_advantage, :168-202:
 * score - mean; the caller whitens adv in place with ragen_whiten from `moments`).  Groups of one use mean 0.
 * moments (may be NULL) = masked moments of adv (double[3], zeroed by the call). */
int ragen_outcome_advantage(ragen_ctx* ctx, int B, int T, const float* rewards, const uint8_t* mask,
                            const int32_t* group, int n_groups, int mode, double epsilon, float* scores, float* adv,
                            double* moments, void* stream);

/* compute_reinforce_plus_plus_outcome_advantage (verl core_algos.py:245-278), first half: discounted returns with the
 * running value reset after masked-out tokens, fp32 in the reference's order, plus their masked moments.  The
 * advantages are ragen_whiten_masked(ret, mask, moments): masked_whiten(returns) * mask (:275-276). */
int ragen_discounted_returns(ragen_ctx* ctx, int B, int T, const float* rewards, const uint8_t* mask, double gamma,
                             float* ret, double* moments, void* stream);
int ragen_whiten_masked(ragen_ctx* ctx, int64_t n, const float* x, const uint8_t* mask, const double* moments,
                        float* out, void* stream);

/* compute_remax_outcome_advantage (verl core_algos.py:281-305): ret = reverse cumulative sum of rewards*mask,
 * adv = ret - baselines[b]*mask. */
int ragen_remax(ragen_ctx* ctx, int B, int T, const float* rewards, const uint8_t* mask, const float* baselines,
                float* adv, float* ret, void* stream);

/* The two estimators above as block scans (the running value of REINFORCE++ is an affine map per token, ReMax a plain suffix sum,
 * still summed in double): equal to ragen_discounted_returns / ragen_remax up to fp32 rounding instead of bit for bit, at HBM
 * speed for any number of rows.  Tiles travel by bulk copies (16-byte-aligned base pointers, B * T % 16 == 0); other tensors are
 * handed to the sequential entry points. */
int ragen_discounted_returns_scan(ragen_ctx* ctx, int B, int T, const float* rewards, const uint8_t* mask, double gamma,
                                  float* ret, double* moments, void* stream);
int ragen_remax_scan(ragen_ctx* ctx, int B, int T, const float* rewards, const uint8_t* mask, const float* baselines,
                     float* adv, float* ret, void* stream);

/* apply_kl_penalty (verl/verl/trainer/ppo/ray_trainer.py:145-175) with kl_penalty (verl core_algos.py:469-501):
 * kld = f(logp, ref_logp) * mask with mode 0 "kl", 1 "abs", 2 "mse", 3 "low_var_kl"; rewards_out = scores - beta*kld;
 * row_kl[b] (may be NULL) = sum_t kld / (sum_t mask + 1e-8), whose mean over rows is the controller's current_kl. */
int ragen_kl_rewards(ragen_ctx* ctx, int B, int T, const float* scores, const float* logp, const float* ref_logp,
                     const uint8_t* mask, int mode, double beta, float* rewards_out, float* row_kl, void* stream);

/* _filter_rollout's statistics (ragen/trainer/agent_trainer.py:466-469): row_sum[b] = sum_t scores[b][t] for
 * b in [0, n_groups*group_size) and, per group of group_size consecutive rows, unbiased std / max / mean. */
int ragen_group_filter_stats(ragen_ctx* ctx, int n_groups, int group_size, int T, const float* scores, float* row_sum,
                             float* gstd, float* gmax, float* gmean, void* stream);

/* Answer scoring of the text environments: n (prediction, label) pairs as two UTF-8 blobs with
 * int64 offsets [n+1].  mode 0: MetaMathQAEnv._check_answer (ragen/env/metamathqa/env.py:55-66); 1, 2, 3:
 * compute_score_exact_match / compute_score_numeric / compute_score_multiple_choice (ragen/env/static/utils.py:69-132).
 * result[i]: bit0 is_correct, bit1 is_valid, bit2 normalised texts equal, bit7 "unsupported": a non-ASCII byte (Python
 * folds case by Unicode tables) or a number of more than 15 significant digits was seen - the other bits are then the
 * ASCII / decimal-string answer. */
int ragen_text_score(ragen_ctx* ctx, int64_t n, int mode, const uint8_t* pred, const int64_t* pred_offsets,
                     const uint8_t* label, const int64_t* label_offsets, uint8_t* result, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RAGEN_B200_H */
