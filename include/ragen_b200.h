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

/* SokobanEnv.render text branch only (env.py:53-57). */
int ragen_sokoban_render(ragen_ctx* ctx, int64_t n, int H, int W, const uint8_t* state,
                         const int16_t* meta, uint8_t* obs, void* stream);

/* Post-reset bookkeeping of SokobanEnv.reset (env.py:37-38) for rooms already in device memory:
 * player = argwhere(state == 5)[0], num_env_steps = 0, boxes_on_target = 0 -> meta.
 * status[i] (may be NULL) = 0 ok, 1 no player cell, 2 border not wall, 3 room_fixed is not the
 * function of room_state described above (such rooms cannot come from the reference generator). */
int ragen_sokoban_adopt(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, const uint8_t* state,
                        const uint8_t* fixed, int16_t* meta, int32_t* status, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RAGEN_B200_H */
