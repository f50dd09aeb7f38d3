/* libb200gate -- profiling aids.  NOT part of the ABI a V-Gate maintainer binds (that is b200gate.h):
 * these entry points install device buffers that the kernels fill with clock stamps while they run; they
 * are used by tools/timeline.py, tools/umma_trace.py and tools/attn_trace.py.  None has a counterpart in
 * the reference.  All are no-ops for the results the kernels compute. */
#ifndef B200GATE_DEBUG_H
#define B200GATE_DEBUG_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Profiling aid: when dev_buf is non-NULL the tcgen05 GEMM writes 8 x uint64 SM-clock stamps per CTA
 * (0 entry, 1 set-up done, 2 first stage complete, 3 last MMA issued, 4 accumulator complete, 5 accumulator
 * staged in shared memory, 6 epilogue done, 7 dependency wait of the activation-TMA warp returned). */
void b200_debug_umma_trace(void* dev_buf);
/* Same, restricted to launches whose grid is (grid_x, *, grid_z) -- to pick one projection inside a step. */
void b200_debug_umma_trace_grid(void* dev_buf, int32_t grid_x, int32_t grid_z);
/* Same for the attention kernel (0 entry, 1 dependency wait returned and tiles checked, 2 prologue done, 3 first
 * K/V tile landed, 4 scores, 5 softmax, 6 chunk loop end, 7 merge + store done). */
void b200_debug_attn_trace(void* dev_buf);
/* In-situ timeline: while `buf` (device, capacity x 4 uint64) and `counter` (device uint32, zeroed by the
 * caller) are installed, block 0 of every kernel appends {kernel id | grid, entry, ready, exit} in
 * %globaltimer nanoseconds.  NULL buf switches it off.  Works under CUDA-graph replay. */
void b200_debug_timeline(void* buf, void* counter, uint32_t capacity);

#ifdef __cplusplus
}
#endif
#endif /* B200GATE_DEBUG_H */
