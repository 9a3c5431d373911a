/*
 * stm_b200_debug.h — measurement-only entry points of libstm_b200.so.  None of them is on a product path and none has a
 * counterpart in the reference: bench.py uses the timing and launch counters for its roofline lines, tools/shift_probe.py the
 * operand-shift probe.  The product ABI is include/stm_b200.h.
 */
#ifndef STM_B200_DEBUG_H
#define STM_B200_DEBUG_H

#include "stm_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Per-kernel device timing for bench.py's roofline: when enabled, a CUDA event pair is recorded on
 * the launch stream around every K1/K2/K3 launch.  stm_timing_read waits for them and returns the
 * summed milliseconds and launch counts by kind {0: clone, 1: step, 2: frames, 3: stack shift (stm_bfs runs it on its own
 * stream, under the step kernel)} — arrays of 4 — then clears. */
int stm_timing_enable(stm_ctx *ctx, int on);
int stm_timing_read(stm_ctx *ctx, float *ms_by_kind, int *launches_by_kind);

/* number of kernels this ctx has launched since creation (bench.py's gpu_launches) */
long long stm_launch_count(const stm_ctx *ctx);
/* 1-based index of the ahead-of-time block-translated step kernel that serves this ctx's cartridge image;
 * 0 = interpreter only (an image the library has no translation for, or STM_XLATE=0). */
int stm_translated_core(const stm_ctx *ctx);
/* (probe, no reference counterpart, not on any product path) runs one 128 x 64 x 64 bf16 tcgen05 MMA whose A operand
 * starts `shift` rows into a 128B-swizzled [rows][64] shared-memory tile, with `base_off` in the descriptor's
 * base-offset field; out = A[shift .. shift + 128) x B^T as f32.  tools/shift_probe.py uses it to show that a
 * 128-byte-aligned operand start is read correctly — the fact the window GEMMs of K4 rest on. */
int stm_debug_shift_probe(stm_ctx *ctx, const void *A, const void *B, float *out, int rows, int shift, int base_off, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* STM_B200_DEBUG_H */
