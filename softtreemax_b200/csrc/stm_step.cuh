// stm_step.cuh — K2, the batched Atari 2600 agent-step kernel, as a header: stm_kernels.cu instantiates it with the
// interpreter (stm::InterpCore), every translated cartridge (csrc/gen/xl_<name>.cu, one translation unit each so that
// they compile in parallel) instantiates it with its own core.  Reference call site replaced: policies/cule_bfs.py
// :152-158 (step + get_data x frameskip).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/stm_b200.h"
#include "../../include/stm_palette.h"
#include "emu_core.cuh"

// arguments of one K2 launch (a named type: it crosses translation units)
struct StmStepArgs {
    uint8_t *states; const uint8_t *actions; float *rewards; uint8_t *done; int32_t *lives;
    const uint32_t *rom; const uint32_t *lut; const uint2 *pre; uint32_t *scratch; uint8_t *stepped;
    const uint32_t *grey;                 // the ctx's colour-index -> grey table, 128 bytes as 32 words (stm_set_palette)
    int n, frameskip, episodic_life, render; uint32_t rom_words; float weight;
    int lock_scope;                       // who advances scanline by scanline together: 0 the warps of a scheduler (default), 1 the block, 2 each warp alone (A/B)
    int deal;                             // 1: warps dealt round-robin to the blocks
    int warps_per_cta;                    // 1..kStepThreads/32: how many of the block's warps carry envs (the launch spreads its warps over the SMs)
    stm::GameParams gp;
    // frame-by-frame driving (stm_step_frame, the torchcule-style surface): `frozen` decides which envs sit the frame out
    // (the done flags as they were when the agent step began) instead of done[], and `actions` holds ALE action codes
    const uint8_t *frozen; int action_codes;
};

namespace stm {
// what stm_kernels.cu knows about a translated cartridge (filled in by csrc/gen/xl_<name>.cu)
struct XlEntry {
    const char *name;
    uint32_t abi;                                                      // kXlAbi of the build that compiled it (plug-ins: must equal the library's)
    uint32_t rom_hash, rom_size, n_cases;
    const uint16_t *offsets;                                           // image offsets of the translated instructions (stamp_cases)
    void (*launch)(int blocks, size_t smem, cudaStream_t st, const StmStepArgs &a);
    cudaError_t (*set_attrs)(size_t smem, int carveout_pct);
};
}  // namespace stm

#ifndef STM_STEP_THREADS
#define STM_STEP_THREADS 512  /* one block per SM: 16 warps, 4 per scheduler, the register file at <= 128 registers a thread */
#endif
namespace stm {
// what a translated core and the library must agree on: the launch-argument record, the shared-memory pitch, the state blob
constexpr uint32_t kXlAbi = 0x02000000u + (uint32_t)sizeof(StmStepArgs) * 4096u + (uint32_t)STM_STEP_THREADS;
}  // namespace stm

namespace {

using namespace stm;
typedef StmStepArgs StepArgs;

constexpr int kTileEnvs = 32;                                   // envs per native-scratch tile (= warp)
// Native-frame scratch of one tile, two frames.  Pixel area: [frame][slot 0..209][quad 10][lane 32] uint4 — a SLOT holds one
// distinct row of 160 grey pixels; slot 0 is never written and stays zero.  Row table: [frame][row 0..209][lane 32] bytes — the
// slot that native row `row` of the lane's frame shows.  K2 stores a row's pixels once, when its content differs from the row
// above (whole-line reuse: a repeated row is ONE byte in the table, not a 160-byte copy), K3 reads rows through the table.
constexpr size_t kTilePixWords = 2ull * STM_VIS_H * 40 * kTileEnvs;
constexpr size_t kTileSlotWords = 2ull * STM_VIS_H * kTileEnvs / 4;
constexpr size_t kTileWords = kTilePixWords + kTileSlotWords;    // u32 per tile
constexpr int kStepThreads = STM_STEP_THREADS;
#ifndef STM_STEP_MIN_CTAS
#define STM_STEP_MIN_CTAS 1
#endif

// ---------------------------------------------------------------------------------------------
// K2 — batched agent step.  One env per thread; one block per SM, of which the first `warps_per_cta` warps carry envs (the
// host spreads a launch's warps evenly over the SMs: 65 536 envs = 14 warps on each of 147 SMs, 16 384 envs = 4 warps on 128).
// The warps of a block that share a scheduler advance scanline by scanline together (emu_core.cuh stm_group_all).
// Shared memory: the block's state blobs, word-interleaved (word k of thread t at [k*512 + t], so lane t always hits
// bank t%32: conflict-free for every access pattern, RAM bytes included), the cartridge image (4 or 8 KB), the decode
// table and the grey table.
// ---------------------------------------------------------------------------------------------

struct DevMem {
    uint8_t *blob;             // &sblob[tid] as bytes
    const uint32_t *rom;       // shared
    const uint32_t *lut;       // shared
    const uint8_t *sgrey;      // shared: colour index -> grey (data of the ctx, not of the build)
    const uint2 *pre;          // global, read-only path (L1-resident: 32 KB for a 4 KB cartridge)
    uint32_t *scr;             // native scratch of this env's tile, + lane
    uint8_t *rowslot;          // row table of this env's tile, + lane
    uint32_t fsel;             // which of the two native frames is being drawn
    uint32_t cur_row, cur_slot, nslots;     // the row opened last, its slot, slots used in this frame

    __device__ __forceinline__ uint32_t boff(uint32_t o) const { return (o >> 2) * (kStepThreads * 4) + (o & 3); }
    __device__ __forceinline__ uint32_t ld8(uint32_t o) const { return blob[boff(o)]; }
    __device__ __forceinline__ void st8(uint32_t o, uint32_t v) { blob[boff(o)] = (uint8_t)v; }
    __device__ __forceinline__ uint32_t ld16(uint32_t o) const { return *(const uint16_t *)(blob + boff(o)); }
    __device__ __forceinline__ void st16(uint32_t o, uint32_t v) { *(uint16_t *)(blob + boff(o)) = (uint16_t)v; }
    __device__ __forceinline__ uint32_t ld32(uint32_t o) const { return *(const uint32_t *)(blob + boff(o)); }
    __device__ __forceinline__ void st32(uint32_t o, uint32_t v) { *(uint32_t *)(blob + boff(o)) = v; }
    __device__ __forceinline__ uint32_t rom8(uint32_t i) const { return (rom[i >> 2] >> ((i & 3) * 8)) & 0xFFu; }
    __device__ __forceinline__ void fetch(uint32_t off, uint32_t &w, uint32_t &ops) const
    { uint2 v = __ldg(pre + off); w = v.x; ops = v.y; }
    __device__ __forceinline__ uint32_t opinfo(uint32_t op) const { return lut[op]; }
    __device__ __forceinline__ uint32_t grey(uint32_t i) const { return sgrey[i & 127]; }
    // scr points at this lane's uint4 of (frame 0, slot 0, quad 0); quads are kTileEnvs uint4 apart.  Every pixel store goes
    // to the slot of the row opened last (open_row precedes the renderer), whatever `row` says.
    __device__ __forceinline__ uint32_t *quad(uint32_t, uint32_t q) const
    { return scr + (size_t)((fsel * STM_VIS_H + cur_slot) * 10 + q) * (kTileEnvs * 4); }
    __device__ __forceinline__ void begin_frame(uint32_t f) { fsel = f; cur_row = 0xFFFFFFFFu; cur_slot = 0; nslots = 0; }
    __device__ __forceinline__ void set_row(uint32_t row, uint32_t slot) { rowslot[(size_t)(fsel * STM_VIS_H + row) * kTileEnvs] = (uint8_t)slot; }
    __device__ __forceinline__ void open_row(uint32_t row)
    { if (row != cur_row) { cur_row = row; cur_slot = ++nslots; set_row(row, cur_slot); } }
    __device__ __forceinline__ void link_row(uint32_t dst, uint32_t) { set_row(dst, cur_slot); }      // src == cur_row: the clean row is the one opened last
    __device__ __forceinline__ void zero_row(uint32_t row) { set_row(row, 0); }
    __device__ __forceinline__ void px16(uint32_t row, uint32_t q, uint32_t a, uint32_t b, uint32_t c, uint32_t d)
    { __stcg(reinterpret_cast<uint4 *>(quad(row, q)), make_uint4(a, b, c, d)); }
    __device__ __forceinline__ void px4(uint32_t row, uint32_t g, uint32_t w) { quad(row, g >> 2)[g & 3] = w; }
    __device__ __forceinline__ void px1(uint32_t row, uint32_t x, uint32_t v)
    { reinterpret_cast<uint8_t *>(quad(row, x >> 4))[x & 15] = (uint8_t)v; }
};



// Core = stm::InterpCore (any cartridge) or a translated cartridge stm::Xl_<name> (csrc/gen/)
template <class Core>
__global__ void __launch_bounds__(kStepThreads, STM_STEP_MIN_CTAS)
stm_step_kernel(const StepArgs a)
{
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t *sblob = smem;                              // 64 words x kStepThreads (the pitch; a smaller block uses the first columns)
    uint32_t *srom = sblob + 64 * kStepThreads;
    uint32_t *slut = srom + a.rom_words;
    const int tid = threadIdx.x;
    // the block is launched with 32 * warps_per_cta threads (kStepThreads is the shared-memory pitch and the upper bound): what
    // a smaller block leaves of the SM's registers is room for K3b's copy blocks beside it
    for (uint32_t i = tid; i < a.rom_words; i += blockDim.x) srom[i] = a.rom[i];
    for (uint32_t i = tid; i < 256; i += blockDim.x) slut[i] = a.lut[i];
    uint32_t *sgrey = slut + 256;
    if (tid < 32) sgrey[tid] = a.grey[tid];
    const int warp = tid >> 5;
    // which 32 envs this warp carries: consecutive warps of the launch go to one block (deal 0), or are dealt round-robin to
    // the blocks (deal 1: every SM gets a sample of the whole launch, whatever order the caller's envs are in)
    const int gwarp = a.deal ? (int)(blockIdx.x + warp * gridDim.x) : (int)(blockIdx.x * a.warps_per_cta + warp);
    const int env = gwarp * 32 + (tid & 31);
    const bool valid = warp < a.warps_per_cta && env < a.n;
    if (valid) {
        const uint4 *g = reinterpret_cast<const uint4 *>(a.states + (size_t)env * STM_STATE_BYTES);
        #pragma unroll
        for (int k = 0; k < 16; k++) {
            uint4 v = g[k];
            sblob[(4 * k + 0) * kStepThreads + tid] = v.x; sblob[(4 * k + 1) * kStepThreads + tid] = v.y;
            sblob[(4 * k + 2) * kStepThreads + tid] = v.z; sblob[(4 * k + 3) * kStepThreads + tid] = v.w;
        }
    }
    __syncthreads();
    // threads of this warp's scheduler group (the warps w of the block with w % 4 == warp % 4 that carry envs)
    const uint32_t gw = a.lock_scope == 1 ? (uint32_t)a.warps_per_cta : a.lock_scope == 2 ? 1u : (uint32_t)((a.warps_per_cta - (warp & 3) + 3) >> 2);
    const uint32_t lanes = 32u * gw | (a.lock_scope == 1 ? 1u : 1u + (uint32_t)(warp & 3)) << 16;
    uint32_t done = valid ? a.done[env] : 1u;
    const uint32_t frozen = valid ? (a.frozen ? (uint32_t)a.frozen[env] : done) : 1u;
    // every warp of a group has to reach every barrier of the lock-step, so dead lanes (past n, or frozen: nothing changes
    // for them) stay and idle through run_frame
    const bool live = valid && !frozen;
    if (valid && a.stepped) a.stepped[env] = frozen ? 0 : 1;

    DevMem m;
    m.blob = reinterpret_cast<uint8_t *>(sblob + tid);
    m.rom = srom; m.lut = slut; m.pre = a.pre; m.sgrey = reinterpret_cast<const uint8_t *>(sgrey);
    m.scr = a.scratch ? a.scratch + (size_t)(env / kTileEnvs) * kTileWords + (env % kTileEnvs) * 4 : nullptr;
    m.rowslot = a.scratch ? reinterpret_cast<uint8_t *>(a.scratch + (size_t)(env / kTileEnvs) * kTileWords + kTilePixWords) + env % kTileEnvs : nullptr;
    m.begin_frame(0);
    Emu<DevMem> e(m, a.gp);
    int lives = 0; float rew = 0.f; uint32_t ale = 0;
    if (live) {
        e.load();
        lives = a.lives[env];
        rew = a.rewards[env];
        uint32_t act = a.actions ? a.actions[env] : (uint32_t)(env % (int)a.gp.n_actions);
        ale = a.action_codes ? act : (uint32_t)a.gp.actions[act];
    }
    for (int f = 0; f < a.frameskip; f++) {
        bool want = a.render && f >= a.frameskip - 2;
        // the newest frame always goes to slot 1, the one before it to slot 0
        m.begin_frame((f == a.frameskip - 1) ? 1u : 0u);
        e.template run_frame<Core>(ale, want, lanes, live);
        if (live) {
            int r = e.get_data(a.episodic_life != 0, done, lives);
            rew = __fadd_rn(rew, __fmul_rn(a.weight, (float)r));
        }
    }
    if (!live) return;
    e.store();
    uint4 *g = reinterpret_cast<uint4 *>(a.states + (size_t)env * STM_STATE_BYTES);
    #pragma unroll
    for (int k = 0; k < 16; k++) {
        uint4 v;
        v.x = sblob[(4 * k + 0) * kStepThreads + tid]; v.y = sblob[(4 * k + 1) * kStepThreads + tid];
        v.z = sblob[(4 * k + 2) * kStepThreads + tid]; v.w = sblob[(4 * k + 3) * kStepThreads + tid];
        g[k] = v;
    }
    a.rewards[env] = rew; a.done[env] = (uint8_t)done; a.lives[env] = lives;
}

template <class Core>
cudaError_t set_step_attrs(size_t smem, int pct)
{
    cudaError_t e = cudaFuncSetAttribute(stm_step_kernel<Core>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(stm_step_kernel<Core>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
}

}  // namespace
