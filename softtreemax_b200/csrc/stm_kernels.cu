// stm_kernels.cu — sm_100a kernels K1 (clone), K2 (batched Atari 2600 agent step), K3 (frames)
// and the C ABI of include/stm_b200.h around them.
//
// Reference call sites replaced (policies/cule_bfs.py): expand :143 / new_expand :81-96 (K1),
// step + get_data :152-158 (K2), generate_frames + max + replicate_state + cat :162-172 (K3),
// the per-depth driver loop :129-173 and epilogue :204-209 (stm_bfs).
//
// Data layout in HBM
//   env pool      : n x 256 B state blobs (include/stm_state.h), rewards f32, done u8, lives i32
//   obs stacks    : n x 4 x 84 x 84 u8
//   native scratch: per tile of 32 envs, [frame 0..1][row 0..209][quad 0..9][lane 0..31] uint4,
//                   one uint4 = 16 horizontally adjacent pixels (7-bit colour indices) of one env.
//                   Lane-minor, so that a warp whose lanes draw the same quad of the same row (the
//                   usual case: K2 wraps scanlines in warp lock-step) issues ONE coalesced 512-byte
//                   store, and K3 reads it back coalesced.
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <cmath>
#include <new>
#include "../../include/stm_b200.h"
#include "../../include/stm_palette.h"
#include "emu_core.cuh"
#include "gen/xl_index.cuh"

namespace {

using namespace stm;

static_assert(sizeof(stm_env_state) == STM_STATE_BYTES, "state blob must be 256 bytes");
static_assert(offsetof(stm_env_state, pc) == O_PC && offsetof(stm_env_state, timer) == O_TIMER &&
              offsetof(stm_env_state, cx) == O_CX && offsetof(stm_env_state, pos) == O_POS &&
              offsetof(stm_env_state, vsync) == O_VSYNC && offsetof(stm_env_state, nusiz0) == O_NUSIZ0 &&
              offsetof(stm_env_state, pf2) == O_PF2 && offsetof(stm_env_state, enam0) == O_ENAM0 &&
              offsetof(stm_env_state, hmp0) == O_HMP0 && offsetof(stm_env_state, vdelbl) == O_VDELBL &&
              offsetof(stm_env_state, resmp1) == O_RESMP1 && offsetof(stm_env_state, score_prev) == O_SCORE_PREV &&
              offsetof(stm_env_state, swcha_out) == O_SWCHA_OUT && offsetof(stm_env_state, paddle) == O_PADDLE,
              "emu_core.cuh offsets out of sync with stm_state.h");

constexpr int kTileEnvs = 32;                                   // envs per native-scratch tile (= warp)
constexpr size_t kTileWords = 2ull * STM_VIS_H * 40 * kTileEnvs; // u32 per tile (2 frames)
constexpr int kStepThreads = 128;
#ifndef STM_STEP_MIN_CTAS
#define STM_STEP_MIN_CTAS 4   /* measured: 4 CTAs/SM without spills beats 5 with */
#endif

thread_local char g_create_error[256] = "";

} // namespace

struct stm_ctx {
    int device = 0;
    stm_game_desc desc{};
    GameParams gp{};
    uint32_t rom_words = 0;
    uint32_t *d_rom = nullptr;       // cartridge image, as words
    uint32_t *d_lut = nullptr;       // 256-entry decode table
    uint2 *d_pre = nullptr;          // predecoded cartridge: one record per image byte (emu_core.cuh build_predecode)
    uint32_t *d_scratch = nullptr;   // native frames
    uint8_t  *d_stepped = nullptr;   // per scratch env: 1 if the env was emulated by the last stm_step
    int scratch_envs = 0;
    int xl = 0;                      // 0: interpreter; k > 0: the k-th translated cartridge of STM_XL_LIST
    int last_render_frames = 0;      // how many native frames the last stm_step(render) produced (1 or 2)
    // intermediate-level pools for stm_bfs (ping-pong)
    uint8_t *pool_state[2] = {nullptr, nullptr};
    float   *pool_reward[2] = {nullptr, nullptr};
    uint8_t *pool_done[2] = {nullptr, nullptr};
    int32_t *pool_lives[2] = {nullptr, nullptr};
    uint8_t *pool_stack[2] = {nullptr, nullptr};
    size_t pool_envs[2] = {0, 0};
    size_t pool_stack_envs[2] = {0, 0};
    // device staging for stm_bfs_host
    uint8_t *h_in_state = nullptr, *h_in_stack = nullptr, *h_out_stack = nullptr;
    float *h_out_reward = nullptr; uint8_t *h_out_done = nullptr; int32_t *h_out_lives = nullptr;
    size_t h_roots = 0, h_leaves = 0;
    // stm_bfs_host uploads the root stacks (28 KB per root) on its own stream while the first level steps
    cudaStream_t up_stream = nullptr;
    cudaEvent_t up_done = nullptr, up_go = nullptr;
    uint32_t reset_pc = 0;
    void *cnn = nullptr;             // K4 state (stm_cnn.cu): tiled weights + activation scratch
    long long launches = 0;
    // optional per-kernel timing (stm_timing_*): CUDA events recorded around every launch
    bool timing = false;
    static constexpr int kMaxTimed = 4096;
    cudaEvent_t *ev = nullptr;          // 2 * kMaxTimed events
    int ev_kind[kMaxTimed];
    int n_timed = 0;
    char err[256] = "";
};

namespace {

int fail(stm_ctx *ctx, int code, const char *fmt, ...)
{
    char *dst = ctx ? ctx->err : g_create_error;
    va_list ap; va_start(ap, fmt); vsnprintf(dst, 256, fmt, ap); va_end(ap);
    return code;
}
#define CUDA_TRY(ctx, expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) \
    return fail(ctx, e_ == cudaErrorMemoryAllocation ? STM_ERR_NOMEM : STM_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); } while (0)

// ---------------------------------------------------------------------------------------------
// K1 — clone.  Persistent CTAs; a tile of kCloneParents parent rows (8 KB) is brought into shared
// memory with one 1-D bulk async copy (TMA, completes on an mbarrier), double buffered, and fanned
// out with 128-bit stores.  HBM traffic = n_src*S read + n_src*fanout*S written (the algorithmic
// minimum); the small per-env arrays ride along.
// ---------------------------------------------------------------------------------------------
constexpr int kCloneParents = 32;
constexpr int kCloneThreads = 256;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{ asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{ asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}"
        :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__global__ void __launch_bounds__(kCloneThreads)
stm_clone_kernel(const uint4 *__restrict__ src, uint4 *__restrict__ dst,
                 const float *__restrict__ src_r, float *__restrict__ dst_r,
                 const uint8_t *__restrict__ src_d, uint8_t *__restrict__ dst_d,
                 const int32_t *__restrict__ src_l, int32_t *__restrict__ dst_l,
                 int n_src, int fanout)
{
    __shared__ __align__(128) uint4 tile[2][kCloneParents * 16];
    __shared__ __align__(8) uint64_t bar[2];
    const int ntiles = (n_src + kCloneParents - 1) / kCloneParents;
    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1); mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](int t, int stage) {
        int p0 = t * kCloneParents;
        int np = min(kCloneParents, n_src - p0);
        uint32_t bytes = (uint32_t)np * STM_STATE_BYTES;
        mbar_expect_tx(&bar[stage], bytes);
        bulk_g2s(tile[stage], src + (size_t)p0 * 16, bytes, &bar[stage]);
    };
    int t = blockIdx.x, it = 0;
    if (t < ntiles && threadIdx.x == 0) issue(t, 0);
    for (; t < ntiles; t += gridDim.x, it++) {
        int stage = it & 1;
        int tn = t + gridDim.x;
        if (tn < ntiles && threadIdx.x == 0) issue(tn, stage ^ 1);
        mbar_wait(&bar[stage], (it >> 1) & 1);
        int p0 = t * kCloneParents;
        int np = min(kCloneParents, n_src - p0);
        int nchunks = np * fanout * 16;
        uint4 *out = dst + (size_t)p0 * fanout * 16;
        for (int i = threadIdx.x; i < nchunks; i += kCloneThreads) {
            int child = i >> 4, piece = i & 15;
            int parent = child / fanout;
            out[i] = tile[stage][parent * 16 + piece];
        }
        if (src_r) {
            for (int c = threadIdx.x; c < np * fanout; c += kCloneThreads) {
                int parent = p0 + c / fanout;
                size_t o = (size_t)p0 * fanout + c;
                dst_r[o] = src_r[parent]; dst_d[o] = src_d[parent]; dst_l[o] = src_l[parent];
            }
        }
        __syncthreads();      // everyone is done reading tile[stage] before it is refilled
    }
}

// ---------------------------------------------------------------------------------------------
// K2 — batched agent step.  One env per thread, 128 envs per CTA.
// Shared memory: the CTA's 128 state blobs, word-interleaved (word k of env t at [k*128 + t], so
// lane t always hits bank t%32: conflict-free for every access pattern, RAM bytes included), the
// cartridge image (4 or 8 KB) and the decode table.
// ---------------------------------------------------------------------------------------------
__constant__ uint8_t c_gray[128] = STM_GRAY_LUT_INIT;

struct DevMem {
    uint8_t *blob;             // &sblob[tid] as bytes
    const uint32_t *rom;       // shared
    const uint32_t *lut;       // shared
    const uint2 *pre;          // global, read-only path (L1-resident: 32 KB for a 4 KB cartridge)
    uint32_t *scr;             // native scratch of this env's tile, + lane
    uint32_t fsel;             // which of the two native frames is being drawn

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
    __device__ __forceinline__ uint32_t grey(uint32_t i) const { return c_gray[i & 127]; }
    // scr points at this lane's uint4 slot of (frame 0, row 0, quad 0); quads are kTileEnvs uint4 apart
    __device__ __forceinline__ uint32_t *quad(uint32_t row, uint32_t q) const
    { return scr + (size_t)((fsel * STM_VIS_H + row) * 10 + q) * (kTileEnvs * 4); }
    __device__ __forceinline__ void px16(uint32_t row, uint32_t q, uint32_t a, uint32_t b, uint32_t c, uint32_t d)
    { __stcg(reinterpret_cast<uint4 *>(quad(row, q)), make_uint4(a, b, c, d)); }
    __device__ __forceinline__ void copy_row(uint32_t dst, uint32_t src)
    {
        const uint4 *s = reinterpret_cast<const uint4 *>(quad(src, 0));
        uint4 *d = reinterpret_cast<uint4 *>(quad(dst, 0));
        #pragma unroll
        for (int h = 0; h < 2; h++) {                    // five loads in flight, then five stores
            uint4 v[5];
            #pragma unroll
            for (int q = 0; q < 5; q++) v[q] = __ldcg(s + (size_t)(5 * h + q) * kTileEnvs);   // L2 only: L1 is for the fetch path
            #pragma unroll
            for (int q = 0; q < 5; q++) __stcg(d + (size_t)(5 * h + q) * kTileEnvs, v[q]);
        }
    }
    __device__ __forceinline__ void px4(uint32_t row, uint32_t g, uint32_t w) { quad(row, g >> 2)[g & 3] = w; }
    __device__ __forceinline__ void px1(uint32_t row, uint32_t x, uint32_t v)
    { reinterpret_cast<uint8_t *>(quad(row, x >> 4))[x & 15] = (uint8_t)v; }
};

struct StepArgs {
    uint8_t *states; const uint8_t *actions; float *rewards; uint8_t *done; int32_t *lives;
    const uint32_t *rom; const uint32_t *lut; const uint2 *pre; uint32_t *scratch; uint8_t *stepped;
    int n, frameskip, episodic_life, render; uint32_t rom_words; float weight;
    GameParams gp;
};

// Core = stm::InterpCore (any cartridge) or a translated cartridge stm::Xl_<name> (csrc/gen/)
template <class Core>
__global__ void __launch_bounds__(kStepThreads, STM_STEP_MIN_CTAS)
stm_step_kernel(const StepArgs a)
{
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t *sblob = smem;                              // 64 words x 128 threads
    uint32_t *srom = sblob + 64 * kStepThreads;
    uint32_t *slut = srom + a.rom_words;
    const int tid = threadIdx.x;
    for (uint32_t i = tid; i < a.rom_words; i += kStepThreads) srom[i] = a.rom[i];
    for (uint32_t i = tid; i < 256; i += kStepThreads) slut[i] = a.lut[i];
    const int env = blockIdx.x * kStepThreads + tid;
    const bool valid = env < a.n;
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
    uint32_t done = valid ? a.done[env] : 1u;
    // lanes that emulate; they advance scanline by scanline in lock-step (Emu::run_frame)
    const uint32_t lanes = __ballot_sync(0xFFFFFFFFu, valid && !done);
    if (!valid) return;
    if (a.stepped) a.stepped[env] = done ? 0 : 1;
    if (done) return;                                    // frozen env: nothing changes

    DevMem m;
    m.blob = reinterpret_cast<uint8_t *>(sblob + tid);
    m.rom = srom; m.lut = slut; m.pre = a.pre;
    m.scr = a.scratch ? a.scratch + (size_t)(env / kTileEnvs) * kTileWords + (env % kTileEnvs) * 4 : nullptr;
    m.fsel = 0;
    Emu<DevMem> e(m, a.gp);
    e.load();
    int lives = a.lives[env];
    float rew = a.rewards[env];
    uint32_t act = a.actions ? a.actions[env] : (uint32_t)(env % (int)a.gp.n_actions);
    uint32_t ale = a.gp.actions[act];
    for (int f = 0; f < a.frameskip; f++) {
        bool want = a.render && f >= a.frameskip - 2;
        // the newest frame always goes to slot 1, the one before it to slot 0
        m.fsel = (f == a.frameskip - 1) ? 1u : 0u;
        e.template run_frame<Core>(ale, want, lanes);
        int r = e.get_data(a.episodic_life != 0, done, lives);
        rew = __fadd_rn(rew, __fmul_rn(a.weight, (float)r));
    }
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

// ---------------------------------------------------------------------------------------------
// K3 — observation kernel.  One CTA per (tile of 32 envs, output row oy).  K2 stores GREY pixels,
// so the 84x84 bilinear resize is pure integer arithmetic here, done separably:
//   vertical   V[x] = wy0 * row_y0[x] + wy1 * row_y1[x]   (wy0 + wy1 = 4, V <= 1020), two pixels per
//              32-bit multiply-add, straight from the coalesced (lane-minor) 16-byte global loads
//              into shared memory as 16-bit values, env-major with an odd word pitch;
//   horizontal out = ((42 - fx) * V[x0] + fx * V[x0 + 1] + 84) / 168, one rounding, as the oracle;
// then the max over the two frames, four output pixels per packed word, 8 threads per env.
// The same CTA moves row oy of the three older stack frames from the parent's stack.
// HBM per leaf: reads 2*168*160 B of native frames (rows 2 mod 5 are never sampled) + 21 168 B of
// parent stack, writes 28 224 B.
// ---------------------------------------------------------------------------------------------
constexpr int kFrameThreads = 256;
constexpr int kVPitch = 83;     // words per (frame, env) row of V in shared memory: 81 used, odd => conflict-free

__global__ void __launch_bounds__(kFrameThreads)
stm_frames_kernel(const uint32_t *__restrict__ scratch, const uint8_t *__restrict__ stepped,
                  const uint8_t *__restrict__ parent, uint8_t *__restrict__ child,
                  int n, int fanout, int nframes)
{
    __shared__ uint32_t sv[2][kTileEnvs * kVPitch];        // [frame][env][x / 2]: V[x] | V[x + 1] << 16
    __shared__ uint8_t sx0[STM_OBS_W], sfx[STM_OBS_W];     // source column and weight of every output column
    const int tile = blockIdx.x, oy = blockIdx.y, tid = threadIdx.x;
    const int env0 = tile * kTileEnvs;
    const int nenv = min(kTileEnvs, n - env0);
    if (tid < STM_OBS_W) { int nx = 80 * tid + 19; sx0[tid] = (uint8_t)(nx / 42); sfx[tid] = (uint8_t)(nx % 42); }
    const int ny = 10 * oy + 3, y0 = ny >> 2, fy = ny & 3;
    const int y1 = min(y0 + 1, STM_VIS_H - 1);
    const uint32_t wy0 = 4 - fy, wy1 = fy;
    const uint4 *tbase = reinterpret_cast<const uint4 *>(scratch + (size_t)tile * kTileWords);
    // vertical pass: item = (frame, quad, lane)
    for (int i = tid; i < 2 * 10 * kTileEnvs; i += kFrameThreads) {
        const int lane = i & 31, q = (i >> 5) % 10, frame = i / (10 * 32);
        uint32_t *d = &sv[frame][lane * kVPitch + 8 * q];
        if (frame >= 2 - nframes) {          // with a single rendered frame only slot 1 holds data
            const uint4 a = __ldcs(tbase + ((size_t)(frame * STM_VIS_H + y0) * 10 + q) * kTileEnvs + lane);
            const uint4 b = __ldcs(tbase + ((size_t)(frame * STM_VIS_H + y1) * 10 + q) * kTileEnvs + lane);
            const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
            #pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint32_t ve = wy0 * (aw[k] & 0x00FF00FFu) + wy1 * (bw[k] & 0x00FF00FFu);            // px 0, 2
                const uint32_t vo = wy0 * ((aw[k] >> 8) & 0x00FF00FFu) + wy1 * ((bw[k] >> 8) & 0x00FF00FFu);   // px 1, 3
                d[2 * k] = __byte_perm(ve, vo, 0x5410);
                d[2 * k + 1] = __byte_perm(ve, vo, 0x7632);
            }
        } else {
            #pragma unroll
            for (int k = 0; k < 8; k++) d[k] = 0;
        }
        if (q == 9) d[8] = d[7] >> 16;       // V[160] = V[159]: the right tap of the last column clamps
    }
    __syncthreads();
    // horizontal pass + max: 8 threads per env, each up to 3 packed words of the new frame
    {
        const int e = tid >> 3, sub = tid & 7;
        if (e < nenv) {
            const bool live = stepped[env0 + e] != 0;
            const uint16_t *v0 = reinterpret_cast<const uint16_t *>(&sv[0][e * kVPitch]);
            const uint16_t *v1 = reinterpret_cast<const uint16_t *>(&sv[1][e * kVPitch]);
            uint32_t *dst = reinterpret_cast<uint32_t *>(child + ((size_t)(env0 + e) * STM_STACK + 3) * STM_OBS_BYTES + oy * STM_OBS_W);
            for (int wq = sub; wq < 21; wq += 8) {
                uint32_t packed = 0;
                if (live) {
                    #pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const int ox = 4 * wq + k;
                        const uint32_t x0 = sx0[ox], fx = sfx[ox];
                        const uint32_t va = ((42u - fx) * v0[x0] + fx * v0[x0 + 1] + 84u) / 168u;
                        const uint32_t vb = ((42u - fx) * v1[x0] + fx * v1[x0 + 1] + 84u) / 168u;
                        packed |= max(va, vb) << (8 * k);
                    }
                }
                dst[wq] = packed;
            }
        }
    }
    // stack shift: child[env][0..2] = parent[env / fanout][1..3], 21 168 contiguous bytes per env; this CTA
    // moves its 1/84 share of the tile's 32 x 1 323 sixteen-byte pieces
    {
        constexpr int kPieces = 3 * STM_OBS_BYTES / 16;                 // 1 323 per env
        constexpr int kShare = (kTileEnvs * kPieces + STM_OBS_H - 1) / STM_OBS_H;   // 504
        for (int k = tid; k < kShare; k += kFrameThreads) {
            int q = oy * kShare + k;
            int e = q / kPieces, off = q - e * kPieces;
            if (e < nenv) {
                size_t pe = (size_t)(env0 + e) / fanout;
                const uint4 *s = reinterpret_cast<const uint4 *>(parent + (pe * STM_STACK + 1) * STM_OBS_BYTES) + off;
                uint4 *d = reinterpret_cast<uint4 *>(child + (size_t)(env0 + e) * STM_STACK * STM_OBS_BYTES) + off;
                *d = *s;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// small kernels
// ---------------------------------------------------------------------------------------------
__global__ void stm_reset_kernel(uint8_t *states, float *rewards, uint8_t *done, int32_t *lives, int n,
                                 uint32_t reset_pc, int lives0)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint4 *g = reinterpret_cast<uint4 *>(states + (size_t)i * STM_STATE_BYTES);
    for (int k = 0; k < 16; k++) g[k] = make_uint4(0, 0, 0, 0);
    stm_env_state *e = reinterpret_cast<stm_env_state *>(states + (size_t)i * STM_STATE_BYTES);
    e->pc = (uint16_t)reset_pc; e->sp = 0xFD; e->p = 0x24; e->tshift = 10;
    e->paddle[0] = e->paddle[1] = STM_PADDLE_INIT;
    if (rewards) rewards[i] = 0.f;
    if (done) done[i] = 0;
    if (lives) lives[i] = lives0;
}

// root injection for stm_bfs: rewards/done zeroed (cule_bfs.py:115-118), lives decoded from RAM
__global__ void stm_root_init_kernel(const uint8_t *states, float *rewards, uint8_t *done, int32_t *lives,
                                     int n, GameParams gp)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    rewards[i] = 0.f; done[i] = 0;
    int lv = 0;
    if (gp.lives_addr != 0xFF) lv = (states[(size_t)i * STM_STATE_BYTES + (gp.lives_addr & 0x7F)] >> gp.lives_shift) & gp.lives_mask;
    lives[i] = lv;
}

__global__ void stm_finalize_kernel(const float *rewards, float *out_rewards, int n, int clip)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float r = rewards[i];
    if (clip) r = (float)((r > 0.f) - (r < 0.f));
    out_rewards[i] = r;
}

enum { T_CLONE = 0, T_STEP = 1, T_FRAMES = 2, T_KINDS = 3 };
struct Timed {
    stm_ctx *ctx; cudaStream_t st; int slot;
    Timed(stm_ctx *c, cudaStream_t s, int kind) : ctx(c), st(s), slot(-1)
    {
        if (c->timing && c->n_timed < stm_ctx::kMaxTimed) {
            slot = c->n_timed++;
            c->ev_kind[slot] = kind;
            cudaEventRecord(c->ev[2 * slot], s);
        }
    }
    ~Timed() { if (slot >= 0) cudaEventRecord(ctx->ev[2 * slot + 1], st); }
};

int ensure_scratch(stm_ctx *ctx, int n_envs)
{
    if (n_envs <= ctx->scratch_envs) return STM_OK;
    int tiles = (n_envs + kTileEnvs - 1) / kTileEnvs;
    if (ctx->d_scratch) { cudaFree(ctx->d_scratch); ctx->d_scratch = nullptr; }
    if (ctx->d_stepped) { cudaFree(ctx->d_stepped); ctx->d_stepped = nullptr; }
    ctx->scratch_envs = 0;
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_scratch, (size_t)tiles * kTileWords * sizeof(uint32_t)));
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_stepped, (size_t)tiles * kTileEnvs));
    ctx->scratch_envs = tiles * kTileEnvs;
    return STM_OK;
}

// the step kernel for this ctx's cartridge: its translated core if the library has one, else the interpreter
void launch_step_kernel(const stm_ctx *ctx, int blocks, size_t smem, cudaStream_t st, const StepArgs &a)
{
    [[maybe_unused]] int k = 0;
#define STM_LAUNCH(name) if (ctx->xl == ++k) { stm_step_kernel<stm::Xl_##name><<<blocks, kStepThreads, smem, st>>>(a); return; }
    STM_XL_LIST(STM_LAUNCH)
#undef STM_LAUNCH
    stm_step_kernel<stm::InterpCore><<<blocks, kStepThreads, smem, st>>>(a);
}

template <class Core>
cudaError_t set_step_attrs(size_t smem, int pct)
{
    cudaError_t e = cudaFuncSetAttribute(stm_step_kernel<Core>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(stm_step_kernel<Core>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
}

int launch_step(stm_ctx *ctx, uint8_t *states, const uint8_t *actions, int n, int frameskip,
                int episodic_life, float weight, float *rewards, uint8_t *done, int32_t *lives,
                int render, cudaStream_t st)
{
    StepArgs a;
    a.states = states; a.actions = actions; a.rewards = rewards; a.done = done; a.lives = lives;
    a.rom = ctx->d_rom; a.lut = ctx->d_lut; a.pre = ctx->d_pre;
    a.scratch = render ? ctx->d_scratch : nullptr;
    a.stepped = render ? ctx->d_stepped : nullptr;
    a.n = n; a.frameskip = frameskip; a.episodic_life = episodic_life; a.render = render;
    a.rom_words = ctx->rom_words; a.weight = weight; a.gp = ctx->gp;
    size_t smem = (size_t)(64 * kStepThreads + ctx->rom_words + 256) * sizeof(uint32_t);
    int blocks = (n + kStepThreads - 1) / kStepThreads;
    { Timed t(ctx, st, T_STEP); launch_step_kernel(ctx, blocks, smem, st, a); }
    ctx->launches++;
    if (render) ctx->last_render_frames = frameskip >= 2 ? 2 : 1;
    CUDA_TRY(ctx, cudaGetLastError());
    return STM_OK;
}

int launch_frames(stm_ctx *ctx, const uint8_t *parent, uint8_t *child, int n, int fanout, cudaStream_t st)
{
    dim3 grid((n + kTileEnvs - 1) / kTileEnvs, STM_OBS_H);
    { Timed t(ctx, st, T_FRAMES);
      stm_frames_kernel<<<grid, kFrameThreads, 0, st>>>(ctx->d_scratch, ctx->d_stepped, parent, child, n, fanout,
                                                       ctx->last_render_frames); }
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    return STM_OK;
}

int launch_clone(stm_ctx *ctx, const uint8_t *ss, const float *sr, const uint8_t *sd, const int32_t *sl,
                 uint8_t *ds, float *dr, uint8_t *dd, int32_t *dl, int n_src, int fanout, cudaStream_t st)
{
    int ntiles = (n_src + kCloneParents - 1) / kCloneParents;
    int grid = ntiles < 148 * 4 ? ntiles : 148 * 4;       // persistent: up to 4 CTAs on each of the 148 SMs
    { Timed t(ctx, st, T_CLONE);
      stm_clone_kernel<<<grid, kCloneThreads, 0, st>>>(reinterpret_cast<const uint4 *>(ss), reinterpret_cast<uint4 *>(ds),
                                                      sr, dr, sd, dd, sl, dl, n_src, fanout); }
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    return STM_OK;
}

template <class T> int grow(stm_ctx *ctx, T *&p, size_t count)
{
    if (p) { cudaFree(p); p = nullptr; }
    CUDA_TRY(ctx, cudaMalloc(&p, count * sizeof(T)));
    return STM_OK;
}

int ensure_pool(stm_ctx *ctx, int which, size_t envs, size_t stack_envs)
{
    int rc;
    if (envs > ctx->pool_envs[which]) {
        ctx->pool_envs[which] = 0;
        if ((rc = grow(ctx, ctx->pool_state[which], envs * STM_STATE_BYTES))) return rc;
        if ((rc = grow(ctx, ctx->pool_reward[which], envs))) return rc;
        if ((rc = grow(ctx, ctx->pool_done[which], envs))) return rc;
        if ((rc = grow(ctx, ctx->pool_lives[which], envs))) return rc;
        ctx->pool_envs[which] = envs;
    }
    if (stack_envs > ctx->pool_stack_envs[which]) {
        ctx->pool_stack_envs[which] = 0;
        if ((rc = grow(ctx, ctx->pool_stack[which], stack_envs * STM_STACK * STM_OBS_BYTES))) return rc;
        ctx->pool_stack_envs[which] = stack_envs;
    }
    return STM_OK;
}

} // namespace

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

// shared with the other translation units of the library (stm_logits.cu, stm_cnn.cu)
int stm_internal_fail(stm_ctx *ctx, int code, const char *msg) { return fail(ctx, code, "%s", msg); }
void stm_internal_count_launch(stm_ctx *ctx) { if (ctx) ctx->launches++; }
int stm_internal_device(const stm_ctx *ctx) { return ctx ? ctx->device : -1; }
void **stm_internal_cnn_slot(stm_ctx *ctx) { return &ctx->cnn; }
void stm_internal_cnn_free(void *p);

const char *stm_last_error(const stm_ctx *ctx) { return ctx ? ctx->err : g_create_error; }
size_t stm_state_bytes(const stm_ctx *) { return STM_STATE_BYTES; }
int stm_num_actions(const stm_ctx *ctx) { return ctx ? ctx->desc.n_actions : 0; }
long long stm_launch_count(const stm_ctx *ctx) { return ctx ? ctx->launches : 0; }
int stm_translated_core(const stm_ctx *ctx) { return ctx ? ctx->xl : 0; }
int stm_max_step_envs(const stm_ctx *ctx) { return ctx ? ctx->scratch_envs : 0; }

int stm_ctx_create(stm_ctx **out, const stm_game_desc *desc, const uint8_t *rom, int device)
{
    if (!out || !desc || !rom) return fail(nullptr, STM_ERR_ARG, "stm_ctx_create: null argument");
    *out = nullptr;
    bool ok_size = (desc->bank_scheme == STM_BANK_NONE && (desc->rom_size == 2048 || desc->rom_size == 4096)) ||
                   (desc->bank_scheme == STM_BANK_F8 && desc->rom_size == 8192);
    if (!ok_size) return fail(nullptr, STM_ERR_ARG, "stm_ctx_create: unsupported rom_size %u / bank_scheme %u", desc->rom_size, desc->bank_scheme);
    if (desc->n_actions < 1 || desc->n_actions > 18) return fail(nullptr, STM_ERR_ARG, "stm_ctx_create: n_actions %u out of range", desc->n_actions);
    if (desc->controller > STM_CTRL_PADDLE) return fail(nullptr, STM_ERR_ARG, "stm_ctx_create: unknown controller");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, STM_ERR_CUDA, "stm_ctx_create: no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(nullptr, STM_ERR_ARG, "stm_ctx_create: device %d out of range", device);
    stm_ctx *ctx = new (std::nothrow) stm_ctx();
    if (!ctx) return fail(nullptr, STM_ERR_NOMEM, "stm_ctx_create: out of host memory");
    ctx->device = device; ctx->desc = *desc;
    GameParams &gp = ctx->gp;
    memset(&gp, 0, sizeof(gp));
    gp.rom_mask = desc->bank_scheme == STM_BANK_F8 ? 0xFFF : desc->rom_size - 1;
    gp.bank_scheme = desc->bank_scheme; gp.controller = desc->controller; gp.n_actions = desc->n_actions;
    memcpy(gp.actions, desc->actions, 18);
    memcpy(gp.score_addr, desc->score_addr, 3); memcpy(gp.score2_addr, desc->score2_addr, 3);
    gp.lives_addr = desc->lives_addr; gp.lives_shift = desc->lives_shift; gp.lives_mask = desc->lives_mask;
    gp.term_addr = desc->term_addr; gp.term_mask = desc->term_mask; gp.term_value = desc->term_value;
    ctx->rom_words = desc->rom_size / 4;
    ctx->reset_pc = rom[0xFFC & gp.rom_mask] | (rom[0xFFD & gp.rom_mask] << 8);   // power-on bank is 0
    uint32_t lut[256];
    build_hot_table(lut);
    int rc = STM_OK;
    auto body = [&]() -> int {
        CUDA_TRY(ctx, cudaSetDevice(device));
        CUDA_TRY(ctx, cudaMalloc(&ctx->d_rom, desc->rom_size));
        CUDA_TRY(ctx, cudaMalloc(&ctx->d_lut, sizeof(lut)));
        CUDA_TRY(ctx, cudaMemcpy(ctx->d_rom, rom, desc->rom_size, cudaMemcpyHostToDevice));
        CUDA_TRY(ctx, cudaMemcpy(ctx->d_lut, lut, sizeof(lut), cudaMemcpyHostToDevice));
        {
            uint32_t *pre = new (std::nothrow) uint32_t[2 * desc->rom_size];
            if (!pre) return fail(ctx, STM_ERR_NOMEM, "stm_ctx_create: out of host memory");
            build_predecode(rom, desc->rom_size, gp.rom_mask, pre);
            // A translated core for this very image?  OFF unless STM_XLATE=1: on B200 the translated kernels
            // execute half the instructions of the interpreter and still lose to it, because their code
            // (one block of SASS per cartridge instruction, 300 KB) misses the instruction caches where the
            // interpreter's 10 KB loop hits; they stay in the library as a measured experiment.
            {
                const char *on = getenv("STM_XLATE");
                const char *no = (on && atoi(on)) ? "0" : "1";
                const uint32_t h = stm::rom_hash(rom, desc->rom_size);
                [[maybe_unused]] int k = 0; (void)h; (void)no;
#define STM_MATCH(name) \
                ++k; \
                if (!(no && atoi(no)) && h == stm::Xl_##name::kRomHash && desc->rom_size == stm::Xl_##name::kRomSize) { \
                    ctx->xl = k; stm::stamp_cases(pre, stm::Xl_##name::offsets(), stm::Xl_##name::kCases); }
                STM_XL_LIST(STM_MATCH)
#undef STM_MATCH
            }
            cudaError_t e1 = cudaMalloc(&ctx->d_pre, 8ull * desc->rom_size);
            if (e1 == cudaSuccess) e1 = cudaMemcpy(ctx->d_pre, pre, 8ull * desc->rom_size, cudaMemcpyHostToDevice);
            delete[] pre;
            CUDA_TRY(ctx, e1);
        }
        // the attribute is per kernel, not per ctx: always allow the largest cartridge (8 KB)
        size_t smem = (size_t)(64 * kStepThreads + 8192 / 4 + 256) * sizeof(uint32_t);
        // Shared memory for the resident CTAs and no more: what is left of the 256 KB is L1, which holds
        // the predecoded cartridge (8 B per image byte) every instruction fetch reads.
        int pct = (int)((STM_STEP_MIN_CTAS * (smem + 1024) * 100 + 228 * 1024 - 1) / (228 * 1024));
        if (const char *e = getenv("STM_STEP_CARVEOUT")) pct = atoi(e);
        if (pct > 100) pct = 100;
        CUDA_TRY(ctx, set_step_attrs<stm::InterpCore>(smem, pct));
        {
            [[maybe_unused]] int k = 0;
#define STM_ATTR(name) if (ctx->xl == ++k) CUDA_TRY(ctx, set_step_attrs<stm::Xl_##name>(smem, pct));
            STM_XL_LIST(STM_ATTR)
#undef STM_ATTR
        }
        return ensure_scratch(ctx, 4096);
    };
    rc = body();
    if (rc != STM_OK) { snprintf(g_create_error, sizeof g_create_error, "%s", ctx->err); stm_ctx_destroy(ctx); return rc; }
    *out = ctx;
    return STM_OK;
}

int stm_ctx_destroy(stm_ctx *ctx)
{
    if (!ctx) return STM_OK;
    cudaSetDevice(ctx->device);
    stm_internal_cnn_free(ctx->cnn);
    cudaFree(ctx->d_rom); cudaFree(ctx->d_lut); cudaFree(ctx->d_pre); cudaFree(ctx->d_scratch); cudaFree(ctx->d_stepped);
    for (int i = 0; i < 2; i++) {
        cudaFree(ctx->pool_state[i]); cudaFree(ctx->pool_reward[i]); cudaFree(ctx->pool_done[i]);
        cudaFree(ctx->pool_lives[i]); cudaFree(ctx->pool_stack[i]);
    }
    cudaFree(ctx->h_in_state); cudaFree(ctx->h_in_stack); cudaFree(ctx->h_out_stack);
    cudaFree(ctx->h_out_reward); cudaFree(ctx->h_out_done); cudaFree(ctx->h_out_lives);
    if (ctx->up_done) cudaEventDestroy(ctx->up_done);
    if (ctx->up_go) cudaEventDestroy(ctx->up_go);
    if (ctx->up_stream) cudaStreamDestroy(ctx->up_stream);
    if (ctx->ev) { for (int i = 0; i < 2 * stm_ctx::kMaxTimed; i++) cudaEventDestroy(ctx->ev[i]); delete[] ctx->ev; }
    delete ctx;
    return STM_OK;
}

int stm_timing_enable(stm_ctx *ctx, int on)
{
    if (!ctx) return fail(ctx, STM_ERR_ARG, "stm_timing_enable: null ctx");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (on && !ctx->ev) {
        ctx->ev = new (std::nothrow) cudaEvent_t[2 * stm_ctx::kMaxTimed];
        if (!ctx->ev) return fail(ctx, STM_ERR_NOMEM, "stm_timing_enable: out of host memory");
        for (int i = 0; i < 2 * stm_ctx::kMaxTimed; i++) CUDA_TRY(ctx, cudaEventCreate(&ctx->ev[i]));
    }
    ctx->timing = on != 0;
    ctx->n_timed = 0;
    return STM_OK;
}

int stm_timing_read(stm_ctx *ctx, float *ms_by_kind, int *launches_by_kind)
{
    if (!ctx || !ms_by_kind || !launches_by_kind) return fail(ctx, STM_ERR_ARG, "stm_timing_read: bad argument");
    for (int k = 0; k < T_KINDS; k++) { ms_by_kind[k] = 0.f; launches_by_kind[k] = 0; }
    for (int i = 0; i < ctx->n_timed; i++) {
        CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev[2 * i + 1]));
        float ms = 0.f;
        CUDA_TRY(ctx, cudaEventElapsedTime(&ms, ctx->ev[2 * i], ctx->ev[2 * i + 1]));
        ms_by_kind[ctx->ev_kind[i]] += ms; launches_by_kind[ctx->ev_kind[i]]++;
    }
    ctx->n_timed = 0;
    return STM_OK;
}

int stm_set_scratch_envs(stm_ctx *ctx, int n_envs)
{
    if (!ctx || n_envs < 1) return fail(ctx, STM_ERR_ARG, "stm_set_scratch_envs: bad argument");
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return ensure_scratch(ctx, n_envs);
}

int stm_reset_states(stm_ctx *ctx, uint8_t *states, float *rewards, uint8_t *done, int32_t *lives, int n, void *stream)
{
    if (!ctx || !states || n < 0) return fail(ctx, STM_ERR_ARG, "stm_reset_states: bad argument");
    if (n == 0) return STM_OK;
    stm_reset_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(states, rewards, done, lives, n, ctx->reset_pc, 0);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    return STM_OK;
}

int stm_expand(stm_ctx *ctx, const uint8_t *src_states, const float *src_rewards, const uint8_t *src_done,
               const int32_t *src_lives, uint8_t *dst_states, float *dst_rewards, uint8_t *dst_done,
               int32_t *dst_lives, int n_src, int fanout, void *stream)
{
    if (!ctx || !src_states || !dst_states || n_src < 0 || fanout < 1) return fail(ctx, STM_ERR_ARG, "stm_expand: bad argument");
    bool all = src_rewards && src_done && src_lives && dst_rewards && dst_done && dst_lives;
    bool none = !src_rewards && !src_done && !src_lives && !dst_rewards && !dst_done && !dst_lives;
    if (!all && !none) return fail(ctx, STM_ERR_ARG, "stm_expand: rewards/done/lives must be all given or all NULL");
    if (((uintptr_t)src_states | (uintptr_t)dst_states) & 15) return fail(ctx, STM_ERR_ARG, "stm_expand: state pointers must be 16-byte aligned");
    if (n_src == 0) return STM_OK;
    return launch_clone(ctx, src_states, src_rewards, src_done, src_lives, dst_states, dst_rewards, dst_done, dst_lives,
                        n_src, fanout, (cudaStream_t)stream);
}

int stm_step(stm_ctx *ctx, uint8_t *states, const uint8_t *actions, int n, int frameskip, int episodic_life,
             float weight, float *rewards, uint8_t *done, int32_t *lives, int render, void *stream)
{
    if (!ctx || !states || !rewards || !done || !lives || n < 0 || frameskip < 1)
        return fail(ctx, STM_ERR_ARG, "stm_step: bad argument");
    if ((uintptr_t)states & 15) return fail(ctx, STM_ERR_ARG, "stm_step: states must be 16-byte aligned");
    if (render && n > ctx->scratch_envs)
        return fail(ctx, STM_ERR_ARG, "stm_step: n=%d exceeds the native-frame scratch (%d envs); call stm_set_scratch_envs", n, ctx->scratch_envs);
    if (n == 0) return STM_OK;
    return launch_step(ctx, states, actions, n, frameskip, episodic_life, weight, rewards, done, lives, render, (cudaStream_t)stream);
}

int stm_frames(stm_ctx *ctx, const uint8_t *parent_stacks, uint8_t *child_stacks, int n, int fanout, void *stream)
{
    if (!ctx || !parent_stacks || !child_stacks || n < 0 || fanout < 1) return fail(ctx, STM_ERR_ARG, "stm_frames: bad argument");
    if (n > ctx->scratch_envs) return fail(ctx, STM_ERR_ARG, "stm_frames: n exceeds the native-frame scratch");
    if (ctx->last_render_frames == 0) return fail(ctx, STM_ERR_ARG, "stm_frames: no rendered stm_step precedes this call");
    if (((uintptr_t)parent_stacks | (uintptr_t)child_stacks) & 15) return fail(ctx, STM_ERR_ARG, "stm_frames: stacks must be 16-byte aligned");
    if (n == 0) return STM_OK;
    return launch_frames(ctx, parent_stacks, child_stacks, n, fanout, (cudaStream_t)stream);
}

// `stacks_ready` (optional): an event the stream must wait for before the first K3 launch reads
// root_stacks — the emulation of the first level does not need the stacks, only the frames kernel does.
static int bfs_impl(stm_ctx *ctx, const uint8_t *root_states, const uint8_t *root_stacks, int n_roots, int depth,
                    float gamma, int frameskip, int episodic_life, int clip_reward, uint8_t *leaf_stacks,
                    float *leaf_rewards, uint8_t *leaf_done, int32_t *leaf_lives, uint8_t *leaf_states, void *stream,
                    cudaEvent_t stacks_ready, const float *node_rewards = nullptr, const uint8_t *node_done = nullptr,
                    const int32_t *node_lives = nullptr, int level0 = 0)
{
    if (!ctx || !root_states || !root_stacks || !leaf_stacks || !leaf_rewards || n_roots < 1 || depth < 1 || frameskip < 1)
        return fail(ctx, STM_ERR_ARG, "stm_bfs: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    const int A = ctx->desc.n_actions;
    // level sizes
    size_t n_level[16];
    if (depth > 15) return fail(ctx, STM_ERR_ARG, "stm_bfs: depth %d too large", depth);
    n_level[0] = (size_t)n_roots;
    for (int k = 1; k <= depth; k++) {
        n_level[k] = n_level[k - 1] * (size_t)A;
        if (n_level[k] > 0x7FFFFFFFull) return fail(ctx, STM_ERR_ARG, "stm_bfs: %d roots x %d^%d leaves overflow int32", n_roots, A, depth);
    }
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    // pools: level k lives in pool k&1.  States/rewards/done/lives of EVERY level (the last one too:
    // leaf_states/done/lives are optional outputs) go through the pools; stacks of the last level are
    // written straight into leaf_stacks.
    size_t need[2] = {0, 0}, need_stack[2] = {0, 0};
    for (int k = 0; k <= depth; k++) {
        if (n_level[k] > need[k & 1]) need[k & 1] = n_level[k];
        if (k >= 1 && k < depth && n_level[k] > need_stack[k & 1]) need_stack[k & 1] = n_level[k];
    }
    int rc;
    for (int w = 0; w < 2; w++) if ((rc = ensure_pool(ctx, w, need[w], need_stack[w]))) return rc;
    int chunk = ctx->scratch_envs;

    // level 0 = the roots
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pool_state[0], root_states, n_level[0] * STM_STATE_BYTES, cudaMemcpyDeviceToDevice, st));
    if (node_rewards) {                                   // resuming from interior nodes: their accumulated state comes in
        CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pool_reward[0], node_rewards, n_level[0] * sizeof(float), cudaMemcpyDeviceToDevice, st));
        CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pool_done[0], node_done, n_level[0], cudaMemcpyDeviceToDevice, st));
        CUDA_TRY(ctx, cudaMemcpyAsync(ctx->pool_lives[0], node_lives, n_level[0] * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    } else {
        stm_root_init_kernel<<<(n_roots + 127) / 128, 128, 0, st>>>(ctx->pool_state[0], ctx->pool_reward[0], ctx->pool_done[0],
                                                                   ctx->pool_lives[0], n_roots, ctx->gp);
        ctx->launches++;
    }
    for (int k = 0; k < depth; k++) {
        int cur = k & 1, nxt = cur ^ 1;
        size_t n_src = n_level[k], n_dst = n_level[k + 1];
        if ((rc = launch_clone(ctx, ctx->pool_state[cur], ctx->pool_reward[cur], ctx->pool_done[cur], ctx->pool_lives[cur],
                               ctx->pool_state[nxt], ctx->pool_reward[nxt], ctx->pool_done[nxt], ctx->pool_lives[nxt],
                               (int)n_src, A, st))) return rc;
        float weight = (float)pow((double)gamma, (double)(level0 + k));
        const uint8_t *pstack = k == 0 ? root_stacks : ctx->pool_stack[cur];
        uint8_t *cstack = k == depth - 1 ? leaf_stacks : ctx->pool_stack[nxt];
        // chunks are multiples of the tile size and of A, so "action = env % A" holds inside a chunk
        size_t step = (size_t)chunk - (size_t)chunk % (size_t)(kTileEnvs * A);
        if (step == 0) return fail(ctx, STM_ERR_ARG, "stm_bfs: scratch too small for %d actions", A);
        for (size_t off = 0; off < n_dst; off += step) {
            int cn = (int)((n_dst - off) < step ? (n_dst - off) : step);
            if ((rc = launch_step(ctx, ctx->pool_state[nxt] + off * STM_STATE_BYTES, nullptr, cn, frameskip, episodic_life, weight,
                                  ctx->pool_reward[nxt] + off, ctx->pool_done[nxt] + off, ctx->pool_lives[nxt] + off, 1, st))) return rc;
            if (stacks_ready) { CUDA_TRY(ctx, cudaStreamWaitEvent(st, stacks_ready, 0)); stacks_ready = nullptr; }
            // parent of env (off+i) is (off+i)/A = off/A + i/A because off % A == 0
            if ((rc = launch_frames(ctx, pstack + (off / A) * STM_STACK * STM_OBS_BYTES, cstack + off * STM_STACK * STM_OBS_BYTES, cn, A, st))) return rc;
        }
    }
    int fin = depth & 1;
    size_t L = n_level[depth];
    stm_finalize_kernel<<<(unsigned)((L + 255) / 256), 256, 0, st>>>(ctx->pool_reward[fin], leaf_rewards, (int)L, clip_reward);
    ctx->launches++;
    if (leaf_done) CUDA_TRY(ctx, cudaMemcpyAsync(leaf_done, ctx->pool_done[fin], L, cudaMemcpyDeviceToDevice, st));
    if (leaf_lives) CUDA_TRY(ctx, cudaMemcpyAsync(leaf_lives, ctx->pool_lives[fin], L * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    if (leaf_states) CUDA_TRY(ctx, cudaMemcpyAsync(leaf_states, ctx->pool_state[fin], L * STM_STATE_BYTES, cudaMemcpyDeviceToDevice, st));
    CUDA_TRY(ctx, cudaGetLastError());
    return STM_OK;
}

int stm_bfs(stm_ctx *ctx, const uint8_t *root_states, const uint8_t *root_stacks, int n_roots, int depth,
            float gamma, int frameskip, int episodic_life, int clip_reward, uint8_t *leaf_stacks,
            float *leaf_rewards, uint8_t *leaf_done, int32_t *leaf_lives, uint8_t *leaf_states, void *stream)
{
    return bfs_impl(ctx, root_states, root_stacks, n_roots, depth, gamma, frameskip, episodic_life, clip_reward,
                    leaf_stacks, leaf_rewards, leaf_done, leaf_lives, leaf_states, stream, nullptr);
}

int stm_bfs_resume(stm_ctx *ctx, const uint8_t *node_states, const uint8_t *node_stacks, const float *node_rewards,
                   const uint8_t *node_done, const int32_t *node_lives, int n_nodes, int level0, int depth_more,
                   float gamma, int frameskip, int episodic_life, int clip_reward, uint8_t *leaf_stacks,
                   float *leaf_rewards, uint8_t *leaf_done, int32_t *leaf_lives, uint8_t *leaf_states, void *stream)
{
    if (!node_rewards || !node_done || !node_lives || level0 < 0)
        return fail(ctx, STM_ERR_ARG, "stm_bfs_resume: bad argument");
    return bfs_impl(ctx, node_states, node_stacks, n_nodes, depth_more, gamma, frameskip, episodic_life, clip_reward,
                    leaf_stacks, leaf_rewards, leaf_done, leaf_lives, leaf_states, stream, nullptr,
                    node_rewards, node_done, node_lives, level0);
}

int stm_bfs_host(stm_ctx *ctx, const uint8_t *root_states, const uint8_t *root_stacks, int n_roots, int depth,
                 float gamma, int frameskip, int episodic_life, int clip_reward, uint8_t *leaf_stacks,
                 int leaf_stacks_on_device, float *leaf_rewards, uint8_t *leaf_done, int32_t *leaf_lives, void *stream)
{
    if (!ctx || !root_states || !root_stacks || !leaf_stacks || !leaf_rewards || n_roots < 1 || depth < 1)
        return fail(ctx, STM_ERR_ARG, "stm_bfs_host: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    size_t L = (size_t)n_roots;
    for (int k = 0; k < depth; k++) L *= (size_t)ctx->desc.n_actions;
    const size_t SB = (size_t)STM_STACK * STM_OBS_BYTES;
    if ((size_t)n_roots > ctx->h_roots) {
        cudaFree(ctx->h_in_state); cudaFree(ctx->h_in_stack); ctx->h_in_state = ctx->h_in_stack = nullptr; ctx->h_roots = 0;
        CUDA_TRY(ctx, cudaMalloc(&ctx->h_in_state, (size_t)n_roots * STM_STATE_BYTES));
        CUDA_TRY(ctx, cudaMalloc(&ctx->h_in_stack, (size_t)n_roots * SB));
        ctx->h_roots = (size_t)n_roots;
    }
    if (L > ctx->h_leaves || (!leaf_stacks_on_device && !ctx->h_out_stack)) {
        cudaFree(ctx->h_out_stack); cudaFree(ctx->h_out_reward); cudaFree(ctx->h_out_done); cudaFree(ctx->h_out_lives);
        ctx->h_out_stack = nullptr; ctx->h_out_reward = nullptr; ctx->h_out_done = nullptr; ctx->h_out_lives = nullptr; ctx->h_leaves = 0;
        if (!leaf_stacks_on_device) CUDA_TRY(ctx, cudaMalloc(&ctx->h_out_stack, L * SB));
        CUDA_TRY(ctx, cudaMalloc(&ctx->h_out_reward, L * sizeof(float)));
        CUDA_TRY(ctx, cudaMalloc(&ctx->h_out_done, L));
        CUDA_TRY(ctx, cudaMalloc(&ctx->h_out_lives, L * sizeof(int32_t)));
        ctx->h_leaves = L;
    }
    if (!ctx->up_stream) {
        CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->up_stream, cudaStreamNonBlocking));
        CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->up_done, cudaEventDisableTiming));
        CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->up_go, cudaEventDisableTiming));
    }
    // The states (256 B per root) go first on the caller's stream; the stacks (28 KB per root, needed
    // only by the first frames kernel) follow on the upload stream, under the first level's emulation.
    CUDA_TRY(ctx, cudaEventRecord(ctx->up_go, st));                      // after whatever the caller queued before us
    CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->up_stream, ctx->up_go, 0));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_in_state, root_states, (size_t)n_roots * STM_STATE_BYTES, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_in_stack, root_stacks, (size_t)n_roots * SB, cudaMemcpyHostToDevice, ctx->up_stream));
    CUDA_TRY(ctx, cudaEventRecord(ctx->up_done, ctx->up_stream));
    int rc = bfs_impl(ctx, ctx->h_in_state, ctx->h_in_stack, n_roots, depth, gamma, frameskip, episodic_life, clip_reward,
                      leaf_stacks_on_device ? leaf_stacks : ctx->h_out_stack, ctx->h_out_reward, ctx->h_out_done,
                      ctx->h_out_lives, nullptr, stream, ctx->up_done);
    if (rc) { cudaStreamSynchronize(ctx->up_stream); return rc; }
    if (!leaf_stacks_on_device) CUDA_TRY(ctx, cudaMemcpyAsync(leaf_stacks, ctx->h_out_stack, L * SB, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(leaf_rewards, ctx->h_out_reward, L * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (leaf_done) CUDA_TRY(ctx, cudaMemcpyAsync(leaf_done, ctx->h_out_done, L, cudaMemcpyDeviceToHost, st));
    if (leaf_lives) CUDA_TRY(ctx, cudaMemcpyAsync(leaf_lives, ctx->h_out_lives, L * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return STM_OK;
}

} // extern "C"

// =============================================================================================
// Allocation-free variant of "one agent step + observation stacks" (stm_step + stm_frames).
//
// The regular path keeps the native 160x210 frames that K2 writes and K3 reads in a buffer the ctx
// obtains from cudaMalloc (and re-obtains when stm_set_scratch_envs grows it).  Some users of this
// library must not trigger a device allocation from inside a call: a program that has handed the
// whole GPU to its own pool at start-up (an inference server with a fixed arena, a PyTorch process
// whose caching allocator owns the card) has nothing left for cudaMalloc to give, and a call made
// from a latency-critical loop cannot afford cudaMalloc's implicit device-wide synchronisation.
// For them the same two kernels run against a scratch area that is part of the library image:
// a file-scope __device__ array, reserved on every GPU that loads the library, from load to exit.
//
// Size: STM_NOALLOC_MAX_ENVS = 128 envs -> 4 tiles x 2 frames x 210 rows x 160 B x 32 lanes
// = 8 601 600 B, plus 128 B of "was stepped" flags.
//
// Rules that follow from sharing one array:
//   - every call uses the same memory, and it still holds the previous call's frames.  Nothing in
//     it has to start at zero — each byte K3 reads (sampled rows of the frames of envs that were
//     stepped, and every env's flag) is written by K2 earlier in the same call — so nothing is
//     cleared;
//   - two calls that overlap on the GPU (different streams, or different threads' contexts) race
//     on it.  Serialise them.
// =============================================================================================
namespace {

constexpr int kNoAllocMaxEnvs = STM_NOALLOC_MAX_ENVS;
static_assert(kNoAllocMaxEnvs % kTileEnvs == 0, "whole tiles");

static __device__ __align__(256) uint32_t g_noalloc_frames[(size_t)(kNoAllocMaxEnvs / kTileEnvs) * kTileWords];
static __device__ __align__(256) uint8_t g_noalloc_stepped[kNoAllocMaxEnvs];

// In host code the two names above are host-side shadows of the device symbols, not device
// addresses: the real addresses are looked up once per device and cached here.
constexpr int kMaxDevices = 64;
uint32_t *g_noalloc_frames_ptr[kMaxDevices];
uint8_t *g_noalloc_stepped_ptr[kMaxDevices];
bool g_noalloc_ready[kMaxDevices];

} // namespace

extern "C" int stm_step_frames_noalloc(stm_ctx *ctx, uint8_t *states, const uint8_t *actions, int n, int frameskip,
                                       int episodic_life, float weight, float *rewards, uint8_t *done, int32_t *lives,
                                       const uint8_t *parent_stacks, uint8_t *child_stacks, int fanout, void *stream)
{
    if (!ctx || !states || !rewards || !done || !lives || !parent_stacks || !child_stacks || n < 0 || frameskip < 1 || fanout < 1)
        return fail(ctx, STM_ERR_ARG, "stm_step_frames_noalloc: bad argument");
    if (n > kNoAllocMaxEnvs)
        return fail(ctx, STM_ERR_ARG, "stm_step_frames_noalloc: n=%d exceeds the built-in scratch of %d envs "
                                      "(use stm_step + stm_frames, whose scratch is allocated and resizable)", n, kNoAllocMaxEnvs);
    if (((uintptr_t)states | (uintptr_t)parent_stacks | (uintptr_t)child_stacks) & 15)
        return fail(ctx, STM_ERR_ARG, "stm_step_frames_noalloc: states and stacks must be 16-byte aligned");
    const int dev = ctx->device;
    if (dev < 0 || dev >= kMaxDevices) return fail(ctx, STM_ERR_ARG, "stm_step_frames_noalloc: device index %d unsupported", dev);
    CUDA_TRY(ctx, cudaSetDevice(dev));
    if (!g_noalloc_ready[dev]) {
        // first call on this device: resolve the device addresses of the built-in scratch
        CUDA_TRY(ctx, cudaGetSymbolAddress((void **)&g_noalloc_frames_ptr[dev], g_noalloc_frames));
        CUDA_TRY(ctx, cudaGetSymbolAddress((void **)&g_noalloc_stepped_ptr[dev], g_noalloc_stepped));
        g_noalloc_ready[dev] = true;
    }
    if (n == 0) return STM_OK;
    cudaStream_t st = (cudaStream_t)stream;
    StepArgs a;
    a.states = states; a.actions = actions; a.rewards = rewards; a.done = done; a.lives = lives;
    a.rom = ctx->d_rom; a.lut = ctx->d_lut; a.pre = ctx->d_pre;
    a.scratch = g_noalloc_frames_ptr[dev];
    a.stepped = g_noalloc_stepped_ptr[dev];
    a.n = n; a.frameskip = frameskip; a.episodic_life = episodic_life; a.render = 1;
    a.rom_words = ctx->rom_words; a.weight = weight; a.gp = ctx->gp;
    const size_t smem = (size_t)(64 * kStepThreads + ctx->rom_words + 256) * sizeof(uint32_t);
    launch_step_kernel(ctx, (n + kStepThreads - 1) / kStepThreads, smem, st, a);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    dim3 grid((n + kTileEnvs - 1) / kTileEnvs, STM_OBS_H);
    stm_frames_kernel<<<grid, kFrameThreads, 0, st>>>(g_noalloc_frames_ptr[dev], g_noalloc_stepped_ptr[dev], parent_stacks,
                                                     child_stacks, n, fanout, frameskip >= 2 ? 2 : 1);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    return STM_OK;
}
