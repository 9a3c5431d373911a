// stm_cnn.cu — K4: NatureCNN + policy/value heads on all leaves as tcgen05 / TMEM implicit GEMMs.
//
// Replaces the library path behind policies/actor_critic_ts.py:199-218 (extract_features ->
// SB3 NatureCNN -> action_net / value_net; cuDNN + cuBLAS fp32 in the reference).  This is the
// only dense contraction on the hot path, so the only place tensor cores are used.
//
// Every layer is one GEMM  D[M, N] = A[M, K] * W[N, K]^T  with bf16 operands and fp32 accumulation:
//   conv1  M = n*400  K = 256 (c,kh,kw)   N = 32    A gathered from the u8 NCHW observation stack
//   conv2  M = n*81   K = 512 (kh,kw,c)   N = 64    A gathered from conv1's NHWC bf16 output
//   conv3  M = n*49   K = 576 (kh,kw,c)   N = 64
//   fc     M = n      K = 3136 (h,w,c)    N = 512   (two N = 256 halves)
//   heads  M = n      K = 512             N = 32    (A action logits + 1 value, zero padded), fp32 out
// "Implicit": the im2col matrix is never written to HBM — each CTA gathers its 128-row x 64-column
// slice of A straight from the activation tensor into shared memory, already in the tensor core's
// 128-byte-swizzled K-major layout, while the previous slice is being
// multiplied.  Weights are pre-tiled once into the same layout, so a slice of W is a linear copy.
//
// Per CTA (persistent: CTAs loop over 128-row output tiles): 9 warps, warp-specialised.
//   warps 0-3  producers: gather the A slice of (tile, k-slice) into a 4-stage shared-memory ring
//              (loads of the next slice are in flight while the current one is converted and
//              stored), fence.proxy.async, arrive on full[stage];
//   warp  8    one thread issues tcgen05.mma (M = 128, N = layer width, K = 16 per instruction, 4 per
//              64-wide slice) into one of TWO TMEM accumulators, tcgen05.commit -> empty[stage],
//              and after the last slice of a tile -> tmem_full[buf];
//   warps 4-7  epilogue: tcgen05.ld (each warp its 32 TMEM lanes), scale + bias (+ ReLU), store;
//              arrive on tmem_empty[buf] — so the epilogue of tile i overlaps the MMAs of tile i+1.
// The whole weight matrix of conv1/conv2/conv3/heads (16-72 KB) stays resident in shared memory
// for the life of the CTA; only fc (3.2 MB) streams its W slices through the ring.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <new>
#include <cstdlib>
#include "../../include/stm_b200.h"

extern "C" int stm_internal_fail(stm_ctx *ctx, int code, const char *msg);
extern "C" void stm_internal_count_launch(stm_ctx *ctx);
extern "C" int stm_internal_device(const stm_ctx *ctx);
extern "C" void **stm_internal_cnn_slot(stm_ctx *ctx);

namespace {

constexpr int kTileM = 128;
constexpr int kChunkK = 64;                         // K elements per shared-memory slice
constexpr int kABytes = kTileM * kChunkK * 2;       // 16 KB

enum { MODE_CONV1 = 0, MODE_NHWC = 1 };

struct LayerGeom {
    int mode;        // MODE_CONV1: u8 NCHW input, K = (c, kh, kw); MODE_NHWC: bf16 input, K = (kh, kw, c)
    int P, OW;       // output positions per image, output width
    int IH, IW, C;   // input height / width / channels
    int KWC;         // KW * C: contiguous K elements per kernel row (MODE_NHWC)
    int S;           // stride
    int K;           // total K (multiple of 64)
    int n_real;      // valid output columns (<= N)
    int relu;
    int out_f32;     // 1: fp32 output with row stride n_real; 0: bf16 output with row stride N
    float scale;     // multiplies the accumulator before the bias (1/255 for conv1)
    int recip_ow;    // ceil(65536 / OW): pos / OW == (pos * recip_ow) >> 16 for every pos < P (checked by the host)
    int f16;         // 1: A and B operands are fp16 (conv1: u8 pixels are exact in fp16 and cheap to make), 0: bf16
    int debug;       // timing experiments only: bit0 producers skip loads, bit1 MMA thread skips the MMAs, bit2 epilogue skips stores, bit3 conv1 only, bit4 conv1 producers skip convert+store, bit5 epilogue skips the TMEM loads
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{ asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count)); }
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

// shared-memory matrix descriptor: K-major, 128-byte swizzle.  A slice is [rows][64 bf16] = one 128-byte
// line per row; the eight 16-byte pieces of row r are stored at piece position (c ^ (r & 7)) — what TMA's
// SWIZZLE_128B would produce — so that both "one lane per row" and "eight lanes per row" writers are
// bank-conflict-light and the tensor core reads conflict-free.  SBO = 1024 B (next 8-row group); LBO is
// unused for a slice exactly one swizzle span wide; tile bases are 1024-byte aligned; successive K = 16
// MMAs advance the start address by 32 bytes inside the span.
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr)
{
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__device__ __forceinline__ uint32_t swz(uint32_t row, uint32_t piece) { return row * 128u + ((piece ^ (row & 7u)) << 4); }
// instruction descriptor, kind::f16: D = f32, A and B both bf16 (format 1) or both fp16 (format 0), K-major, M = 128
__device__ __forceinline__ uint32_t make_idesc(int N, int f16)
{
    const uint32_t fmt = f16 ? 0u : ((1u << 7) | (1u << 10));
    return (1u << 4) | fmt | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
}
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t *bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                 :: "r"(smem_u32(bar)) : "memory");
}

// 8 u8 (two words) -> 8 fp16, exact, on the integer / half2 pipes only (no I2F: the conversion unit runs at a
// quarter of the ALU rate and was conv1's limiter): byte b under a 0x64 high byte is the fp16 number 1024 + b;
// subtracting 1024 leaves b.
__device__ __forceinline__ uint4 u8x8_to_f16x8(uint32_t lo, uint32_t hi)
{
    const __half2 k1024 = __half2half2(__ushort_as_half((unsigned short)0x6400));
    auto two = [&](uint32_t w, uint32_t sel) -> uint32_t {
        uint32_t biased = __byte_perm(w, 0x64646464u, sel);
        __half2 h = __hsub2(*reinterpret_cast<__half2 *>(&biased), k1024);
        return *reinterpret_cast<uint32_t *>(&h);
    };
    uint4 r;
    r.x = two(lo, 0x5140); r.y = two(lo, 0x7362);
    r.z = two(hi, 0x5140); r.w = two(hi, 0x7362);
    return r;
}

constexpr int kEpilogueThreads = 128;
// threads of a CTA with kProdWarps producer warps: producers, 4 epilogue warps, the MMA warp.  The epilogue
// warps must sit at warp indices = 0..3 mod 4 (a warp reads the TMEM lanes 32 * (warp % 4) ...), so kProdWarps % 4 == 0.
__host__ __device__ constexpr int cta_threads(int prod_warps) { return prod_warps * 32 + kEpilogueThreads + 32; }

__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{ asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory"); }

template <int N, bool kResidentW, int kStages, int kProdWarps>
__global__ void __launch_bounds__(cta_threads(kProdWarps), 2)
stm_cnn_gemm_kernel(const void *__restrict__ in, const __nv_bfloat16 *__restrict__ wt, const float *__restrict__ bias,
                    void *__restrict__ out, long long M_total, LayerGeom g, int n_off, int N_total)
{
    constexpr int kBBytes = N * kChunkK * 2;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);      // swizzled tiles need 1024-byte bases
    uint8_t *sA = smem;                                   // kStages x 16 KB
    uint8_t *sB = smem + kStages * kABytes;               // resident: nchunks x kBBytes; streamed: kStages x kBBytes
    __shared__ __align__(8) uint64_t full[kStages], empty[kStages], tfull[2], tempty[2];
    __shared__ uint32_t tmem_base_s;
    static_assert(kProdWarps % 4 == 0, "epilogue warps must start at a multiple of 4");
    constexpr int kProducerThreads = kProdWarps * 32, kThreads2 = cta_threads(kProdWarps), kMmaWarp = kProdWarps + 4;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int nchunks = g.K / kChunkK;
    const long long ntiles = (M_total + kTileM - 1) / kTileM;
    constexpr uint32_t kCols = (2 * N < 32) ? 32 : 2 * N;             // two accumulators

    if (warp == kMmaWarp) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32(&tmem_base_s)), "r"(kCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0) {
        for (int i = 0; i < kStages; i++) { mbar_init(&full[i], kProducerThreads / 32); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], kEpilogueThreads / 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (kResidentW) {
        // whole (pre-tiled) weight matrix of this N range: nchunks slices of kBBytes
        for (int kc = 0; kc < nchunks; kc++) {
            const uint4 *wsrc = reinterpret_cast<const uint4 *>(wt + (size_t)kc * N_total * kChunkK + (size_t)n_off * kChunkK);
            uint4 *wdst = reinterpret_cast<uint4 *>(sB + (size_t)kc * kBBytes);
            for (int i = tid; i < kBBytes / 16; i += kThreads2) wdst[i] = __ldg(wsrc + i);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;

    if (warp < kProdWarps && g.mode != MODE_CONV1) {
        // ===================== producers, bf16 NHWC input: cp.async, kDepth slices in flight =====================
        // The stream of (tile, k-slice) work items is issued kDepth items ahead of the one being
        // published: 16-byte cp.async copies land straight in the ring (no registers), and an item is
        // handed to the MMA thread once its group has completed and been fenced to the async proxy.
        constexpr int kDepth = kStages - 1;
        // eight consecutive lanes fetch the eight 16-byte pieces of one row's 128-byte K slice: every
        // warp-level copy touches four whole lines of global memory and four whole lines of shared memory
        constexpr int kRowsPerPass = kProducerThreads / 8, kPasses = kTileM / kRowsPerPass;
        const uint32_t piece = (uint32_t)tid & 7u, row0 = (uint32_t)tid >> 3;          // rows row0 + kRowsPerPass j
        // All index arithmetic is 32-bit (M_total < 2^31 is checked by the host) and incremental: one division
        // per tile, not two 64-bit ones per row — the address generation, not the copies, was the producers' cost.
        const uint32_t Mt = (uint32_t)M_total, P = (uint32_t)g.P, OW = (uint32_t)g.OW, rOW = (uint32_t)g.recip_ow;
        const int my_tiles = (long long)blockIdx.x < ntiles ? (int)((ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
        const int items = my_tiles * nchunks;
        uint32_t tile = blockIdx.x;
        int kc = 0;
        const uint8_t *rowp[kPasses];
        uint32_t okmask = 0, st = 0, ph = 0, st_pub = 0;
        for (int i = 0; i < items + kDepth; i++) {
            if (i < items) {
                if (kc == 0) {
                    okmask = 0;
                    uint32_t m = tile * kTileM + row0;
                    uint32_t img = m / P, pos = m - img * P;
                    #pragma unroll
                    for (int j = 0; j < kPasses; j++) {
                        const bool ok = m < Mt;
                        const uint32_t oy = (pos * rOW) >> 16, ox = pos - oy * OW;
                        rowp[j] = reinterpret_cast<const uint8_t *>(in)
                                  + ((size_t)img * (uint32_t)(g.IH * g.IW) + (oy * (uint32_t)g.IW + ox) * (uint32_t)g.S) * (uint32_t)(g.C * 2)
                                  + piece * 16;
                        okmask |= (ok ? 1u : 0u) << j;
                        m += kRowsPerPass;
                        if (P == 1) img += kRowsPerPass;
                        else { pos += kRowsPerPass; if (pos >= P) { pos -= P; img++; } }
                    }
                }
                mbar_wait(&empty[st], ph ^ 1);
                const int k0 = kc * kChunkK;
                const int kh = k0 / g.KWC, koff = k0 - kh * g.KWC;
                const uint32_t goff = (uint32_t)((kh * g.IW * g.C + koff) * 2);
                const uint32_t dst0 = smem_u32(sA + st * kABytes);
                if (!(g.debug & 1)) {
                #pragma unroll
                for (int j = 0; j < kPasses; j++) {
                    const uint32_t r = row0 + kRowsPerPass * j;
                    const uint32_t nbytes = (okmask >> j) & 1u ? 16u : 0u;         // 0 => zero fill
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;"
                                 :: "r"(dst0 + swz(r, piece)), "l"(rowp[j] + goff), "r"(nbytes) : "memory");
                }
                }
                if (!kResidentW) {
                    const uint8_t *wsrc = reinterpret_cast<const uint8_t *>(wt + (size_t)kc * N_total * kChunkK + (size_t)n_off * kChunkK);
                    const uint32_t wdst = smem_u32(sB + st * kBBytes);
                    #pragma unroll 4
                    for (int j = tid; j < kBBytes / 16; j += kProducerThreads)
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(wdst + j * 16), "l"(wsrc + (size_t)j * 16) : "memory");
                }
                if (++kc == nchunks) { kc = 0; tile += gridDim.x; }
                if (++st == kStages) { st = 0; ph ^= 1; }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");              // (an empty group past the end keeps the count regular)
            if (i >= kDepth) {
                asm volatile("cp.async.wait_group %0;" :: "n"(kDepth) : "memory");
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();                                             // every lane's copies are fenced: one arrival per warp
                if ((tid & 31) == 0) mbar_arrive(&full[st_pub]);
                if (++st_pub == kStages) st_pub = 0;
            }
        }
    } else if (warp < kProdWarps) {
        // ===================== producers, u8 NCHW input (conv1): registers, converted to fp16 =====================
        // kSplit threads share a row of the tile: thread t owns row t % 128 and the patch rows
        // [8 / kSplit * (t / 128), ...) of it, so one warp covers 32 consecutive rows (coalesced 4-byte loads).
        // The loads of work item i+1 (the next channel, or the first channel of the next tile) are issued before
        // item i is widened to fp16 and stored, so their latency hides under the conversion.  (A cp.async raw-byte
        // ring with two items in flight per thread measured 8 % slower: more instructions, same latency.)
        constexpr int kSplit = kProducerThreads / kTileM, kMyRows = 8 / kSplit;                 // patch rows per thread
        static_assert(kProducerThreads % kTileM == 0 && 8 % kSplit == 0, "whole rows per producer thread");
        const int row = tid % kTileM, kg0 = (tid / kTileM) * kMyRows;
        const uint32_t Mt = (uint32_t)M_total, P = (uint32_t)g.P, OW = (uint32_t)g.OW, rOW = (uint32_t)g.recip_ow;
        const int my_tiles = (long long)blockIdx.x < ntiles ? (int)((ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
        const int items = my_tiles * nchunks;
        uint32_t tile = blockIdx.x, st = 0, ph = 0;
        int kc = 0;
        const uint8_t *src8 = reinterpret_cast<const uint8_t *>(in);
        bool row_ok = false;
        uint2 cur[kMyRows], nxt[kMyRows];
        auto load_next = [&](uint2 *r) {
            if (kc == 0) {
                const uint32_t m = tile * kTileM + row;
                row_ok = m < Mt;
                const uint32_t img = row_ok ? m / P : 0u;
                const uint32_t pos = row_ok ? m - img * P : 0u;
                const uint32_t oy = (pos * rOW) >> 16, ox = pos - oy * OW;
                src8 = reinterpret_cast<const uint8_t *>(in) + (size_t)img * (uint32_t)(g.C * g.IH * g.IW)
                       + (oy * (uint32_t)g.S + kg0) * (uint32_t)g.IW + ox * (uint32_t)g.S;
            }
            const uint8_t *p8 = src8 + (uint32_t)(kc * g.IH * g.IW);
            #pragma unroll
            for (int kg = 0; kg < kMyRows; kg++) {
                const uint32_t *p = reinterpret_cast<const uint32_t *>(p8 + kg * g.IW);
                const bool ld = row_ok && !(g.debug & 1);
                r[kg].x = ld ? __ldg(p) : 0u; r[kg].y = ld ? __ldg(p + 1) : 0u;
            }
            if (++kc == nchunks) { kc = 0; tile += gridDim.x; }
        };
        if (items > 0) load_next(nxt);
        for (int i = 0; i < items; i++) {
            #pragma unroll
            for (int kg = 0; kg < kMyRows; kg++) cur[kg] = nxt[kg];
            if (i + 1 < items) load_next(nxt);
            mbar_wait(&empty[st], ph ^ 1);                          // slot free (first pass: returns at once)
            uint8_t *abase = sA + st * kABytes;
            if (!(g.debug & 16)) {
            #pragma unroll
            for (int kg = 0; kg < kMyRows; kg++)
                *reinterpret_cast<uint4 *>(abase + swz((uint32_t)row, (uint32_t)(kg0 + kg))) = u8x8_to_f16x8(cur[kg].x, cur[kg].y);
            }
            if (!(g.debug & 64)) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if ((tid & 31) == 0) mbar_arrive(&full[st]);
            if (++st == kStages) { st = 0; ph ^= 1; }
        }
    } else if (warp == kMmaWarp) {
        // ===================== MMA issuer =====================
        if ((tid & 31) == 0) {
            const uint32_t idesc = make_idesc(N, g.f16);
            uint32_t it = 0, t = 0;
            for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, t++) {
                const uint32_t buf = t & 1;
                mbar_wait(&tempty[buf], ((t >> 1) & 1) ^ 1);        // the epilogue has drained this accumulator
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_d = tmem_base + buf * N;
                for (int kc = 0; kc < nchunks; kc++, it++) {
                    const uint32_t st = it % kStages, ph = (it / kStages) & 1;
                    mbar_wait(&full[st], ph);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t a0 = smem_u32(sA + st * kABytes);
                    const uint32_t b0 = smem_u32(sB + (kResidentW ? (size_t)kc : (size_t)st) * kBBytes);
                    if (!(g.debug & 2)) {
                    #pragma unroll
                    for (int j = 0; j < kChunkK / 16; j++)
                        umma(tmem_d, make_desc(a0 + j * 32), make_desc(b0 + j * 32), idesc, (kc | j) ? 1u : 0u);
                    }
                    if (g.debug & 128) mbar_arrive(&empty[st]); else
                    umma_commit(&empty[st]);                        // frees the slot when these MMAs are done
                }
                if (g.debug & 128) mbar_arrive(&tfull[buf]); else
                umma_commit(&tfull[buf]);                           // accumulator complete
            }
        }
    } else {
        // ===================== epilogue =====================
        const int et = tid - kProducerThreads;                     // 0..127 = TMEM lane = row in tile
        const int ew = et >> 5;
        uint32_t t = 0;
        for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, t++) {
            const uint32_t buf = t & 1;
            mbar_wait(&tfull[buf], (t >> 1) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const long long m = tile * kTileM + et;
            const bool row_ok = m < M_total;
            const uint32_t taddr = tmem_base + buf * N + ((uint32_t)(ew * 32) << 16);
            #pragma unroll 1
            for (int c0 = 0; c0 < N; c0 += 16) {
                if (g.debug & 32) break;
                uint32_t r[16];
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                             "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                               "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                             : "r"(taddr + (uint32_t)c0));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (row_ok && !(g.debug & 4)) {
                    float v[16];
                    #pragma unroll
                    for (int j = 0; j < 16; j++) {
                        int n = n_off + c0 + j;
                        float x = __uint_as_float(r[j]) * g.scale + (n < g.n_real ? __ldg(bias + n) : 0.f);
                        v[j] = g.relu ? fmaxf(x, 0.f) : x;
                    }
                    if (g.out_f32) {
                        float *o = reinterpret_cast<float *>(out) + (size_t)m * g.n_real;
                        #pragma unroll
                        for (int j = 0; j < 16; j++) if (n_off + c0 + j < g.n_real) o[n_off + c0 + j] = v[j];
                    } else {
                        uint32_t pk[8];
                        #pragma unroll
                        for (int j = 0; j < 8; j++) {
                            __nv_bfloat162 t2 = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
                            pk[j] = *reinterpret_cast<uint32_t *>(&t2);
                        }
                        uint4 *o = reinterpret_cast<uint4 *>(reinterpret_cast<__nv_bfloat16 *>(out) + (size_t)m * N_total + n_off + c0);
                        o[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                        o[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                    }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if ((tid & 31) == 0) mbar_arrive(&tempty[buf]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == kMmaWarp)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"(kCols) : "memory");
}

// fp32 torch-layout weights (n_real, K) -> bf16, K reordered (k = s*C + c <- column c*KHW + s), zero padded
// to N rows, tiled per 64-wide K slice in the 128-byte-swizzled K-major layout the kernel uses:
//   element (n, k): slice k/64, byte offset n*128 + ((((k%64)/8) ^ (n%8)) * 16) + (k%8)*2
__global__ void stm_tile_weights_kernel(const float *__restrict__ w, __nv_bfloat16 *__restrict__ dst, int n_real, int N,
                                        int K, int KHW, int C, int f16)
{
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= N * K) return;
    int n = idx / K, k = idx % K;
    int s = k / C, c = k % C;
    float v = n < n_real ? w[(size_t)n * K + (size_t)c * KHW + s] : 0.f;
    int kc = k / 64, kk = k % 64;
    size_t off = (size_t)kc * N * 64 + (size_t)n * 64 + (((kk / 8) ^ (n % 8)) * 8) + (kk % 8);
    if (f16) reinterpret_cast<__half *>(dst)[off] = __float2half_rn(v);
    else dst[off] = __float2bfloat16_rn(v);
}

struct CnnState {
    int A = 0;
    __nv_bfloat16 *w[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    float *b[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    __nv_bfloat16 *act1 = nullptr, *act2 = nullptr, *act3 = nullptr, *feat = nullptr;
    long long cap_images = 0;
};

template <int N, bool kResidentW, int kStages, int kProdWarps>
int launch_layer(stm_ctx *ctx, const void *in, const __nv_bfloat16 *w, const float *b, void *out, long long M,
                 const LayerGeom &g, int n_off, int N_total, cudaStream_t st)
{
    static bool attr_done = false;
    const int nchunks = g.K / kChunkK;
    size_t smem = (size_t)kStages * kABytes + (size_t)(kResidentW ? nchunks : kStages) * N * kChunkK * 2 + 1024;
    if (!attr_done) {
        cudaFuncSetAttribute(stm_cnn_gemm_kernel<N, kResidentW, kStages, kProdWarps>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 2048);
        cudaFuncSetAttribute(stm_cnn_gemm_kernel<N, kResidentW, kStages, kProdWarps>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        attr_done = true;
    }
    if (M >= (1LL << 31) - kTileM) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn: too many rows for one pass");
    for (int pos = 0; pos < g.P; pos++)
        if (((pos * g.recip_ow) >> 16) != pos / g.OW) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn: reciprocal of OW is not exact");
    long long tiles = (M + kTileM - 1) / kTileM;
    int per_sm = smem <= 110 * 1024 ? 2 : 1;                         // persistent: fill every SM, CTAs loop over tiles
    long long grid = tiles < 148LL * per_sm ? tiles : 148LL * per_sm;
    stm_cnn_gemm_kernel<N, kResidentW, kStages, kProdWarps><<<(unsigned)grid, cta_threads(kProdWarps), smem, st>>>(in, w, b, out, M, g, n_off, N_total);
    stm_internal_count_launch(ctx);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? STM_OK : stm_internal_fail(ctx, STM_ERR_CUDA, cudaGetErrorString(e));
}

void free_state(CnnState *s)
{
    if (!s) return;
    for (int i = 0; i < 5; i++) { cudaFree(s->w[i]); cudaFree(s->b[i]); }
    cudaFree(s->act1); cudaFree(s->act2); cudaFree(s->act3); cudaFree(s->feat);
    delete s;
}

} // namespace

extern "C" {

void stm_internal_cnn_free(void *p) { free_state(reinterpret_cast<CnnState *>(p)); }

int stm_cnn_set_weights(stm_ctx *ctx, const float *conv1_w, const float *conv1_b, const float *conv2_w,
                        const float *conv2_b, const float *conv3_w, const float *conv3_b, const float *fc_w,
                        const float *fc_b, const float *action_w, const float *action_b, const float *value_w,
                        const float *value_b, int n_actions, void *stream)
{
    if (!ctx || !conv1_w || !conv1_b || !conv2_w || !conv2_b || !conv3_w || !conv3_b || !fc_w || !fc_b || !action_w ||
        !action_b || !value_w || !value_b || n_actions < 1 || n_actions > 18)
        return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn_set_weights: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaSetDevice(stm_internal_device(ctx)) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_CUDA, "cudaSetDevice");
    void **slot = stm_internal_cnn_slot(ctx);
    CnnState *s = reinterpret_cast<CnnState *>(*slot);
    if (!s) {
        s = new (std::nothrow) CnnState();
        if (!s) return stm_internal_fail(ctx, STM_ERR_NOMEM, "stm_cnn_set_weights: out of host memory");
        *slot = s;
    }
    s->A = n_actions;
    struct Spec { const float *w, *b; int n_real, N, K, KHW, C; } spec[4] = {
        { conv1_w, conv1_b, 32, 32, 256, 1, 256 },       // (c, kh, kw): torch order already
        { conv2_w, conv2_b, 64, 64, 512, 16, 32 },       // (kh, kw, c)
        { conv3_w, conv3_b, 64, 64, 576, 9, 64 },
        { fc_w, fc_b, 512, 512, 3136, 49, 64 },          // columns (c, h, w) -> (h, w, c)
    };
    for (int i = 0; i < 4; i++) {
        if (!s->w[i] && cudaMalloc(&s->w[i], (size_t)spec[i].N * spec[i].K * 2) != cudaSuccess)
            return stm_internal_fail(ctx, STM_ERR_NOMEM, "stm_cnn_set_weights: cudaMalloc");
        if (!s->b[i] && cudaMalloc(&s->b[i], (size_t)spec[i].N * 4) != cudaSuccess)
            return stm_internal_fail(ctx, STM_ERR_NOMEM, "stm_cnn_set_weights: cudaMalloc");
        int total = spec[i].N * spec[i].K;
        stm_tile_weights_kernel<<<(total + 255) / 256, 256, 0, st>>>(spec[i].w, s->w[i], spec[i].n_real, spec[i].N,
                                                                    spec[i].K, spec[i].KHW, spec[i].C, i == 0 ? 1 : 0);
        stm_internal_count_launch(ctx);
        cudaMemcpyAsync(s->b[i], spec[i].b, (size_t)spec[i].n_real * 4, cudaMemcpyDeviceToDevice, st);
    }
    // heads: rows 0..A-1 = action_net, row A = value_net, zero padded to 32 rows
    if (!s->w[4] && cudaMalloc(&s->w[4], 32 * 512 * 2) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_NOMEM, "cudaMalloc");
    if (!s->b[4] && cudaMalloc(&s->b[4], 32 * 4) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_NOMEM, "cudaMalloc");
    float *tmp = nullptr;
    if (cudaMalloc(&tmp, 32 * 512 * 4) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_NOMEM, "cudaMalloc");
    cudaMemsetAsync(tmp, 0, 32 * 512 * 4, st);
    cudaMemsetAsync(s->b[4], 0, 32 * 4, st);
    cudaMemcpyAsync(tmp, action_w, (size_t)n_actions * 512 * 4, cudaMemcpyDeviceToDevice, st);
    cudaMemcpyAsync(tmp + (size_t)n_actions * 512, value_w, 512 * 4, cudaMemcpyDeviceToDevice, st);
    cudaMemcpyAsync(s->b[4], action_b, (size_t)n_actions * 4, cudaMemcpyDeviceToDevice, st);
    cudaMemcpyAsync(s->b[4] + n_actions, value_b, 4, cudaMemcpyDeviceToDevice, st);
    stm_tile_weights_kernel<<<(32 * 512 + 255) / 256, 256, 0, st>>>(tmp, s->w[4], n_actions + 1, 32, 512, 1, 512, 0);
    stm_internal_count_launch(ctx);
    cudaStreamSynchronize(st);
    cudaFree(tmp);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? STM_OK : stm_internal_fail(ctx, STM_ERR_CUDA, cudaGetErrorString(e));
}

int stm_cnn_forward(stm_ctx *ctx, const uint8_t *obs, long long n, float *heads, void *stream)
{
    if (!ctx || !obs || !heads || n < 0) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn_forward: bad argument");
    CnnState *s = reinterpret_cast<CnnState *>(*stm_internal_cnn_slot(ctx));
    if (!s || !s->w[0]) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn_forward: call stm_cnn_set_weights first");
    if (n == 0) return STM_OK;
    if (cudaSetDevice(stm_internal_device(ctx)) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_CUDA, "cudaSetDevice");
    cudaStream_t st = (cudaStream_t)stream;
    const long long chunk = n < 32768 ? n : 32768;                   // images per pass: bounds the activation scratch
    if (chunk > s->cap_images) {
        cudaFree(s->act1); cudaFree(s->act2); cudaFree(s->act3); cudaFree(s->feat);
        s->act1 = s->act2 = s->act3 = s->feat = nullptr; s->cap_images = 0;
        if (cudaMalloc(&s->act1, (size_t)chunk * 400 * 32 * 2) != cudaSuccess || cudaMalloc(&s->act2, (size_t)chunk * 81 * 64 * 2) != cudaSuccess ||
            cudaMalloc(&s->act3, (size_t)chunk * 49 * 64 * 2) != cudaSuccess || cudaMalloc(&s->feat, (size_t)chunk * 512 * 2) != cudaSuccess)
            return stm_internal_fail(ctx, STM_ERR_NOMEM, "stm_cnn_forward: activation scratch");
        s->cap_images = chunk;
    }
    const int A1 = s->A + 1;
    const char *dbg_env = getenv("STM_CNN_DEBUG");
    const int dbg = dbg_env ? atoi(dbg_env) : 0;
    for (long long i0 = 0; i0 < n; i0 += chunk) {
        long long c = n - i0 < chunk ? n - i0 : chunk;
        int rc;
        LayerGeom g1 = { MODE_CONV1, 400, 20, 84, 84, 4, 0, 4, 256, 32, 1, 0, 1.0f / 255.0f, (65536 + 20 - 1) / 20, 1, dbg };
        if ((rc = launch_layer<32, true, 4, 8>(ctx, obs + (size_t)i0 * 4 * 84 * 84, s->w[0], s->b[0], s->act1, c * 400, g1, 0, 32, st))) return rc;
        if (dbg & 8) continue;                                       // timing experiments: conv1 alone
        LayerGeom g2 = { MODE_NHWC, 81, 9, 20, 20, 32, 4 * 32, 2, 512, 64, 1, 0, 1.0f, (65536 + 9 - 1) / 9, 0, dbg };
        if ((rc = launch_layer<64, true, 8, 4>(ctx, s->act1, s->w[1], s->b[1], s->act2, c * 81, g2, 0, 64, st))) return rc;
        LayerGeom g3 = { MODE_NHWC, 49, 7, 9, 9, 64, 3 * 64, 1, 576, 64, 1, 0, 1.0f, (65536 + 7 - 1) / 7, 0, dbg };
        if ((rc = launch_layer<64, true, 8, 4>(ctx, s->act2, s->w[2], s->b[2], s->act3, c * 49, g3, 0, 64, st))) return rc;
        LayerGeom g4 = { MODE_NHWC, 1, 1, 1, 1, 3136, 3136, 1, 3136, 512, 1, 0, 1.0f, 65536, 0, dbg };
        for (int half = 0; half < 2; half++)
            if ((rc = launch_layer<256, false, 4, 4>(ctx, s->act3, s->w[3], s->b[3], s->feat, c, g4, half * 256, 512, st))) return rc;
        LayerGeom g5 = { MODE_NHWC, 1, 1, 1, 1, 512, 512, 1, 512, A1, 0, 1, 1.0f, 65536, 0, dbg };
        if ((rc = launch_layer<32, true, 4, 4>(ctx, s->feat, s->w[4], s->b[4], heads + (size_t)i0 * A1, c, g5, 0, 32, st))) return rc;
    }
    return STM_OK;
}

} // extern "C"
