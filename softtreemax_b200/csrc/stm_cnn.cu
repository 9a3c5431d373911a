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
#include <cuda.h>                 // CUtensorMap and its enums (types only: the encoder is fetched through the runtime, no libcuda link)
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <new>
#include <cstdlib>
#include "../../include/stm_b200.h"
#include "../../include/stm_b200_debug.h"

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
    int split_chunks; // split-K (weight-gradient GEMMs, whose K is batch x positions): > 0 = CTA column blockIdx.y handles the
                      // K slices [y * split_chunks, (y + 1) * split_chunks) and writes its fp32 partial to out + y * M * n_real
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{ asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count)); }
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 0x40;\n\t"
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
// The same with the descriptors as (low word, high word) pairs.  The high words are constants and the low words
// differ between MMAs only by small immediates (start address in 16-byte units), so an unrolled issue loop costs an
// add or two per MMA instead of rebuilding 64-bit descriptors — the issuing thread is ONE lane and every instruction
// it spends per tile is serial time of the whole CTA (it was ~800 instructions per 36-MMA tile: the bottleneck).
constexpr uint32_t kDescHi = (uint32_t)(1024 >> 4) | (1u << 14) | (2u << 29);      // SBO, version, 128B swizzle
__device__ __forceinline__ uint32_t desc_lo(uint32_t saddr) { return ((saddr & 0x3FFFF) >> 4) | (1u << 16); }
template <bool kAccumulate>
__device__ __forceinline__ void umma_lo(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t idesc)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 da, {%1, %4};\n\t"
        "mov.b64 db, {%2, %4};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t}"
        :: "r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(kDescHi), "n"(kAccumulate ? 1 : 0) : "memory");
}
__device__ __forceinline__ void umma_lo_rt(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 da, {%1, %4};\n\t"
        "mov.b64 db, {%2, %4};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t}"
        :: "r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(kDescHi), "r"(accumulate) : "memory");
}
// one lane of a fully converged warp (elect.sync): the form ptxas recognises, so that tcgen05.mma is issued without
// the per-instruction election loop it wraps around an MMA that sits under an ordinary divergent branch
__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
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
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{ asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory"); }
// TMA: one 2-D tile (box of the tensor map, 128B-swizzled on the way in) from THIS GPU's global memory into shared memory;
// completion is signalled on the mbarrier as transaction bytes.  No multicast: one CTA, one destination.
__device__ __forceinline__ void tma_load_2d(void *dst_smem, const CUtensorMap *map, int c0, int c1, uint64_t *bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 :: "r"(smem_u32(dst_smem)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}

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
    const int kc0 = g.split_chunks > 0 ? (int)blockIdx.y * g.split_chunks : 0;
    const int nchunks = g.split_chunks > 0 ? min(g.split_chunks, g.K / kChunkK - kc0) : g.K / kChunkK;
    const long long ntiles = nchunks > 0 ? (M_total + kTileM - 1) / kTileM : 0;
    // two accumulators (the epilogue of tile i overlaps the MMAs of tile i + 1) — or ONE for the wide tile with a short ring:
    // that block needs half the tensor memory and half the shared memory, so two of them share an SM and overlap each other
    constexpr int kAccG = (N == 256 && kStages <= 2) ? 1 : 2;
    constexpr uint32_t kCols = (kAccG * N < 32) ? 32 : kAccG * N;
    if (g.split_chunks > 0) out = reinterpret_cast<float *>(out) + (size_t)blockIdx.y * (size_t)M_total * (size_t)g.n_real;

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
                const int k0 = (kc0 + kc) * kChunkK;
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
                    const uint8_t *wsrc = reinterpret_cast<const uint8_t *>(wt + (size_t)(kc0 + kc) * N_total * kChunkK + (size_t)n_off * kChunkK);
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
            const uint32_t a_lo0 = desc_lo(smem_u32(sA)), b_lo0 = desc_lo(smem_u32(sB));          // + start address in 16-byte units
            uint32_t it = 0, t = 0;
            for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, t++) {
                const uint32_t buf = t % kAccG;
                mbar_wait(&tempty[buf], ((t / kAccG) & 1) ^ 1);     // the epilogue has drained this accumulator
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_d = tmem_base + buf * N;
                for (int kc = 0; kc < nchunks; kc++, it++) {
                    const uint32_t st = it % kStages, ph = (it / kStages) & 1;
                    mbar_wait(&full[st], ph);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t a_lo = a_lo0 + st * (uint32_t)(kABytes >> 4);
                    const uint32_t b_lo = b_lo0 + (kResidentW ? (uint32_t)kc : st) * (uint32_t)(kBBytes >> 4);
                    if (!(g.debug & 2)) {
                        umma_lo_rt(tmem_d, a_lo, b_lo, idesc, kc ? 1u : 0u);
                        #pragma unroll
                        for (int j = 1; j < kChunkK / 16; j++)
                            umma_lo<true>(tmem_d, a_lo + 2u * j, b_lo + 2u * j, idesc);
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
            const uint32_t buf = t % kAccG;
            mbar_wait(&tfull[buf], (t / kAccG) & 1);
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

// ---------------------------------------------------------------------------------------------------------
// Window GEMM for the stride-1 / stride-2 convolutions on bf16 NHWC activations (conv2, conv3).
//
// View the input of an image as rows of 128 bytes (conv3: one pixel = 64 channels; conv2: two pixels of 32
// channels) and number the output positions on the INPUT grid ("virtual" rows: conv3 9x9 per image of which
// 7x7 are real outputs, conv2 10x20 of which 9x9): then the im2col row of virtual output v for kernel tap t is
// input row v + off[t], for every v — the A operand of tap t is the SAME shared-memory tile as for every other
// tap, started off[t] rows further in.  The tensor core takes such a start address (128-byte aligned inside a
// 128B-swizzled tile whose pieces are keyed by the absolute row: probed, tools/shift_probe.py), so a tile of 128
// virtual rows needs ONE load of 160 contiguous input rows (20 KB) instead of 8 or 9 gathered 16 KB slices: the
// producers issue 10 copies per thread and tile instead of 64-72, and the L2 -> shared traffic drops 6-7x.
// The price is tensor work on the virtual rows that are not outputs (x2.5 for conv2, x1.65 for conv3) — the
// tensor pipe was 9 % busy — and an epilogue that skips them.
// ---------------------------------------------------------------------------------------------------------
struct WinGeom {
    int V, VW, OH, OW;       // virtual rows per image, virtual grid width, real output height / width
    int recip_vw;            // ceil(65536 / VW)
    int nchunks;             // kernel taps = 64-wide K slices
    int off[9];              // input-row offset of every tap
    int relu;
    int debug;               // timing experiments only (STM_CNN_DEBUG): 1 no loads, 2 no MMAs, 4 no epilogue math/stores, 32 no TMEM loads
};
constexpr int kWinRows = 160, kWinBytes = kWinRows * 128;          // 128 virtual rows + 32 rows of halo
// rows of the window tile of conv<kConv> (a multiple of 8: whole swizzle atoms): conv3 needs 128 + 20, and the 1 KB a stage
// saves is what lets two blocks with its 72 KB of weights share an SM
template <int kConv> __host__ __device__ constexpr int win_rows() { return kConv == 3 ? 152 : kWinRows; }

// tap -> input-row offset, known at compile time so that the issue loop adds immediates (kConv = 2: conv2, 3: conv3)
template <int kConv> __host__ __device__ constexpr int win_taps() { return kConv == 2 ? 8 : 9; }
template <int kConv> __host__ __device__ constexpr int win_off(int kc) { return kConv == 2 ? (kc >> 1) * 10 + (kc & 1) : (kc / 3) * 9 + kc % 3; }

template <int kStages, int kConv>
__global__ void __launch_bounds__(288, 2)
stm_cnn_window_kernel(const __grid_constant__ CUtensorMap in_map, const __nv_bfloat16 *__restrict__ wt, const float *__restrict__ bias,
                      __nv_bfloat16 *__restrict__ out, uint32_t M_total, WinGeom g)
{
    constexpr int N = 64, kBBytes = N * kChunkK * 2;
    constexpr int kWinBytes = win_rows<kConv>() * 128;     // shadows the 160-row default: conv3's taps reach 20 rows, its window is 152
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t *sW = smem;                                    // kStages windows of 19 / 20 KB
    uint8_t *sB = smem + kStages * kWinBytes;              // nchunks x 8 KB, resident
    // kAcc accumulators in tensor memory (64 columns each): the MMA thread may run kAcc tiles ahead of the epilogue,
    // which hides the latency of the tfull -> epilogue -> tempty hand-off (measured: ~1.5 us per round trip)
    constexpr int kAcc = kStages <= 3 ? 4 : 8;            // a short ring = the two-blocks-per-SM configuration: half the tensor memory each
    __shared__ __align__(8) uint64_t full[kStages], empty[kStages], tfull[kAcc], tempty[kAcc];
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(16) float sbias[N];
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t ntiles = (M_total + kTileM - 1) / kTileM;
    constexpr uint32_t kCols = kAcc * N;                  // 512 (the whole tensor memory of the SM: one block per SM) or 256
    if (tid < N) sbias[tid] = bias[tid];

    if (warp == 8) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32(&tmem_base_s)), "r"(kCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0) {
        for (int i = 0; i < kStages; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }     // full: one expect_tx arrival + the TMA's bytes
        for (int i = 0; i < kAcc; i++) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&in_map)) : "memory");
    }
    for (int kc = 0; kc < g.nchunks; kc++) {
        const uint4 *wsrc = reinterpret_cast<const uint4 *>(wt + (size_t)kc * N * kChunkK);
        uint4 *wdst = reinterpret_cast<uint4 *>(sB + (size_t)kc * kBBytes);
        for (int i = tid; i < kBBytes / 16; i += 288) wdst[i] = __ldg(wsrc + i);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;

    if (warp < 4) {
        // ===================== producer: one TMA per tile =====================
        // The window of a tile is 160 consecutive 128-byte rows of the activation matrix: ONE cp.async.bulk.tensor (box 64 x 160,
        // SWIZZLE_128B — the layout the MMA descriptors expect, written by the copy engine itself) issued by one thread; rows past
        // the end of the matrix are zero-filled by the hardware.  The ring runs kStages tiles ahead of the MMA warp.
        if (tid == 0) {
            uint32_t st = 0, ph = 0;
            for (uint32_t tile = blockIdx.x; tile < ntiles && !(g.debug & 2048); tile += gridDim.x) {
                mbar_wait(&empty[st], ph ^ 1);
                mbar_arrive_expect_tx(&full[st], kWinBytes);
                tma_load_2d(sW + st * kWinBytes, &in_map, 0, (int)(tile * kTileM), &full[st]);
                if (++st == kStages) { st = 0; ph ^= 1; }
            }
        }
    } else if (warp == 8) {
        // ===================== MMA issuer: every tap reads the same window, started off[tap] rows in =====================
        // The whole warp walks the tile loop converged (all lanes wait on the barriers); one elected lane issues.
        {
            const uint32_t idesc = make_idesc(N, 0);
            const uint32_t b_lo0 = desc_lo(smem_u32(sB));
            uint32_t st = 0, ph = 0, t = 0;
            for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, t++) {
                const uint32_t buf = t % kAcc;
                if (!(g.debug & 1024)) mbar_wait(&tempty[buf], ((t / kAcc) & 1) ^ 1);
                if (!(g.debug & 2048)) mbar_wait(&full[st], ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_d = tmem_base + buf * N;
                const uint32_t a_lo0 = desc_lo(smem_u32(sW + st * kWinBytes));
                if (elect_one()) {
                    if (!(g.debug & 2)) {
                        #pragma unroll
                        for (int kc = 0; kc < win_taps<kConv>(); kc++) {
                            #pragma unroll
                            for (int j = 0; j < kChunkK / 16; j++) {
                                const uint32_t a_lo = a_lo0 + (uint32_t)(win_off<kConv>(kc) * 8 + 2 * j), b_lo = b_lo0 + (uint32_t)(kc * (kBBytes >> 4) + 2 * j);
                                if (kc == 0 && j == 0) umma_lo<false>(tmem_d, a_lo, b_lo, idesc);
                                else umma_lo<true>(tmem_d, a_lo, b_lo, idesc);
                            }
                        }
                    }
                    if (!(g.debug & 2048)) umma_commit(&empty[st]);
                    if (!(g.debug & 1024)) umma_commit(&tfull[buf]);
                }
                __syncwarp();
                if (++st == kStages) { st = 0; ph ^= 1; }
            }
        }
    } else {
        // ===================== epilogue: only the virtual rows that are real outputs are stored =====================
        const int et = tid - 128, ew = et >> 5;
        uint32_t t = 0;
        const uint32_t V = (uint32_t)g.V, VW = (uint32_t)g.VW, P = (uint32_t)(g.OH * g.OW);
        for (uint32_t tile = blockIdx.x; tile < ntiles && !(g.debug & 1024); tile += gridDim.x, t++) {
            const uint32_t buf = t % kAcc;
            mbar_wait(&tfull[buf], (t / kAcc) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t v = tile * kTileM + et;
            const uint32_t img = v / V, q = v - img * V;
            const uint32_t oy = (q * (uint32_t)g.recip_vw) >> 16, ox = q - oy * VW;
            const bool valid = v < M_total && oy < (uint32_t)g.OH && ox < (uint32_t)g.OW && !(g.debug & 4);
            __nv_bfloat16 *orow = out + ((size_t)img * P + oy * (uint32_t)g.OW + ox) * N;
            const uint32_t taddr = tmem_base + buf * N + ((uint32_t)(ew * 32) << 16);
            // the whole 64-column accumulator row in two 32-column TMEM loads and ONE wait (it was four loads, each waited for)
            uint32_t r[64];
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(taddr));
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                         : "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]), "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]), "=r"(r[48]), "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]), "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63]) : "r"(taddr + 32u));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            // the accumulator is in registers: hand the buffer back to the MMA warp before the arithmetic and the stores
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if ((tid & 31) == 0) mbar_arrive(&tempty[buf]);
            if (valid) {
                #pragma unroll
                for (int c0 = 0; c0 < N; c0 += 16) {
                    uint32_t pk[8];
                    float bb[16];                               // sixteen biases as four 16-byte shared loads
                    #pragma unroll
                    for (int j = 0; j < 4; j++) *reinterpret_cast<float4 *>(&bb[4 * j]) = *reinterpret_cast<const float4 *>(&sbias[c0 + 4 * j]);
                    #pragma unroll
                    for (int j = 0; j < 8; j++) {
                        float x0 = __uint_as_float(r[c0 + 2 * j]) + bb[2 * j];
                        float x1 = __uint_as_float(r[c0 + 2 * j + 1]) + bb[2 * j + 1];
                        if (g.relu) { x0 = fmaxf(x0, 0.f); x1 = fmaxf(x1, 0.f); }
                        __nv_bfloat162 t2 = __floats2bfloat162_rn(x0, x1);
                        pk[j] = *reinterpret_cast<uint32_t *>(&t2);
                    }
                    uint4 *o = reinterpret_cast<uint4 *>(orow + c0);
                    o[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    o[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 8)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"(kCols) : "memory");
}

// ---------------------------------------------------------------------------------------------------------
// conv1 as a window GEMM.  The 8x8 stride-4 convolution over the 4 x 84 x 84 u8 stack is a 2x2 stride-1
// convolution over the 21 x 21 grid of 4x4-pixel blocks with 64 "channels" (4 frames x 4 x 4 pixels): virtual
// output v = block index, tap (dy, dx) = window row v + 21 dy + dx.  The producers BUILD the window: thread
// (row r, half h) fetches the eight 4-byte pixel runs (2 frames x 4 image rows) of block r that make up half of
// its 128-byte fp16 row, widens them on the integer / half2 pipes and stores them swizzled — every pixel is
// loaded and converted ONCE per tile (the gather kernel did it four times), the loads of the next tile are in
// flight while this one is converted.  320 of the 384 producer threads work (160 window rows x 2 halves).
// ---------------------------------------------------------------------------------------------------------
constexpr int kC1Threads = 12 * 32 + 128 + 32;
#ifndef STM_W2_STAGES
#define STM_W2_STAGES 2     /* conv2: window tiles in the ring; 2 x 2 blocks per SM measured 2.62 ms per 65 536 images for K4, 6 x 1 2.77 */
#endif
#ifndef STM_W2_PER_SM
#define STM_W2_PER_SM 2
#endif
#ifndef STM_FC_STAGES
#define STM_FC_STAGES 2     /* fc: stages of the A/B ring; <= 2 selects the one-accumulator block that fits an SM twice (K4 2.56 -> 2.48 ms per 65 536 images against 4 stages, two accumulators, one block per SM) */
#endif
#ifndef STM_W3_STAGES
#define STM_W3_STAGES 2     /* conv3: 2 x 2 blocks per SM (152-row windows, 72 KB of weights each): K4 2.62 -> 2.57 ms per 65 536 images */
#endif
#ifndef STM_W3_PER_SM
#define STM_W3_PER_SM 2
#endif
#ifndef STM_C1_STAGES
#define STM_C1_STAGES 4     /* window tiles (20 KB each) in the shared-memory ring: 4 x 2 blocks per SM measured 2.78 ms per 65 536 images, 6 x 1 2.86, 10 x 1 2.95 */
#endif
#ifndef STM_C1_PER_SM
#define STM_C1_PER_SM 2     /* persistent blocks per SM (each 8 x 32 TMEM columns, 56 registers x 544 threads) */
#endif
#ifndef STM_C1_AHEAD
#define STM_C1_AHEAD 2      /* tiles the producers' global loads run ahead of the conversion */
#endif
template <int kStages>
__global__ void __launch_bounds__(kC1Threads, STM_C1_PER_SM)
stm_cnn_conv1_window_kernel(const uint8_t *__restrict__ in, const __nv_bfloat16 *__restrict__ wt, const float *__restrict__ bias,
                            __nv_bfloat16 *__restrict__ out, uint32_t M_total, int debug)
{
    constexpr int N = 32, kBBytes = N * kChunkK * 2, kTaps = 4, kAcc = 8, kProd = 384, kMmaWarp = 16;
    constexpr int V = 441, VW = 21, OH = 20, OW = 20, P = 400;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t *sW = smem;                                    // kStages windows of 20 KB
    uint8_t *sB = smem + kStages * kWinBytes;              // 4 taps x 4 KB, resident
    __shared__ __align__(8) uint64_t full[kStages], empty[kStages], tfull[kAcc], tempty[kAcc];
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(16) float sbias[N];
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t ntiles = (M_total + kTileM - 1) / kTileM;
    constexpr uint32_t kCols = kAcc * N;                   // 256 columns
    if (tid < N) sbias[tid] = bias[tid];
    if (warp == kMmaWarp) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32(&tmem_base_s)), "r"(kCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0) {
        for (int i = 0; i < kStages; i++) { mbar_init(&full[i], 10); mbar_init(&empty[i], 1); }
        for (int i = 0; i < kAcc; i++) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < kTaps * kBBytes / 16; i += kC1Threads)
        reinterpret_cast<uint4 *>(sB)[i] = __ldg(reinterpret_cast<const uint4 *>(wt) + i);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;

    if (warp < 10) {
        // ===================== producers =====================
        const uint32_t r = (uint32_t)tid % kWinRows, half = (uint32_t)tid / kWinRows;
        const int my_tiles = blockIdx.x < ntiles ? (int)((ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
        uint32_t tile = blockIdx.x, st = 0, ph = 0;
        // the global loads run kAhead tiles ahead of the conversion (one tile, ~0.9 us, does not cover DRAM latency under load:
        // the producers sat on the scoreboard, ncu long-scoreboard 6.1 per issue)
        constexpr int kAhead = STM_C1_AHEAD;
        uint32_t cur[8], pre[kAhead][8];
        auto load_next = [&](uint32_t *w) {
            const uint32_t v = tile * kTileM + r;
            const bool ok = v < M_total && !(debug & 1);
            const uint32_t img = v / V, q = v - img * V, by = q / VW, bx = q - by * VW;
            const uint8_t *p = in + (size_t)img * (4 * 84 * 84) + (2 * half) * (84 * 84) + (4 * by) * 84 + 4 * bx;
            #pragma unroll
            for (int cc = 0; cc < 2; cc++)
                #pragma unroll
                for (int iy = 0; iy < 4; iy++)
                    w[cc * 4 + iy] = ok ? __ldg(reinterpret_cast<const uint32_t *>(p + cc * (84 * 84) + iy * 84)) : 0u;
            tile += gridDim.x;
        };
        #pragma unroll
        for (int d = 0; d < kAhead; d++) if (my_tiles > d) load_next(pre[d]);
        const __half2 k1024 = __half2half2(__ushort_as_half((unsigned short)0x6400));
        auto widen = [&](uint32_t w, uint32_t sel) -> uint32_t {
            uint32_t biased = __byte_perm(w, 0x64646464u, sel);
            __half2 h = __hsub2(*reinterpret_cast<__half2 *>(&biased), k1024);
            return *reinterpret_cast<uint32_t *>(&h);
        };
        for (int i = 0; i < my_tiles; i++) {
            #pragma unroll
            for (int k = 0; k < 8; k++) {
                cur[k] = pre[0][k];
                #pragma unroll
                for (int d = 0; d + 1 < kAhead; d++) pre[d][k] = pre[d + 1][k];
            }
            if (i + kAhead < my_tiles) load_next(pre[kAhead - 1]);
            mbar_wait(&empty[st], ph ^ 1);
            uint8_t *wbase = sW + st * kWinBytes;
            if (!(debug & 16)) {
            #pragma unroll
            for (int k = 0; k < 4; k++) {                           // 16-byte piece 4 half + k = (frame 2 half + k / 2, image rows 2 (k & 1), +1)
                const uint32_t w0 = cur[(k >> 1) * 4 + 2 * (k & 1)], w1 = cur[(k >> 1) * 4 + 2 * (k & 1) + 1];
                *reinterpret_cast<uint4 *>(wbase + swz(r, 4 * half + k)) =
                    make_uint4(widen(w0, 0x5140), widen(w0, 0x7362), widen(w1, 0x5140), widen(w1, 0x7362));
            }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if ((tid & 31) == 0) mbar_arrive(&full[st]);
            if (++st == kStages) { st = 0; ph ^= 1; }
        }
    } else if (warp == kMmaWarp) {
        // the whole warp walks the tile loop converged; one elected lane issues
        const uint32_t idesc = make_idesc(N, 1);
        const uint32_t b_lo0 = desc_lo(smem_u32(sB));
        uint32_t st = 0, ph = 0, t = 0;
        for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, t++) {
            const uint32_t buf = t % kAcc;
            mbar_wait(&tempty[buf], ((t / kAcc) & 1) ^ 1);
            mbar_wait(&full[st], ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t tmem_d = tmem_base + buf * N;
            const uint32_t a_lo0 = desc_lo(smem_u32(sW + st * kWinBytes));
            if (elect_one()) {
                if (!(debug & 2)) {
                    #pragma unroll
                    for (int kc = 0; kc < kTaps; kc++) {
                        #pragma unroll
                        for (int j = 0; j < kChunkK / 16; j++) {
                            const uint32_t a_lo = a_lo0 + (uint32_t)(((kc >> 1) * VW + (kc & 1)) * 8 + 2 * j);
                            const uint32_t b_lo = b_lo0 + (uint32_t)(kc * (kBBytes >> 4) + 2 * j);
                            if (kc == 0 && j == 0) umma_lo<false>(tmem_d, a_lo, b_lo, idesc);
                            else umma_lo<true>(tmem_d, a_lo, b_lo, idesc);
                        }
                    }
                }
                umma_commit(&empty[st]);
                umma_commit(&tfull[buf]);
            }
            __syncwarp();
            if (++st == kStages) { st = 0; ph ^= 1; }
        }
    } else if (warp >= 12 && warp < 16) {
        // ===================== epilogue =====================
        const int et = tid - kProd, ew = et >> 5;
        uint32_t t = 0;
        for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, t++) {
            const uint32_t buf = t % kAcc;
            mbar_wait(&tfull[buf], (t / kAcc) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t v = tile * kTileM + et;
            const uint32_t img = v / V, q = v - img * V, oy = q / VW, ox = q - oy * VW;
            const bool valid = v < M_total && oy < OH && ox < OW && !(debug & 4);
            __nv_bfloat16 *orow = out + ((size_t)img * P + oy * OW + ox) * N;
            const uint32_t taddr = tmem_base + buf * N + ((uint32_t)(ew * 32) << 16);
            uint32_t rr[32];                                    // the 32-column accumulator row in one TMEM load
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                         : "=r"(rr[0]), "=r"(rr[1]), "=r"(rr[2]), "=r"(rr[3]), "=r"(rr[4]), "=r"(rr[5]), "=r"(rr[6]), "=r"(rr[7]), "=r"(rr[8]), "=r"(rr[9]), "=r"(rr[10]), "=r"(rr[11]), "=r"(rr[12]), "=r"(rr[13]), "=r"(rr[14]), "=r"(rr[15]), "=r"(rr[16]), "=r"(rr[17]), "=r"(rr[18]), "=r"(rr[19]), "=r"(rr[20]), "=r"(rr[21]), "=r"(rr[22]), "=r"(rr[23]), "=r"(rr[24]), "=r"(rr[25]), "=r"(rr[26]), "=r"(rr[27]), "=r"(rr[28]), "=r"(rr[29]), "=r"(rr[30]), "=r"(rr[31]) : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");     // registers hold it: the buffer goes back before the arithmetic
            __syncwarp();
            if ((tid & 31) == 0) mbar_arrive(&tempty[buf]);
            if (valid) {
                #pragma unroll
                for (int c0 = 0; c0 < N; c0 += 16) {
                    uint32_t pk[8];
                    float bb[16];
                    #pragma unroll
                    for (int j = 0; j < 4; j++) *reinterpret_cast<float4 *>(&bb[4 * j]) = *reinterpret_cast<const float4 *>(&sbias[c0 + 4 * j]);
                    #pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const float x0 = fmaxf(__uint_as_float(rr[c0 + 2 * j]) * (1.0f / 255.0f) + bb[2 * j], 0.f);
                        const float x1 = fmaxf(__uint_as_float(rr[c0 + 2 * j + 1]) * (1.0f / 255.0f) + bb[2 * j + 1], 0.f);
                        __nv_bfloat162 t2 = __floats2bfloat162_rn(x0, x1);
                        pk[j] = *reinterpret_cast<uint32_t *>(&t2);
                    }
                    uint4 *o = reinterpret_cast<uint4 *>(orow + c0);
                    o[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    o[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == kMmaWarp)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"(kCols) : "memory");
}

// Probe for the "shifted window" operand trick: a [rows][64] bf16 matrix sits in shared memory in the 128-byte
// swizzled K-major layout (piece position keyed by the ABSOLUTE row, i.e. by the address); the MMA's A operand is
// the 128-row window that starts `shift` rows in.  Does the tensor core read it correctly when the start address
// is only 128-byte aligned, and what must the descriptor's base-offset field (bits 49..51) hold?
__global__ void __launch_bounds__(128)
stm_shift_probe_kernel(const __nv_bfloat16 *__restrict__ A, const __nv_bfloat16 *__restrict__ B, float *__restrict__ out,
                       int rows, int shift, int base_off)
{
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t *sA = smem;                                  // rows x 128 B
    uint8_t *sB = smem + 32 * 1024;                      // 64 x 128 B
    __shared__ __align__(8) uint64_t done;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < rows * 8; i += 128) {
        int r = i >> 3, pc = i & 7;
        *reinterpret_cast<uint4 *>(sA + swz(r, pc)) = *reinterpret_cast<const uint4 *>(A + (size_t)r * 64 + pc * 8);
    }
    for (int i = tid; i < 64 * 8; i += 128) {
        int r = i >> 3, pc = i & 7;
        *reinterpret_cast<uint4 *>(sB + swz(r, pc)) = *reinterpret_cast<const uint4 *>(B + (size_t)r * 64 + pc * 8);
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(&tmem_base_s)), "r"(64) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0) { mbar_init(&done, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_s;
    if (tid == 0) {
        const uint32_t idesc = make_idesc(64, 0);
        const uint32_t a0 = smem_u32(sA) + (uint32_t)shift * 128u, b0 = smem_u32(sB);
        for (int j = 0; j < 4; j++) {
            uint64_t ad = make_desc(a0 + j * 32) | ((uint64_t)(base_off & 7) << 49);
            umma(tmem, ad, make_desc(b0 + j * 32), idesc, j ? 1u : 0u);
        }
        umma_commit(&done);
    }
    mbar_wait(&done, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
    for (int c0 = 0; c0 < 64; c0 += 16) {
        uint32_t r[16];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                     "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                       "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                     : "r"(taddr + (uint32_t)c0));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int j = 0; j < 16; j++) out[(size_t)tid * 64 + c0 + j] = __uint_as_float(r[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(64) : "memory");
}

// fp32 torch-layout weights (n_real, K) -> bf16, K reordered (k = s*C + c <- column c*KHW + s), zero padded
// to N rows, tiled per 64-wide K slice in the 128-byte-swizzled K-major layout the kernel uses:
//   element (n, k): slice k/64, byte offset n*128 + ((((k%64)/8) ^ (n%8)) * 16) + (k%8)*2
__global__ void stm_tile_weights_kernel(const float *__restrict__ w, __nv_bfloat16 *__restrict__ dst, int n_real, int N,
                                        int K, int KHW, int C, int f16, int s2d = 0)
{
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= N * K) return;
    int n = idx / K, k = idx % K;
    int s = k / C, c = k % C;
    size_t col = (size_t)c * KHW + s;
    if (s2d) {          // conv1 on 4x4-pixel blocks: k = tap (dy, dx) * 64 + frame * 16 + iy * 4 + ix  <-  W[n][frame][4 dy + iy][4 dx + ix]
        int t = k / 64, kk = k % 64, fr = kk / 16, iy = (kk / 4) % 4, ix = kk % 4;
        col = (size_t)fr * 64 + (size_t)(4 * (t / 2) + iy) * 8 + 4 * (t % 2) + ix;
    }
    float v = n < n_real ? w[(size_t)n * K + col] : 0.f;
    int kc = k / 64, kk = k % 64;
    size_t off = (size_t)kc * N * 64 + (size_t)n * 64 + (((kk / 8) ^ (n % 8)) * 8) + (kk % 8);
    if (f16) reinterpret_cast<__half *>(dst)[off] = __float2half_rn(v);
    else dst[off] = __float2bfloat16_rn(v);
}

// The same layout for a TRANSPOSED weight matrix: row nn = (s, c) of the result (s = spatial tap / position, c = input channel: the
// order the activations are stored in), column k = output channel, taken from the fp32 torch tensor W[k][c][s]:
//   element (nn, k) <- w[k * (KHW * C) + (nn % C) * KHW + nn / C]
__global__ void stm_tile_weights_T_kernel(const float *__restrict__ w, __nv_bfloat16 *__restrict__ dst, int rows, int K, int KHW, int C)
{
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= rows * K) return;
    int nn = idx / K, k = idx % K;
    float v = w[(size_t)k * rows + (size_t)(nn % C) * KHW + nn / C];
    int kc = k / 64, kk = k % 64;
    size_t off = (size_t)kc * rows * 64 + (size_t)nn * 64 + (((kk / 8) ^ (nn % 8)) * 8) + (kk % 8);
    dst[off] = __float2bfloat16_rn(v);
}

struct CnnState {
    int A = 0;
    __nv_bfloat16 *w[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    __nv_bfloat16 *w_c1 = nullptr;                     // conv1 weights in the block (space-to-depth) order of the window kernel, fp16
    float *b[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    __nv_bfloat16 *act1 = nullptr, *act2 = nullptr, *act3 = nullptr, *feat = nullptr;
    long long cap_images = 0;
    float *head_tmp = nullptr;                         // 32 x 512 fp32 staging of the head weights (set_weights)
    // ---- update path (stm_cnn_forward_train / stm_cnn_backward) ----
    __nv_bfloat16 *wT[3] = {nullptr, nullptr, nullptr};   // conv2, conv3, fc weights TRANSPOSED and tiled: the B operands of the data-gradient GEMMs
    float *wh_f32 = nullptr;                           // (A + 1) x 512 head weights in fp32 (rows: action_net, then value_net)
    float *zeros = nullptr;                            // 3 136 zeros: the "bias" of the gradient GEMMs
    __nv_bfloat16 *t_act1 = nullptr, *t_act2 = nullptr, *t_act3 = nullptr, *t_feat = nullptr;   // activations kept from forward_train
    long long t_cap = 0, t_n = 0;
    uint8_t *bwd = nullptr; size_t bwd_bytes = 0;      // backward workspace (gradients of activations, transposed operands, split-K partials)
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

template <int kConv>
int launch_window(stm_ctx *ctx, const __nv_bfloat16 *in, const __nv_bfloat16 *w, const float *b, __nv_bfloat16 *out,
                  long long n_images, const WinGeom &g, cudaStream_t st)
{
    // conv2 (64 KB of resident weights) fits two blocks per SM with a short window ring; conv3 (72 KB) does not
    constexpr int kStages = kConv == 2 ? STM_W2_STAGES : STM_W3_STAGES, kPerSm = kConv == 2 ? STM_W2_PER_SM : STM_W3_PER_SM;
    static bool attr_done = false;
    if (!attr_done) {
        cudaFuncSetAttribute(stm_cnn_window_kernel<kStages, kConv>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 2048);
        cudaFuncSetAttribute(stm_cnn_window_kernel<kStages, kConv>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        attr_done = true;
    }
    const long long M = n_images * g.V;
    if (M >= (1LL << 31) - kTileM) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn: too many rows for one pass");
    for (int q = 0; q < g.V; q++)
        if (((q * g.recip_vw) >> 16) != q / g.VW) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn: reciprocal of VW is not exact");
    if (g.nchunks != win_taps<kConv>()) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn: tap count");
    for (int c = 0; c < g.nchunks; c++)
        if (g.off[c] != win_off<kConv>(c) || g.off[c] + kTileM > win_rows<kConv>()) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn: tap offsets");
    const size_t smem = (size_t)kStages * win_rows<kConv>() * 128 + (size_t)g.nchunks * 64 * kChunkK * 2 + 1024;
    const long long tiles = (M + kTileM - 1) / kTileM;
    const long long grid = tiles < 148 * kPerSm ? tiles : 148 * kPerSm;
    // the activation matrix [M rows][64 bf16] as a 2-D tensor map; box = one window (64 x 160), 128-byte swizzle
    CUtensorMap map;
    {
        typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                     const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        static EncodeFn encode = nullptr;
        if (!encode) {
            cudaDriverEntryPointQueryResult q;
            void *fn = nullptr;
            if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn)
                return stm_internal_fail(ctx, STM_ERR_CUDA, "stm_cnn: cuTensorMapEncodeTiled is not available from this driver");
            encode = reinterpret_cast<EncodeFn>(fn);
        }
        const cuuint64_t dims[2] = { 64, (cuuint64_t)M }, strides[1] = { 128 };
        const cuuint32_t box[2] = { 64, (cuuint32_t)win_rows<kConv>() }, estr[2] = { 1, 1 };
        if (encode(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16 *>(in), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return stm_internal_fail(ctx, STM_ERR_CUDA, "stm_cnn: cuTensorMapEncodeTiled failed");
    }
    stm_cnn_window_kernel<kStages, kConv><<<(unsigned)grid, 288, smem, st>>>(map, w, b, out, (uint32_t)M, g);
    stm_internal_count_launch(ctx);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? STM_OK : stm_internal_fail(ctx, STM_ERR_CUDA, cudaGetErrorString(e));
}

int launch_conv1_window(stm_ctx *ctx, const uint8_t *obs, const __nv_bfloat16 *w, const float *b, __nv_bfloat16 *out,
                        long long n_images, int debug, cudaStream_t st)
{
    constexpr int kStages = STM_C1_STAGES;
    static bool attr_done = false;
    if (!attr_done) {
        cudaFuncSetAttribute(stm_cnn_conv1_window_kernel<kStages>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 2048);
        cudaFuncSetAttribute(stm_cnn_conv1_window_kernel<kStages>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        attr_done = true;
    }
    const long long M = n_images * 441;
    if (M >= (1LL << 31) - kTileM) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn: too many rows for one pass");
    const size_t smem = (size_t)kStages * kWinBytes + 4 * 32 * kChunkK * 2 + 1024;
    const long long tiles = (M + kTileM - 1) / kTileM;
    const long long grid = tiles < 148 * STM_C1_PER_SM ? tiles : 148 * STM_C1_PER_SM;
    stm_cnn_conv1_window_kernel<kStages><<<(unsigned)grid, kC1Threads, smem, st>>>(obs, w, b, out, (uint32_t)M, debug);
    stm_internal_count_launch(ctx);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? STM_OK : stm_internal_fail(ctx, STM_ERR_CUDA, cudaGetErrorString(e));
}

void free_state(CnnState *s)
{
    if (!s) return;
    for (int i = 0; i < 5; i++) { cudaFree(s->w[i]); cudaFree(s->b[i]); }
    for (int i = 0; i < 3; i++) cudaFree(s->wT[i]);
    cudaFree(s->w_c1); cudaFree(s->head_tmp); cudaFree(s->wh_f32); cudaFree(s->zeros);
    cudaFree(s->act1); cudaFree(s->act2); cudaFree(s->act3); cudaFree(s->feat);
    cudaFree(s->t_act1); cudaFree(s->t_act2); cudaFree(s->t_act3); cudaFree(s->t_feat); cudaFree(s->bwd);
    delete s;
}


// =========================================================================================================
// Backward pass (update path).  Every contraction is the same tcgen05 GEMM as the forward's fully connected layer,
//     D[M, N] = A[M, K] * B[N, K]^T,   A row-major bf16 streamed by the producers, B pre-tiled, fp32 accumulation in TMEM,
// fed with explicitly laid-out operands (the update batch is 544 - 7 000 images: the extra passes over HBM are a few
// hundred MB, the GEMMs stay on the tensor cores):
//   data gradients    dX_col[rows, taps*C] = dY[rows, Cout] * W^T          (B = the transposed, tiled weights made by set_weights)
//                     followed by a col2im gather (+ the ReLU mask of the layer below)
//   weight gradients  dW^T[taps*C, Cout]  = col^T[taps*C, rows] * dY^T      (K = rows = batch x positions: split-K over CTA columns,
//                     fp32 partials reduced and permuted into torch's (Cout, C, kh, kw) order by the finalize kernel)
// The heads (A + 1 outputs) are too narrow for a tile and run on the CUDA cores.
// =========================================================================================================
constexpr long long kMaxTrainImages = 32768;

struct SrcPlain {                     // (r, k) of a row-major [rows, ld] bf16 matrix
    const __nv_bfloat16 *p; int ld;
    __device__ __forceinline__ __nv_bfloat16 at(long long r, int k) const { return p[(size_t)r * ld + k]; }
};
struct SrcIm2colNhwc {                // (r = (img, oy, ox), k = (ky, kx, c)) of a bf16 NHWC activation
    const __nv_bfloat16 *p; int IH, IW, C, KW, S, OW, P;
    __device__ __forceinline__ __nv_bfloat16 at(long long r, int k) const
    {
        const long long img = r / P; const int pos = (int)(r - img * P), oy = pos / OW, ox = pos - oy * OW;
        const int t = k / C, c = k - t * C, ky = t / KW, kx = t - ky * KW;
        return p[(((size_t)img * IH + oy * S + ky) * IW + ox * S + kx) * C + c];
    }
};
struct SrcIm2colU8 {                  // conv1: (r = (img, oy, ox), k = (c, ky, kx)) of the u8 NCHW observation (pixel values are exact in bf16)
    const uint8_t *p;
    __device__ __forceinline__ __nv_bfloat16 at(long long r, int k) const
    {
        const long long img = r / 400; const int pos = (int)(r - img * 400), oy = pos / 20, ox = pos - oy * 20;
        const int c = k >> 6, ky = (k >> 3) & 7, kx = k & 7;
        return __float2bfloat16_rn((float)p[(((size_t)img * 4 + c) * 84 + oy * 4 + ky) * 84 + ox * 4 + kx]);
    }
};
struct DstRowMajor {                  // dst[k][r], row length rows_pad
    __nv_bfloat16 *p; long long rows_pad;
    __device__ __forceinline__ void put(int k, long long r, __nv_bfloat16 v) const { p[(size_t)k * rows_pad + r] = v; }
};
struct DstTiled {                     // dst = B operand [N rows = k][K = r] in the tiled, swizzled layout (see stm_tile_weights_kernel)
    __nv_bfloat16 *p; int N;
    __device__ __forceinline__ void put(int k, long long r, __nv_bfloat16 v) const
    {
        const long long kc = r >> 6; const int kk = (int)(r & 63);
        p[(size_t)kc * N * 64 + (size_t)k * 64 + (((kk >> 3) ^ (k & 7)) << 3) + (kk & 7)] = v;
    }
};
// dst(k, r) = src(r, k) for k < Kdim, r < rows_pad (zero beyond `rows`): a 32 x 32 shared-memory transpose, reads coalesced along k,
// writes along r
template <class Src, class Dst>
__global__ void __launch_bounds__(256) stm_transpose_kernel(Src src, Dst dst, int Kdim, long long rows, long long rows_pad)
{
    __shared__ __nv_bfloat16 tile[32][34];
    const long long r0 = (long long)blockIdx.x * 32;
    const int k0 = blockIdx.y * 32, tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    #pragma unroll
    for (int j = 0; j < 4; j++) {
        const long long r = r0 + ty + 8 * j; const int k = k0 + tx;
        tile[ty + 8 * j][tx] = (r < rows && k < Kdim) ? src.at(r, k) : __float2bfloat16_rn(0.f);
    }
    __syncthreads();
    #pragma unroll
    for (int j = 0; j < 4; j++) {
        const int k = k0 + ty + 8 * j; const long long r = r0 + tx;
        if (k < Kdim && r < rows_pad) dst.put(k, r, tile[tx][ty + 8 * j]);
    }
}

// heads backward: dfeat[i][k] = feat[i][k] > 0 ? sum_a dheads[i][a] * Wh[a][k] : 0     (ReLU of the fully connected layer folded in)
__global__ void stm_heads_dgrad_kernel(const float *__restrict__ dheads, const float *__restrict__ wh, const __nv_bfloat16 *__restrict__ feat,
                                       __nv_bfloat16 *__restrict__ dfeat, long long n, int A1)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * 512) return;
    const long long i = idx >> 9; const int k = (int)(idx & 511);
    float acc = 0.f;
    for (int a = 0; a < A1; a++) acc += dheads[i * A1 + a] * __ldg(wh + a * 512 + k);
    dfeat[idx] = __float2bfloat16_rn(__bfloat162float(feat[idx]) > 0.f ? acc : 0.f);
}
// dWh[a][k] = sum_i dheads[i][a] * feat[i][k]; rows a < A go to the action head, row A to the value head.  grid (A1, 4), block 128.
__global__ void stm_heads_wgrad_kernel(const float *__restrict__ dheads, const __nv_bfloat16 *__restrict__ feat, float *__restrict__ g_action_w,
                                       float *__restrict__ g_value_w, float *__restrict__ g_action_b, float *__restrict__ g_value_b,
                                       long long n, int A1)
{
    const int a = blockIdx.x, k = blockIdx.y * 128 + threadIdx.x;
    float acc = 0.f, bsum = 0.f;
    for (long long i = 0; i < n; i++) {
        const float d = dheads[i * A1 + a];
        acc += d * __bfloat162float(feat[i * 512 + k]);
        bsum += d;
    }
    if (a < A1 - 1) g_action_w[a * 512 + k] = acc; else g_value_w[k] = acc;
    if (k == 0) { if (a < A1 - 1) g_action_b[a] = bsum; else g_value_b[0] = bsum; }
}
// out[c] += sum_r x[r][c]  (out zeroed by the caller); cols is 32, 64 or 512
__global__ void stm_colsum_kernel(const __nv_bfloat16 *__restrict__ x, float *__restrict__ out, long long rows, int cols)
{
    const int per = 256 / (cols < 256 ? cols : 256);                 // row lanes per block (1 when cols >= 256)
    for (int c = threadIdx.x % (cols < 256 ? cols : 256); c < cols; c += 256) {
        const int lane = cols < 256 ? threadIdx.x / cols : 0;
        float acc = 0.f;
        for (long long r = (long long)blockIdx.x * per + lane; r < rows; r += (long long)gridDim.x * per) acc += __bfloat162float(x[(size_t)r * cols + c]);
        atomicAdd(out + c, acc);
    }
}
// d = act > 0 ? d : 0, eight elements per thread
__global__ void stm_relu_mask_kernel(__nv_bfloat16 *__restrict__ d, const __nv_bfloat16 *__restrict__ act, long long n8)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n8) return;
    uint4 dv = reinterpret_cast<uint4 *>(d)[i];
    const uint4 av = reinterpret_cast<const uint4 *>(act)[i];
    uint32_t *dw = reinterpret_cast<uint32_t *>(&dv); const uint32_t *aw = reinterpret_cast<const uint32_t *>(&av);
    #pragma unroll
    for (int j = 0; j < 4; j++) {                                   // a post-ReLU bf16 is > 0 iff its bits are non-zero and the sign is clear
        const uint32_t lo = aw[j] & 0xFFFFu, hi = aw[j] >> 16;
        const uint32_t keep = ((lo != 0 && !(lo & 0x8000u)) ? 0x0000FFFFu : 0u) | ((hi != 0 && !(hi & 0x8000u)) ? 0xFFFF0000u : 0u);
        dw[j] &= keep;
    }
    reinterpret_cast<uint4 *>(d)[i] = dv;
}
// col2im as a gather: dact[img][iy][ix][c] = (act > 0) * sum over the taps (ky, kx) that reach (iy, ix) of dcol[(img, oy, ox)][(ky, kx, c)]
__global__ void stm_col2im_kernel(const __nv_bfloat16 *__restrict__ dcol, const __nv_bfloat16 *__restrict__ act, __nv_bfloat16 *__restrict__ dact,
                                  long long n, int IH, int IW, int C, int KH, int KW, int S, int OH, int OW)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * IH * IW * C) return;
    const int c = (int)(idx % C); long long t = idx / C;
    const int ix = (int)(t % IW); t /= IW;
    const int iy = (int)(t % IH); const long long img = t / IH;
    float acc = 0.f;
    if (__bfloat162float(act[idx]) > 0.f) {
        const int Kc = KH * KW * C;
        for (int ky = iy % S; ky < KH; ky += S) {
            const int oy = (iy - ky) / S;
            if (iy - ky < 0 || oy >= OH) continue;
            for (int kx = ix % S; kx < KW; kx += S) {
                const int ox = (ix - kx) / S;
                if (ix - kx < 0 || ox >= OW) continue;
                acc += __bfloat162float(dcol[((size_t)img * OH * OW + oy * OW + ox) * Kc + (ky * KW + kx) * C + c]);
            }
        }
    }
    dact[idx] = __float2bfloat16_rn(acc);
}
// grad[nn][c * KHW + t] = scale * sum_s part[s][m = t * C + c][nn]     (fp32 split-K partials -> torch's (Cout, C, kh, kw) order)
__global__ void stm_wgrad_finalize_kernel(const float *__restrict__ part, float *__restrict__ grad, int S, int M, int N, int KHW, int C, float scale)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= M * N) return;
    const int m = idx / N, nn = idx - m * N;
    float acc = 0.f;
    for (int s2 = 0; s2 < S; s2++) acc += part[(size_t)s2 * M * N + idx];
    const int t = m / C, c = m - t * C;
    grad[(size_t)nn * M + (size_t)c * KHW + t] = acc * scale;
}

// D[M, N range] = A[M, K] (row-major bf16) * B[N_total, K]^T (tiled bf16).  splits > 1: split-K, fp32 partials out[split][M][N_total].
template <int N>
int launch_rm(stm_ctx *ctx, const __nv_bfloat16 *A, const __nv_bfloat16 *Bt, const float *zeros, void *out, long long M, int K,
              int n_off, int N_total, int out_f32, int splits, cudaStream_t st)
{
    static bool attr_done = false;
    constexpr int kStages = 4, kProdWarps = 4;
    const size_t smem = (size_t)kStages * kABytes + (size_t)kStages * N * kChunkK * 2 + 1024;
    if (!attr_done) {
        cudaFuncSetAttribute(stm_cnn_gemm_kernel<N, false, kStages, kProdWarps>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 2048);
        cudaFuncSetAttribute(stm_cnn_gemm_kernel<N, false, kStages, kProdWarps>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        attr_done = true;
    }
    if (K % kChunkK || M >= (1LL << 31) - kTileM) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn_backward: GEMM shape");
    const int total = K / kChunkK;
    int split_chunks = 0, S = 1;
    if (splits > 1) { split_chunks = (total + splits - 1) / splits; S = (total + split_chunks - 1) / split_chunks; }
    LayerGeom g = { MODE_NHWC, 1, 1, 1, 1, K, K, 1, K, N_total, 0, out_f32, 1.0f, 65536, 0, 0, split_chunks };
    const long long tiles = (M + kTileM - 1) / kTileM;
    const int per_sm = smem <= 110 * 1024 ? 2 : 1;
    long long gx = tiles < 148LL * per_sm ? tiles : 148LL * per_sm;
    stm_cnn_gemm_kernel<N, false, kStages, kProdWarps><<<dim3((unsigned)gx, (unsigned)S), cta_threads(kProdWarps), smem, st>>>(
        A, Bt, zeros, out, M, g, n_off, N_total);
    stm_internal_count_launch(ctx);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? S : stm_internal_fail(ctx, STM_ERR_CUDA, cudaGetErrorString(e));       // > 0: the number of partials written
}
// all N_total columns in pieces of 256 / 64 / 32
int gemm_rm(stm_ctx *ctx, const __nv_bfloat16 *A, const __nv_bfloat16 *Bt, const float *zeros, void *out, long long M, int K, int N_total,
            int out_f32, int splits, cudaStream_t st)
{
    int S = 1, n_off = 0;
    while (n_off < N_total) {
        int rc;
        if (N_total - n_off >= 256) { rc = launch_rm<256>(ctx, A, Bt, zeros, out, M, K, n_off, N_total, out_f32, splits, st); n_off += 256; }
        else if (N_total - n_off >= 64) { rc = launch_rm<64>(ctx, A, Bt, zeros, out, M, K, n_off, N_total, out_f32, splits, st); n_off += 64; }
        else { rc = launch_rm<32>(ctx, A, Bt, zeros, out, M, K, n_off, N_total, out_f32, splits, st); n_off += 32; }
        if (rc < 0) return rc;
        S = rc;
    }
    return S;
}

template <class Src, class Dst>
void transpose_to(stm_ctx *ctx, Src src, Dst dst, int Kdim, long long rows, long long rows_pad, cudaStream_t st)
{
    dim3 grid((unsigned)((rows_pad + 31) / 32), (unsigned)((Kdim + 31) / 32));
    stm_transpose_kernel<<<grid, 256, 0, st>>>(src, dst, Kdim, rows, rows_pad);
    stm_internal_count_launch(ctx);
}

inline long long pad64(long long v) { return (v + 63) / 64 * 64; }

int backward_impl(stm_ctx *ctx, CnnState *s, const uint8_t *obs, long long n, const float *dheads, float *const *grads, cudaStream_t st)
{
    const int A1 = s->A + 1;
    const long long rows1 = n * 400, rows2 = n * 81, rows3 = n * 49;
    const long long r1p = pad64(rows1), r2p = pad64(rows2), r3p = pad64(rows3), np = pad64(n);
    // workspace carve-up (bytes, 256-byte aligned pieces)
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    const size_t o_dfeat = take((size_t)n * 512 * 2), o_dfeatT = take((size_t)512 * np * 2), o_act3T = take((size_t)3136 * np * 2);
    const size_t o_dact3 = take((size_t)n * 3136 * 2), o_dY3T = take((size_t)64 * r3p * 2), o_col3T = take((size_t)576 * r3p * 2);
    const size_t o_dcol3 = take((size_t)rows3 * 576 * 2), o_dact2 = take((size_t)rows2 * 64 * 2), o_dY2T = take((size_t)64 * r2p * 2);
    const size_t o_col2T = take((size_t)512 * r2p * 2), o_dcol2 = take((size_t)rows2 * 512 * 2), o_dact1 = take((size_t)rows1 * 32 * 2);
    const size_t o_dY1T = take((size_t)32 * r1p * 2), o_col1T = take((size_t)256 * r1p * 2);
    const size_t o_part = take((size_t)8 * 3136 * 512 * 4);
    if (off > s->bwd_bytes) {
        cudaFree(s->bwd); s->bwd = nullptr; s->bwd_bytes = 0;
        if (cudaMalloc(&s->bwd, off) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_NOMEM, "stm_cnn_backward: workspace");
        s->bwd_bytes = off;
    }
    auto B16 = [&](size_t o) { return reinterpret_cast<__nv_bfloat16 *>(s->bwd + o); };
    float *part = reinterpret_cast<float *>(s->bwd + o_part);
    float *g_c1w = grads[0], *g_c1b = grads[1], *g_c2w = grads[2], *g_c2b = grads[3], *g_c3w = grads[4], *g_c3b = grads[5],
          *g_fcw = grads[6], *g_fcb = grads[7], *g_aw = grads[8], *g_ab = grads[9], *g_vw = grads[10], *g_vb = grads[11];
    int rc, S;
    auto colsum = [&](const __nv_bfloat16 *x, float *out, long long rows, int cols) {
        cudaMemsetAsync(out, 0, (size_t)cols * 4, st);
        stm_colsum_kernel<<<296, 256, 0, st>>>(x, out, rows, cols);
        stm_internal_count_launch(ctx);
    };
    auto finalize = [&](float *grad, int S2, int M, int N, int KHW, int C, float scale) {
        stm_wgrad_finalize_kernel<<<(M * N + 255) / 256, 256, 0, st>>>(part, grad, S2, M, N, KHW, C, scale);
        stm_internal_count_launch(ctx);
    };
    // ---- heads (CUDA cores) + the ReLU of the fully connected layer ----
    stm_heads_dgrad_kernel<<<(unsigned)((n * 512 + 255) / 256), 256, 0, st>>>(dheads, s->wh_f32, s->t_feat, B16(o_dfeat), n, A1);
    stm_heads_wgrad_kernel<<<dim3(A1, 4), 128, 0, st>>>(dheads, s->t_feat, g_aw, g_vw, g_ab, g_vb, n, A1);
    stm_internal_count_launch(ctx); stm_internal_count_launch(ctx);
    colsum(B16(o_dfeat), g_fcb, n, 512);
    // ---- fully connected: data gradient [n, 3136] = dfeat [n, 512] * Wfc, then conv3's ReLU ----
    if ((rc = gemm_rm(ctx, B16(o_dfeat), s->wT[2], s->zeros, B16(o_dact3), n, 512, 3136, 0, 1, st)) < 0) return rc;
    stm_relu_mask_kernel<<<(unsigned)((n * 3136 / 8 + 255) / 256), 256, 0, st>>>(B16(o_dact3), s->t_act3, n * 3136 / 8);
    stm_internal_count_launch(ctx);
    // weight gradient: [3136, 512] = act3^T [3136, n] * dfeat^T
    transpose_to(ctx, SrcPlain{ s->t_act3, 3136 }, DstRowMajor{ B16(o_act3T), np }, 3136, n, np, st);
    transpose_to(ctx, SrcPlain{ B16(o_dfeat), 512 }, DstTiled{ B16(o_dfeatT), 512 }, 512, n, np, st);
    if ((S = gemm_rm(ctx, B16(o_act3T), B16(o_dfeatT), s->zeros, part, 3136, (int)np, 512, 1, 8, st)) < 0) return S;
    finalize(g_fcw, S, 3136, 512, 49, 64, 1.0f);
    // ---- conv3 ----
    colsum(B16(o_dact3), g_c3b, rows3, 64);
    transpose_to(ctx, SrcIm2colNhwc{ s->t_act2, 9, 9, 64, 3, 1, 7, 49 }, DstRowMajor{ B16(o_col3T), r3p }, 576, rows3, r3p, st);
    transpose_to(ctx, SrcPlain{ B16(o_dact3), 64 }, DstTiled{ B16(o_dY3T), 64 }, 64, rows3, r3p, st);
    if ((S = gemm_rm(ctx, B16(o_col3T), B16(o_dY3T), s->zeros, part, 576, (int)r3p, 64, 1, 48, st)) < 0) return S;
    finalize(g_c3w, S, 576, 64, 9, 64, 1.0f);
    if ((rc = gemm_rm(ctx, B16(o_dact3), s->wT[1], s->zeros, B16(o_dcol3), rows3, 64, 576, 0, 1, st)) < 0) return rc;
    stm_col2im_kernel<<<(unsigned)((rows2 * 64 + 255) / 256), 256, 0, st>>>(B16(o_dcol3), s->t_act2, B16(o_dact2), n, 9, 9, 64, 3, 3, 1, 7, 7);
    stm_internal_count_launch(ctx);
    // ---- conv2 ----
    colsum(B16(o_dact2), g_c2b, rows2, 64);
    transpose_to(ctx, SrcIm2colNhwc{ s->t_act1, 20, 20, 32, 4, 2, 9, 81 }, DstRowMajor{ B16(o_col2T), r2p }, 512, rows2, r2p, st);
    transpose_to(ctx, SrcPlain{ B16(o_dact2), 64 }, DstTiled{ B16(o_dY2T), 64 }, 64, rows2, r2p, st);
    if ((S = gemm_rm(ctx, B16(o_col2T), B16(o_dY2T), s->zeros, part, 512, (int)r2p, 64, 1, 48, st)) < 0) return S;
    finalize(g_c2w, S, 512, 64, 16, 32, 1.0f);
    if ((rc = gemm_rm(ctx, B16(o_dact2), s->wT[0], s->zeros, B16(o_dcol2), rows2, 64, 512, 0, 1, st)) < 0) return rc;
    stm_col2im_kernel<<<(unsigned)((rows1 * 32 + 255) / 256), 256, 0, st>>>(B16(o_dcol2), s->t_act1, B16(o_dact1), n, 20, 20, 32, 4, 4, 2, 9, 9);
    stm_internal_count_launch(ctx);
    // ---- conv1 (no data gradient: the input is the observation) ----
    colsum(B16(o_dact1), g_c1b, rows1, 32);
    transpose_to(ctx, SrcIm2colU8{ obs }, DstRowMajor{ B16(o_col1T), r1p }, 256, rows1, r1p, st);
    transpose_to(ctx, SrcPlain{ B16(o_dact1), 32 }, DstTiled{ B16(o_dY1T), 32 }, 32, rows1, r1p, st);
    if ((S = gemm_rm(ctx, B16(o_col1T), B16(o_dY1T), s->zeros, part, 256, (int)r1p, 32, 1, 96, st)) < 0) return S;
    finalize(g_c1w, S, 256, 32, 1, 256, 1.0f / 255.0f);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? STM_OK : stm_internal_fail(ctx, STM_ERR_CUDA, cudaGetErrorString(e));
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
    if (!s->w_c1 && cudaMalloc(&s->w_c1, 32 * 256 * 2) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_NOMEM, "stm_cnn_set_weights: cudaMalloc");
    stm_tile_weights_kernel<<<(32 * 256 + 255) / 256, 256, 0, st>>>(conv1_w, s->w_c1, 32, 32, 256, 1, 256, 1, 1);
    stm_internal_count_launch(ctx);
    // heads: rows 0..A-1 = action_net, row A = value_net, zero padded to 32 rows
    if (!s->w[4] && cudaMalloc(&s->w[4], 32 * 512 * 2) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_NOMEM, "cudaMalloc");
    if (!s->b[4] && cudaMalloc(&s->b[4], 32 * 4) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_NOMEM, "cudaMalloc");
    if (!s->head_tmp && cudaMalloc(&s->head_tmp, 32 * 512 * 4) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_NOMEM, "cudaMalloc");
    float *tmp = s->head_tmp;                                        // allocated once: every later call is allocation-free
    cudaMemsetAsync(tmp, 0, 32 * 512 * 4, st);
    cudaMemsetAsync(s->b[4], 0, 32 * 4, st);
    cudaMemcpyAsync(tmp, action_w, (size_t)n_actions * 512 * 4, cudaMemcpyDeviceToDevice, st);
    cudaMemcpyAsync(tmp + (size_t)n_actions * 512, value_w, 512 * 4, cudaMemcpyDeviceToDevice, st);
    cudaMemcpyAsync(s->b[4], action_b, (size_t)n_actions * 4, cudaMemcpyDeviceToDevice, st);
    cudaMemcpyAsync(s->b[4] + n_actions, value_b, 4, cudaMemcpyDeviceToDevice, st);
    stm_tile_weights_kernel<<<(32 * 512 + 255) / 256, 256, 0, st>>>(tmp, s->w[4], n_actions + 1, 32, 512, 1, 512, 0);
    stm_internal_count_launch(ctx);
    // update path: transposed, tiled copies of conv2 / conv3 / fc (B operands of the data-gradient GEMMs) and the fp32 heads
    {
        struct SpecT { const float *w; int rows, K, KHW, C; } st3[3] = { { conv2_w, 512, 64, 16, 32 }, { conv3_w, 576, 64, 9, 64 }, { fc_w, 3136, 512, 49, 64 } };
        for (int i = 0; i < 3; i++) {
            if (!s->wT[i] && cudaMalloc(&s->wT[i], (size_t)st3[i].rows * st3[i].K * 2) != cudaSuccess)
                return stm_internal_fail(ctx, STM_ERR_NOMEM, "stm_cnn_set_weights: cudaMalloc");
            const int total = st3[i].rows * st3[i].K;
            stm_tile_weights_T_kernel<<<(total + 255) / 256, 256, 0, st>>>(st3[i].w, s->wT[i], st3[i].rows, st3[i].K, st3[i].KHW, st3[i].C);
            stm_internal_count_launch(ctx);
        }
        if (!s->wh_f32 && cudaMalloc(&s->wh_f32, 19 * 512 * 4) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_NOMEM, "cudaMalloc");
        cudaMemcpyAsync(s->wh_f32, tmp, (size_t)(n_actions + 1) * 512 * 4, cudaMemcpyDeviceToDevice, st);
        if (!s->zeros) {
            if (cudaMalloc(&s->zeros, 3136 * 4) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_NOMEM, "cudaMalloc");
            cudaMemsetAsync(s->zeros, 0, 3136 * 4, st);
        }
    }
    cudaStreamSynchronize(st);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? STM_OK : stm_internal_fail(ctx, STM_ERR_CUDA, cudaGetErrorString(e));
}

int stm_debug_shift_probe(stm_ctx *ctx, const void *A, const void *B, float *out, int rows, int shift, int base_off, void *stream)
{
    if (!ctx || !A || !B || !out || rows < 128 || rows > 250 || shift < 0 || shift + 128 > rows)
        return stm_internal_fail(ctx, STM_ERR_ARG, "stm_debug_shift_probe: bad argument");
    cudaFuncSetAttribute(stm_shift_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024);
    stm_shift_probe_kernel<<<1, 128, 41 * 1024 + 1024, (cudaStream_t)stream>>>(
        reinterpret_cast<const __nv_bfloat16 *>(A), reinterpret_cast<const __nv_bfloat16 *>(B), out, rows, shift, base_off);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? STM_OK : stm_internal_fail(ctx, STM_ERR_CUDA, cudaGetErrorString(e));
}

// one pass of the network over c images (c <= 32 768) with the given activation buffers
static int forward_chunk(stm_ctx *ctx, CnnState *s, const uint8_t *obs, long long c, __nv_bfloat16 *act1, __nv_bfloat16 *act2,
                         __nv_bfloat16 *act3, __nv_bfloat16 *feat, float *heads, int dbg, cudaStream_t st)
{
    const int A1 = s->A + 1;
    int rc;
    LayerGeom g1 = { MODE_CONV1, 400, 20, 84, 84, 4, 0, 4, 256, 32, 1, 0, 1.0f / 255.0f, (65536 + 20 - 1) / 20, 1, dbg };
    if (!(dbg & 512)) {
        if (!(dbg & 256)) rc = launch_conv1_window(ctx, obs, s->w_c1, s->b[0], act1, c, dbg, st);
        else rc = launch_layer<32, true, 4, 8>(ctx, obs, s->w[0], s->b[0], act1, c * 400, g1, 0, 32, st);
        if (rc) return rc;
    }
    if (dbg & 8) return STM_OK;                                   // timing experiments: conv1 alone
    if (!(dbg & 256)) {
        // conv2: rows of two pixels on the 10 x 20 virtual grid; tap (ky, kx in {0, 2}) = input row + 10 ky + kx / 2
        WinGeom w2 = { 200, 20, 9, 9, (65536 + 20 - 1) / 20, 8, { 0, 1, 10, 11, 20, 21, 30, 31, 0 }, 1, dbg };
        if ((rc = launch_window<2>(ctx, act1, s->w[1], s->b[1], act2, c, w2, st))) return rc;
        // conv3: rows of one pixel on the 9 x 9 input grid; tap (ky, kx) = input row + 9 ky + kx
        WinGeom w3 = { 81, 9, 7, 7, (65536 + 9 - 1) / 9, 9, { 0, 1, 2, 9, 10, 11, 18, 19, 20 }, 1, dbg };
        if ((rc = launch_window<3>(ctx, act2, s->w[2], s->b[2], act3, c, w3, st))) return rc;
    } else {                                                      // the gather kernels (kept for A/B timing)
        LayerGeom g2 = { MODE_NHWC, 81, 9, 20, 20, 32, 4 * 32, 2, 512, 64, 1, 0, 1.0f, (65536 + 9 - 1) / 9, 0, dbg };
        if ((rc = launch_layer<64, true, 8, 4>(ctx, act1, s->w[1], s->b[1], act2, c * 81, g2, 0, 64, st))) return rc;
        LayerGeom g3 = { MODE_NHWC, 49, 7, 9, 9, 64, 3 * 64, 1, 576, 64, 1, 0, 1.0f, (65536 + 7 - 1) / 7, 0, dbg };
        if ((rc = launch_layer<64, true, 8, 4>(ctx, act2, s->w[2], s->b[2], act3, c * 49, g3, 0, 64, st))) return rc;
    }
    if (dbg & 512) return STM_OK;                                 // timing experiments: conv2 + conv3 alone (with bit 9: no conv1 either)
    LayerGeom g4 = { MODE_NHWC, 1, 1, 1, 1, 3136, 3136, 1, 3136, 512, 1, 0, 1.0f, 65536, 0, dbg };
    for (int half = 0; half < 2; half++)
        if ((rc = launch_layer<256, false, STM_FC_STAGES, 4>(ctx, act3, s->w[3], s->b[3], feat, c, g4, half * 256, 512, st))) return rc;
    LayerGeom g5 = { MODE_NHWC, 1, 1, 1, 1, 512, 512, 1, 512, A1, 0, 1, 1.0f, 65536, 0, dbg };
    return launch_layer<32, true, 4, 4>(ctx, feat, s->w[4], s->b[4], heads, c, g5, 0, 32, st);
}

// timing experiments inside K4 (tools/cnn_probe.py) exist only in builds made with -DSTM_CNN_PROBES; the product
// library never reads the environment on this path
static int cnn_debug_bits()
{
#ifdef STM_CNN_PROBES
    const char *e = getenv("STM_CNN_DEBUG");
    return e ? atoi(e) : 0;
#else
    return 0;
#endif
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
    const int A1 = s->A + 1, dbg = cnn_debug_bits();
    for (long long i0 = 0; i0 < n; i0 += chunk) {
        long long c = n - i0 < chunk ? n - i0 : chunk;
        int rc = forward_chunk(ctx, s, obs + (size_t)i0 * 4 * 84 * 84, c, s->act1, s->act2, s->act3, s->feat, heads + (size_t)i0 * A1, dbg, st);
        if (rc) return rc;
    }
    return STM_OK;
}

// ---------------------------------------------------------------------------------------------------------
// Update path: the forward pass that KEEPS its activations, and the backward pass (policies/actor_critic_ts.py:139
// runs extract_features on the cached leaves with autograd; SB3 1.4.0 PPO.train backpropagates through it).
// ---------------------------------------------------------------------------------------------------------
int stm_cnn_forward_train(stm_ctx *ctx, const uint8_t *obs, long long n, float *heads, void *stream)
{
    if (!ctx || !obs || !heads || n < 1) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn_forward_train: bad argument");
    if (n > kMaxTrainImages) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn_forward_train: more than 32 768 images in one minibatch");
    CnnState *s = reinterpret_cast<CnnState *>(*stm_internal_cnn_slot(ctx));
    if (!s || !s->w[0]) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn_forward_train: call stm_cnn_set_weights first");
    if (cudaSetDevice(stm_internal_device(ctx)) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_CUDA, "cudaSetDevice");
    cudaStream_t st = (cudaStream_t)stream;
    if (n > s->t_cap) {
        cudaFree(s->t_act1); cudaFree(s->t_act2); cudaFree(s->t_act3); cudaFree(s->t_feat);
        s->t_act1 = s->t_act2 = s->t_act3 = s->t_feat = nullptr; s->t_cap = 0;
        const long long cap = (n + 63) / 64 * 64;
        if (cudaMalloc(&s->t_act1, (size_t)cap * 400 * 32 * 2) != cudaSuccess || cudaMalloc(&s->t_act2, (size_t)cap * 81 * 64 * 2) != cudaSuccess ||
            cudaMalloc(&s->t_act3, (size_t)cap * 49 * 64 * 2) != cudaSuccess || cudaMalloc(&s->t_feat, (size_t)cap * 512 * 2) != cudaSuccess)
            return stm_internal_fail(ctx, STM_ERR_NOMEM, "stm_cnn_forward_train: activation buffers");
        s->t_cap = cap;
    }
    s->t_n = n;
    return forward_chunk(ctx, s, obs, n, s->t_act1, s->t_act2, s->t_act3, s->t_feat, heads, 0, st);
}

int stm_cnn_backward(stm_ctx *ctx, const uint8_t *obs, long long n, const float *dheads, float *g_conv1_w, float *g_conv1_b,
                     float *g_conv2_w, float *g_conv2_b, float *g_conv3_w, float *g_conv3_b, float *g_fc_w, float *g_fc_b,
                     float *g_action_w, float *g_action_b, float *g_value_w, float *g_value_b, void *stream)
{
    if (!ctx || !obs || !dheads || !g_conv1_w || !g_conv1_b || !g_conv2_w || !g_conv2_b || !g_conv3_w || !g_conv3_b || !g_fc_w ||
        !g_fc_b || !g_action_w || !g_action_b || !g_value_w || !g_value_b)
        return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn_backward: bad argument");
    CnnState *s = reinterpret_cast<CnnState *>(*stm_internal_cnn_slot(ctx));
    if (!s || !s->wT[0] || s->t_n != n || n < 1)
        return stm_internal_fail(ctx, STM_ERR_ARG, "stm_cnn_backward: must follow stm_cnn_forward_train of the same n images");
    if (cudaSetDevice(stm_internal_device(ctx)) != cudaSuccess) return stm_internal_fail(ctx, STM_ERR_CUDA, "cudaSetDevice");
    float *grads[12] = { g_conv1_w, g_conv1_b, g_conv2_w, g_conv2_b, g_conv3_w, g_conv3_b, g_fc_w, g_fc_b, g_action_w, g_action_b,
                         g_value_w, g_value_b };
    return backward_impl(ctx, s, obs, n, dheads, grads, (cudaStream_t)stream);
}

} // extern "C"
