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
// canonical K-major layout (8x8 core matrices, no swizzle), while the previous slice is being
// multiplied.  Weights are pre-tiled once into the same layout, so a slice of W is a linear copy.
//
// Per CTA: 128 threads.  All threads gather; thread 0 issues tcgen05.mma (M = 128, N = layer width,
// K = 16 per instruction, 4 per 64-wide slice) into a TMEM accumulator and commits to an mbarrier;
// two shared-memory stages so that gathering slice i+1 overlaps the MMAs of slice i; the epilogue
// reads TMEM with tcgen05.ld (each warp its 32 lanes), applies scale + bias (+ ReLU) and stores.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <new>
#include "../../include/stm_b200.h"

extern "C" int stm_internal_fail(stm_ctx *ctx, int code, const char *msg);
extern "C" void stm_internal_count_launch(stm_ctx *ctx);
extern "C" int stm_internal_device(const stm_ctx *ctx);
extern "C" void **stm_internal_cnn_slot(stm_ctx *ctx);

namespace {

constexpr int kThreads = 128;
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

// shared-memory matrix descriptor: K-major, no swizzle; LBO = 128 B (next core matrix along K),
// SBO = 1024 B (next 8-row group: 8 core matrices of one 64-wide slice lie between)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr)
{
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)(128 >> 4) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
// instruction descriptor, kind::f16: D = f32, A = B = bf16, both K-major, M = 128
__device__ __forceinline__ uint32_t make_idesc(int N)
{
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
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

// pack 8 u8 (two words) into 8 bf16 (exact: 0..255 are representable)
__device__ __forceinline__ uint4 u8x8_to_bf16x8(uint32_t lo, uint32_t hi)
{
    auto two = [](uint32_t a, uint32_t b) -> uint32_t {
        __nv_bfloat162 v = __floats2bfloat162_rn((float)a, (float)b);
        return *reinterpret_cast<uint32_t *>(&v);
    };
    uint4 r;
    r.x = two(lo & 0xFF, (lo >> 8) & 0xFF);  r.y = two((lo >> 16) & 0xFF, lo >> 24);
    r.z = two(hi & 0xFF, (hi >> 8) & 0xFF);  r.w = two((hi >> 16) & 0xFF, hi >> 24);
    return r;
}

template <int N>
__global__ void __launch_bounds__(kThreads)
stm_cnn_gemm_kernel(const void *__restrict__ in, const __nv_bfloat16 *__restrict__ wt, const float *__restrict__ bias,
                    void *__restrict__ out, long long M_total, LayerGeom g, int n_off, int N_total)
{
    constexpr int kBBytes = N * kChunkK * 2;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *sA[2] = { smem, smem + kABytes };
    uint8_t *sB[2] = { smem + 2 * kABytes, smem + 2 * kABytes + kBBytes };
    __shared__ __align__(8) uint64_t bar[2];
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32(&tmem_base_s)), "r"((uint32_t)(N < 32 ? 32 : N)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0) {
        mbar_init(&bar[0], 1); mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_s;

    // this thread's output row
    const long long m = (long long)blockIdx.x * kTileM + tid;
    const bool row_ok = m < M_total;
    const long long img = row_ok ? m / g.P : 0;
    const int pos = row_ok ? (int)(m % g.P) : 0;
    const int oy = pos / g.OW, ox = pos % g.OW;
    const uint32_t a_row_off = (uint32_t)(tid >> 3) * 1024u + (uint32_t)(tid & 7) * 16u;

    const int nchunks = g.K / kChunkK;
    const uint32_t idesc = make_idesc(N);
    const __nv_bfloat16 *wbase = wt + (size_t)n_off * kChunkK;     // tiles of this N range inside a chunk

    for (int kc = 0; kc < nchunks; kc++) {
        const int st = kc & 1;
        if (kc >= 2) mbar_wait(&bar[st], ((kc >> 1) - 1) & 1);      // MMAs that read this stage are done
        // ---- gather A slice: row `tid`, 64 K elements = 8 pieces of 16 bytes ----
        uint8_t *arow = sA[st] + a_row_off;
        if (g.mode == MODE_CONV1) {
            // slice kc = input channel; piece kg = kernel row kh; 8 u8 along kw
            const uint8_t *src = reinterpret_cast<const uint8_t *>(in) + ((size_t)img * g.C + kc) * g.IH * g.IW
                                 + (size_t)(oy * g.S) * g.IW + ox * g.S;
            #pragma unroll
            for (int kg = 0; kg < 8; kg++) {
                uint4 v = make_uint4(0, 0, 0, 0);
                if (row_ok) {
                    const uint32_t *p = reinterpret_cast<const uint32_t *>(src + kg * g.IW);
                    v = u8x8_to_bf16x8(__ldg(p), __ldg(p + 1));
                }
                *reinterpret_cast<uint4 *>(arow + kg * 128) = v;
            }
        } else {
            const int k0 = kc * kChunkK;
            const int kh = k0 / g.KWC, koff = k0 % g.KWC;
            const __nv_bfloat16 *src = reinterpret_cast<const __nv_bfloat16 *>(in)
                + (((size_t)img * g.IH + (oy * g.S + kh)) * g.IW + ox * g.S) * g.C + koff;
            #pragma unroll
            for (int kg = 0; kg < 8; kg++) {
                uint4 v = make_uint4(0, 0, 0, 0);
                if (row_ok) v = __ldg(reinterpret_cast<const uint4 *>(src) + kg);
                *reinterpret_cast<uint4 *>(arow + kg * 128) = v;
            }
        }
        // ---- copy W slice (pre-tiled: linear) ----
        {
            const uint4 *wsrc = reinterpret_cast<const uint4 *>(wbase + (size_t)kc * N_total * kChunkK);
            uint4 *wdst = reinterpret_cast<uint4 *>(sB[st]);
            #pragma unroll 4
            for (int i = tid; i < kBBytes / 16; i += kThreads) wdst[i] = __ldg(wsrc + i);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the tensor core
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t a0 = smem_u32(sA[st]), b0 = smem_u32(sB[st]);
            #pragma unroll
            for (int j = 0; j < kChunkK / 16; j++)
                umma(tmem_d, make_desc(a0 + j * 256), make_desc(b0 + j * 256), idesc, (kc | j) ? 1u : 0u);
            umma_commit(&bar[st]);
        }
    }
    // wait for the last commit of each stage (a commit covers every earlier MMA)
    {
        const int last = nchunks - 1;
        mbar_wait(&bar[last & 1], (last >> 1) & 1);
        if (nchunks >= 2) { const int prev = last - 1; mbar_wait(&bar[prev & 1], (prev >> 1) & 1); }
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // ---- epilogue: TMEM -> registers -> scale, bias, ReLU -> global ----
    const uint32_t taddr = tmem_d + ((uint32_t)(warp * 32) << 16);
    #pragma unroll 1
    for (int c0 = 0; c0 < N; c0 += 8) {
        uint32_t r[8];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                     : "r"(taddr + (uint32_t)c0));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (row_ok) {
            float v[8];
            #pragma unroll
            for (int j = 0; j < 8; j++) {
                int n = n_off + c0 + j;
                float x = __uint_as_float(r[j]) * g.scale + (n < g.n_real ? bias[n] : 0.f);
                v[j] = g.relu ? fmaxf(x, 0.f) : x;
            }
            if (g.out_f32) {
                float *o = reinterpret_cast<float *>(out) + (size_t)m * g.n_real;
                #pragma unroll
                for (int j = 0; j < 8; j++) if (n_off + c0 + j < g.n_real) o[n_off + c0 + j] = v[j];
            } else {
                uint4 pk;
                __nv_bfloat162 t0 = __floats2bfloat162_rn(v[0], v[1]), t1 = __floats2bfloat162_rn(v[2], v[3]);
                __nv_bfloat162 t2 = __floats2bfloat162_rn(v[4], v[5]), t3 = __floats2bfloat162_rn(v[6], v[7]);
                pk.x = *reinterpret_cast<uint32_t *>(&t0); pk.y = *reinterpret_cast<uint32_t *>(&t1);
                pk.z = *reinterpret_cast<uint32_t *>(&t2); pk.w = *reinterpret_cast<uint32_t *>(&t3);
                *reinterpret_cast<uint4 *>(reinterpret_cast<__nv_bfloat16 *>(out) + (size_t)m * N_total + n_off + c0) = pk;
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_d), "r"((uint32_t)(N < 32 ? 32 : N)) : "memory");
}

// fp32 torch-layout weights (n_real, K) -> bf16, K reordered (k = s*C + c <- column c*KHW + s), zero padded
// to N rows, tiled per 64-wide K slice in the tensor core's canonical K-major layout:
//   element (n, k): slice k/64, byte offset (n/8)*1024 + ((k%64)/8)*128 + (n%8)*16 + (k%8)*2
__global__ void stm_tile_weights_kernel(const float *__restrict__ w, __nv_bfloat16 *__restrict__ dst, int n_real, int N,
                                        int K, int KHW, int C)
{
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= N * K) return;
    int n = idx / K, k = idx % K;
    int s = k / C, c = k % C;
    float v = n < n_real ? w[(size_t)n * K + (size_t)c * KHW + s] : 0.f;
    int kc = k / 64, kk = k % 64;
    size_t off = (size_t)kc * N * 64 + (size_t)(n / 8) * 512 + (kk / 8) * 64 + (n % 8) * 8 + (kk % 8);
    dst[off] = __float2bfloat16_rn(v);
}

struct CnnState {
    int A = 0;
    __nv_bfloat16 *w[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    float *b[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    __nv_bfloat16 *act1 = nullptr, *act2 = nullptr, *act3 = nullptr, *feat = nullptr;
    long long cap_images = 0;
};

template <int N>
int launch_layer(stm_ctx *ctx, const void *in, const __nv_bfloat16 *w, const float *b, void *out, long long M,
                 const LayerGeom &g, int n_off, int N_total, cudaStream_t st)
{
    static bool attr_done = false;
    size_t smem = 2 * kABytes + 2 * (size_t)N * kChunkK * 2 + 1024;
    if (!attr_done) {
        cudaFuncSetAttribute(stm_cnn_gemm_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attr_done = true;
    }
    long long tiles = (M + kTileM - 1) / kTileM;
    stm_cnn_gemm_kernel<N><<<(unsigned)tiles, kThreads, smem, st>>>(in, w, b, out, M, g, n_off, N_total);
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
                                                                    spec[i].K, spec[i].KHW, spec[i].C);
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
    stm_tile_weights_kernel<<<(32 * 512 + 255) / 256, 256, 0, st>>>(tmp, s->w[4], n_actions + 1, 32, 512, 1, 512);
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
    for (long long i0 = 0; i0 < n; i0 += chunk) {
        long long c = n - i0 < chunk ? n - i0 : chunk;
        int rc;
        LayerGeom g1 = { MODE_CONV1, 400, 20, 84, 84, 4, 0, 4, 256, 32, 1, 0, 1.0f / 255.0f };
        if ((rc = launch_layer<32>(ctx, obs + (size_t)i0 * 4 * 84 * 84, s->w[0], s->b[0], s->act1, c * 400, g1, 0, 32, st))) return rc;
        LayerGeom g2 = { MODE_NHWC, 81, 9, 20, 20, 32, 4 * 32, 2, 512, 64, 1, 0, 1.0f };
        if ((rc = launch_layer<64>(ctx, s->act1, s->w[1], s->b[1], s->act2, c * 81, g2, 0, 64, st))) return rc;
        LayerGeom g3 = { MODE_NHWC, 49, 7, 9, 9, 64, 3 * 64, 1, 576, 64, 1, 0, 1.0f };
        if ((rc = launch_layer<64>(ctx, s->act2, s->w[2], s->b[2], s->act3, c * 49, g3, 0, 64, st))) return rc;
        LayerGeom g4 = { MODE_NHWC, 1, 1, 1, 1, 3136, 3136, 1, 3136, 512, 1, 0, 1.0f };
        for (int half = 0; half < 2; half++)
            if ((rc = launch_layer<256>(ctx, s->act3, s->w[3], s->b[3], s->feat, c, g4, half * 256, 512, st))) return rc;
        LayerGeom g5 = { MODE_NHWC, 1, 1, 1, 1, 512, 512, 1, 512, A1, 0, 1, 1.0f };
        if ((rc = launch_layer<32>(ctx, s->feat, s->w[4], s->b[4], heads + (size_t)i0 * A1, c, g5, 0, 32, st))) return rc;
    }
    return STM_OK;
}

} // extern "C"
