// stm_logits.cu — K5: SoftTreeMax logits, forward and backward, fused.
//
// Replaces policies/actor_critic_ts.py:60-95 (forward) and :147-188 (evaluate_actions, which loops
// over the batch in Python with ~10 tiny torch kernels per root) plus utils.py:52-57
// (add_regularization_logits).  Full-width trees only (max_width == -1): every root has L = A^d
// leaves, leaf l belongs to root action l / (L / A), and each leaf contributes its A head outputs,
// so a root action aggregates cnt = L entries (actor_critic_ts.py:66-69).
//
//   m[l,a']  = val_coef * head[l,a'] + reward[l]                                     (:66, :149)
//   z[g]     = logsumexp_{l in g, a'}(beta * m) - log(cnt)        (E-SoftTreeMax, :84-85)
//            = beta * sum(m) / cnt                                (C-SoftTreeMax, :81)
//   NaN in z -> 0                                                 (:88-90)
//   x[g]     = alpha * z[g] + (1 - alpha) * depth0[g]             (:94)
//   y[g]     = log((1 - eps) * softmax(x)[g] + eps / A)           (utils.py:52-57)
//   vmax     = max_l (val_coef * value[l] + reward[l])            (use_leaves_v, :62-63, :157-158)
//
// Width-limited trees (max_width != -1, actor_critic_ts.py:70-79,163-171): leaves carry an explicit root
// action (`group`, the reference's first_action), a root action aggregates count = members*A entries,
// an action with no surviving leaf gets -inf, and — reproduced on purpose — the reference scatters the
// members into a ZERO-initialised (A, n, A) tensor before the logsumexp, so every non-member slot
// contributes exp(beta*0) = 1 to the sum of its row (the "# - 1e6" comment at :166 is the fix the
// authors left out).  C-mode is unaffected (zeros add nothing to a sum).
//
// One warp per root: the L*A inputs of a root are a few KB, so the kernel is latency/HBM bound
// (reads L*A + L floats, writes A per root); lanes stride over a group's entries with an online
// max/sum and combine by shuffles.  fp32 throughout, expf/logf (not the fast intrinsics).
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>
#include "../../include/stm_b200.h"

namespace {

constexpr int kMaxA = 18;
constexpr int kWarpsPerBlock = 4;

__device__ __forceinline__ float warp_max(float v)
{
    for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xFFFFFFFFu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum(float v)
{
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}

// z[g] for one root, all lanes return the same value.  grp == nullptr: full-width tree, group g is the
// contiguous block of `per` leaves.  grp != nullptr: leaf l belongs to root action grp[l]; *cnt_out
// receives members*A; the E-mode sum includes one exp(0) per non-member slot (see the header).
__device__ float group_z(const float *head, const float *rew, const int *grp, int g, int per, int L, int A,
                         float val_coef, float beta, int cumulative, int lane, float *lse_out, int *cnt_out)
{
    const int l0 = grp ? 0 : g * per, nl = grp ? L : per, n = nl * A;
    const float *h = head + (size_t)l0 * A;
    int members = 0;
    if (grp) {
        for (int l = lane; l < L; l += 32) members += grp[l] == g;
        for (int o = 16; o; o >>= 1) members += __shfl_xor_sync(0xFFFFFFFFu, members, o);
    } else members = per;
    const int cnt = members * A;
    if (cnt_out) *cnt_out = cnt;
    if (cnt == 0) { if (lse_out) *lse_out = 0.f; return -CUDART_INF_F; }
    if (cumulative) {
        float s = 0.f;
        for (int i = lane; i < n; i += 32)
            if (!grp || grp[i / A] == g) s += val_coef * h[i] + rew[l0 + i / A];
        s = warp_sum(s);
        if (lse_out) *lse_out = 0.f;
        return beta * s / (float)cnt;
    }
    const bool zeros = grp && cnt < n;                       // non-member slots hold beta * 0
    float mx = zeros ? 0.f : -CUDART_INF_F;
    for (int i = lane; i < n; i += 32)
        if (!grp || grp[i / A] == g) mx = fmaxf(mx, beta * (val_coef * h[i] + rew[l0 + i / A]));
    mx = warp_max(mx);
    float s = 0.f;
    for (int i = lane; i < n; i += 32)
        if (!grp || grp[i / A] == g) s += expf(beta * (val_coef * h[i] + rew[l0 + i / A]) - mx);
    s = warp_sum(s);
    if (zeros) s += (float)(n - cnt) * expf(-mx);
    float lse = mx + logf(s);
    if (lse_out) *lse_out = lse;
    return lse - logf((float)cnt);
}

__global__ void __launch_bounds__(32 * kWarpsPerBlock)
stm_logits_fwd_kernel(const float *__restrict__ head, const float *__restrict__ rew, const float *__restrict__ depth0,
                      const float *__restrict__ values, const int *__restrict__ group, int R, int L, int A, float val_coef, float beta, float alpha,
                      float eps, int cumulative, float *__restrict__ logits, float *__restrict__ raw,
                      float *__restrict__ vmax)
{
    const int r = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r >= R) return;
    const float *h = head + (size_t)r * L * A, *rw = rew + (size_t)r * L;
    const int *gr = group ? group + (size_t)r * L : nullptr;
    const int per = L / A;
    float x[kMaxA];
    float xm = -CUDART_INF_F;
    for (int g = 0; g < A; g++) {
        float z = group_z(h, rw, gr, g, per, L, A, val_coef, beta, cumulative, lane, nullptr, nullptr);
        if (z != z) z = 0.f;                                            // NaN scrub
        if (raw && lane == 0) raw[(size_t)r * A + g] = z;
        float d0 = depth0 ? depth0[(size_t)r * A + g] : 0.f;
        if (d0 != d0) d0 = 0.f;
        x[g] = alpha * z + (1.f - alpha) * d0;
        xm = fmaxf(xm, x[g]);
    }
    float s = 0.f;
    for (int g = 0; g < A; g++) s += expf(x[g] - xm);
    const float lse = xm + logf(s);
    if (lane == 0)
        for (int g = 0; g < A; g++)
            logits[(size_t)r * A + g] = logf((1.f - eps) * expf(x[g] - lse) + eps / (float)A);
    if (vmax) {
        float m = -CUDART_INF_F;
        const float *v = values + (size_t)r * L;
        for (int l = lane; l < L; l += 32) m = fmaxf(m, val_coef * v[l] + rw[l]);
        m = warp_max(m);
        if (lane == 0) vmax[r] = m;
    }
}

// Backward.  Given dy = dLoss/dlogits:
//   p = softmax(x), q = (1-eps) p + eps/A, dq = dy / q, dp = (1-eps) dq, dx = p * (dp - sum(p dp))
//   dz = alpha dx, ddepth0 = (1-alpha) dx, dalpha = sum dx (z - depth0)
//   E: w = exp(beta m - lse_g): dhead = val_coef beta dz[g] w, dbeta = sum dz[g] w m
//   C: dhead = val_coef beta dz[g] / cnt, dbeta = sum dz[g] m / cnt
__global__ void __launch_bounds__(32 * kWarpsPerBlock)
stm_logits_bwd_kernel(const float *__restrict__ head, const float *__restrict__ rew, const float *__restrict__ depth0,
                      const int *__restrict__ group, int R, int L, int A, float val_coef, float beta, float alpha, float eps, int cumulative,
                      const float *__restrict__ dlogits, float *__restrict__ dhead, float *__restrict__ ddepth0,
                      float *__restrict__ dbeta, float *__restrict__ dalpha)
{
    const int r = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r >= R) return;
    const float *h = head + (size_t)r * L * A, *rw = rew + (size_t)r * L;
    const int *gr = group ? group + (size_t)r * L : nullptr;
    const int per = L / A;
    float z[kMaxA], lseg[kMaxA], x[kMaxA], dz[kMaxA];
    int cntg[kMaxA];
    bool nan[kMaxA];
    float xm = -CUDART_INF_F;
    for (int g = 0; g < A; g++) {
        z[g] = group_z(h, rw, gr, g, per, L, A, val_coef, beta, cumulative, lane, &lseg[g], &cntg[g]);
        nan[g] = z[g] != z[g];
        if (nan[g]) z[g] = 0.f;
        float d0 = depth0 ? depth0[(size_t)r * A + g] : 0.f;
        if (d0 != d0) d0 = 0.f;
        x[g] = alpha * z[g] + (1.f - alpha) * d0;
        xm = fmaxf(xm, x[g]);
    }
    float s = 0.f;
    for (int g = 0; g < A; g++) s += expf(x[g] - xm);                    // exp(-inf - xm) = 0 for empty groups
    const float lse = xm + logf(s);
    float pdp = 0.f, p[kMaxA], dp[kMaxA];
    for (int g = 0; g < A; g++) {
        p[g] = expf(x[g] - lse);
        float q = (1.f - eps) * p[g] + eps / (float)A;
        dp[g] = (1.f - eps) * dlogits[(size_t)r * A + g] / q;
        pdp += p[g] * dp[g];
    }
    float da = 0.f;
    for (int g = 0; g < A; g++) {
        float dx = p[g] * (dp[g] - pdp);
        float d0 = depth0 ? depth0[(size_t)r * A + g] : 0.f;
        if (d0 != d0) d0 = 0.f;
        if (cntg[g] > 0) da += dx * (z[g] - d0);
        dz[g] = (nan[g] || cntg[g] == 0) ? 0.f : alpha * dx;
        if (ddepth0 && lane == 0) ddepth0[(size_t)r * A + g] = (1.f - alpha) * dx;
    }
    float db = 0.f;
    float *dh = dhead + (size_t)r * L * A;
    if (gr) {
        for (int i = lane; i < L * A; i += 32) {
            const int g = gr[i / A];
            float m = val_coef * h[i] + rw[i / A];
            float w = cntg[g] == 0 ? 0.f : (cumulative ? 1.f / (float)cntg[g] : expf(beta * m - lseg[g]));
            dh[i] = val_coef * beta * dz[g] * w;
            db += dz[g] * w * m;
        }
    } else {
        const int n = per * A;
        for (int g = 0; g < A; g++) {
            const int l0 = g * per;
            for (int i = lane; i < n; i += 32) {
                float m = val_coef * h[(size_t)l0 * A + i] + rw[l0 + i / A];
                float w = cumulative ? 1.f / (float)n : expf(beta * m - lseg[g]);
                dh[(size_t)l0 * A + i] = val_coef * beta * dz[g] * w;
                db += dz[g] * w * m;
            }
        }
    }
    db = warp_sum(db);
    if (lane == 0) { dbeta[r] = db; dalpha[r] = da; }
}

} // namespace

extern "C" {

// declared in stm_kernels.cu
int stm_internal_fail(stm_ctx *ctx, int code, const char *msg);
void stm_internal_count_launch(stm_ctx *ctx);

static int check_args(stm_ctx *ctx, const void *head, const void *rew, int R, int L, int A, const char *who)
{
    if (!ctx || !head || !rew || R < 1 || L < 1 || A < 1) return stm_internal_fail(ctx, STM_ERR_ARG, who);
    if (A > kMaxA) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_logits: more than 18 actions");
    return STM_OK;
}

int stm_logits_fwd(stm_ctx *ctx, const float *head, const float *rewards, const float *depth0, const float *values,
                   const int *group, int R, int L, int A, float val_coef, float beta, float alpha, float eps, int cumulative,
                   float *logits, float *raw, float *vmax, void *stream)
{
    int rc = check_args(ctx, head, rewards, R, L, A, "stm_logits_fwd: bad argument");
    if (rc) return rc;
    if (!group && L % A != 0) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_logits_fwd: without group ids L must be a multiple of A (full-width tree)");
    if (!logits || (vmax && !values)) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_logits_fwd: bad argument");
    int blocks = (R + kWarpsPerBlock - 1) / kWarpsPerBlock;
    stm_logits_fwd_kernel<<<blocks, 32 * kWarpsPerBlock, 0, (cudaStream_t)stream>>>(
        head, rewards, depth0, values, group, R, L, A, val_coef, beta, alpha, eps, cumulative, logits, raw, vmax);
    stm_internal_count_launch(ctx);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? STM_OK : stm_internal_fail(ctx, STM_ERR_CUDA, cudaGetErrorString(e));
}

int stm_logits_bwd(stm_ctx *ctx, const float *head, const float *rewards, const float *depth0, const int *group,
                   int R, int L, int A, float val_coef, float beta, float alpha, float eps, int cumulative,
                   const float *dlogits, float *dhead, float *ddepth0, float *dbeta, float *dalpha, void *stream)
{
    int rc = check_args(ctx, head, rewards, R, L, A, "stm_logits_bwd: bad argument");
    if (rc) return rc;
    if (!group && L % A != 0) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_logits_bwd: without group ids L must be a multiple of A (full-width tree)");
    if (!dlogits || !dhead || !dbeta || !dalpha) return stm_internal_fail(ctx, STM_ERR_ARG, "stm_logits_bwd: bad argument");
    int blocks = (R + kWarpsPerBlock - 1) / kWarpsPerBlock;
    stm_logits_bwd_kernel<<<blocks, 32 * kWarpsPerBlock, 0, (cudaStream_t)stream>>>(
        head, rewards, depth0, group, R, L, A, val_coef, beta, alpha, eps, cumulative, dlogits, dhead, ddepth0, dbeta, dalpha);
    stm_internal_count_launch(ctx);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? STM_OK : stm_internal_fail(ctx, STM_ERR_CUDA, cudaGetErrorString(e));
}

} // extern "C"
