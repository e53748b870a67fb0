// Diffusion-process kernels: one CTA per graph, all reductions in a fixed order (bit-identical reruns), the
// elementwise updates written with explicitly rounded fp32 operations in the reference's operation order so that
// given identical inputs they reproduce torch's CPU results bit for bit.
//
//   remove_mean                      diffusion_x_h.py:5-14 (per-graph branch writes in place)
//   diffuse_batch                    diffusion_x_h.py:51-59 inside parts/train_per_iretation.py:52-72,133-145
//   loss_fwd_bwd                     parts/train_per_iretation.py:161-169 (+ its gradient)
//   sampler_step                     diffusion_x_h.py:61-90 inside parts/train_per_iretation.py:335-428
//   adam_step                        torch.optim.Adam as selected by parts/def_for_main.py:133-134
//   radam_sf_step, lerp              schedulefree.RAdamScheduleFree as selected by parts/def_for_main.py:137-138
#include "common.cuh"

namespace {

// column means of v[a..b)[C0..C0+3) with row stride ld, minus nothing; fixed order; result in all threads
__device__ __forceinline__ void graph_mean3(const float* __restrict__ v, const float* __restrict__ w, int ld, int a, int b,
                                            float* red, float out[3]) {
    // mean over nodes of (v - w) (w may be NULL)
    float s[3] = {0.f, 0.f, 0.f};
    for (int i = a + threadIdx.x; i < b; i += 256) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float q = v[(size_t)i * ld + c];
            if (w) q = __fsub_rn(q, w[(size_t)i * ld + c]);
            s[c] += q;
        }
    }
    const float n = (float)(b - a);
#pragma unroll
    for (int c = 0; c < 3; ++c) out[c] = __fdiv_rn(block_sum_256(s[c], red), n);
}

__global__ void __launch_bounds__(256) remove_mean_kernel(float* __restrict__ x, int C, const int32_t* __restrict__ node_ptr) {
    __shared__ float red[8];
    const int a = node_ptr[blockIdx.x], b = node_ptr[blockIdx.x + 1];
    if (a >= b) return;
    const float n = (float)(b - a);
    for (int c = 0; c < C; ++c) {
        float s = 0.f;
        for (int i = a + threadIdx.x; i < b; i += 256) s += x[(size_t)i * C + c];
        const float mean = __fdiv_rn(block_sum_256(s, red), n);
        for (int i = a + threadIdx.x; i < b; i += 256) x[(size_t)i * C + c] = __fsub_rn(x[(size_t)i * C + c], mean);
    }
}

struct DiffuseArgs {
    const float *pos, *x0, *noise_x, *noise_h, *cond;
    const int32_t* t;
    const float *alpha, *sigma;
    int T;
    const int32_t* node_ptr;
    int A, C;
    float *z_x, *eps_x, *h_in;
};
__global__ void __launch_bounds__(256) diffuse_batch_kernel(DiffuseArgs p) {
    __shared__ float red[8];
    const int g = blockIdx.x;
    const int a = p.node_ptr[g], b = p.node_ptr[g + 1];
    if (a >= b) return;
    const int t = p.t[g];
    const float at = p.alpha[t], st = p.sigma[t];
    float mean[3];
    graph_mean3(p.noise_x, nullptr, 3, a, b, red, mean);
    for (int idx = 3 * a + threadIdx.x; idx < 3 * b; idx += 256) {
        const float e = __fsub_rn(p.noise_x[idx], mean[idx % 3]);
        p.eps_x[idx] = e;
        p.z_x[idx] = __fadd_rn(__fmul_rn(at, p.pos[idx]), __fmul_rn(st, e));
    }
    const int H = p.A + p.C + 1;
    const float tcol = __fdiv_rn((float)t, (float)p.T);
    for (int idx = a * H + threadIdx.x; idx < b * H; idx += 256) {
        const int i = idx / H, c = idx - i * H;
        float v;
        if (c < p.A) v = __fadd_rn(__fmul_rn(at, p.x0[(size_t)i * p.A + c]), __fmul_rn(st, p.noise_h[(size_t)i * p.A + c]));
        else if (c < p.A + p.C) v = p.cond[(size_t)i * p.C + (c - p.A)];
        else v = tcol;
        p.h_in[idx] = v;
    }
}

struct LossArgs {
    const float *x_out, *h_out, *z_x, *eps_x, *eps_h;
    const int32_t* node_ptr;
    int G, H, A;
    float *loss_g, *gx_out, *gh_out;
};
__global__ void __launch_bounds__(256) loss_kernel(LossArgs p) {
    __shared__ float red[8];
    const int g = blockIdx.x;
    const int a = p.node_ptr[g], b = p.node_ptr[g + 1];
    if (a >= b) {
        if (threadIdx.x == 0) p.loss_g[g] = 0.f;
        return;
    }
    float mean[3];
    graph_mean3(p.x_out, p.z_x, 3, a, b, red, mean);  // COM of the predicted position noise
    // residual d = P(x' - z_x) - eps ; its own mean is needed for the projected gradient
    float ds[3] = {0.f, 0.f, 0.f};
    float sq = 0.f;
    for (int i = a + threadIdx.x; i < b; i += 256) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float d = (p.x_out[3 * i + c] - p.z_x[3 * i + c] - mean[c]) - p.eps_x[3 * i + c];
            ds[c] += d;
            sq += d * d;
        }
    }
    const float n = (float)(b - a);
    float dmean[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) dmean[c] = block_sum_256(ds[c], red) / n;
    const float scale = 2.f / (float)p.G;
    for (int idx = 3 * a + threadIdx.x; idx < 3 * b; idx += 256) {
        const int c = idx % 3;
        const float d = (p.x_out[idx] - p.z_x[idx] - mean[c]) - p.eps_x[idx];
        p.gx_out[idx] = scale * (d - dmean[c]);  // P is symmetric: grad wrt x' = P(2 d / G)
    }
    for (int idx = a * p.H + threadIdx.x; idx < b * p.H; idx += 256) {
        const int i = idx / p.H, c = idx - i * p.H;
        float gval = 0.f;
        if (c < p.A) {
            const float d = p.h_out[idx] - p.eps_h[(size_t)i * p.A + c];
            sq += d * d;
            gval = scale * d;
        }
        p.gh_out[idx] = gval;
    }
    const float tot = block_sum_256(sq, red);
    if (threadIdx.x == 0) p.loss_g[g] = tot / (float)p.G;
}

struct StepArgs {
    float *z_x, *h_in;
    const float *x_out, *h_out, *noise_x, *noise_h;
    const float* coef;     // device [T+1][8]: a_ts, s2_ts, sig_t, std, t_next column, alpha_0, sigma_0, -
    const int32_t* t_dev;  // current t (device), NULL when final
    int t_base, period;
    long long stride_x, stride_h;  // noise of step t lives at noise + ((t_base - t) % period) * stride
    float onehot_scale;
    const int32_t* node_ptr;
    int H, A, final_step;
    float* z_h_out;
    int32_t* finite_flag;
};
__global__ void __launch_bounds__(256) sampler_step_kernel(StepArgs p) {
    __shared__ float red[8];
    pdl_launch_dependents();
    pdl_wait();  // common.cuh: PDL -- x_out / h_out come from the denoiser's last kernels
    const int g = blockIdx.x;
    const int a = p.node_ptr[g], b = p.node_ptr[g + 1];
    if (a >= b) return;
    const int t = p.final_step ? 0 : *p.t_dev;
    const float* cf = p.coef + 8 * t;
    int slot = p.final_step ? 0 : p.t_base - t;
    if (p.period > 0) slot %= p.period;
    const float* nx = p.noise_x + (long long)slot * p.stride_x;
    const float* nh = p.noise_h + (long long)slot * p.stride_h;
    float dmean[3], nmean[3];
    graph_mean3(p.x_out, p.z_x, 3, a, b, red, dmean);  // eps_x = P(x' - z_x)
    graph_mean3(nx, nullptr, 3, a, b, red, nmean);     // noise is made COM-free for positions
    bool ok = true;
    const int H = p.H, A = p.A;
    if (!p.final_step) {
        const float a_ts = cf[0], s2_ts = cf[1], sig_t = cf[2], sd = cf[3], tnext = cf[4];
        for (int idx = 3 * a + threadIdx.x; idx < 3 * b; idx += 256) {
            const int c = idx % 3;
            const float z = p.z_x[idx];
            const float eps = __fsub_rn(__fsub_rn(p.x_out[idx], z), dmean[c]);
            const float mu = __fsub_rn(__fdiv_rn(z, a_ts), __fdiv_rn(__fdiv_rn(__fmul_rn(s2_ts, eps), a_ts), sig_t));
            const float v = __fadd_rn(mu, __fmul_rn(sd, __fsub_rn(nx[idx], nmean[c])));
            ok = ok && isfinite(v);
            p.z_x[idx] = v;
        }
        for (int idx = a * H + threadIdx.x; idx < b * H; idx += 256) {
            const int i = idx / H, c = idx - i * H;
            if (c < A) {
                const float z = p.h_in[idx];
                const float eps = p.h_out[idx];
                const float mu = __fsub_rn(__fdiv_rn(z, a_ts), __fdiv_rn(__fdiv_rn(__fmul_rn(s2_ts, eps), a_ts), sig_t));
                const float v = __fadd_rn(mu, __fmul_rn(sd, nh[(size_t)i * A + c]));
                ok = ok && isfinite(v);
                p.h_in[idx] = __fmul_rn(p.onehot_scale, v);
            } else if (c == H - 1) {
                p.h_in[idx] = tnext;
            }
        }
    } else {
        const float a0 = cf[5], s0 = cf[6];
        for (int idx = 3 * a + threadIdx.x; idx < 3 * b; idx += 256) {
            const int c = idx % 3;
            const float z = p.z_x[idx];
            const float eps = __fsub_rn(__fsub_rn(p.x_out[idx], z), dmean[c]);
            const float mu = __fsub_rn(__fdiv_rn(z, a0), __fdiv_rn(__fmul_rn(s0, eps), a0));
            const float v = __fadd_rn(mu, __fdiv_rn(__fmul_rn(s0, __fsub_rn(nx[idx], nmean[c])), a0));
            ok = ok && isfinite(v);
            p.z_x[idx] = v;
        }
        for (int idx = a * A + threadIdx.x; idx < b * A; idx += 256) {
            const int i = idx / A, c = idx - i * A;
            const float z = p.h_in[(size_t)i * H + c];
            const float eps = p.h_out[(size_t)i * H + c];
            const float mu = __fsub_rn(__fdiv_rn(z, a0), __fdiv_rn(__fmul_rn(s0, eps), a0));
            const float v = __fadd_rn(mu, __fdiv_rn(__fmul_rn(s0, nh[idx]), a0));
            ok = ok && isfinite(v);
            p.z_h_out[idx] = v;
        }
    }
    const int all_ok = __syncthreads_and(ok ? 1 : 0);
    if (threadIdx.x == 0 && !all_ok) p.finite_flag[g] = 0;
}

__global__ void counter_add_kernel(int32_t* ctr, int32_t delta) {
    pdl_wait();  // the kernel in front (sampler step, or the last backward kernel) may still be reading the counter
    *ctr += delta;
}

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                   float* __restrict__ v, long long n, float lr, float b1, float b2,
                                                   float eps, float wd, float bc1, float bc2_sqrt, float gscale) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    float gi = g[i] * gscale;
    const float pi = p[i];
    if (wd != 0.f) gi = fmaf(wd, pi, gi);
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = pi - (lr / bc1) * (mi / denom);
}

// Adam with the step number in DEVICE memory: a CUDA-graph replay of a training step must not freeze the bias corrections
// at the step it was captured on.  *step_dev is the 1-based number of THIS update.
__global__ void __launch_bounds__(256) adam_dev_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                       float* __restrict__ v, long long n, float lr, float b1, float b2,
                                                       float eps, float wd, const int32_t* __restrict__ step_dev, float gscale) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const float step = (float)(*step_dev);
    const float bc1 = 1.f - powf(b1, step), bc2_sqrt = sqrtf(1.f - powf(b2, step));
    float gi = g[i] * gscale;
    const float pi = p[i];
    if (wd != 0.f) gi = fmaf(wd, pi, gi);
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = pi - (lr / bc1) * (mi / denom);
}

// torch.lerp with a scalar weight (ATen's two-sided form, so that weight = 1 returns `end` exactly)
__device__ __forceinline__ float lerp_f(float a, float b, float w) {
    const float d = b - a;
    return w < 0.5f ? fmaf(w, d, a) : b - d * (1.f - w);
}

// One RAdamScheduleFree update of a flat parameter range (see egnn_radam_sf_step in the header); scalars from the host.
__global__ void __launch_bounds__(256) radam_sf_kernel(float* __restrict__ y, const float* __restrict__ g, float* __restrict__ z,
                                                       float* __restrict__ v, long long n, float lr, float b2, float eps,
                                                       float wd, float ckp1, float y_lr, float bc2, int adam, float gscale) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    float gi = g[i] * gscale;
    const float vi = fmaf(1.f - b2, gi * gi, b2 * v[i]);
    v[i] = vi;
    if (adam) gi = gi / (sqrtf(vi / bc2) + eps);
    float yi = y[i];
    if (wd != 0.f) gi = fmaf(wd, yi, gi);  // weight decay evaluated at y
    const float zi = z[i];
    yi = lerp_f(yi, zi, ckp1);
    y[i] = fmaf(y_lr, gi, yi);
    z[i] = fmaf(-lr, gi, zi);
}

__global__ void __launch_bounds__(256) lerp_kernel(float* __restrict__ p, const float* __restrict__ end, long long n, float w) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i < n) p[i] = lerp_f(p[i], end[i], w);
}

}  // namespace

extern "C" int egnn_remove_mean(float* x, int32_t C, const int32_t* node_ptr, int32_t G, void* stream) {
    EGNN_REQUIRE(x && node_ptr && C >= 1 && G >= 0, "egnn_remove_mean: bad arguments");
    if (G == 0) return EGNN_OK;
    remove_mean_kernel<<<G, 256, 0, (cudaStream_t)stream>>>(x, C, node_ptr);
    EGNN_CHECK_LAUNCH("remove_mean_kernel");
    return EGNN_OK;
}

extern "C" int egnn_diffuse_batch(const float* pos, const float* x0, const float* noise_x, const float* noise_h,
                                  const float* cond, const int32_t* t, const float* alpha, const float* sigma, int32_t T,
                                  const int32_t* node_ptr, int32_t G, int32_t A, int32_t C, float* z_x, float* eps_x,
                                  float* h_in, void* stream) {
    EGNN_REQUIRE(pos && x0 && noise_x && noise_h && t && alpha && sigma && node_ptr && z_x && eps_x && h_in,
                 "egnn_diffuse_batch: NULL pointer");
    EGNN_REQUIRE(T >= 1 && G >= 1 && A >= 1 && C >= 0 && (C == 0 || cond), "egnn_diffuse_batch: bad sizes");
    DiffuseArgs p{pos, x0, noise_x, noise_h, cond, t, alpha, sigma, T, node_ptr, A, C, z_x, eps_x, h_in};
    diffuse_batch_kernel<<<G, 256, 0, (cudaStream_t)stream>>>(p);
    EGNN_CHECK_LAUNCH("diffuse_batch_kernel");
    return EGNN_OK;
}

extern "C" int egnn_loss_fwd_bwd(const float* x_out, const float* h_out, const float* z_x, const float* eps_x,
                                 const float* eps_h, const int32_t* node_ptr, int32_t G, int32_t H, int32_t A,
                                 float* loss_g, float* gx_out, float* gh_out, void* stream) {
    EGNN_REQUIRE(x_out && h_out && z_x && eps_x && eps_h && node_ptr && loss_g && gx_out && gh_out,
                 "egnn_loss_fwd_bwd: NULL pointer");
    EGNN_REQUIRE(G >= 1 && H >= 1 && A >= 0 && A <= H, "egnn_loss_fwd_bwd: bad sizes");
    LossArgs p{x_out, h_out, z_x, eps_x, eps_h, node_ptr, G, H, A, loss_g, gx_out, gh_out};
    loss_kernel<<<G, 256, 0, (cudaStream_t)stream>>>(p);
    EGNN_CHECK_LAUNCH("loss_kernel");
    return EGNN_OK;
}

extern "C" int egnn_sampler_step(float* z_x, float* h_in, const float* x_out, const float* h_out, const float* noise_x,
                                 const float* noise_h, const float* coef, const int32_t* t_dev, int32_t t_base,
                                 int64_t noise_stride_x, int64_t noise_stride_h, int32_t noise_period,
                                 float onehot_scale, const int32_t* node_ptr, int32_t G, int32_t H, int32_t A, int32_t final_step,
                                 float* z_h_out, int32_t* finite_flag, void* stream) {
    EGNN_REQUIRE(z_x && h_in && x_out && h_out && noise_x && noise_h && coef && node_ptr && finite_flag,
                 "egnn_sampler_step: NULL pointer");
    EGNN_REQUIRE(final_step ? z_h_out != nullptr : t_dev != nullptr, "egnn_sampler_step: t_dev / z_h_out missing");
    EGNN_REQUIRE(G >= 1 && H >= 1 && A >= 1 && A < H, "egnn_sampler_step: bad sizes");
    StepArgs p{z_x, h_in, x_out, h_out, noise_x, noise_h, coef, t_dev, t_base, noise_period, noise_stride_x, noise_stride_h,
               onehot_scale, node_ptr, H, A, final_step, z_h_out, finite_flag};
    EGNN_CUDA(egnn_launch(sampler_step_kernel, dim3(G), dim3(256), 0, (cudaStream_t)stream, true, p));
    EGNN_CHECK_LAUNCH("sampler_step_kernel");
    return EGNN_OK;
}

extern "C" int egnn_counter_add(int32_t* counter, int32_t delta, void* stream) {
    EGNN_REQUIRE(counter, "egnn_counter_add: NULL pointer");
    EGNN_CUDA(egnn_launch(counter_add_kernel, dim3(1), dim3(1), 0, (cudaStream_t)stream, true, counter, delta));
    EGNN_CHECK_LAUNCH("counter_add_kernel");
    return EGNN_OK;
}

extern "C" int egnn_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                              float beta1, float beta2, float eps, float weight_decay, int32_t step, float grad_scale,
                              void* stream) {
    EGNN_REQUIRE(params && grads && exp_avg && exp_avg_sq && n >= 0 && step >= 1, "egnn_adam_step: bad arguments");
    if (n == 0) return EGNN_OK;
    const float bc1 = 1.f - powf(beta1, (float)step);
    const float bc2s = sqrtf(1.f - powf(beta2, (float)step));
    adam_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(params, grads, exp_avg, exp_avg_sq, n, lr,
                                                                              beta1, beta2, eps, weight_decay, bc1, bc2s,
                                                                              grad_scale);
    EGNN_CHECK_LAUNCH("adam_kernel");
    return EGNN_OK;
}

extern "C" int egnn_radam_sf_step(float* y, const float* grads, float* z, float* exp_avg_sq, int64_t n, float lr, float beta2,
                                  float eps, float weight_decay, float ckp1, float adaptive_y_lr, float bias_correction2,
                                  int32_t adam_phase, float grad_scale, void* stream) {
    EGNN_REQUIRE(y && grads && z && exp_avg_sq && n >= 0 && bias_correction2 > 0.f, "egnn_radam_sf_step: bad arguments");
    if (n == 0) return EGNN_OK;
    radam_sf_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(y, grads, z, exp_avg_sq, n, lr, beta2, eps,
                                                                                  weight_decay, ckp1, adaptive_y_lr,
                                                                                  bias_correction2, adam_phase, grad_scale);
    EGNN_CHECK_LAUNCH("radam_sf_kernel");
    return EGNN_OK;
}

extern "C" int egnn_lerp(float* p, const float* end, int64_t n, float weight, void* stream) {
    EGNN_REQUIRE(p && end && n >= 0, "egnn_lerp: bad arguments");
    if (n == 0) return EGNN_OK;
    lerp_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(p, end, n, weight);
    EGNN_CHECK_LAUNCH("lerp_kernel");
    return EGNN_OK;
}

extern "C" int egnn_adam_step_dev(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                                  float beta1, float beta2, float eps, float weight_decay, const int32_t* step_dev,
                                  float grad_scale, void* stream) {
    EGNN_REQUIRE(params && grads && exp_avg && exp_avg_sq && step_dev && n >= 0, "egnn_adam_step_dev: bad arguments");
    if (n == 0) return EGNN_OK;
    adam_dev_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(params, grads, exp_avg, exp_avg_sq, n, lr,
                                                                                  beta1, beta2, eps, weight_decay, step_dev,
                                                                                  grad_scale);
    EGNN_CHECK_LAUNCH("adam_dev_kernel");
    return EGNN_OK;
}
