// Structure-analysis kernels for "sample -> evaluate" (SURVEY.md 8f rank 2): the reference's per-structure python loops
// (evaluate_RDF.py:39-60, evaluate_Si-O-Si.py:23-53 + CN2_evaluate.py:12-21, evaluate_rmsd_for_pos_generate.py:11-51) as
// one CTA per structure.  Integer results (histogram counts, neighbour indices) are bit-exact against the reference;
// fp64 curves follow its operation order; no floating-point atomics, every reduction has a fixed order.
// All three kernels are HBM-bound streaming reads of pos [N,3] (12 B per atom) with a few hundred bytes of output per
// structure.
#include "common.cuh"

namespace {

constexpr int AT = 256;  // threads per structure

// |p_i - p_0| exactly as torch.norm evaluates a 3-vector on the CPU (pinned in oracle/make_golden.py against 2e5 vectors)
__device__ __forceinline__ float dist3(float dx, float dy, float dz) {
    return __fsqrt_rn(__fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx))));
}

// fixed-order block sum of doubles (tree over shared memory; the same order for every launch)
__device__ double block_sum(double v, double* red) {
    const int t = threadIdx.x;
    __syncthreads();
    red[t] = v;
    __syncthreads();
#pragma unroll
    for (int s = AT / 2; s > 0; s >>= 1) {
        if (t < s) red[t] += red[t + s];
        __syncthreads();
    }
    return red[0];
}

// ---------------------------------------------------------------------------------------------------------------------
// Local RDF around atom 0 of each structure (evaluate_RDF.py:48-60).
//   counts[k] = #{ i >= 1 : lo[k] < d_i < hi[k] }     lo/hi = fp32(r_k), fp32(r_k + dR): the reference's chained
//                                                     comparison runs in fp32 (oracle/eval_ref.py: rdf_edges)
//   raw[k]    = counts[k] / (((4 pi * rho) * r_k^2) * dR),  rho = n / vol      (float64, the reference's association)
//   smooth    = scipy.ndimage.gaussian_filter1d(raw, sigma): correlation with host-computed weights w[-rad..rad],
//               boundary mode 'reflect' (d c b a | a b c d | d c b a)
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(AT) rdf_local_kernel(const float* __restrict__ pos, const int32_t* __restrict__ node_ptr,
                                                        const float* __restrict__ lo, const float* __restrict__ hi,
                                                        const double* __restrict__ r2, int nb, float inv_dr,
                                                        double four_pi, double vol, double dR,
                                                        const double* __restrict__ w, int rad,
                                                        int32_t* __restrict__ counts, double* __restrict__ smooth,
                                                        float* __restrict__ dists) {
    extern __shared__ __align__(16) unsigned char sm[];
    double* raw = reinterpret_cast<double*>(sm);
    int* cnt = reinterpret_cast<int*>(sm + (size_t)nb * 8);
    const int s = blockIdx.x, t = threadIdx.x;
    const int n0 = node_ptr[s], n = node_ptr[s + 1] - n0;
    for (int k = t; k < nb; k += AT) cnt[k] = 0;
    __syncthreads();
    if (n > 0) {
        const float x0 = pos[3 * (size_t)n0], y0 = pos[3 * (size_t)n0 + 1], z0 = pos[3 * (size_t)n0 + 2];
        if (t == 0 && dists) dists[n0] = 0.f;
        for (int i = 1 + t; i < n; i += AT) {
            const float* p = pos + 3 * (size_t)(n0 + i);
            const float d = dist3(__fsub_rn(p[0], x0), __fsub_rn(p[1], y0), __fsub_rn(p[2], z0));
            if (dists) dists[n0 + i] = d;
            // bin k covers about (k+1) dR .. (k+2) dR; fp32 edge images may overlap or gap by an ulp: test the neighbours too
            const float g = d * inv_dr;
            if (g < (float)(nb + 4)) {  // also false for NaN
                const int kc = (int)g - 1;
#pragma unroll
                for (int k = kc - 2; k <= kc + 2; ++k)
                    if (k >= 0 && k < nb && lo[k] < d && d < hi[k]) atomicAdd(&cnt[k], 1);  // integer: order-free
            }
        }
    }
    __syncthreads();
    const double rho = (double)n / vol;
    const double c0 = __dmul_rn(four_pi, rho);
    for (int k = t; k < nb; k += AT) {
        counts[(size_t)s * nb + k] = cnt[k];
        raw[k] = __ddiv_rn((double)cnt[k], __dmul_rn(__dmul_rn(c0, r2[k]), dR));
    }
    __syncthreads();
    for (int k = t; k < nb; k += AT) {
        // scipy's correlate1d sums the centre tap first, then symmetric pairs outwards from the far end
        auto at = [&](int j) {  // 'reflect' extension, valid for any offset
            const int period = 2 * nb;
            j %= period;
            if (j < 0) j += period;
            return raw[j < nb ? j : period - 1 - j];
        };
        double acc = __dmul_rn(at(k), w[rad]);
        for (int j = -rad; j < 0; ++j) acc = __dadd_rn(acc, __dmul_rn(__dadd_rn(at(k + j), at(k - j)), w[j + rad]));
        smooth[(size_t)s * nb + k] = acc;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// CN2 site of atom 0 (evaluate_Si-O-Si.py:28-39, CN2_evaluate.py:12-21).
//   radius > 0 : neighbours = atoms i >= 1 with |p_i - p_0| < radius; count[s] of them; idx = the first two in index order
//   radius <= 0: idx = the two atoms nearest to atom 0 (in index order), count = min(2, n - 1)
//   angle (degrees, float64(acosf(cos)) * 180/pi) at atom 0 between idx[0] and idx[1], and the two bond lengths; written
//   whenever two atoms were selected (the caller filters on count == 2 like :37)
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(AT) cn2_kernel(const float* __restrict__ pos, const int32_t* __restrict__ node_ptr,
                                                  float radius, int32_t* __restrict__ count, int32_t* __restrict__ idx,
                                                  double* __restrict__ angle, float* __restrict__ bonds) {
    __shared__ int s_cnt;
    __shared__ unsigned long long s_key[2];
    const int s = blockIdx.x, t = threadIdx.x;
    const int n0 = node_ptr[s], n = node_ptr[s + 1] - n0;
    const bool by_radius = radius > 0.f;
    if (t == 0) {
        s_cnt = 0;
        s_key[0] = s_key[1] = ~0ull;
    }
    __syncthreads();
    float x0 = 0.f, y0 = 0.f, z0 = 0.f;
    if (n > 0) {
        x0 = pos[3 * (size_t)n0];
        y0 = pos[3 * (size_t)n0 + 1];
        z0 = pos[3 * (size_t)n0 + 2];
    }
    // key = (ordering value << 32) | index: index order under the radius rule, distance order otherwise
    auto key_of = [&](int i, bool& take) {
        const float* p = pos + 3 * (size_t)(n0 + i);
        const float d = dist3(__fsub_rn(p[0], x0), __fsub_rn(p[1], y0), __fsub_rn(p[2], z0));
        take = by_radius ? d < radius : d == d;
        return ((unsigned long long)(by_radius ? 0u : __float_as_uint(d)) << 32) | (unsigned)i;
    };
    for (int i = 1 + t; i < n; i += AT) {
        bool take;
        const unsigned long long k = key_of(i, take);
        if (take) {
            atomicAdd(&s_cnt, 1);
            atomicMin(&s_key[0], k);
        }
    }
    __syncthreads();
    const unsigned long long first = s_key[0];
    for (int i = 1 + t; i < n; i += AT) {
        bool take;
        const unsigned long long k = key_of(i, take);
        if (take && k > first) atomicMin(&s_key[1], k);
    }
    __syncthreads();
    if (t == 0) {
        int a = s_key[0] == ~0ull ? -1 : (int)(s_key[0] & 0xffffffffu);
        int b = s_key[1] == ~0ull ? -1 : (int)(s_key[1] & 0xffffffffu);
        if (a >= 0 && b >= 0 && b < a) {  // report in index order (position[[0] + index_list], :43)
            const int c = a;
            a = b;
            b = c;
        }
        count[s] = by_radius ? s_cnt : min(s_cnt, 2);
        idx[2 * s] = a;
        idx[2 * s + 1] = b;
        double ang = __longlong_as_double(0x7ff8000000000000ll);
        float b1 = __int_as_float(0x7fc00000), b2 = b1;
        if (a >= 0 && b >= 0) {
            const float* pa = pos + 3 * (size_t)(n0 + a);
            const float* pb = pos + 3 * (size_t)(n0 + b);
            const float ax = __fsub_rn(pa[0], x0), ay = __fsub_rn(pa[1], y0), az = __fsub_rn(pa[2], z0);
            const float bx = __fsub_rn(pb[0], x0), by = __fsub_rn(pb[1], y0), bz = __fsub_rn(pb[2], z0);
            b1 = dist3(ax, ay, az);
            b2 = dist3(bx, by, bz);
            const float dot = __fmaf_rn(az, bz, __fmaf_rn(ay, by, __fmul_rn(ax, bx)));
            const float c = __fdiv_rn(dot, __fmul_rn(b1, b2));
            ang = (double)acosf(c) * (180.0 / 3.14159265358979323846);
        }
        angle[s] = ang;
        bonds[2 * s] = b1;
        bonds[2 * s + 1] = b2;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Kabsch alignment of P onto Q per structure (evaluate_rmsd_for_pos_generate.py:11-51): R, t, RMSD.
// H = p^T q (centred), R0 = V U^T = orthogonal polar factor of H^T, from the eigen-decomposition of H^T H (cyclic Jacobi,
// float64).  Kept quirk (:42-45): when det(R0) < 0 the reference negates the last COLUMN of Vt, i.e. R = diag(1,1,-1) R0.
// Accumulation in float64 (the reference's float32 sums differ from this by ~1e-6 relative).
// ---------------------------------------------------------------------------------------------------------------------
__device__ void jacobi3(double A[3][3], double V[3][3]) {
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) V[i][j] = i == j ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 12; ++sweep) {
        const double off = fabs(A[0][1]) + fabs(A[0][2]) + fabs(A[1][2]);
        if (off < 1e-300) break;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                if (fabs(A[p][q]) < 1e-300) continue;
                const double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
                const double tt = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(tt * tt + 1.0), sn = tt * c;
                for (int k = 0; k < 3; ++k) {  // A <- A J
                    const double akp = A[k][p], akq = A[k][q];
                    A[k][p] = c * akp - sn * akq;
                    A[k][q] = sn * akp + c * akq;
                }
                for (int k = 0; k < 3; ++k) {  // A <- J^T A
                    const double apk = A[p][k], aqk = A[q][k];
                    A[p][k] = c * apk - sn * aqk;
                    A[q][k] = sn * apk + c * aqk;
                }
                for (int k = 0; k < 3; ++k) {
                    const double vkp = V[k][p], vkq = V[k][q];
                    V[k][p] = c * vkp - sn * vkq;
                    V[k][q] = sn * vkp + c * vkq;
                }
            }
    }
}

__global__ void __launch_bounds__(AT) kabsch_kernel(const float* __restrict__ P, const float* __restrict__ Q,
                                                     const int32_t* __restrict__ node_ptr, float* __restrict__ Rout,
                                                     float* __restrict__ tout, float* __restrict__ rmsd) {
    __shared__ double red[AT];
    __shared__ double sR[9];
    const int s = blockIdx.x, t = threadIdx.x;
    const int n0 = node_ptr[s], n = node_ptr[s + 1] - n0;
    double c[6] = {0, 0, 0, 0, 0, 0};
    for (int i = t; i < n; i += AT)
        for (int k = 0; k < 3; ++k) {
            c[k] += (double)P[3 * (size_t)(n0 + i) + k];
            c[3 + k] += (double)Q[3 * (size_t)(n0 + i) + k];
        }
    for (int k = 0; k < 6; ++k) c[k] = block_sum(c[k], red) / (double)max(n, 1);
    double h[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = t; i < n; i += AT) {
        double p[3], q[3];
        for (int k = 0; k < 3; ++k) {
            p[k] = (double)P[3 * (size_t)(n0 + i) + k] - c[k];
            q[k] = (double)Q[3 * (size_t)(n0 + i) + k] - c[3 + k];
        }
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) h[3 * a + b] += p[a] * q[b];
    }
    for (int k = 0; k < 9; ++k) h[k] = block_sum(h[k], red);
    if (t == 0) {
        double Hm[3][3], K[3][3], V[3][3], U[3][3], sg[3];
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) Hm[a][b] = h[3 * a + b];
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) K[a][b] = Hm[0][a] * Hm[0][b] + Hm[1][a] * Hm[1][b] + Hm[2][a] * Hm[2][b];
        jacobi3(K, V);  // H^T H = V diag(sigma^2) V^T
        int ord[3] = {0, 1, 2};  // descending singular values
        for (int a = 0; a < 2; ++a)
            for (int b = a + 1; b < 3; ++b)
                if (K[ord[b]][ord[b]] > K[ord[a]][ord[a]]) {
                    const int tmp = ord[a];
                    ord[a] = ord[b];
                    ord[b] = tmp;
                }
        double Vs[3][3];
        for (int j = 0; j < 3; ++j) {
            sg[j] = sqrt(fmax(K[ord[j]][ord[j]], 0.0));
            for (int k = 0; k < 3; ++k) Vs[k][j] = V[k][ord[j]];
        }
        for (int j = 0; j < 3; ++j) {  // u_j = H v_j / sigma_j
            const bool tiny = sg[j] <= 1e-12 * fmax(sg[0], 1e-300);
            if (!tiny || j < 2) {
                const double inv = sg[j] > 0 ? 1.0 / sg[j] : 0.0;
                for (int k = 0; k < 3; ++k) U[k][j] = (Hm[k][0] * Vs[0][j] + Hm[k][1] * Vs[1][j] + Hm[k][2] * Vs[2][j]) * inv;
            } else {  // coplanar point sets: the third pair is not determined by H; take the proper completion
                const double dv = Vs[0][0] * (Vs[1][1] * Vs[2][2] - Vs[1][2] * Vs[2][1]) -
                                  Vs[0][1] * (Vs[1][0] * Vs[2][2] - Vs[1][2] * Vs[2][0]) +
                                  Vs[0][2] * (Vs[1][0] * Vs[2][1] - Vs[1][1] * Vs[2][0]);
                const double sgn = dv >= 0 ? 1.0 : -1.0;
                U[0][2] = sgn * (U[1][0] * U[2][1] - U[2][0] * U[1][1]);
                U[1][2] = sgn * (U[2][0] * U[0][1] - U[0][0] * U[2][1]);
                U[2][2] = sgn * (U[0][0] * U[1][1] - U[1][0] * U[0][1]);
            }
        }
        double R[3][3];
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) R[a][b] = Vs[a][0] * U[b][0] + Vs[a][1] * U[b][1] + Vs[a][2] * U[b][2];
        const double det = R[0][0] * (R[1][1] * R[2][2] - R[1][2] * R[2][1]) - R[0][1] * (R[1][0] * R[2][2] - R[1][2] * R[2][0]) +
                           R[0][2] * (R[1][0] * R[2][1] - R[1][1] * R[2][0]);
        if (det < 0.0)
            for (int b = 0; b < 3; ++b) R[2][b] = -R[2][b];  // Vt[:, -1] *= -1  ==  last row of V U^T negated
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) {
                sR[3 * a + b] = R[a][b];
                Rout[9 * (size_t)s + 3 * a + b] = (float)R[a][b];
            }
        for (int k = 0; k < 3; ++k) tout[3 * (size_t)s + k] = (float)(c[3 + k] - c[k]);
    }
    __syncthreads();
    double e = 0.0;
    for (int i = t; i < n; i += AT) {
        double p[3];
        for (int k = 0; k < 3; ++k) p[k] = (double)P[3 * (size_t)(n0 + i) + k] - c[k];
        for (int a = 0; a < 3; ++a) {  // (p R^T)_a - q_a
            const double v = p[0] * sR[3 * a] + p[1] * sR[3 * a + 1] + p[2] * sR[3 * a + 2] -
                             ((double)Q[3 * (size_t)(n0 + i) + a] - c[3 + a]);
            e += v * v;
        }
    }
    e = block_sum(e, red);
    if (t == 0) rmsd[s] = (float)sqrt(e / (double)max(n, 1));
}

}  // namespace

extern "C" int egnn_rdf_local(const float* pos, const int32_t* node_ptr, int32_t S, const float* edge_lo,
                              const float* edge_hi, const double* r_sq, int32_t nbins, double dR, double vol,
                              const double* weights, int32_t radius, int32_t* counts, double* rdf, float* dists,
                              void* stream) {
    if (S == 0) return EGNN_OK;  // empty batch: nothing to read or write
    EGNN_REQUIRE(pos && node_ptr && edge_lo && edge_hi && r_sq && weights && counts && rdf, "egnn_rdf_local: NULL pointer");
    EGNN_REQUIRE(S >= 0 && nbins >= 1 && nbins <= 4096 && radius >= 0 && radius <= 4096 && dR > 0 && vol > 0,
                 "egnn_rdf_local: bad sizes (S=%d nbins=%d radius=%d)", S, nbins, radius);
    const size_t smem = (size_t)nbins * 12;
    EGNN_SMEM_ATTR(rdf_local_kernel, 4096 * 12);
    rdf_local_kernel<<<S, AT, smem, (cudaStream_t)stream>>>(pos, node_ptr, edge_lo, edge_hi, r_sq, nbins, (float)(1.0 / dR),
                                                            4 * 3.141592653589793, vol, dR, weights, radius, counts, rdf,
                                                            dists);
    EGNN_CHECK_LAUNCH("rdf_local_kernel");
    return EGNN_OK;
}

extern "C" int egnn_cn2_sites(const float* pos, const int32_t* node_ptr, int32_t S, float radius, int32_t* count,
                              int32_t* idx, double* angle_deg, float* bonds, void* stream) {
    if (S == 0) return EGNN_OK;  // empty batch: nothing to read or write
    EGNN_REQUIRE(pos && node_ptr && count && idx && angle_deg && bonds, "egnn_cn2_sites: NULL pointer");
    EGNN_REQUIRE(S >= 0, "egnn_cn2_sites: negative S");
    cn2_kernel<<<S, AT, 0, (cudaStream_t)stream>>>(pos, node_ptr, radius, count, idx, angle_deg, bonds);
    EGNN_CHECK_LAUNCH("cn2_kernel");
    return EGNN_OK;
}

extern "C" int egnn_kabsch(const float* P, const float* Q, const int32_t* node_ptr, int32_t S, float* R, float* t,
                           float* rmsd, void* stream) {
    if (S == 0) return EGNN_OK;  // empty batch: nothing to read or write
    EGNN_REQUIRE(P && Q && node_ptr && R && t && rmsd, "egnn_kabsch: NULL pointer");
    EGNN_REQUIRE(S >= 0, "egnn_kabsch: negative S");
    kabsch_kernel<<<S, AT, 0, (cudaStream_t)stream>>>(P, Q, node_ptr, R, t, rmsd);
    EGNN_CHECK_LAUNCH("kabsch_kernel");
    return EGNN_OK;
}
