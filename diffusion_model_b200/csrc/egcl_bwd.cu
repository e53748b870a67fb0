// Backward of the EGNN denoiser, fp32 (CUDA-core) path.  Per layer, in reverse order:
//
//   egcl_xupdate_bwd  one CTA per graph: gdx = gx'/(S+1);  coefS = gS/S with gS = -sum_i gx'_i.dxu_i/(S+1)^2
//   egcl_node_bwd     32 nodes per tile: back through mlp_h -> gh (node path), gM; wgrads of mlp_h
//   egcl_edge_bwd     TEB edges per tile: recompute both edge MLPs, back through gate / phi / SiLU / Linear,
//                     wgrads of mlp_m, mlp_x, attention; writes per-edge receiver and sender contributions
//   egcl_gather_bwd   per node: gh_in, gx_in = node path + sum over its receiver edges + sum over its sender edges
//   reduce_partials   grad[l] = sum over CTA-private partial rows in CTA order
//
// Only the layer inputs and per-node aggregates are kept from the forward; edge activations are recomputed.
// No atomics: every CTA adds into its own partial row, rows are summed in a fixed order -> bit-identical reruns.
#include <algorithm>

#include "tile_gemm.cuh"
#include "tile_wgrad.cuh"

namespace {

// nodes per tile of the FFMA node backward kernel: 32 where its shared memory fits (every BASELINE width), 16 or 8 for wide
// mlp_h layers (parameters.yaml: h_hidden_size 1024 needs 306 KB at 32 nodes) -- template parameter of egcl_node_bwd
constexpr int TNB_MAX = 32;

// one CTA per graph.  Whole-call scope: every CTA reduces its own graph's share of sum_i gx'_i . dxu_i, the last CTA to
// finish (integer ticket, self-resetting) adds the per-graph shares in graph order -> same bits on every run.
__global__ void __launch_bounds__(256) egcl_xupdate_bwd(const float* __restrict__ gx_out, const float* __restrict__ dxu,
                                                        const float* __restrict__ S, const int32_t* __restrict__ node_ptr,
                                                        int N, int scope, float* __restrict__ gdx,
                                                        float* __restrict__ coefS, float* __restrict__ dotp,
                                                        unsigned int* __restrict__ ticket) {
    __shared__ float red[8];
    __shared__ bool is_last;
    const int g = blockIdx.x, G = gridDim.x;
    const int a = node_ptr[g], b = node_ptr[g + 1];
    float s = 0.f;
    for (int idx = 3 * a + threadIdx.x; idx < 3 * b; idx += 256) s = fmaf(gx_out[idx], dxu[idx], s);
    const float dot = block_sum_256(s, red);
    const float Sg = S[g], den = Sg + 1.f;
    for (int idx = 3 * a + threadIdx.x; idx < 3 * b; idx += 256) gdx[idx] = gx_out[idx] / den;
    if (scope != EGNN_NORM_CALL) {
        if (threadIdx.x == 0) coefS[g] = Sg > 0.f ? -dot / (den * den) / Sg : 0.f;
        return;
    }
    if (threadIdx.x == 0) {
        dotp[g] = dot;
        __threadfence();
        is_last = atomicAdd(ticket, 1u) == (unsigned)G - 1;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    float t = 0.f;
    for (int k = threadIdx.x; k < G; k += 256) t += __ldcg(dotp + k);
    const float tot = block_sum_256(t, red);
    const float c = Sg > 0.f ? -tot / (den * den) / Sg : 0.f;  // S is the same for every graph in this scope
    for (int k = threadIdx.x; k < G; k += 256) coefS[k] = c;
    if (threadIdx.x == 0) *ticket = 0u;
}

struct NodeBwdArgs {
    egnn_net_t net;
    int N, tiles;
    const float *h, *msum;  // layer input h [N,H], combined message sums [N,Ms]
    const float *P, *Pk;
    const float* gh_out;    // [N,H]
    float *gh_node, *gM;    // [N,H], [N,Ms]
    float* part;            // [gridDim.x][layer param count] private rows
};

template <int TNB>
__global__ void __launch_bounds__(256, 2) egcl_node_bwd(NodeBwdArgs a) {
    extern __shared__ __align__(16) float smem[];
    const egnn_net_t& net = a.net;
    const int H = net.H, Ms = net.Ms, Hh = net.Hh, K = H + Ms;
    const int lda = round4(K) + 4, ldz = round4(Hh) + 4, ldh = round4(H) + 4;
    float* A = smem;             // [TNB][lda]  [h, M]
    float* Sz = A + TNB * lda;   // [TNB][ldz]  SiLU(z1)
    float* GZ = Sz + TNB * ldz;  // [TNB][ldz]  dSiLU(z1) -> gz1
    float* GH = GZ + TNB * ldz;  // [TNB][ldh]  gh'
    const LayerOff lo = layer_off(net);
    const PackOff po = pack_off(net);
    float* part = a.part + (size_t)blockIdx.x * lo.total;
    const int tid = threadIdx.x;
    for (int tile = blockIdx.x; tile < a.tiles; tile += gridDim.x) {
        const int n0 = tile * TNB;
        const int rows = min(TNB, a.N - n0);
        __syncthreads();
        for (int idx = tid; idx < TNB * lda; idx += 256) {
            const int r = idx / lda, c = idx - r * lda;
            float v = 0.f;
            if (r < rows) {
                if (c < H) v = a.h[(size_t)(n0 + r) * H + c];
                else if (c < K) v = a.msum[(size_t)(n0 + r) * Ms + (c - H)];
            }
            A[idx] = v;
        }
        for (int idx = tid; idx < TNB * ldh; idx += 256) {
            const int r = idx / ldh, c = idx - r * ldh;
            GH[idx] = (r < rows && c < H) ? a.gh_out[(size_t)(n0 + r) * H + c] : 0.f;
        }
        zero_pad_cols(Sz, ldz, TNB, Hh);
        zero_pad_cols(GZ, ldz, TNB, Hh);
        __syncthreads();
        const float* b1 = a.P + lo.bh1;
        tile_gemm<TNB / 4, 4>(A, lda, a.Pk + po.wh1t, Hh, K, Hh, [&](int r, int c, float acc) {
            const float z = acc + __ldg(b1 + c);
            Sz[r * ldz + c] = silu_f(z);
            GZ[r * ldz + c] = dsilu_f(z);
        });
        __syncthreads();
        // gz1[r][k] = dSiLU * sum_o gh'[r][o] Wh2[o][k]
        const float* W2 = a.P + lo.wh2;
        for (int idx = tid; idx < TNB * Hh; idx += 256) {
            const int r = idx / Hh, k = idx - r * Hh;
            float s = 0.f;
            for (int o = 0; o < H; ++o) s = fmaf(GH[r * ldh + o], __ldg(W2 + o * Hh + k), s);
            GZ[r * ldz + k] = r < rows ? s * GZ[r * ldz + k] : 0.f;
        }
        __syncthreads();
        // weight gradients of mlp_h (rows beyond `rows` are zero in GH / GZ)
        const bool add = tile != (int)blockIdx.x;  // the first tile stores: the mlp_h entries of the row need no zero-fill
        tile_wgrad(GH, ldh, Sz, ldz, rows, H, Hh, part + lo.wh2, Hh, add);
        tile_colsum(GH, ldh, rows, H, nullptr, part + lo.bh2, add);
        tile_wgrad(GZ, ldz, A, lda, rows, Hh, K, part + lo.wh1, K, add);
        tile_colsum(GZ, ldz, rows, Hh, nullptr, part + lo.bh1, add);
        // g[h, M] = gz1 . Wh1
        tile_gemm<TNB / 4, 4>(GZ, ldz, a.P + lo.wh1, K, Hh, K, [&](int r, int c, float acc) {
            if (r < rows) {
                if (c < H) a.gh_node[(size_t)(n0 + r) * H + c] = acc;
                else a.gM[(size_t)(n0 + r) * Ms + (c - H)] = acc;
            }
        });
    }
}

// grad[p] = sum of the partial rows that hold parameter p, in row order.  rows_m > 0: the tensor-core edge backward
// gives rows [0, rows_m) to the mlp_m + attention parameters and rows [rows_m, rowsE) to the mlp_x parameters (other
// entries of a row are never written, so they must not be read); otherwise every edge row holds every edge parameter.
// 64 parameters x 4 row groups per CTA: group g sums rows r0+g, r0+g+4, ... (eight loads in flight), the four group
// sums are then added in group order -- a fixed order, so reruns are bit-identical.
// d = 256 (tc_edge256_bwd.cu): rows [0, nM) / [nM, nM + nX) hold the layer-1 and row-dot parameters of the two branches,
// rows 148 + (2 branch + n / 128) ns + [0, ns) the K splits of row n of [W2 | b2]
struct RowMap256 {
    int on, nM, nX, ns, wm2, wx2, D;
};
__global__ void __launch_bounds__(256) reduce_partials_kernel(const float* __restrict__ part, int rowsE, int rowsN, int n,
                                                              int x_begin, int h_begin, int h_end, int rows_m,
                                                              RowMap256 m256, float* __restrict__ out) {
    __shared__ float red[4][64];
    const int pc = threadIdx.x & 63, g = threadIdx.x >> 6;
    const int p = blockIdx.x * 64 + pc;
    float s = 0.f;
    if (p < n) {
        int r0 = 0, r1 = rowsE;
        if (p >= h_begin && p < h_end) {
            r0 = rowsE;
            r1 = rowsE + rowsN;
        } else if (m256.on) {
            const int branch = p >= x_begin && p < h_begin;
            const int q = p - (branch ? m256.wx2 : m256.wm2);  // offset inside [W2 | b2]
            if (q >= 0 && q < m256.D * m256.D + m256.D) {
                const int row = q < m256.D * m256.D ? q / m256.D : q - m256.D * m256.D;
                r0 = 148 + (2 * branch + (row >> 7)) * m256.ns;
                r1 = r0 + m256.ns;
            } else {
                r0 = branch ? m256.nM : 0;
                r1 = branch ? m256.nM + m256.nX : m256.nM;
            }
        } else if (rows_m > 0) {
            if (p >= x_begin && p < h_begin) r0 = rows_m;
            else r1 = rows_m;
        }
        const float* src = part + p;
        int r = r0 + g;
        for (; r + 28 < r1; r += 32) {
            float v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = src[(size_t)(r + 4 * k) * n];
#pragma unroll
            for (int k = 0; k < 8; ++k) s += v[k];
        }
        for (; r < r1; r += 4) s += src[(size_t)r * n];
    }
    red[g][pc] = s;
    __syncthreads();
    if (g == 0 && p < n) out[p] = ((red[0][pc] + red[1][pc]) + red[2][pc]) + red[3][pc];
}

struct GatherArgs {
    int N, H;
    const int32_t *row_ptr, *col_ptr, *col_edge;
    const float *gh_node, *gx_out, *Gr, *Gs;
    const float *Gr2, *Gs2;  // second per-branch slab of the tensor-core edge backward (nullptr on the FFMA path)
    float *gh_in, *gx_in;
};
__global__ void __launch_bounds__(256) egcl_gather_bwd(GatherArgs a) {
    const int W = a.H + 3;
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    if (idx >= (long long)a.N * W) return;
    const int i = (int)(idx / W), c = (int)(idx - (long long)i * W);
    float s = c < a.H ? a.gh_node[(size_t)i * a.H + c] : a.gx_out[(size_t)i * 3 + (c - a.H)];
    if (a.Gr2) {
        for (int e = a.row_ptr[i]; e < a.row_ptr[i + 1]; ++e) s += a.Gr[(size_t)e * W + c] + a.Gr2[(size_t)e * W + c];
        for (int q = a.col_ptr[i]; q < a.col_ptr[i + 1]; ++q) {
            const size_t o = (size_t)a.col_edge[q] * W + c;
            s += a.Gs[o] + a.Gs2[o];
        }
    } else {
        for (int e = a.row_ptr[i]; e < a.row_ptr[i + 1]; ++e) s += a.Gr[(size_t)e * W + c];
        for (int q = a.col_ptr[i]; q < a.col_ptr[i + 1]; ++q) s += a.Gs[(size_t)a.col_edge[q] * W + c];
    }
    if (c < a.H) a.gh_in[(size_t)i * a.H + c] = s;
    else a.gx_in[(size_t)i * 3 + (c - a.H)] = s;
}


struct EdgeBwdArgs {
    egnn_net_t net;
    int N, E, tiles;
    const int32_t *recv, *send, *node_graph;
    const float *h, *x;      // layer input
    const float *P, *Pk;
    const float *gM;         // [N,Ms]  grad of the message sums
    const float *gdx;        // [N,3]   gx'/(S+1)
    const float *coefS;      // [G]     gS/S
    float *Gr, *Gs;          // [E,H+3] receiver / sender contributions of every edge
    float* part;             // [gridDim.x][layer param count]
};

// One branch (mlp_m or mlp_x) of the edge backward for the current tile.  On entry U holds u_e; on exit GU has been
// (over)written / accumulated with gz1 . W1 and the partial row with this branch's weight gradients.
//   IS_M: message branch (gate + gM), else coordinate branch (phi + gdx.r)
template <int RT, bool IS_M>
__device__ __forceinline__ void edge_branch_bwd(const EdgeBwdArgs& a, const LayerOff& lo, const PackOff& po, int nrows,
                                                const int* sI, const float* sR, float* sPhi, float* sCoef, float* U,
                                                float* GU, int ldu, float* bufA, float* bufD1, float* bufB, float* bufD2,
                                                int ld, float* part) {
    constexpr int TE = 4 * RT;
    const egnn_net_t& net = a.net;
    const int K0 = 2 * net.H + 1;
    const int D1 = IS_M ? net.Mh : net.Xh;  // hidden width
    const int D2 = IS_M ? net.Ms : net.Xh;  // output width of the second Linear
    const int w1 = IS_M ? lo.wm1 : lo.wx1, b1o = IS_M ? lo.bm1 : lo.bx1;
    const int w2 = IS_M ? lo.wm2 : lo.wx2, b2o = IS_M ? lo.bm2 : lo.bx2;
    const int w1t = IS_M ? po.wm1t : po.wx1t, w2t = IS_M ? po.wm2t : po.wx2t;
    const int wv = IS_M ? lo.wa : lo.wx3, bv = IS_M ? lo.ba : lo.bx3;  // the row-dot vector (attention / mlp_x.4)
    const int tid = threadIdx.x;

    zero_pad_cols(bufA, ld, TE, D1);
    zero_pad_cols(bufD1, ld, TE, D1);
    zero_pad_cols(bufB, ld, TE, D2);
    zero_pad_cols(bufD2, ld, TE, D2);
    __syncthreads();
    {   // recompute layer 1: a1 = SiLU(z1), d1 = SiLU'(z1)
        const float* b1 = a.P + b1o;
        tile_gemm<RT, EGNN_EDGE_KU(RT)>(U, ldu, a.Pk + w1t, D1, K0, D1, [&](int r, int c, float acc) {
            const float z = acc + __ldg(b1 + c);
            bufA[r * ld + c] = silu_f(z);
            bufD1[r * ld + c] = dsilu_f(z);
        });
    }
    __syncthreads();
    {   // recompute layer 2: out = SiLU(z2), d2 = SiLU'(z2)
        const float* b2 = a.P + b2o;
        tile_gemm<RT, EGNN_EDGE_KU(RT)>(bufA, ld, a.Pk + w2t, D2, D1, D2, [&](int r, int c, float acc) {
            const float z = acc + __ldg(b2 + c);
            bufB[r * ld + c] = silu_f(z);
            bufD2[r * ld + c] = dsilu_f(z);
        });
    }
    __syncthreads();
    const float* vec = a.P + wv;
    if (IS_M) {
        // gate g = sigmoid(wa.m + ba);  t = gM[i].m;  gs = t g (1-g);  gm = gM[i] g + gs wa;  gz2 = gm * d2
        const int lane = tid & 31, warp = tid >> 5;
        const float ba = __ldg(a.P + bv);
        for (int r = warp; r < TE; r += 8) {
            float g = 0.f, gs = 0.f;
            if (r < nrows) {
                const float* gMi = a.gM + (size_t)sI[r] * net.Ms;
                float s = 0.f, t = 0.f;
                for (int c = lane; c < D2; c += 32) {
                    const float m = bufB[r * ld + c];
                    s = fmaf(__ldg(vec + c), m, s);
                    t = fmaf(__ldg(gMi + c), m, t);
                }
                s = warp_sum(s);
                t = warp_sum(t);
                g = sigmoid_f(s + ba);
                gs = t * g * (1.f - g);
                for (int c = lane; c < D2; c += 32)
                    bufD2[r * ld + c] *= fmaf(__ldg(gMi + c), g, gs * __ldg(vec + c));
            } else {
                for (int c = lane; c < D2; c += 32) bufD2[r * ld + c] = 0.f;
            }
            if (lane == 0) sCoef[r] = gs;
        }
    } else {
        // phi = wx3.q + bx3;  gphi = gdx[i].r;  gz2 = gphi wx3 * d2
        const int lane = tid & 31, warp = tid >> 5;
        const float b3 = __ldg(a.P + bv);
        for (int r = warp; r < TE; r += 8) {
            float gphi = 0.f, phi = 0.f;
            if (r < nrows) {
                float s = 0.f;
                for (int c = lane; c < D2; c += 32) s = fmaf(__ldg(vec + c), bufB[r * ld + c], s);
                phi = warp_sum(s) + b3;
                const float* gd = a.gdx + (size_t)sI[r] * 3;
                gphi = gd[0] * sR[4 * r + 0] + gd[1] * sR[4 * r + 1] + gd[2] * sR[4 * r + 2];
                for (int c = lane; c < D2; c += 32) bufD2[r * ld + c] *= gphi * __ldg(vec + c);
            } else {
                for (int c = lane; c < D2; c += 32) bufD2[r * ld + c] = 0.f;
            }
            if (lane == 0) {
                sCoef[r] = gphi;
                sPhi[r] = phi;
            }
        }
    }
    __syncthreads();
    // gradients of the row-dot vector and its bias: d vec[c] = sum_r coef[r] out[r][c], d bias = sum_r coef[r]
    tile_colsum(bufB, ld, nrows, D2, sCoef, part + wv);
    if (tid == 0) {
        float s = 0.f;
        for (int r = 0; r < nrows; ++r) s += sCoef[r];
        part[bv] += s;
    }
    // second Linear: dW2 += gz2^T a1, db2 += colsum gz2
    tile_wgrad(bufD2, ld, bufA, ld, nrows, D2, D1, part + w2, D1);
    tile_colsum(bufD2, ld, nrows, D2, nullptr, part + b2o);
    // ga1 = gz2 . W2 ; gz1 = ga1 * d1   (in place in bufD1; each element is read and written by the same thread)
    tile_gemm<RT, EGNN_EDGE_KU(RT)>(bufD2, ld, a.P + w2, D1, D2, D1,
                  [&](int r, int c, float acc) { bufD1[r * ld + c] = r < nrows ? acc * bufD1[r * ld + c] : 0.f; });
    __syncthreads();
    // first Linear: dW1 += gz1^T u, db1 += colsum gz1
    tile_wgrad(bufD1, ld, U, ldu, nrows, D1, K0, part + w1, K0);
    tile_colsum(bufD1, ld, nrows, D1, nullptr, part + b1o);
    // gu (+)= gz1 . W1
    tile_gemm<RT, EGNN_EDGE_KU(RT)>(bufD1, ld, a.P + w1, K0, D1, K0, [&](int r, int c, float acc) {
        if (IS_M) GU[r * ldu + c] += acc;
        else GU[r * ldu + c] = acc;
    });
    __syncthreads();
}

template <int RT>
__global__ void __launch_bounds__(256) egcl_edge_bwd(EdgeBwdArgs a) {
    constexpr int TE = 4 * RT;
    extern __shared__ __align__(16) float smem[];
    const egnn_net_t& net = a.net;
    const int H = net.H, K0 = 2 * H + 1;
    const int ldu = round4(K0) + 4;
    int dmax = net.Mh > net.Ms ? net.Mh : net.Ms;
    dmax = dmax > net.Xh ? dmax : net.Xh;
    const int ld = round4(dmax) + 4;
    int* sI = reinterpret_cast<int*>(smem);
    int* sJ = sI + TE;
    float* sR = smem + 2 * TE;
    float* sPhi = sR + 4 * TE;
    float* sCoef = sPhi + TE;
    float* U = sCoef + TE;
    float* GU = U + TE * ldu;
    float* bufA = GU + TE * ldu;
    float* bufD1 = bufA + TE * ld;
    float* bufB = bufD1 + TE * ld;
    float* bufD2 = bufB + TE * ld;
    const LayerOff lo = layer_off(net);
    const PackOff po = pack_off(net);
    float* part = a.part + (size_t)blockIdx.x * lo.total;
    const int tid = threadIdx.x;
    const int W = H + 3;

    for (int tile = blockIdx.x; tile < a.tiles; tile += gridDim.x) {
        const int e0 = tile * TE;
        const int nrows = min(TE, a.E - e0);
        __syncthreads();
        for (int r = tid; r < TE; r += 256) {
            int i = 0, j = 0;
            float rx = 0.f, ry = 0.f, rz = 0.f, dot = 0.f;
            if (r < nrows) {
                i = a.recv[e0 + r];
                j = a.send[e0 + r];
                rx = a.x[3 * i + 0] - a.x[3 * j + 0];
                ry = a.x[3 * i + 1] - a.x[3 * j + 1];
                rz = a.x[3 * i + 2] - a.x[3 * j + 2];
                dot = rx * rx + ry * ry + rz * rz;
            }
            sI[r] = i;
            sJ[r] = j;
            sR[4 * r + 0] = rx;
            sR[4 * r + 1] = ry;
            sR[4 * r + 2] = rz;
            sR[4 * r + 3] = dot;
        }
        __syncthreads();
        for (int idx = tid; idx < TE * ldu; idx += 256) {
            const int r = idx / ldu, c = idx - r * ldu;
            float v = 0.f;
            if (r < nrows) {
                if (c < H) v = a.h[(size_t)sI[r] * H + c];
                else if (c < 2 * H) v = a.h[(size_t)sJ[r] * H + (c - H)];
                else if (c == 2 * H) {
                    const float nrm = sqrtf(sR[4 * r + 3]);
                    v = nrm * nrm;
                }
            }
            U[idx] = v;
            GU[idx] = 0.f;
        }
        __syncthreads();
        edge_branch_bwd<RT, false>(a, lo, po, nrows, sI, sR, sPhi, sCoef, U, GU, ldu, bufA, bufD1, bufB, bufD2, ld, part);
        // keep gphi-independent pieces of the coordinate path before sCoef is reused: gr = gdx[i] phi + coefS r (+ 2 gd r later)
        edge_branch_bwd<RT, true>(a, lo, po, nrows, sI, sR, sPhi, sCoef, U, GU, ldu, bufA, bufD1, bufB, bufD2, ld, part);
        // per-edge outputs
        for (int idx = tid; idx < nrows * W; idx += 256) {
            const int r = idx / W, c = idx - r * W;
            const size_t e = (size_t)(e0 + r);
            if (c < H) {
                a.Gr[e * W + c] = GU[r * ldu + c];
                a.Gs[e * W + c] = GU[r * ldu + H + c];
            } else {
                const int k = c - H;
                const int i = sI[r];
                const float gd = GU[r * ldu + 2 * H];  // grad of d_e = |r_e|^2
                const float rk = sR[4 * r + k];
                const float gr = a.gdx[(size_t)i * 3 + k] * sPhi[r] + (a.coefS[a.node_graph[i]] + 2.f * gd) * rk;
                a.Gr[e * W + c] = gr;
                a.Gs[e * W + c] = -gr;
            }
        }
    }
}

}  // namespace

// ---- host side -------------------------------------------------------------------------------------------
size_t egnn_edge_fwd_smem(const egnn_net_t& net, int TE);
int egnn_validate_net(const egnn_net_t* net);
int egnn_validate_graph(const egnn_graph_t* g, bool need_sender_csr);
int egnn_tc_bwd_grid(int64_t E);
int egnn_tc_bwd256_rows(int64_t E);
int64_t egnn_tc_bwd256_img_bytes(int64_t E);
void egnn_tc_bwd256_split(int64_t E, int* nM, int* nX, int* ns);
int egnn_tc_edge_backward256(const egnn_net_t& net, const egnn_graph_t& g, const float* packed_layer, const float* h,
                             const float* x, const float* gM, const float* gdx, const float* coefS, float* Gr, float* Gs,
                             float* part, uint8_t* img, int precision, cudaStream_t st);
void egnn_tc_bwd_split(int64_t E, int* nM, int* nX);
int egnn_tc_node_bwd_grid(int64_t N);
int egnn_tc_node_backward(const egnn_net_t& net, int N, const float* params_layer, const float* packed_layer, const float* h,
                          const float* msum, const float* gh_out, float* gh_node, float* gM, float* part, int precision,
                          cudaStream_t st);
int egnn_tc_edge_backward(const egnn_net_t& net, const egnn_graph_t& g, const float* packed_layer, const float* h,
                          const float* x, const float* gM, const float* gdx, const float* coefS, float* Gr, float* Gs,
                          float* part, int precision, cudaStream_t st);

// tc_wide.cu: tile-GEMM backward of the edge phase for the networks the fused kernels do not cover
int egnn_wide_bwd_rows(const egnn_net_t& net, int64_t E);
int64_t egnn_wide_bwd_bytes(const egnn_net_t& net, int64_t E);
int egnn_wide_edge_backward(const egnn_net_t& net, const egnn_graph_t& g, const float* params_layer, const float* packed_layer,
                            const float* h, const float* x, const float* gM, const float* gdx, const float* coefS, float* Gr,
                            float* Gs, float* part, float* scratch, int precision, cudaStream_t st);
int egnn_wide_node_bwd_rows(const egnn_net_t& net, int64_t N);
int64_t egnn_wide_node_bwd_bytes(const egnn_net_t& net, int64_t N);
int egnn_wide_node_backward(const egnn_net_t& net, int N, const float* params_layer, const float* packed_layer, const float* h,
                            const float* msum, const float* gh_out, float* gh_node, float* gM, float* part, float* scratch,
                            int precision, cudaStream_t st);

namespace {

constexpr int kNumSM = 148;  // B200

// d = 256 has its own tcgen05 backward (tc_edge256_bwd.cu: two kernels, operand tile images in the workspace)
static inline bool tc_bwd256_supported(const egnn_net_t& n) { return n.Mh == 256 && tc_fwd_supported(n); }

// FFMA node backward: shared memory of one CTA, and its grid = private partial rows (two CTAs per SM when they fit)
inline size_t node_bwd_smem_for(const egnn_net_t& n, int tnb) {
    return sizeof(float) * (size_t)tnb * ((round4(n.H + n.Ms) + 4) + 2 * (round4(n.Hh) + 4) + (round4(n.H) + 4));
}
inline int node_tile_bwd(const egnn_net_t& n) {  // largest node tile whose shared memory fits; 8 even if that does not
    int tnb = TNB_MAX;
    while (tnb > 8 && node_bwd_smem_for(n, tnb) > 200 * 1024) tnb >>= 1;
    return tnb;
}
inline size_t node_bwd_smem(const egnn_net_t& n) { return node_bwd_smem_for(n, node_tile_bwd(n)); }
inline int node_bwd_rows(const egnn_net_t& n, int64_t N) {
    const int tnb = node_tile_bwd(n);
    const int64_t tilesN = (N + tnb - 1) / tnb;
    const int cap = 2 * (node_bwd_smem(n) + 1024) <= 227 * 1024 ? 2 * kNumSM : kNumSM;
    return (int)(tilesN < cap ? tilesN : cap);
}

struct BwdWs {
    int64_t gdx, coefS, dotp, gM, gh_node, Gr, Gs, ghb, gxb, part, img, img_node, total;
    int gridE, gridN, TEB;
};
inline int edge_tile_bwd(const egnn_net_t& n) {
    const int te = edge_tile_for(n) / 2;
    return te < 8 ? 8 : te;
}
inline BwdWs bwd_ws(const egnn_net_t& n, int64_t N, int64_t E, int64_t G) {
    BwdWs w;
    w.TEB = edge_tile_bwd(n);
    const int64_t tilesE = (E + w.TEB - 1) / w.TEB;
    w.gridE = (int)(tilesE < kNumSM ? tilesE : kNumSM);
    if (tc_supported(n) && E > 0 && egnn_tc_bwd_grid(E) > w.gridE) w.gridE = egnn_tc_bwd_grid(E);
    if (tc_bwd256_supported(n) && egnn_tc_bwd256_rows(E) > w.gridE) w.gridE = egnn_tc_bwd256_rows(E);
    if (wide_fwd_supported(n) && egnn_wide_bwd_rows(n, E) > w.gridE) w.gridE = egnn_wide_bwd_rows(n, E);
    w.gridN = node_bwd_rows(n, N);
    if (tc_node_supported(n) && egnn_tc_node_bwd_grid(N) > w.gridN) w.gridN = egnn_tc_node_bwd_grid(N);
    if (wide_node_supported(n) && egnn_wide_node_bwd_rows(n, N) > w.gridN) w.gridN = egnn_wide_node_bwd_rows(n, N);
    const int64_t W = n.H + 3;
    int64_t p = 0;
    auto take = [&](int64_t cnt) { int64_t o = p; p += round_up64(cnt, 4); return o; };
    w.gdx = take(N * 3);
    w.coefS = take(G);
    w.dotp = take(G + 4);  // per-graph shares of the whole-call dot, then the CTA ticket
    w.gM = take(N * n.Ms);
    w.gh_node = take(N * n.H);
    w.Gr = take(2 * E * W);  // two slabs: the tensor-core path writes one per branch
    w.Gs = take(2 * E * W);
    w.ghb = take(2 * N * n.H);
    w.gxb = take(2 * N * 3);
    w.part = take((int64_t)(w.gridE + w.gridN) * layer_off(n).total);
    w.img = take(tc_bwd256_supported(n) ? egnn_tc_bwd256_img_bytes(E) / 4      // 16-byte aligned: every take() is rounded to 4 floats
                 : wide_fwd_supported(n) && E > 0 ? egnn_wide_bwd_bytes(n, E) / 4 : 0);
    w.img_node = take(wide_node_supported(n) ? egnn_wide_node_bwd_bytes(n, N) / 4 : 0);
    w.total = p;
    return w;
}
inline size_t edge_bwd_smem(const egnn_net_t& net, int TE) {
    const int K0 = 2 * net.H + 1;
    int dmax = net.Mh > net.Ms ? net.Mh : net.Ms;
    dmax = dmax > net.Xh ? dmax : net.Xh;
    const size_t ld = round4(dmax) + 4, ldu = round4(K0) + 4;
    return sizeof(float) * ((size_t)TE * 8 + 2 * (size_t)TE * ldu + 4 * (size_t)TE * ld);
}

template <int RT>
int launch_edge_bwd(const EdgeBwdArgs& a, int grid, size_t smem, cudaStream_t st) {
    EGNN_SMEM_ATTR(egcl_edge_bwd<RT>, 200 * 1024);
    egnn_prof_begin(EGNN_K_EDGE_BWD, st);
    egcl_edge_bwd<RT><<<grid, 256, smem, st>>>(a);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("egcl_edge_bwd");
    return EGNN_OK;
}

}  // namespace

extern "C" int64_t egnn_backward_workspace_floats(const egnn_net_t* net, int32_t N, int32_t E, int32_t G) {
    return bwd_ws(*net, N, E, G).total;
}

static int net_backward_impl(const egnn_net_t* net, const egnn_graph_t* g, const float* params, const float* packed,
                             const float* h_in, const float* x_in, const float* fwd_ws_ptr, const float* gh_out,
                             const float* gx_out, float* gh_in, float* gx_in, float* grad_params, float* ws,
                             int64_t ws_floats, int32_t norm_scope, int32_t precision, void* const* layer_events,
                             void* stream);

extern "C" int egnn_net_backward(const egnn_net_t* net, const egnn_graph_t* g, const float* params, const float* packed,
                                 const float* h_in, const float* x_in, const float* fwd_ws_ptr, const float* gh_out,
                                 const float* gx_out, float* gh_in, float* gx_in, float* grad_params, float* ws,
                                 int64_t ws_floats, int32_t norm_scope, int32_t precision, void* stream) {
    return net_backward_impl(net, g, params, packed, h_in, x_in, fwd_ws_ptr, gh_out, gx_out, gh_in, gx_in, grad_params, ws,
                             ws_floats, norm_scope, precision, nullptr, stream);
}

extern "C" int egnn_net_backward_events(const egnn_net_t* net, const egnn_graph_t* g, const float* params,
                                        const float* packed, const float* h_in, const float* x_in, const float* fwd_ws_ptr,
                                        const float* gh_out, const float* gx_out, float* gh_in, float* gx_in,
                                        float* grad_params, float* ws, int64_t ws_floats, int32_t norm_scope,
                                        int32_t precision, void* const* layer_events_host, void* stream) {
    EGNN_REQUIRE(layer_events_host != nullptr, "egnn_net_backward_events: NULL event array (use egnn_net_backward)");
    return net_backward_impl(net, g, params, packed, h_in, x_in, fwd_ws_ptr, gh_out, gx_out, gh_in, gx_in, grad_params, ws,
                             ws_floats, norm_scope, precision, layer_events_host, stream);
}

static int net_backward_impl(const egnn_net_t* net, const egnn_graph_t* g, const float* params, const float* packed,
                             const float* h_in, const float* x_in, const float* fwd_ws_ptr, const float* gh_out,
                             const float* gx_out, float* gh_in, float* gx_in, float* grad_params, float* ws,
                             int64_t ws_floats, int32_t norm_scope, int32_t precision, void* const* layer_events,
                             void* stream) {
    if (int rc = egnn_validate_net(net)) return rc;
    if (int rc = egnn_validate_graph(g, true)) return rc;
    EGNN_REQUIRE(params && packed && h_in && x_in && fwd_ws_ptr && gh_out && gx_out && gh_in && gx_in && grad_params && ws,
                 "egnn_net_backward: NULL tensor pointer");
    EGNN_REQUIRE(norm_scope == EGNN_NORM_CALL || norm_scope == EGNN_NORM_GRAPH, "bad norm_scope %d", norm_scope);
    EGNN_REQUIRE(g->node_graph != nullptr, "egnn_net_backward: node_graph is NULL");
    EGNN_REQUIRE(precision >= EGNN_PREC_FP32 && precision <= EGNN_PREC_BF16, "bad precision %d", precision);
    if (precision != EGNN_PREC_FP32)
        EGNN_REQUIRE(tc_fwd_supported(*net) || wide_fwd_supported(*net),
                     "tcgen05 path needs Mh = Ms = Xh in {64,128,256} and 2H+1 <= 16, or the tile-GEMM limits; use precision fp32");
    // edge phase on tcgen05 (bf16x3 split products or plain bf16): d <= 128 one fused kernel with the weight gradients
    // persistent in TMEM, d = 256 the two-kernel form of tc_edge256_bwd.cu
    const bool use_tc = precision != EGNN_PREC_FP32 && tc_supported(*net);
    const bool use_tc256 = precision != EGNN_PREC_FP32 && tc_bwd256_supported(*net);
    const bool use_wide = precision != EGNN_PREC_FP32 && wide_fwd_supported(*net);
    const FwdWs fw = fwd_ws(*net, g->N, g->E, g->G);
    const BwdWs bw = bwd_ws(*net, g->N, g->E, g->G);
    if (ws_floats < bw.total) {
        egnn_set_error("egnn_net_backward: workspace too small (%lld < %lld floats)", (long long)ws_floats,
                       (long long)bw.total);
        return EGNN_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const LayerOff lo = layer_off(*net);
    const PackOff po = pack_off(*net);
    const int N = g->N, E = g->E, H = net->H;
    const size_t smemE = edge_bwd_smem(*net, bw.TEB);
    const size_t smemN = node_bwd_smem(*net);
    EGNN_REQUIRE(smemE <= 200 * 1024 && smemN <= 200 * 1024, "network too wide for the shared-memory tiling");
    EGNN_SMEM_ATTR(egcl_node_bwd<32>, 200 * 1024);  // per device (the macro keeps its own per-device flag)
    EGNN_SMEM_ATTR(egcl_node_bwd<16>, 200 * 1024);
    EGNN_SMEM_ATTR(egcl_node_bwd<8>, 200 * 1024);
    const bool node_tc = use_tc && tc_node_supported(*net);
    // rows actually used by this call's kernels (the workspace is sized for the larger of the two paths)
    const int rowsE = E > 0 ? (use_tc ? egnn_tc_bwd_grid(E) : use_tc256 ? egnn_tc_bwd256_rows(E) : use_wide ? egnn_wide_bwd_rows(*net, E)
                                                            : (int)std::min<int64_t>((E + bw.TEB - 1) / bw.TEB, kNumSM)) : 0;
    const bool node_wide = precision != EGNN_PREC_FP32 && wide_node_supported(*net);
    const int rowsN = node_tc ? egnn_tc_node_bwd_grid(N) : node_wide ? egnn_wide_node_bwd_rows(*net, N) : node_bwd_rows(*net, N);
    RowMap256 m256{};
    if (use_tc256 && E > 0) {
        m256.on = 1;
        egnn_tc_bwd256_split(E, &m256.nM, &m256.nX, &m256.ns);
        m256.wm2 = lo.wm2, m256.wx2 = lo.wx2, m256.D = net->Mh;
    }
    EGNN_CUDA(cudaMemsetAsync(ws + bw.dotp + g->G, 0, 4 * sizeof(float), st));  // the ticket of egcl_xupdate_bwd
    int rows_m = 0;
    if (use_tc && E > 0) {
        int nX;
        egnn_tc_bwd_split(E, &rows_m, &nX);
    }
    for (int l = net->L - 1; l >= 0; --l) {
        const float* P = params + (size_t)l * lo.total;
        const float* Pk = packed + (size_t)l * po.total;
        const float* hl = l == 0 ? h_in : fwd_ws_ptr + fw.hl + (size_t)l * N * H;
        const float* xl = l == 0 ? x_in : fwd_ws_ptr + fw.xl + (size_t)l * N * 3;
        const float* gho = l == net->L - 1 ? gh_out : ws + bw.ghb + (size_t)((l + 1) & 1) * N * H;
        const float* gxo = l == net->L - 1 ? gx_out : ws + bw.gxb + (size_t)((l + 1) & 1) * N * 3;
        float* ghi = l == 0 ? gh_in : ws + bw.ghb + (size_t)(l & 1) * N * H;
        float* gxi = l == 0 ? gx_in : ws + bw.gxb + (size_t)(l & 1) * N * 3;

        egcl_xupdate_bwd<<<g->G, 256, 0, st>>>(gxo, fwd_ws_ptr + fw.dxu + (size_t)l * N * 3,
                                               fwd_ws_ptr + fw.S + (size_t)l * g->G, g->node_ptr, N, norm_scope,
                                               ws + bw.gdx, ws + bw.coefS, ws + bw.dotp,
                                               reinterpret_cast<unsigned int*>(ws + bw.dotp + g->G));
        EGNN_CHECK_LAUNCH("egcl_xupdate_bwd");
        // the tensor-core kernels overwrite exactly the entries the reduction reads; the FFMA kernels accumulate
        // (the FFMA node kernel stores every mlp_h entry of its row on its first tile: no zero-fill of the node rows)
        if (!use_tc && !use_tc256 && rowsE > 0)
            EGNN_CUDA(cudaMemsetAsync(ws + bw.part, 0, sizeof(float) * (size_t)rowsE * lo.total, st));
        if (node_wide) {
            if (int rc = egnn_wide_node_backward(*net, N, P, Pk, hl, fwd_ws_ptr + fw.msum + (size_t)l * N * net->Ms, gho,
                                                 ws + bw.gh_node, ws + bw.gM, ws + bw.part + (size_t)rowsE * lo.total,
                                                 ws + bw.img_node, precision, st))
                return rc;
        } else if (node_tc) {
            if (int rc = egnn_tc_node_backward(*net, N, P, Pk, hl, fwd_ws_ptr + fw.msum + (size_t)l * N * net->Ms, gho,
                                               ws + bw.gh_node, ws + bw.gM, ws + bw.part + (size_t)rowsE * lo.total,
                                               precision, st))
                return rc;
        } else {
            const int tnb = node_tile_bwd(*net);
            NodeBwdArgs na{*net, N, (N + tnb - 1) / tnb, hl, fwd_ws_ptr + fw.msum + (size_t)l * N * net->Ms, P, Pk, gho,
                           ws + bw.gh_node, ws + bw.gM, ws + bw.part + (size_t)rowsE * lo.total};
            egnn_prof_begin(EGNN_K_NODE_BWD, st);
            if (tnb == 32) egcl_node_bwd<32><<<rowsN, 256, smemN, st>>>(na);
            else if (tnb == 16) egcl_node_bwd<16><<<rowsN, 256, smemN, st>>>(na);
            else egcl_node_bwd<8><<<rowsN, 256, smemN, st>>>(na);
            egnn_prof_end(st);
            EGNN_CHECK_LAUNCH("egcl_node_bwd");
        }
        if (E > 0 && use_tc256) {
            if (int rc = egnn_tc_edge_backward256(*net, *g, Pk, hl, xl, ws + bw.gM, ws + bw.gdx, ws + bw.coefS, ws + bw.Gr,
                                                  ws + bw.Gs, ws + bw.part, reinterpret_cast<uint8_t*>(ws + bw.img),
                                                  precision, st))
                return rc;
        } else if (E > 0 && use_wide) {
            if (int rc = egnn_wide_edge_backward(*net, *g, P, Pk, hl, xl, ws + bw.gM, ws + bw.gdx, ws + bw.coefS, ws + bw.Gr,
                                                 ws + bw.Gs, ws + bw.part, ws + bw.img, precision, st))
                return rc;
        } else if (E > 0 && use_tc) {
            if (int rc = egnn_tc_edge_backward(*net, *g, Pk, hl, xl, ws + bw.gM, ws + bw.gdx, ws + bw.coefS, ws + bw.Gr,
                                               ws + bw.Gs, ws + bw.part, precision, st))
                return rc;
        } else if (E > 0) {
            EdgeBwdArgs ea{*net, N, E, (E + bw.TEB - 1) / bw.TEB, g->recv, g->send, g->node_graph, hl, xl, P, Pk,
                           ws + bw.gM, ws + bw.gdx, ws + bw.coefS, ws + bw.Gr, ws + bw.Gs, ws + bw.part};
            int rc;
            switch (bw.TEB) {
                case 64: rc = launch_edge_bwd<16>(ea, rowsE, smemE, st); break;
                case 32: rc = launch_edge_bwd<8>(ea, rowsE, smemE, st); break;
                case 16: rc = launch_edge_bwd<4>(ea, rowsE, smemE, st); break;
                default: rc = launch_edge_bwd<2>(ea, rowsE, smemE, st); break;
            }
            if (rc) return rc;
        }
        {
            GatherArgs ga{N, H, g->row_ptr, g->col_ptr, g->col_edge, ws + bw.gh_node, gxo, ws + bw.Gr, ws + bw.Gs,
                          (use_tc || use_tc256) ? ws + bw.Gr + (size_t)E * (H + 3) : nullptr,
                          (use_tc || use_tc256) ? ws + bw.Gs + (size_t)E * (H + 3) : nullptr, ghi, gxi};
            const long long tot = (long long)N * (H + 3);
            egnn_prof_begin(EGNN_K_GATHER_BWD, st);
            egcl_gather_bwd<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(ga);
            egnn_prof_end(st);
            EGNN_CHECK_LAUNCH("egcl_gather_bwd");
        }
        reduce_partials_kernel<<<(lo.total + 63) / 64, 256, 0, st>>>(ws + bw.part, rowsE, rowsN, lo.total, lo.wx1, lo.wh1,
                                                                      lo.wa, rows_m, m256,
                                                                      grad_params + (size_t)l * lo.total);
        EGNN_CHECK_LAUNCH("reduce_partials_kernel");
        // layer l's slice of grad_params is final: a communication stream waiting on this event can all-reduce it
        // while the layers below are still in backward
        if (layer_events && layer_events[l]) EGNN_CUDA(cudaEventRecord((cudaEvent_t)layer_events[l], st));
    }
    return EGNN_OK;
}
