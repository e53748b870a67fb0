// Forward of the EGNN denoiser, fp32 (CUDA-core) path.  Three kernels per EGCL layer:
//
//   egcl_edge_fwd   one CTA per tile of TE receiver-sorted edges: gather u_e = [h_i, h_j, |x_i-x_j|^2], both edge MLPs
//                   (mlp_x then mlp_m + gate) in shared memory, then a fixed-order segmented sum per receiver of
//                   [ g_e*m_e (Ms) | r_e*phi_e (3) | r_e.r_e (1) ].  Complete segments go to agg[i], the (at most two)
//                   segments cut by the tile boundary go to part[tile][0|1].       (EquivariantGraphNeuralNetwork.py:55-65)
//   egcl_node_fwd   32 nodes per CTA: combine agg/partials in tile order, h' = mlp_h([h, M])   (:69, no residual)
//   egcl_xupdate    one CTA per graph: S = sqrt(sum D_i) over the norm scope, x' = x + dxu/(S+1)    (:64, :70)
//
// No atomics anywhere: reruns are bit-identical.  Within a tile edges are summed in ascending edge id, the order the
// reference's CPU scatter_add_ uses.
#include "tile_gemm.cuh"
#ifndef EGNN_NODE_FWD_KU
#define EGNN_NODE_FWD_KU 2
#endif

namespace {

struct EdgeFwdArgs {
    egnn_net_t net;
    int N, E;
    const int32_t *recv, *send, *row_ptr;
    const float *h, *x;  // layer input [N,H], [N,3]
    const float* P;      // this layer's parameters (flat layout)
    const float* Pk;     // this layer's packed transposes
    float *agg, *part;
};

__global__ void __launch_bounds__(256) pack_weights_kernel(egnn_net_t net, const float* __restrict__ params,
                                                           float* __restrict__ packed) {
    const LayerOff lo = layer_off(net);
    const PackOff po = pack_off(net);
    const int K0 = 2 * net.H + 1;
    const float* P = params + (size_t)blockIdx.y * lo.total;
    float* Q = packed + (size_t)blockIdx.y * po.total;
    const int stride = gridDim.x * blockDim.x;
    const int t0 = blockIdx.x * blockDim.x + threadIdx.x;
    auto tr = [&](int src, int dst, int rows /*n*/, int cols /*k*/) {
        for (int idx = t0; idx < rows * cols; idx += stride) {
            const int k = idx / rows, n = idx % rows;  // write coalesced: dst[k][n]
            Q[dst + idx] = P[src + n * cols + k];
        }
    };
    tr(lo.wm1, po.wm1t, net.Mh, K0);
    tr(lo.wm2, po.wm2t, net.Ms, net.Mh);
    tr(lo.wx1, po.wx1t, net.Xh, K0);
    tr(lo.wx2, po.wx2t, net.Xh, net.Xh);
    tr(lo.wh1, po.wh1t, net.Hh, net.H + net.Ms);
}

template <int RT>
__global__ void __launch_bounds__(256) egcl_edge_fwd(EdgeFwdArgs a) {
    constexpr int TE = 4 * RT;
    extern __shared__ __align__(16) float smem[];
    const egnn_net_t& net = a.net;
    const int H = net.H, K0 = 2 * H + 1, Mh = net.Mh, Ms = net.Ms, Xh = net.Xh;
    const int ldu = round4(K0) + 4;
    int dmax = Mh > Ms ? Mh : Ms;
    dmax = dmax > Xh ? dmax : Xh;
    const int ld = round4(dmax) + 4;
    int* sI = reinterpret_cast<int*>(smem);
    int* sJ = sI + TE;
    float* sR = smem + 2 * TE;  // [TE][4] = r_x, r_y, r_z, r.r
    float* sPhi = sR + 4 * TE;
    float* sGate = sPhi + TE;
    float* U = sGate + TE;
    float* buf1 = U + TE * ldu;
    float* buf2 = buf1 + TE * ld;

    const LayerOff lo = layer_off(net);
    const PackOff po = pack_off(net);
    const int tid = threadIdx.x;
    const int e0 = blockIdx.x * TE;
    const int nrows = min(TE, a.E - e0);

    // ---- gather
    for (int r = tid; r < TE; r += 256) {
        int i = -1, j = -1;
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
                const float nrm = sqrtf(sR[4 * r + 3]);  // :56  torch.norm(...)**2 -- sqrt, then square
                v = nrm * nrm;
            }
        }
        U[idx] = v;
    }
    zero_pad_cols(buf1, ld, TE, Xh);
    __syncthreads();

    // ---- coordinate branch: phi_e = wx3 . SiLU(Wx2 SiLU(Wx1 u + b) + b) + b        (:19-25, :63-64)
    {
        const float* b1 = a.P + lo.bx1;
        tile_gemm<RT, EGNN_EDGE_KU(RT)>(U, ldu, a.Pk + po.wx1t, Xh, K0, Xh,
                      [&](int r, int c, float acc) { buf1[r * ld + c] = silu_f(acc + __ldg(b1 + c)); });
        __syncthreads();
        const float* b2 = a.P + lo.bx2;
        tile_gemm<RT, EGNN_EDGE_KU(RT)>(buf1, ld, a.Pk + po.wx2t, Xh, Xh, Xh,
                      [&](int r, int c, float acc) { buf2[r * ld + c] = silu_f(acc + __ldg(b2 + c)); });
        __syncthreads();
        row_dots(buf2, ld, TE, Xh, a.P + lo.wx3, __ldg(a.P + lo.bx3), [&](int r, float v) { sPhi[r] = v; });
        zero_pad_cols(buf1, ld, TE, Mh);
        __syncthreads();
    }
    // ---- message branch: m_e = SiLU(Wm2 SiLU(Wm1 u + b) + b),  gate g_e = sigmoid(wa . m_e + ba)   (:13-18, :31-34, :57-61)
    {
        const float* b1 = a.P + lo.bm1;
        tile_gemm<RT, EGNN_EDGE_KU(RT)>(U, ldu, a.Pk + po.wm1t, Mh, K0, Mh,
                      [&](int r, int c, float acc) { buf1[r * ld + c] = silu_f(acc + __ldg(b1 + c)); });
        __syncthreads();
        const float* b2 = a.P + lo.bm2;
        tile_gemm<RT, EGNN_EDGE_KU(RT)>(buf1, ld, a.Pk + po.wm2t, Ms, Mh, Ms,
                      [&](int r, int c, float acc) { buf2[r * ld + c] = silu_f(acc + __ldg(b2 + c)); });
        __syncthreads();
        row_dots(buf2, ld, TE, Ms, a.P + lo.wa, __ldg(a.P + lo.ba), [&](int r, float v) { sGate[r] = sigmoid_f(v); });
        __syncthreads();
    }
    // ---- fixed-order segmented sum over receivers
    const int AGGW = Ms + 4;
    const int e_end = e0 + TE;
    float* part = a.part + (size_t)blockIdx.x * 2 * AGGW;
    for (int c = tid; c < AGGW; c += 256) {
        float sum = 0.f;
        int cur = sI[0];
        for (int r = 0; r <= nrows; ++r) {
            const int i = r < nrows ? sI[r] : -2;
            if (i != cur) {
                const int ra = a.row_ptr[cur], rb = a.row_ptr[cur + 1];
                if (ra < e0) part[c] = sum;
                else if (rb > e_end) part[AGGW + c] = sum;
                else a.agg[(size_t)cur * AGGW + c] = sum;
                sum = 0.f;
                cur = i;
            }
            if (r < nrows) {
                float v;
                if (c < Ms) v = sGate[r] * buf2[r * ld + c];
                else if (c < Ms + 3) v = sR[4 * r + (c - Ms)] * sPhi[r];
                else v = sR[4 * r + 3];
                sum += v;
            }
        }
    }
}

struct NodeFwdArgs {
    egnn_net_t net;
    int N, TE;
    const int32_t* row_ptr;
    const float* h;  // layer input
    const float *P, *Pk;
    const float *agg, *part;
    float *h_out, *msum, *dxu, *D;
    // tcgen05 path: the edge kernel leaves phi_e per edge instead of the 4 extra aggregate columns; dxu / D are then
    // summed here over the receiver's edges (8 lanes per node, fixed order)
    const float* phi_edge;
    const int32_t* send;
    const float* x;
};

// combined per-receiver sum of column c (see header comment)
__device__ __forceinline__ float combine_agg(const float* __restrict__ agg, const float* __restrict__ part,
                                             const int32_t* __restrict__ row_ptr, int i, int c, int AGGW, int TE) {
    const int ra = row_ptr[i], rb = row_ptr[i + 1];
    if (ra == rb) return 0.f;
    const int t0 = ra / TE, t1 = (rb - 1) / TE;
    if (t0 == t1) return agg[(size_t)i * AGGW + c];
    // the node is cut by tile boundaries: tail partial of t0 (or head, when it starts exactly at the tile start
    // it still spans past the end, which the edge kernel files under "tail"), then heads of t0+1..t1
    float s = part[((size_t)t0 * 2 + 1) * AGGW + c];
    for (int t = t0 + 1; t <= t1; ++t) s += part[((size_t)t * 2) * AGGW + c];
    return s;
}

constexpr int TN = 32;  // nodes per CTA of the node kernel (RT = 8)

__global__ void __launch_bounds__(256) egcl_node_fwd(NodeFwdArgs a) {
    extern __shared__ __align__(16) float smem[];
    const egnn_net_t& net = a.net;
    const int H = net.H, Ms = net.Ms, Hh = net.Hh, K = H + Ms, AGGW = Ms + 4;
    const int lda = round4(K) + 4, ldz = round4(Hh) + 4;
    float* A = smem;
    float* Z = A + TN * lda;
    const LayerOff lo = layer_off(net);
    const PackOff po = pack_off(net);
    const int tid = threadIdx.x;
    const int n0 = blockIdx.x * TN;
    for (int idx = tid; idx < TN * lda; idx += 256) {
        const int r = idx / lda, c = idx - r * lda;
        const int i = n0 + r;
        float v = 0.f;
        if (i < a.N) {
            if (c < H) v = a.h[(size_t)i * H + c];
            else if (c < K) {
                v = combine_agg(a.agg, a.part, a.row_ptr, i, c - H, AGGW, a.TE);
                a.msum[(size_t)i * Ms + (c - H)] = v;
            }
        }
        A[idx] = v;
    }
    if (a.phi_edge == nullptr) {
        for (int idx = tid; idx < TN * 4; idx += 256) {
            const int r = idx >> 2, c = idx & 3;
            const int i = n0 + r;
            if (i < a.N) {
                const float v = combine_agg(a.agg, a.part, a.row_ptr, i, Ms + c, AGGW, a.TE);
                if (c < 3) a.dxu[(size_t)i * 3 + c] = v;
                else a.D[i] = v;
            }
        }
    } else {
        const int r = tid >> 3, sub = tid & 7;  // 8 lanes per node
        const int i = n0 + r;
        float sx = 0.f, sy = 0.f, sz = 0.f, sd = 0.f;
        if (i < a.N) {
            const float xi = a.x[3 * i], yi = a.x[3 * i + 1], zi = a.x[3 * i + 2];
            for (int e = a.row_ptr[i] + sub; e < a.row_ptr[i + 1]; e += 8) {
                const int j = a.send[e];
                const float rx = xi - a.x[3 * j], ry = yi - a.x[3 * j + 1], rz = zi - a.x[3 * j + 2];
                const float ph = a.phi_edge[e];
                sx = fmaf(rx, ph, sx);
                sy = fmaf(ry, ph, sy);
                sz = fmaf(rz, ph, sz);
                sd += rx * rx + ry * ry + rz * rz;
            }
        }
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) {
            sx += __shfl_xor_sync(0xffffffffu, sx, o);
            sy += __shfl_xor_sync(0xffffffffu, sy, o);
            sz += __shfl_xor_sync(0xffffffffu, sz, o);
            sd += __shfl_xor_sync(0xffffffffu, sd, o);
        }
        if (i < a.N && sub == 0) {
            a.dxu[(size_t)i * 3 + 0] = sx;
            a.dxu[(size_t)i * 3 + 1] = sy;
            a.dxu[(size_t)i * 3 + 2] = sz;
            a.D[i] = sd;
        }
    }
    __syncthreads();
    const float* b1 = a.P + lo.bh1;
    tile_gemm<TN / 4, EGNN_NODE_FWD_KU>(A, lda, a.Pk + po.wh1t, Hh, K, Hh,
                      [&](int r, int c, float acc) { Z[r * ldz + c] = silu_f(acc + __ldg(b1 + c)); });
    __syncthreads();
    const float* W2 = a.P + lo.wh2;
    const float* b2 = a.P + lo.bh2;
    // one warp per (node, output) pair, lanes stride over k, fixed shuffle order
    for (int idx = tid >> 5; idx < TN * H; idx += 8) {
        const int r = idx / H, o = idx - r * H;
        const int i = n0 + r;
        if (i >= a.N) continue;
        float s = 0.f;
        for (int k = tid & 31; k < Hh; k += 32) s = fmaf(__ldg(W2 + o * Hh + k), Z[r * ldz + k], s);
        s = warp_sum(s);
        if ((tid & 31) == 0) a.h_out[(size_t)i * H + o] = s + __ldg(b2 + o);
    }
}

// one CTA per graph.  scope: EGNN_NORM_CALL -> every CTA reduces D over all N nodes in the same fixed order
__global__ void __launch_bounds__(256) egcl_xupdate(const float* __restrict__ x, const float* __restrict__ dxu,
                                                    const float* __restrict__ D, const int32_t* __restrict__ node_ptr,
                                                    int N, int scope, float* __restrict__ x_out, float* __restrict__ S_out) {
    __shared__ float red[8];
    pdl_launch_dependents();
    pdl_wait();  // common.cuh: PDL -- D / dxu come from the node kernel
    const int g = blockIdx.x;
    const int a = node_ptr[g], b = node_ptr[g + 1];
    const int lo = scope == EGNN_NORM_CALL ? 0 : a, hi = scope == EGNN_NORM_CALL ? N : b;
    float s = 0.f;
    for (int i = lo + threadIdx.x; i < hi; i += 256) s += D[i];
    const float S = sqrtf(block_sum_256(s, red));
    if (threadIdx.x == 0) S_out[g] = S;
    const float den = S + 1.f;
    for (int idx = 3 * a + threadIdx.x; idx < 3 * b; idx += 256) x_out[idx] = x[idx] + dxu[idx] / den;
}

template <int RT>
int launch_edge_fwd(const EdgeFwdArgs& a, size_t smem, cudaStream_t st) {
    EGNN_SMEM_ATTR(egcl_edge_fwd<RT>, 200 * 1024);
    const int TE = 4 * RT;
    const int tiles = (a.E + TE - 1) / TE;
    egnn_prof_begin(EGNN_K_EDGE_FWD, st);
    egcl_edge_fwd<RT><<<tiles, 256, smem, st>>>(a);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("egcl_edge_fwd");
    return EGNN_OK;
}

}  // namespace

size_t egnn_edge_fwd_smem(const egnn_net_t& net, int TE) {
    const int K0 = 2 * net.H + 1;
    int dmax = net.Mh > net.Ms ? net.Mh : net.Ms;
    dmax = dmax > net.Xh ? dmax : net.Xh;
    const size_t ld = round4(dmax) + 4, ldu = round4(K0) + 4;
    return sizeof(float) * ((size_t)TE * (2 + 4 + 1 + 1) + (size_t)TE * ldu + 2 * (size_t)TE * ld);
}

int egnn_validate_net(const egnn_net_t* net) {
    EGNN_REQUIRE(net != nullptr, "net is NULL");
    EGNN_REQUIRE(net->L >= 1 && net->H >= 1 && net->Mh >= 1 && net->Ms >= 1 && net->Xh >= 1 && net->Hh >= 1,
                 "network sizes must be positive (L=%d H=%d Mh=%d Ms=%d Xh=%d Hh=%d)", net->L, net->H, net->Mh, net->Ms,
                 net->Xh, net->Hh);
    EGNN_REQUIRE(net->Mh <= 4096 && net->Ms <= 4096 && net->Xh <= 4096 && net->Hh <= 4096 && net->H <= 1024,
                 "network sizes too large for the shared-memory tiling");
    return EGNN_OK;
}

extern "C" int64_t egnn_param_count(const egnn_net_t* net) { return (int64_t)net->L * layer_off(*net).total; }
extern "C" int64_t egnn_packed_count(const egnn_net_t* net) { return (int64_t)net->L * pack_off(*net).total; }
extern "C" int32_t egnn_edge_tile(const egnn_net_t* net) { return edge_tile_for(*net); }
extern "C" int64_t egnn_forward_workspace_floats(const egnn_net_t* net, int32_t N, int32_t E, int32_t G) {
    return fwd_ws(*net, N, E, G).total;
}

int egnn_tc_pack(const egnn_net_t& net, const float* params, float* packed, cudaStream_t st);
int egnn_tc_edge_forward(const egnn_net_t& net, const egnn_graph_t& g, const float* packed_layer, const float* h,
                         const float* x, float* agg, float* part, float* phi_edge, int precision, cudaStream_t st, bool pdl);

int egnn_tc_node_forward(const egnn_net_t& net, const egnn_graph_t& g, const float* params_layer, const float* packed_layer,
                         const float* h, const float* x, const float* agg, const float* part, const float* phi_edge,
                         float* h_out, float* msum, float* dxu, float* Dn, int precision, cudaStream_t st);
// tc_wide.cu: the generic tcgen05 forward of the networks the fused kernels do not cover
int egnn_wide_pack(const egnn_net_t& net, const float* params, float* packed, cudaStream_t st);
int egnn_wide_edge_forward(const egnn_net_t& net, const egnn_graph_t& g, const float* params_layer, const float* packed_layer,
                           const float* h, const float* x, float* agg, float* phi_edge, float* scratch, int precision,
                           cudaStream_t st);
int egnn_wide_node_forward(const egnn_net_t& net, const egnn_graph_t& g, const float* params_layer, const float* packed_layer,
                           const float* h, const float* x, const float* agg, const float* part, int TE, const float* phi_edge,
                           float* h_out, float* msum, float* dxu, float* Dn, float* scratch, int precision, cudaStream_t st);

extern "C" int egnn_pack_weights(const egnn_net_t* net, const float* params, float* packed, void* stream) {
    if (int rc = egnn_validate_net(net)) return rc;
    pack_weights_kernel<<<dim3(32, net->L), 256, 0, (cudaStream_t)stream>>>(*net, params, packed);
    EGNN_CHECK_LAUNCH("pack_weights_kernel");
    if (int rc = egnn_wide_pack(*net, params, packed, (cudaStream_t)stream)) return rc;
    return egnn_tc_pack(*net, params, packed, (cudaStream_t)stream);
}

// one EGCL layer forward (fp32 path); shared with the backward's recompute-free bookkeeping
int egnn_layer_forward_f32(const egnn_net_t& net, const egnn_graph_t& g, const float* P, const float* Pk, const float* h,
                           const float* x, float* h_out, float* x_out, float* ws, const FwdWs& w, int l, int norm_scope,
                           int precision, cudaStream_t st) {
    const bool tensor = precision != EGNN_PREC_FP32;
    const bool wide = tensor && wide_fwd_supported(net);
    // wide path: agg holds complete per-receiver sums (no tile partials): a tile size no edge range can straddle
    const int TE = wide ? 1 << 30 : tensor ? part_tile_for(net) : edge_tile_for(net);
    if (g.E > 0 && wide) {
        if (int rc = egnn_wide_edge_forward(net, g, P, Pk, h, x, ws + w.agg, ws + w.phi, ws + w.wide, precision, st)) return rc;
    } else if (g.E > 0 && tensor) {
        // layer 0 is launched normally: its prologue reads the packed weights, which the pack kernels right in front of a
        // training forward write; every later kernel of the chain may overlap its predecessor's tail (common.cuh: PDL)
        if (int rc = egnn_tc_edge_forward(net, g, Pk, h, x, ws + w.agg, ws + w.part, ws + w.phi, precision, st, l > 0)) return rc;
    } else if (g.E > 0) {
        EdgeFwdArgs ea{net, g.N, g.E, g.recv, g.send, g.row_ptr, h, x, P, Pk, ws + w.agg, ws + w.part};
        const size_t smem = egnn_edge_fwd_smem(net, TE);
        int rc;
        switch (TE) {
            case 128: rc = launch_edge_fwd<32>(ea, smem, st); break;
            case 64: rc = launch_edge_fwd<16>(ea, smem, st); break;
            case 32: rc = launch_edge_fwd<8>(ea, smem, st); break;
            default: rc = launch_edge_fwd<4>(ea, smem, st); break;
        }
        if (rc) return rc;
    }
    if (tensor && g.E > 0 && wide_node_supported(net)) {
        if (int rc = egnn_wide_node_forward(net, g, P, Pk, h, x, ws + w.agg, ws + w.part, TE, ws + w.phi, h_out, ws + w.msum + (size_t)l * g.N * net.Ms,
                                            ws + w.dxu + (size_t)l * g.N * 3, ws + w.D, ws + w.wide, precision, st))
            return rc;
    } else if (tensor && tc_node_supported(net)) {
        if (int rc = egnn_tc_node_forward(net, g, P, Pk, h, x, ws + w.agg, ws + w.part, ws + w.phi, h_out,
                                          ws + w.msum + (size_t)l * g.N * net.Ms, ws + w.dxu + (size_t)l * g.N * 3, ws + w.D,
                                          precision, st))
            return rc;
    } else {
        NodeFwdArgs na{net, g.N, TE, g.row_ptr, h, P, Pk, ws + w.agg, ws + w.part, h_out,
                       ws + w.msum + (size_t)l * g.N * net.Ms, ws + w.dxu + (size_t)l * g.N * 3, ws + w.D,
                       tensor ? ws + w.phi : nullptr, g.send, x};
        const size_t smem = sizeof(float) * (size_t)TN * ((round4(net.H + net.Ms) + 4) + (round4(net.Hh) + 4));
        EGNN_SMEM_ATTR(egcl_node_fwd, 200 * 1024);
        egnn_prof_begin(EGNN_K_NODE_FWD, st);
        egcl_node_fwd<<<(g.N + TN - 1) / TN, 256, smem, st>>>(na);
        egnn_prof_end(st);
        EGNN_CHECK_LAUNCH("egcl_node_fwd");
    }
    egnn_prof_begin(EGNN_K_XUPDATE, st);
    EGNN_CUDA(egnn_launch(egcl_xupdate, dim3(g.G), dim3(256), 0, st, true, x, (const float*)(ws + w.dxu + (size_t)l * g.N * 3),
                          (const float*)(ws + w.D), g.node_ptr, g.N, norm_scope, x_out, ws + w.S + (size_t)l * g.G));
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("egcl_xupdate");
    return EGNN_OK;
}

int egnn_validate_graph(const egnn_graph_t* g, bool need_sender_csr) {
    EGNN_REQUIRE(g != nullptr, "graph is NULL");
    EGNN_REQUIRE(g->N >= 1 && g->E >= 0 && g->G >= 1, "graph sizes out of range (N=%d E=%d G=%d)", g->N, g->E, g->G);
    EGNN_REQUIRE(g->row_ptr && g->node_ptr && (g->E == 0 || (g->recv && g->send)), "graph index pointers are NULL");
    if (need_sender_csr) EGNN_REQUIRE(g->col_ptr && (g->E == 0 || g->col_edge), "sender CSR pointers are NULL");
    return EGNN_OK;
}

extern "C" int egnn_net_forward(const egnn_net_t* net, const egnn_graph_t* g, const float* params, const float* packed,
                                const float* h_in, const float* x_in, float* h_out, float* x_out, float* ws,
                                int64_t ws_floats, int32_t norm_scope, int32_t precision, void* stream) {
    if (int rc = egnn_validate_net(net)) return rc;
    if (int rc = egnn_validate_graph(g, false)) return rc;
    EGNN_REQUIRE(params && packed && h_in && x_in && h_out && x_out && ws, "egnn_net_forward: NULL tensor pointer");
    EGNN_REQUIRE(norm_scope == EGNN_NORM_CALL || norm_scope == EGNN_NORM_GRAPH, "bad norm_scope %d", norm_scope);
    EGNN_REQUIRE(precision == EGNN_PREC_FP32 || precision == EGNN_PREC_BF16X3 || precision == EGNN_PREC_BF16,
                 "bad precision %d", precision);
    EGNN_REQUIRE(precision == EGNN_PREC_FP32 || tc_fwd_supported(*net) || wide_fwd_supported(*net),
                 "tcgen05 precisions need Mh = Ms = Xh in {64,128,256} and 2H+1 <= 16 (fused kernels) or Mh, Xh <= 1024, Ms <= 256, "
                 "2H+1 <= 128 (tile-GEMM path); use precision fp32 for this network");
    const FwdWs w = fwd_ws(*net, g->N, g->E, g->G);
    if (ws_floats < w.total) {
        egnn_set_error("egnn_net_forward: workspace too small (%lld < %lld floats)", (long long)ws_floats,
                       (long long)w.total);
        return EGNN_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const LayerOff lo = layer_off(*net);
    const PackOff po = pack_off(*net);
    const float* h = h_in;
    const float* x = x_in;
    for (int l = 0; l < net->L; ++l) {
        float* ho = l == net->L - 1 ? h_out : ws + w.hl + (size_t)(l + 1) * g->N * net->H;
        float* xo = l == net->L - 1 ? x_out : ws + w.xl + (size_t)(l + 1) * g->N * 3;
        if (int rc = egnn_layer_forward_f32(*net, *g, params + (size_t)l * lo.total, packed + (size_t)l * po.total, h, x,
                                            ho, xo, ws, w, l, norm_scope, precision, st))
            return rc;
        h = ho;
        x = xo;
    }
    return EGNN_OK;
}
