// Shared helpers for libegnn_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/egnn_b200.h"

void egnn_set_error(const char* fmt, ...);
void egnn_count_launch();
bool egnn_prof_on();
void egnn_prof_begin(int kind, cudaStream_t st);
void egnn_prof_end(cudaStream_t st);
// kernel families the profiler can bracket (egnn_profile_read)
enum { EGNN_K_EDGE_FWD = 0, EGNN_K_NODE_FWD = 1, EGNN_K_XUPDATE = 2, EGNN_K_EDGE_BWD = 3, EGNN_K_NODE_BWD = 4,
       EGNN_K_GATHER_BWD = 5, EGNN_K_REDUCE = 6, EGNN_K_TC_EDGE_FWD = 7, EGNN_K_TC_EDGE_BWD = 8, EGNN_K_TC_WGRAD = 9 };

#define EGNN_REQUIRE(cond, ...)              \
    do {                                     \
        if (!(cond)) {                       \
            egnn_set_error(__VA_ARGS__);     \
            return EGNN_ERR_ARG;             \
        }                                    \
    } while (0)

#define EGNN_CHECK_LAUNCH(name)                                       \
    do {                                                              \
        cudaError_t e__ = cudaGetLastError();                         \
        egnn_count_launch();                                          \
        if (e__ != cudaSuccess) {                                     \
            egnn_set_error("%s: %s", name, cudaGetErrorString(e__));  \
            return EGNN_ERR_CUDA;                                     \
        }                                                             \
    } while (0)

#define EGNN_CUDA(call)                                                   \
    do {                                                                  \
        cudaError_t e__ = (call);                                         \
        if (e__ != cudaSuccess) {                                         \
            egnn_set_error("%s: %s", #call, cudaGetErrorString(e__));     \
            return EGNN_ERR_CUDA;                                         \
        }                                                                 \
    } while (0)

// cudaFuncAttributeMaxDynamicSharedMemorySize is a PER-DEVICE setting: remember it per device ordinal, not per process
// (a process driving a second GPU would otherwise launch kernels needing > 48 KB of shared memory unconfigured).
#define EGNN_SMEM_ATTR(kernel, bytes)                                                                          \
    do {                                                                                                       \
        static bool done__[64] = {};                                                                           \
        int dev__ = 0;                                                                                         \
        EGNN_CUDA(cudaGetDevice(&dev__));                                                                      \
        if (dev__ < 0 || dev__ >= 64 || !done__[dev__]) {                                                      \
            EGNN_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes))); \
            if (dev__ >= 0 && dev__ < 64) done__[dev__] = true;                                                \
        }                                                                                                      \
    } while (0)
// Grids are sized for the 148 SMs of a B200 (one persistent CTA per SM).  This is a tuning constant, not a correctness
// requirement: no kernel synchronises across CTAs, so a device with another SM count runs the same grids correctly.
constexpr int kB200SMs = 148;

// ---- programmatic dependent launch (PDL) --------------------------------------------------------------------------------
// The forward path of a small graph is a chain of ~3 L short kernels (a 24-atom structure is 5 edge tiles): launch latency
// and each kernel's prologue (weight-image load, TMEM allocation, barrier set-up) cost more than the math.  A kernel
// launched with egnn_launch(..., pdl = true) may START while its predecessor in the stream is still running; it does its
// prologue -- which must touch only memory no kernel of the chain writes (packed weights, the graph index) -- and then
// calls pdl_wait(), which returns once the predecessor has completed and its writes are visible.  Every kernel of the
// chain calls pdl_launch_dependents() at its top so that its successor can be scheduled early.  A kernel that was
// launched normally sees pdl_wait() return at once.  Works under CUDA-graph capture (programmatic edges).
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <typename... KArgs, typename... Args>
static inline cudaError_t egnn_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, bool pdl,
                                      Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif

static inline int64_t round_up64(int64_t a, int64_t b) { return (a + b - 1) / b * b; }
__host__ __device__ static inline int round4(int a) { return (a + 3) & ~3; }

// ---- flat parameter layout (reference named_parameters() order, see include/egnn_b200.h) --------------------
struct LayerOff {
    int wm1, bm1, wm2, bm2, wx1, bx1, wx2, bx2, wx3, bx3, wh1, bh1, wh2, bh2, wa, ba, total;
};
__host__ __device__ static inline LayerOff layer_off(const egnn_net_t& n) {
    LayerOff o;
    const int K0 = 2 * n.H + 1;
    int p = 0;
    o.wm1 = p; p += n.Mh * K0;
    o.bm1 = p; p += n.Mh;
    o.wm2 = p; p += n.Ms * n.Mh;
    o.bm2 = p; p += n.Ms;
    o.wx1 = p; p += n.Xh * K0;
    o.bx1 = p; p += n.Xh;
    o.wx2 = p; p += n.Xh * n.Xh;
    o.bx2 = p; p += n.Xh;
    o.wx3 = p; p += n.Xh;
    o.bx3 = p; p += 1;
    o.wh1 = p; p += n.Hh * (n.H + n.Ms);
    o.bh1 = p; p += n.Hh;
    o.wh2 = p; p += n.H * n.Hh;
    o.bh2 = p; p += n.H;
    o.wa = p; p += n.Ms;
    o.ba = p; p += 1;
    o.total = p;
    return o;
}
// packed = K-major transposes used as the B operand of the forward FFMA GEMMs: Wt[k][n] = W[n][k]
struct PackOff {
    int wm1t, wm2t, wx1t, wx2t, wh1t, tc, wide, total;  // wide: the tile-GEMM weight images (tc_wide.cu), 16-byte aligned
};
// the tcgen05 edge kernels cover the uniform-width nets of BASELINE.json: Mh = Ms = Xh = d in {64,128}, 2H+1 <= 16
__host__ __device__ static inline bool tc_supported(const egnn_net_t& n) {
    return n.Mh == n.Ms && n.Ms == n.Xh && (n.Mh == 64 || n.Mh == 128) && 2 * n.H + 1 <= 16;
}
// the tcgen05 edge FORWARD additionally covers d = 256 (tc_edge256.cu: one branch per CTA, W2 streamed through shared
// memory in K slices by bulk-async copies); its backward still runs on the FFMA kernels
__host__ __device__ static inline bool tc_fwd_supported(const egnn_net_t& n) {
    return n.Mh == n.Ms && n.Ms == n.Xh && (n.Mh == 64 || n.Mh == 128 || n.Mh == 256) && 2 * n.H + 1 <= 16;
}
// ---- generic ("wide") tcgen05 forward path: tc_wide.cu -------------------------------------------------------------------------
// Networks the fused edge kernels do not cover -- the reference's own parameters.yaml shape (K0 = 73, m/x hidden 1024,
// m_size 256), ragged widths, conditional models -- run their edge MLPs as a chain of tile GEMMs [128 edges x 256 outputs]
// with K streamed in 64-wide slices.  Operands are bf16 hi + lo "tile images": a [R rows x 64] slice is 8 chunk columns of
// R x 16 bytes (hi) followed by the same for lo -- exactly the SWIZZLE_NONE K-major layout of tc_common.cuh, so one
// cp.async.bulk per slice lands it in shared memory ready for tcgen05.mma.  Widths are zero-padded to the tile sizes.
constexpr int WIDE_TM = 128, WIDE_TN = 256, WIDE_TK = 64;
constexpr int WIDE_A_SLICE = WIDE_TM * WIDE_TK * 2 * 2;  // bytes of one activation slice (hi | lo): 32 KB
constexpr int WIDE_B_SLICE = WIDE_TN * WIDE_TK * 2 * 2;  // bytes of one weight slice (hi | lo): 64 KB
__host__ __device__ static inline int wide_cdiv(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ static inline bool tc_fwd_supported(const egnn_net_t& n);
// forward only; the backward of such a network runs the fp32 FFMA kernels
__host__ __device__ static inline bool wide_fwd_supported(const egnn_net_t& n) {
    return !tc_fwd_supported(n) && n.Mh <= 1024 && n.Xh <= 1024 && n.Ms <= WIDE_TN && 2 * n.H + 1 <= 128;
}
struct WideOff {  // per layer, in units of WIDE_B_SLICE blocks: block (nt, ks) of a matrix sits at first + nt * KS + ks
    int ks0, ksm, ksx, kss;      // K slices: K0, Mh, Xh, Ms
    int ntm1, ntm2, ntx1, ntx2;  // N tiles: Mh, Ms, Xh, Xh
    int m1, m2, x1, x2;          // forward: first block of each matrix
    // backward (dgrad) images, rows = INPUT features: Wm2^T [Mh x Ms], Wx2^T [Xh x Xh], [Wm1^T | Wx1^T] [K0 x (Mh + Xh)]
    int m2t, x2t, w1t;
    // node MLP (mlp_h): Wh1 [Hh x (H + Ms)], Wh2 [H x Hh] and their transposes
    int ksn, ksh, ksH, nth, ntn;  // K slices: H + Ms, Hh, H;  N tiles: Hh, H + Ms
    int h1, h2, h2t, h1t, blocks;
};
__host__ __device__ static inline WideOff wide_off(const egnn_net_t& n) {
    WideOff o;
    o.ks0 = wide_cdiv(2 * n.H + 1, WIDE_TK);
    o.ksm = wide_cdiv(n.Mh, WIDE_TK);
    o.ksx = wide_cdiv(n.Xh, WIDE_TK);
    o.kss = wide_cdiv(n.Ms, WIDE_TK);
    o.ntm1 = wide_cdiv(n.Mh, WIDE_TN);
    o.ntm2 = wide_cdiv(n.Ms, WIDE_TN);
    o.ntx1 = wide_cdiv(n.Xh, WIDE_TN);
    o.ntx2 = wide_cdiv(n.Xh, WIDE_TN);
    o.m1 = 0;
    o.m2 = o.m1 + o.ntm1 * o.ks0;
    o.x1 = o.m2 + o.ntm2 * o.ksm;
    o.x2 = o.x1 + o.ntx1 * o.ks0;
    o.m2t = o.x2 + o.ntx2 * o.ksx;
    o.x2t = o.m2t + o.ntm1 * o.kss;
    o.w1t = o.x2t + o.ntx2 * o.ksx;
    o.ksn = wide_cdiv(n.H + n.Ms, WIDE_TK);
    o.ksh = wide_cdiv(n.Hh, WIDE_TK);
    o.ksH = wide_cdiv(n.H, WIDE_TK);
    o.nth = wide_cdiv(n.Hh, WIDE_TN);
    o.ntn = wide_cdiv(n.H + n.Ms, WIDE_TN);
    o.h1 = o.w1t + 1 * (o.ksm + o.ksx);
    o.h2 = o.h1 + o.nth * o.ksn;
    o.h2t = o.h2 + 1 * o.ksh;
    o.h1t = o.h2t + o.nth * o.ksH;
    o.blocks = o.h1t + o.ntn * o.ksh;
    return o;
}
// forward workspace of the wide path (bytes, each part 1 KB aligned): U image | A1 image | g*m fp32 [E][256] | phi parts
struct WideWs {
    int64_t u, a1, gm, phip, ni, a1n, total;
    int mtiles, a1_ks;  // a1_ks: slices per tile of the A1 image (its row stride)
    int mtiles_n;       // node phase: row tiles of the [h | M] image (ni) and of the hidden image (a1n, nth * 4 slices per tile)
};
// the node MLP runs on the tile GEMMs too when its hidden width fits the same limit (else the FFMA node kernels)
// -- also for the d = 64 / 128 / 256 networks of the fused edge kernels whose node MLP has no fused tcgen05 kernel (tc_node.cu
// covers d <= 128 with Hh = d): at d = 256 the FFMA node kernels were 18 % of the cfg4 step
__host__ __device__ static inline bool wide_node_supported(const egnn_net_t& n) {
    const bool fused_node = tc_supported(n) && n.Hh == n.Ms && n.H + 1 <= 16;  // = tc_node_supported (defined below)
    return n.Hh <= 1024 && n.Ms <= WIDE_TN && 2 * n.H + 1 <= 128 && (wide_fwd_supported(n) || (tc_fwd_supported(n) && !fused_node));
}
__host__ __device__ static inline bool wide_any(const egnn_net_t& n) { return wide_fwd_supported(n) || wide_node_supported(n); }
static inline WideWs wide_ws(const egnn_net_t& n, int64_t E, int64_t N) {
    WideWs w;
    const WideOff o = wide_off(n);
    w.mtiles = (int)((E + WIDE_TM - 1) / WIDE_TM);
    const int npm = o.ntm1 > o.ntx1 ? o.ntm1 : o.ntx1;
    w.a1_ks = npm * (WIDE_TN / WIDE_TK);
    int64_t p = 0;
    auto take = [&](int64_t b) { int64_t q = p; p += (b + 1023) / 1024 * 1024; return q; };
    const bool edges = wide_fwd_supported(n);
    w.u = take(edges ? (int64_t)w.mtiles * o.ks0 * WIDE_A_SLICE : 0);
    w.a1 = take(edges ? (int64_t)w.mtiles * w.a1_ks * WIDE_A_SLICE : 0);
    w.gm = take(edges ? E * WIDE_TN * 4 : 0);
    w.phip = take(edges ? (int64_t)o.ntx2 * E * 4 : 0);
    w.mtiles_n = (int)((N + WIDE_TM - 1) / WIDE_TM);
    w.ni = take(wide_node_supported(n) ? (int64_t)w.mtiles_n * o.ksn * WIDE_A_SLICE : 0);
    w.a1n = take(wide_node_supported(n) ? (int64_t)w.mtiles_n * o.nth * (WIDE_TN / WIDE_TK) * WIDE_A_SLICE : 0);
    w.total = p;
    return w;
}
// backward workspace of the wide path (tc_wide.cu: egnn_wide_edge_backward).  "T" images hold the transpose (rows = features,
// K = edges in 64-edge slices), the operands of the weight-gradient GEMMs.  Ep = edges padded to the 128-edge tile.
struct WideBwdWs {
    int64_t u, ut, a1, a1t, d1, g2, g2t, g1m, g1x, g1t, rows, q, gphi, phip, phi, gu, total;
    int mtiles, kst, np, w_ks;  // kst: 64-edge slices; np: padded hidden width; w_ks: slices per tile row of the np-wide images
    int splits, per;            // split-K of the weight gradients over the edge slices = private rows of the partial buffer
};
// node phase of the backward (rows = nodes): [h | M] image + transpose, hidden image + transpose, SiLU' rows, gh_out image +
// transpose, gz1 image + transpose, g[h | M] rows
struct WideNodeBwdWs {
    int64_t ni, nit, a1, a1t, d1, go, got, g1, g1t, gin, total;
    int mtiles, kst, hp, h_ks, gin_ld, splits, per;
};
static inline WideNodeBwdWs wide_node_bwd_ws(const egnn_net_t& n, int64_t N) {
    WideNodeBwdWs w;
    const WideOff o = wide_off(n);
    w.mtiles = (int)((N + WIDE_TM - 1) / WIDE_TM);
    w.kst = 2 * w.mtiles;
    w.hp = o.nth * WIDE_TN;
    w.h_ks = w.hp / WIDE_TK;
    w.gin_ld = o.ntn * WIDE_TN;
    w.splits = w.kst < 16 ? (w.kst > 0 ? w.kst : 1) : 16;
    w.per = wide_cdiv(w.kst > 0 ? w.kst : 1, w.splits);
    w.splits = wide_cdiv(w.kst > 0 ? w.kst : 1, w.per);
    const int64_t Np = (int64_t)w.mtiles * WIDE_TM;
    const int64_t img = (int64_t)w.mtiles * w.h_ks * WIDE_A_SLICE;
    int64_t p = 0;
    auto take = [&](int64_t b) { int64_t q = p; p += (b + 1023) / 1024 * 1024; return q; };
    w.ni = take((int64_t)w.mtiles * o.ksn * WIDE_A_SLICE);
    w.nit = take((int64_t)o.ntn * w.kst * WIDE_B_SLICE);
    w.a1 = take(img);
    w.a1t = take(img);
    w.d1 = take(Np * w.hp * 4);
    w.go = take((int64_t)w.mtiles * o.ksH * WIDE_A_SLICE);
    w.got = take((int64_t)wide_cdiv(n.H, WIDE_TM) * w.kst * WIDE_A_SLICE);
    w.g1 = take(img);
    w.g1t = take(img);
    w.gin = take(Np * w.gin_ld * 4);
    w.total = p;
    return w;
}
static inline WideBwdWs wide_bwd_ws(const egnn_net_t& n, int64_t E) {
    WideBwdWs w;
    w.mtiles = (int)((E + WIDE_TM - 1) / WIDE_TM);
    w.kst = 2 * w.mtiles;
    const WideOff o = wide_off(n);
    const int npm = o.ntm1 > o.ntx1 ? o.ntm1 : o.ntx1;
    w.np = npm * WIDE_TN;
    w.w_ks = w.np / WIDE_TK;
    w.splits = w.kst < 16 ? (w.kst > 0 ? w.kst : 1) : 16;
    w.per = wide_cdiv(w.kst > 0 ? w.kst : 1, w.splits);
    w.splits = wide_cdiv(w.kst > 0 ? w.kst : 1, w.per);
    const int64_t Ep = (int64_t)w.mtiles * WIDE_TM;
    const int64_t img = (int64_t)w.mtiles * w.w_ks * WIDE_A_SLICE;  // [Ep x np] hi | lo = the same bytes in either orientation
    int64_t p = 0;
    auto take = [&](int64_t b) { int64_t q = p; p += (b + 1023) / 1024 * 1024; return q; };
    w.u = take((int64_t)w.mtiles * o.ks0 * WIDE_A_SLICE);
    w.ut = take((int64_t)w.kst * WIDE_B_SLICE);
    w.a1 = take(img);
    w.a1t = take(img);
    w.d1 = take(Ep * w.np * 4);
    w.g2 = take(img);
    w.g2t = take(img);
    w.g1m = take(img);
    w.g1x = take(img);
    w.g1t = take(img);
    w.rows = take(Ep * w.np * 4);
    w.q = take(Ep * 4);
    w.gphi = take(Ep * 4);
    w.phip = take((int64_t)o.ntx2 * Ep * 4);
    w.phi = take(Ep * 4);
    w.gu = take(Ep * 128 * 4);
    w.total = p;
    return w;
}

// bytes of one layer's tcgen05 weight image (see tc_edge.cu: W1 hi/lo [2d x 16], W2m hi/lo, W2x hi/lo [d x d] bf16, fp32 vectors)
__host__ __device__ static inline int tc_blob_bytes(int d) { return 2 * (2 * d * 32) + 4 * (d * d * 2) + (6 * d + 4) * 4; }
// d = 256 only: K-major images of W2^T / 2 (hi | lo per branch), the streamed B operand of the dgrad in tc_edge256_bwd.cu;
// appended to the edge part of the image (d = 256 has no node part)
__host__ __device__ static inline int tc_w2t_offset(int d) { return (tc_blob_bytes(d) + 15) / 16 * 16; }
__host__ __device__ static inline int tc_w2t_bytes(int d) { return d == 256 ? 4 * d * d * 2 : 0; }
// the tcgen05 node backward additionally needs Hh = d (mlp_h hidden width) and room for the constant-1 column
__host__ __device__ static inline bool tc_node_supported(const egnn_net_t& n) {
    return tc_supported(n) && n.Hh == n.Ms && n.H + 1 <= 16;
}
// node part of the image, appended to the edge part: Wh1 hi/lo as a K-major [d x (d+16)] bf16 tile (columns >= H+Ms zero)
__host__ __device__ static inline int tc_node_blob_bytes(int d) { return 2 * (d * (d + 16) * 2); }
__host__ __device__ static inline PackOff pack_off(const egnn_net_t& n) {
    PackOff o;
    const int K0 = 2 * n.H + 1;
    int p = 0;
    o.wm1t = p; p += K0 * n.Mh;
    o.wm2t = p; p += n.Mh * n.Ms;
    o.wx1t = p; p += K0 * n.Xh;
    o.wx2t = p; p += n.Xh * n.Xh;
    o.wh1t = p; p += (n.H + n.Ms) * n.Hh;
    p = round4(p);
    o.tc = p;
    if (tc_fwd_supported(n)) p += (tc_blob_bytes(n.Mh) + 15) / 16 * 4 + tc_w2t_bytes(n.Mh) / 4;
    if (tc_node_supported(n)) p += tc_node_blob_bytes(n.Mh) / 4;
    p = round4(p);
    o.wide = p;
    if (wide_any(n)) p += wide_off(n).blocks * (WIDE_B_SLICE / 4);
    o.total = round4(p);
    return o;
}

// ---- forward workspace layout (floats).  Kept by the caller between forward and backward. -----------------
struct FwdWs {
    int64_t agg;    // [N][Ms+4]   per-receiver sums written directly by edge tiles
    int64_t part;   // [T][2][Ms+4] head / tail partial sums of tiles
    int64_t D;      // [N]         per-node sum of |r_e|^2
    int64_t hl;     // [L][N][H]   input h of layer l (l >= 1; slot 0 unused)
    int64_t xl;     // [L][N][3]
    int64_t msum;   // [L][N][Ms]  combined message sums (input of mlp_h)
    int64_t dxu;    // [L][N][3]   un-normalised coordinate sums
    int64_t S;      // [L][G]      normaliser per graph
    int64_t phi;    // [E]         per-edge phi of the current layer (tcgen05 path)
    int64_t wide;   // wide-path scratch (WideWs, in floats)
    int64_t total;
};
static inline int edge_tile_for(const egnn_net_t& n) {
    int dmax = n.Mh;
    if (n.Ms > dmax) dmax = n.Ms;
    if (n.Xh > dmax) dmax = n.Xh;
    int te = 128;
    while (te > 16 && (int64_t)te * (round4(dmax) + 4) > 128 * 132) te >>= 1;
    return te;
}
// rows of the tile-partial buffer are sized for the finest segmentation any path uses (tcgen05: 32-edge segments)
static inline int part_tile_for(const egnn_net_t& n) { return tc_fwd_supported(n) ? 32 : edge_tile_for(n); }
static inline FwdWs fwd_ws(const egnn_net_t& n, int64_t N, int64_t E, int64_t G) {
    FwdWs w;
    const int64_t AGGW = n.Ms + 4;
    const int64_t T = (E + part_tile_for(n) - 1) / part_tile_for(n);
    int64_t p = 0;
    auto take = [&](int64_t cnt) { int64_t o = p; p += round_up64(cnt, 4); return o; };
    w.agg = take(N * AGGW);
    w.part = take((T + 1) * 2 * AGGW);
    w.D = take(N);
    w.hl = take((int64_t)n.L * N * n.H);
    w.xl = take((int64_t)n.L * N * 3);
    w.msum = take((int64_t)n.L * N * n.Ms);
    w.dxu = take((int64_t)n.L * N * 3);
    w.S = take((int64_t)n.L * G);
    w.phi = take(tc_fwd_supported(n) || wide_fwd_supported(n) ? E : 0);
    w.wide = take(wide_any(n) ? 256 + wide_ws(n, E, N).total / 4 : 0);  // + 256 floats: the caller aligns to 1 KB
    w.total = p;
    return w;
}

// ---- device math ---------------------------------------------------------------------------------------
// sigmoid via two MUFU ops (ex2, rcp; ~3 ulp): 1 / (1 + 2^(-z log2 e)).  z -> -inf gives rcp(inf) = 0, z -> +inf gives 1.
__device__ __forceinline__ float sigmoid_f(float z) {
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(z * -1.4426950408889634f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f + e));
    return r;
}
__device__ __forceinline__ float silu_f(float z) { return z * sigmoid_f(z); }
// d/dz SiLU(z) = s (1 + z (1 - s))
__device__ __forceinline__ float dsilu_f(float z) {
    const float s = sigmoid_f(z);
    return s * (1.f + z * (1.f - s));
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// fixed-order block reduction (blockDim.x == 256); result valid in every thread
__device__ __forceinline__ float block_sum_256(float v, float* red /* [8] smem */) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += red[w];
    return t;
}
