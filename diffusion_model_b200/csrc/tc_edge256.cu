// Fused EGCL edge kernel on tcgen05 for hidden width d = 256 (forward): BASELINE configs[3] / the d = 256 points of the
// layer sweep.  Same maths and the same outputs (agg / part / phi_edge) as tc_edge_fwd in tc_edge.cu; what changes is
// where the operands live, because one branch's layer-2 weights alone (W2 [256 x 256] as bf16 hi + lo = 256 KB) exceed
// the 227 KB of shared memory:
//   * each CTA owns ONE branch (blockIdx.x < nM: mlp_m + gate, else mlp_x) and walks its 128-edge tiles;
//   * the activation tile A1 = SiLU(z1) [128 x 256] (hi + lo, 128 KB) stays resident for the tile;
//   * W2 is STREAMED: its K-major image is cut into 16 K-slices of [256 rows x 16] (8 KB hi + 8 KB lo, contiguous in the
//     packed image) that one elected thread pulls from L2 with cp.async.bulk (TMA, 1-D) into a 4-slot ring; every slot
//     has a "full" mbarrier (complete_tx bytes) and an "empty" mbarrier (tcgen05.commit of the MMAs that read it).  The
//     slice sequence is the same for every tile, so the ring runs ahead across tile boundaries: the first slices of the
//     next tile are already in shared memory when its layer-1 epilogue finishes;
//   * TMEM: z1 in columns [0, 256), z2 in [256, 512): all 512 columns.
// Per tile:  GEMM1 (issued one tile ahead) -> SiLU -> A1 -> 16 x (slice wait, NPASS MMAs of K = 16, commit) -> layer-2
// epilogue: mlp_x: phi_e = wx3 . SiLU(z2 + b2) + bx3;  mlp_m: m = SiLU(z2 + b2), g = sigmoid(wa . m + ba), g * m summed per
// receiver in edge order (fixed order, no atomics) exactly like the d <= 128 kernel.
#include <cstdlib>

#include "tc_bulk.cuh"
#include "tc_tile.cuh"

#ifdef EGNN_DBG_TIMING  // phase time stamps of tile #3 of CTA 0 and CTA 147 (timing experiment, scripts/dbg_phases256.py)
__device__ long long g_dbg_stamp[2][16];
#define DBG_STAMP(k) do { if (it == 3 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) g_dbg_stamp[blockIdx.x ? 1 : 0][k] = clock64(); } while (0)
extern "C" int egnn_dbg_read_stamps(long long* host16x2) {
    return cudaMemcpyFromSymbol(host16x2, g_dbg_stamp, sizeof(g_dbg_stamp)) == cudaSuccess ? 0 : -2;
}
#else
#define DBG_STAMP(k) do { } while (0)
#endif

namespace {

constexpr int D = 256, NT = 512;
constexpr int KS = 16;            // K extent of one streamed W2 slice (one tcgen05 k-step)
constexpr int NSLICE = D / KS;    // slices per tile
constexpr int NSLOT = 4;          // ring depth
constexpr int SLICE_B = D * KS * 2;  // bytes of one slice of the hi (or lo) image

struct L256 {
    static constexpr int w1_hi = 0, w1_lo = D * 32;            // this branch's [256 x 16] layer-1 image
    static constexpr int u_hi = 2 * D * 32, u_lo = u_hi + 128 * 32;
    static constexpr int a1_hi = u_lo + 128 * 32, a1_lo = a1_hi + 128 * D * 2;
    static constexpr int stage = a1_hi;                        // fp32 [128][256] staging of m aliases A1 (same size)
    static constexpr int ring = a1_lo + 128 * D * 2;           // NSLOT x (hi slice | lo slice)
    static constexpr int vec = ring + NSLOT * 2 * SLICE_B;     // b2/2 [256] | wv [256] | bv, 0, 0, 0
    static constexpr int rowPart = vec + (2 * D + 4) * 4;      // float [4][128]
    static constexpr int sGate = rowPart + 4 * 512;            // float [128]
    static constexpr int sDst = sGate + 512;                   // int [128]
    static constexpr int sMask = sDst + 512;                   // u32 [4]
    static constexpr int bars = sMask + 16;                    // [0] GEMM1, [1] GEMM2, [2..5] full, [6..9] empty
    static constexpr int slot = bars + 96;
    static constexpr int TOTAL = slot + 16;
    static_assert(TOTAL <= 227 * 1024, "tc_edge_fwd256 shared memory");
    static_assert(ring % 128 == 0 && a1_hi % 128 == 0, "operand tiles must be 128-byte aligned");
};

struct Tc256Args {
    egnn_net_t net;
    int N, E, tiles;
    const int32_t *recv, *send, *row_ptr;
    const float *h, *x;
    const uint8_t* blob;
    float *agg, *part, *phi_edge;
    int nM;
};

template <int NPASS>
__global__ void __launch_bounds__(NT, 1) tc_edge_fwd256(Tc256Args a) {
    using L = L256;
    constexpr bool LO = NPASS == 3;
    constexpr int TS = 32;  // edges per segment == partial-buffer granularity (part_tile_for)
    constexpr uint32_t CS = 128 * 16;
    extern __shared__ __align__(1024) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int q = warp & 3, cg = warp >> 2;
    const int r = q * 32 + lane;  // this thread's edge row == TMEM lane
    const int c0 = cg * 64;       // this thread's 64 columns (two passes of 32)
    const int branch = (int)blockIdx.x < a.nM ? 0 : 1;
    const int H = a.net.H, AGGW = D + 4;
    float* vec = reinterpret_cast<float*>(smem + L::vec);
    float* rowPart = reinterpret_cast<float*>(smem + L::rowPart);
    float* sGate = reinterpret_cast<float*>(smem + L::sGate);
    int* sDst = reinterpret_cast<int*>(smem + L::sDst);
    uint32_t* sMask = reinterpret_cast<uint32_t*>(smem + L::sMask);
    float* stage = reinterpret_cast<float*>(smem + L::stage);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::bars);
    uint32_t* slot = reinterpret_cast<uint32_t*>(smem + L::slot);

    // packed image (TcLayout in tc_edge.cu): W1 hi | W1 lo ([2D x 16] each), W2m hi | lo, W2x hi | lo, fp32 vectors
    constexpr int W1B = 2 * D * 32, W2B = D * D * 2;
    {
        for (int o = tid * 16; o < 4 * D * 16; o += NT * 16) {  // [hi|lo][chunk j][this branch's D rows * 16 bytes]
            const int half = o / (2 * D * 16), j = (o / (D * 16)) & 1, off = o % (D * 16);
            *reinterpret_cast<uint4*>(smem + (half ? L::w1_lo : L::w1_hi) + j * D * 16 + off) =
                __ldg(reinterpret_cast<const uint4*>(a.blob + half * W1B + j * 2 * D * 16 + branch * D * 16 + off));
        }
        const float* vsrc = reinterpret_cast<const float*>(a.blob + 2 * W1B + 4 * W2B);
        for (int c = tid; c < D; c += NT) {
            vec[c] = vsrc[2 * D + branch * D + c];      // b2 / 2
            vec[D + c] = vsrc[4 * D + branch * D + c];  // wa | wx3
        }
        if (tid == 0) vec[2 * D] = vsrc[6 * D + branch];  // ba | bx3
    }
    const uint8_t* w2_hi_g = a.blob + 2 * W1B + (size_t)branch * 2 * W2B;
    const uint8_t* w2_lo_g = w2_hi_g + W2B;
    if (warp == 0) tc::tmem_alloc(slot, 512);
    if (tid == 0) {
        for (int b = 0; b < 2 + 2 * NSLOT; ++b) tc::mbar_init(&bars[b], 1);
        tc::fence_mbar_init();
    }
    const uint32_t sb = tc::smem_u32(smem);
    const TcOperand opU{sb + L::u_hi, sb + L::u_lo, CS, false};
    const TcOperand opW1{sb + L::w1_hi, sb + L::w1_lo, D * 16, false};

    const int tstride = branch ? (int)gridDim.x - a.nM : a.nM;
    const int tile0 = branch ? (int)blockIdx.x - a.nM : (int)blockIdx.x;
    const int nit = tile0 < a.tiles ? (a.tiles - tile0 + tstride - 1) / tstride : 0;
    const int total_slices = nit * NSLICE;

    struct Row {
        int i, inext, j;
        float xi[3], xj[3];
        float u[14];
        int rp0, rp1;
    };
    auto gather_ids = [&](int tile, Row& g) {
        const int e = min(tile * 128 + tid, a.E - 1);  // rows past the end repeat the last edge; they are never flushed
        g.i = a.recv[e];
        g.j = a.send[e];
        g.inext = a.recv[min(e + 1, a.E - 1)];
    };
    auto gather_nodes = [&](Row& g) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            g.xi[k] = a.x[3 * g.i + k];
            g.xj[k] = a.x[3 * g.j + k];
        }
        const float* hi_ = a.h + (size_t)g.i * H;
        const float* hjm = a.h + (size_t)g.j * H - H;
#pragma unroll
        for (int k = 0; k < 14; ++k) {
            const float* p = k < H ? hi_ : hjm;
            g.u[k] = k < 2 * H ? p[k] : 0.f;
        }
        g.rp0 = a.row_ptr[g.i];
        g.rp1 = a.row_ptr[g.i + 1];
    };
    auto store_u = [&](int tile, const Row& g) -> int {
        const int e0 = tile * 128, nrows = min(128, a.E - e0);
        const bool valid = tid < nrows;
        const float rx = g.xi[0] - g.xj[0], ry = g.xi[1] - g.xj[1], rz = g.xi[2] - g.xj[2];
        const float nrm = sqrtf(rx * rx + ry * ry + rz * rz);
        const float d2 = nrm * nrm;
        float u[16];
#pragma unroll
        for (int k = 0; k < 14; ++k) u[k] = k == 2 * H ? d2 : g.u[k];
        u[14] = 2 * H == 14 ? d2 : 0.f;
        u[15] = 1.f;  // bias slot (layer-1 biases travel in K slot 15 of W1)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int p = 0; p < 4; ++p) tc::split2(u[8 * j + 2 * p], u[8 * j + 2 * p + 1], hi[p], lo[p]);
            *reinterpret_cast<uint4*>(smem + L::u_hi + tc::chunk_off(128, tid, j)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(smem + L::u_lo + tc::chunk_off(128, tid, j)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
        int dst = -3;  // >= 0: agg[node]; -1: head partial; -2: tail partial; -3: the run continues in the next row
        if (valid) {
            const bool last = tid == nrows - 1 || (tid + 1) % TS == 0 || g.inext != g.i;
            if (last) {
                const int es = e0 + (tid / TS) * TS;
                dst = g.rp0 < es ? -1 : (g.rp1 > es + TS ? -2 : g.i);
            }
        }
        return dst;
    };
    auto store_dst = [&](int dst) {
        sDst[tid] = dst;
        const uint32_t m = __ballot_sync(0xffffffffu, dst != -3);
        if (lane == 0) sMask[warp] = m;
    };
    // the elected issuer thread: pull slice number g (tile-independent content: slice g % NSLICE) into ring slot g % NSLOT
    auto load_slice = [&](int g) {
#ifdef EGNN_DBG_NOSTREAM
        return;
#endif
        const int sl = g % NSLICE, st = g % NSLOT;
        const uint32_t full = tc::smem_u32(&bars[2 + st]);
        const uint32_t dst = sb + L::ring + st * 2 * SLICE_B;
        tc::mbar_expect_tx(full, LO ? 2 * SLICE_B : SLICE_B);
        tc::bulk_g2s(dst, w2_hi_g + (size_t)sl * SLICE_B, SLICE_B, full);
        if (LO) tc::bulk_g2s(dst + SLICE_B, w2_lo_g + (size_t)sl * SLICE_B, SLICE_B, full);
    };

    if (tid < 128 && nit > 0) {
        Row g;
        gather_ids(tile0, g);
        gather_nodes(g);
        store_dst(store_u(tile0, g));
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *slot;
    const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
    const bool issuer_warp = warp == NT / 32 - 1;
    // The W2 ring has its own producer thread (lane 0 of the warp before the issuer's): the MMA issuer only ever waits for
    // "slice landed", never for "slot free", so several slices' MMAs stay queued on the tensor pipe.  (With one thread
    // doing both, every slice waited for the previous slice's MMAs to drain: the GEMM ran at one commit round trip per slice.)
    const bool producer = warp == NT / 32 - 2 && lane == 0;
    int gload = 0;  // producer: next slice to pull
    bool ok = true;
    auto produce = [&](int limit) {
        for (; gload < limit; ++gload) {
            // slots are released in PAIRS (one tcgen05.commit per two slices: the commit is the expensive part of an
            // issue iteration): the pair of slices g, g+1 (g even) waits for the commit that followed slice g - 3
            if (gload >= NSLOT && (gload & 1) == 0)
                ok = tc::mbar_wait(&bars[2 + NSLOT + (gload / 2) % 2], ((gload - NSLOT) / NSLOT) & 1) && ok;
            load_slice(gload);
        }
    };
    if (producer && nit > 0) produce(min(total_slices, NSLOT));
    if (issuer_warp && nit > 0 && tc::elect_one()) {
        tc_issue_op<NPASS>(tmem, 0, opU, opW1, 16, D, false);
        tc::mma_commit(&bars[0]);
    }

    for (int it = 0; it < nit; ++it) {
        const int tile = tile0 + it * tstride;
        const int e0 = tile * 128;
        const int nrows = min(128, a.E - e0);
        const uint32_t phase = it & 1;
        const bool has_next = it + 1 < nit;
        Row gnext;
        int dst_next = -3;
        if (tid < 128 && has_next) gather_ids(tile + tstride, gnext);
        if (tid == 0) DBG_STAMP(0);  // tile start
        ok = tc::mbar_wait(&bars[0], phase) && ok;  // z1 = U . W1^T
        tc::fence_after_sync();
        if (tid == 0) DBG_STAMP(1);  // GEMM1 observed
        if (tid < 128 && has_next) gather_nodes(gnext);
        // ---- layer-1 epilogue: a1 = SiLU(z1) -> tile A1 (bf16 hi / lo), this thread's 64 columns
        tc_epi_store_chunked(trow + c0, smem + L::a1_hi, smem + L::a1_lo, r, c0, LO);
        tc_epi_store_chunked(trow + c0 + 32, smem + L::a1_hi, smem + L::a1_lo, r, c0 + 32, LO);
        if (tid < 128 && has_next) dst_next = store_u(tile + tstride, gnext);  // GEMM1 is done with the U tile
        if (tid == 0) DBG_STAMP(2);  // warp 0 finished its layer-1 epilogue
        tc::fence_async_smem();
        tc::fence_before_sync();
        __syncthreads();
        if (tid == 0) DBG_STAMP(3);  // barrier passed
        if (issuer_warp && tc::elect_one()) {
            tc::fence_after_sync();
            DBG_STAMP(8);  // issuer starts
            const uint32_t idesc = tc::make_idesc_bf16(128, D);
            for (int s = 0; s < NSLICE; ++s) {
                const int g = it * NSLICE + s, st = g % NSLOT;
#ifndef EGNN_DBG_NOSTREAM  // timing experiment only (wrong results): MMAs without the W2 stream
                ok = tc::mbar_wait(&bars[2 + st], (g / NSLOT) & 1) && ok;  // slice s has landed
#endif
                const uint32_t wb = sb + L::ring + st * 2 * SLICE_B;
                const uint64_t a_hi = tc::make_desc(sb + L::a1_hi + s * 2 * CS, CS, 128);
                const uint64_t a_lo = tc::make_desc(sb + L::a1_lo + s * 2 * CS, CS, 128);
                const uint64_t b_hi = tc::make_desc(wb, D * 16, 128), b_lo = tc::make_desc(wb + SLICE_B, D * 16, 128);
                tc::mma_bf16(tmem + D, a_hi, b_hi, idesc, s > 0);
                if (LO) {
                    tc::mma_bf16(tmem + D, a_lo, b_hi, idesc, true);
                    tc::mma_bf16(tmem + D, a_hi, b_lo, idesc, true);
                }
                if (g & 1) tc::mma_commit(&bars[2 + NSLOT + (g / 2) % 2]);  // slots of slices g-1, g are free once these MMAs are done
            }
            tc::mma_commit(&bars[1]);
            DBG_STAMP(9);  // last MMA of GEMM2 issued and committed
            if (has_next) {  // z1 has been consumed by every warp (barrier above): next tile's GEMM1 queues behind GEMM2
                tc_issue_op<NPASS>(tmem, 0, opU, opW1, 16, D, false);
                tc::mma_commit(&bars[0]);
            }
        } else if (producer) {
            produce(min(total_slices, (it + 1) * NSLICE + NSLOT));  // this tile's slices, then the first ones of the next tile
        }
        __syncwarp();
        ok = tc::mbar_wait(&bars[1], phase) && ok;
        tc::fence_after_sync();
        if (tid == 0) DBG_STAMP(4);  // GEMM2 observed
        // ---- layer-2 epilogue
        if (branch == 1) {
            float s = 0.f;
#pragma unroll 1
            for (int hh = 0; hh < 2; ++hh) {
                float v[32];
                tc::tmem_ld32(trow + D + c0 + 32 * hh, v);
                const float* b2 = vec + c0 + 32 * hh;
                const float* wv = vec + D + c0 + 32 * hh;
#pragma unroll
                for (int i = 0; i < 32; ++i) s = fmaf(wv[i], tc_silu_h(v[i] + b2[i]), s);
            }
            rowPart[cg * 128 + r] = s;
            tc::fence_before_sync();
            __syncthreads();
            if (cg == 0 && r < nrows)
                a.phi_edge[e0 + r] = vec[2 * D] + rowPart[r] + rowPart[128 + r] + rowPart[256 + r] + rowPart[384 + r];
        } else {
            float s = 0.f;
#pragma unroll 1
            for (int hh = 0; hh < 2; ++hh) {
                float v[32];
                tc::tmem_ld32(trow + D + c0 + 32 * hh, v);
                const float* b2 = vec + c0 + 32 * hh;
                const float* wv = vec + D + c0 + 32 * hh;
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float m = tc_silu_h(v[i] + b2[i]);
                    s = fmaf(wv[i], m, s);
                    stage[r * D + ((c0 + 32 * hh + i) ^ lane)] = m;  // GEMM2 is complete: A1 may be overwritten
                }
            }
            rowPart[cg * 128 + r] = s;
            tc::fence_before_sync();
            __syncthreads();
            if (cg == 0) sGate[r] = sigmoid_f(vec[2 * D] + rowPart[r] + rowPart[128 + r] + rowPart[256 + r] + rowPart[384 + r]);
            __syncthreads();
            // fixed-order segmented sums of g_e * m_e over receivers: thread (segment, column) walks its 32 rows in edge
            // order and flushes at the rows marked in the segment's mask word
#pragma unroll 1
            for (int k = 0; k < 2; ++k) {
                const int seg = (tid >> 8) + 2 * k, c = tid & 255;
                const int r0 = seg * TS;
                float* part = a.part + (size_t)(tile * 4 + seg) * 2 * AGGW;
                uint32_t mask = sMask[seg];
                float sum = 0.f;
                int rr = r0;
                while (mask) {
                    const int stop = r0 + __ffs(mask) - 1;
                    mask &= mask - 1;
#pragma unroll 4
                    for (; rr <= stop; ++rr) sum += __fmul_rn(sGate[rr], stage[rr * D + (c ^ (rr & 31))]);
                    const int dst = sDst[stop];
                    if (dst >= 0) a.agg[(size_t)dst * AGGW + c] = sum;
                    else part[(dst == -1 ? 0 : AGGW) + c] = sum;
                    sum = 0.f;
                }
            }
        }
        __syncthreads();  // staging (aliases A1), rowPart and the flush table are free
        if (tid == 0) DBG_STAMP(5);  // tile end
        if (tid < 128 && has_next) store_dst(dst_next);
    }
    if (!ok) asm volatile("trap;");  // an mbarrier wait timed out: fail loudly instead of returning garbage
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

template <int NPASS>
int launch256(const Tc256Args& a, int grid, cudaStream_t st) {
    EGNN_SMEM_ATTR(tc_edge_fwd256<NPASS>, L256::TOTAL);
    egnn_prof_begin(EGNN_K_TC_EDGE_FWD, st);
    tc_edge_fwd256<NPASS><<<grid, NT, L256::TOTAL, st>>>(a);
    egnn_prof_end(st);
    EGNN_CHECK_LAUNCH("tc_edge_fwd256");
    return EGNN_OK;
}

}  // namespace

int egnn_tc_edge_forward256(const egnn_net_t& net, const egnn_graph_t& g, const float* packed_layer, const float* h,
                            const float* x, float* agg, float* part, float* phi_edge, int precision, cudaStream_t st) {
    EGNN_REQUIRE(net.Mh == 256 && tc_fwd_supported(net), "egnn_tc_edge_forward256: needs Mh = Ms = Xh = 256 and 2H+1 <= 16");
    const PackOff po = pack_off(net);
    const int tiles = (g.E + 127) / 128;
    // one CTA per SM; the message branch does more per tile (gate, staging, segmented sums: 19.0 k cycles against 13.7 k for
    // the coordinate branch, scripts/dbg_phases256.py), so it gets the larger share of the CTAs
    static int share = -1;
    if (share < 0) {
#ifdef EGNN_DBG  // tuning knob of debug builds only
        const char* env = getenv("EGNN_TC256_FWD_MCTAS");
        share = env ? atoi(env) : 86;
#else
        share = 86;
#endif
        if (share < 16 || share > 132) share = 86;
    }
    int nM = share, nX = 148 - share;
    if (tiles < nM) nM = tiles;
    if (tiles < nX) nX = tiles;
    Tc256Args a{net, g.N, g.E, tiles, g.recv, g.send, g.row_ptr, h, x,
                reinterpret_cast<const uint8_t*>(packed_layer + po.tc), agg, part, phi_edge, nM};
    return precision == EGNN_PREC_BF16X3 ? launch256<3>(a, nM + nX, st) : launch256<1>(a, nM + nX, st);
}
