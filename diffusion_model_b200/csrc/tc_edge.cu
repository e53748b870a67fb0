// tcgen05 (5th-generation tensor core) path of the edge MLPs: operands staged in shared memory as bf16 hi/lo splits,
// fp32 accumulators in TMEM, SiLU epilogues out of TMEM.  This file also holds the one-tile GEMM self-test that pins
// the shared-memory descriptor conventions on real hardware (tests/test_gpu_tc.py).
#include "tc_common.cuh"

namespace {

// D[128 x N] = A[128 x K] . B[N x K]^T   (one CTA, 128 threads; K % 16 == 0, N % 16 == 0, N <= 256)
//   npass 1: bf16(A) bf16(B);  npass 3: hi.hi + lo.hi + hi.lo  (fp32-class accuracy)
//   swap != 0 exchanges the LBO / SBO fields of the descriptors (convention probe)
__global__ void __launch_bounds__(128) tc_selftest_kernel(const float* __restrict__ A, const float* __restrict__ B,
                                                          float* __restrict__ D, int K, int N, int npass, int swap,
                                                          int* __restrict__ status) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int KC = K / 8;
    uint8_t* a_hi = smem;
    uint8_t* a_lo = a_hi + 128 * K * 2;
    uint8_t* b_hi = a_lo + 128 * K * 2;
    uint8_t* b_lo = b_hi + N * K * 2;
    for (int j = 0; j < KC; ++j) {
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) tc::split2(A[tid * K + j * 8 + 2 * q], A[tid * K + j * 8 + 2 * q + 1], hi[q], lo[q]);
        *reinterpret_cast<uint4*>(a_hi + tc::chunk_off(128, tid, j)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(a_lo + tc::chunk_off(128, tid, j)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
    for (int n = tid; n < N; n += 128) {
        for (int j = 0; j < KC; ++j) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) tc::split2(B[n * K + j * 8 + 2 * q], B[n * K + j * 8 + 2 * q + 1], hi[q], lo[q]);
            *reinterpret_cast<uint4*>(b_hi + tc::chunk_off(N, n, j)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(b_lo + tc::chunk_off(N, n, j)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
    }
    if (warp == 0) tc::tmem_alloc(&tmem_slot, 256);
    if (tid == 0) {
        tc::mbar_init(&bar, 1);
        tc::fence_mbar_init();
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = tmem_slot;
    if (tid == 0) {
        const uint32_t idesc = tc::make_idesc_bf16(128, N);
        const uint32_t lboA = 128 * 16, lboB = N * 16, sbo = 128;
        bool acc = false;
        for (int p = 0; p < npass; ++p) {
            const uint8_t* ta = p == 1 ? a_lo : a_hi;
            const uint8_t* tb = p == 2 ? b_lo : b_hi;
            for (int s = 0; s < K / 16; ++s) {
                const uint32_t aa = tc::smem_u32(ta) + s * 2 * lboA, ba = tc::smem_u32(tb) + s * 2 * lboB;
                const uint64_t da = swap ? tc::make_desc(aa, sbo, lboA) : tc::make_desc(aa, lboA, sbo);
                const uint64_t db = swap ? tc::make_desc(ba, sbo, lboB) : tc::make_desc(ba, lboB, sbo);
                tc::mma_bf16(tmem, da, db, idesc, acc);
                acc = true;
            }
        }
        tc::mma_commit(&bar);
    }
    const bool ok = tc::mbar_wait(&bar, 0);
    tc::fence_after_sync();
    if (!ok && tid == 0) *status = 1;
    if (ok) {
        for (int c0 = 0; c0 < N; c0 += 32) {
            float v[32];
            tc::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (c0 + i < N) D[tid * N + c0 + i] = v[i];
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 256);
}

}  // namespace

extern "C" int egnn_tc_selftest(const float* A, const float* B, float* D, int32_t K, int32_t N, int32_t npass,
                                int32_t swap_lbo_sbo, int32_t* status_dev, void* stream) {
    EGNN_REQUIRE(A && B && D && status_dev, "egnn_tc_selftest: NULL pointer");
    EGNN_REQUIRE(K >= 16 && K % 16 == 0 && K <= 128 && N >= 16 && N % 16 == 0 && N <= 256 && (npass == 1 || npass == 3),
                 "egnn_tc_selftest: unsupported shape K=%d N=%d npass=%d", K, N, npass);
    const size_t smem = 200 * 1024;  // fixed and ample: the swapped-descriptor probe must stay inside the allocation
    EGNN_CUDA(cudaFuncSetAttribute(tc_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    tc_selftest_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(A, B, D, K, N, npass, swap_lbo_sbo, status_dev);
    EGNN_CHECK_LAUNCH("tc_selftest_kernel");
    return EGNN_OK;
}
