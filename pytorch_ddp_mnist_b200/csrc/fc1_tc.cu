// fc1 forward on the 5th-generation tensor cores (tcgen05, TF32 inputs, FP32 accumulation in TMEM).
//
//   part[s][b][n] = sum_{k in split s} X[b][k] * W1[n][k]        M = batch, N = 128, K = 784
//
// Opt-in (DDPM_FWD_TF32): TF32 rounds both operands to 10 mantissa bits, so results agree with the
// fp32 reference to ~1e-3 relative (the tolerance BASELINE.json:north_star states), not 2e-5 like the
// FFMA kernels.  It pays off only where fc1 is FLOP-bound, i.e. large batches (4096: 0.82 GFLOP).
//
// One CTA = one 128-row batch tile x one of 4 K splits (6-7 chunks of 32 k).  The A operand does not
// exist in memory: the four warps gather the uint8 pixel rows through `idx`, decode them with a
// TF32-pre-rounded 256-entry table and write them straight into shared memory in the canonical
// K-major SWIZZLE_128B layout the UMMA descriptor expects; W1 chunks go in the same way.  After one
// proxy fence a single thread issues the chain of tcgen05.mma (M128 N128 K8) into a 128-column TMEM
// accumulator, commits to an mbarrier, and the warps read their 32 TMEM lanes back with tcgen05.ld
// and store the split's partial sums, which rows_kernel adds up like any other split-K partial.
#include "tc_common.cuh"

namespace ddpm {

constexpr int TC_BM = 128;                 // batch rows per CTA (UMMA M)
constexpr int TC_BK = 32;                  // k per chunk: 32 fp32 = one 128-byte swizzle row
constexpr int TC_CHUNKS = (IN + TC_BK - 1) / TC_BK;      // 25 (k 784..799 are zero padding)
constexpr int TC_KSPLIT = 4;
constexpr int TC_MAX_CHUNKS = 7;           // chunks per CTA: 6, 6, 6, 7
constexpr int TC_TILE_BYTES = TC_BM * 128; // 16 KB per operand per chunk
constexpr int TC_THREADS = 128;
constexpr size_t TC_SMEM = (size_t)TC_MAX_CHUNKS * 2 * TC_TILE_BYTES + 1024 /*lut*/ + 64 /*barrier, tmem ptr*/ + 1024 /*align*/;

int fc1_tc_ksplit() { return TC_KSPLIT; }

// Instruction descriptor (cute::UMMA::InstrDescriptor) for kind::tf32, FP32 accumulate, both operands
// K-major: c_format[4,6)=1 (F32), a_format[7,10)=2 (TF32), b_format[10,13)=2, n_dim[17,23)=N>>3, m_dim[24,29)=M>>4
constexpr uint32_t TC_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(HID >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);

__global__ void __launch_bounds__(TC_THREADS, 1)
fc1_fwd_tc_kernel(const float* __restrict__ W1, const uint8_t* __restrict__ images, int64_t n_samples,
                  const int64_t* __restrict__ idx, int B, float* __restrict__ part) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    float* lut = reinterpret_cast<float*>(smem + (size_t)TC_MAX_CHUNKS * 2 * TC_TILE_BYTES);
    uint64_t* mbar = reinterpret_cast<uint64_t*>(lut + 256);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbar + 1);

    const int tid = threadIdx.x, warp = tid >> 5;
    const int mtile = blockIdx.x / TC_KSPLIT, split = blockIdx.x - mtile * TC_KSPLIT;
    const int c0 = (split * TC_CHUNKS) / TC_KSPLIT, c1 = ((split + 1) * TC_CHUNKS) / TC_KSPLIT;
    const int m0 = mtile * TC_BM;

    // ---- set-up: decode table (TF32-rounded), mbarrier, TMEM (128 fp32 columns x 128 lanes)
    lut[tid] = to_tf32(decode_px(tid));
    lut[tid + 128] = to_tf32(decode_px(tid + 128));
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(mbar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" :: "r"(smem_u32(tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    // ---- producers: every chunk of this CTA's K range goes to shared memory, swizzled
    {
        // A: thread t owns batch row m0 + t: 32 pixels per chunk = two 16-byte loads
        const int row = m0 + tid;
        int64_t sidx = -1;
        if (row < B) {
            const int64_t v = __ldg(idx + row);
            if (v >= 0 && v < n_samples) sidx = v;
        }
        const uint8_t* src = images + (sidx < 0 ? 0 : sidx) * IN;
        for (int c = c0; c < c1; ++c) {
            uint8_t* At = smem + (size_t)(c - c0) * 2 * TC_TILE_BYTES;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int k = c * TC_BK + h * 16;
                uint4 v = make_uint4(0, 0, 0, 0);
                const bool ok = sidx >= 0 && k < IN;
                if (ok) v = __ldg(reinterpret_cast<const uint4*>(src + k));
                const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float4 o;
                    o.x = ok ? lut[w[q] & 0xff] : 0.f;          o.y = ok ? lut[(w[q] >> 8) & 0xff] : 0.f;
                    o.z = ok ? lut[(w[q] >> 16) & 0xff] : 0.f;  o.w = ok ? lut[w[q] >> 24] : 0.f;
                    *reinterpret_cast<float4*>(At + sw128(tid, h * 4 + q)) = o;
                }
            }
        }
        // B: W1 rows; 8 lanes cover one 128-byte k row (coalesced), 16 rows per pass
        const int q = tid & 7, nr = tid >> 3;
        for (int c = c0; c < c1; ++c) {
            uint8_t* Bt = smem + (size_t)(c - c0) * 2 * TC_TILE_BYTES + TC_TILE_BYTES;
            const int k = c * TC_BK + q * 4;
#pragma unroll
            for (int j = 0; j < HID / 16; ++j) {
                const int n = j * 16 + nr;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (k < IN) v = __ldg(reinterpret_cast<const float4*>(W1 + (size_t)n * IN + k));
                v.x = to_tf32(v.x); v.y = to_tf32(v.y); v.z = to_tf32(v.z); v.w = to_tf32(v.w);
                *reinterpret_cast<float4*>(Bt + sw128(n, q)) = v;
            }
        }
    }
    // generic-proxy writes -> visible to the async proxy that tcgen05.mma reads shared memory through
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();

    // ---- MMA: one thread issues (c1-c0) x 4 instructions of K = 8, accumulating in TMEM
    if (tid == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t a0 = smem_u32(smem);
        for (int i = 0; i < c1 - c0; ++i) {
#pragma unroll
            for (int j = 0; j < TC_BK / 8; ++j) {
                const uint64_t ad = umma_desc_sw128(a0 + i * 2 * TC_TILE_BYTES + j * 32);
                const uint64_t bd = umma_desc_sw128(a0 + i * 2 * TC_TILE_BYTES + TC_TILE_BYTES + j * 32);
                const uint32_t acc = (i | j) ? 1u : 0u;
                asm volatile(
                    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                    "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                    :: "r"(tmem_base), "l"(ad), "l"(bd), "r"(TC_IDESC), "r"(acc) : "memory");
            }
        }
        // arrives on the mbarrier when every MMA above has completed (implies fence::before_thread_sync)
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                     :: "r"(smem_u32(mbar)) : "memory");
    }
    {
        uint32_t done = 0;
        while (!done) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\t"
                         "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(mbar)) : "memory");
        }
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // ---- epilogue: warp w owns TMEM lanes 32w..32w+31 = batch rows m0+32w..; 32 columns per load
    {
        const int row = m0 + tid;
        float* dst = part + ((size_t)split * B + row) * HID;
#pragma unroll
        for (int cb = 0; cb < HID; cb += 32) {
            uint32_t r[32];
            const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)cb;
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                  "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                  "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                : "r"(taddr) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (row < B) {
#pragma unroll
                for (int i = 0; i < 32; i += 4)
                    *reinterpret_cast<float4*>(dst + cb + i) = make_float4(__uint_as_float(r[i]), __uint_as_float(r[i + 1]),
                                                                           __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" :: "r"(tmem_base) : "memory");
}

// ------------------------------------------------------------------------------------------
// fc1 weight gradient on the tensor cores:  dW1[m][n] = sum_b dz1[b][m] * X[b][n]   (K = batch)
//
// Both operands are stored batch-major in memory (the contraction index is the slow one), so the
// producers transpose while staging: each CTA takes one 112-pixel column tile (7 of them) and one
// batch slice of up to 7 chunks of 32 rows, and writes A = dz1^T [128 units][32 b] and
// B = X^T [112 px][32 b] per chunk in the same K-major SWIZZLE_128B layout as the forward kernel
// (lanes walk along b, so the transposing stores fill one 128-byte swizzle row each: conflict-free).
// 4 x tcgen05.mma (M128 N112 K8) per chunk accumulate in 112 TMEM columns; the slice's partial
// tile goes to gpart[slice] and grads_finalize_kernel sums the slices in fixed order.
// ------------------------------------------------------------------------------------------
constexpr int DW_BN = 112;                         // pixels per column tile (UMMA N)
constexpr int DW_NT = IN / DW_BN;                  // 7
constexpr int DW_MAX_CHUNKS = 7;
constexpr int DW_A_BYTES = HID * 128, DW_B_BYTES = DW_BN * 128;      // 16 KB + 14 KB per chunk
constexpr size_t DW_SMEM = (size_t)DW_MAX_CHUNKS * (DW_A_BYTES + DW_B_BYTES) + 1024 + 64 + 1024;
constexpr uint32_t DW_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(DW_BN >> 3) << 17) | ((uint32_t)(HID >> 4) << 24);

int dw1_tc_slices(int batch) {
    const int chunks = (batch + 31) / 32;
    return (chunks + DW_MAX_CHUNKS - 1) / DW_MAX_CHUNKS;
}

constexpr int DW_THREADS = 512;                    // 16 producer warps: 4 chunks in flight x 4 sub-ranges; warps 0-3 also do the epilogue
__global__ void __launch_bounds__(DW_THREADS, 1)
dw1_tc_kernel(const float* __restrict__ dz1, const uint8_t* __restrict__ images, int64_t n_samples,
              const int64_t* __restrict__ idx, int B, int nslices, float* __restrict__ gpart) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    float* lut = reinterpret_cast<float*>(smem + (size_t)DW_MAX_CHUNKS * (DW_A_BYTES + DW_B_BYTES));
    uint64_t* mbar = reinterpret_cast<uint64_t*>(lut + 256);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbar + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nt = blockIdx.x % DW_NT, slice = blockIdx.x / DW_NT;
    const int chunks = (B + 31) / 32;
    const int c0 = (slice * chunks) / nslices, c1 = ((slice + 1) * chunks) / nslices;
    const int n0 = nt * DW_BN;

    if (tid < 256) lut[tid] = to_tf32(decode_px(tid));
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(mbar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" :: "r"(smem_u32(tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    // ---- producers.  Lane = batch row within the chunk (the K index); warp -> (chunk in flight, group of units / pixels).
    const int pw = warp & 3;                                   // sub-range within a chunk
    for (int c = c0 + (warp >> 2); c < c1; c += DW_THREADS / 128) {
        uint8_t* At = smem + (size_t)(c - c0) * (DW_A_BYTES + DW_B_BYTES);
        uint8_t* Bt = At + DW_A_BYTES;
        const int b = c * 32 + lane;
        const uint32_t kq = (uint32_t)(lane >> 2), kr = (uint32_t)(lane & 3) * 4u;       // 16-byte chunk / byte within it
        // A: dz1[b][m] -> row m, k = lane.  Warp w covers units 32w .. 32w+31, 4 at a time.
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int m = pw * 32 + i * 4;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (b < B) v = __ldcg(reinterpret_cast<const float4*>(dz1 + (size_t)b * HID + m));
            const float e[4] = {to_tf32(v.x), to_tf32(v.y), to_tf32(v.z), to_tf32(v.w)};
#pragma unroll
            for (int j = 0; j < 4; ++j)
                *reinterpret_cast<float*>(At + (uint32_t)(m + j) * 128u + ((kq ^ (uint32_t)((m + j) & 7)) << 4) + kr) = e[j];
        }
        // B: X[b][n0 + n] -> row n, k = lane.  112 pixels = 7 x 16; warp w takes 16-pixel groups w, w+4
        int64_t sidx = -1;
        if (b < B) {
            const int64_t v = __ldg(idx + b);
            if (v >= 0 && v < n_samples) sidx = v;
        }
        for (int g = pw; g < DW_BN / 16; g += 4) {
            uint4 v = make_uint4(0, 0, 0, 0);
            if (sidx >= 0) v = __ldg(reinterpret_cast<const uint4*>(images + sidx * IN + n0 + g * 16));
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int n = g * 16 + q * 4 + j;
                    const float px = sidx >= 0 ? lut[(w[q] >> (8 * j)) & 0xff] : 0.f;
                    *reinterpret_cast<float*>(Bt + (uint32_t)n * 128u + ((kq ^ (uint32_t)(n & 7)) << 4) + kr) = px;
                }
        }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t a0 = smem_u32(smem);
        for (int i = 0; i < c1 - c0; ++i) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint64_t ad = umma_desc_sw128(a0 + i * (DW_A_BYTES + DW_B_BYTES) + j * 32);
                const uint64_t bd = umma_desc_sw128(a0 + i * (DW_A_BYTES + DW_B_BYTES) + DW_A_BYTES + j * 32);
                const uint32_t acc = (i | j) ? 1u : 0u;
                asm volatile(
                    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                    "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                    :: "r"(tmem_base), "l"(ad), "l"(bd), "r"(DW_IDESC), "r"(acc) : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                     :: "r"(smem_u32(mbar)) : "memory");
    }
    {
        uint32_t done = 0;
        while (!done) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\t"
                         "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(mbar)) : "memory");
        }
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // ---- epilogue (warps 0-3): TMEM lane = hidden unit m = tid; 112 columns = 7 loads of 16
    if (warp < 4) {
        float* dst = gpart + (size_t)slice * NP + DDPM_OFF_W1 + (size_t)tid * IN + n0;
#pragma unroll
        for (int cb = 0; cb < DW_BN; cb += 16) {
            uint32_t r[16];
            const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)cb;
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                : "r"(taddr) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int i = 0; i < 16; i += 4)
                *reinterpret_cast<float4*>(dst + cb + i) = make_float4(__uint_as_float(r[i]), __uint_as_float(r[i + 1]),
                                                                       __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" :: "r"(tmem_base) : "memory");
}

int launch_dw1_tc(const float* dz1, const uint8_t* images, int64_t n_samples, const int64_t* idx, int B, float* gpart,
                  cudaStream_t st) {
    static bool configured[64] = {};
    int dev = 0;
    DDPM_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        DDPM_CUDA(cudaFuncSetAttribute(dw1_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DW_SMEM));
        configured[dev] = true;
    }
    const int nslices = dw1_tc_slices(B);
    dw1_tc_kernel<<<DW_NT * nslices, DW_THREADS, DW_SMEM, st>>>(dz1, images, n_samples, idx, B, nslices, gpart);
    DDPM_CUDA(cudaGetLastError());
    return 0;
}

int launch_fc1_tc(const float* W1, const uint8_t* images, int64_t n_samples, const int64_t* idx, int B, float* part,
                  cudaStream_t st) {
    static bool configured[64] = {};
    int dev = 0;
    DDPM_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        DDPM_CUDA(cudaFuncSetAttribute(fc1_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM));
        configured[dev] = true;
    }
    const int mtiles = (B + TC_BM - 1) / TC_BM;
    fc1_fwd_tc_kernel<<<mtiles * TC_KSPLIT, TC_THREADS, TC_SMEM, st>>>(W1, images, n_samples, idx, B, part);
    DDPM_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace ddpm
