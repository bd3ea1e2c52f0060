// How long does it take N SMs to pull the SAME freshly written 64 KB (the fc2 weight matrix) out of L2 at the same
// moment?  (Round-2 design question for the persistent kernel's row CTAs.)  Variants: number of reading CTAs,
// register loads (16 x LDG.128.cg per thread) vs one TMA bulk copy per CTA, with / without the other CTAs polling L2.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o bcast_bench bcast_bench.cu && ./bcast_bench
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;
constexpr int N = 128 * 128;

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

template <int MODE>   // 0 = register loads, 1 = TMA bulk copy
__global__ void __launch_bounds__(256, 1) k(float* w, int readers, int pollers, volatile unsigned* flag, long long* out, int iters) {
    extern __shared__ __align__(128) unsigned char smem[];
    float* sw = reinterpret_cast<float*>(smem);
    unsigned long long* mbar = reinterpret_cast<unsigned long long*>(smem + N * 4);
    cg::grid_group grid = cg::this_grid();
    const int tid = threadIdx.x, cta = blockIdx.x;
    if (MODE == 1 && tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(mbar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    long long tot = 0;
    float sink = 0.f;
    for (int it = 0; it < iters; ++it) {
        // everybody rewrites a slice of w (fresh lines in L2, as after an SGD update)
        for (int i = cta * 256 + tid; i < N; i += gridDim.x * 256) w[i] = (float)(it + i);
        if (cta == 0 && tid == 0) *flag = 0;
        grid.sync();
        if (cta < readers) {
            long long t0 = clock64();
            if (MODE == 0) {
                float4 v[16];
                const int jb = tid >> 4, kb = tid & 15;
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    v[2 * r] = __ldcg(reinterpret_cast<const float4*>(w + (jb * 8 + r) * 128 + kb * 8));
                    v[2 * r + 1] = __ldcg(reinterpret_cast<const float4*>(w + (jb * 8 + r) * 128 + kb * 8) + 1);
                }
#pragma unroll
                for (int r = 0; r < 16; ++r) sink += v[r].x + v[r].y + v[r].z + v[r].w;
            } else {
                if (tid == 0) {
                    asm volatile("fence.proxy.async;" ::: "memory");
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(mbar)), "r"(N * 4) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 :: "r"(smem_u32(sw)), "l"(w), "r"(N * 4), "r"(smem_u32(mbar)) : "memory");
                }
                unsigned done;
                do {
                    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                                 : "=r"(done) : "r"(smem_u32(mbar)), "r"((unsigned)(it & 1)) : "memory");
                } while (!done);
                sink += sw[tid];
            }
            __syncthreads();
            long long t1 = clock64();
            if (tid == 0) { tot += t1 - t0; }
            __syncthreads();
            if (cta == 0 && tid == 0) { __threadfence(); *flag = 1; }
        } else if (cta < readers + pollers) {
            // background: poll 16-byte words of L2 like the waiting CTAs of the persistent kernel do
            const unsigned* p = reinterpret_cast<const unsigned*>(w) + ((cta * 256 + tid) * 4) % N;
            while (*flag == 0) {
                unsigned a, b, c, d;
                asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p) : "memory");
                sink += (float)(a & 1);
            }
        }
        grid.sync();
    }
    if (tid == 0 && cta < readers) out[cta] = tot / iters;
    if (sink == 12345.678f) out[200] = 1;
}

template <int MODE>
void run(int readers, int pollers) {
    float* w; unsigned* flag; long long* out;
    cudaMalloc(&w, N * 4); cudaMalloc(&flag, 4); cudaMalloc(&out, 256 * 8);
    cudaMemset(out, 0, 256 * 8);
    int iters = 20;
    size_t smem = N * 4 + 64;
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    void* args[] = {&w, &readers, &pollers, &flag, &out, &iters};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)k<MODE>, dim3(148), dim3(256), args, smem, 0);
    e = cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, out, 148 * 8, cudaMemcpyDeviceToHost);
    long long mx = 0, sum = 0;
    for (int i = 0; i < readers; ++i) { mx = h[i] > mx ? h[i] : mx; sum += h[i]; }
    printf("%s readers %3d pollers %3d: mean %6lld max %6lld cycles  (%s)\n", MODE ? "tma " : "regs", readers, pollers,
           sum / readers, mx, cudaGetErrorString(e));
    cudaFree(w); cudaFree(flag); cudaFree(out);
}

int main() {
    for (int pollers : {0, 84}) {
        for (int readers : {1, 8, 16, 32, 64}) { run<0>(readers, pollers); run<1>(readers, pollers); }
    }
    return 0;
}
