// Microbenchmark: cost of one grid-wide barrier on B200 for 148 co-resident CTAs x 256 threads,
// several implementations.  Each iteration every thread does one global store (so the release
// has something to publish), then the barrier.  Prints cycles per barrier (clock64 on CTA 0).
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long* p) {
    unsigned long long v; asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ unsigned long long ld_acquire(const unsigned long long* p) {
    unsigned long long v; asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }

template <int MODE>
__global__ void __launch_bounds__(256, 1) k(unsigned long long* ctr, float* buf, long long* out, int iters) {
    cg::grid_group grid = cg::this_grid();
    unsigned long long target = 0;
    const int tid = threadIdx.x;
    long long t0 = 0;
    for (int it = 0; it < iters + 10; ++it) {
        if (it == 10 && tid == 0) t0 = clock64();
        buf[(size_t)blockIdx.x * 256 + tid] = (float)it;                 // something to publish
        if (MODE == 0) {                                                   // red.release + relaxed poll + fence
            __syncthreads();
            if (tid == 255) {
                target += gridDim.x;
                asm volatile("red.release.gpu.global.add.u64 [%0], 1;" :: "l"(ctr) : "memory");
                while (ld_relaxed(ctr) < target) { }
                asm volatile("fence.acq_rel.gpu;" ::: "memory");
            }
            __syncthreads();
        } else if (MODE == 1) {                                            // red.release + acquire poll
            __syncthreads();
            if (tid == 255) {
                target += gridDim.x;
                asm volatile("red.release.gpu.global.add.u64 [%0], 1;" :: "l"(ctr) : "memory");
                while (ld_acquire(ctr) < target) { }
            }
            __syncthreads();
        } else if (MODE == 2) {                                            // cooperative groups
            grid.sync();
        } else if (MODE == 3) {                                            // __threadfence + atomicAdd + volatile poll
            __syncthreads();
            if (tid == 255) {
                target += gridDim.x;
                __threadfence();
                atomicAdd(ctr, 1ull);
                while (*((volatile unsigned long long*)ctr) < target) { }
                __threadfence();
            }
            __syncthreads();
        } else if (MODE == 4) {                                            // relaxed red, no fences at all (lower bound, NOT correct)
            __syncthreads();
            if (tid == 255) {
                target += gridDim.x;
                asm volatile("red.relaxed.gpu.global.add.u64 [%0], 1;" :: "l"(ctr) : "memory");
                while (ld_relaxed(ctr) < target) { }
            }
            __syncthreads();
        } else if (MODE == 5) {                                            // per-CTA flags: each CTA sets its own word, CTA polls all (no atomics)
            __syncthreads();
            target += 1;
            if (tid == 255) asm volatile("st.release.gpu.global.u64 [%0], %1;" :: "l"(ctr + 16 * (blockIdx.x + 1)), "l"(target) : "memory");
            if (tid < gridDim.x) { while (ld_relaxed(ctr + 16 * (tid + 1)) < target) { } }
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
            __syncthreads();
        } else if (MODE == 6 || MODE == 7) {                               // 2 / 4 counters (CTA index mod), arrive on one, poll all
            constexpr int NC = MODE == 6 ? 2 : 4;
            __syncthreads();
            if (tid == 255) {
                target += 1;                                               // every counter reaches (#CTAs mapped to it) x target
                asm volatile("red.release.gpu.global.add.u64 [%0], 1;" :: "l"(ctr + 16 * (blockIdx.x % NC)) : "memory");
                bool done = false;
                while (!done) {
                    unsigned long long v[NC];
#pragma unroll
                    for (int c = 0; c < NC; ++c) v[c] = ld_acquire(ctr + 16 * c);
                    done = true;
#pragma unroll
                    for (int c = 0; c < NC; ++c) {
                        const unsigned long long members = (gridDim.x - c + NC - 1) / NC;
                        done = done && (v[c] >= members * target);
                    }
                }
            }
            __syncthreads();
        }
    }
    if (blockIdx.x == 0 && tid == 0) out[0] = (clock64() - t0) / iters;
}

template <int MODE> void run(const char* name, unsigned long long* ctr, float* buf, long long* out) {
    cudaMemset(ctr, 0, 16 * 256 * 8);
    int iters = 2000;
    void* args[] = {&ctr, &buf, &out, &iters};
    cudaLaunchCooperativeKernel((void*)k<MODE>, dim3(148), dim3(256), args, 0, 0);
    cudaError_t e = cudaDeviceSynchronize();
    long long h = 0;
    cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
    printf("%-55s %6lld cycles/barrier  (%s)\n", name, h, cudaGetErrorString(e));
}

int main() {
    unsigned long long* ctr; float* buf; long long* out;
    cudaMalloc(&ctr, 16 * 256 * 8); cudaMalloc(&buf, 148 * 256 * 4); cudaMalloc(&out, 8);
    run<0>("red.release + relaxed poll + fence.acq_rel", ctr, buf, out);
    run<1>("red.release + acquire poll", ctr, buf, out);
    run<2>("cooperative_groups grid.sync()", ctr, buf, out);
    run<3>("__threadfence + atomicAdd + volatile poll + __threadfence", ctr, buf, out);
    run<4>("relaxed red + relaxed poll, no fences (lower bound)", ctr, buf, out);
    run<5>("per-CTA release flags, 148 threads poll one flag each", ctr, buf, out);
    run<6>("red.release on 2 counters (CTA parity), acquire-poll both", ctr, buf, out);
    run<7>("red.release on 4 counters, acquire-poll all four", ctr, buf, out);
    return 0;
}
