// TMEM -> register load throughput per SM (tcgen05.ld shapes x warps), and MUFU.EX2 throughput.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/micro/tmem_probe.bin profiles/micro/tmem_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int X>
__device__ __forceinline__ void ldtm(uint32_t taddr, uint32_t* r);
template <>
__device__ __forceinline__ void ldtm<32>(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
template <>
__device__ __forceinline__ void ldtm<16>(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
template <>
__device__ __forceinline__ void ldtm<64>(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
        "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
        "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]), "=r"(r[32]),
          "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]), "=r"(r[40]),
          "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]), "=r"(r[48]),
          "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]), "=r"(r[56]),
          "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
        : "r"(taddr)
        : "memory");
}

template <int X>
__global__ void __launch_bounds__(512) tmem_kernel(int iters, unsigned long long* cyc, uint32_t* sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t r[X], acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int c = 0; c < 256; c += X) {
            ldtm<X>(base + ((warp >> 2) & 1) * 256 + c, r);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int i = 0; i < X; i++) acc ^= r[i];
        }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = (unsigned long long)(t1 - t0);
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(slot) : "memory");
}

// two loads in flight before the wait
template <int X>
__global__ void __launch_bounds__(512) tmem_kernel2(int iters, unsigned long long* cyc, uint32_t* sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t r[2][X], acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int c = 0; c < 256; c += 2 * X) {
            ldtm<X>(base + ((warp >> 2) & 1) * 256 + c, r[0]);
            ldtm<X>(base + ((warp >> 2) & 1) * 256 + c + X, r[1]);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int i = 0; i < X; i++) acc ^= r[0][i] ^ r[1][i];
        }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = (unsigned long long)(t1 - t0);
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(slot) : "memory");
}

__global__ void __launch_bounds__(512) ex2_kernel(int iters, unsigned long long* cyc, float* sink) {
    float x[16];
#pragma unroll
    for (int i = 0; i < 16; i++) x[i] = -0.001f * (threadIdx.x + i);
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = (unsigned long long)(t1 - t0);
    float s = 0.f;
    for (int i = 0; i < 16; i++) s += x[i];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static double run(F launch, unsigned long long* dc, int nblk) {
    launch();
    cudaDeviceSynchronize();
    launch();
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        printf("error %s\n", cudaGetErrorString(e));
        return -1;
    }
    unsigned long long h[148];
    cudaMemcpy(h, dc, sizeof(unsigned long long) * nblk, cudaMemcpyDeviceToHost);
    double s = 0;
    for (int i = 0; i < nblk; i++) s += (double)h[i];
    return s / nblk;
}

int main() {
    unsigned long long* dc;
    uint32_t* sink;
    cudaMalloc(&dc, 148 * 8);
    cudaMalloc(&sink, 148 * 512 * 4);
    const int iters = 2000;
    for (int warps : {1, 4, 8, 16}) {
        const double bytes = (double)iters * warps * 32 * 256 * 4;
        double c;
        c = run([&] { tmem_kernel<16><<<148, warps * 32>>>(iters, dc, sink); }, dc, 148);
        printf("tcgen05.ld 32x32b.x16, %2d warps, wait after each      : %7.1f B/clk/SM\n", warps, bytes / c);
        c = run([&] { tmem_kernel<32><<<148, warps * 32>>>(iters, dc, sink); }, dc, 148);
        printf("tcgen05.ld 32x32b.x32, %2d warps, wait after each      : %7.1f B/clk/SM\n", warps, bytes / c);
        c = run([&] { tmem_kernel<64><<<148, warps * 32>>>(iters, dc, sink); }, dc, 148);
        printf("tcgen05.ld 32x32b.x64, %2d warps, wait after each      : %7.1f B/clk/SM\n", warps, bytes / c);
        c = run([&] { tmem_kernel2<32><<<148, warps * 32>>>(iters, dc, sink); }, dc, 148);
        printf("tcgen05.ld 32x32b.x32, %2d warps, two loads per wait   : %7.1f B/clk/SM\n", warps, bytes / c);
    }
    for (int warps : {4, 8, 16}) {
        const double n = (double)iters * warps * 32 * 16;
        double c = run([&] { ex2_kernel<<<148, warps * 32>>>(iters, dc, (float*)sink); }, dc, 148);
        printf("ex2.approx.ftz.f32, %2d warps                           : %7.2f results/clk/SM\n", warps, n / c);
    }
    return 0;
}
