// ubench.cu — instruction-throughput and VMM probes used to choose designs (run on the B200 via gpurun).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench ubench.cu
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

constexpr int ITERS = 4096, ILP = 8;

template <int OP>
__global__ void __launch_bounds__(1024) k_op(float *out, int n_iter, float seed) {
    float f[ILP]; double d[ILP]; int i32[ILP]; unsigned short h[ILP];
#pragma unroll
    for (int j = 0; j < ILP; j++) { f[j] = seed + j + threadIdx.x; d[j] = seed + j; i32[j] = (int)seed + j + threadIdx.x; h[j] = (unsigned short)(0x3c00 + j + threadIdx.x); }
    for (int it = 0; it < n_iter; it++) {
#pragma unroll
        for (int j = 0; j < ILP; j++) {
            if (OP == 0) f[j] = fmaf(f[j], 1.0001f, 0.5f);                       // FFMA
            if (OP == 1) d[j] = d[j] + 1.0000001;                                 // DADD
            if (OP == 2) d[j] = fma(d[j], 1.0000001, 0.5);                        // DFMA
            if (OP == 3) { d[j] += (double)f[j]; }                                // F2F.F64.F32 + DADD
            if (OP == 4) { double t; asm volatile("cvt.f64.f16 %0, %1;" : "=d"(t) : "h"(h[j])); d[j] += t; }  // F2F.F64.F16 + DADD
            if (OP == 5) { f[j] += (float)i32[j]; i32[j] += 3; }                  // I2F + FADD + IADD
            if (OP == 6) { f[j] = (float)d[j]; d[j] += 1.0; }                     // F2F.F32.F64 + DADD
            if (OP == 7) { i32[j] = __dp4a(i32[j], 0x01020304, i32[j]); }         // IDP4A
            if (OP == 8) { f[j] = __half2float(__ushort_as_half(h[j])) + f[j]; }  // F2F.F32.F16 (HADD2.F32) + FADD
            if (OP == 9) { i32[j] = i32[j] * 3 + 7; }                             // IMAD
            if (OP == 10) { f[j] = __int_as_float(0x4B400000 + i32[j]) - 12582912.0f + f[j]; i32[j] ^= 5; } // magic I2F
        }
    }
    float s = 0; 
#pragma unroll
    for (int j = 0; j < ILP; j++) s += f[j] + (float)d[j] + (float)i32[j] + (float)h[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
static void run(const char *name, int extra_ops) {
    int dev; CK(cudaGetDevice(&dev));
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, dev));
    const int grid = p.multiProcessorCount * 2, block = 1024;
    float *out; CK(cudaMalloc(&out, (size_t)grid * block * 4));
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    k_op<OP><<<grid, block>>>(out, 16, 1.0f);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a));
    k_op<OP><<<grid, block>>>(out, ITERS, 1.0f);
    CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    const double ops = (double)grid * block * ITERS * ILP;
    int clk_khz; CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, dev));
    printf("%-34s %8.3f ms  %8.1f Gop/s  %6.1f op/clk/SM @%d MHz nominal (%d extra op(s) per iteration)\n", name, ms, ops / ms / 1e6,
           ops / (ms * 1e-3) / p.multiProcessorCount / (clk_khz * 1e3), clk_khz / 1000, extra_ops);
    cudaFree(out);
}

static void vmm_probe() {
    printf("---- VMM probe\n");
    CUresult (*pGran)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*pReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*pCreate)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long) = nullptr;
    CUresult (*pMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*pSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t) = nullptr;
    CUresult (*pUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*pRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*pFree)(CUdeviceptr, size_t) = nullptr;
    cudaDriverEntryPointQueryResult qr;
#define GET(ptr, name) do { void *f = nullptr; cudaError_t e = cudaGetDriverEntryPoint(name, &f, cudaEnableDefault, &qr); \
        if (e != cudaSuccess || !f) { printf("  %s: not found (%s)\n", name, cudaGetErrorString(e)); return; } *(void **)(&ptr) = f; } while (0)
    GET(pGran, "cuMemGetAllocationGranularity"); GET(pReserve, "cuMemAddressReserve"); GET(pCreate, "cuMemCreate");
    GET(pMap, "cuMemMap"); GET(pSetAccess, "cuMemSetAccess"); GET(pUnmap, "cuMemUnmap"); GET(pRelease, "cuMemRelease");
    GET(pFree, "cuMemAddressFree");
    int dev; CK(cudaGetDevice(&dev));
    CUmemAllocationProp prop = {};
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = dev;
    size_t gran = 0;
    CUresult r = pGran(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM);
    printf("  granularity (minimum): rc %d, %zu bytes\n", (int)r, gran);
    size_t gran2 = 0; pGran(&gran2, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
    printf("  granularity (recommended): %zu bytes\n", gran2);
    if (r != CUDA_SUCCESS || !gran) return;
    const int n_pages = 64;
    CUdeviceptr va = 0;
    r = pReserve(&va, gran * n_pages, 0, 0, 0);
    printf("  reserve %d pages: rc %d va %p\n", n_pages, (int)r, (void *)va);
    if (r != CUDA_SUCCESS) return;
    CUmemGenericAllocationHandle hnd[n_pages];
    CUmemAccessDesc acc = {};
    acc.location = prop.location; acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int i = 0; i < n_pages; i++) { r = pCreate(&hnd[i], gran, &prop, 0); if (r) { printf("  create rc %d at %d\n", (int)r, i); return; } }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    printf("  cuMemCreate: %.1f us per page\n", ((t1.tv_sec - t0.tv_sec) * 1e9 + (t1.tv_nsec - t0.tv_nsec)) / 1e3 / n_pages);
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int i = 0; i < n_pages; i++) {
        r = pMap(va + (size_t)i * gran, gran, 0, hnd[i], 0); if (r) { printf("  map rc %d\n", (int)r); return; }
        r = pSetAccess(va + (size_t)i * gran, gran, &acc, 1); if (r) { printf("  setaccess rc %d\n", (int)r); return; }
    }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    printf("  cuMemMap + cuMemSetAccess: %.1f us per page\n", ((t1.tv_sec - t0.tv_sec) * 1e9 + (t1.tv_nsec - t0.tv_nsec)) / 1e3 / n_pages);
    CK(cudaMemset((void *)va, 0x5a, gran * n_pages));
    unsigned char hb[4];
    CK(cudaMemcpy(hb, (void *)(va + gran * (n_pages - 1) + 100), 4, cudaMemcpyDeviceToHost));
    printf("  write/read through the mapping: %02x (want 5a)\n", hb[0]);
    // map the same physical page at a second address (prefix sharing)
    CUdeviceptr va2 = 0;
    r = pReserve(&va2, gran, 0, 0, 0);
    r = pMap(va2, gran, 0, hnd[3], 0);
    printf("  second mapping of one handle: rc %d\n", (int)r);
    if (!r) { pSetAccess(va2, gran, &acc, 1); CK(cudaMemcpy(hb, (void *)va2, 1, cudaMemcpyDeviceToHost)); printf("  aliased read: %02x\n", hb[0]); pUnmap(va2, gran); }
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int i = 0; i < n_pages; i++) pUnmap(va + (size_t)i * gran, gran);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    printf("  cuMemUnmap: %.1f us per page\n", ((t1.tv_sec - t0.tv_sec) * 1e9 + (t1.tv_nsec - t0.tv_nsec)) / 1e3 / n_pages);
    for (int i = 0; i < n_pages; i++) pRelease(hnd[i]);
    pFree(va, gran * n_pages); pFree(va2, gran);
    printf("  VMM ok\n");
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    printf("%s, %d SMs, L2 %d MB, smem/SM %zu\n", p.name, p.multiProcessorCount, p.l2CacheSize >> 20, p.sharedMemPerMultiprocessor);
    run<0>("FFMA", 0); run<1>("DADD", 0); run<2>("DFMA", 0); run<3>("F2F.F64.F32 + DADD", 1); run<4>("F2F.F64.F16 + DADD", 1);
    run<5>("I2F.F32.S32 + FADD + IADD", 2); run<6>("F2F.F32.F64 + DADD", 1); run<7>("IDP4A", 0); run<8>("F2F.F32.F16 + FADD", 1);
    run<9>("IMAD", 0); run<10>("magic I2F (IADD+FADD) + FADD + LOP", 3);
    vmm_probe();
    return 0;
}
