// kvpool.cu — see kvpool.h.  The driver entry points are resolved through the runtime (cudaGetDriverEntryPoint), so the
// library has no link-time dependency on libcuda and still loads on a machine without a driver.
#include "kvpool.h"

#include <string.h>

#include <string>

#include "common.cuh"

namespace aios {

namespace {
struct Vmm {
    CUresult (*GetGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*AddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*AddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*Create)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long) = nullptr;
    CUresult (*Release)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*Map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*Unmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*SetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t) = nullptr;
    CUresult (*GetErrorString)(CUresult, const char **) = nullptr;
};

const Vmm &vmm() {
    static Vmm v;
    static std::once_flag once;
    std::call_once(once, [] {
        auto get = [](const char *name, void *slot) {
            void *f = nullptr;
            cudaDriverEntryPointQueryResult qr;
            cudaError_t e = cudaGetDriverEntryPoint(name, &f, cudaEnableDefault, &qr);
            if (e != cudaSuccess || !f) {
                cudaGetLastError();
                throw CudaError(std::string("driver entry point ") + name + " is not available");
            }
            memcpy(slot, &f, sizeof(f));
        };
        get("cuMemGetAllocationGranularity", &v.GetGranularity);
        get("cuMemAddressReserve", &v.AddressReserve);
        get("cuMemAddressFree", &v.AddressFree);
        get("cuMemCreate", &v.Create);
        get("cuMemRelease", &v.Release);
        get("cuMemMap", &v.Map);
        get("cuMemUnmap", &v.Unmap);
        get("cuMemSetAccess", &v.SetAccess);
        get("cuGetErrorString", &v.GetErrorString);
    });
    return v;
}

void cu_check(CUresult r, const char *what) {
    if (r == CUDA_SUCCESS) return;
    const char *msg = nullptr;
    if (vmm().GetErrorString) vmm().GetErrorString(r, &msg);
    throw CudaError(std::string(what) + ": " + (msg ? msg : "CUDA driver error") + " (" + std::to_string((int)r) + ")");
}

CUmemAllocationProp device_prop(int device) {
    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = device;
    return prop;
}
}  // namespace

void KvPool::init(int dev, size_t want_page_bytes, size_t max_bytes) {
    device = dev;
    const CUmemAllocationProp prop = device_prop(dev);
    size_t gran = 0;
    cu_check(vmm().GetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM), "cuMemGetAllocationGranularity");
    if (!gran) throw CudaError("KV pool: zero allocation granularity");
    page_bytes = (want_page_bytes + gran - 1) / gran * gran;
    max_pages = max_bytes ? (max_bytes + page_bytes - 1) / page_bytes : 0;
}

KvPool::~KvPool() {
    for (auto h : free_pages) vmm().Release(h);
}

CUmemGenericAllocationHandle KvPool::take() {
    std::lock_guard<std::mutex> lk(mu);
    if (!free_pages.empty()) {
        CUmemGenericAllocationHandle h = free_pages.back();
        free_pages.pop_back();
        return h;
    }
    if (max_pages && n_created >= max_pages) throw CudaError("KV pool exhausted (the cap on KV pages is reached)");
    const CUmemAllocationProp prop = device_prop(device);
    CUmemGenericAllocationHandle h;
    CUresult r = vmm().Create(&h, page_bytes, &prop, 0);
    if (r == CUDA_ERROR_OUT_OF_MEMORY) throw CudaError("KV pool exhausted (device memory is full)");
    cu_check(r, "cuMemCreate");
    n_created++;
    return h;
}

void KvPool::give(CUmemGenericAllocationHandle h) {
    std::lock_guard<std::mutex> lk(mu);
    free_pages.push_back(h);
}

size_t KvPool::pages_in_use() {
    std::lock_guard<std::mutex> lk(mu);
    return n_created - free_pages.size();
}

size_t KvPool::pages_available() {
    std::lock_guard<std::mutex> lk(mu);
    if (!max_pages) return SIZE_MAX;
    return free_pages.size() + (max_pages - n_created);
}

void KvSpace::init(KvPool *p, size_t bytes) {
    pool = p;
    va_bytes = (bytes + p->page_bytes - 1) / p->page_bytes * p->page_bytes;
    cu_check(vmm().AddressReserve(&va, va_bytes, 0, 0, 0), "cuMemAddressReserve");
}

KvSpace::~KvSpace() {
    if (!pool) return;
    try { trim(0); } catch (...) {}
    if (va) vmm().AddressFree(va, va_bytes);
}

void KvSpace::ensure(size_t bytes) {
    if (bytes > va_bytes) throw CudaError("KV range overflow");
    const size_t pb = pool->page_bytes;
    while (pages.size() * pb < bytes) {
        CUmemGenericAllocationHandle h = pool->take();
        const CUdeviceptr at = va + pages.size() * pb;
        CUresult r = vmm().Map(at, pb, 0, h, 0);
        if (r != CUDA_SUCCESS) { pool->give(h); cu_check(r, "cuMemMap"); }
        CUmemAccessDesc acc;
        memset(&acc, 0, sizeof(acc));
        acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc.location.id = pool->device;
        acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        r = vmm().SetAccess(at, pb, &acc, 1);
        if (r != CUDA_SUCCESS) { vmm().Unmap(at, pb); pool->give(h); cu_check(r, "cuMemSetAccess"); }
        pages.push_back(h);
    }
}

void KvSpace::trim(size_t bytes) {
    const size_t pb = pool->page_bytes;
    const size_t keep = (bytes + pb - 1) / pb;
    while (pages.size() > keep) {
        const CUdeviceptr at = va + (pages.size() - 1) * pb;
        cu_check(vmm().Unmap(at, pb), "cuMemUnmap");
        pool->give(pages.back());
        pages.pop_back();
    }
}

}  // namespace aios
