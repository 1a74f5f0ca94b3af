// kvpool.h — paged KV-cache memory (north_star "KV-cache allocator").
//
// llama-server sizes its KV cache once, at start, for --ctx-size x --parallel slots (the reference passes
// --ctx-size only, runtime/src/model_manager.rs:187-199).  With 64 concurrent agent requests per GPU most of such a
// cache would sit empty: 64 x 4096 positions of Mistral-7B are 34 GB.  Here the cache of a model is a POOL of
// fixed-size physical pages (8 MB), and every sequence owns a contiguous VIRTUAL range big enough for its whole
// context into which pages are mapped as its position grows (CUDA virtual memory management: cuMemCreate /
// cuMemAddressReserve / cuMemMap).  The sequence's page table is the list of handles mapped into its range; the
// GPU's own MMU does the translation, so the attention kernels keep plain `base + position * stride` addressing and
// CUDA graphs keep stable pointers.  Pages return to the pool when a sequence is reset or freed.
//
// Layout inside a range is POSITION-MAJOR: [position][layer][K|V][n_kv*head_dim] f16, so a page holds whole
// positions (16 per 2 MB for Mistral-7B) of every layer and a sequence of n positions holds ceil(n * bytes_per_position
// / page) pages, whatever the layer count.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <vector>

namespace aios {

struct KvPool {
    int device = 0;
    size_t page_bytes = 0;     // multiple of the allocation granularity
    size_t max_pages = 0;      // cap on pages ever created (0 = until the device runs out)
    size_t n_created = 0;
    std::vector<CUmemGenericAllocationHandle> free_pages;
    std::mutex mu;

    KvPool() = default;
    ~KvPool();
    void init(int device, size_t want_page_bytes, size_t max_bytes);
    CUmemGenericAllocationHandle take();           // a free page, or a newly created one; throws CudaError when exhausted
    void give(CUmemGenericAllocationHandle h);
    size_t pages_in_use();
    // pages that can still be handed out without exceeding the cap (SIZE_MAX when uncapped)
    size_t pages_available();
};

// one sequence's virtual range + the pages mapped into it
struct KvSpace {
    KvPool *pool = nullptr;
    CUdeviceptr va = 0;
    size_t va_bytes = 0;
    std::vector<CUmemGenericAllocationHandle> pages;  // page i is mapped at va + i * page_bytes

    KvSpace() = default;
    ~KvSpace();
    KvSpace(const KvSpace &) = delete;
    KvSpace &operator=(const KvSpace &) = delete;
    void init(KvPool *pool, size_t bytes);   // reserve the range (no memory yet)
    void ensure(size_t bytes);               // map pages until [0, bytes) is backed
    void trim(size_t bytes);                 // unmap pages wholly beyond `bytes` (0: all) and return them to the pool
    size_t mapped_bytes() const { return pages.size() * (pool ? pool->page_bytes : 0); }
    template <typename T>
    T *as() const { return reinterpret_cast<T *>(va); }
};

}  // namespace aios
