// Device arrays that grow in place (CUDA virtual memory management).
//
// faiss.IndexFlat keeps its rows in a std::vector: index.add (semantic-indexer/indexer.py:41, one row per call in
// the reference) grows it by reallocating and copying.  Doing the same in HBM means that a regrowth needs the old
// and the new storage resident at once -- a 77 GB index (BASELINE config 5 on one GPU) could not grow on a 180 GB
// part -- and costs a device-to-device copy of everything stored so far plus a stream wait.  Here the index
// reserves a virtual address range once, large enough for as many rows as the GPU could ever hold, and maps
// physical memory behind it in chunks as rows arrive: growth is a few driver calls, no copy, no wait, and the
// base addresses (hence the cached tensor maps) never change.
//
// The driver entry points are resolved at run time (the library does not link libcuda); when they are missing, or
// the reservation fails, the index falls back to cudaMalloc + copy (bfk_index.cu).
#include <cuda.h>

#include <cstdlib>
#include <vector>

#include "bfk_common.cuh"

namespace bfk {

struct VmApi {
    CUresult (*AddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*AddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*Create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*Release)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*Map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*Unmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*SetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*GetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
    bool ok = false;
};

static const VmApi& vm_api() {
    static const VmApi api = [] {
        VmApi a;
        const char* off = getenv("BFK_VMEM");
        if (off && atoi(off) == 0) return a;  // BFK_VMEM=0: cudaMalloc + copy, as in round 1
        auto get = [](const char* name, void** fn) {
            cudaDriverEntryPointQueryResult q;
            if (cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
                cudaGetLastError();
                *fn = nullptr;
            }
            return *fn != nullptr;
        };
        bool all = get("cuMemAddressReserve", (void**)&a.AddressReserve);
        all = get("cuMemAddressFree", (void**)&a.AddressFree) && all;
        all = get("cuMemCreate", (void**)&a.Create) && all;
        all = get("cuMemRelease", (void**)&a.Release) && all;
        all = get("cuMemMap", (void**)&a.Map) && all;
        all = get("cuMemUnmap", (void**)&a.Unmap) && all;
        all = get("cuMemSetAccess", (void**)&a.SetAccess) && all;
        all = get("cuMemGetAllocationGranularity", (void**)&a.GetGranularity) && all;
        a.ok = all;
        return a;
    }();
    return api;
}

struct VmArray {
    int device = 0;
    CUdeviceptr base = 0;
    size_t va_bytes = 0;   // reserved
    size_t mapped = 0;     // usable from base, a multiple of `gran`
    size_t gran = 0;
    std::vector<std::pair<CUmemGenericAllocationHandle, size_t>> chunks;  // physical pieces, in address order
};

static CUmemAllocationProp vm_prop(int device) {
    CUmemAllocationProp prop = {};
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = device;
    return prop;
}

// reserves `va_bytes` of address space on `device` (the device must be current); nothing is mapped yet
int vm_create(int device, size_t va_bytes, VmArray** out) {
    *out = nullptr;
    const VmApi& api = vm_api();
    if (!api.ok) return 1;
    const CUmemAllocationProp prop = vm_prop(device);
    size_t gran = 0;
    if (api.GetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM) != CUDA_SUCCESS || gran == 0) return 1;
    VmArray* a = new VmArray();
    a->device = device;
    a->gran = gran;
    a->va_bytes = (size_t)round_up((int64_t)std::max<size_t>(va_bytes, gran), (int64_t)gran);
    if (api.AddressReserve(&a->base, a->va_bytes, 0, 0, 0) != CUDA_SUCCESS) {
        delete a;
        return 1;
    }
    *out = a;
    return 0;
}

// makes at least `bytes` usable behind the base address: maps more physical memory, touches nothing already
// there.  Returns non-zero (and leaves the array as it was) when the GPU has no memory left for it.
int vm_grow(VmArray* a, size_t bytes) {
    if (bytes <= a->mapped) return 0;
    const VmApi& api = vm_api();
    const size_t target = (size_t)round_up((int64_t)bytes, (int64_t)a->gran);
    if (target > a->va_bytes) return 1;
    const CUmemAllocationProp prop = vm_prop(a->device);
    CUmemAccessDesc acc = {};
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = a->device;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    const size_t before = a->mapped;
    const size_t n_before = a->chunks.size();
    bool failed = false;
    while (a->mapped < target && !failed) {
        const size_t piece = std::min<size_t>(target - a->mapped, (size_t)1 << 30);  // <= 1 GiB per physical handle
        CUmemGenericAllocationHandle h;
        if (api.Create(&h, piece, &prop, 0) != CUDA_SUCCESS) {
            failed = true;
            break;
        }
        if (api.Map(a->base + a->mapped, piece, 0, h, 0) != CUDA_SUCCESS) {
            api.Release(h);
            failed = true;
            break;
        }
        if (api.SetAccess(a->base + a->mapped, piece, &acc, 1) != CUDA_SUCCESS) {
            api.Unmap(a->base + a->mapped, piece);
            api.Release(h);
            failed = true;
            break;
        }
        a->chunks.emplace_back(h, piece);
        a->mapped += piece;
    }
    if (failed) {  // undo this call's pieces
        while (a->chunks.size() > n_before) {
            const auto c = a->chunks.back();
            a->mapped -= c.second;
            api.Unmap(a->base + a->mapped, c.second);
            api.Release(c.first);
            a->chunks.pop_back();
        }
        a->mapped = before;
        return 1;
    }
    return 0;
}

void* vm_ptr(const VmArray* a) { return a ? reinterpret_cast<void*>(a->base) : nullptr; }
size_t vm_mapped(const VmArray* a) { return a ? a->mapped : 0; }

// the caller has made sure no work that touches the array is in flight
void vm_destroy(VmArray* a) {
    if (!a) return;
    const VmApi& api = vm_api();
    size_t off = 0;
    for (const auto& c : a->chunks) {
        api.Unmap(a->base + off, c.second);
        api.Release(c.first);
        off += c.second;
    }
    api.AddressFree(a->base, a->va_bytes);
    delete a;
}

}  // namespace bfk
