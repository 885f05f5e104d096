// Symmetric slab on the CUDA virtual-memory-management API, with an NVSwitch multicast (NVLS) view -- the library's OWN path to
// peer-mapped + multicast memory (cuMemCreate with a shareable handle, cuMemMap of every peer's allocation,
// cuMulticastCreate / AddDevice / BindMem).  Round 1 obtained the multicast mapping only through torch's symmetric-memory
// allocator; that remains the fallback, and cudaMalloc + CUDA IPC (no multicast) the last resort.
//
// One VmmSlab per rank.  The host side (parallel/native_comm.py) drives the collective protocol and moves the POSIX file
// descriptors between the processes (SCM_RIGHTS over a unix socket); this file only talks to the driver:
//   picd_vmm_create          allocate `bytes` of device memory as a shareable VMM allocation, map it, return base + export fd
//   picd_vmm_import_peer     import a peer's allocation from a (now local) fd and map it into this process
//   picd_vmm_mc_create       rank 0: create the multicast object for `world` devices, return its export fd
//   picd_vmm_mc_import       other ranks: import the multicast object from a local fd
//   picd_vmm_mc_add_device   every rank: add its device to the multicast object (all ranks must have done so before any bind)
//   picd_vmm_mc_bind_map     every rank: bind its allocation into the multicast object and map the multicast view
//   picd_vmm_destroy         unmap / release everything
// All driver entry points are resolved through cudaGetDriverEntryPoint, so the library keeps no link-time dependency on libcuda.
#include "../common/picd_common.cuh"

#include <unistd.h>

#include <cstring>

namespace picd {

constexpr int VMM_MAX_PEERS = 8;

struct VmmSlab {
    size_t bytes = 0;                    // padded to the allocation / multicast granularity
    int dev = 0;
    CUmemGenericAllocationHandle own = 0;
    CUdeviceptr own_va = 0;
    CUmemGenericAllocationHandle peer_h[VMM_MAX_PEERS] = {};
    CUdeviceptr peer_va[VMM_MAX_PEERS] = {};
    CUmemGenericAllocationHandle mc = 0;
    CUdeviceptr mc_va = 0;
    bool mc_bound = false;
};

struct DriverApi {
    CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
    CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
    CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                                 unsigned long long) = nullptr;
    CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
    CUresult (*DeviceGet)(CUdevice*, int) = nullptr;
    CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
    bool ok = false;
};

template <typename F>
static bool resolve(F& fn, const char* name) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || p == nullptr)
        return false;
    fn = reinterpret_cast<F>(p);
    return true;
}

static const DriverApi& driver() {
    static DriverApi api = [] {
        DriverApi a;
        a.ok = resolve(a.MemGetAllocationGranularity, "cuMemGetAllocationGranularity") && resolve(a.MemCreate, "cuMemCreate") &&
               resolve(a.MemExportToShareableHandle, "cuMemExportToShareableHandle") &&
               resolve(a.MemImportFromShareableHandle, "cuMemImportFromShareableHandle") &&
               resolve(a.MemAddressReserve, "cuMemAddressReserve") && resolve(a.MemMap, "cuMemMap") &&
               resolve(a.MemSetAccess, "cuMemSetAccess") && resolve(a.MemUnmap, "cuMemUnmap") &&
               resolve(a.MemRelease, "cuMemRelease") && resolve(a.MemAddressFree, "cuMemAddressFree") &&
               resolve(a.MulticastCreate, "cuMulticastCreate") && resolve(a.MulticastAddDevice, "cuMulticastAddDevice") &&
               resolve(a.MulticastBindMem, "cuMulticastBindMem") && resolve(a.MulticastGetGranularity, "cuMulticastGetGranularity") &&
               resolve(a.DeviceGet, "cuDeviceGet") && resolve(a.DeviceGetAttribute, "cuDeviceGetAttribute");
        return a;
    }();
    return api;
}

#define PICD_CU(expr)                                                                            \
    do {                                                                                         \
        CUresult _r = (expr);                                                                    \
        if (_r != CUDA_SUCCESS) {                                                                \
            ::picd::set_error("%s:%d: %s -> CUresult %d", __FILE__, __LINE__, #expr, (int)_r);   \
            return 1;                                                                            \
        }                                                                                        \
    } while (0)

static CUmemAllocationProp alloc_prop(int dev) {
    CUmemAllocationProp prop;
    std::memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = dev;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return prop;
}

static int map_rw(const DriverApi& d, CUdeviceptr* va, size_t bytes, size_t align, CUmemGenericAllocationHandle h, int dev) {
    PICD_CU(d.MemAddressReserve(va, bytes, align, 0, 0));
    PICD_CU(d.MemMap(*va, bytes, 0, h, 0));
    CUmemAccessDesc acc;
    std::memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = dev;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    PICD_CU(d.MemSetAccess(*va, bytes, &acc, 1));
    return 0;
}

}  // namespace picd

using namespace picd;

// Allocates `bytes` (rounded up to the larger of the allocation and the multicast granularity for `world` devices) on the
// current device as a VMM allocation exportable as a POSIX fd, maps and zeroes it.  *fd_out is the export fd (the caller closes
// it after the peers have received it); *padded_out the size every rank must use for imports and the multicast object.
extern "C" int picd_vmm_create(size_t bytes, int world, void** slab_out, void** base_out, int* fd_out, size_t* padded_out) {
    const DriverApi& d = driver();
    PICD_REQUIRE(d.ok, "vmm: the driver does not expose the cuMem* / cuMulticast* entry points");
    PICD_REQUIRE(world >= 1 && world <= VMM_MAX_PEERS, "vmm: world=%d unsupported", world);
    int dev = 0;
    PICD_CUDA(cudaGetDevice(&dev));
    PICD_CUDA(cudaFree(nullptr));                               // make sure the primary context is current
    CUmemAllocationProp prop = alloc_prop(dev);
    size_t gran = 0;
    PICD_CU(d.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    if (world > 1) {
        CUmulticastObjectProp mp;
        std::memset(&mp, 0, sizeof(mp));
        mp.numDevices = static_cast<unsigned>(world);
        mp.size = bytes;
        mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        size_t mgran = 0;
        if (d.MulticastGetGranularity(&mgran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mgran > gran) gran = mgran;
    }
    const size_t padded = (bytes + gran - 1) / gran * gran;
    VmmSlab* s = new VmmSlab();
    s->bytes = padded;
    s->dev = dev;
    if (d.MemCreate(&s->own, padded, &prop, 0) != CUDA_SUCCESS) {
        delete s;
        set_error("vmm: cuMemCreate(%zu bytes, POSIX fd handle) failed", padded);
        return 1;
    }
    if (map_rw(d, &s->own_va, padded, gran, s->own, dev)) return 1;
    PICD_CUDA(cudaMemset(reinterpret_cast<void*>(s->own_va), 0, padded));
    PICD_CUDA(cudaDeviceSynchronize());
    int fd = -1;
    PICD_CU(d.MemExportToShareableHandle(&fd, s->own, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    *slab_out = s;
    *base_out = reinterpret_cast<void*>(s->own_va);
    *fd_out = fd;
    *padded_out = padded;
    return 0;
}

// Imports the allocation behind `fd` (a descriptor of THIS process, received from peer `peer_rank`) and maps it read-write.
extern "C" int picd_vmm_import_peer(void* slab, int peer_rank, int fd, void** base_out) {
    const DriverApi& d = driver();
    VmmSlab* s = static_cast<VmmSlab*>(slab);
    PICD_REQUIRE(s != nullptr && peer_rank >= 0 && peer_rank < VMM_MAX_PEERS, "vmm_import_peer: bad arguments");
    PICD_CU(d.MemImportFromShareableHandle(&s->peer_h[peer_rank], reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                           CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    if (map_rw(d, &s->peer_va[peer_rank], s->bytes, 0, s->peer_h[peer_rank], s->dev)) return 1;
    *base_out = reinterpret_cast<void*>(s->peer_va[peer_rank]);
    return 0;
}

// 1 if the current device can take part in NVSwitch multicast objects.
extern "C" int picd_vmm_multicast_supported() {
    const DriverApi& d = driver();
    if (!d.ok) return 0;
    int dev = 0, val = 0;
    CUdevice cu;
    if (cudaGetDevice(&dev) != cudaSuccess || d.DeviceGet(&cu, dev) != CUDA_SUCCESS) return 0;
    if (d.DeviceGetAttribute(&val, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cu) != CUDA_SUCCESS) return 0;
    return val ? 1 : 0;
}

extern "C" int picd_vmm_mc_create(void* slab, int world, int* fd_out) {
    const DriverApi& d = driver();
    VmmSlab* s = static_cast<VmmSlab*>(slab);
    CUmulticastObjectProp mp;
    std::memset(&mp, 0, sizeof(mp));
    mp.numDevices = static_cast<unsigned>(world);
    mp.size = s->bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    PICD_CU(d.MulticastCreate(&s->mc, &mp));
    int fd = -1;
    PICD_CU(d.MemExportToShareableHandle(&fd, s->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    *fd_out = fd;
    return 0;
}

extern "C" int picd_vmm_mc_import(void* slab, int fd) {
    const DriverApi& d = driver();
    VmmSlab* s = static_cast<VmmSlab*>(slab);
    PICD_CU(d.MemImportFromShareableHandle(&s->mc, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                           CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    return 0;
}

extern "C" int picd_vmm_mc_add_device(void* slab) {
    const DriverApi& d = driver();
    VmmSlab* s = static_cast<VmmSlab*>(slab);
    CUdevice cu;
    PICD_CU(d.DeviceGet(&cu, s->dev));
    PICD_CU(d.MulticastAddDevice(s->mc, cu));
    return 0;
}

// Every rank, after ALL ranks have added their device: bind this rank's allocation at offset 0 of the multicast object and
// map the multicast view.  A multimem.* access through *mc_base_out then reaches every rank's copy of the slab.
extern "C" int picd_vmm_mc_bind_map(void* slab, void** mc_base_out) {
    const DriverApi& d = driver();
    VmmSlab* s = static_cast<VmmSlab*>(slab);
    PICD_CU(d.MulticastBindMem(s->mc, 0, s->own, 0, s->bytes, 0));
    s->mc_bound = true;
    if (map_rw(d, &s->mc_va, s->bytes, 0, s->mc, s->dev)) return 1;
    *mc_base_out = reinterpret_cast<void*>(s->mc_va);
    return 0;
}

extern "C" int picd_vmm_destroy(void* slab) {
    const DriverApi& d = driver();
    VmmSlab* s = static_cast<VmmSlab*>(slab);
    if (s == nullptr || !d.ok) return 0;
    cudaDeviceSynchronize();
    if (s->mc_va) { d.MemUnmap(s->mc_va, s->bytes); d.MemAddressFree(s->mc_va, s->bytes); }
    if (s->mc) d.MemRelease(s->mc);
    for (int i = 0; i < VMM_MAX_PEERS; ++i) {
        if (s->peer_va[i]) { d.MemUnmap(s->peer_va[i], s->bytes); d.MemAddressFree(s->peer_va[i], s->bytes); }
        if (s->peer_h[i]) d.MemRelease(s->peer_h[i]);
    }
    if (s->own_va) { d.MemUnmap(s->own_va, s->bytes); d.MemAddressFree(s->own_va, s->bytes); }
    if (s->own) d.MemRelease(s->own);
    delete s;
    return 0;
}
