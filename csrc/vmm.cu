// Symmetric memory owned by this library: CUDA virtual-memory-management allocations that every
// rank of the NVSwitch domain maps, plus the NVLS multicast object bound over them.
//
// SURVEY §5 "Symmetric arena": "each rank allocates its arenas with the CUDA VMM API (cuMemCreate
// with a shareable handle), every peer's arena is mapped into every rank's VA (cuMemMap +
// cuMemSetAccess), a multicast object (cuMulticastCreate / AddDevice / BindMem) over the gradient
// arena gives the NVLS path".  Replaces what the reference gets from c10d behind
// `dist.init_process_group` / `DDP(model)` (ref train.py:71,233) and what round 1 borrowed from
// torch.distributed._symmetric_memory for the multicast arena.
//
// The handles cross process boundaries as POSIX file descriptors; sending a descriptor to another
// process (SCM_RIGHTS over an AF_UNIX socket) is host plumbing done by the caller
// (b200ddp/parallel/symm.py).  Everything that touches the driver is here:
//
//   b200_vmm_create        cuMemCreate (device-local, fd-exportable) + reserve + map + access
//   b200_vmm_export        cuMemExportToShareableHandle                       -> fd
//   b200_vmm_import        cuMemImportFromShareableHandle + reserve + map + access (a peer's memory)
//   b200_mc_create         cuMulticastCreate for `ranks` devices               -> object + fd  (one rank)
//   b200_mc_import         the other ranks' handle on that object
//   b200_mc_add_device     cuMulticastAddDevice        (every rank, before anyone binds)
//   b200_mc_bind_map       cuMulticastBindMem of the local allocation + map the object -> multicast VA
//
// Driver entry points are resolved with cudaGetDriverEntryPoint (the library links cudart only).
#include "common.cuh"

#include <mutex>
#include <new>
#include <unistd.h>

namespace {

struct Drv {
    CUresult (*GetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
    CUresult (*Create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
    CUresult (*Release)(CUmemGenericAllocationHandle);
    CUresult (*AddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
    CUresult (*AddressFree)(CUdeviceptr, size_t);
    CUresult (*Map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
    CUresult (*Unmap)(CUdeviceptr, size_t);
    CUresult (*SetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
    CUresult (*ExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
    CUresult (*ImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                                 unsigned long long);
    CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
    CUresult (*DeviceGet)(CUdevice*, int);
    CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
    CUresult (*GetErrorString)(CUresult, const char**);
    bool ok;
};

template <typename F>
bool load(F& fn, const char* name) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p)
        return false;
    fn = reinterpret_cast<F>(p);
    return true;
}

const Drv& drv() {
    static Drv d{};
    static std::once_flag once;
    std::call_once(once, [] {
        d.ok = load(d.GetAllocationGranularity, "cuMemGetAllocationGranularity") && load(d.Create, "cuMemCreate") &&
               load(d.Release, "cuMemRelease") && load(d.AddressReserve, "cuMemAddressReserve") &&
               load(d.AddressFree, "cuMemAddressFree") && load(d.Map, "cuMemMap") && load(d.Unmap, "cuMemUnmap") &&
               load(d.SetAccess, "cuMemSetAccess") && load(d.ExportToShareableHandle, "cuMemExportToShareableHandle") &&
               load(d.ImportFromShareableHandle, "cuMemImportFromShareableHandle") &&
               load(d.MulticastCreate, "cuMulticastCreate") && load(d.MulticastAddDevice, "cuMulticastAddDevice") &&
               load(d.MulticastBindMem, "cuMulticastBindMem") &&
               load(d.MulticastGetGranularity, "cuMulticastGetGranularity") && load(d.DeviceGet, "cuDeviceGet") &&
               load(d.DeviceGetAttribute, "cuDeviceGetAttribute") && load(d.GetErrorString, "cuGetErrorString");
    });
    return d;
}

const char* cu_err(CUresult r) {
    const char* s = nullptr;
    if (drv().ok && drv().GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
    return "?";
}

#define B200_CU_OK(expr)                                                                         \
    do {                                                                                         \
        CUresult _r = (expr);                                                                    \
        if (_r != CUDA_SUCCESS) {                                                                \
            b200_set_error("%s:%d: %s -> %s (%d)", __FILE__, __LINE__, #expr, cu_err(_r), (int)_r); \
            return 1000 + (int)_r;                                                               \
        }                                                                                        \
    } while (0)

// one mapping: a physical allocation (ours or a peer's) or a multicast object, at a reserved VA
struct Mapping {
    CUmemGenericAllocationHandle handle;
    CUdeviceptr va;
    size_t bytes;
    bool mapped;
};

int current_device(int* dev, CUdevice* cudev) {
    B200_CUDA_OK(cudaGetDevice(dev));
    B200_CUDA_OK(cudaFree(0));   // the primary context must exist before driver calls
    B200_CU_OK(drv().DeviceGet(cudev, *dev));
    return 0;
}

CUmemAllocationProp local_prop(int dev) {
    CUmemAllocationProp prop{};
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = dev;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return prop;
}

int map_handle(Mapping* m, int dev) {
    B200_CU_OK(drv().AddressReserve(&m->va, m->bytes, 0, 0, 0));
    B200_CU_OK(drv().Map(m->va, m->bytes, 0, m->handle, 0));
    m->mapped = true;
    CUmemAccessDesc acc{};
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = dev;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    B200_CU_OK(drv().SetAccess(m->va, m->bytes, &acc, 1));
    return 0;
}

}  // namespace

// out[0] = VMM with fd handles usable, out[1] = multicast supported by the current device
extern "C" int b200_vmm_support(int* out2) {
    B200_REQUIRE(out2 != nullptr, "vmm_support: null argument");
    out2[0] = out2[1] = 0;
    if (!drv().ok) return 0;
    int dev;
    CUdevice cudev;
    if (int rc = current_device(&dev, &cudev)) return rc;
    int v = 0;
    if (drv().DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, cudev) == CUDA_SUCCESS)
        out2[0] = v;
    v = 0;
    if (drv().DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cudev) == CUDA_SUCCESS) out2[1] = v;
    return 0;
}

// Size every rank must allocate so that the memory can be bound to a multicast object of `ranks`
// devices (the larger of the allocation and the multicast granularity); ranks <= 1: allocation only.
extern "C" int b200_vmm_padded_size(int64_t bytes, int ranks, int64_t* out) {
    B200_REQUIRE(drv().ok, "vmm: the CUDA driver lacks the virtual memory management entry points");
    B200_REQUIRE(bytes > 0 && out, "vmm_padded_size: bad arguments");
    int dev;
    CUdevice cudev;
    if (int rc = current_device(&dev, &cudev)) return rc;
    const CUmemAllocationProp prop = local_prop(dev);
    size_t g = 0;
    B200_CU_OK(drv().GetAllocationGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM));
    if (ranks > 1) {
        CUmulticastObjectProp mp{};
        mp.numDevices = (unsigned)ranks;
        mp.size = (size_t)bytes;
        mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        size_t mg = 0;
        if (drv().MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_MINIMUM) == CUDA_SUCCESS && mg > g) g = mg;
    }
    *out = (int64_t)(((size_t)bytes + g - 1) / g * g);
    return 0;
}

// Allocates `bytes` (already padded) of device memory, maps it, zero-fills it.  *mapping is the
// object the other calls take; *ptr the device address.
extern "C" int b200_vmm_create(int64_t bytes, void** mapping, void** ptr) {
    B200_REQUIRE(drv().ok, "vmm: the CUDA driver lacks the virtual memory management entry points");
    B200_REQUIRE(bytes > 0 && mapping && ptr, "vmm_create: bad arguments");
    int dev;
    CUdevice cudev;
    if (int rc = current_device(&dev, &cudev)) return rc;
    Mapping* m = new (std::nothrow) Mapping{};
    B200_REQUIRE(m != nullptr, "vmm_create: out of host memory");
    m->bytes = (size_t)bytes;
    const CUmemAllocationProp prop = local_prop(dev);
    CUresult r = drv().Create(&m->handle, m->bytes, &prop, 0);
    if (r != CUDA_SUCCESS) {
        delete m;
        b200_set_error("vmm_create: cuMemCreate(%lld bytes) -> %s", (long long)bytes, cu_err(r));
        return 1000 + (int)r;
    }
    if (int rc = map_handle(m, dev)) return rc;
    B200_CUDA_OK(cudaMemset((void*)m->va, 0, m->bytes));
    B200_CUDA_OK(cudaDeviceSynchronize());
    *mapping = m;
    *ptr = (void*)m->va;
    return 0;
}

// POSIX file descriptor of the allocation (or multicast object) behind `mapping`; the caller sends
// it to the peers and closes it.
extern "C" int b200_vmm_export(void* mapping, int* fd) {
    Mapping* m = (Mapping*)mapping;
    B200_REQUIRE(m && fd, "vmm_export: bad arguments");
    int out = -1;
    B200_CU_OK(drv().ExportToShareableHandle(&out, m->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    *fd = out;
    return 0;
}

// Maps a peer's allocation (descriptor received from it) into this process for the current device.
extern "C" int b200_vmm_import(int fd, int64_t bytes, void** mapping, void** ptr) {
    B200_REQUIRE(drv().ok && fd >= 0 && bytes > 0 && mapping && ptr, "vmm_import: bad arguments");
    int dev;
    CUdevice cudev;
    if (int rc = current_device(&dev, &cudev)) return rc;
    Mapping* m = new (std::nothrow) Mapping{};
    B200_REQUIRE(m != nullptr, "vmm_import: out of host memory");
    m->bytes = (size_t)bytes;
    B200_CU_OK(drv().ImportFromShareableHandle(&m->handle, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    if (int rc = map_handle(m, dev)) return rc;
    *mapping = m;
    *ptr = (void*)m->va;
    return 0;
}

extern "C" int b200_vmm_release(void* mapping) {
    Mapping* m = (Mapping*)mapping;
    if (!m) return 0;
    if (m->mapped) drv().Unmap(m->va, m->bytes);
    if (m->va) drv().AddressFree(m->va, m->bytes);
    drv().Release(m->handle);
    delete m;
    return 0;
}

// ---------------------------------------------------------------- NVLS multicast object
// One rank creates the object for `ranks` devices and exports it (b200_vmm_export); the others import.
extern "C" int b200_mc_create(int64_t bytes, int ranks, void** mapping) {
    B200_REQUIRE(drv().ok && bytes > 0 && ranks > 1 && mapping, "mc_create: bad arguments");
    Mapping* m = new (std::nothrow) Mapping{};
    B200_REQUIRE(m != nullptr, "mc_create: out of host memory");
    m->bytes = (size_t)bytes;
    CUmulticastObjectProp mp{};
    mp.numDevices = (unsigned)ranks;
    mp.size = m->bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUresult r = drv().MulticastCreate(&m->handle, &mp);
    if (r != CUDA_SUCCESS) {
        delete m;
        b200_set_error("mc_create: cuMulticastCreate(%lld bytes, %d devices) -> %s", (long long)bytes, ranks, cu_err(r));
        return 1000 + (int)r;
    }
    *mapping = m;
    return 0;
}

extern "C" int b200_mc_import(int fd, int64_t bytes, void** mapping) {
    B200_REQUIRE(drv().ok && fd >= 0 && bytes > 0 && mapping, "mc_import: bad arguments");
    Mapping* m = new (std::nothrow) Mapping{};
    B200_REQUIRE(m != nullptr, "mc_import: out of host memory");
    m->bytes = (size_t)bytes;
    B200_CU_OK(drv().ImportFromShareableHandle(&m->handle, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    *mapping = m;
    return 0;
}

// Every rank adds its device; ALL ranks must have done so before any rank binds memory.
extern "C" int b200_mc_add_device(void* mc) {
    Mapping* m = (Mapping*)mc;
    B200_REQUIRE(m != nullptr, "mc_add_device: null argument");
    int dev;
    CUdevice cudev;
    if (int rc = current_device(&dev, &cudev)) return rc;
    B200_CU_OK(drv().MulticastAddDevice(m->handle, cudev));
    return 0;
}

// Binds the local allocation at offset 0 of the object and maps the object: a store to *mc_ptr + x
// lands at offset x of every rank's allocation, multimem.ld_reduce there returns the sum over ranks.
extern "C" int b200_mc_bind_map(void* mc, void* local_mapping, void** mc_ptr) {
    Mapping* m = (Mapping*)mc;
    Mapping* loc = (Mapping*)local_mapping;
    B200_REQUIRE(m && loc && mc_ptr && loc->bytes == m->bytes, "mc_bind_map: the allocation and the object must have the same size");
    int dev;
    CUdevice cudev;
    if (int rc = current_device(&dev, &cudev)) return rc;
    B200_CU_OK(drv().MulticastBindMem(m->handle, 0, loc->handle, 0, m->bytes, 0));
    if (int rc = map_handle(m, dev)) return rc;
    *mc_ptr = (void*)m->va;
    return 0;
}

extern "C" int b200_close_fd(int fd) {
    if (fd >= 0) close(fd);
    return 0;
}
