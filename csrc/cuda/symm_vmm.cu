// Own symmetric-memory allocator on the CUDA virtual-memory-management API (SURVEY 5.9 / 7.1 "symm_mem"): physical
// allocations that are exported as POSIX file descriptors, imported and mapped by every peer process, and - where the
// fabric supports it - bound to an NVSwitch MULTICAST object whose mapping lets one `multimem.ld_reduce` / `multimem.st`
// address all replicas.  This replaces the two borrowed pieces of round 1: `cudaIpc*` handles for the P2P arenas and
// `torch.distributed._symmetric_memory` for the NVLS mapping.  The reference's counterpart is the MPI library underneath
// `comm.Bcast` / `comm.Allreduce` (Module_3/TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:74-98): here the "transport" is a
// set of mappings, made once, that kernels then load from and store to.
//
// Host code only (driver entry points are fetched through the runtime, so the library still links without -lcuda).  The file
// descriptors travel between the ranks of a node over AF_UNIX sockets (SCM_RIGHTS) - that part is Python
// (`crossscale/parallel/vmm.py`), as is the ordering of the collective steps (create -> add device -> bind -> map).
#include "common/cs_common.cuh"

#include <cuda.h>
#include <string.h>
#include <unistd.h>

namespace cs {

#define CS_DRV_FN(name) \
    static decltype(&name) p_##name() { \
        static decltype(&name) fn = nullptr; \
        if (!fn) { \
            void* p = nullptr; cudaDriverEntryPointQueryResult q; \
            if (cudaGetDriverEntryPoint(#name, &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) \
                fn = reinterpret_cast<decltype(&name)>(p); \
        } \
        return fn; \
    }
CS_DRV_FN(cuMemCreate)
CS_DRV_FN(cuMemRelease)
CS_DRV_FN(cuMemExportToShareableHandle)
CS_DRV_FN(cuMemImportFromShareableHandle)
CS_DRV_FN(cuMemAddressReserve)
CS_DRV_FN(cuMemAddressFree)
CS_DRV_FN(cuMemMap)
CS_DRV_FN(cuMemUnmap)
CS_DRV_FN(cuMemSetAccess)
CS_DRV_FN(cuMemGetAllocationGranularity)
CS_DRV_FN(cuMulticastCreate)
CS_DRV_FN(cuMulticastAddDevice)
CS_DRV_FN(cuMulticastBindMem)
CS_DRV_FN(cuMulticastGetGranularity)
CS_DRV_FN(cuDeviceGetAttribute)
#undef CS_DRV_FN

#define CS_DRV_OK(call, what) do { CUresult r_ = (call); CS_REQUIRE(r_ == CUDA_SUCCESS, "vmm: %s failed (CUresult %d)", what, (int)r_); } while (0)

static bool have_vmm() {
    return p_cuMemCreate() && p_cuMemRelease() && p_cuMemExportToShareableHandle() && p_cuMemImportFromShareableHandle() &&
           p_cuMemAddressReserve() && p_cuMemAddressFree() && p_cuMemMap() && p_cuMemUnmap() && p_cuMemSetAccess() &&
           p_cuMemGetAllocationGranularity() && p_cuDeviceGetAttribute();
}
static bool have_mc() { return p_cuMulticastCreate() && p_cuMulticastAddDevice() && p_cuMulticastBindMem() && p_cuMulticastGetGranularity(); }

static CUmemAllocationProp device_prop(int dev) {
    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = dev;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return prop;
}

// reserve a virtual range, map `handle` into it and give the current device read/write access
static int map_handle(CUmemGenericAllocationHandle handle, size_t bytes, size_t align, int dev, void** ptr) {
    CUdeviceptr va = 0;
    CS_DRV_OK(p_cuMemAddressReserve()(&va, bytes, align, 0, 0), "cuMemAddressReserve");
    CUresult r = p_cuMemMap()(va, bytes, 0, handle, 0);
    if (r != CUDA_SUCCESS) { p_cuMemAddressFree()(va, bytes); CS_REQUIRE(false, "vmm: cuMemMap failed (CUresult %d)", (int)r); }
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE; acc.location.id = dev; acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = p_cuMemSetAccess()(va, bytes, &acc, 1);
    if (r != CUDA_SUCCESS) { p_cuMemUnmap()(va, bytes); p_cuMemAddressFree()(va, bytes); CS_REQUIRE(false, "vmm: cuMemSetAccess failed (CUresult %d)", (int)r); }
    *ptr = (void*)va;
    return CS_OK;
}

}  // namespace cs

using namespace cs;

extern "C" {

/* bit 0: VMM allocations with POSIX-fd export on the current device; bit 1: multicast objects (NVLS) */
int cs_vmm_support(void) {
    if (!have_vmm()) return 0;
    int dev = 0, v = 0, out = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
    if (p_cuDeviceGetAttribute()(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev) == CUDA_SUCCESS && v) out |= 1;
    v = 0;
    if ((out & 1) && have_mc() && p_cuDeviceGetAttribute()(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) == CUDA_SUCCESS && v) out |= 2;
    return out;
}

/* Physical allocation of >= `bytes` on the current device (rounded up to the allocation - and, if `for_multicast`, the
 * multicast - granularity), mapped read/write for this device and zero-filled.  *fd is the exportable descriptor (the
 * caller closes it after it has been sent to the peers). */
int cs_vmm_alloc(size_t bytes, int for_multicast, int world, void** ptr, size_t* alloc_bytes, unsigned long long* handle, int* fd) {
    CS_REQUIRE(have_vmm(), "vmm: the driver offers no virtual-memory-management entry points");
    CS_REQUIRE(bytes > 0 && ptr && alloc_bytes && handle && fd, "vmm_alloc: bad arguments");
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    CS_CUDA_OK(cudaFree(0));                                 // make sure the primary context exists
    CUmemAllocationProp prop = device_prop(dev);
    size_t gran = 0;
    CS_DRV_OK(p_cuMemGetAllocationGranularity()(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "cuMemGetAllocationGranularity");
    if (for_multicast) {
        CS_REQUIRE(have_mc(), "vmm: no multicast entry points");
        CUmulticastObjectProp mp;
        memset(&mp, 0, sizeof(mp));
        mp.numDevices = (unsigned)world; mp.size = bytes; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        size_t mg = 0;
        CS_DRV_OK(p_cuMulticastGetGranularity()(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED), "cuMulticastGetGranularity");
        if (mg > gran) gran = mg;
    }
    const size_t size = (bytes + gran - 1) / gran * gran;
    CUmemGenericAllocationHandle h = 0;
    CS_DRV_OK(p_cuMemCreate()(&h, size, &prop, 0), "cuMemCreate");
    int rc = map_handle(h, size, gran, dev, ptr);
    if (rc) { p_cuMemRelease()(h); return rc; }
    CS_CUDA_OK(cudaMemset(*ptr, 0, size));
    CS_CUDA_OK(cudaDeviceSynchronize());
    int f = -1;
    CUresult r = p_cuMemExportToShareableHandle()(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    CS_REQUIRE(r == CUDA_SUCCESS, "vmm: cuMemExportToShareableHandle failed (CUresult %d)", (int)r);
    *alloc_bytes = size; *handle = (unsigned long long)h; *fd = f;
    return CS_OK;
}

/* Map a peer's allocation (its descriptor received over a socket) into this process, read/write for the current device. */
int cs_vmm_import(int fd, size_t alloc_bytes, void** ptr, unsigned long long* handle) {
    CS_REQUIRE(have_vmm() && fd >= 0 && ptr && handle, "vmm_import: bad arguments");
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    CUmemGenericAllocationHandle h = 0;
    CS_DRV_OK(p_cuMemImportFromShareableHandle()(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR), "cuMemImportFromShareableHandle");
    int rc = map_handle(h, alloc_bytes, 0, dev, ptr);
    if (rc) { p_cuMemRelease()(h); return rc; }
    *handle = (unsigned long long)h;
    return CS_OK;
}

/* unmap + release one mapping made by cs_vmm_alloc / cs_vmm_import / cs_vmm_mc_map */
int cs_vmm_free(void* ptr, size_t alloc_bytes, unsigned long long handle) {
    if (!have_vmm()) return CS_OK;
    if (ptr) { p_cuMemUnmap()((CUdeviceptr)ptr, alloc_bytes); p_cuMemAddressFree()((CUdeviceptr)ptr, alloc_bytes); }
    if (handle) p_cuMemRelease()((CUmemGenericAllocationHandle)handle);
    return CS_OK;
}

/* Multicast object for `world` devices and `alloc_bytes` (the size cs_vmm_alloc returned); created by ONE rank, which
 * sends *fd to the others (they call cs_vmm_mc_import). */
int cs_vmm_mc_create(int world, size_t alloc_bytes, unsigned long long* mc_handle, int* fd) {
    CS_REQUIRE(have_vmm() && have_mc(), "vmm: no multicast support");
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)world; mp.size = alloc_bytes; mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUmemGenericAllocationHandle h = 0;
    CS_DRV_OK(p_cuMulticastCreate()(&h, &mp), "cuMulticastCreate");
    int f = -1;
    CUresult r = p_cuMemExportToShareableHandle()(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) { p_cuMemRelease()(h); CS_REQUIRE(false, "vmm: export of the multicast object failed (CUresult %d)", (int)r); }
    *mc_handle = (unsigned long long)h; *fd = f;
    return CS_OK;
}

int cs_vmm_mc_import(int fd, unsigned long long* mc_handle) {
    CS_REQUIRE(have_vmm() && have_mc() && fd >= 0, "vmm_mc_import: bad arguments");
    CUmemGenericAllocationHandle h = 0;
    CS_DRV_OK(p_cuMemImportFromShareableHandle()(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR), "cuMemImportFromShareableHandle (multicast)");
    *mc_handle = (unsigned long long)h;
    return CS_OK;
}

/* step 1 on every rank (all ranks must have done it before anyone binds) */
int cs_vmm_mc_add_device(unsigned long long mc_handle) {
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    CS_DRV_OK(p_cuMulticastAddDevice()((CUmemGenericAllocationHandle)mc_handle, dev), "cuMulticastAddDevice");
    return CS_OK;
}

/* step 2 on every rank: bind this rank's physical allocation at offset 0 and map the multicast object; loads / stores through
 * *mc_ptr must be multimem instructions */
int cs_vmm_mc_bind_map(unsigned long long mc_handle, unsigned long long mem_handle, size_t alloc_bytes, void** mc_ptr) {
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    CS_DRV_OK(p_cuMulticastBindMem()((CUmemGenericAllocationHandle)mc_handle, 0, (CUmemGenericAllocationHandle)mem_handle, 0, alloc_bytes, 0), "cuMulticastBindMem");
    return map_handle((CUmemGenericAllocationHandle)mc_handle, alloc_bytes, 0, dev, mc_ptr);
}

int cs_vmm_close_fd(int fd) { return fd >= 0 ? close(fd) : 0; }

}  // extern "C"
