// Symmetric (peer-mapped) memory for one NVSwitch domain.
//
// The reference moved every byte through host sockets
// (/root/reference/srcs/go/rchannel/connection/connection.go:57-101) or handed
// it to NCCL (/root/reference/srcs/cpp/src/nccl/gpu_collective.cpp:59-128).
// Here a rank's communication buffers are CUDA VMM allocations exported as
// POSIX file descriptors; the fds travel over the control plane's unix
// sockets (SCM_RIGHTS), every peer maps them, and from then on kernels simply
// load/store remote HBM through NVLink. A multicast object over the same
// memory gives the NVLS (in-switch reduce / multicast store) address.
//
// The CUDA driver API is reached through cudaGetDriverEntryPoint so the
// library has no link-time dependency on libcuda.so (it must dlopen fine on a
// CPU-only build box).
#include <cuda.h>
#include <cuda_runtime.h>
#include <unistd.h>

#include <cstdio>
#include <cstring>
#include <mutex>

#include "../kernels/common.cuh"

namespace {

struct Driver {
    bool ok = false;
    CUresult (*MemCreate)(CUmemGenericAllocationHandle *, size_t,
                          const CUmemAllocationProp *, unsigned long long);
    CUresult (*MemRelease)(CUmemGenericAllocationHandle);
    CUresult (*MemAddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr,
                                  unsigned long long);
    CUresult (*MemAddressFree)(CUdeviceptr, size_t);
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t,
                       CUmemGenericAllocationHandle, unsigned long long);
    CUresult (*MemUnmap)(CUdeviceptr, size_t);
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *,
                             size_t);
    CUresult (*MemExportToShareableHandle)(void *, CUmemGenericAllocationHandle,
                                           CUmemAllocationHandleType,
                                           unsigned long long);
    CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle *,
                                             void *, CUmemAllocationHandleType);
    CUresult (*MemGetAllocationGranularity)(size_t *, const CUmemAllocationProp *,
                                            CUmemAllocationGranularity_flags);
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle *,
                                const CUmulticastObjectProp *);
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t,
                                 CUmemGenericAllocationHandle, size_t, size_t,
                                 unsigned long long);
    CUresult (*MulticastGetGranularity)(size_t *, const CUmulticastObjectProp *,
                                        CUmulticastGranularity_flags);
    CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
    CUresult (*DeviceGetAttribute)(int *, CUdevice_attribute, CUdevice);
    CUresult (*GetErrorString)(CUresult, const char **);
};

Driver g_drv;
std::once_flag g_drv_once;

template <typename F> bool load_sym(const char *name, F *fn)
{
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess || !p)
        return false;
    *fn = reinterpret_cast<F>(p);
    return true;
}

const Driver *driver()
{
    std::call_once(g_drv_once, [] {
        Driver &d = g_drv;
        bool ok   = true;
        ok &= load_sym("cuMemCreate", &d.MemCreate);
        ok &= load_sym("cuMemRelease", &d.MemRelease);
        ok &= load_sym("cuMemAddressReserve", &d.MemAddressReserve);
        ok &= load_sym("cuMemAddressFree", &d.MemAddressFree);
        ok &= load_sym("cuMemMap", &d.MemMap);
        ok &= load_sym("cuMemUnmap", &d.MemUnmap);
        ok &= load_sym("cuMemSetAccess", &d.MemSetAccess);
        ok &= load_sym("cuMemExportToShareableHandle", &d.MemExportToShareableHandle);
        ok &= load_sym("cuMemImportFromShareableHandle", &d.MemImportFromShareableHandle);
        ok &= load_sym("cuMemGetAllocationGranularity", &d.MemGetAllocationGranularity);
        ok &= load_sym("cuDeviceGetAttribute", &d.DeviceGetAttribute);
        ok &= load_sym("cuGetErrorString", &d.GetErrorString);
        // multicast is optional (capability probed per device)
        load_sym("cuMulticastCreate", &d.MulticastCreate);
        load_sym("cuMulticastAddDevice", &d.MulticastAddDevice);
        load_sym("cuMulticastBindMem", &d.MulticastBindMem);
        load_sym("cuMulticastGetGranularity", &d.MulticastGetGranularity);
        load_sym("cuMulticastUnbind", &d.MulticastUnbind);
        d.ok = ok;
    });
    return g_drv.ok ? &g_drv : nullptr;
}

#define KFB_DRV_OK(expr)                                                       \
    do {                                                                       \
        CUresult _r = (expr);                                                  \
        if (_r != CUDA_SUCCESS) {                                              \
            const char *_m = "?";                                              \
            if (g_drv.GetErrorString) g_drv.GetErrorString(_r, &_m);           \
            kfb::set_last_error("%s:%d %s -> %d (%s)", __FILE__, __LINE__,     \
                                #expr, (int)_r, _m);                           \
            return KFB_ERR_DRIVER;                                             \
        }                                                                      \
    } while (0)

CUmemAllocationProp make_prop(int device)
{
    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.type                 = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type        = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id          = device;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return prop;
}

int map_handle(const Driver *d, int device, CUmemGenericAllocationHandle h,
               size_t bytes, size_t align, CUdeviceptr *va_out)
{
    CUdeviceptr va = 0;
    KFB_DRV_OK(d->MemAddressReserve(&va, bytes, align, 0, 0));
    CUresult r = d->MemMap(va, bytes, 0, h, 0);
    if (r != CUDA_SUCCESS) {
        d->MemAddressFree(va, bytes);
        KFB_DRV_OK(r);
    }
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id   = device;
    acc.flags         = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r                 = d->MemSetAccess(va, bytes, &acc, 1);
    if (r != CUDA_SUCCESS) {
        d->MemUnmap(va, bytes);
        d->MemAddressFree(va, bytes);
        KFB_DRV_OK(r);
    }
    *va_out = va;
    return KFB_OK;
}

}  // namespace

struct kfb_symm_block {
    CUmemGenericAllocationHandle handle;
    CUdeviceptr va;
    size_t bytes;
    int device;
    bool imported;
};

struct kfb_mc_object {
    CUmemGenericAllocationHandle handle;
    CUdeviceptr va;
    size_t bytes;
    int device;
    int np;
    bool mapped;
};

extern "C" size_t kfb_symm_granularity(int device, int multicast)
{
    const Driver *d = driver();
    if (!d) return 0;
    size_t g = 0;
    if (multicast) {
        if (!d->MulticastGetGranularity) return 0;
        CUmulticastObjectProp mp;
        memset(&mp, 0, sizeof(mp));
        mp.numDevices  = 2;
        mp.size        = 2u << 20;
        mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        if (d->MulticastGetGranularity(&g, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS)
            return 0;
        return g;
    }
    CUmemAllocationProp prop = make_prop(device);
    if (d->MemGetAllocationGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS)
        return 0;
    return g;
}

extern "C" int kfb_multicast_supported(int device)
{
    const Driver *d = driver();
    if (!d || !d->MulticastCreate) return 0;
    int v = 0;
    if (d->DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device) != CUDA_SUCCESS)
        return 0;
    return v;
}

extern "C" int kfb_symm_alloc(int device, size_t bytes, kfb_symm_block **out)
{
    const Driver *d = driver();
    if (!d) {
        kfb::set_last_error("CUDA driver VMM entry points unavailable");
        return KFB_ERR_DRIVER;
    }
    KFB_CUDA_OK(cudaSetDevice(device));
    KFB_CUDA_OK(cudaFree(0));  // make sure the primary context exists
    CUmemAllocationProp prop = make_prop(device);
    size_t gran = 0;
    KFB_DRV_OK(d->MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    // multicast binding wants its own (larger) granularity; take the max
    size_t mg = kfb_multicast_supported(device) ? kfb_symm_granularity(device, 1) : 0;
    if (mg > gran) gran = mg;
    const size_t sz = (bytes + gran - 1) / gran * gran;
    kfb_symm_block *b = new kfb_symm_block();
    b->device   = device;
    b->bytes    = sz;
    b->imported = false;
    CUresult r  = d->MemCreate(&b->handle, sz, &prop, 0);
    if (r != CUDA_SUCCESS) {
        delete b;
        KFB_DRV_OK(r);
    }
    int rc = map_handle(d, device, b->handle, sz, gran, &b->va);
    if (rc) {
        d->MemRelease(b->handle);
        delete b;
        return rc;
    }
    KFB_CUDA_OK(cudaMemset(reinterpret_cast<void *>(b->va), 0, sz));
    *out = b;
    return KFB_OK;
}

extern "C" int kfb_symm_export_fd(kfb_symm_block *b, int *fd_out)
{
    const Driver *d = driver();
    int fd          = -1;
    KFB_DRV_OK(d->MemExportToShareableHandle(&fd, b->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
    *fd_out = fd;
    return KFB_OK;
}

extern "C" int kfb_symm_import_fd(int device, int fd, size_t bytes,
                                  kfb_symm_block **out)
{
    const Driver *d = driver();
    if (!d) {
        kfb::set_last_error("CUDA driver VMM entry points unavailable");
        return KFB_ERR_DRIVER;
    }
    KFB_CUDA_OK(cudaSetDevice(device));
    kfb_symm_block *b = new kfb_symm_block();
    b->device   = device;
    b->bytes    = bytes;
    b->imported = true;
    CUresult r  = d->MemImportFromShareableHandle(&b->handle, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (r != CUDA_SUCCESS) {
        delete b;
        KFB_DRV_OK(r);
    }
    CUmemAllocationProp prop = make_prop(device);
    size_t gran = 0;
    d->MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
    int rc = map_handle(d, device, b->handle, bytes, gran, &b->va);
    if (rc) {
        d->MemRelease(b->handle);
        delete b;
        return rc;
    }
    *out = b;
    return KFB_OK;
}

extern "C" void *kfb_symm_ptr(kfb_symm_block *b)
{
    return reinterpret_cast<void *>(b->va);
}
extern "C" size_t kfb_symm_size(kfb_symm_block *b) { return b->bytes; }

extern "C" int kfb_symm_free(kfb_symm_block *b)
{
    const Driver *d = driver();
    if (!b) return KFB_OK;
    if (d) {
        d->MemUnmap(b->va, b->bytes);
        d->MemAddressFree(b->va, b->bytes);
        d->MemRelease(b->handle);
    }
    delete b;
    return KFB_OK;
}

// ---- legacy CUDA IPC fallback ---------------------------------------------------
extern "C" int kfb_ipc_alloc(int device, size_t bytes, void **ptr_out,
                             unsigned char handle_out[64])
{
    KFB_CUDA_OK(cudaSetDevice(device));
    void *p = nullptr;
    KFB_CUDA_OK(cudaMalloc(&p, bytes));
    KFB_CUDA_OK(cudaMemset(p, 0, bytes));
    cudaIpcMemHandle_t h;
    KFB_CUDA_OK(cudaIpcGetMemHandle(&h, p));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle_out, &h, 64);
    *ptr_out = p;
    return KFB_OK;
}

extern "C" int kfb_ipc_open(int device, const unsigned char handle[64],
                            void **ptr_out)
{
    KFB_CUDA_OK(cudaSetDevice(device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    KFB_CUDA_OK(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
    return KFB_OK;
}

extern "C" int kfb_ipc_close(void *ptr)
{
    KFB_CUDA_OK(cudaIpcCloseMemHandle(ptr));
    return KFB_OK;
}

extern "C" int kfb_ipc_free(void *ptr)
{
    KFB_CUDA_OK(cudaFree(ptr));
    return KFB_OK;
}

// ---- multicast (NVLS) ---------------------------------------------------------------
static CUmulticastObjectProp mc_prop(int np, size_t bytes)
{
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices  = (unsigned)np;
    mp.size        = bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return mp;
}

extern "C" int kfb_mc_create(int device, int np, size_t bytes,
                             kfb_mc_object **out, int *fd_out)
{
    const Driver *d = driver();
    if (!d || !d->MulticastCreate || !kfb_multicast_supported(device)) {
        kfb::set_last_error("multicast not supported on device %d", device);
        return KFB_ERR_UNSUPP;
    }
    KFB_CUDA_OK(cudaSetDevice(device));
    CUmulticastObjectProp mp = mc_prop(np, bytes);
    kfb_mc_object *mc = new kfb_mc_object();
    mc->device = device;
    mc->np     = np;
    mc->bytes  = bytes;
    mc->mapped = false;
    CUresult r = d->MulticastCreate(&mc->handle, &mp);
    if (r != CUDA_SUCCESS) {
        delete mc;
        KFB_DRV_OK(r);
    }
    int fd = -1;
    r = d->MemExportToShareableHandle(&fd, mc->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        d->MemRelease(mc->handle);
        delete mc;
        KFB_DRV_OK(r);
    }
    *fd_out = fd;
    *out    = mc;
    return KFB_OK;
}

extern "C" int kfb_mc_import(int device, int np, int fd, size_t bytes,
                             kfb_mc_object **out)
{
    const Driver *d = driver();
    if (!d || !d->MulticastCreate) {
        kfb::set_last_error("multicast not supported");
        return KFB_ERR_UNSUPP;
    }
    KFB_CUDA_OK(cudaSetDevice(device));
    kfb_mc_object *mc = new kfb_mc_object();
    mc->device = device;
    mc->np     = np;
    mc->bytes  = bytes;
    mc->mapped = false;
    CUresult r = d->MemImportFromShareableHandle(&mc->handle, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (r != CUDA_SUCCESS) {
        delete mc;
        KFB_DRV_OK(r);
    }
    *out = mc;
    return KFB_OK;
}

extern "C" int kfb_mc_add_device(kfb_mc_object *mc)
{
    const Driver *d = driver();
    KFB_DRV_OK(d->MulticastAddDevice(mc->handle, mc->device));
    return KFB_OK;
}

// Call after EVERY rank has added its device (barrier in between).
extern "C" int kfb_mc_bind(kfb_mc_object *mc, kfb_symm_block *mem,
                           void **mc_ptr_out)
{
    const Driver *d = driver();
    if (mem->bytes < mc->bytes) {
        kfb::set_last_error("multicast bind: block smaller than the object");
        return KFB_ERR_ARG;
    }
    KFB_DRV_OK(d->MulticastBindMem(mc->handle, 0, mem->handle, 0, mc->bytes, 0));
    size_t gran = kfb_symm_granularity(mc->device, 1);
    int rc = map_handle(d, mc->device, mc->handle, mc->bytes, gran, &mc->va);
    if (rc) return rc;
    mc->mapped  = true;
    *mc_ptr_out = reinterpret_cast<void *>(mc->va);
    return KFB_OK;
}

extern "C" int kfb_mc_free(kfb_mc_object *mc)
{
    const Driver *d = driver();
    if (!mc) return KFB_OK;
    if (d) {
        if (mc->mapped) {
            d->MemUnmap(mc->va, mc->bytes);
            d->MemAddressFree(mc->va, mc->bytes);
            // give this device's physical pages back to the switch group before the handle goes
            if (d->MulticastUnbind) d->MulticastUnbind(mc->handle, mc->device, 0, mc->bytes);
        }
        d->MemRelease(mc->handle);
    }
    delete mc;
    return KFB_OK;
}

extern "C" int kfb_close_fd(int fd) { return close(fd); }
