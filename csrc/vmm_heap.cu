// Symmetric heap on the CUDA virtual-memory-management API, with an NVLS multicast mapping.
//
// The default heap (api.cu: cudaMalloc + cudaIpc handles) cannot be bound to a multicast object -- the switch can only
// reduce / broadcast over memory created with cuMemCreate.  This file is that second heap: every rank creates a
// physical allocation, exports it as a POSIX file descriptor (the Python side passes the descriptors between the
// towers over a UNIX socket, parallel/p2p.py), maps every peer's allocation, and binds its own allocation into ONE
// multicast object whose mapping is the address multimem.ld_reduce / multimem.st use (csrc/p2p_allreduce.cu, modes
// 3 / 4).  It replaces the allocation the reference gets implicitly from TensorFlow's NCCL all-reduce
// (Test.ipynb:200, model.py:462-467) and removes the dependency on torch's private symmetric-memory module.
//
// The library links cudart only, so every driver entry point is fetched with cudaGetDriverEntryPoint.
#include <cuda.h>
#include <cuda_runtime.h>
#include <unistd.h>

#include <cstdint>
#include <cstdio>
#include <cstring>

#include "tdl_common.cuh"

namespace {

struct Drv {
  decltype(&cuMemCreate) memCreate = nullptr;
  decltype(&cuMemRelease) memRelease = nullptr;
  decltype(&cuMemExportToShareableHandle) memExport = nullptr;
  decltype(&cuMemImportFromShareableHandle) memImport = nullptr;
  decltype(&cuMemAddressReserve) addrReserve = nullptr;
  decltype(&cuMemAddressFree) addrFree = nullptr;
  decltype(&cuMemMap) memMap = nullptr;
  decltype(&cuMemUnmap) memUnmap = nullptr;
  decltype(&cuMemSetAccess) memSetAccess = nullptr;
  decltype(&cuMemGetAllocationGranularity) memGran = nullptr;
  decltype(&cuMulticastCreate) mcCreate = nullptr;
  decltype(&cuMulticastAddDevice) mcAddDevice = nullptr;
  decltype(&cuMulticastBindMem) mcBindMem = nullptr;
  decltype(&cuMulticastGetGranularity) mcGran = nullptr;
  decltype(&cuDeviceGetAttribute) devAttr = nullptr;
  bool vmm = false, mc = false, tried = false;
};

template <class F>
bool load(const char* name, F* out) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
      p == nullptr)
    return false;
  *out = reinterpret_cast<F>(p);
  return true;
}

Drv& drv() {
  static Drv d;
  if (!d.tried) {
    d.tried = true;
    cudaFree(nullptr);  // the driver calls below need this thread's primary context
    d.vmm = load("cuMemCreate", &d.memCreate) && load("cuMemRelease", &d.memRelease) &&
            load("cuMemExportToShareableHandle", &d.memExport) &&
            load("cuMemImportFromShareableHandle", &d.memImport) && load("cuMemAddressReserve", &d.addrReserve) &&
            load("cuMemAddressFree", &d.addrFree) && load("cuMemMap", &d.memMap) && load("cuMemUnmap", &d.memUnmap) &&
            load("cuMemSetAccess", &d.memSetAccess) && load("cuMemGetAllocationGranularity", &d.memGran) &&
            load("cuDeviceGetAttribute", &d.devAttr);
    d.mc = d.vmm && load("cuMulticastCreate", &d.mcCreate) && load("cuMulticastAddDevice", &d.mcAddDevice) &&
           load("cuMulticastBindMem", &d.mcBindMem) && load("cuMulticastGetGranularity", &d.mcGran);
  }
  return d;
}

int drv_fail(const char* what, long long code) {
  snprintf(tdl::err_buf(), 512, "%s failed (CUresult %lld)", what, code);
  return -5;
}

#define TDL_DRV_OK(expr)                                      \
  do {                                                        \
    CUresult _r = (expr);                                     \
    if (_r != CUDA_SUCCESS) return drv_fail(#expr, (int)_r);  \
  } while (0)

struct VmmAlloc {
  CUmemGenericAllocationHandle handle = 0;
  CUdeviceptr va = 0;
  size_t size = 0;
  int fd = -1;
};

struct VmmMulticast {
  CUmemGenericAllocationHandle handle = 0;
  CUdeviceptr va = 0;
  size_t size = 0;
  int fd = -1;
};

int current_device(int* dev) {
  if (cudaGetDevice(dev) != cudaSuccess) return tdl::fail(-5, "%s%s", "cudaGetDevice failed", "");
  return 0;
}

CUmemAllocationProp alloc_prop(int dev) {
  CUmemAllocationProp p;
  memset(&p, 0, sizeof(p));
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = dev;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

int map_rw(Drv& d, CUmemGenericAllocationHandle h, size_t size, size_t align, int dev, CUdeviceptr* va) {
  TDL_DRV_OK(d.addrReserve(va, size, align, 0, 0));
  CUresult r = d.memMap(*va, size, 0, h, 0);
  if (r != CUDA_SUCCESS) {
    d.addrFree(*va, size);
    return drv_fail("cuMemMap", (int)r);
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d.memSetAccess(*va, size, &acc, 1);
  if (r != CUDA_SUCCESS) {
    d.memUnmap(*va, size);
    d.addrFree(*va, size);
    return drv_fail("cuMemSetAccess", (int)r);
  }
  return 0;
}

}  // namespace

// What this device / driver offers: vmm_ok = cuMem* with POSIX-fd handles, mc_ok = NVLS multicast objects;
// *granularity = the size every allocation of `world` towers must be a multiple of (the larger of the allocation and
// the recommended multicast granularity).
TDL_API int tdl_vmm_query(int world, int* vmm_ok, int* mc_ok, int64_t* granularity) {
  Drv& d = drv();
  int dev = 0;
  if (int rc = current_device(&dev)) return rc;
  *vmm_ok = 0;
  *mc_ok = 0;
  *granularity = 2 << 20;
  if (!d.vmm) return 0;
  int v = 0;
  TDL_DRV_OK(d.devAttr(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev));
  *vmm_ok = v;
  if (!v) return 0;
  CUmemAllocationProp p = alloc_prop(dev);
  size_t g = 0;
  TDL_DRV_OK(d.memGran(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  *granularity = (int64_t)g;
  if (d.mc && world > 1) {
    int m = 0;
    TDL_DRV_OK(d.devAttr(&m, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev));
    if (m) {
      CUmulticastObjectProp mp;
      memset(&mp, 0, sizeof(mp));
      mp.numDevices = (unsigned)world;
      mp.size = g;
      mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
      size_t mg = 0;
      if (d.mcGran(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > 0) {
        *mc_ok = 1;
        if ((int64_t)mg > *granularity) *granularity = (int64_t)mg;
      }
    }
  }
  return 0;
}

// Physical allocation of `padded` bytes (a multiple of tdl_vmm_query's granularity) on the current device, mapped
// read/write; *fd_out is its shareable descriptor (owned by the allocation, closed by tdl_vmm_free).
TDL_API int tdl_vmm_alloc(int64_t padded, void** alloc_out, void** ptr_out, int* fd_out) {
  Drv& d = drv();
  TDL_REQUIRE(d.vmm, "CUDA virtual memory management entry points unavailable");
  TDL_REQUIRE(padded > 0, "size must be positive");
  int dev = 0;
  if (int rc = current_device(&dev)) return rc;
  CUmemAllocationProp p = alloc_prop(dev);
  VmmAlloc* a = new VmmAlloc();
  a->size = (size_t)padded;
  CUresult r = d.memCreate(&a->handle, a->size, &p, 0);
  if (r != CUDA_SUCCESS) {
    delete a;
    return drv_fail("cuMemCreate", (int)r);
  }
  if (int rc = map_rw(d, a->handle, a->size, 0, dev, &a->va)) {
    d.memRelease(a->handle);
    delete a;
    return rc;
  }
  r = d.memExport(&a->fd, a->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    d.memUnmap(a->va, a->size);
    d.addrFree(a->va, a->size);
    d.memRelease(a->handle);
    delete a;
    return drv_fail("cuMemExportToShareableHandle", (int)r);
  }
  *alloc_out = a;
  *ptr_out = reinterpret_cast<void*>(a->va);
  *fd_out = a->fd;
  return 0;
}

// Map a peer tower's allocation (its descriptor received over a UNIX socket) read/write for the current device.
TDL_API int tdl_vmm_import(int fd, int64_t padded, void** alloc_out, void** ptr_out) {
  Drv& d = drv();
  TDL_REQUIRE(d.vmm, "CUDA virtual memory management entry points unavailable");
  int dev = 0;
  if (int rc = current_device(&dev)) return rc;
  VmmAlloc* a = new VmmAlloc();
  a->size = (size_t)padded;
  a->fd = fd;  // ours now
  CUresult r = d.memImport(&a->handle, reinterpret_cast<void*>((uintptr_t)fd), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) {
    delete a;
    return drv_fail("cuMemImportFromShareableHandle", (int)r);
  }
  if (int rc = map_rw(d, a->handle, a->size, 0, dev, &a->va)) {
    d.memRelease(a->handle);
    delete a;
    return rc;
  }
  *alloc_out = a;
  *ptr_out = reinterpret_cast<void*>(a->va);
  return 0;
}

TDL_API int tdl_vmm_free(void* alloc) {
  if (!alloc) return 0;
  Drv& d = drv();
  VmmAlloc* a = static_cast<VmmAlloc*>(alloc);
  if (a->va) {
    d.memUnmap(a->va, a->size);
    d.addrFree(a->va, a->size);
  }
  if (a->handle) d.memRelease(a->handle);
  if (a->fd >= 0) close(a->fd);
  delete a;
  return 0;
}

// Tower 0: the multicast object for `world` devices and `padded` bytes; *fd_out goes to every other tower.
TDL_API int tdl_vmm_mc_create(int world, int64_t padded, void** mc_out, int* fd_out) {
  Drv& d = drv();
  TDL_REQUIRE(d.mc, "multicast entry points unavailable");
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)world;
  mp.size = (size_t)padded;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  VmmMulticast* m = new VmmMulticast();
  m->size = (size_t)padded;
  CUresult r = d.mcCreate(&m->handle, &mp);
  if (r != CUDA_SUCCESS) {
    delete m;
    return drv_fail("cuMulticastCreate", (int)r);
  }
  r = d.memExport(&m->fd, m->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) {
    d.memRelease(m->handle);
    delete m;
    return drv_fail("cuMemExportToShareableHandle(multicast)", (int)r);
  }
  *mc_out = m;
  *fd_out = m->fd;
  return 0;
}

TDL_API int tdl_vmm_mc_import(int fd, int64_t padded, void** mc_out) {
  Drv& d = drv();
  TDL_REQUIRE(d.mc, "multicast entry points unavailable");
  VmmMulticast* m = new VmmMulticast();
  m->size = (size_t)padded;
  m->fd = fd;
  CUresult r = d.memImport(&m->handle, reinterpret_cast<void*>((uintptr_t)fd), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) {
    delete m;
    return drv_fail("cuMemImportFromShareableHandle(multicast)", (int)r);
  }
  *mc_out = m;
  return 0;
}

// Every tower, once it holds the object: register the current device.  ALL towers must have done this before any of
// them binds memory (barrier in between).
TDL_API int tdl_vmm_mc_add_device(void* mc) {
  Drv& d = drv();
  VmmMulticast* m = static_cast<VmmMulticast*>(mc);
  int dev = 0;
  if (int rc = current_device(&dev)) return rc;
  TDL_DRV_OK(d.mcAddDevice(m->handle, dev));
  return 0;
}

// Bind this tower's allocation at offset 0 of the multicast object and map the object: *mc_ptr_out is the address
// multimem instructions take (a store reaches every tower's allocation, a ld_reduce sums over them).
TDL_API int tdl_vmm_mc_bind_map(void* mc, void* alloc, void** mc_ptr_out) {
  Drv& d = drv();
  VmmMulticast* m = static_cast<VmmMulticast*>(mc);
  VmmAlloc* a = static_cast<VmmAlloc*>(alloc);
  TDL_REQUIRE(m->size == a->size, "allocation and multicast object differ in size");
  int dev = 0;
  if (int rc = current_device(&dev)) return rc;
  TDL_DRV_OK(d.mcBindMem(m->handle, 0, a->handle, 0, a->size, 0));
  if (int rc = map_rw(d, m->handle, m->size, 0, dev, &m->va)) return rc;
  *mc_ptr_out = reinterpret_cast<void*>(m->va);
  return 0;
}

TDL_API int tdl_vmm_mc_free(void* mc) {
  if (!mc) return 0;
  Drv& d = drv();
  VmmMulticast* m = static_cast<VmmMulticast*>(mc);
  if (m->va) {
    d.memUnmap(m->va, m->size);
    d.addrFree(m->va, m->size);
  }
  if (m->handle) d.memRelease(m->handle);
  if (m->fd >= 0) close(m->fd);
  delete m;
  return 0;
}
