// NVSwitch multicast pages: see multicast.h.  Driver API (virtual memory management + multicast
// objects) reached through cudaGetDriverEntryPoint, so nothing links against libcuda.
#include "multicast.h"

#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>

namespace dmlp {
namespace gpu {

namespace {

using FnDeviceGet = CUresult (*)(CUdevice *, int);
using FnDeviceGetAttribute = CUresult (*)(int *, CUdevice_attribute, CUdevice);
using FnGetErrorName = CUresult (*)(CUresult, const char **);
using FnMcCreate = CUresult (*)(CUmemGenericAllocationHandle *, const CUmulticastObjectProp *);
using FnMcAddDevice = CUresult (*)(CUmemGenericAllocationHandle, CUdevice);
using FnMcBindMem = CUresult (*)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                                 unsigned long long);
using FnMcGranularity = CUresult (*)(size_t *, const CUmulticastObjectProp *, CUmulticastGranularity_flags);
using FnMemGranularity = CUresult (*)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags);
using FnMemCreate = CUresult (*)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long);
using FnMemRelease = CUresult (*)(CUmemGenericAllocationHandle);
using FnMemReserve = CUresult (*)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long);
using FnMemFree = CUresult (*)(CUdeviceptr, size_t);
using FnMemMap = CUresult (*)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
using FnMemUnmap = CUresult (*)(CUdeviceptr, size_t);
using FnMemSetAccess = CUresult (*)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t);

struct Api {
  FnDeviceGet device_get = nullptr;
  FnDeviceGetAttribute device_attr = nullptr;
  FnGetErrorName error_name = nullptr;
  FnMcCreate mc_create = nullptr;
  FnMcAddDevice mc_add = nullptr;
  FnMcBindMem mc_bind = nullptr;
  FnMcGranularity mc_gran = nullptr;
  FnMemGranularity mem_gran = nullptr;
  FnMemCreate mem_create = nullptr;
  FnMemRelease mem_release = nullptr;
  FnMemReserve mem_reserve = nullptr;
  FnMemFree mem_free = nullptr;
  FnMemMap mem_map = nullptr;
  FnMemUnmap mem_unmap = nullptr;
  FnMemSetAccess mem_access = nullptr;
  std::string missing;
};

template <class Fn>
void load(const char *name, Fn *out, std::string *missing) {
  void *p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess &&
      p != nullptr) {
    *out = reinterpret_cast<Fn>(p);
  } else {
    cudaGetLastError();
    if (missing->empty()) *missing = name;
  }
}

const Api &api() {
  static Api a = [] {
    Api x;
    load("cuDeviceGet", &x.device_get, &x.missing);
    load("cuDeviceGetAttribute", &x.device_attr, &x.missing);
    load("cuGetErrorName", &x.error_name, &x.missing);
    load("cuMulticastCreate", &x.mc_create, &x.missing);
    load("cuMulticastAddDevice", &x.mc_add, &x.missing);
    load("cuMulticastBindMem", &x.mc_bind, &x.missing);
    load("cuMulticastGetGranularity", &x.mc_gran, &x.missing);
    load("cuMemGetAllocationGranularity", &x.mem_gran, &x.missing);
    load("cuMemCreate", &x.mem_create, &x.missing);
    load("cuMemRelease", &x.mem_release, &x.missing);
    load("cuMemAddressReserve", &x.mem_reserve, &x.missing);
    load("cuMemAddressFree", &x.mem_free, &x.missing);
    load("cuMemMap", &x.mem_map, &x.missing);
    load("cuMemUnmap", &x.mem_unmap, &x.missing);
    load("cuMemSetAccess", &x.mem_access, &x.missing);
    return x;
  }();
  return a;
}

std::string describe(const Api &a, const char *what, CUresult rc) {
  const char *name = nullptr;
  if (a.error_name) a.error_name(rc, &name);
  char buf[160];
  std::snprintf(buf, sizeof buf, "%s: %s (%d)", what, name ? name : "error", static_cast<int>(rc));
  return buf;
}

size_t round_up(size_t v, size_t m) { return m ? (v + m - 1) / m * m : v; }

}  // namespace

void McPages::destroy() {
  const Api &a = api();
  if (!a.missing.empty()) return;
  if (mc != nullptr) {
    a.mem_unmap(reinterpret_cast<CUdeviceptr>(mc), bytes);
    a.mem_free(reinterpret_cast<CUdeviceptr>(mc), bytes);
    mc = nullptr;
  }
  for (int d = 0; d < kMcMaxDev; ++d) {
    if (local[d] != nullptr) {
      a.mem_unmap(reinterpret_cast<CUdeviceptr>(local[d]), bytes);
      a.mem_free(reinterpret_cast<CUdeviceptr>(local[d]), bytes);
      local[d] = nullptr;
    }
    if (mem_handle_[d] != 0) {
      a.mem_release(mem_handle_[d]);
      mem_handle_[d] = 0;
    }
  }
  if (mc_handle_ != 0) {
    a.mem_release(mc_handle_);
    mc_handle_ = 0;
  }
  ok = false;
}

McPages McPages::create(int ndev, size_t min_bytes) {
  McPages p;
  p.ndev = ndev;
  const Api &a = api();
  auto fail = [&](const std::string &why) {
    p.destroy();
    p.why = why;
    return p;
  };
  if (ndev < 2 || ndev > kMcMaxDev) return fail("needs 2 to 8 GPUs");
  if (!a.missing.empty()) return fail("driver entry point " + a.missing + " not available");
  CUdevice dev[kMcMaxDev];
  for (int d = 0; d < ndev; ++d) {
    if (cudaSetDevice(d) != cudaSuccess || cudaFree(nullptr) != cudaSuccess) {  // (primary context up)
      cudaGetLastError();
      return fail("cudaSetDevice failed");
    }
    CUresult rc = a.device_get(&dev[d], d);
    if (rc != CUDA_SUCCESS) return fail(describe(a, "cuDeviceGet", rc));
    int sup = 0;
    rc = a.device_attr(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev[d]);
    if (rc != CUDA_SUCCESS) return fail(describe(a, "cuDeviceGetAttribute", rc));
    if (!sup) return fail("GPU " + std::to_string(d) + " reports no multicast support");
  }
  CUmulticastObjectProp mprop;
  std::memset(&mprop, 0, sizeof mprop);
  mprop.numDevices = static_cast<unsigned>(ndev);
  mprop.size = min_bytes ? min_bytes : 1;
  mprop.handleTypes = 0;
  size_t gran = 0;
  CUresult rc = a.mc_gran(&gran, &mprop, CU_MULTICAST_GRANULARITY_RECOMMENDED);
  if (rc != CUDA_SUCCESS || gran == 0) return fail(describe(a, "cuMulticastGetGranularity", rc));
  CUmemAllocationProp aprop;
  std::memset(&aprop, 0, sizeof aprop);
  aprop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  aprop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  aprop.location.id = 0;
  size_t agran = 0;
  rc = a.mem_gran(&agran, &aprop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  if (rc != CUDA_SUCCESS || agran == 0) return fail(describe(a, "cuMemGetAllocationGranularity", rc));
  const size_t g = gran > agran ? gran : agran;
  p.bytes = round_up(round_up(mprop.size, gran), agran);
  mprop.size = p.bytes;

  CUmemGenericAllocationHandle mch = 0;
  rc = a.mc_create(&mch, &mprop);
  if (rc != CUDA_SUCCESS) return fail(describe(a, "cuMulticastCreate", rc));
  p.mc_handle_ = mch;
  for (int d = 0; d < ndev; ++d) {
    rc = a.mc_add(mch, dev[d]);
    if (rc != CUDA_SUCCESS) return fail(describe(a, "cuMulticastAddDevice", rc));
  }
  CUmemAccessDesc access[kMcMaxDev];
  for (int d = 0; d < ndev; ++d) {
    std::memset(&access[d], 0, sizeof access[d]);
    access[d].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    access[d].location.id = d;
    access[d].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  }
  for (int d = 0; d < ndev; ++d) {
    aprop.location.id = d;
    CUmemGenericAllocationHandle mh = 0;
    rc = a.mem_create(&mh, p.bytes, &aprop, 0);
    if (rc != CUDA_SUCCESS) return fail(describe(a, "cuMemCreate", rc));
    p.mem_handle_[d] = mh;
    rc = a.mc_bind(mch, 0, mh, 0, p.bytes, 0);
    if (rc != CUDA_SUCCESS) return fail(describe(a, "cuMulticastBindMem", rc));
    CUdeviceptr va = 0;
    rc = a.mem_reserve(&va, p.bytes, g, 0, 0);
    if (rc != CUDA_SUCCESS) return fail(describe(a, "cuMemAddressReserve", rc));
    rc = a.mem_map(va, p.bytes, 0, mh, 0);
    if (rc != CUDA_SUCCESS) {
      a.mem_free(va, p.bytes);
      return fail(describe(a, "cuMemMap", rc));
    }
    p.local[d] = reinterpret_cast<unsigned char *>(va);
    rc = a.mem_access(va, p.bytes, access, static_cast<size_t>(ndev));
    if (rc != CUDA_SUCCESS) return fail(describe(a, "cuMemSetAccess", rc));
  }
  {
    CUdeviceptr va = 0;
    rc = a.mem_reserve(&va, p.bytes, g, 0, 0);
    if (rc != CUDA_SUCCESS) return fail(describe(a, "cuMemAddressReserve (multicast)", rc));
    rc = a.mem_map(va, p.bytes, 0, mch, 0);
    if (rc != CUDA_SUCCESS) {
      a.mem_free(va, p.bytes);
      return fail(describe(a, "cuMemMap (multicast)", rc));
    }
    p.mc = reinterpret_cast<unsigned char *>(va);
    rc = a.mem_access(va, p.bytes, access, static_cast<size_t>(ndev));
    if (rc != CUDA_SUCCESS) return fail(describe(a, "cuMemSetAccess (multicast)", rc));
  }
  for (int d = 0; d < ndev; ++d) {
    if (cudaSetDevice(d) != cudaSuccess || cudaMemset(p.local[d], 0, p.bytes) != cudaSuccess ||
        cudaDeviceSynchronize() != cudaSuccess) {
      cudaGetLastError();
      return fail("clearing the pages failed");
    }
  }
  p.ok = true;
  return p;
}

}  // namespace gpu
}  // namespace dmlp
