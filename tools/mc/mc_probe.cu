// tools/mc/mc_probe.cu -- does this box allow NVSwitch multicast (NVLS) objects?  (dev tool)
// Build: nvcc -o mc_probe mc_probe.cu -lcuda    Run on the GPU box.
#include <cstdio>
#include <cstring>
#include <cuda.h>
#include <cuda_runtime.h>
#define DRV(x) do { CUresult r = (x); if (r != CUDA_SUCCESS) { const char *s = nullptr; cuGetErrorString(r, &s); printf("%s -> %d (%s)\n", #x, (int)r, s ? s : "?"); return 1; } } while (0)
int main() {
  int n = 0;
  cudaGetDeviceCount(&n);
  printf("devices=%d\n", n);
  DRV(cuInit(0));
  for (int d = 0; d < n; ++d) {
    int mc = -1, fabric = -1, posix = -1, vmm = -1;
    cuDeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d);
    cuDeviceGetAttribute(&fabric, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, d);
    cuDeviceGetAttribute(&posix, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, d);
    cuDeviceGetAttribute(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, d);
    printf("dev %d multicast=%d fabric_handles=%d posix_fd_handles=%d vmm=%d\n", d, mc, fabric, posix, vmm);
  }
  if (n < 2) return 0;
  for (int d = 0; d < n; ++d) { cudaSetDevice(d); cudaFree(0); }
  CUmulticastObjectProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.numDevices = n;
  prop.size = 64ull << 20;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gmin = 0, grec = 0;
  DRV(cuMulticastGetGranularity(&gmin, &prop, CU_MULTICAST_GRANULARITY_MINIMUM));
  DRV(cuMulticastGetGranularity(&grec, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  printf("multicast granularity min=%zu recommended=%zu\n", gmin, grec);
  prop.size = ((prop.size + grec - 1) / grec) * grec;
  CUmemGenericAllocationHandle mch;
  DRV(cuMulticastCreate(&mch, &prop));
  printf("cuMulticastCreate ok (size %zu)\n", (size_t)prop.size);
  for (int d = 0; d < n; ++d) DRV(cuMulticastAddDevice(mch, d));
  printf("cuMulticastAddDevice x%d ok\n", n);
  // bind one physical allocation per device and map the multicast VA on device 0
  for (int d = 0; d < n; ++d) {
    CUmemAllocationProp ap;
    memset(&ap, 0, sizeof(ap));
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = d;
    ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUmemGenericAllocationHandle h;
    DRV(cuMemCreate(&h, prop.size, &ap, 0));
    DRV(cuMulticastBindMem(mch, 0, h, 0, prop.size, 0));
  }
  printf("cuMulticastBindMem x%d ok\n", n);
  CUdeviceptr va = 0;
  DRV(cuMemAddressReserve(&va, prop.size, grec, 0, 0));
  DRV(cuMemMap(va, prop.size, 0, mch, 0));
  CUmemAccessDesc ad;
  memset(&ad, 0, sizeof(ad));
  ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ad.location.id = 0;
  ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  DRV(cuMemSetAccess(va, prop.size, &ad, 1));
  printf("multicast VA mapped on device 0: %p\nMULTICAST_OK\n", (void *)va);
  return 0;
}
