"""Probe: does this box support NVLink multicast (NVLS) and which handle types?"""
import os, subprocess
from cuda.bindings import driver as cu
def ck(r):
    if r[0] != cu.CUresult.CUDA_SUCCESS:
        raise RuntimeError(str(r[0]))
    return r[1] if len(r) == 2 else r[1:]
ck(cu.cuInit(0))
n = ck(cu.cuDeviceGetCount())
print("devices", n)
A = cu.CUdevice_attribute
for d in range(n):
    dev = ck(cu.cuDeviceGet(d))
    vals = {}
    for name in ("CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED", "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED",
                 "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED", "CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED",
                 "CU_DEVICE_ATTRIBUTE_GPU_DIRECT_RDMA_WITH_CUDA_VMM_SUPPORTED"):
        vals[name.replace("CU_DEVICE_ATTRIBUTE_", "")] = ck(cu.cuDeviceGetAttribute(getattr(A, name), dev))
    print(d, vals)
# multicast granularity for 2..n devices
try:
    prop = cu.CUmulticastObjectProp()
    prop.numDevices = n
    prop.size = 1 << 30
    prop.handleTypes = cu.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR
    g_min = ck(cu.cuMulticastGetGranularity(prop, cu.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_MINIMUM))
    g_rec = ck(cu.cuMulticastGetGranularity(prop, cu.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED))
    print("multicast granularity min/rec", g_min, g_rec)
    dev0 = ck(cu.cuDeviceGet(0))
    ctx = ck(cu.cuDevicePrimaryCtxRetain(dev0))
    ck(cu.cuCtxSetCurrent(ctx))
    mc = ck(cu.cuMulticastCreate(prop))
    print("cuMulticastCreate ok", mc)
    fd = ck(cu.cuMemExportToShareableHandle(mc, cu.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0))
    print("export fd ok", fd)
except Exception as e:
    print("multicast probe failed:", e)
print("ptrace_scope", open("/proc/sys/kernel/yama/ptrace_scope").read().strip() if os.path.exists("/proc/sys/kernel/yama/ptrace_scope") else "n/a")
print(subprocess.run("grep Cap /proc/self/status; uname -r", shell=True, capture_output=True, text=True).stdout)
