"""Does this box support NVSwitch multicast (NVLS) and POSIX-fd shareable VMM allocations?"""
import ctypes, os
cu = ctypes.CDLL("libcuda.so.1")
def chk(r, what):
    if r != 0:
        s = ctypes.c_char_p(); cu.cuGetErrorString(r, ctypes.byref(s)); print(f"{what}: CUresult {r} {s.value}"); return False
    return True
chk(cu.cuInit(0), "cuInit")
n = ctypes.c_int(); cu.cuDeviceGetCount(ctypes.byref(n)); print("devices", n.value)
for d in range(n.value):
    dev = ctypes.c_int(); cu.cuDeviceGet(ctypes.byref(dev), d)
    for name, attr in (("MULTICAST_SUPPORTED", 132), ("HANDLE_TYPE_POSIX_FD_SUPPORTED", 103), ("VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED", 102), ("HANDLE_TYPE_FABRIC_SUPPORTED", 128)):
        v = ctypes.c_int(-1); r = cu.cuDeviceGetAttribute(ctypes.byref(v), attr, dev); print(f"dev{d} {name} = {v.value} (rc {r})")
ctx = ctypes.c_void_p(); dev = ctypes.c_int(); cu.cuDeviceGet(ctypes.byref(dev), 0)
chk(cu.cuDevicePrimaryCtxRetain(ctypes.byref(ctx), dev), "ctx"); chk(cu.cuCtxSetCurrent(ctx), "setcurrent")
class Loc(ctypes.Structure): _fields_ = [("type", ctypes.c_int), ("id", ctypes.c_int)]
class AllocFlags(ctypes.Structure): _fields_ = [("compressionType", ctypes.c_ubyte), ("gpuDirectRDMACapable", ctypes.c_ubyte), ("usage", ctypes.c_ushort), ("reserved", ctypes.c_ubyte * 4)]
class Prop(ctypes.Structure): _fields_ = [("type", ctypes.c_int), ("requestedHandleTypes", ctypes.c_int), ("location", Loc), ("win32HandleMetaData", ctypes.c_void_p), ("allocFlags", AllocFlags)]
prop = Prop(); prop.type = 1; prop.requestedHandleTypes = 1; prop.location = Loc(1, 0)
gran = ctypes.c_size_t(); chk(cu.cuMemGetAllocationGranularity(ctypes.byref(gran), ctypes.byref(prop), 1), "granularity"); print("alloc granularity (recommended)", gran.value)
h = ctypes.c_ulonglong(); ok = chk(cu.cuMemCreate(ctypes.byref(h), ctypes.c_size_t(gran.value), ctypes.byref(prop), ctypes.c_ulonglong(0)), "cuMemCreate(POSIX_FD)")
if ok:
    fd = ctypes.c_int(-1); chk(cu.cuMemExportToShareableHandle(ctypes.byref(fd), h, 1, ctypes.c_ulonglong(0)), "export fd"); print("exported fd", fd.value)
class McProp(ctypes.Structure): _fields_ = [("numDevices", ctypes.c_uint), ("size", ctypes.c_size_t), ("handleTypes", ctypes.c_ulonglong), ("flags", ctypes.c_ulonglong)]
mp = McProp(max(n.value, 1), 2 << 20, 1, 0)
g2 = ctypes.c_size_t(); 
if chk(cu.cuMulticastGetGranularity(ctypes.byref(g2), ctypes.byref(mp), 1), "mc granularity"): print("multicast granularity (recommended)", g2.value)
mc = ctypes.c_ulonglong(); mp.size = max(g2.value, 2 << 20)
if chk(cu.cuMulticastCreate(ctypes.byref(mc), ctypes.byref(mp)), "cuMulticastCreate"): print("multicast object created")
