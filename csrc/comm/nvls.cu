// NVLS path (K13 / W6): in-switch reduction and multicast through NVSwitch.
//
//   multimem.ld_reduce.add.f32   one load returns the sum of the same address on every GPU - the
//                                reduction the reference does with 8 fadd lanes per NIC
//                                (hw/all_reduce.sv:1090-1101) happens inside the switch;
//   multimem.st                  one store lands in every GPU's copy - the all-gather half of the
//                                ring (FORWARD_OUTPUT / OUTPUT states, hw/all_reduce.sv:996-1086).
// Per GPU and direction about S bytes cross NVLink instead of 2(N-1)/N * S for a P2P two-shot.
// The switch cannot add BFP/MX payloads, so this path carries uncompressed fp32 gradients.
//
// Needs a multicast object (cuMulticastCreate) bound to one physical allocation per GPU, so the
// gradient buckets for this path live in a separate VMM-backed "NVLS heap":
//   stage 1  every rank: cuMemCreate + map its slice; rank 0 creates the multicast object and
//            hands its POSIX fd to the other ranks over an abstract unix socket (SCM_RIGHTS);
//            everyone adds its device;                                   <barrier>
//   stage 2  everyone binds its physical memory to the multicast object; <barrier>
//   stage 3  everyone maps the multicast object -> mc pointer.
// Driver entry points are resolved at run time (cudaGetDriverEntryPoint) so the library still
// loads on machines without libcuda.  If anything is unsupported the caller gets
// AINIC_ERR_UNSUPPORTED with a reason and falls back to the P2P kernels.
#include <cuda.h>
#include <cuda_runtime.h>
#include <errno.h>
#include <poll.h>
#include <stddef.h>
#include <stdio.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <string>

#include "allreduce.cuh"

using namespace ainic;
using namespace ainic::comm;

extern "C" void ainic_set_error_(const char* msg);

namespace {

struct Drv {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*DeviceGet)(CUdevice*, int);
  bool ok = false;
};

template <class F>
bool resolve(F& fn, const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) return false;
  fn = reinterpret_cast<F>(p);
  return true;
}

Drv& drv() {
  static Drv d;
  static bool tried = false;
  if (!tried) {
    tried = true;
    d.ok = resolve(d.MemCreate, "cuMemCreate") && resolve(d.MemRelease, "cuMemRelease") &&
           resolve(d.MemAddressReserve, "cuMemAddressReserve") && resolve(d.MemAddressFree, "cuMemAddressFree") &&
           resolve(d.MemMap, "cuMemMap") && resolve(d.MemUnmap, "cuMemUnmap") && resolve(d.MemSetAccess, "cuMemSetAccess") &&
           resolve(d.MemGetAllocationGranularity, "cuMemGetAllocationGranularity") &&
           resolve(d.MemExportToShareableHandle, "cuMemExportToShareableHandle") &&
           resolve(d.MemImportFromShareableHandle, "cuMemImportFromShareableHandle") &&
           resolve(d.MulticastCreate, "cuMulticastCreate") && resolve(d.MulticastAddDevice, "cuMulticastAddDevice") &&
           resolve(d.MulticastBindMem, "cuMulticastBindMem") && resolve(d.MulticastGetGranularity, "cuMulticastGetGranularity") &&
           resolve(d.DeviceGetAttribute, "cuDeviceGetAttribute") && resolve(d.DeviceGet, "cuDeviceGet");
  }
  return d;
}

int unsupported(const std::string& why) {
  ainic_set_error_(("nvls: " + why).c_str());
  return AINIC_ERR_UNSUPPORTED;
}

// ---- fd passing over an abstract unix socket -----------------------------------------------
sockaddr_un abstract_addr(const char* name, socklen_t* len) {
  sockaddr_un a;
  memset(&a, 0, sizeof(a));
  a.sun_family = AF_UNIX;
  const size_t n = strnlen(name, sizeof(a.sun_path) - 2);
  memcpy(a.sun_path + 1, name, n);  // leading NUL = abstract namespace, nothing to unlink
  *len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
  return a;
}

int send_fd_to_peers(const char* name, int fd, int npeers) {
  int ls = socket(AF_UNIX, SOCK_STREAM, 0);
  if (ls < 0) return -1;
  socklen_t len;
  sockaddr_un a = abstract_addr(name, &len);
  if (bind(ls, (sockaddr*)&a, len) < 0 || listen(ls, npeers) < 0) { close(ls); return -1; }
  int rc = 0;
  for (int i = 0; i < npeers && rc == 0; ++i) {
    pollfd pf{ls, POLLIN, 0};
    if (poll(&pf, 1, 30000) <= 0) { rc = -1; errno = ETIMEDOUT; break; }  // a peer gave up before connecting
    int cs = accept(ls, nullptr, nullptr);
    if (cs < 0) { rc = -1; break; }
    char byte = 'x';
    iovec io{&byte, 1};
    char ctrl[CMSG_SPACE(sizeof(int))];
    memset(ctrl, 0, sizeof(ctrl));
    msghdr msg{};
    msg.msg_iov = &io;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof(ctrl);
    cmsghdr* c = CMSG_FIRSTHDR(&msg);
    c->cmsg_level = SOL_SOCKET;
    c->cmsg_type = SCM_RIGHTS;
    c->cmsg_len = CMSG_LEN(sizeof(int));
    memcpy(CMSG_DATA(c), &fd, sizeof(int));
    if (sendmsg(cs, &msg, 0) < 0) rc = -1;
    close(cs);
  }
  close(ls);
  return rc;
}

int recv_fd_from_root(const char* name) {
  socklen_t len;
  sockaddr_un a = abstract_addr(name, &len);
  int s = -1;
  for (int attempt = 0; attempt < 3000; ++attempt) {  // the root may not be listening yet: retry for ~30 s
    s = socket(AF_UNIX, SOCK_STREAM, 0);
    if (s < 0) return -1;
    if (connect(s, (sockaddr*)&a, len) == 0) break;
    close(s);
    s = -1;
    usleep(10000);
  }
  if (s < 0) return -1;
  char byte;
  iovec io{&byte, 1};
  char ctrl[CMSG_SPACE(sizeof(int))];
  msghdr msg{};
  msg.msg_iov = &io;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  int fd = -1;
  if (recvmsg(s, &msg, 0) > 0) {
    cmsghdr* c = CMSG_FIRSTHDR(&msg);
    if (c && c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) memcpy(&fd, CMSG_DATA(c), sizeof(int));
  }
  close(s);
  return fd;
}

}  // namespace

struct ainic_nvls {
  int rank = 0, nranks = 1, device = 0;
  size_t bytes = 0, bump = 0;
  CUmemGenericAllocationHandle phys = 0, mc = 0;
  CUdeviceptr uc_va = 0, mc_va = 0;
  bool bound = false, mapped = false;
};

extern "C" int ainic_nvls_supported(void) {
  Drv& d = drv();
  if (!d.ok) return 0;
  int dev = 0, v = 0;
  CUdevice cud;
  if (cudaGetDevice(&dev) != cudaSuccess || d.DeviceGet(&cud, dev) != CUDA_SUCCESS) return 0;
  if (d.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cud) != CUDA_SUCCESS) return 0;
  return v;
}

extern "C" int ainic_nvls_stage1(ainic_nvls_t** out, int rank, int nranks, size_t bytes, const char* sock_name) {
  if (!out || nranks < 2 || rank < 0 || rank >= nranks || !sock_name) return AINIC_ERR_ARG;
  Drv& d = drv();
  if (!d.ok) return unsupported("driver entry points for cuMem*/cuMulticast* not available");
  if (!ainic_nvls_supported()) return unsupported("CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED is 0 on this device");
  cudaFree(0);  // make sure the primary context exists
  ainic_nvls* h = new ainic_nvls();
  h->rank = rank;
  h->nranks = nranks;
  cudaGetDevice(&h->device);
  CUmulticastObjectProp mp{};
  mp.numDevices = (unsigned)nranks;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  mp.flags = 0;
  size_t gran_mc = 0, gran_mem = 0;
  mp.size = bytes;
  if (d.MulticastGetGranularity(&gran_mc, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) { delete h; return unsupported("cuMulticastGetGranularity failed"); }
  CUmemAllocationProp ap{};
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = h->device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  if (d.MemGetAllocationGranularity(&gran_mem, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) { delete h; return unsupported("cuMemGetAllocationGranularity failed"); }
  const size_t gran = gran_mc > gran_mem ? gran_mc : gran_mem;
  h->bytes = (bytes + gran - 1) / gran * gran;
  mp.size = h->bytes;
  CUresult r;
  if ((r = d.MemCreate(&h->phys, h->bytes, &ap, 0)) != CUDA_SUCCESS) { delete h; return unsupported("cuMemCreate failed (" + std::to_string((int)r) + ")"); }
  if ((r = d.MemAddressReserve(&h->uc_va, h->bytes, gran, 0, 0)) != CUDA_SUCCESS ||
      (r = d.MemMap(h->uc_va, h->bytes, 0, h->phys, 0)) != CUDA_SUCCESS) { delete h; return unsupported("mapping the local slice failed (" + std::to_string((int)r) + ")"); }
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = h->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  if ((r = d.MemSetAccess(h->uc_va, h->bytes, &acc, 1)) != CUDA_SUCCESS) { delete h; return unsupported("cuMemSetAccess failed"); }
  cudaMemset(reinterpret_cast<void*>(h->uc_va), 0, h->bytes);
  cudaDeviceSynchronize();
  // multicast object: created by rank 0, imported by everyone else
  if (rank == 0) {
    if ((r = d.MulticastCreate(&h->mc, &mp)) != CUDA_SUCCESS) { delete h; return unsupported("cuMulticastCreate failed (" + std::to_string((int)r) + ")"); }
    int fd = -1;
    if ((r = d.MemExportToShareableHandle(&fd, h->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0)) != CUDA_SUCCESS) { delete h; return unsupported("exporting the multicast handle failed"); }
    const int src = send_fd_to_peers(sock_name, fd, nranks - 1);
    close(fd);
    if (src != 0) { delete h; return unsupported(std::string("passing the multicast fd failed: ") + strerror(errno)); }
  } else {
    const int fd = recv_fd_from_root(sock_name);
    if (fd < 0) { delete h; return unsupported("did not receive the multicast fd from rank 0"); }
    r = d.MemImportFromShareableHandle(&h->mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    close(fd);
    if (r != CUDA_SUCCESS) { delete h; return unsupported("importing the multicast handle failed (" + std::to_string((int)r) + ")"); }
  }
  CUdevice cud;
  d.DeviceGet(&cud, h->device);
  if ((r = d.MulticastAddDevice(h->mc, cud)) != CUDA_SUCCESS) { delete h; return unsupported("cuMulticastAddDevice failed (" + std::to_string((int)r) + ")"); }
  *out = h;
  return AINIC_OK;
}

// All ranks must have returned from stage 1 (every device added) before anyone binds.
extern "C" int ainic_nvls_stage2(ainic_nvls_t* h) {
  if (!h) return AINIC_ERR_ARG;
  CUresult r = drv().MulticastBindMem(h->mc, 0, h->phys, 0, h->bytes, 0);
  if (r != CUDA_SUCCESS) return unsupported("cuMulticastBindMem failed (" + std::to_string((int)r) + ")");
  h->bound = true;
  return AINIC_OK;
}

// All ranks must have bound before the multicast mapping is used.
extern "C" int ainic_nvls_stage3(ainic_nvls_t* h) {
  if (!h || !h->bound) return AINIC_ERR_ARG;
  Drv& d = drv();
  CUresult r;
  if ((r = d.MemAddressReserve(&h->mc_va, h->bytes, 2u << 20, 0, 0)) != CUDA_SUCCESS ||
      (r = d.MemMap(h->mc_va, h->bytes, 0, h->mc, 0)) != CUDA_SUCCESS)
    return unsupported("mapping the multicast object failed (" + std::to_string((int)r) + ")");
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = h->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  if ((r = d.MemSetAccess(h->mc_va, h->bytes, &acc, 1)) != CUDA_SUCCESS) return unsupported("cuMemSetAccess(mc) failed");
  h->mapped = true;
  return AINIC_OK;
}

extern "C" void* ainic_nvls_alloc(ainic_nvls_t* h, size_t bytes) {
  if (!h || !bytes) return nullptr;
  const size_t off = (h->bump + 4095) / 4096 * 4096;
  if (off + bytes > h->bytes) return nullptr;
  h->bump = off + bytes;
  return reinterpret_cast<void*>(h->uc_va + off);
}
extern "C" void* ainic_nvls_mc_ptr(ainic_nvls_t* h, const void* uc_ptr) {
  if (!h || !h->mapped) return nullptr;
  const CUdeviceptr p = reinterpret_cast<CUdeviceptr>(uc_ptr);
  if (p < h->uc_va || p >= h->uc_va + h->bytes) return nullptr;
  return reinterpret_cast<void*>(h->mc_va + (p - h->uc_va));
}
extern "C" int ainic_nvls_destroy(ainic_nvls_t* h) {
  if (!h) return AINIC_ERR_ARG;
  Drv& d = drv();
  cudaDeviceSynchronize();
  if (h->mc_va) { d.MemUnmap(h->mc_va, h->bytes); d.MemAddressFree(h->mc_va, h->bytes); }
  if (h->uc_va) { d.MemUnmap(h->uc_va, h->bytes); d.MemAddressFree(h->uc_va, h->bytes); }
  if (h->mc) d.MemRelease(h->mc);
  if (h->phys) d.MemRelease(h->phys);
  delete h;
  return AINIC_OK;
}

// ------------------------------------------------------------------------------------------ kernel
namespace ainic {
namespace comm {

struct NvlsCtx {
  RankCtx base;        // flags / counters / weights / count / lr / epoch (flags live in the IPC heap)
  float* grad_mc;      // multicast alias of base.grad
};

__device__ __forceinline__ void fence_alias() { asm volatile("fence.proxy.alias;\n" ::: "memory"); }

// Start barrier -> owner: in-switch reduce of its shard (multimem.ld_reduce), broadcast of the sum
// into every rank's gradient buffer (multimem.st), SGD on its shard -> barrier -> everyone applies
// the other shards from its LOCAL gradient buffer.
__global__ void __launch_bounds__(kCommThreads) allreduce_nvls_kernel(const NvlsCtx x) {
  const RankCtx& c = x.base;
  __shared__ WaitStats ws;
  if (threadIdx.x == 0) begin_call(c, ws);
  const unsigned long long t0 = ptx::globaltimer_ns();
  __syncthreads();
  const int cta = blockIdx.x, nctas = gridDim.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int N = c.nranks;
  const size_t frames = c.count / kFrameVals;       // whole frames only (host enforces count % 512 == 0)
  const size_t shard = (frames + N - 1) / N;
  signal_all(c, 0, cta, ws.epoch);
  wait_all(c, 0, cta, ws.epoch, ws);
  fence_alias();
  float v[16];
  for (size_t j = (size_t)cta + (size_t)warp * nctas; j < shard; j += (size_t)nctas * wpb) {
    const size_t f = (size_t)c.rank * shard + j;
    if (f >= frames) break;
    const float* src = x.grad_mc + f * kFrameVals;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 s = ptx::multimem_ld_reduce_f32x4(src + i * 128 + lane * 4);
      v[4 * i] = s.x; v[4 * i + 1] = s.y; v[4 * i + 2] = s.z; v[4 * i + 3] = s.w;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
      ptx::multimem_st_v4(x.grad_mc + f * kFrameVals + i * 128 + lane * 4,
                          make_uint4(__float_as_uint(v[4 * i]), __float_as_uint(v[4 * i + 1]), __float_as_uint(v[4 * i + 2]),
                                     __float_as_uint(v[4 * i + 3])));
    apply_frame(c, f, v, lane);
  }
  fence_alias();
  signal_all(c, 1, cta, ws.epoch);
  wait_all(c, 1, cta, ws.epoch, ws);
  fence_alias();
  for (int dp = 1; dp < N; ++dp) {
    const int p = (c.rank + dp) % N;
    for (size_t j = (size_t)cta + (size_t)warp * nctas; j < shard; j += (size_t)nctas * wpb) {
      const size_t f = (size_t)p * shard + j;
      if (f >= frames) break;
      frame_load_f32(c.grad + f * kFrameVals, kFrameVals, v, lane);
      apply_frame(c, f, v, lane);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned long long t1 = ptx::globaltimer_ns();
    if (ws.timeouts) atomicAdd(c.counters + CTR_TIMEOUTS, (unsigned long long)ws.timeouts);
    if (cta == 0) {
      atomicAdd(c.counters + CTR_BUSY_NS, t1 - t0);
      atomicAdd(c.counters + CTR_WAIT_NS, ws.wait_ns);
      atomicAdd(c.counters + CTR_LAUNCHES, 1ull);
      c.counters[CTR_LAST_NS] = t1 - t0;
    }
    end_call(c, ws, nctas);
  }
}

// ---- bf16 gradients (two bytes per value on the wire and in HBM, no encode / decode work) --------------
// The gradient bucket holds bf16 values (written by a cast kernel or a bf16 GEMM epilogue).  Unit of
// work = 256 values = one 16-byte multimem operation per lane; lane l owns values 8 l .. 8 l + 7 of the
// unit, so its weights are two adjacent float4 and its shadow one 16-byte store.
__device__ __forceinline__ void apply_bf16x8(const RankCtx& c, size_t idx, uint4 g) {
  if (!c.weight) return;
  float4 w0 = *reinterpret_cast<const float4*>(c.weight + idx);
  float4 w1 = *reinterpret_cast<const float4*>(c.weight + idx + 4);
  const float nlr = -c.lr;
  w0.x = fmaf(nlr, __uint_as_float(g.x << 16), w0.x);
  w0.y = fmaf(nlr, __uint_as_float(g.x & 0xFFFF0000u), w0.y);
  w0.z = fmaf(nlr, __uint_as_float(g.y << 16), w0.z);
  w0.w = fmaf(nlr, __uint_as_float(g.y & 0xFFFF0000u), w0.w);
  w1.x = fmaf(nlr, __uint_as_float(g.z << 16), w1.x);
  w1.y = fmaf(nlr, __uint_as_float(g.z & 0xFFFF0000u), w1.y);
  w1.z = fmaf(nlr, __uint_as_float(g.w << 16), w1.z);
  w1.w = fmaf(nlr, __uint_as_float(g.w & 0xFFFF0000u), w1.w);
  *reinterpret_cast<float4*>(c.weight + idx) = w0;
  *reinterpret_cast<float4*>(c.weight + idx + 4) = w1;
  if (c.wshadow) {
    __nv_bfloat162 a = __floats2bfloat162_rn(w0.x, w0.y), b = __floats2bfloat162_rn(w0.z, w0.w);
    __nv_bfloat162 d = __floats2bfloat162_rn(w1.x, w1.y), e = __floats2bfloat162_rn(w1.z, w1.w);
    uint4 o;
    o.x = *reinterpret_cast<uint32_t*>(&a);
    o.y = *reinterpret_cast<uint32_t*>(&b);
    o.z = *reinterpret_cast<uint32_t*>(&d);
    o.w = *reinterpret_cast<uint32_t*>(&e);
    *reinterpret_cast<uint4*>(c.wshadow + idx) = o;
  }
}

// Start barrier -> owner: in-switch sum of its shard (multimem.ld_reduce, fp32 accumulation, bf16 result),
// multicast of the sum into every rank's bucket, SGD on the shard -> barrier -> everyone applies the other
// shards from its local bucket.  Every rank applies the same bf16 bits, so replicas stay identical.
__global__ void __launch_bounds__(kCommThreads) allreduce_nvls_bf16_kernel(const NvlsCtx x) {
  const RankCtx& c = x.base;
  __shared__ WaitStats ws;
  if (threadIdx.x == 0) begin_call(c, ws);
  const unsigned long long t0 = ptx::globaltimer_ns();
  __syncthreads();
  const int cta = blockIdx.x, nctas = gridDim.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int N = c.nranks;
  constexpr size_t kUnit = 256;                       // values per warp-wide 16-byte operation
  constexpr int kBatch = 4;                           // units in flight per warp
  const size_t units = c.count / kUnit;               // host enforces count % 512 == 0
  const size_t shard = ((units + N - 1) / N + kBatch - 1) / kBatch * kBatch;
  const __nv_bfloat16* gl = reinterpret_cast<const __nv_bfloat16*>(c.grad);
  __nv_bfloat16* gmc = reinterpret_cast<__nv_bfloat16*>(x.grad_mc);
  signal_all(c, 0, cta, ws.epoch);
  wait_all(c, 0, cta, ws.epoch, ws);
  fence_alias();
  const size_t stride = (size_t)nctas * wpb * kBatch;
  for (size_t j = ((size_t)cta + (size_t)warp * nctas) * kBatch; j < shard; j += stride) {
    const size_t u0 = (size_t)c.rank * shard + j;
    uint4 s[kBatch];
#pragma unroll
    for (int i = 0; i < kBatch; ++i)
      if (u0 + i < units) s[i] = ptx::multimem_ld_reduce_bf16x8(gmc + (u0 + i) * kUnit + lane * 8);
#pragma unroll
    for (int i = 0; i < kBatch; ++i)
      if (u0 + i < units) ptx::multimem_st_bf16x8(gmc + (u0 + i) * kUnit + lane * 8, s[i]);
#pragma unroll
    for (int i = 0; i < kBatch; ++i)
      if (u0 + i < units) apply_bf16x8(c, (u0 + i) * kUnit + lane * 8, s[i]);
  }
  fence_alias();
  signal_all(c, 1, cta, ws.epoch);
  wait_all(c, 1, cta, ws.epoch, ws);
  fence_alias();
  for (int dp = 1; dp < N; ++dp) {
    const int p = (c.rank + dp) % N;
    for (size_t j = ((size_t)cta + (size_t)warp * nctas) * kBatch; j < shard; j += stride) {
      const size_t u0 = (size_t)p * shard + j;
      uint4 s[kBatch];
#pragma unroll
      for (int i = 0; i < kBatch; ++i)
        if (u0 + i < units) s[i] = ptx::ld_stream_v4(gl + (u0 + i) * kUnit + lane * 8);
#pragma unroll
      for (int i = 0; i < kBatch; ++i)
        if (u0 + i < units) apply_bf16x8(c, (u0 + i) * kUnit + lane * 8, s[i]);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned long long t1 = ptx::globaltimer_ns();
    if (ws.timeouts) atomicAdd(c.counters + CTR_TIMEOUTS, (unsigned long long)ws.timeouts);
    if (cta == 0) {
      atomicAdd(c.counters + CTR_BUSY_NS, t1 - t0);
      atomicAdd(c.counters + CTR_WAIT_NS, ws.wait_ns);
      atomicAdd(c.counters + CTR_LAUNCHES, 1ull);
      c.counters[CTR_LAST_NS] = t1 - t0;
    }
    end_call(c, ws, nctas);
  }
}

}  // namespace comm
}  // namespace ainic

// Implemented in comm.cu: fills the communicator-owned fields of a RankCtx for a new fused call.
extern "C" int ainic_comm_fill_ctx_(ainic_comm_t* c, RankCtx* ctx, const ainic_allreduce_args_t* a);

extern "C" int ainic_allreduce_nvls(ainic_comm_t* c, ainic_nvls_t* h, const ainic_allreduce_args_t* a, void* stream) {
  if (!c || !h || !a || !a->grad || !h->mapped) return AINIC_ERR_ARG;
  if (a->count == 0 || (a->count % kFrameVals)) { ainic_set_error_("nvls: count must be a multiple of 512"); return AINIC_ERR_ARG; }
  void* mc = ainic_nvls_mc_ptr(h, a->grad);
  if (!mc) { ainic_set_error_("nvls: the gradient bucket must be allocated from the NVLS heap"); return AINIC_ERR_ARG; }
  NvlsCtx x{};
  int rc = ainic_comm_fill_ctx_(c, &x.base, a);
  if (rc) return rc;
  x.grad_mc = static_cast<float*>(mc);
  const size_t frames = a->count / kFrameVals;
  size_t ctas = (frames + 15) / 16;
  if (a->max_ctas > 0) ctas = (size_t)a->max_ctas;
  if (ctas < 1) ctas = 1;
  if (ctas > 148) ctas = 148;
  if (a->codec == AINIC_CODEC_BF16) {   // the bucket holds bf16 values: half the wire and HBM bytes of the fp32 form
    if (a->grad_out) { ainic_set_error_("nvls: grad_out is not offered with a bf16 bucket"); return AINIC_ERR_UNSUPPORTED; }
    allreduce_nvls_bf16_kernel<<<(unsigned)ctas, kCommThreadsOverlap, 0, (cudaStream_t)stream>>>(x);
  } else {
    allreduce_nvls_kernel<<<(unsigned)ctas, kCommThreadsOverlap, 0, (cudaStream_t)stream>>>(x);
  }
  return cudaGetLastError() == cudaSuccess ? AINIC_OK : AINIC_ERR_CUDA;
}
