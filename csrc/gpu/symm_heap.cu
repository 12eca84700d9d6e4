#include "symm_heap.h"

#include <cuda.h>
#include <errno.h>
#include <stdlib.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include "../host/process_comm.h"
#include "cuda_check.h"

namespace mp4j {

namespace {

// Driver API through the runtime's entry-point lookup: no link-time libcuda dependency, so the
// library still loads on a box without a driver (CPU tests of the host plane).
template <typename Fn>
Fn drv(const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult st;
  cudaError_t e = cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &st);
  if (e != cudaSuccess || st != cudaDriverEntryPointSuccess || !p)
    throw Error(std::string("driver entry point not available: ") + name);
  return (Fn)p;
}

std::string cu_err(CUresult r) {
  const char* s = nullptr;
  auto f = drv<CUresult (*)(CUresult, const char**)>("cuGetErrorString");
  f(r, &s);
  return s ? s : "unknown CUDA driver error";
}

#define CU_TRY(call_expr, what)                                                  \
  do {                                                                           \
    CUresult _r = (call_expr);                                                   \
    if (_r != CUDA_SUCCESS) throw Error(std::string(what) + ": " + cu_err(_r)); \
  } while (0)

struct Drv {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemGetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*McCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*McAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*McBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                        unsigned long long);
  CUresult (*McUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
  CUresult (*McGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  Drv() {
    MemCreate = drv<decltype(MemCreate)>("cuMemCreate");
    MemRelease = drv<decltype(MemRelease)>("cuMemRelease");
    MemExport = drv<decltype(MemExport)>("cuMemExportToShareableHandle");
    MemImport = drv<decltype(MemImport)>("cuMemImportFromShareableHandle");
    MemAddressReserve = drv<decltype(MemAddressReserve)>("cuMemAddressReserve");
    MemAddressFree = drv<decltype(MemAddressFree)>("cuMemAddressFree");
    MemMap = drv<decltype(MemMap)>("cuMemMap");
    MemUnmap = drv<decltype(MemUnmap)>("cuMemUnmap");
    MemSetAccess = drv<decltype(MemSetAccess)>("cuMemSetAccess");
    MemGetGranularity = drv<decltype(MemGetGranularity)>("cuMemGetAllocationGranularity");
    DeviceGetAttribute = drv<decltype(DeviceGetAttribute)>("cuDeviceGetAttribute");
    McCreate = drv<decltype(McCreate)>("cuMulticastCreate");
    McAddDevice = drv<decltype(McAddDevice)>("cuMulticastAddDevice");
    McBindMem = drv<decltype(McBindMem)>("cuMulticastBindMem");
    McUnbind = drv<decltype(McUnbind)>("cuMulticastUnbind");
    McGetGranularity = drv<decltype(McGetGranularity)>("cuMulticastGetGranularity");
  }
};
Drv& D() {
  static Drv d;
  return d;
}

CUmemAllocationProp alloc_prop(int device) {
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

char* map_handle(CUmemGenericAllocationHandle h, size_t bytes, int device, size_t align) {
  CUdeviceptr va = 0;
  CU_TRY(D().MemAddressReserve(&va, bytes, align, 0, 0), "cuMemAddressReserve");
  CU_TRY(D().MemMap(va, bytes, 0, h, 0), "cuMemMap");
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CU_TRY(D().MemSetAccess(va, bytes, &acc, 1), "cuMemSetAccess");
  return (char*)va;
}

// ---- fd passing over abstract-namespace Unix-domain sockets
std::string uds_name(int master_port_tag, int rank) {
  return "mp4j-heap-" + std::to_string(master_port_tag) + "-" + std::to_string(rank);
}
int uds_listen(const std::string& name) {
  int fd = ::socket(AF_UNIX, SOCK_STREAM, 0);
  if (fd < 0) throw Error("uds socket failed");
  sockaddr_un a{};
  a.sun_family = AF_UNIX;
  memcpy(a.sun_path + 1, name.data(), name.size());  // leading NUL = abstract namespace
  socklen_t len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + name.size());
  if (::bind(fd, (sockaddr*)&a, len) < 0 || ::listen(fd, 64) < 0) {
    ::close(fd);
    throw Error("uds bind/listen failed for " + name + ": " + strerror(errno));
  }
  return fd;
}
int uds_connect(const std::string& name) {
  for (int attempt = 0; attempt < 600; ++attempt) {
    int fd = ::socket(AF_UNIX, SOCK_STREAM, 0);
    sockaddr_un a{};
    a.sun_family = AF_UNIX;
    memcpy(a.sun_path + 1, name.data(), name.size());
    socklen_t len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + name.size());
    if (::connect(fd, (sockaddr*)&a, len) == 0) return fd;
    ::close(fd);
    usleep(50 * 1000);
  }
  throw Error("uds connect failed for " + name);
}
void send_fds(int sock, int src_rank, const std::vector<int>& fds) {
  int32_t hdr[2] = {src_rank, (int32_t)fds.size()};
  iovec iov{hdr, sizeof(hdr)};
  char ctrl[CMSG_SPACE(sizeof(int) * 8)] = {};
  msghdr msg{};
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  if (!fds.empty()) {  // a rank with nothing to share still announces itself (header only)
    msg.msg_control = ctrl;
    msg.msg_controllen = CMSG_SPACE(sizeof(int) * fds.size());
    cmsghdr* c = CMSG_FIRSTHDR(&msg);
    c->cmsg_level = SOL_SOCKET;
    c->cmsg_type = SCM_RIGHTS;
    c->cmsg_len = CMSG_LEN(sizeof(int) * fds.size());
    memcpy(CMSG_DATA(c), fds.data(), sizeof(int) * fds.size());
  }
  if (::sendmsg(sock, &msg, 0) < 0) throw Error(std::string("sendmsg(SCM_RIGHTS): ") + strerror(errno));
}
void recv_fds(int sock, int* src_rank, std::vector<int>* fds) {
  int32_t hdr[2] = {-1, 0};
  iovec iov{hdr, sizeof(hdr)};
  char ctrl[CMSG_SPACE(sizeof(int) * 8)] = {};
  msghdr msg{};
  msg.msg_iov = &iov;
  msg.msg_iovlen = 1;
  msg.msg_control = ctrl;
  msg.msg_controllen = sizeof(ctrl);
  ssize_t k = ::recvmsg(sock, &msg, MSG_WAITALL);
  if (k != (ssize_t)sizeof(hdr)) throw Error("recvmsg(SCM_RIGHTS) failed");
  *src_rank = hdr[0];
  fds->assign(hdr[1], -1);
  if (hdr[1] == 0) return;
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (!c || c->cmsg_type != SCM_RIGHTS) throw Error("no fds in control message");
  memcpy(fds->data(), CMSG_DATA(c), sizeof(int) * hdr[1]);
}

bool all_ok(ProcessComm* pc, bool mine) {
  char b = mine ? 1 : 0;
  auto parts = pc->master().gather_all(pc->rank(), &b, 1);
  for (auto& p : parts)
    if (p.size() != 1 || p[0] != 1) return false;
  return true;
}

}  // namespace

SymmHeap::~SymmHeap() { destroy(); }

void SymmHeap::exchange_fds(ProcessComm* pc, const std::vector<int>& send, std::vector<std::vector<int>>* recv) {
  // tag = port of rank 0's data listener: unique per job on this host
  const std::string& a0 = pc->addresses()[0];
  int tag = atoi(a0.c_str() + a0.rfind('#') + 1);
  static int round = 0;  // several exchanges per job (heap, multicast)
  int my_round = round++;
  int lfd = uds_listen(uds_name(tag * 16 + my_round, rank_));
  all_ok(pc, true);  // everyone is listening before anyone connects
  for (int q = 0; q < world_; ++q) {
    if (q == rank_) continue;
    int s = uds_connect(uds_name(tag * 16 + my_round, q));
    send_fds(s, rank_, send);
    ::close(s);
  }
  recv->assign(world_, std::vector<int>());
  for (int i = 0; i < world_ - 1; ++i) {
    int s = ::accept(lfd, nullptr, nullptr);
    if (s < 0) throw Error("uds accept failed");
    int src;
    std::vector<int> fds;
    recv_fds(s, &src, &fds);
    ::close(s);
    if (src < 0 || src >= world_) throw Error("bad rank in fd exchange");
    (*recv)[src] = fds;
  }
  ::close(lfd);
}

bool SymmHeap::try_vmm_local(size_t bytes) {
  try {
    CUmemAllocationProp prop = alloc_prop(device_);
    CUmemGenericAllocationHandle h;
    if (D().MemCreate(&h, bytes, &prop, 0) != CUDA_SUCCESS) return false;
    local_handle_ = h;
    return true;
  } catch (const Error&) {
    return false;
  }
}

void SymmHeap::init_real(ProcessComm* pc, int device, size_t bytes, bool want_multicast) {
  world_ = pc->size();
  rank_ = pc->rank();
  device_ = device;
  if (world_ > kMaxRanks) throw Error("GPU plane supports at most 8 ranks per node");
  MP4J_CUDA(cudaSetDevice(device));
  MP4J_CUDA(cudaFree(0));
  // granularity: multicast wants (much) coarser alignment than plain VMM
  size_t gran = 2 << 20;
  bool mc_possible = false;
  try {
    CUmemAllocationProp prop = alloc_prop(device);
    D().MemGetGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
    int mc = 0;
    D().DeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device);
    mc_possible = want_multicast && mc && world_ > 1;
    if (mc_possible) {
      CUmulticastObjectProp mp{};
      mp.numDevices = world_;
      mp.size = bytes;
      mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
      size_t mg = 0;
      if (D().McGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran)
        gran = mg;
    }
  } catch (const Error& e) {
    notes_ += std::string("vmm probe: ") + e.what() + "; ";
  }
  bytes_ = (bytes + gran - 1) / gran * gran;

  bool vmm_env = !(getenv("MP4J_NO_VMM") && atoi(getenv("MP4J_NO_VMM")));
  bool local_ok = vmm_env && try_vmm_local(bytes_);
  int my_fd = -1;
  if (local_ok) {
    if (D().MemExport(&my_fd, local_handle_, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) != CUDA_SUCCESS) {
      local_ok = false;
      my_fd = -1;
    }
  }
  vmm_ = all_ok(pc, local_ok);
  if (vmm_) {
    base_[rank_] = map_handle(local_handle_, bytes_, device_, gran);
    std::vector<std::vector<int>> got;
    exchange_fds(pc, {my_fd}, &got);
    ::close(my_fd);
    for (int q = 0; q < world_; ++q) {
      if (q == rank_) continue;
      CUmemGenericAllocationHandle h;
      CU_TRY(D().MemImport(&h, (void*)(uintptr_t)got[q][0], CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
             "cuMemImportFromShareableHandle");
      ::close(got[q][0]);
      peer_handles_[q] = h;
      base_[q] = map_handle(h, bytes_, device_, gran);
    }
    // ---- multicast (NVLS)
    bool mc_ok = all_ok(pc, mc_possible);
    if (mc_ok) {
      CUmulticastObjectProp mp{};
      mp.numDevices = world_;
      mp.size = bytes_;
      mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
      bool ok = true;
      int mc_fd = -1;
      CUmemGenericAllocationHandle mch = 0;
      if (rank_ == 0) {
        ok = D().McCreate(&mch, &mp) == CUDA_SUCCESS &&
             D().MemExport(&mc_fd, mch, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) == CUDA_SUCCESS;
      }
      if (all_ok(pc, ok)) {
        std::vector<std::vector<int>> got2;
        exchange_fds(pc, rank_ == 0 ? std::vector<int>{mc_fd} : std::vector<int>{}, &got2);
        if (rank_ == 0) ::close(mc_fd);
        else {
          ok = D().MemImport(&mch, (void*)(uintptr_t)got2[0][0], CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) ==
               CUDA_SUCCESS;
          ::close(got2[0][0]);
        }
        if (ok) ok = D().McAddDevice(mch, device_) == CUDA_SUCCESS;
        if (all_ok(pc, ok)) {  // every device added before anyone binds
          ok = D().McBindMem(mch, 0, local_handle_, 0, bytes_, 0) == CUDA_SUCCESS;
          mc_bound_ = ok;
          if (all_ok(pc, ok)) {
            try {
              mc_base_ = map_handle(mch, bytes_, device_, gran);
            } catch (const Error& e) {
              notes_ += std::string("multicast map failed: ") + e.what() + "; ";
              mc_base_ = nullptr;
            }
            if (!all_ok(pc, mc_base_ != nullptr)) mc_base_ = nullptr;
          } else notes_ += "cuMulticastBindMem failed on some rank; ";
        } else notes_ += "cuMulticastAddDevice failed on some rank; ";
        mc_handle_ = mch;
      } else notes_ += "cuMulticastCreate/export failed; ";
    } else if (want_multicast && world_ > 1) {
      notes_ += "multicast not supported on some rank; ";
    }
  } else {
    // ---- cudaMalloc + CUDA IPC fallback
    if (local_handle_) {
      D().MemRelease(local_handle_);
      local_handle_ = 0;
    }
    notes_ += "VMM path unavailable, using cudaIpc; ";
    void* p = nullptr;
    MP4J_CUDA(cudaMalloc(&p, bytes_));
    base_[rank_] = (char*)p;
    cudaIpcMemHandle_t h;
    MP4J_CUDA(cudaIpcGetMemHandle(&h, p));
    auto parts = pc->master().gather_all(rank_, &h, sizeof(h));
    for (int q = 0; q < world_; ++q) {
      if (q == rank_) continue;
      cudaIpcMemHandle_t ph;
      memcpy(&ph, parts[q].data(), sizeof(ph));
      void* pp = nullptr;
      MP4J_CUDA(cudaIpcOpenMemHandle(&pp, ph, cudaIpcMemLazyEnablePeerAccess));
      base_[q] = (char*)pp;
      ipc_opened_[q] = true;
    }
  }
  MP4J_CUDA(cudaMemset(base_[rank_], 0, kSigPadBytes));
  MP4J_CUDA(cudaDeviceSynchronize());
  all_ok(pc, true);  // pads are zero everywhere before any kernel signals
}

void SymmHeap::init_emulated(int world, int device, size_t bytes) {
  if (world < 1 || world > kMaxRanks) throw Error("emulated world must be in [1, 8]");
  world_ = world;
  rank_ = 0;
  device_ = device;
  emulated_ = true;
  bytes_ = (bytes + 255) / 256 * 256;
  MP4J_CUDA(cudaSetDevice(device));
  for (int q = 0; q < world; ++q) {
    void* p = nullptr;
    MP4J_CUDA(cudaMalloc(&p, bytes_));
    base_[q] = (char*)p;
    MP4J_CUDA(cudaMemset(p, 0, kSigPadBytes));
  }
  MP4J_CUDA(cudaDeviceSynchronize());
}

void SymmHeap::destroy() {
  if (world_ == 0) return;
  cudaDeviceSynchronize();
  if (emulated_) {
    for (int q = 0; q < world_; ++q)
      if (base_[q]) cudaFree(base_[q]);
  } else if (vmm_) {
    try {
      if (mc_base_) {
        D().MemUnmap((CUdeviceptr)mc_base_, bytes_);
        D().MemAddressFree((CUdeviceptr)mc_base_, bytes_);
      }
      if (mc_handle_) {
        if (mc_bound_) D().McUnbind(mc_handle_, device_, 0, bytes_);
        D().MemRelease(mc_handle_);
      }
      for (int q = 0; q < world_; ++q) {
        if (!base_[q]) continue;
        D().MemUnmap((CUdeviceptr)base_[q], bytes_);
        D().MemAddressFree((CUdeviceptr)base_[q], bytes_);
        if (q != rank_ && peer_handles_[q]) D().MemRelease(peer_handles_[q]);
      }
      if (local_handle_) D().MemRelease(local_handle_);
    } catch (const Error&) {
    }
  } else {
    for (int q = 0; q < world_; ++q) {
      if (q == rank_) {
        if (base_[q]) cudaFree(base_[q]);
      } else if (ipc_opened_[q]) {
        cudaIpcCloseMemHandle(base_[q]);
      }
    }
  }
  world_ = 0;
}

}  // namespace mp4j
