// B1 (NVLS form) — symmetric memory from the CUDA virtual-memory API with an NVSwitch multicast
// mapping on top.
//
// cudaMalloc + CUDA IPC (bootstrap.cpp) is enough for peer loads / stores, but memory can only be
// bound to a multicast object (so that `multimem.ld_reduce` / `multimem.st` reduce and broadcast
// INSIDE the switch) if it was created with cuMemCreate.  So this allocator
//   1. creates each rank's region with cuMemCreate (POSIX-fd exportable) and maps it locally,
//   2. passes the fds between the ranks over abstract-namespace Unix datagram sockets (SCM_RIGHTS),
//      imports and maps every peer's region (unicast peer pointers, same role as the IPC mapping),
//   3. lets rank 0 create the multicast object, passes its fd around, adds every device, binds every
//      rank's region at offset 0 and maps the object: one extra pointer per rank whose stores reach
//      all ranks and whose `multimem.ld_reduce` loads return the sum over all ranks.
// All driver entry points are fetched with cudaGetDriverEntryPoint (no link-time libcuda).
// The reference has no counterpart (CPU tensors over MPI); this is the "NVLS-multicast reduction"
// option of the gradient all-reduce.
#include <cuda.h>
#include <errno.h>
#include <poll.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <vector>

#include "../common.cuh"

extern "C" int b200rnn_boot_allgather(struct b200rnn_boot* b, const void* send, int32_t bytes, void* recv);
extern "C" int b200rnn_boot_barrier(struct b200rnn_boot* b);

namespace {

using b200rnn::set_error;

struct Api {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*McCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*McAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*McBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*McGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*GetErrorString)(CUresult, const char**);
  bool ok = false;
};

Api& api() {
  static Api a;
  if (a.ok) return a;
  auto get = [](const char* name, void** out) {
    cudaDriverEntryPointQueryResult q;
    return cudaGetDriverEntryPoint(name, out, cudaEnableDefault, &q) == cudaSuccess && *out != nullptr;
  };
  bool ok = true;
  ok &= get("cuMemCreate", (void**)&a.MemCreate);
  ok &= get("cuMemRelease", (void**)&a.MemRelease);
  ok &= get("cuMemAddressReserve", (void**)&a.MemAddressReserve);
  ok &= get("cuMemAddressFree", (void**)&a.MemAddressFree);
  ok &= get("cuMemMap", (void**)&a.MemMap);
  ok &= get("cuMemUnmap", (void**)&a.MemUnmap);
  ok &= get("cuMemSetAccess", (void**)&a.MemSetAccess);
  ok &= get("cuMemExportToShareableHandle", (void**)&a.MemExport);
  ok &= get("cuMemImportFromShareableHandle", (void**)&a.MemImport);
  ok &= get("cuMulticastCreate", (void**)&a.McCreate);
  ok &= get("cuMulticastAddDevice", (void**)&a.McAddDevice);
  ok &= get("cuMulticastBindMem", (void**)&a.McBindMem);
  ok &= get("cuMulticastGetGranularity", (void**)&a.McGetGranularity);
  ok &= get("cuGetErrorString", (void**)&a.GetErrorString);
  a.ok = ok;
  return a;
}

const char* cu_err(CUresult r) {
  const char* s = nullptr;
  if (api().GetErrorString && api().GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
  return "unknown CUDA driver error";
}

#define CU_TRY(expr)                                                                              \
  do {                                                                                            \
    CUresult _r = (expr);                                                                         \
    if (_r != CUDA_SUCCESS) {                                                                     \
      set_error("%s:%d: %s failed: %s (%d)", __FILE__, __LINE__, #expr, cu_err(_r), (int)_r);     \
      return B200RNN_ECUDA;                                                                       \
    }                                                                                             \
  } while (0)

// ---- fd passing over abstract-namespace Unix datagram sockets ----
void sock_name(sockaddr_un* a, socklen_t* len, int tag, int rank) {
  memset(a, 0, sizeof(*a));
  a->sun_family = AF_UNIX;
  const int n = snprintf(a->sun_path + 1, sizeof(a->sun_path) - 1, "b200rnn-%d-%d", tag, rank);  // leading NUL = abstract
  *len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
}

int send_fd(int sock, int tag, int dst, int my_rank, int fd) {
  sockaddr_un to;
  socklen_t tl;
  sock_name(&to, &tl, tag, dst);
  int32_t payload = my_rank;
  iovec iov{&payload, sizeof(payload)};
  char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msghdr msg{};
  msg.msg_name = &to; msg.msg_namelen = tl;
  msg.msg_iov = &iov; msg.msg_iovlen = 1;
  msg.msg_control = ctrl; msg.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET; c->cmsg_type = SCM_RIGHTS; c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  for (int tries = 0; tries < 200; ++tries) {
    if (sendmsg(sock, &msg, 0) >= 0) return 0;
    if (errno != ECONNREFUSED && errno != ENOENT && errno != EAGAIN) break;
    usleep(10000);
  }
  return -1;
}

int recv_fd(int sock, int* from_rank, int* fd, int timeout_ms) {
  pollfd pf{sock, POLLIN, 0};
  if (poll(&pf, 1, timeout_ms) <= 0) return -1;
  int32_t payload = -1;
  iovec iov{&payload, sizeof(payload)};
  char ctrl[CMSG_SPACE(sizeof(int))];
  msghdr msg{};
  msg.msg_iov = &iov; msg.msg_iovlen = 1;
  msg.msg_control = ctrl; msg.msg_controllen = sizeof(ctrl);
  if (recvmsg(sock, &msg, 0) < 0) return -1;
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (!c || c->cmsg_type != SCM_RIGHTS) return -1;
  memcpy(fd, CMSG_DATA(c), sizeof(int));
  *from_rank = payload;
  return 0;
}

}  // namespace

struct b200rnn_symm_vmm {
  int rank = 0, world = 1, device = 0;
  size_t bytes = 0;
  CUmemGenericAllocationHandle local = 0, mc = 0;
  std::vector<CUmemGenericAllocationHandle> peer_h;
  std::vector<CUdeviceptr> va;   // unicast mapping of every rank's region
  CUdeviceptr mc_va = 0;
  bool has_mc = false;
};

extern "C" int b200rnn_symm_vmm_destroy(b200rnn_symm_vmm* s);

// Collective over the bootstrap.  `tag` must be unique per node and job (the bootstrap port works).
// want_multicast = 0 skips step 3.
extern "C" int b200rnn_symm_vmm_create(struct b200rnn_boot* boot, int32_t rank, int32_t world, int32_t device, int64_t bytes,
                                       int32_t tag, int32_t want_multicast, b200rnn_symm_vmm** out) {
  B200_REQUIRE(boot && out && bytes > 0 && world >= 1 && world <= B200RNN_MAX_PEERS, B200RNN_EINVAL, "symm_vmm_create: bad argument");
  Api& A = api();
  B200_REQUIRE(A.ok, B200RNN_ECUDA, "symm_vmm_create: CUDA virtual-memory / multicast driver entry points not available");
  B200_CHECK_CUDA(cudaSetDevice(device));
  B200_CHECK_CUDA(cudaFree(0));   // make sure the primary context exists
  auto* s = new b200rnn_symm_vmm();
  s->rank = rank; s->world = world; s->device = device;
  s->peer_h.assign(world, 0);
  s->va.assign(world, 0);

  CUmulticastObjectProp mcp{};
  mcp.numDevices = (unsigned)world;
  mcp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 2u << 20;
  if (want_multicast && world > 1) {
    mcp.size = (size_t)bytes;
    size_t g = 0;
    if (A.McGetGranularity(&g, &mcp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && g > gran) gran = g;
  }
  const size_t size = ((size_t)bytes + gran - 1) / gran * gran;
  s->bytes = size;
  mcp.size = size;

  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;

#define VMM_FAIL_CLEANUP(expr)                                                                    \
  do {                                                                                            \
    CUresult _r = (expr);                                                                         \
    if (_r != CUDA_SUCCESS) {                                                                     \
      set_error("%s:%d: %s failed: %s (%d)", __FILE__, __LINE__, #expr, cu_err(_r), (int)_r);     \
      b200rnn_symm_vmm_destroy(s);                                                                \
      return B200RNN_ECUDA;                                                                       \
    }                                                                                             \
  } while (0)

  // 1. my region
  VMM_FAIL_CLEANUP(A.MemCreate(&s->local, size, &prop, 0));
  s->peer_h[rank] = s->local;
  VMM_FAIL_CLEANUP(A.MemAddressReserve(&s->va[rank], size, gran, 0, 0));
  VMM_FAIL_CLEANUP(A.MemMap(s->va[rank], size, 0, s->local, 0));
  VMM_FAIL_CLEANUP(A.MemSetAccess(s->va[rank], size, &acc, 1));
  if (cudaMemset((void*)s->va[rank], 0, size) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
    set_error("symm_vmm_create: memset of the new region failed");
    b200rnn_symm_vmm_destroy(s);
    return B200RNN_ECUDA;
  }

  if (world > 1) {
    // 2. exchange region fds
    int sock = ::socket(AF_UNIX, SOCK_DGRAM, 0);
    sockaddr_un me;
    socklen_t ml;
    sock_name(&me, &ml, tag, rank);
    if (sock < 0 || ::bind(sock, (sockaddr*)&me, ml) != 0) {
      set_error("symm_vmm_create: unix socket: %s", strerror(errno));
      if (sock >= 0) ::close(sock);
      b200rnn_symm_vmm_destroy(s);
      return B200RNN_ESYS;
    }
    int rc = b200rnn_boot_barrier(boot);   // every rank's socket is bound
    int my_fd = -1;
    if (rc == B200RNN_OK && A.MemExport(&my_fd, s->local, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) != CUDA_SUCCESS) {
      set_error("symm_vmm_create: cuMemExportToShareableHandle failed");
      rc = B200RNN_ECUDA;
    }
    for (int p = 0; rc == B200RNN_OK && p < world; ++p)
      if (p != rank && send_fd(sock, tag, p, rank, my_fd) != 0) { set_error("symm_vmm_create: fd send to rank %d failed: %s", p, strerror(errno)); rc = B200RNN_ESYS; }
    for (int i = 0; rc == B200RNN_OK && i < world - 1; ++i) {
      int from = -1, fd = -1;
      if (recv_fd(sock, &from, &fd, 30000) != 0 || from < 0 || from >= world) { set_error("symm_vmm_create: fd receive failed"); rc = B200RNN_ESYS; break; }
      CUresult r = A.MemImport(&s->peer_h[from], (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
      ::close(fd);
      if (r == CUDA_SUCCESS) r = A.MemAddressReserve(&s->va[from], size, gran, 0, 0);
      if (r == CUDA_SUCCESS) r = A.MemMap(s->va[from], size, 0, s->peer_h[from], 0);
      if (r == CUDA_SUCCESS) r = A.MemSetAccess(s->va[from], size, &acc, 1);
      if (r != CUDA_SUCCESS) { set_error("symm_vmm_create: mapping rank %d's region failed: %s", from, cu_err(r)); rc = B200RNN_ECUDA; }
    }
    if (my_fd >= 0) ::close(my_fd);
    if (rc == B200RNN_OK) rc = b200rnn_boot_barrier(boot);

    // 3. multicast object
    if (rc == B200RNN_OK && want_multicast) {
      int mc_fd = -1;
      if (rank == 0) {
        CUresult r = A.McCreate(&s->mc, &mcp);
        if (r == CUDA_SUCCESS) r = A.MemExport(&mc_fd, s->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
        if (r != CUDA_SUCCESS) { set_error("symm_vmm_create: cuMulticastCreate/export failed: %s", cu_err(r)); rc = B200RNN_ECUDA; }
        for (int p = 1; rc == B200RNN_OK && p < world; ++p)
          if (send_fd(sock, tag, p, 0, mc_fd) != 0) { set_error("symm_vmm_create: multicast fd send failed"); rc = B200RNN_ESYS; }
        if (mc_fd >= 0) ::close(mc_fd);
      } else {
        int from = -1;
        if (recv_fd(sock, &from, &mc_fd, 30000) != 0) { set_error("symm_vmm_create: multicast fd receive failed"); rc = B200RNN_ESYS; }
        else {
          CUresult r = A.MemImport(&s->mc, (void*)(uintptr_t)mc_fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
          ::close(mc_fd);
          if (r != CUDA_SUCCESS) { set_error("symm_vmm_create: multicast import failed: %s", cu_err(r)); rc = B200RNN_ECUDA; }
        }
      }
      // every rank reports success/failure so that nobody blocks in the collective steps below
      int32_t okflag = rc == B200RNN_OK ? 1 : 0;
      std::vector<int32_t> all(world);
      b200rnn_boot_allgather(boot, &okflag, 4, all.data());
      bool all_ok = true;
      for (int v : all) all_ok &= (v == 1);
      if (all_ok) {
        CUresult r = A.McAddDevice(s->mc, (CUdevice)device);
        okflag = r == CUDA_SUCCESS;
        if (!okflag) set_error("symm_vmm_create: cuMulticastAddDevice failed: %s", cu_err(r));
        b200rnn_boot_allgather(boot, &okflag, 4, all.data());
        for (int v : all) all_ok &= (v == 1);
      }
      if (all_ok) {
        CUresult r = A.McBindMem(s->mc, 0, s->local, 0, size, 0);
        if (r == CUDA_SUCCESS) r = A.MemAddressReserve(&s->mc_va, size, gran, 0, 0);
        if (r == CUDA_SUCCESS) r = A.MemMap(s->mc_va, size, 0, s->mc, 0);
        if (r == CUDA_SUCCESS) r = A.MemSetAccess(s->mc_va, size, &acc, 1);
        okflag = r == CUDA_SUCCESS;
        if (!okflag) set_error("symm_vmm_create: multicast bind/map failed: %s", cu_err(r));
        b200rnn_boot_allgather(boot, &okflag, 4, all.data());
        for (int v : all) all_ok &= (v == 1);
      }
      s->has_mc = all_ok;   // without multicast the unicast mappings are still fully usable
      if (!all_ok) s->mc_va = 0;
      rc = B200RNN_OK;
    }
    ::close(sock);
    if (rc != B200RNN_OK) { b200rnn_symm_vmm_destroy(s); return rc; }
  }
  *out = s;
  return B200RNN_OK;
}

extern "C" int b200rnn_symm_vmm_pointers(b200rnn_symm_vmm* s, void** out, void** mc_out) {
  B200_REQUIRE(s && out, B200RNN_EINVAL, "symm_vmm_pointers: null");
  for (int r = 0; r < s->world; ++r) out[r] = (void*)s->va[r];
  if (mc_out) *mc_out = s->has_mc ? (void*)s->mc_va : nullptr;
  return B200RNN_OK;
}

extern "C" int b200rnn_symm_vmm_destroy(b200rnn_symm_vmm* s) {
  if (!s) return B200RNN_OK;
  Api& A = api();
  if (A.ok) {
    if (s->mc_va) { A.MemUnmap(s->mc_va, s->bytes); A.MemAddressFree(s->mc_va, s->bytes); }
    for (int r = 0; r < s->world; ++r) {
      if (s->va[r]) { A.MemUnmap(s->va[r], s->bytes); A.MemAddressFree(s->va[r], s->bytes); }
      if (s->peer_h[r]) A.MemRelease(s->peer_h[r]);
    }
    if (s->mc) A.MemRelease(s->mc);
  }
  delete s;
  return B200RNN_OK;
}
