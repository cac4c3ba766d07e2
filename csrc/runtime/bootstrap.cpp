// B1 — process rendezvous, NVLink topology probe and symmetric-memory setup (host C++).
//
// Replaces the launch/rendezvous half of what the reference delegates to OpenMPI
// (mpirun + MPI_Init behind init_process_group('mpi'), src/motion/trainer/ddp.py:18), to
// horovodrun/hvd.init() (trainer/horovod.py:17) and to torch RPC's TCP rendezvous
// (param_server/__init__.py:41-42): ranks meet over a star of TCP connections to rank 0,
// exchange CUDA IPC handles of their symmetric regions, and map every peer's region so that
// kernels can address it directly over NVLink.  Single node (one process per GPU).
#include <arpa/inet.h>
#include <errno.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <string.h>
#include <sys/socket.h>
#include <time.h>
#include <unistd.h>

#include <vector>

#include "../common.cuh"

struct b200rnn_boot {
  int rank = 0, world = 1;
  int listen_fd = -1;
  std::vector<int> peers;  // rank 0: fd per rank (peers[0] unused); others: peers[0] = fd to rank 0
};

namespace {

using b200rnn::set_error;

long long now_ms() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (long long)ts.tv_sec * 1000 + ts.tv_nsec / 1000000;
}

bool send_all(int fd, const void* buf, size_t n) {
  const char* p = (const char*)buf;
  while (n > 0) {
    ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
    if (k < 0) { if (errno == EINTR) continue; return false; }
    p += k; n -= (size_t)k;
  }
  return true;
}

bool recv_all(int fd, void* buf, size_t n) {
  char* p = (char*)buf;
  while (n > 0) {
    ssize_t k = ::recv(fd, p, n, 0);
    if (k < 0) { if (errno == EINTR) continue; return false; }
    if (k == 0) return false;
    p += k; n -= (size_t)k;
  }
  return true;
}

void tune(int fd) {
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
}

}  // namespace

extern "C" int b200rnn_boot_listen(int32_t* port, int32_t* listen_fd) {
  B200_REQUIRE(port && listen_fd, B200RNN_EINVAL, "boot_listen: null argument");
  int fd = ::socket(AF_INET, SOCK_STREAM, 0);
  B200_REQUIRE(fd >= 0, B200RNN_ESYS, "boot_listen: socket(): %s", strerror(errno));
  int one = 1;
  setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in a{};
  a.sin_family = AF_INET;
  a.sin_addr.s_addr = htonl(INADDR_ANY);
  a.sin_port = htons((uint16_t)*port);
  if (::bind(fd, (sockaddr*)&a, sizeof(a)) != 0 || ::listen(fd, 64) != 0) {
    set_error("boot_listen: bind/listen on port %d: %s", *port, strerror(errno));
    ::close(fd);
    return B200RNN_ESYS;
  }
  socklen_t len = sizeof(a);
  getsockname(fd, (sockaddr*)&a, &len);
  *port = ntohs(a.sin_port);
  *listen_fd = fd;
  return B200RNN_OK;
}

extern "C" int b200rnn_boot_init(int32_t rank, int32_t world, const char* addr, int32_t port, int32_t listen_fd,
                                 int32_t timeout_ms, b200rnn_boot** out) {
  B200_REQUIRE(out && world >= 1 && world <= B200RNN_MAX_PEERS && rank >= 0 && rank < world, B200RNN_EINVAL,
               "boot_init: bad rank/world %d/%d", rank, world);
  auto* b = new b200rnn_boot();
  b->rank = rank; b->world = world;
  const long long deadline = now_ms() + (timeout_ms > 0 ? timeout_ms : 60000);
  if (world == 1) { *out = b; return B200RNN_OK; }
  if (rank == 0) {
    if (listen_fd < 0) { delete b; set_error("boot_init: rank 0 needs a listening socket"); return B200RNN_EINVAL; }
    b->listen_fd = listen_fd;
    b->peers.assign(world, -1);
    int have = 0;
    while (have < world - 1) {
      pollfd pf{listen_fd, POLLIN, 0};
      const long long left = deadline - now_ms();
      if (left <= 0 || ::poll(&pf, 1, (int)left) <= 0) {
        set_error("boot_init: timed out waiting for %d of %d ranks to connect", world - 1 - have, world - 1);
        b200rnn_boot_destroy(b);
        return B200RNN_ESYS;
      }
      int fd = ::accept(listen_fd, nullptr, nullptr);
      if (fd < 0) continue;
      tune(fd);
      int32_t r = -1;
      if (!recv_all(fd, &r, 4) || r <= 0 || r >= world || b->peers[r] != -1) { ::close(fd); continue; }
      b->peers[r] = fd;
      ++have;
    }
  } else {
    B200_REQUIRE(addr != nullptr, B200RNN_EINVAL, "boot_init: null address");
    addrinfo hints{}, *res = nullptr;
    hints.ai_family = AF_INET; hints.ai_socktype = SOCK_STREAM;
    char ports[16];
    snprintf(ports, sizeof(ports), "%d", port);
    if (getaddrinfo(addr, ports, &hints, &res) != 0 || !res) {
      delete b; set_error("boot_init: cannot resolve %s", addr); return B200RNN_ESYS;
    }
    int fd = -1;
    while (now_ms() < deadline) {
      fd = ::socket(AF_INET, SOCK_STREAM, 0);
      if (fd >= 0 && ::connect(fd, res->ai_addr, res->ai_addrlen) == 0) break;
      if (fd >= 0) ::close(fd);
      fd = -1;
      usleep(20000);
    }
    freeaddrinfo(res);
    if (fd < 0) { delete b; set_error("boot_init: rank %d could not reach %s:%d", rank, addr, port); return B200RNN_ESYS; }
    tune(fd);
    int32_t r = rank;
    if (!send_all(fd, &r, 4)) { ::close(fd); delete b; set_error("boot_init: handshake failed"); return B200RNN_ESYS; }
    b->peers.assign(1, fd);
  }
  *out = b;
  return B200RNN_OK;
}

extern "C" int b200rnn_boot_allgather(b200rnn_boot* b, const void* send, int32_t bytes, void* recv) {
  B200_REQUIRE(b && send && recv && bytes > 0, B200RNN_EINVAL, "boot_allgather: bad argument");
  char* out = (char*)recv;
  if (b->world == 1) { memcpy(out, send, bytes); return B200RNN_OK; }
  if (b->rank == 0) {
    memcpy(out, send, bytes);
    for (int r = 1; r < b->world; ++r)
      B200_REQUIRE(recv_all(b->peers[r], out + (size_t)r * bytes, bytes), B200RNN_ESYS, "boot_allgather: lost rank %d", r);
    for (int r = 1; r < b->world; ++r)
      B200_REQUIRE(send_all(b->peers[r], out, (size_t)b->world * bytes), B200RNN_ESYS, "boot_allgather: lost rank %d", r);
  } else {
    B200_REQUIRE(send_all(b->peers[0], send, bytes) && recv_all(b->peers[0], out, (size_t)b->world * bytes), B200RNN_ESYS,
                 "boot_allgather: lost rank 0");
  }
  return B200RNN_OK;
}

extern "C" int b200rnn_boot_barrier(b200rnn_boot* b) {
  B200_REQUIRE(b, B200RNN_EINVAL, "boot_barrier: null");
  int32_t x = b->rank;
  std::vector<int32_t> all(b->world);
  return b200rnn_boot_allgather(b, &x, 4, all.data());
}

extern "C" void b200rnn_boot_destroy(b200rnn_boot* b) {
  if (!b) return;
  for (int fd : b->peers) if (fd >= 0) ::close(fd);
  if (b->listen_fd >= 0) ::close(b->listen_fd);
  delete b;
}

// ---------------------------------------------------------------------------
// Topology
// ---------------------------------------------------------------------------
extern "C" int b200rnn_topology_probe(b200rnn_topology* t) {
  B200_REQUIRE(t, B200RNN_EINVAL, "topology_probe: null");
  memset(t, 0, sizeof(*t));
  int n = 0;
  B200_CHECK_CUDA(cudaGetDeviceCount(&n));
  if (n > B200RNN_MAX_PEERS) n = B200RNN_MAX_PEERS;
  t->n_devices = n;
  // CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED = 132 (driver API), fetched without linking libcuda
  typedef int (*cuDeviceGetAttribute_t)(int*, int, int);
  cuDeviceGetAttribute_t getattr = nullptr;
  {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuDeviceGetAttribute", &fn, cudaEnableDefault, &q) == cudaSuccess && fn != nullptr)
      getattr = (cuDeviceGetAttribute_t)fn;
    else
      (void)cudaGetLastError();
  }
  for (int i = 0; i < n; ++i) {
    cudaDeviceProp p;
    B200_CHECK_CUDA(cudaGetDeviceProperties(&p, i));
    snprintf(t->name[i], sizeof(t->name[i]), "%s", p.name);
    t->sm_count[i] = p.multiProcessorCount;
    int mc = 0;
    if (getattr) getattr(&mc, 132, i);
    t->multicast_supported[i] = mc;
    for (int j = 0; j < n; ++j) {
      if (i == j) { t->can_access[i][j] = 1; continue; }
      int v = 0;
      B200_CHECK_CUDA(cudaDeviceCanAccessPeer(&v, i, j));
      t->can_access[i][j] = v;
      if (v) {
        int pr = 0, at = 0;
        if (cudaDeviceGetP2PAttribute(&pr, cudaDevP2PAttrPerformanceRank, i, j) == cudaSuccess) t->perf_rank[i][j] = pr;
        if (cudaDeviceGetP2PAttribute(&at, cudaDevP2PAttrNativeAtomicSupported, i, j) == cudaSuccess) t->native_atomics[i][j] = at;
        (void)cudaGetLastError();
      }
    }
  }
  return B200RNN_OK;
}

// ---------------------------------------------------------------------------
// Symmetric memory over CUDA IPC
// ---------------------------------------------------------------------------
struct b200rnn_symm {
  int rank = 0, world = 1, device = 0;
  int64_t bytes = 0;
  void* local = nullptr;
  std::vector<void*> ptrs;
};

extern "C" int b200rnn_symm_create(b200rnn_boot* b, int32_t device, int64_t bytes, b200rnn_symm** out) {
  B200_REQUIRE(b && out && bytes > 0, B200RNN_EINVAL, "symm_create: bad argument");
  B200_CHECK_CUDA(cudaSetDevice(device));
  auto* s = new b200rnn_symm();
  s->rank = b->rank; s->world = b->world; s->device = device; s->bytes = bytes;
  s->ptrs.assign(b->world, nullptr);
  cudaError_t e = cudaMalloc(&s->local, (size_t)bytes);
  if (e != cudaSuccess) { delete s; set_error("symm_create: cudaMalloc(%lld): %s", (long long)bytes, cudaGetErrorString(e)); return B200RNN_ECUDA; }
  e = cudaMemset(s->local, 0, (size_t)bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { cudaFree(s->local); delete s; set_error("symm_create: memset: %s", cudaGetErrorString(e)); return B200RNN_ECUDA; }
  s->ptrs[s->rank] = s->local;
  if (b->world > 1) {
    struct Msg { cudaIpcMemHandle_t h; int32_t device; int32_t pad; };
    Msg mine{};
    e = cudaIpcGetMemHandle(&mine.h, s->local);
    if (e != cudaSuccess) { cudaFree(s->local); delete s; set_error("symm_create: cudaIpcGetMemHandle: %s", cudaGetErrorString(e)); return B200RNN_ECUDA; }
    mine.device = device;
    std::vector<Msg> all(b->world);
    int rc = b200rnn_boot_allgather(b, &mine, (int32_t)sizeof(Msg), all.data());
    if (rc != B200RNN_OK) { cudaFree(s->local); delete s; return rc; }
    for (int r = 0; r < b->world; ++r) {
      if (r == s->rank) continue;
      if (all[r].device != device) {
        int can = 0;
        cudaDeviceCanAccessPeer(&can, device, all[r].device);
        if (!can) {
          set_error("symm_create: GPU %d cannot access peer GPU %d (no NVLink/P2P path)", device, all[r].device);
          b200rnn_symm_destroy(s);
          return B200RNN_ECUDA;
        }
      }
      e = cudaIpcOpenMemHandle(&s->ptrs[r], all[r].h, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) {
        set_error("symm_create: cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
        s->ptrs[r] = nullptr;
        b200rnn_symm_destroy(s);
        return B200RNN_ECUDA;
      }
    }
    rc = b200rnn_boot_barrier(b);  // nobody proceeds (or frees) before everyone has mapped everyone
    if (rc != B200RNN_OK) { b200rnn_symm_destroy(s); return rc; }
  }
  *out = s;
  return B200RNN_OK;
}

extern "C" int b200rnn_symm_pointers(b200rnn_symm* s, void** out) {
  B200_REQUIRE(s && out, B200RNN_EINVAL, "symm_pointers: null");
  for (int r = 0; r < s->world; ++r) out[r] = s->ptrs[r];
  return B200RNN_OK;
}

extern "C" int b200rnn_symm_destroy(b200rnn_symm* s) {
  if (!s) return B200RNN_OK;
  for (int r = 0; r < s->world; ++r)
    if (r != s->rank && s->ptrs[r] != nullptr) cudaIpcCloseMemHandle(s->ptrs[r]);
  if (s->local) cudaFree(s->local);
  delete s;
  return B200RNN_OK;
}
