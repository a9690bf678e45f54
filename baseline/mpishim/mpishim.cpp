// mpishim.cpp -- see mpi.h.  One process per rank, one node, POSIX shared memory.
#include "mpi.h"

#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#ifdef PB_SHIM_WITH_CUDA
#include <cuda_runtime.h>
#endif

namespace {

constexpr int kMaxRanks = 32;
constexpr int kTags = 8;
constexpr size_t kSlotBytes = 1u << 19;  // 512 KiB payload per (src, dst, tag) mailbox
constexpr int kRedMax = 16;              // elements per collective
constexpr uint64_t kMagic = 0x70625f6d70697368ull;

struct Mailbox {
  std::atomic<int> full;
  int bytes;
  char pad[56];
  char data[kSlotBytes];
};

struct Shared {
  std::atomic<uint64_t> magic;
  int size;
  std::atomic<int> attached;
  // sense-reversing barrier
  std::atomic<int> bar_count;
  std::atomic<int> bar_sense;
  // collectives
  double red[kMaxRanks][kRedMax];
  Mailbox box[1];  // [size][size][kTags], allocated to measure
};

Shared* g_sh = nullptr;
size_t g_bytes = 0;
int g_rank = 0, g_size = 1;
int g_local_sense = 0;
std::string g_name;

int env_int(const char* a, const char* b, int dflt) {
  const char* v = getenv(a);
  if (!v && b) v = getenv(b);
  return v ? atoi(v) : dflt;
}

Mailbox& box(int src, int dst, int tag) {
  return g_sh->box[((size_t)src * g_size + dst) * kTags + (size_t)(tag & (kTags - 1))];
}

size_t elem_size(MPI_Datatype t) { return t == MPI_INT ? sizeof(int) : sizeof(double); }

void relax() { sched_yield(); }

void barrier() {
  if (g_size == 1) return;
  g_local_sense ^= 1;
  if (g_sh->bar_count.fetch_add(1, std::memory_order_acq_rel) + 1 == g_size) {
    g_sh->bar_count.store(0, std::memory_order_relaxed);
    g_sh->bar_sense.store(g_local_sense, std::memory_order_release);
  } else {
    int spins = 0;
    while (g_sh->bar_sense.load(std::memory_order_acquire) != g_local_sense)
      if (++spins > 2000) relax();
  }
}

bool try_send(const void* buf, int bytes, int dest, int tag) {
  Mailbox& m = box(g_rank, dest, tag);
  if (m.full.load(std::memory_order_acquire)) return false;
  if ((size_t)bytes > kSlotBytes) {
    fprintf(stderr, "mpishim: message of %d bytes exceeds the mailbox\n", bytes);
    abort();
  }
  memcpy(m.data, buf, bytes);
  m.bytes = bytes;
  m.full.store(1, std::memory_order_release);
  return true;
}

bool try_recv(void* buf, int bytes, int src, int tag) {
  Mailbox& m = box(src, g_rank, tag);
  if (!m.full.load(std::memory_order_acquire)) return false;
  memcpy(buf, m.data, bytes < m.bytes ? bytes : m.bytes);
  m.full.store(0, std::memory_order_release);
  return true;
}

}  // namespace

struct pb_mpi_request {
  bool is_send, done;
  void* buf;
  int bytes, peer, tag;
};

extern "C" {

int MPI_Init(int*, char***) {
  g_rank = env_int("PB_MPI_RANK", "RANK", 0);
  g_size = env_int("PB_MPI_SIZE", "WORLD_SIZE", 1);
  if (g_size > kMaxRanks) {
    fprintf(stderr, "mpishim: at most %d ranks\n", kMaxRanks);
    exit(1);
  }
  const char* job = getenv("PB_MPI_JOB");
  const char* port = getenv("MASTER_PORT");
  g_name = std::string("/pb_mpishim_") + (job ? job : (port ? port : "default"));
  g_bytes = sizeof(Shared) + sizeof(Mailbox) * ((size_t)g_size * g_size * kTags);
  if (g_size > 1) {
    int fd = -1;
    if (g_rank == 0) {
      shm_unlink(g_name.c_str());
      fd = shm_open(g_name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
      if (fd < 0 || ftruncate(fd, (off_t)g_bytes) != 0) {
        perror("mpishim: shm_open/ftruncate");
        exit(1);
      }
    } else {
      for (int tries = 0; fd < 0; ++tries) {
        fd = shm_open(g_name.c_str(), O_RDWR, 0600);
        struct stat st;
        if (fd >= 0 && (fstat(fd, &st) != 0 || (size_t)st.st_size < g_bytes)) {
          close(fd);
          fd = -1;
        }
        if (fd < 0) {
          if (tries > 600000) {
            fprintf(stderr, "mpishim: rank %d could not attach to %s\n", g_rank, g_name.c_str());
            exit(1);
          }
          usleep(100);
        }
      }
    }
    g_sh = (Shared*)mmap(nullptr, g_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (g_sh == MAP_FAILED) {
      perror("mpishim: mmap");
      exit(1);
    }
    if (g_rank == 0) {
      g_sh->size = g_size;
      g_sh->attached.store(0);
      g_sh->bar_count.store(0);
      g_sh->bar_sense.store(0);
      g_sh->magic.store(kMagic, std::memory_order_release);
    } else {
      while (g_sh->magic.load(std::memory_order_acquire) != kMagic) usleep(50);
    }
    g_sh->attached.fetch_add(1);
    while (g_sh->attached.load() < g_size) usleep(50);
  }
#ifdef PB_SHIM_WITH_CUDA
  cudaFree(0);  // create the CUDA context outside the program's own timers
#endif
  barrier();
  return MPI_SUCCESS;
}

int MPI_Finalize(void) {
  if (g_size > 1) {
    barrier();
    if (g_rank == 0) shm_unlink(g_name.c_str());
    munmap(g_sh, g_bytes);
    g_sh = nullptr;
  }
  return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm, int* rank) {
  *rank = g_rank;
  return MPI_SUCCESS;
}
int MPI_Comm_size(MPI_Comm, int* size) {
  *size = g_size;
  return MPI_SUCCESS;
}
double MPI_Wtime(void) {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}
int MPI_Barrier(MPI_Comm) {
  barrier();
  return MPI_SUCCESS;
}

int MPI_Send(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm) {
  if (dest == MPI_PROC_NULL) return MPI_SUCCESS;
  int spins = 0;
  while (!try_send(buf, (int)(count * elem_size(type)), dest, tag))
    if (++spins > 2000) relax();
  return MPI_SUCCESS;
}
int MPI_Recv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm, MPI_Status*) {
  if (source == MPI_PROC_NULL) return MPI_SUCCESS;
  int spins = 0;
  while (!try_recv(buf, (int)(count * elem_size(type)), source, tag))
    if (++spins > 2000) relax();
  return MPI_SUCCESS;
}
int MPI_Sendrecv(const void* sendbuf, int sendcount, MPI_Datatype sendtype, int dest, int sendtag, void* recvbuf,
                 int recvcount, MPI_Datatype recvtype, int source, int recvtag, MPI_Comm comm, MPI_Status* st) {
  MPI_Send(sendbuf, sendcount, sendtype, dest, sendtag, comm);
  MPI_Recv(recvbuf, recvcount, recvtype, source, recvtag, comm, st);
  return MPI_SUCCESS;
}

int MPI_Isend(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm, MPI_Request* req) {
  pb_mpi_request* r = new pb_mpi_request{true, dest == MPI_PROC_NULL, const_cast<void*>(buf),
                                         (int)(count * elem_size(type)), dest, tag};
  if (!r->done) r->done = try_send(r->buf, r->bytes, r->peer, r->tag);
  *req = r;
  return MPI_SUCCESS;
}
int MPI_Irecv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm, MPI_Request* req) {
  *req = new pb_mpi_request{false, source == MPI_PROC_NULL, buf, (int)(count * elem_size(type)), source, tag};
  return MPI_SUCCESS;
}
int MPI_Waitall(int count, MPI_Request* reqs, MPI_Status*) {
  int pending = 0, spins = 0;
  do {
    pending = 0;
    for (int i = 0; i < count; ++i) {
      pb_mpi_request* r = reqs[i];
      if (!r || r->done) continue;
      r->done = r->is_send ? try_send(r->buf, r->bytes, r->peer, r->tag) : try_recv(r->buf, r->bytes, r->peer, r->tag);
      if (!r->done) ++pending;
    }
    if (pending && ++spins > 2000) relax();
  } while (pending);
  for (int i = 0; i < count; ++i) {
    delete reqs[i];
    reqs[i] = MPI_REQUEST_NULL;
  }
  return MPI_SUCCESS;
}

static int reduce_impl(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op, int root) {
  if (type != MPI_DOUBLE || count > kRedMax) {
    fprintf(stderr, "mpishim: unsupported reduction (type %d, count %d)\n", type, count);
    abort();
  }
  const double* in = (const double*)sendbuf;
  double* out = (double*)recvbuf;
  if (g_size == 1) {
    for (int i = 0; i < count; ++i) out[i] = in[i];
    return MPI_SUCCESS;
  }
  for (int i = 0; i < count; ++i) g_sh->red[g_rank][i] = in[i];
  barrier();
  if (root < 0 || root == g_rank) {
    for (int i = 0; i < count; ++i) {
      double acc = g_sh->red[0][i];
      for (int r = 1; r < g_size; ++r) {
        const double v = g_sh->red[r][i];
        acc = op == MPI_SUM ? acc + v : (op == MPI_MAX ? (v > acc ? v : acc) : (v < acc ? v : acc));
      }
      out[i] = acc;
    }
  }
  barrier();
  return MPI_SUCCESS;
}

int MPI_Allreduce(const void* s, void* r, int count, MPI_Datatype type, MPI_Op op, MPI_Comm) {
  return reduce_impl(s, r, count, type, op, -1);
}
int MPI_Reduce(const void* s, void* r, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm) {
  return reduce_impl(s, r, count, type, op, root);
}
int MPI_Abort(MPI_Comm, int code) {
  _exit(code);
  return 0;
}

}  // extern "C"
