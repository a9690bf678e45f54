// shm_comm.cpp -- see shm_comm.h.
#include "shm_comm.h"

#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <cstdint>
#include <cstring>
#include <stdexcept>

namespace pb {

namespace {
constexpr int kMaxShmRanks = 64;
constexpr int kRedSlots = 8;
constexpr uint64_t kReady = 0x706273686d636f6dull;

double mono_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}
// spin, then yield: ranks may share cores (the box has fewer cores than a large run has ranks)
template <class Pred>
void wait_until(Pred&& ready, const char* what) {
  const double t0 = mono_s();
  for (unsigned spins = 0; !ready(); ++spins) {
    if (spins > 200) sched_yield();
    if ((spins & 0xfffff) == 0xfffff && mono_s() - t0 > 120.0) throw std::runtime_error(std::string("ShmComm: timed out in ") + what);
  }
}
}  // namespace

struct ShmComm::Header {
  std::atomic<uint64_t> ready;
  std::atomic<int> attached;
  std::atomic<int> bar_count;
  std::atomic<int> bar_sense;
  char pad0[40];
  struct alignas(64) RedSlot {
    std::atomic<unsigned long long> seq;
    double v[kRedSlots];
  } red[2][kMaxShmRanks];
  struct alignas(64) LineSeq {
    std::atomic<unsigned long long> seq;
  } line_seq[kMaxShmRanks][4][2];
};

ShmComm::ShmComm(const std::string& name, int rank, int size, size_t line_doubles) : rank_(rank), size_(size), name_("/pbshm_" + name) {
  if (size < 1 || size > kMaxShmRanks || rank < 0 || rank >= size) throw std::runtime_error("ShmComm: bad rank / size");
  line_bytes_ = ((line_doubles * sizeof(double) + 63) / 64) * 64;
  const size_t hdr_bytes = ((sizeof(Header) + 4095) / 4096) * 4096;
  bytes_ = hdr_bytes + (size_t)size * 4 * 2 * line_bytes_;
  int fd = -1;
  if (rank == 0) {
    shm_unlink(name_.c_str());
    fd = shm_open(name_.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0 || ftruncate(fd, (off_t)bytes_) != 0) throw std::runtime_error("ShmComm: cannot create " + name_);
  } else {
    const double t0 = mono_s();
    while (true) {
      fd = shm_open(name_.c_str(), O_RDWR, 0600);
      struct stat st;
      if (fd >= 0 && fstat(fd, &st) == 0 && (size_t)st.st_size >= bytes_) break;
      if (fd >= 0) close(fd);
      if (mono_s() - t0 > 120.0) throw std::runtime_error("ShmComm: rank 0 never created " + name_);
      usleep(1000);
    }
  }
  base_ = static_cast<char*>(mmap(nullptr, bytes_, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0));
  close(fd);
  if (base_ == MAP_FAILED) throw std::runtime_error("ShmComm: mmap failed");
  hdr_ = reinterpret_cast<Header*>(base_);
  base_ += hdr_bytes;
  if (rank == 0) {
    std::memset(static_cast<void*>(hdr_), 0, hdr_bytes);  // a fresh segment is zero anyway; be explicit
    hdr_->ready.store(kReady, std::memory_order_release);
  } else {
    wait_until([&] { return hdr_->ready.load(std::memory_order_acquire) == kReady; }, "attach");
  }
  hdr_->attached.fetch_add(1, std::memory_order_acq_rel);
  wait_until([&] { return hdr_->attached.load(std::memory_order_acquire) >= size_; }, "attach (all ranks)");
  barrier();
  if (rank == 0) shm_unlink(name_.c_str());  // everyone is mapped: the name can go, the memory lives until the last unmap
}

ShmComm::~ShmComm() {
  if (hdr_) {
    munmap(static_cast<void*>(hdr_), bytes_);  // the mapping starts at the header and is bytes_ long
  }
}

double* ShmComm::line(int rank, int dir, int parity) const {
  return reinterpret_cast<double*>(base_ + (((size_t)rank * 4 + dir) * 2 + parity) * line_bytes_);
}

void ShmComm::barrier() {
  sense_ ^= 1;
  if (hdr_->bar_count.fetch_add(1, std::memory_order_acq_rel) + 1 == size_) {
    hdr_->bar_count.store(0, std::memory_order_relaxed);
    hdr_->bar_sense.store(sense_, std::memory_order_release);
  } else {
    wait_until([&] { return hdr_->bar_sense.load(std::memory_order_acquire) == sense_; }, "barrier");
  }
}

void ShmComm::allreduce_sum(double* v, int n) {
  if (n > kRedSlots) throw std::runtime_error("ShmComm::allreduce: at most 8 values");
  if (size_ == 1) return;
  const unsigned long long seq = ++red_round_;
  auto* mine = &hdr_->red[seq & 1][rank_];
  for (int k = 0; k < n; ++k) mine->v[k] = v[k];
  mine->seq.store(seq, std::memory_order_release);
  double acc[kRedSlots] = {0};
  for (int r = 0; r < size_; ++r) {  // rank order: the same bits on every rank
    auto* s = &hdr_->red[seq & 1][r];
    wait_until([&] { return s->seq.load(std::memory_order_acquire) == seq; }, "allreduce");
    for (int k = 0; k < n; ++k) acc[k] += s->v[k];
  }
  for (int k = 0; k < n; ++k) v[k] = acc[k];
}

void ShmComm::allreduce_max(double* v, int n) {
  if (n > kRedSlots) throw std::runtime_error("ShmComm::allreduce: at most 8 values");
  if (size_ == 1) return;
  const unsigned long long seq = ++red_round_;
  auto* mine = &hdr_->red[seq & 1][rank_];
  for (int k = 0; k < n; ++k) mine->v[k] = v[k];
  mine->seq.store(seq, std::memory_order_release);
  for (int r = 0; r < size_; ++r) {
    auto* s = &hdr_->red[seq & 1][r];
    wait_until([&] { return s->seq.load(std::memory_order_acquire) == seq; }, "allreduce");
    for (int k = 0; k < n; ++k) v[k] = s->v[k] > v[k] ? s->v[k] : v[k];
  }
}

void ShmComm::exchange(const int nbr[4], const double* const send[4], double* const recv[4], const int count[4]) {
  const unsigned long long seq = ++round_;
  const int par = (int)(seq & 1);
  // A line of parity `par` was last used in round seq - 2; its reader finished that round before it entered round
  // seq - 1, whose lines we have already read -- so it is free.
  for (int d = 0; d < 4; ++d) {
    if (nbr[d] < 0) continue;
    if ((size_t)count[d] * sizeof(double) > line_bytes_) throw std::runtime_error("ShmComm::exchange: line too long");
    std::memcpy(line(rank_, d, par), send[d], (size_t)count[d] * sizeof(double));
    hdr_->line_seq[rank_][d][par].seq.store(seq, std::memory_order_release);
  }
  for (int d = 0; d < 4; ++d) {
    if (nbr[d] < 0) continue;
    const int from = nbr[d], their_dir = d ^ 1;  // what they sent towards us left through their opposite direction
    auto& s = hdr_->line_seq[from][their_dir][par].seq;
    wait_until([&] { return s.load(std::memory_order_acquire) == seq; }, "halo exchange");
    std::memcpy(recv[d], line(from, their_dir, par), (size_t)count[d] * sizeof(double));
  }
}

}  // namespace pb
