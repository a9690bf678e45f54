// shm_comm.h -- the CPU transport of the distributed CPU solver: P processes of one node, one POSIX shared-memory
// segment, no library.
//
// Replaces what MPI is to the reference's stages 2 and 3: MPI_Init / Comm_rank / Comm_size (S2:464-468), the
// non-blocking halo exchange (MPI_Irecv / MPI_Isend / MPI_Waitall, S2:283-327), the scalar MPI_Allreduce (S2:396,
// S2:412, S2:435, S2:439) and MPI_Barrier (S2:483, S2:489).  The image has no MPI; torch.distributed's gloo remains
// available as the other transport (parallel/collective.py) and is the one that crosses node boundaries.
//
// Every rank owns one outbox line per direction (double-buffered by the parity of the exchange round, a sequence
// number per line): a halo exchange is "write my four lines, publish their sequence numbers, read my neighbours'
// lines when their numbers arrive" -- one copy per side, no mailbox handshake.  Sums are taken in rank order from
// per-rank slots, so every rank computes bit-identical results.
#pragma once
#include <cstddef>
#include <string>

namespace pb {

class ShmComm {
 public:
  // Collective over the `size` ranks of job `name` (any string unique to the run): rank 0 creates the segment, the
  // others attach.  line_doubles = capacity of one halo line.
  ShmComm(const std::string& name, int rank, int size, size_t line_doubles);
  ~ShmComm();
  ShmComm(const ShmComm&) = delete;
  ShmComm& operator=(const ShmComm&) = delete;

  int rank() const { return rank_; }
  int size() const { return size_; }
  void barrier();
  // in-place sum of n <= 8 doubles over all ranks, rank-ordered
  void allreduce_sum(double* v, int n);
  void allreduce_max(double* v, int n);
  // Halo exchange with up to four neighbours: send[d] (count[d] doubles) goes to rank nbr[d] (skipped when < 0) and
  // arrives there as its recv[d ^ 1]; recv[d] receives what nbr[d] sent towards us.
  void exchange(const int nbr[4], const double* const send[4], double* const recv[4], const int count[4]);

 private:
  struct Header;
  Header* hdr_ = nullptr;
  char* base_ = nullptr;
  size_t bytes_ = 0, line_bytes_ = 0;
  int rank_, size_;
  std::string name_;
  unsigned long long round_ = 0, red_round_ = 0;
  int sense_ = 0;
  double* line(int rank, int dir, int parity) const;
};

}  // namespace pb
