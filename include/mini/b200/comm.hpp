// comm.hpp — what the reference's drivers and `mesh::part::Part` use MPI for, between the rank processes of ONE node
// (one process per GPU, or several sharing a GPU): rank and size, a barrier, MPI_Allreduce(MIN | MAX | SUM) of doubles,
// MPI_Wtime, and an all-gather of small byte blocks that carries the CUDA IPC handles of the device-side ghost exchange
// (mcfd_peer_*) once at set-up.  The ghost data itself never comes through here: it moves between the GPUs
// (include/minicfd_b200.h, "peer" section).  There is no MPI on the target image; `nompi/mpi.h` maps the handful of MPI
// calls of demo/euler/sourceless.cpp:60-101, 187-215 onto this class so that the driver's text compiles unchanged.
//
// Transport: one POSIX shared-memory object per session.  Launch the ranks with
//     MCFD_SESSION=<unique name> MCFD_WORLD_SIZE=<n> MCFD_RANK=<r> ./driver ...
// (torchrun's RANK / WORLD_SIZE are accepted too).  Rank 0 removes the object when it finalises.
#ifndef MINI_B200_COMM_HPP_
#define MINI_B200_COMM_HPP_

#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>

namespace mini {
namespace b200 {

class Comm {
 public:
  static constexpr int kMaxRanks = 64, kBlockBytes = 1024;
  static constexpr double kTimeoutSeconds = 120.0;

  static Comm &World() {
    static Comm world;
    return world;
  }
  int rank() const { return rank_; }
  int size() const { return size_; }

  void Barrier() {
    if (size_ == 1) return;
    const int sense = local_sense_ ^= 1;
    if (shared_->arrived.fetch_add(1, std::memory_order_acq_rel) + 1 == size_) {
      shared_->arrived.store(0, std::memory_order_relaxed);
      shared_->sense.store(sense, std::memory_order_release);
    } else {
      const auto t0 = Clock::now();
      while (shared_->sense.load(std::memory_order_acquire) != sense) {
        sched_yield();
        if (Seconds(t0) > kTimeoutSeconds) throw std::runtime_error("mini::b200::Comm: a rank did not reach the barrier");
      }
    }
  }
  // every rank's block of `bytes` (<= kBlockBytes) bytes, in rank order
  void Allgather(const void *mine, size_t bytes, void *all) {
    if (bytes > (size_t)kBlockBytes) throw std::runtime_error("mini::b200::Comm::Allgather: block too large");
    if (size_ == 1) { std::memcpy(all, mine, bytes); return; }
    std::memcpy(shared_->block[rank_], mine, bytes);
    Barrier();
    for (int r = 0; r < size_; ++r) std::memcpy(static_cast<char *>(all) + r * bytes, shared_->block[r], bytes);
    Barrier();
  }
  enum Op { kMin, kMax, kSum };
  double Allreduce(double mine, Op op) {
    double all[kMaxRanks];
    Allgather(&mine, sizeof(double), all);
    double v = all[0];   // rank order on every rank: the same bits everywhere
    for (int r = 1; r < size_; ++r) v = op == kMin ? (all[r] < v ? all[r] : v) : op == kMax ? (all[r] > v ? all[r] : v) : v + all[r];
    return v;
  }
  double Wtime() const { return Seconds(start_); }
  void Finalize() {
    if (!shared_) return;
    Barrier();
    munmap(shared_, sizeof(Shared));
    shared_ = nullptr;
    if (rank_ == 0) shm_unlink(name_.c_str());
  }
  ~Comm() {
    if (shared_) { munmap(shared_, sizeof(Shared)); if (rank_ == 0) shm_unlink(name_.c_str()); }
  }

 private:
  using Clock = std::chrono::steady_clock;
  struct Shared {
    std::atomic<int> arrived, sense;
    alignas(64) char block[kMaxRanks][kBlockBytes];
  };
  static double Seconds(Clock::time_point t0) { return std::chrono::duration<double>(Clock::now() - t0).count(); }
  static int Env(const char *a, const char *b, int fallback) {
    const char *v = std::getenv(a);
    if (!v) v = std::getenv(b);
    return v ? std::atoi(v) : fallback;
  }
  Comm() : start_(Clock::now()) {
    rank_ = Env("MCFD_RANK", "RANK", 0);
    size_ = Env("MCFD_WORLD_SIZE", "WORLD_SIZE", 1);
    if (size_ < 1 || size_ > kMaxRanks || rank_ < 0 || rank_ >= size_) throw std::runtime_error("mini::b200::Comm: bad rank / world size");
    if (size_ == 1) return;
    const char *session = std::getenv("MCFD_SESSION");
    if (!session) throw std::runtime_error("mini::b200::Comm: MCFD_SESSION must name the run when there is more than one rank");
    name_ = std::string("/mcfd_") + session;
    const int fd = shm_open(name_.c_str(), O_CREAT | O_RDWR, 0600);
    if (fd < 0 || ftruncate(fd, sizeof(Shared)) != 0) throw std::runtime_error("mini::b200::Comm: shm_open failed for " + name_);
    void *p = mmap(nullptr, sizeof(Shared), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);   // a new object reads as zeros
    close(fd);
    if (p == MAP_FAILED) throw std::runtime_error("mini::b200::Comm: mmap failed");
    shared_ = static_cast<Shared *>(p);
    Barrier();
  }

  int rank_ = 0, size_ = 1, local_sense_ = 0;
  std::string name_;
  Shared *shared_ = nullptr;
  Clock::time_point start_;
};

}  // namespace b200
}  // namespace mini
#endif  // MINI_B200_COMM_HPP_
