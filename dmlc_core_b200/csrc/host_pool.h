/*
 * HostPool: a handful of persistent host threads with a fork-join Run().  Shared by the two
 * places of this package that do per-byte work on the host: the C-ABI library, which rebuilds
 * the column ids of wide CSV rows from offset[] instead of bringing them back over PCIe
 * (b200p_next), and TextSplit, which fans one chunk's pread out over the pool.
 *
 * Run(parts, fn) calls fn(0..parts-1) once each, the calling thread included, and returns when
 * all are done.  One Run at a time per pool.
 */
#ifndef DMLC_CORE_B200_HOST_POOL_H_
#define DMLC_CORE_B200_HOST_POOL_H_

#include <condition_variable>
#include <cstddef>
#include <cstdint>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace b200p {

class HostPool {
 public:
  explicit HostPool(int threads) : fn_(NULL), parts_(0), next_(0), left_(0), epoch_(0), stop_(false) {
    for (int i = 1; i < threads; ++i) workers_.emplace_back([this] { Work(); });
  }
  ~HostPool() {
    {
      std::lock_guard<std::mutex> g(m_);
      stop_ = true;
    }
    wake_.notify_all();
    for (std::thread &t : workers_) t.join();
  }
  int threads() const { return static_cast<int>(workers_.size()) + 1; }
  void Run(int parts, const std::function<void(int)> &fn) {
    std::unique_lock<std::mutex> g(m_);
    fn_ = &fn;
    parts_ = parts;
    next_ = 0;
    left_ = parts;
    ++epoch_;
    wake_.notify_all();
    Drain(&g);
    idle_.wait(g, [this] { return left_ == 0; });
    fn_ = NULL;
  }

 private:
  void Drain(std::unique_lock<std::mutex> *g) { /* called with the lock held */
    while (fn_ != NULL && next_ < parts_) {
      const int mine = next_++;
      const std::function<void(int)> *fn = fn_;
      g->unlock();
      (*fn)(mine);
      g->lock();
      if (--left_ == 0) idle_.notify_all();
    }
  }
  void Work() {
    std::unique_lock<std::mutex> g(m_);
    uint64_t seen = 0;
    for (;;) {
      wake_.wait(g, [&] { return stop_ || epoch_ != seen; });
      if (stop_) return;
      seen = epoch_;
      Drain(&g);
    }
  }
  std::vector<std::thread> workers_;
  std::mutex m_;
  std::condition_variable wake_, idle_;
  const std::function<void(int)> *fn_;
  int parts_, next_, left_;
  uint64_t epoch_;
  bool stop_;
};

}  // namespace b200p
#endif  // DMLC_CORE_B200_HOST_POOL_H_
