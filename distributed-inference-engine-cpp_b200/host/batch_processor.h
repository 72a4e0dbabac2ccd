// Dynamic batching front of the engine: many request threads, one consumer thread that hands
// batches to a callback.
//
// Same public surface and default behaviour as the reference's BatchProcessor
// (reference: include/batch_processor.h:14-195): process() blocks the caller until its own
// response (or exception) is ready; the consumer wakes as soon as one request is queued and drains
// whatever is there, up to max_batch_size -- it does not wait to fill a batch (:111-121), so
// `timeout_batches` stays 0 and every processed batch counts as "full" (:160-164).  A callback
// exception reaches every request of that batch (:171-180); a callback that returns too few
// responses fails the tail with "No response for batched request" (:148-153).
//
// Differences in mechanism only: requests are moved, not copied, through the queue and into the
// batch (the reference copies each 602 KB payload at :98 and again at :142), and stop() fails the
// requests still queued instead of leaving their promises broken.
//
// Optional policy (off by default, because it changes the /health counters): setFillWait(true)
// makes the consumer wait up to `timeout` for the batch to fill once the first request arrived;
// a batch closed by that deadline is counted in timeout_batches.
// setInFlight(n) (before start(); default 1 = the reference's single consumer, :79) runs n consumer threads on
// the one queue, so that batch i+1 is packed and crosses PCIe while batch i computes (the engine's two transfer
// slots); batches then complete concurrently and the counters count each of them as before.
#pragma once
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <deque>
#include <functional>
#include <future>
#include <mutex>
#include <stdexcept>
#include <thread>
#include <vector>

template <typename Request, typename Response>
class BatchProcessor {
 public:
  using BatchCallback = std::function<std::vector<Response>(const std::vector<Request>&)>;

  struct Metrics {
    int64_t total_requests;
    int64_t total_batches;
    int64_t timeout_batches;
    int64_t full_batches;
    double avg_batch_size;
  };

  BatchProcessor(size_t max_batch_size, std::chrono::milliseconds timeout, BatchCallback callback)
      : max_batch_(max_batch_size), timeout_(timeout), callback_(std::move(callback)) {}
  ~BatchProcessor() { stop(); }

  void setFillWait(bool on) { fill_wait_ = on; }
  void setInFlight(int n) { in_flight_ = n < 1 ? 1 : n; }

  void start() {
    if (running_.exchange(true)) return;
    for (int i = 0; i < in_flight_; ++i) consumers_.emplace_back([this] { consume(); });
  }

  void stop() {
    running_ = false;
    cv_.notify_all();
    for (auto& t : consumers_)
      if (t.joinable()) t.join();
    consumers_.clear();
    std::deque<Pending> left;
    {
      std::lock_guard<std::mutex> g(mu_);
      left.swap(queue_);
    }
    for (auto& p : left)
      p.promise.set_exception(std::make_exception_ptr(std::runtime_error("BatchProcessor stopped")));
  }

  Response process(const Request& request) { return process(Request(request)); }

  // the request is moved into the queue (no copy of its payload)
  Response process(Request&& request) {
    Pending p{std::move(request), std::promise<Response>()};
    std::future<Response> fut = p.promise.get_future();
    {
      std::lock_guard<std::mutex> g(mu_);
      queue_.push_back(std::move(p));
      ++total_requests_;
    }
    cv_.notify_one();
    return fut.get();
  }

  Metrics getMetrics() const {
    std::lock_guard<std::mutex> g(metrics_mu_);
    return Metrics{total_requests_.load(), total_batches_, timeout_batches_, full_batches_, avg_batch_};
  }

 private:
  struct Pending {
    Request request;
    std::promise<Response> promise;
  };

  void consume() {
    std::vector<Pending> batch;
    batch.reserve(max_batch_);
    while (running_) {
      bool by_deadline = false;
      {
        std::unique_lock<std::mutex> lk(mu_);
        const bool woke = cv_.wait_for(lk, timeout_, [this] { return !queue_.empty() || !running_; });
        if (!running_) break;
        if (!woke) continue;                       // idle poll expired with nothing queued
        if (fill_wait_ && queue_.size() < max_batch_) {
          const auto deadline = std::chrono::steady_clock::now() + timeout_;
          by_deadline = !cv_.wait_until(lk, deadline, [this] { return queue_.size() >= max_batch_ || !running_; });
          if (!running_) break;
        }
        batch.clear();
        while (!queue_.empty() && batch.size() < max_batch_) {
          batch.push_back(std::move(queue_.front()));
          queue_.pop_front();
        }
      }
      if (!batch.empty()) run(batch, by_deadline);
    }
  }

  void run(std::vector<Pending>& batch, bool by_deadline) {
    std::vector<Request> reqs;
    reqs.reserve(batch.size());
    for (auto& p : batch) reqs.push_back(std::move(p.request));
    try {
      std::vector<Response> resp = callback_(reqs);
      for (size_t i = 0; i < batch.size(); ++i) {
        if (i < resp.size()) batch[i].promise.set_value(std::move(resp[i]));
        else batch[i].promise.set_exception(std::make_exception_ptr(std::runtime_error("No response for batched request")));
      }
      std::lock_guard<std::mutex> g(metrics_mu_);
      ++total_batches_;
      if (by_deadline) ++timeout_batches_; else ++full_batches_;
      avg_batch_ = (avg_batch_ * double(total_batches_ - 1) + double(batch.size())) / double(total_batches_);
    } catch (const std::exception&) {
      const std::exception_ptr ep = std::current_exception();
      for (auto& p : batch) {
        try { p.promise.set_exception(ep); } catch (...) { /* already satisfied */ }
      }
    }
  }

  size_t max_batch_;
  std::chrono::milliseconds timeout_;
  BatchCallback callback_;
  bool fill_wait_ = false;
  std::deque<Pending> queue_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::vector<std::thread> consumers_;
  int in_flight_ = 1;
  std::atomic<bool> running_{false};
  std::atomic<int64_t> total_requests_{0};
  mutable std::mutex metrics_mu_;
  int64_t total_batches_ = 0, timeout_batches_ = 0, full_batches_ = 0;
  double avg_batch_ = 0.0;
};
