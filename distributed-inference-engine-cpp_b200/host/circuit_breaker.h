// Per-worker circuit breaker of the gateway.
//
// Observable behaviour of the reference's CircuitBreaker (include/circuit_breaker.h, src/circuit_breaker.cpp:12-69):
//   CLOSED   : requests pass; a success clears the failure count; the 5th failure in a row opens the breaker
//   OPEN     : requests are refused until `timeout` (30 s) has passed since the last failure; the first request after
//              that is let through as a probe and the breaker becomes HALF_OPEN with its success count cleared
//   HALF_OPEN: requests pass; 2 successes close the breaker (failure count cleared), any failure re-opens it
// getFailureCount / getSuccessCount / getStateString feed the gateway's /stats (gateway.cpp:63-77); the success
// count is only cleared on OPEN -> HALF_OPEN, so it keeps its last value after closing, as in the reference.
//
// One mutex guards the whole state here (the reference mixes atomics and a mutex and reads some fields unlocked,
// SURVEY 5); the timeout is a steady_clock duration so tests can use milliseconds.
#pragma once
#include <chrono>
#include <mutex>
#include <string>

enum class CircuitState { CLOSED, OPEN, HALF_OPEN };

class CircuitBreaker {
 public:
  using Clock = std::chrono::steady_clock;
  explicit CircuitBreaker(int failure_threshold = 5, int success_threshold = 2,
                          Clock::duration timeout = std::chrono::seconds(30))
      : fail_limit_(failure_threshold), ok_limit_(success_threshold), timeout_(timeout), last_failure_(Clock::now()) {}

  bool allowRequest() {
    std::lock_guard<std::mutex> g(mu_);
    if (state_ != CircuitState::OPEN) return true;
    if (Clock::now() - last_failure_ < timeout_) return false;
    state_ = CircuitState::HALF_OPEN;       // this caller is the probe
    successes_ = 0;
    return true;
  }
  void recordSuccess() {
    std::lock_guard<std::mutex> g(mu_);
    if (state_ == CircuitState::HALF_OPEN) {
      if (++successes_ >= ok_limit_) { state_ = CircuitState::CLOSED; failures_ = 0; }
    } else {
      failures_ = 0;
    }
  }
  void recordFailure() {
    std::lock_guard<std::mutex> g(mu_);
    ++failures_;
    last_failure_ = Clock::now();
    if (state_ == CircuitState::HALF_OPEN || failures_ >= fail_limit_) state_ = CircuitState::OPEN;
  }
  CircuitState getState() const { std::lock_guard<std::mutex> g(mu_); return state_; }
  std::string getStateString() const {
    switch (getState()) {
      case CircuitState::CLOSED: return "CLOSED";
      case CircuitState::OPEN: return "OPEN";
      case CircuitState::HALF_OPEN: return "HALF_OPEN";
    }
    return "UNKNOWN";
  }
  int getFailureCount() const { std::lock_guard<std::mutex> g(mu_); return failures_; }
  int getSuccessCount() const { std::lock_guard<std::mutex> g(mu_); return successes_; }

 private:
  mutable std::mutex mu_;
  CircuitState state_ = CircuitState::CLOSED;
  int failures_ = 0, successes_ = 0;
  const int fail_limit_, ok_limit_;
  const Clock::duration timeout_;
  Clock::time_point last_failure_;
};
