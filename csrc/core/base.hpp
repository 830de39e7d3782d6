// base.hpp — L1 vocabulary of the host runtime: keys, ranges, nodes, messages, the shared
// zero-copy array, the blocking queue and the actor base.
//
// Counterparts in the reference (behaviour, not code): base/magic.hpp:8-13 (Key, typedefs),
// base/third_party/range.h:14-29 (Range), base/node.hpp:8-22 (Node), base/message.hpp:15-73
// (Flag / Meta / Message), base/third_party/sarray.h:43-304 (SArray), base/threadsafe_queue.hpp
// :25-42 (ThreadsafeQueue), base/actor_model.hpp:13-34 (Actor), base/utils.hpp:15-53.
#pragma once

#include <atomic>
#include <cassert>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <deque>
#include <functional>
#include <iostream>
#include <memory>
#include <mutex>
#include <sstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <utility>
#include <vector>

namespace minips {

// ------------------------------------------------------------------ errors / logging
// The reference aborts through glog CHECK; a library that is also loaded into Python must
// throw instead so the host can report and unwind.
struct CheckError : std::runtime_error {
  using std::runtime_error::runtime_error;
};

#define MPS_STR2(x) #x
#define MPS_STR(x) MPS_STR2(x)
#define MINIPS_CHECK(cond)                                                                  \
  if (!(cond))                                                                              \
  ::minips::detail::CheckThrower(__FILE__ ":" MPS_STR(__LINE__) " CHECK(" #cond ") failed ")
#define MINIPS_CHECK_EQ(a, b) MINIPS_CHECK((a) == (b)) << "(" << (a) << " vs " << (b) << ") "
#define MINIPS_CHECK_LT(a, b) MINIPS_CHECK((a) < (b)) << "(" << (a) << " vs " << (b) << ") "
#define MINIPS_CHECK_LE(a, b) MINIPS_CHECK((a) <= (b)) << "(" << (a) << " vs " << (b) << ") "
#define MINIPS_CHECK_GE(a, b) MINIPS_CHECK((a) >= (b)) << "(" << (a) << " vs " << (b) << ") "

int LogLevel();            // MINIPS_VLOG env: 0 = warnings only, 1 = info, 2 = verbose
void SetLogLevel(int lvl);
#define MINIPS_LOG(lvl) \
  if (::minips::LogLevel() >= (lvl)) ::minips::detail::LogLine(lvl)

namespace detail {
struct CheckThrower {
  std::ostringstream os;
  explicit CheckThrower(const char* head) { os << head; }
  template <typename T>
  CheckThrower& operator<<(const T& v) {
    os << v;
    return *this;
  }
  [[noreturn]] ~CheckThrower() noexcept(false) { throw CheckError(os.str()); }
};
struct LogLine {
  std::ostringstream os;
  explicit LogLine(int lvl) { os << (lvl == 0 ? "[W] " : "[I] "); }
  template <typename T>
  LogLine& operator<<(const T& v) {
    os << v;
    return *this;
  }
  ~LogLine() {
    os << '\n';
    std::cerr << os.str();
  }
};
}  // namespace detail

// ------------------------------------------------------------------ keys / ranges / nodes
using Key = uint32_t;

struct Range {
  uint64_t begin_ = 0, end_ = 0;
  Range() = default;
  Range(uint64_t b, uint64_t e) : begin_(b), end_(e) {}
  uint64_t begin() const { return begin_; }
  uint64_t end() const { return end_; }
  uint64_t size() const { return end_ - begin_; }
  bool operator==(const Range& o) const { return begin_ == o.begin_ && end_ == o.end_; }
};

struct Node {
  uint32_t id = 0;
  std::string hostname;
  int port = 0;
  bool is_master = false;
  std::string DebugString() const {
    std::ostringstream os;
    os << "{ id: " << id << " hostname: " << hostname << " port: " << port << " }";
    return os.str();
  }
  bool operator==(const Node& o) const { return id == o.id && hostname == o.hostname && port == o.port; }
};

// ------------------------------------------------------------------ SArray
// Reference-counted array whose storage can be shared between differently typed views and
// sub-ranges without copying; what travels in Message::data.
template <typename V>
class SArray {
 public:
  SArray() = default;
  explicit SArray(size_t n, V init = V()) { resize(n, init); }
  SArray(const std::vector<V>& v) { CopyFrom(v.data(), v.size()); }  // NOLINT: implicit like the reference
  SArray(std::initializer_list<V> l) { CopyFrom(l.begin(), l.size()); }
  // adopt foreign memory; `deleter` runs when the last view dies
  SArray(V* data, size_t n, std::function<void(V*)> deleter) {
    store_ = std::shared_ptr<char>(reinterpret_cast<char*>(data),
                                   [deleter](char* p) { deleter(reinterpret_cast<V*>(p)); });
    data_ = data;
    size_ = capacity_ = n;
  }
  // zero-copy reinterpretation of another element type's storage
  template <typename W>
  SArray(const SArray<W>& o) {  // NOLINT
    *this = o;
  }
  template <typename W>
  SArray& operator=(const SArray<W>& o) {
    store_ = o.store();
    data_ = reinterpret_cast<V*>(const_cast<W*>(o.data()));
    size_ = o.size() * sizeof(W) / sizeof(V);
    capacity_ = o.capacity() * sizeof(W) / sizeof(V);
    return *this;
  }

  void CopyFrom(const V* src, size_t n) {
    alloc(n);
    if (n) std::memcpy(data_, src, n * sizeof(V));
    size_ = n;
  }
  void CopyFrom(const SArray<V>& o) {
    if (this == &o) return;
    CopyFrom(o.data(), o.size());
  }
  template <typename It>
  void CopyFrom(It first, It last) {
    size_t n = static_cast<size_t>(std::distance(first, last));
    alloc(n);
    size_ = n;
    V* p = data_;
    for (; first != last; ++first) *p++ = *first;
  }

  // Grows geometrically.  New elements take `init`; unlike the reference (sarray.h:189-208)
  // shrinking then growing never resurrects stale values.
  void resize(size_t n, V init = V()) {
    if (n > capacity_) {
      size_t cap = n + 5 > capacity_ * 2 ? n + 5 : capacity_ * 2;
      auto old_store = store_;
      V* old = data_;
      size_t old_size = size_;
      alloc(cap);
      if (old_size) std::memcpy(data_, old, old_size * sizeof(V));
      size_ = old_size;
      (void)old_store;
    }
    for (size_t i = size_; i < n; ++i) data_[i] = init;
    size_ = n;
  }
  void reserve(size_t n) {
    if (n > capacity_) {
      size_t keep = size_;
      resize(n);
      size_ = keep;
    }
  }
  void clear() { size_ = 0; }
  void push_back(const V& v) {
    if (size_ == capacity_) reserve(size_ * 2 + 5);
    data_[size_++] = v;
  }
  void pop_back() {
    if (size_) --size_;
  }
  void append(const SArray<V>& o) {
    if (o.empty()) return;
    size_t old = size_;
    resize(old + o.size());
    std::memcpy(data_ + old, o.data(), o.size() * sizeof(V));
  }
  // zero-copy sub-range [b, e)
  SArray<V> segment(size_t b, size_t e) const {
    MINIPS_CHECK(b <= e && e <= size_) << "segment " << b << ".." << e << " of " << size_;
    SArray<V> r;
    r.store_ = store_;
    r.data_ = data_ + b;
    r.size_ = r.capacity_ = e - b;
    return r;
  }

  bool empty() const { return size_ == 0; }
  size_t size() const { return size_; }
  size_t capacity() const { return capacity_; }
  V* data() { return data_; }
  const V* data() const { return data_; }
  V* begin() { return data_; }
  V* end() { return data_ + size_; }
  const V* begin() const { return data_; }
  const V* end() const { return data_ + size_; }
  V& operator[](size_t i) { return data_[i]; }
  const V& operator[](size_t i) const { return data_[i]; }
  V& front() { return data_[0]; }
  V& back() { return data_[size_ - 1]; }
  const V& back() const { return data_[size_ - 1]; }
  const std::shared_ptr<char>& store() const { return store_; }

  std::string DebugStr(size_t limit = 10) const {
    std::ostringstream os;
    os << "[" << size_ << "]{";
    for (size_t i = 0; i < size_ && i < limit; ++i) os << (i ? "," : "") << data_[i];
    if (size_ > limit) os << ",...";
    os << "}";
    return os.str();
  }

 private:
  void alloc(size_t n) {
    if (n == 0) {
      store_.reset();
      data_ = nullptr;
      capacity_ = 0;
      return;
    }
    store_ = std::shared_ptr<char>(new char[n * sizeof(V)], std::default_delete<char[]>());
    data_ = reinterpret_cast<V*>(store_.get());
    capacity_ = n;
  }
  std::shared_ptr<char> store_;
  V* data_ = nullptr;
  size_t size_ = 0, capacity_ = 0;
  template <typename W>
  friend class SArray;
};

// position range [lo, hi) of sorted `arr` whose values fall in `range` (sarray.h:324-330)
template <typename V>
Range FindRange(const SArray<V>& arr, V lower, V upper) {
  if (upper <= lower) return Range(0, 0);
  auto lb = std::lower_bound(arr.begin(), arr.end(), lower);
  auto ub = std::lower_bound(arr.begin(), arr.end(), upper);
  return Range(lb - arr.begin(), ub - arr.begin());
}

template <typename V>
std::vector<V> SArrayToVector(const SArray<V>& a) {
  return std::vector<V>(a.begin(), a.end());
}

using Keys = SArray<Key>;
using KVPairs = std::pair<SArray<Key>, SArray<double>>;
using SVMItem = std::pair<std::vector<std::pair<int, double>>, double>;  // (idx,val)*, label

// ------------------------------------------------------------------ messages
enum class Flag : char {
  kExit,
  kBarrier,
  kResetWorkerInModel,
  kClock,
  kAdd,
  kGet,
  kForceQuit,
  kCheckpoint,
  kHeartBeat,
  kQuitHeartBeat,
  kRollBack,
  kScale,
  kScaleRollback
};
const char* FlagName(Flag f);
inline bool IsControlFlag(Flag f) {
  // control flags address a NODE id in Meta::recver, data flags a THREAD id (mailbox.cpp:235-243)
  switch (f) {
    case Flag::kBarrier: case Flag::kForceQuit: case Flag::kHeartBeat: case Flag::kQuitHeartBeat:
    case Flag::kRollBack: case Flag::kScale: case Flag::kScaleRollback: case Flag::kExit:
      return true;
    default:
      return false;
  }
}

struct Meta {
  int sender = 0;
  int recver = 0;
  int model_id = 0;
  int failed_node_id = -1;
  Flag flag = Flag::kExit;
  std::string DebugString() const;
};

struct Message {
  Meta meta;
  std::vector<SArray<char>> data;
  template <typename V>
  void AddData(const SArray<V>& v) {
    data.push_back(SArray<char>(v));
  }
  std::string DebugString() const;
};

// ------------------------------------------------------------------ queue / actor
template <typename T>
class ThreadsafeQueue {
 public:
  void Push(T v) {
    // notify under the lock: a consumer that owns the queue (Engine::InitTable keeps one on its stack) may
    // destroy it as soon as it has popped, which it cannot do before this thread lets go of the mutex
    std::lock_guard<std::mutex> lk(mu_);
    q_.push_back(std::move(v));
    cv_.notify_all();
  }
  void WaitAndPop(T* out) {
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [this] { return !q_.empty(); });
    *out = std::move(q_.front());
    q_.pop_front();
  }
  // extensions over the reference: non-blocking and timed pops (used by tests / shutdown)
  bool TryPop(T* out) {
    std::lock_guard<std::mutex> lk(mu_);
    if (q_.empty()) return false;
    *out = std::move(q_.front());
    q_.pop_front();
    return true;
  }
  bool WaitAndPopFor(T* out, int timeout_ms) {
    std::unique_lock<std::mutex> lk(mu_);
    if (!cv_.wait_for(lk, std::chrono::milliseconds(timeout_ms), [this] { return !q_.empty(); })) return false;
    *out = std::move(q_.front());
    q_.pop_front();
    return true;
  }
  int Size() {
    std::lock_guard<std::mutex> lk(mu_);
    return static_cast<int>(q_.size());
  }

 private:
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<T> q_;
};

class Actor {
 public:
  explicit Actor(uint32_t id) : id_(id) {}
  virtual ~Actor() {
    if (th_.joinable()) th_.join();
  }
  void Start() {
    th_ = std::thread([this] { Main(); });
  }
  void Stop() {
    if (th_.joinable()) th_.join();
  }
  ThreadsafeQueue<Message>* GetWorkQueue() { return &work_queue_; }
  uint32_t GetId() const { return id_; }

 protected:
  virtual void Main() = 0;
  uint32_t id_;
  ThreadsafeQueue<Message> work_queue_;
  std::thread th_;
};

// Timestamped fault-tolerance phase marker (base/utils.hpp:24-53): phases 2..5 =
// failure detected / process restarted / process recovered / all nodes recovered.
std::string CheckFaultTolerance(int phase);

}  // namespace minips
