// lib.hpp — application-level helpers.
//
//   Parser<Sample>::parse_libsvm  lib/parser.hpp:17-47  (label then (index-1, value) pairs)
//   LabeledSample                  lib/labeled_sample.hpp:8-13
//   DataLoader                     lib/abstract_data_loader.hpp:27-51 (block-assigned parallel load)
//   AsyncReadBuffer / AsyncDataLoader  lib/abstract_aync_data_loader.hpp:35-92 — an empty skeleton in
//                                  the reference; implemented here as a bounded producer/consumer
//                                  prefetcher of parsed batches
//   BatchDataSampler<T>            lib/batch_data_sampler.cpp:24-86 (consecutive wrap-around batches +
//                                  the sorted unique key set of each batch)
//   SVMDumper                      lib/svm_dumper.hpp:22-198 (worker data / config / scale file dumps)
#pragma once

#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <deque>
#include <functional>
#include <mutex>
#include <random>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "core/base.hpp"
#include "core/io.hpp"

namespace minips {
namespace lib {

template <typename Feature, typename Label>
struct LabeledSample {
  Feature x_;
  Label y_;
  LabeledSample() = default;
  LabeledSample(Feature x, Label y) : x_(std::move(x)), y_(y) {}
};

// SVMItem = (vector<(index, value)>, label)
struct Parser {
  // "label idx:val idx:val ..." with 1-based idx in the text, 0-based in memory.
  // Returns false for blank / comment lines.
  static bool parse_libsvm(const std::string& line, SVMItem* item);
  static SVMItem parse_libsvm(const std::string& line) {
    SVMItem it;
    parse_libsvm(line, &it);
    return it;
  }
  // MNIST is a stub in the reference too (lib/parser.hpp:49-52)
  static SVMItem parse_mnist(const std::string&) { return SVMItem(); }
};

// Parallel load of this node's share of `config.url` through the block assigner.
template <typename Sample, typename DataStore>
struct DataLoader {
  static void load(const DataManager::Config& config, const Node& my_node, const std::vector<Node>& nodes,
                   const std::function<bool(const std::string&, Sample*)>& parse, DataStore* datastore) {
    DataManager manager(my_node, nodes, config);
    manager.Start();
    std::mutex mu;
    manager.Run([&](const std::string& line, int) {
      Sample s;
      if (!parse(line, &s)) return;
      std::lock_guard<std::mutex> lk(mu);
      datastore->push_back(std::move(s));
    });
    manager.Stop();
  }
};

// Bounded queue of ready batches filled by a background thread.
template <typename Sample>
class AsyncReadBuffer {
 public:
  using Batch = std::vector<Sample>;
  AsyncReadBuffer() = default;
  ~AsyncReadBuffer() { stop(); }

  // `produce(batch*)` fills the next batch and returns false at end of data.
  void init(std::function<bool(Batch*)> produce, int batch_size, int batch_num) {
    produce_ = std::move(produce);
    batch_size_ = batch_size;
    capacity_ = std::max(1, batch_num);
    eof_ = false;
    running_ = true;
    th_ = std::thread([this] { main(); });
  }
  // blocks until a batch is ready; false when the producer hit end of data and the queue drained
  bool get_batch(Batch* out) {
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [this] { return !q_.empty() || eof_ || !running_; });
    if (q_.empty()) return false;
    *out = std::move(q_.front());
    q_.pop_front();
    cv_.notify_all();
    return true;
  }
  int ask() {  // number of ready batches
    std::lock_guard<std::mutex> lk(mu_);
    return static_cast<int>(q_.size());
  }
  bool end_of_file() {
    std::lock_guard<std::mutex> lk(mu_);
    return eof_ && q_.empty();
  }
  void stop() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      running_ = false;
    }
    cv_.notify_all();
    if (th_.joinable()) th_.join();
  }

 private:
  void main() {
    while (true) {
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [this] { return !running_ || static_cast<int>(q_.size()) < capacity_; });
        if (!running_) return;
      }
      Batch b;
      b.reserve(batch_size_);
      const bool more = produce_(&b);
      {
        std::lock_guard<std::mutex> lk(mu_);
        if (!b.empty()) q_.push_back(std::move(b));
        if (!more) eof_ = true;
      }
      cv_.notify_all();
      if (!more) return;
    }
  }
  std::function<bool(Batch*)> produce_;
  int batch_size_ = 0, capacity_ = 1;
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<Batch> q_;
  bool eof_ = false, running_ = false;
  std::thread th_;
};

// Reads lines through a LineInputFormat on a background thread, parses them and hands out
// batches of `batch_size` samples, at most `batch_num` batches ahead of the consumer.
template <typename Sample>
class AsyncDataLoader {
 public:
  AsyncDataLoader(LineInputFormat* in, std::function<bool(const std::string&, Sample*)> parse, int batch_size,
                  int batch_num) {
    buffer_.init(
        [in, parse, batch_size](std::vector<Sample>* b) {
          std::string line;
          while (static_cast<int>(b->size()) < batch_size) {
            if (!in->next(&line)) return false;
            Sample s;
            if (parse(line, &s)) b->push_back(std::move(s));
          }
          return true;
        },
        batch_size, batch_num);
  }
  bool get_data(std::vector<Sample>* batch) { return buffer_.get_batch(batch); }
  bool end_of_file() { return buffer_.end_of_file(); }

 private:
  AsyncReadBuffer<Sample> buffer_;
};

// T must look like SVMItem: .first = vector<pair<int,double>>
template <typename T>
class BatchDataSampler {
 public:
  BatchDataSampler(const std::vector<T>& data, int batch_size, uint64_t seed = 0)
      : data_(data), batch_size_(batch_size), rng_(seed ? seed : std::random_device{}()) {
    MINIPS_CHECK(!data_.empty()) << "empty dataset";
    batch_ptrs_.resize(batch_size_);
  }
  int get_batch_size() const { return batch_size_; }
  size_t size() const { return data_.size(); }
  void random_start_point() { start_ = static_cast<size_t>(rng_() % data_.size()); }
  void set_start_point(size_t s) { start_ = s % data_.size(); }
  // takes batch_size consecutive samples (wrapping around) and returns the sorted unique
  // feature ids they touch — the key list of the iteration's Get/Add
  SArray<Key> prepare_next_batch() {
    std::set<Key> keys;
    for (int i = 0; i < batch_size_; ++i) {
      const T* s = &data_[start_];
      batch_ptrs_[i] = s;
      for (const auto& f : s->first) keys.insert(static_cast<Key>(f.first));
      start_ = (start_ + 1) % data_.size();
    }
    SArray<Key> out;
    out.CopyFrom(keys.begin(), keys.end());
    return out;
  }
  const std::vector<const T*>& get_data_ptrs() const { return batch_ptrs_; }

 private:
  std::vector<T> data_;  // own copy, like the reference (batch_data_sampler.cpp:28-32)
  int batch_size_;
  size_t start_ = 0;
  std::mt19937_64 rng_;
  std::vector<const T*> batch_ptrs_;
};

struct SVMDumper {
  // <prefix>worker_<my_id>: "label idx:val ..." lines with 1-based indices, so the dump is valid
  // libsvm and reloads through parse_libsvm unchanged (the reference writes 0-based indices and
  // shifts them again on reload — SURVEY 2.8-5).
  static void DumpSVMData(const std::vector<SVMItem>& data, const std::string& path = "");
  static std::vector<SVMItem> LoadSVMData(const std::string& path = "");
  static void DumpConfigData(const std::map<int, int>& iter_map, const std::string& path = "");
  static std::map<int, int> LoadConfigData(const std::string& path = "");
  static void DumpScaleFile(const Node& node, const std::string& path);   // "id:host:port"
  static Node LoadScaleFile(const std::string& path);
};

}  // namespace lib
}  // namespace minips
