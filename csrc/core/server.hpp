// server.hpp — the server side of the host runtime: storages, progress tracking, the three
// consistency models and the server actor.
//
// Reference counterparts (behaviour):
//   AbstractStorage   server/abstract_storage.hpp:13-48    MapStorage     server/map_storage.hpp:13-36
//   VectorStorage     server/vector_storage.hpp:22-94      ProgressTracker server/util/progress_tracker.{hpp,cpp}
//   PendingBuffer     server/util/pending_buffer.cpp:5-37   AbstractModel  server/abstract_model.hpp:12-26
//   BSP/SSP/ASPModel  server/consistency/*.cpp              ServerThread   server/server_thread.{hpp,cpp}
// These classes are the executable specification of the consistency semantics that the
// device kernels (csrc/kernels/clock.cu) implement with flags in peer memory, and the
// servers of the CPU backend.
#pragma once

#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "core/base.hpp"
#include "core/context.hpp"

namespace minips {

// ------------------------------------------------------------------ storage
class AbstractStorage {
 public:
  virtual ~AbstractStorage() = default;

  void Add(Message& msg) {
    MINIPS_CHECK(msg.data.size() == 2) << "kAdd carries [keys, vals]";
    SubAdd(SArray<Key>(msg.data[0]), msg.data[1]);
  }
  // Builds the reply: endpoints swapped, same flag / model, keys echoed, values gathered.
  Message Get(Message& msg) {
    MINIPS_CHECK(msg.data.size() == 1) << "kGet carries [keys]";
    SArray<Key> keys(msg.data[0]);
    Message reply;
    reply.meta.recver = msg.meta.sender;
    reply.meta.sender = msg.meta.recver;
    reply.meta.flag = msg.meta.flag;
    reply.meta.model_id = msg.meta.model_id;
    reply.data.push_back(msg.data[0]);
    reply.data.push_back(SubGet(keys));
    return reply;
  }

  virtual void SubAdd(const SArray<Key>& keys, const SArray<char>& vals) = 0;
  virtual SArray<char> SubGet(const SArray<Key>& keys) = 0;
  virtual void FinishIter() = 0;
  virtual void Dump() {}
  virtual void Restore() {}
  virtual void Update(const Range& /*range*/) {}
  virtual size_t Size() const = 0;
};

template <typename Val>
class MapStorage : public AbstractStorage {
 public:
  void SubAdd(const SArray<Key>& keys, const SArray<char>& vals) override {
    SArray<Val> v(vals);
    MINIPS_CHECK_EQ(keys.size(), v.size());
    for (size_t i = 0; i < keys.size(); ++i) storage_[keys[i]] += v[i];
  }
  SArray<char> SubGet(const SArray<Key>& keys) override {
    SArray<Val> out(keys.size());
    for (size_t i = 0; i < keys.size(); ++i) out[i] = storage_[keys[i]];  // inserts 0 like the reference
    return SArray<char>(out);
  }
  void FinishIter() override {}
  size_t Size() const override { return storage_.size(); }
  const std::map<Key, Val>& Raw() const { return storage_; }

 private:
  std::map<Key, Val> storage_;
};

std::string ServerParamsPath();  // <checkpoint_file_prefix>server_params_<my_id>

template <typename Val>
class VectorStorage : public AbstractStorage {
 public:
  explicit VectorStorage(const Range& range) : range_(range), storage_(range.size(), Val()) {}

  void SubAdd(const SArray<Key>& keys, const SArray<char>& vals) override {
    SArray<Val> v(vals);
    MINIPS_CHECK_EQ(keys.size(), v.size());
    for (size_t i = 0; i < keys.size(); ++i) storage_[Index(keys[i])] += v[i];
  }
  SArray<char> SubGet(const SArray<Key>& keys) override {
    SArray<Val> out(keys.size());
    for (size_t i = 0; i < keys.size(); ++i) out[i] = storage_[Index(keys[i])];
    return SArray<char>(out);
  }
  void FinishIter() override {}
  size_t Size() const override { return storage_.size(); }
  std::vector<Val>& Raw() { return storage_; }

  // "<local_idx>:<val> " for every non-zero entry, one line (vector_storage.hpp:54-73).
  // Values are written with round-trip precision; the reference's 6 digits lose weights.
  void Dump() override {
    if (!Context::get_instance().get_bool("checkpoint_toggle")) return;
    DumpTo(ServerParamsPath());
  }
  void Restore() override {
    if (!Context::get_instance().get_bool("checkpoint_toggle")) return;
    RestoreFrom(ServerParamsPath());
  }
  void DumpTo(const std::string& path) const;
  void RestoreFrom(const std::string& path);
  void Update(const Range& range) override { range_ = range; }
  const Range& GetRange() const { return range_; }

 private:
  size_t Index(Key k) const {
    MINIPS_CHECK(k >= range_.begin() && k < range_.end())
        << "key " << k << " outside [" << range_.begin() << "," << range_.end() << ")";
    return static_cast<size_t>(k - range_.begin());
  }
  Range range_;
  std::vector<Val> storage_;
};

// ------------------------------------------------------------------ tracker / pending
class ProgressTracker {
 public:
  void Init(const std::vector<uint32_t>& tids);
  // progress[tid]++ ; returns the new min clock if tid was the unique minimum, else -1
  int AdvanceAndGetChangedMinClock(int tid);
  int GetNumThreads() const { return static_cast<int>(progresses_.size()); }
  int GetProgress(int tid) const;
  int GetMinClock() const { return min_clock_; }
  bool IsUniqueMin(int tid) const;
  bool CheckThreadValid(int tid) const { return progresses_.count(tid) > 0; }
  // drop the workers of a node ([node*1000, (node+1)*1000)); returns the new min if it moved, else -1
  int DeleteNode(uint32_t node_id);
  // <checkpoint_file_prefix>server_progress_<my_id>: "min_clock:<v> <tid>:<v> ..." with v rounded to
  // the checkpoint interval (progress_tracker.hpp:64-122)
  void Dump() const;
  void Restore();
  void DumpTo(const std::string& path) const;
  void RestoreFrom(const std::string& path);
  std::string DebugString() const;
  const std::map<int, int>& Raw() const { return progresses_; }

 private:
  std::map<int, int> progresses_;
  int min_clock_ = 0;
};

class PendingBuffer {
 public:
  void Push(int clock, Message& msg) { buffer_[clock].push_back(msg); }
  std::vector<Message> Pop(int clock);
  std::vector<Message> PopAll();
  void EraseAll() { buffer_.clear(); }
  int Size(int progress) const;

 private:
  std::unordered_map<int, std::vector<Message>> buffer_;
};

// ------------------------------------------------------------------ consistency models
class AbstractModel {
 public:
  virtual ~AbstractModel() = default;
  virtual void Clock(Message& msg) = 0;
  virtual void Add(Message& msg) = 0;
  virtual void Get(Message& msg) = 0;
  virtual int GetProgress(int tid) = 0;
  virtual void ResetWorker(Message& msg) = 0;
  virtual void Dump(Message& msg) = 0;
  virtual void Restore() = 0;
  virtual void Update(int failed_node_id, const std::vector<Node>& nodes, const Range& range) = 0;
};

class ModelBase : public AbstractModel {
 public:
  ModelBase(uint32_t model_id, std::unique_ptr<AbstractStorage>&& storage, ThreadsafeQueue<Message>* reply_queue)
      : model_id_(model_id), storage_(std::move(storage)), reply_queue_(reply_queue) {}
  int GetProgress(int tid) override { return progress_tracker_.GetProgress(tid); }
  void ResetWorker(Message& msg) override;
  // Dumps storage + tracker + the iteration map, then acks with kCheckpoint.  (The reference
  // only implements this for SSP; its BSP/ASP never ack and the caller hangs — SURVEY 2.8-3.)
  void Dump(Message& msg) override;
  void Restore() override;
  void Update(int failed_node_id, const std::vector<Node>& nodes, const Range& range) override;
  AbstractStorage* GetStorage() { return storage_.get(); }
  ProgressTracker& GetTracker() { return progress_tracker_; }

 protected:
  virtual void AfterRestore() {}
  uint32_t model_id_;
  std::unique_ptr<AbstractStorage> storage_;
  ThreadsafeQueue<Message>* reply_queue_;
  ProgressTracker progress_tracker_;
};

class BSPModel : public ModelBase {
 public:
  using ModelBase::ModelBase;
  void Clock(Message& msg) override;
  void Add(Message& msg) override;
  void Get(Message& msg) override;
  int GetGetPendingSize() const { return static_cast<int>(get_buffer_.size()); }
  int GetAddPendingSize() const { return static_cast<int>(add_buffer_.size()); }

 protected:
  void AfterRestore() override;

 private:
  std::vector<Message> get_buffer_;
  std::vector<Message> add_buffer_;
};

class SSPModel : public ModelBase {
 public:
  SSPModel(uint32_t model_id, std::unique_ptr<AbstractStorage>&& storage, int staleness,
           ThreadsafeQueue<Message>* reply_queue);
  void Clock(Message& msg) override;
  void Add(Message& msg) override;
  void Get(Message& msg) override;
  void ResetWorker(Message& msg) override;
  int GetPendingSize(int progress) const { return buffer_.Size(progress); }
  int GetStaleness() const { return staleness_; }

 protected:
  void AfterRestore() override;

 private:
  void Flush(int updated_min_clock);
  int staleness_;
  PendingBuffer buffer_;
};

class ASPModel : public ModelBase {
 public:
  using ModelBase::ModelBase;
  void Clock(Message& msg) override;
  void Add(Message& msg) override;
  void Get(Message& msg) override;
};

// ------------------------------------------------------------------ server actor
class ServerThread : public Actor {
 public:
  explicit ServerThread(uint32_t server_id) : Actor(server_id) {}
  void RegisterModel(uint32_t model_id, std::unique_ptr<AbstractModel>&& model);
  AbstractModel* GetModel(uint32_t model_id);
  void UpdateModel(int failed_node_id, const std::vector<Node>& nodes, const Range& range);
  void RollbackModel();

 protected:
  void Main() override;

 private:
  std::mutex mu_;  // the receiver thread may roll models back while the actor is serving
  std::unordered_map<uint32_t, std::unique_ptr<AbstractModel>> models_;
};

}  // namespace minips
