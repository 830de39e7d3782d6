// worker.hpp — the worker side of the host runtime.
//
//   AbstractCallbackRunner  worker/abstract_callback_runner.hpp:14-42
//   WorkerThread            worker/worker_thread.{hpp,cpp} ("worker helper": receives pull replies
//                           and checkpoint acks for every user thread of the node)
//   KVClientTable<Val>      worker/kv_client_table.hpp:45-208 (Add / Get / Clock / CheckPoint)
//
// Deviations: replies are interpreted as Val (the reference always reads them as double,
// kv_client_table.hpp:98-99, so only Val=double round-trips there); handler maps are
// guarded by the runner's mutex (SURVEY 5.2 lists the reference's unsynchronised reads).
#pragma once

#include <algorithm>
#include <condition_variable>
#include <functional>
#include <map>
#include <mutex>
#include <utility>
#include <vector>

#include "core/base.hpp"
#include "core/partition.hpp"

namespace minips {

class AbstractCallbackRunner {
 public:
  virtual ~AbstractCallbackRunner() = default;
  virtual void RegisterRecvHandle(uint32_t app_thread_id, uint32_t model_id,
                                  const std::function<void(Message&)>& recv_handle) = 0;
  virtual void RegisterRecvFinishHandle(uint32_t app_thread_id, uint32_t model_id,
                                        const std::function<void()>& recv_finish_handle) = 0;
  virtual void NewRequest(uint32_t app_thread_id, uint32_t model_id, uint32_t expected_responses) = 0;
  virtual void WaitRequest(uint32_t app_thread_id, uint32_t model_id) = 0;
  virtual void AddResponse(uint32_t app_thread_id, uint32_t model_id, Message& msg) = 0;
  virtual void NewCheckPoint(uint32_t app_thread_id, uint32_t model_id, uint32_t expected_responses) = 0;
  virtual void WaitCheckPoint(uint32_t app_thread_id, uint32_t model_id) = 0;
  virtual void CheckPointResponse(uint32_t app_thread_id, uint32_t model_id) = 0;
};

class WorkerThread : public Actor, public AbstractCallbackRunner {
 public:
  explicit WorkerThread(uint32_t worker_id) : Actor(worker_id) {}

  void RegisterRecvHandle(uint32_t app_thread_id, uint32_t model_id,
                          const std::function<void(Message&)>& recv_handle) override;
  void RegisterRecvFinishHandle(uint32_t app_thread_id, uint32_t model_id,
                                const std::function<void()>& recv_finish_handle) override;
  void NewRequest(uint32_t app_thread_id, uint32_t model_id, uint32_t expected_responses) override;
  void WaitRequest(uint32_t app_thread_id, uint32_t model_id) override;
  void AddResponse(uint32_t app_thread_id, uint32_t model_id, Message& msg) override;
  void NewCheckPoint(uint32_t app_thread_id, uint32_t model_id, uint32_t expected_responses) override;
  void WaitCheckPoint(uint32_t app_thread_id, uint32_t model_id) override;
  void CheckPointResponse(uint32_t app_thread_id, uint32_t model_id) override;

  // rollback: complete every outstanding request so blocked Gets return (worker_thread.cpp:137-160)
  void RollBackWorker();
  // a node was lost: stop expecting its shards' replies (worker_thread.cpp:162-190)
  void Update(uint32_t lost_replies_per_request);

 protected:
  void Main() override;

 private:
  using Slot = std::pair<uint32_t, uint32_t>;  // (app thread, model)
  struct Tracker {
    uint32_t expected = 0, current = 0;
  };
  std::mutex mu_;
  std::condition_variable cond_;
  std::map<Slot, Tracker> tracker_, ckpt_tracker_;
  std::map<Slot, std::function<void(Message&)>> recv_handles_;
  std::map<Slot, std::function<void()>> finish_handles_;
};

template <typename Val>
class KVClientTable {
 public:
  KVClientTable(uint32_t app_thread_id, uint32_t model_id, ThreadsafeQueue<Message>* sender_queue,
                const AbstractPartitionManager* partition_manager, AbstractCallbackRunner* callback_runner)
      : app_thread_id_(app_thread_id),
        model_id_(model_id),
        sender_queue_(sender_queue),
        partition_manager_(partition_manager),
        callback_runner_(callback_runner) {
    callback_runner_->RegisterRecvHandle(app_thread_id_, model_id_, [this](Message& m) { HandleMsg(m); });
  }

  // keys must be sorted ascending (binary-search slicing)
  void Add(const std::vector<Key>& keys, const std::vector<Val>& vals) { Add(SArray<Key>(keys), SArray<Val>(vals)); }
  void Add(const SArray<Key>& keys, const SArray<Val>& vals) {
    // fire-and-forget: one kAdd [keys, vals] per owning server, zero-copy segments
    std::vector<TypedSlice> sliced;
    SliceTyped(keys, vals, &sliced);
    for (auto& s : sliced) {
      Message m = MakeMsg(s.server, Flag::kAdd);
      m.AddData(s.keys);
      m.AddData(s.vals);
      sender_queue_->Push(m);
    }
  }

  void Get(const std::vector<Key>& keys, std::vector<Val>* vals) {
    SArray<Val> out;
    Get(SArray<Key>(keys), &out);
    vals->assign(out.begin(), out.end());
  }
  void Get(const SArray<Key>& keys, SArray<Val>* vals) {
    std::vector<std::pair<int, SArray<Key>>> sliced;
    partition_manager_->Slice(keys, &sliced);
    {
      std::lock_guard<std::mutex> lk(mu_);
      recv_kvs_.clear();
    }
    if (sliced.empty()) {
      vals->clear();
      return;
    }
    callback_runner_->RegisterRecvFinishHandle(app_thread_id_, model_id_, [this, vals]() { HandleFinish(vals); });
    callback_runner_->NewRequest(app_thread_id_, model_id_, static_cast<uint32_t>(sliced.size()));
    for (auto& s : sliced) {
      Message m = MakeMsg(static_cast<uint32_t>(s.first), Flag::kGet);
      m.AddData(s.second);
      sender_queue_->Push(m);
    }
    callback_runner_->WaitRequest(app_thread_id_, model_id_);
  }

  // kClock to EVERY server thread of the table (kv_client_table.hpp:123-134)
  void Clock() {
    for (uint32_t server : partition_manager_->GetServerThreadIds()) sender_queue_->Push(MakeMsg(server, Flag::kClock));
  }

  // blocking: every shard dumps and acks (kv_client_table.hpp:136-154)
  void CheckPoint() {
    const auto& servers = partition_manager_->GetServerThreadIds();
    callback_runner_->NewCheckPoint(app_thread_id_, model_id_, static_cast<uint32_t>(servers.size()));
    for (uint32_t server : servers) sender_queue_->Push(MakeMsg(server, Flag::kCheckpoint));
    callback_runner_->WaitCheckPoint(app_thread_id_, model_id_);
  }

  // liveness ping to the master node (kv_client_table.hpp:156-166)
  void HeartBeat(uint32_t node_id, bool quit = false) {
    const int master = partition_manager_->GetMasterNodeId();
    if (master < 0) return;
    Message m;
    m.meta.sender = static_cast<int>(node_id);
    m.meta.recver = master;
    m.meta.model_id = static_cast<int>(model_id_);
    m.meta.flag = quit ? Flag::kQuitHeartBeat : Flag::kHeartBeat;
    sender_queue_->Push(m);
  }

 private:
  struct TypedSlice {
    uint32_t server;
    SArray<Key> keys;
    SArray<Val> vals;
  };

  Message MakeMsg(uint32_t server, Flag flag) const {
    Message m;
    m.meta.sender = static_cast<int>(app_thread_id_);
    m.meta.recver = static_cast<int>(server);
    m.meta.model_id = static_cast<int>(model_id_);
    m.meta.flag = flag;
    return m;
  }

  void SliceTyped(const SArray<Key>& keys, const SArray<Val>& vals, std::vector<TypedSlice>* out) {
    out->clear();
    std::vector<std::pair<int, SArray<Key>>> ks;
    partition_manager_->Slice(keys, &ks);
    if (keys.empty()) return;
    MINIPS_CHECK(vals.size() % keys.size() == 0) << "values must be a multiple of keys";
    const size_t ratio = vals.size() / keys.size();
    for (auto& s : ks) {
      const size_t b = s.second.data() - keys.data(), e = b + s.second.size();
      out->push_back({static_cast<uint32_t>(s.first), s.second, vals.segment(b * ratio, e * ratio)});
    }
  }

  void HandleMsg(Message& msg) {
    MINIPS_CHECK_EQ(msg.data.size(), 2u);
    std::lock_guard<std::mutex> lk(mu_);
    recv_kvs_.emplace_back(SArray<Key>(msg.data[0]), SArray<Val>(msg.data[1]));
  }

  // order the per-shard replies by first key and concatenate (kv_client_table.hpp:74-94)
  void HandleFinish(SArray<Val>* vals) {
    std::lock_guard<std::mutex> lk(mu_);
    std::sort(recv_kvs_.begin(), recv_kvs_.end(), [](const auto& a, const auto& b) {
      if (a.first.empty() || b.first.empty()) return !a.first.empty() && b.first.empty() ? true : false;
      return a.first[0] < b.first[0];
    });
    size_t total = 0;
    for (auto& kv : recv_kvs_) total += kv.second.size();
    vals->resize(total);
    size_t off = 0;
    for (auto& kv : recv_kvs_) {
      if (kv.second.size()) std::memcpy(vals->data() + off, kv.second.data(), kv.second.size() * sizeof(Val));
      off += kv.second.size();
    }
    recv_kvs_.clear();
  }

  uint32_t app_thread_id_;
  uint32_t model_id_;
  ThreadsafeQueue<Message>* sender_queue_;
  const AbstractPartitionManager* partition_manager_;
  AbstractCallbackRunner* callback_runner_;
  std::mutex mu_;
  std::vector<std::pair<SArray<Key>, SArray<Val>>> recv_kvs_;
};

}  // namespace minips
