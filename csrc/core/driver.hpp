// driver.hpp — the user-facing driver of the host runtime.
//
//   SimpleIdMapper  driver/simple_id_mapper.{hpp,cpp}   (tid/1000 = node; servers [0,50), helper 50,
//                                                       user workers [100,1000))
//   WorkerSpec      driver/worker_spec.{hpp,cpp}         WorkerAlloc / MLTask  driver/ml_task.hpp
//   Info            driver/info.hpp                      Engine  driver/engine.{hpp,cpp}
//
// The Engine owns id mapper, mailbox, sender, one worker helper, N server threads, the
// partition managers and the heartbeat / rollback state of one process ("node").
#pragma once

#include <atomic>
#include <functional>
#include <map>
#include <memory>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "core/base.hpp"
#include "core/context.hpp"
#include "core/mailbox.hpp"
#include "core/partition.hpp"
#include "core/server.hpp"
#include "core/worker.hpp"

namespace minips {

// ------------------------------------------------------------------ id mapper
class SimpleIdMapper : public AbstractIdMapper {
 public:
  static constexpr uint32_t kMaxNodeId = 1000;
  static constexpr uint32_t kMaxThreadsPerNode = 1000;
  static constexpr uint32_t kMaxBgThreadsPerNode = 100;  // [0,100) reserved for background threads
  static constexpr uint32_t kWorkerHelperThreadId = 50;  // servers occupy [0,50)

  SimpleIdMapper(const Node& node, const std::vector<Node>& nodes) : node_(node), nodes_(nodes) {}

  uint32_t GetNodeIdForThread(uint32_t tid) override { return tid / kMaxThreadsPerNode; }
  void Init(int num_server_threads_per_node = 1);
  void Update(const std::vector<Node>& nodes, int num_server_threads_per_node = 1);
  uint32_t AllocateWorkerThread(uint32_t node_id);
  void DeallocateWorkerThread(uint32_t node_id, uint32_t tid);

  std::vector<uint32_t> GetServerThreadsForId(uint32_t node_id) const;
  std::vector<uint32_t> GetWorkerHelperThreadsForId(uint32_t node_id) const;
  std::vector<uint32_t> GetWorkerThreadsForId(uint32_t node_id) const;
  std::vector<uint32_t> GetAllServerThreads() const;

 private:
  Node node_;
  std::vector<Node> nodes_;
  std::map<uint32_t, std::vector<uint32_t>> node2server_, node2helper_;
  std::map<uint32_t, std::set<uint32_t>> node2worker_;
};

// ------------------------------------------------------------------ task description
struct WorkerAlloc {
  uint32_t node_id;
  uint32_t num_workers;
};

class WorkerSpec {
 public:
  WorkerSpec() = default;
  explicit WorkerSpec(const std::vector<WorkerAlloc>& worker_alloc);
  bool HasLocalWorkers(uint32_t node_id) const;
  const std::vector<uint32_t>& GetLocalWorkers(uint32_t node_id) const;
  const std::vector<uint32_t>& GetLocalThreads(uint32_t node_id) const;
  std::map<uint32_t, std::vector<uint32_t>> GetNodeToWorkers() const { return node_to_workers_; }
  std::vector<uint32_t> GetAllThreadIds() const;
  void InsertWorkerIdThreadId(uint32_t worker_id, uint32_t thread_id);
  uint32_t GetNumWorkers() const { return num_workers_; }
  const std::map<uint32_t, uint32_t>& GetWorkerToThreadMapper() const { return worker_to_thread_; }
  const std::map<uint32_t, uint32_t>& GetThreadToWorkerMapper() const { return thread_to_worker_; }
  const std::map<uint32_t, uint32_t>& GetWorkerToNodeMapper() const { return worker_to_node_; }

 private:
  std::map<uint32_t, uint32_t> worker_to_node_, worker_to_thread_, thread_to_worker_;
  std::map<uint32_t, std::vector<uint32_t>> node_to_workers_, node_to_threads_;
  std::set<uint32_t> thread_ids_;
  uint32_t num_workers_ = 0;
};

class Engine;

// per-user-thread context handed to the task lambda
struct Info {
  uint32_t thread_id = 0;
  uint32_t worker_id = 0;
  ThreadsafeQueue<Message>* send_queue = nullptr;
  std::map<uint32_t, AbstractPartitionManager*> partition_manager_map;
  AbstractCallbackRunner* callback_runner = nullptr;
  Engine* engine = nullptr;

  template <typename Val>
  std::unique_ptr<KVClientTable<Val>> CreateKVClientTable(uint32_t table_id) const {
    auto it = partition_manager_map.find(table_id);
    MINIPS_CHECK(it != partition_manager_map.end()) << "table " << table_id << " is not part of this task";
    return std::unique_ptr<KVClientTable<Val>>(
        new KVClientTable<Val>(thread_id, table_id, send_queue, it->second, callback_runner));
  }
  std::string DebugString() const {
    std::ostringstream os;
    os << "thread_id: " << thread_id << " worker_id: " << worker_id;
    return os.str();
  }
};

class MLTask {
 public:
  void SetLambda(const std::function<void(const Info&)>& func) { func_ = func; }
  void RunLambda(const Info& info) const { func_(info); }
  void SetWorkerAlloc(const std::vector<WorkerAlloc>& worker_alloc) { worker_alloc_ = worker_alloc; }
  const std::vector<WorkerAlloc>& GetWorkerAlloc() const { return worker_alloc_; }
  void SetTables(const std::vector<uint32_t>& tables) { tables_ = tables; }
  const std::vector<uint32_t>& GetTables() const { return tables_; }
  bool IsSetup() const { return func_ && !worker_alloc_.empty() && !tables_.empty(); }

 private:
  std::function<void(const Info&)> func_;
  std::vector<WorkerAlloc> worker_alloc_;
  std::vector<uint32_t> tables_;
};

enum class ModelType { SSP, BSP, ASP };
enum class StorageType { Map, Vector };

// ------------------------------------------------------------------ engine
class Engine : public MailboxListener {
 public:
  Engine(const Node& node, const std::vector<Node>& nodes, const Node& master_node = Node(),
         const Node& scale_node = Node());
  ~Engine() override;

  // lifecycle
  void StartEverything(int num_server_threads_per_node = 1);
  void StopEverything();
  void CreateIdMapper(int num_server_threads_per_node = 1);
  void CreateMailbox();
  void StartMailbox();
  void StartSender();
  void StartServerThreads();
  void StartWorkerThreads();
  void StartHeartbeatThread();
  void StopHeartbeatThread();
  void AbandonHeartbeatThread();  // stop without the goodbye message (the node is not leaving for good)
  void StopServerThreads();
  void StopWorkerThreads();
  void StopSender();
  void StopMailbox();

  // tables
  std::vector<Range> getRanges();  // even split of `num_dims` over all server threads
  template <typename Val>
  uint32_t CreateTable(const std::vector<Range>& ranges, ModelType model_type, StorageType storage_type,
                       int model_staleness = 0);
  template <typename Val>
  uint32_t CreateTable(std::unique_ptr<AbstractPartitionManager> partition_manager, ModelType model_type,
                       StorageType storage_type, int model_staleness = 0);
  void InitTable(uint32_t table_id, const std::vector<uint32_t>& worker_ids);

  // execution
  void Barrier();
  WorkerSpec AllocateWorkers(const std::vector<WorkerAlloc>& worker_alloc);
  void Run(const MLTask& task);
  void ForceQuit();  // leave the cluster before training (no local data)

  // control messages (engine.hpp:70-114)
  void HeartBeat(uint32_t node_id, bool quit = false);
  void Scale(uint32_t node_id);
  void RollBack(uint32_t failed_node_id);       // master only: broadcast kRollBack
  void ScaleRollBack(uint32_t scale_node_id);   // master only

  // rollback state shared with user threads (engine.hpp:234-258)
  void SetNeedRollBack(bool v) { need_rollback_ = v; }
  bool IsNeedRollBack() const { return need_rollback_; }
  int IncRollBackCount();       // returns the new count
  void RecoverEnd();            // first worker of the node calls this after the fault barrier
  void WaitRecover();           // the others wait for it
  void RollBackServer();
  void RollBackWorker();

  // membership updates
  void UpdateNodes(const std::vector<Node>& nodes);
  void UpdateAndRestart(int failed_node_id, const std::vector<Node>& nodes);
  void SetRestarter(std::function<void()> f) { restarter_ = std::move(f); }
  void SetDumpCallback(std::function<void()> f) { dump_callback_ = std::move(f); }
  void RunDumpCallback() { if (dump_callback_) dump_callback_(); }

  // MailboxListener
  void OnForceQuit(uint32_t node_id) override;
  void OnRollBack(int failed_node_id) override;
  void OnScaleRollBack(int scale_node_id) override;
  void OnCheckpointRequest() override { RunDumpCallback(); }

  // accessors
  SimpleIdMapper* GetIdMapper() { return id_mapper_.get(); }
  Mailbox* GetMailbox() { return mailbox_.get(); }
  Sender* GetSender() { return sender_.get(); }
  const Node& GetNode() const { return node_; }
  const std::vector<Node>& GetNodes() const { return nodes_; }
  bool HasMaster() const { return master_node_.is_master; }
  std::vector<ServerThread*> GetServerThreads();
  WorkerThread* GetWorkerThread() { return worker_thread_.get(); }
  AbstractPartitionManager* GetPartitionManager(uint32_t table_id);
  int GetRollBackIteration() const { return rollback_iteration_; }
  int GetRollBackSignals() const { return rollback_signals_; }  // kRollBack messages seen so far

 private:
  void RegisterPartitionManager(uint32_t table_id, std::unique_ptr<AbstractPartitionManager> pm);
  template <typename Val>
  std::unique_ptr<AbstractStorage> MakeStorage(StorageType type, const Range& range);
  std::unique_ptr<AbstractModel> MakeModel(ModelType type, uint32_t model_id, std::unique_ptr<AbstractStorage> st,
                                           int staleness);

  Node node_;
  std::vector<Node> nodes_;
  Node master_node_;
  Node scale_node_;
  int num_server_threads_per_node_ = 1;

  std::unique_ptr<SimpleIdMapper> id_mapper_;
  std::unique_ptr<Mailbox> mailbox_;
  std::unique_ptr<Sender> sender_;
  std::vector<std::unique_ptr<ServerThread>> server_threads_;
  std::unique_ptr<WorkerThread> worker_thread_;
  std::map<uint32_t, std::unique_ptr<AbstractPartitionManager>> partition_managers_;
  uint32_t model_count_ = 0;

  std::thread heartbeat_thread_;
  std::atomic<bool> heartbeat_running_{false};

  std::atomic<bool> need_rollback_{false};
  std::atomic<int> rollback_signals_{0};
  std::atomic<bool> heartbeat_goodbye_{true};
  std::mutex recover_mu_;
  std::condition_variable recover_cv_;
  int rollback_counter_ = 0;
  bool recover_end_ = false;
  int rollback_iteration_ = 0;

  std::function<void()> restarter_;
  std::function<void()> dump_callback_;
};

// ---- template members
template <typename Val>
std::unique_ptr<AbstractStorage> Engine::MakeStorage(StorageType type, const Range& range) {
  if (type == StorageType::Map) return std::unique_ptr<AbstractStorage>(new MapStorage<Val>());
  return std::unique_ptr<AbstractStorage>(new VectorStorage<Val>(range));
}

template <typename Val>
uint32_t Engine::CreateTable(std::unique_ptr<AbstractPartitionManager> partition_manager, ModelType model_type,
                             StorageType storage_type, int model_staleness) {
  MINIPS_CHECK(id_mapper_) << "StartEverything first";
  auto* rpm = dynamic_cast<RangePartitionManager*>(partition_manager.get());
  const auto server_tids = partition_manager->GetServerThreadIds();
  const uint32_t table_id = model_count_++;
  // one storage + one consistency model per LOCAL server thread (engine.hpp:301-329)
  for (auto& server : server_threads_) {
    Range range;
    if (rpm) {
      for (size_t i = 0; i < server_tids.size(); ++i)
        if (server_tids[i] == server->GetId()) range = rpm->GetRanges()[i];
    }
    auto storage = MakeStorage<Val>(storage_type, range);
    server->RegisterModel(table_id, MakeModel(model_type, table_id, std::move(storage), model_staleness));
  }
  RegisterPartitionManager(table_id, std::move(partition_manager));
  return table_id;
}

template <typename Val>
uint32_t Engine::CreateTable(const std::vector<Range>& ranges, ModelType model_type, StorageType storage_type,
                             int model_staleness) {
  MINIPS_CHECK(id_mapper_) << "StartEverything first";
  const auto server_tids = id_mapper_->GetAllServerThreads();
  MINIPS_CHECK_EQ(ranges.size(), server_tids.size());
  std::unique_ptr<AbstractPartitionManager> pm(
      new RangePartitionManager(server_tids, ranges, HasMaster() ? static_cast<int>(master_node_.id) : -1));
  return CreateTable<Val>(std::move(pm), model_type, storage_type, model_staleness);
}

}  // namespace minips
