#include "core/driver.hpp"

#include <chrono>

#include "core/checkpoint_io.hpp"

namespace minips {

// ------------------------------------------------------------------ SimpleIdMapper
void SimpleIdMapper::Init(int num_server_threads_per_node) {
  MINIPS_CHECK(num_server_threads_per_node > 0 &&
               num_server_threads_per_node <= static_cast<int>(kWorkerHelperThreadId));
  node2server_.clear();
  node2helper_.clear();
  node2worker_.clear();
  const bool has_scale = Context::get_instance().get_bool("has_scale_node");
  const int scale_id = Context::get_instance().get_int32("scale_node_id");
  for (const auto& n : nodes_) {
    MINIPS_CHECK_LT(n.id, kMaxNodeId);
    // a scale-out node hosts workers only (simple_id_mapper.cpp:25-29)
    if (!(has_scale && static_cast<int>(n.id) == scale_id))
      for (int i = 0; i < num_server_threads_per_node; ++i)
        node2server_[n.id].push_back(n.id * kMaxThreadsPerNode + static_cast<uint32_t>(i));
    node2helper_[n.id].push_back(n.id * kMaxThreadsPerNode + kWorkerHelperThreadId);
    node2worker_[n.id];
  }
}

void SimpleIdMapper::Update(const std::vector<Node>& nodes, int num_server_threads_per_node) {
  nodes_ = nodes;
  Init(num_server_threads_per_node);
}

uint32_t SimpleIdMapper::AllocateWorkerThread(uint32_t node_id) {
  MINIPS_CHECK(node2helper_.count(node_id)) << "unknown node " << node_id;
  auto& used = node2worker_[node_id];
  MINIPS_CHECK_LE(used.size(), kMaxThreadsPerNode - kMaxBgThreadsPerNode);
  for (uint32_t i = kMaxBgThreadsPerNode; i < kMaxThreadsPerNode; ++i) {
    const uint32_t tid = node_id * kMaxThreadsPerNode + i;
    if (!used.count(tid)) {
      used.insert(tid);
      return tid;
    }
  }
  MINIPS_CHECK(false) << "no free worker thread id on node " << node_id;
  return 0;
}

void SimpleIdMapper::DeallocateWorkerThread(uint32_t node_id, uint32_t tid) {
  auto it = node2worker_.find(node_id);
  MINIPS_CHECK(it != node2worker_.end() && it->second.count(tid)) << "tid " << tid << " not allocated";
  it->second.erase(tid);
}

std::vector<uint32_t> SimpleIdMapper::GetServerThreadsForId(uint32_t node_id) const {
  auto it = node2server_.find(node_id);
  return it == node2server_.end() ? std::vector<uint32_t>() : it->second;
}
std::vector<uint32_t> SimpleIdMapper::GetWorkerHelperThreadsForId(uint32_t node_id) const {
  auto it = node2helper_.find(node_id);
  return it == node2helper_.end() ? std::vector<uint32_t>() : it->second;
}
std::vector<uint32_t> SimpleIdMapper::GetWorkerThreadsForId(uint32_t node_id) const {
  auto it = node2worker_.find(node_id);
  return it == node2worker_.end() ? std::vector<uint32_t>() : std::vector<uint32_t>(it->second.begin(),
      it->second.end());
}
std::vector<uint32_t> SimpleIdMapper::GetAllServerThreads() const {
  std::vector<uint32_t> out;
  for (const auto& kv : node2server_) out.insert(out.end(), kv.second.begin(), kv.second.end());
  return out;
}

// ------------------------------------------------------------------ WorkerSpec
WorkerSpec::WorkerSpec(const std::vector<WorkerAlloc>& worker_alloc) {
  for (const auto& a : worker_alloc) {
    MINIPS_CHECK(!node_to_workers_.count(a.node_id)) << "node " << a.node_id << " listed twice in worker_alloc";
    for (uint32_t i = 0; i < a.num_workers; ++i) {
      worker_to_node_[num_workers_] = a.node_id;
      node_to_workers_[a.node_id].push_back(num_workers_);
      ++num_workers_;
    }
  }
}

bool WorkerSpec::HasLocalWorkers(uint32_t node_id) const { return node_to_workers_.count(node_id) > 0; }

const std::vector<uint32_t>& WorkerSpec::GetLocalWorkers(uint32_t node_id) const {
  auto it = node_to_workers_.find(node_id);
  MINIPS_CHECK(it != node_to_workers_.end()) << "no workers on node " << node_id;
  return it->second;
}

const std::vector<uint32_t>& WorkerSpec::GetLocalThreads(uint32_t node_id) const {
  auto it = node_to_threads_.find(node_id);
  MINIPS_CHECK(it != node_to_threads_.end()) << "no threads on node " << node_id;
  return it->second;
}

std::vector<uint32_t> WorkerSpec::GetAllThreadIds() const {
  return std::vector<uint32_t>(thread_ids_.begin(), thread_ids_.end());
}

void WorkerSpec::InsertWorkerIdThreadId(uint32_t worker_id, uint32_t thread_id) {
  MINIPS_CHECK(worker_to_node_.count(worker_id)) << "unknown worker " << worker_id;
  MINIPS_CHECK(!worker_to_thread_.count(worker_id)) << "worker " << worker_id << " already has a thread";
  MINIPS_CHECK(!thread_to_worker_.count(thread_id)) << "thread " << thread_id << " already used";
  worker_to_thread_[worker_id] = thread_id;
  thread_to_worker_[thread_id] = worker_id;
  node_to_threads_[worker_to_node_[worker_id]].push_back(thread_id);
  thread_ids_.insert(thread_id);
}

// ------------------------------------------------------------------ Engine
Engine::Engine(const Node& node, const std::vector<Node>& nodes, const Node& master_node, const Node& scale_node)
    : node_(node), nodes_(nodes), master_node_(master_node), scale_node_(scale_node) {}

Engine::~Engine() {
  try {
    StopHeartbeatThread();
  } catch (...) {
  }
}

void Engine::StartEverything(int num_server_threads_per_node) {
  CreateIdMapper(num_server_threads_per_node);
  CreateMailbox();
  StartSender();
  StartServerThreads();
  StartWorkerThreads();
  StartMailbox();
  StartHeartbeatThread();
}

void Engine::StopEverything() {
  StopHeartbeatThread();  // queues the goodbye heartbeat on the sender
  // drains the sender queue (the goodbye must leave before the sockets close; engine.hpp:174-180)
  StopSender();
  StopMailbox();          // barrier, stop receiver, kExit into every registered queue (servers, helper)
  StopServerThreads();
  StopWorkerThreads();
}

void Engine::CreateIdMapper(int num_server_threads_per_node) {
  num_server_threads_per_node_ = num_server_threads_per_node;
  id_mapper_.reset(new SimpleIdMapper(node_, nodes_));
  id_mapper_->Init(num_server_threads_per_node);
}

void Engine::CreateMailbox() {
  mailbox_.reset(new Mailbox(node_, nodes_, id_mapper_.get(), this, master_node_.is_master ? &master_node_ : nullptr));
}

void Engine::StartMailbox() { mailbox_->Start(); }

void Engine::StartSender() {
  sender_.reset(new Sender(mailbox_.get()));
  sender_->Start();
}

void Engine::StartServerThreads() {
  for (uint32_t tid : id_mapper_->GetServerThreadsForId(node_.id)) {
    std::unique_ptr<ServerThread> st(new ServerThread(tid));
    mailbox_->RegisterQueue(tid, st->GetWorkQueue());
    st->Start();
    server_threads_.push_back(std::move(st));
  }
}

void Engine::StartWorkerThreads() {
  const auto helpers = id_mapper_->GetWorkerHelperThreadsForId(node_.id);
  MINIPS_CHECK_EQ(helpers.size(), 1u);
  worker_thread_.reset(new WorkerThread(helpers[0]));
  mailbox_->RegisterQueue(helpers[0], worker_thread_->GetWorkQueue());
  worker_thread_->Start();
}

void Engine::StartHeartbeatThread() {
  const int interval = Context::get_instance().get_int32("heartbeat_interval");
  if (!HasMaster() || interval <= 0) return;
  heartbeat_running_ = true;
  heartbeat_thread_ = std::thread([this, interval] {
    const bool scale = Context::get_instance().get_bool("scale") && scale_node_.id == node_.id &&
                       !scale_node_.hostname.empty();
    if (scale) Scale(node_.id);
    while (heartbeat_running_) {
      HeartBeat(node_.id);
      for (int i = 0; i < interval * 10 && heartbeat_running_; ++i)
        std::this_thread::sleep_for(std::chrono::milliseconds(100));
    }
    if (heartbeat_goodbye_) HeartBeat(node_.id, /*quit=*/true);
  });
}

void Engine::StopHeartbeatThread() {
  heartbeat_running_ = false;
  if (heartbeat_thread_.joinable()) heartbeat_thread_.join();
}

void Engine::AbandonHeartbeatThread() {
  heartbeat_goodbye_ = false;
  StopHeartbeatThread();
}

void Engine::StopServerThreads() {
  for (auto& s : server_threads_) s->Stop();
  server_threads_.clear();
}
void Engine::StopWorkerThreads() {
  if (worker_thread_) worker_thread_->Stop();
  worker_thread_.reset();
}
void Engine::StopSender() {
  if (sender_) sender_->Stop();
}
void Engine::StopMailbox() {
  if (mailbox_) mailbox_->Stop(true);
}

std::vector<Range> Engine::getRanges() {
  auto& ctx = Context::get_instance();
  const uint64_t num_dims = static_cast<uint64_t>(ctx.get_int64("num_dims"));
  const uint64_t shards = id_mapper_ ? id_mapper_->GetAllServerThreads().size()
                                     : nodes_.size() * std::max(1, ctx.get_int32("num_servers_per_node"));
  MINIPS_CHECK(shards > 0);
  std::vector<Range> ranges;
  const uint64_t w = num_dims / shards;
  for (uint64_t i = 0; i + 1 < shards; ++i) ranges.emplace_back(w * i, w * (i + 1));
  ranges.emplace_back(w * (shards - 1), num_dims);
  return ranges;
}

std::unique_ptr<AbstractModel> Engine::MakeModel(ModelType type, uint32_t model_id,
                                                 std::unique_ptr<AbstractStorage> st, int staleness) {
  auto* q = sender_->GetMessageQueue();
  switch (type) {
    case ModelType::SSP: return std::unique_ptr<AbstractModel>(new SSPModel(model_id, std::move(st), staleness, q));
    case ModelType::BSP: return std::unique_ptr<AbstractModel>(new BSPModel(model_id, std::move(st), q));
    case ModelType::ASP: return std::unique_ptr<AbstractModel>(new ASPModel(model_id, std::move(st), q));
  }
  MINIPS_CHECK(false) << "unknown model type";
  return nullptr;
}

void Engine::RegisterPartitionManager(uint32_t table_id, std::unique_ptr<AbstractPartitionManager> pm) {
  MINIPS_CHECK(!partition_managers_.count(table_id));
  partition_managers_[table_id] = std::move(pm);
}

AbstractPartitionManager* Engine::GetPartitionManager(uint32_t table_id) {
  auto it = partition_managers_.find(table_id);
  MINIPS_CHECK(it != partition_managers_.end()) << "unknown table " << table_id;
  return it->second.get();
}

std::vector<ServerThread*> Engine::GetServerThreads() {
  std::vector<ServerThread*> v;
  for (auto& s : server_threads_) v.push_back(s.get());
  return v;
}

void Engine::InitTable(uint32_t table_id, const std::vector<uint32_t>& worker_ids) {
  const auto local_servers = id_mapper_->GetServerThreadsForId(node_.id);
  if (local_servers.empty()) return;
  // a temporary mailbox identity collects the acks
  const uint32_t tmp_tid = id_mapper_->AllocateWorkerThread(node_.id);
  ThreadsafeQueue<Message> acks;
  mailbox_->RegisterQueue(tmp_tid, &acks);
  for (uint32_t server : local_servers) {
    Message m;
    m.meta.flag = Flag::kResetWorkerInModel;
    m.meta.model_id = static_cast<int>(table_id);
    m.meta.sender = static_cast<int>(tmp_tid);
    m.meta.recver = static_cast<int>(server);
    m.AddData(SArray<uint32_t>(worker_ids));
    sender_->GetMessageQueue()->Push(m);
  }
  for (size_t i = 0; i < local_servers.size(); ++i) {
    Message ack;
    acks.WaitAndPop(&ack);
    MINIPS_CHECK(ack.meta.flag == Flag::kResetWorkerInModel && ack.meta.model_id == static_cast<int>(table_id));
  }
  mailbox_->DeregisterQueue(tmp_tid);
  id_mapper_->DeallocateWorkerThread(node_.id, tmp_tid);
}

void Engine::Barrier() { mailbox_->Barrier(); }

WorkerSpec Engine::AllocateWorkers(const std::vector<WorkerAlloc>& worker_alloc) {
  WorkerSpec spec(worker_alloc);
  for (const auto& kv : spec.GetNodeToWorkers())
    for (uint32_t w : kv.second) spec.InsertWorkerIdThreadId(w, id_mapper_->AllocateWorkerThread(kv.first));
  return spec;
}

void Engine::Run(const MLTask& task) {
  MINIPS_CHECK(task.IsSetup()) << "MLTask needs a lambda, a worker allocation and tables";
  WorkerSpec spec = AllocateWorkers(task.GetWorkerAlloc());
  for (uint32_t table : task.GetTables()) InitTable(table, spec.GetAllThreadIds());
  if (!Context::get_instance().get_bool("use_weight_file")) Barrier();

  std::vector<std::thread> threads;
  std::vector<std::string> errors;
  std::mutex err_mu;
  if (spec.HasLocalWorkers(node_.id)) {
    const auto& workers = spec.GetLocalWorkers(node_.id);
    const auto& tids = spec.GetLocalThreads(node_.id);
    for (size_t i = 0; i < workers.size(); ++i) {
      // replies for every user thread of this node land in the worker helper's queue
      mailbox_->RegisterQueue(tids[i], worker_thread_->GetWorkQueue());
      Info info;
      info.thread_id = tids[i];
      info.worker_id = workers[i];
      info.send_queue = sender_->GetMessageQueue();
      info.callback_runner = worker_thread_.get();
      info.engine = this;
      for (uint32_t table : task.GetTables()) info.partition_manager_map[table] = GetPartitionManager(table);
      threads.emplace_back([&task, info, &errors, &err_mu] {
        try {
          task.RunLambda(info);
        } catch (const std::exception& e) {
          std::lock_guard<std::mutex> lk(err_mu);
          errors.push_back(e.what());
        }
      });
    }
  }
  for (auto& t : threads) t.join();
  if (spec.HasLocalWorkers(node_.id)) {
    for (uint32_t tid : spec.GetLocalThreads(node_.id)) {
      mailbox_->DeregisterQueue(tid);
      id_mapper_->DeallocateWorkerThread(node_.id, tid);
    }
  }
  // remote workers' thread ids were reserved in this node's mapper as well
  for (const auto& kv : spec.GetNodeToWorkers())
    if (kv.first != node_.id)
      for (uint32_t w : kv.second) id_mapper_->DeallocateWorkerThread(kv.first, spec.GetWorkerToThreadMapper().at(w));
  Barrier();  // let in-flight messages land (driver/engine.cpp:295-296)
  MINIPS_CHECK(errors.empty()) << "worker lambda failed: " << errors.front();
}

void Engine::ForceQuit() { mailbox_->ForceQuit(node_.id); }

void Engine::HeartBeat(uint32_t node_id, bool quit) {
  if (!HasMaster()) return;
  Message m;
  m.meta.flag = quit ? Flag::kQuitHeartBeat : Flag::kHeartBeat;
  m.meta.sender = static_cast<int>(node_id);
  m.meta.recver = static_cast<int>(master_node_.id);
  sender_->GetMessageQueue()->Push(m);
}

void Engine::Scale(uint32_t node_id) {
  if (!HasMaster()) return;
  Message m;
  m.meta.flag = Flag::kScale;
  m.meta.sender = static_cast<int>(node_id);
  m.meta.recver = static_cast<int>(master_node_.id);
  sender_->GetMessageQueue()->Push(m);
}

void Engine::RollBack(uint32_t failed_node_id) {
  for (const auto& n : nodes_) {
    Message m;
    m.meta.flag = Flag::kRollBack;
    m.meta.sender = static_cast<int>(failed_node_id);
    m.meta.failed_node_id = static_cast<int>(failed_node_id);
    m.meta.recver = static_cast<int>(n.id);
    sender_->GetMessageQueue()->Push(m);
  }
}

void Engine::ScaleRollBack(uint32_t scale_node_id) {
  for (const auto& n : nodes_) {
    Message m;
    m.meta.flag = Flag::kScaleRollback;
    m.meta.sender = static_cast<int>(scale_node_id);
    m.meta.failed_node_id = static_cast<int>(scale_node_id);
    m.meta.recver = static_cast<int>(n.id);
    sender_->GetMessageQueue()->Push(m);
  }
}

int Engine::IncRollBackCount() {
  std::lock_guard<std::mutex> lk(recover_mu_);
  return ++rollback_counter_;
}

void Engine::RecoverEnd() {
  {
    std::lock_guard<std::mutex> lk(recover_mu_);
    recover_end_ = true;
    need_rollback_ = false;
    rollback_counter_ = 0;
  }
  recover_cv_.notify_all();
}

void Engine::WaitRecover() {
  std::unique_lock<std::mutex> lk(recover_mu_);
  recover_cv_.wait(lk, [this] { return recover_end_ || !need_rollback_; });
}

void Engine::RollBackServer() {
  for (auto& s : server_threads_) s->RollbackModel();
}

void Engine::RollBackWorker() {
  if (worker_thread_) worker_thread_->RollBackWorker();
}

void Engine::UpdateNodes(const std::vector<Node>& nodes) {
  nodes_ = nodes;
  if (id_mapper_) id_mapper_->Update(nodes_, num_server_threads_per_node_);
}

void Engine::UpdateAndRestart(int failed_node_id, const std::vector<Node>& nodes) {
  UpdateNodes(nodes);
  const auto ranges = getRanges();
  const auto tids = id_mapper_->GetAllServerThreads();
  for (auto& kv : partition_managers_) kv.second->Update(ranges, tids);
  for (size_t i = 0; i < server_threads_.size(); ++i) {
    Range r;
    for (size_t j = 0; j < tids.size(); ++j)
      if (tids[j] == server_threads_[i]->GetId()) r = ranges[j];
    server_threads_[i]->UpdateModel(failed_node_id, nodes_, r);
  }
  if (worker_thread_) worker_thread_->Update(static_cast<uint32_t>(num_server_threads_per_node_));
  if (restarter_) restarter_();
}

void Engine::OnForceQuit(uint32_t node_id) {
  std::vector<Node> left;
  for (const auto& n : nodes_)
    if (n.id != node_id) left.push_back(n);
  UpdateNodes(left);
}

void Engine::OnRollBack(int failed_node_id) {
  // the socket to the failed node died with it; the relaunched process listens on the same address
  // (ZMQ DEALER sockets of the reference reconnect by themselves)
  if (mailbox_ && failed_node_id >= 0 && static_cast<uint32_t>(failed_node_id) != node_.id)
    for (const auto& n : nodes_)
      if (n.id == static_cast<uint32_t>(failed_node_id)) mailbox_->ConnectTo(n);
  ++rollback_signals_;
  // every node returns to the last checkpoint (comm/mailbox.cpp:172-190)
  const auto iters = checkpoint::LoadConfigData();
  Context::get_instance().update_iteration(iters);
  rollback_iteration_ = iters.empty() ? 0 : iters.begin()->second;
  {
    std::lock_guard<std::mutex> lk(recover_mu_);
    recover_end_ = false;
  }
  SetNeedRollBack(true);
  RollBackServer();
  RollBackWorker();
}

void Engine::OnScaleRollBack(int scale_node_id) {
  (void)scale_node_id;
  if (mailbox_ && !scale_node_.hostname.empty()) {
    mailbox_->ConnectTo(scale_node_);
    mailbox_->SetHasScaleNode(true);
  }
  OnRollBack(-1);
}

}  // namespace minips
