#include "core/master.hpp"

#include <chrono>
#include <cstdlib>

#include "core/lib.hpp"

namespace minips {

// ------------------------------------------------------------------ MasterThread
void MasterThread::Init(const std::vector<Node>& nodes) {
  std::lock_guard<std::mutex> lk(mu_);
  heartbeats_.clear();
  const time_t now = std::time(nullptr);
  for (const auto& n : nodes) heartbeats_[n.id] = now;
  serving_nodes_ = nodes.size();
  quit_count_ = 0;
}

void MasterThread::AddNode(uint32_t id) {
  std::lock_guard<std::mutex> lk(mu_);
  if (heartbeats_.find(id) == heartbeats_.end()) ++serving_nodes_;
  heartbeats_[id] = std::time(nullptr);
}

std::map<uint32_t, time_t> MasterThread::GetHeartBeats() {
  std::lock_guard<std::mutex> lk(mu_);
  return heartbeats_;
}

void MasterThread::Handle(Message& msg) {
  const uint32_t sender = static_cast<uint32_t>(msg.meta.sender);
  switch (msg.meta.flag) {
    case Flag::kQuitHeartBeat: {
      bool all = false;
      {
        std::lock_guard<std::mutex> lk(mu_);
        heartbeats_.erase(sender);
        ++quit_count_;
        all = quit_count_ >= serving_nodes_;
      }
      if (all) all_quit_ = true;
      break;
    }
    case Flag::kHeartBeat: {
      {
        std::lock_guard<std::mutex> lk(mu_);
        heartbeats_[sender] = std::time(nullptr);
      }
      // the relaunched process's first heartbeat triggers the global rollback
      if (recovering_node_id_ == static_cast<int>(sender)) {
        recovering_node_id_ = -1;
        if (master_) master_->RollBack(sender);
      }
      break;
    }
    case Flag::kScale:
      if (master_) master_->ScaleRollBack(sender);
      break;
    default:
      break;
  }
}

void MasterThread::Main() {
  while (true) {
    Message msg;
    work_queue_.WaitAndPop(&msg);
    if (msg.meta.flag == Flag::kExit) break;
    Handle(msg);
    if (all_quit_) break;
  }
}

// ------------------------------------------------------------------ HeartBeatCheckThread
std::vector<uint32_t> HeartBeatCheckThread::CheckOnce(time_t now) {
  std::vector<uint32_t> failed;
  if (master_thread_->GetRecoveringNodeId() >= 0) return failed;  // one recovery at a time
  for (const auto& kv : master_thread_->GetHeartBeats()) {
    if (now - kv.second > 3 * interval_) {
      CheckFaultTolerance(2);
      const std::string cmd = relaunch_cmd_ + " " + std::to_string(kv.first);
      master_thread_->SetRecoveringNodeId(static_cast<int>(kv.first));
      if (launcher_) launcher_(cmd);
      else if (!relaunch_cmd_.empty()) {
        const int rc = std::system(cmd.c_str());
        if (rc != 0) MINIPS_LOG(0) << "relaunch command failed (" << rc << "): " << cmd;
      }
      failed.push_back(kv.first);
      break;
    }
  }
  return failed;
}

void HeartBeatCheckThread::Start() {
  if (interval_ <= 0) return;
  running_ = true;
  th_ = std::thread([this] {
    while (running_) {
      for (int i = 0; i < interval_ * 10 && running_; ++i)
        std::this_thread::sleep_for(std::chrono::milliseconds(100));
      if (!running_) break;
      CheckOnce(std::time(nullptr));
    }
  });
}

void HeartBeatCheckThread::Stop() {
  running_ = false;
  if (th_.joinable()) th_.join();
}

// ------------------------------------------------------------------ Master
Master::Master(const Node& master_node, const std::vector<Node>& nodes) : master_node_(master_node), nodes_(nodes) {
  master_node_.is_master = true;
}

Master::~Master() {
  try {
    Shutdown();
  } catch (...) {
  }
}

void Master::StartMaster() {
  // the master hosts no server / worker threads: only the transport (master.hpp:22-28)
  engine_.reset(new Engine(master_node_, nodes_, master_node_));
  engine_->CreateIdMapper();
  engine_->CreateMailbox();
  engine_->StartSender();
  master_thread_.reset(new MasterThread(master_node_.id, this));
  master_thread_->Init(nodes_);
  engine_->GetMailbox()->RegisterQueue(master_node_.id, master_thread_->GetWorkQueue());
  engine_->GetMailbox()->BindAndListen();
  engine_->GetMailbox()->StartReceiving();
  engine_->GetMailbox()->ConnectAll();
  master_thread_->Start();
  auto& ctx = Context::get_instance();
  heartbeat_check_.reset(new HeartBeatCheckThread(master_thread_.get(), ctx.get_int32("heartbeat_interval"),
                                                  ctx.get_string("relaunch_cmd")));
  heartbeat_check_->Start();
  started_ = true;
}

void Master::StopMaster() {
  if (!started_) return;
  master_thread_->Stop();  // returns when every node sent kQuitHeartBeat
  Shutdown();
}

void Master::Shutdown() {
  if (!started_) return;
  started_ = false;
  if (heartbeat_check_) heartbeat_check_->Stop();
  if (master_thread_ && !master_thread_->AllQuit()) {
    Message m;
    m.meta.flag = Flag::kExit;
    master_thread_->GetWorkQueue()->Push(m);
    master_thread_->Stop();
  }
  engine_->GetMailbox()->StopReceiving();
  engine_->StopSender();
  engine_->GetMailbox()->CloseSockets();
}

void Master::RollBack(uint32_t failed_node_id) {
  CheckFaultTolerance(4);
  // the socket to the failed node died with it; the relaunched process listens on the same address
  for (const auto& n : nodes_)
    if (n.id == failed_node_id) engine_->GetMailbox()->ConnectTo(n);
  if (rollback_hook_) rollback_hook_(failed_node_id);
  engine_->RollBack(failed_node_id);
}

void Master::ScaleRollBack(uint32_t scale_node_id) {
  // the new node announced its address in the scale file (lib/svm_dumper.hpp:95-126); the master joins
  // it to its node list so that kScaleRollback reaches it too and its heartbeats / goodbye are counted
  // (the reference's peers Connect() to it the same way, comm/mailbox.cpp:197-219)
  bool known = false;
  for (const auto& n : nodes_) known = known || n.id == scale_node_id;
  if (!known) {
    const std::string path = Context::get_instance().get_string("scale_file");
    if (!path.empty()) {
      const Node n = lib::SVMDumper::LoadScaleFile(path);
      if (n.id == scale_node_id) {
        nodes_.push_back(n);
        engine_->UpdateNodes(nodes_);
        engine_->GetMailbox()->ConnectTo(n);
        master_thread_->AddNode(n.id);
      } else {
        MINIPS_LOG(0) << "scale file " << path << " names node " << n.id << ", kScale came from " << scale_node_id;
      }
    }
  }
  if (scale_hook_) scale_hook_(scale_node_id);
  engine_->ScaleRollBack(scale_node_id);
}

// ------------------------------------------------------------------ ControlAgent
ControlAgent::ControlAgent(const Node& node, const std::vector<Node>& nodes, const Node& master_node)
    : node_(node), master_node_(master_node), nodes_(nodes) {
  master_node_.is_master = true;
}

ControlAgent::~ControlAgent() {
  try {
    Stop(false);
  } catch (...) {
  }
}

void ControlAgent::Start() {
  engine_.reset(new Engine(node_, nodes_, master_node_));
  engine_->CreateIdMapper();
  engine_->CreateMailbox();
  engine_->StartSender();
  engine_->GetMailbox()->BindAndListen();
  engine_->GetMailbox()->StartReceiving();
  engine_->GetMailbox()->ConnectAll();
  engine_->StartHeartbeatThread();
  started_ = true;
}

void ControlAgent::Scale() {
  if (engine_) engine_->Scale(node_.id);
}

void ControlAgent::Stop(bool quit_heartbeat) {
  if (!started_) return;
  started_ = false;
  if (quit_heartbeat) engine_->StopHeartbeatThread();  // its last message is kQuitHeartBeat
  else engine_->AbandonHeartbeatThread();
  engine_->StopSender();
  engine_->GetMailbox()->StopReceiving();
  engine_->GetMailbox()->CloseSockets();
}

}  // namespace minips
