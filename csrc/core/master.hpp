// master.hpp — failure detection and recovery coordination (host control plane).
//
//   Master                master/master.hpp:14-87   runs on the master node (id 1): only id mapper,
//                         mailbox and sender of an Engine, plus the two actors below
//   MasterThread          master/master_thread.{hpp,cpp}: records heartbeats, triggers RollBack when
//                         the recovering node reports back, ScaleRollBack on kScale, exits when every
//                         node sent kQuitHeartBeat
//   HeartBeatCheckThread  master/heartbeat_check_thread.{hpp,cpp}: a node silent for more than
//                         3 x heartbeat_interval is declared failed and relaunched (relaunch_cmd + id)
#pragma once

#include <atomic>
#include <ctime>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "core/base.hpp"
#include "core/driver.hpp"

namespace minips {

class Master;

class MasterThread : public Actor {
 public:
  MasterThread(uint32_t id, Master* master) : Actor(id), master_(master) {}
  void Init(const std::vector<Node>& nodes);
  std::map<uint32_t, time_t> GetHeartBeats();
  // a node joined the cluster (scale-out): its heartbeats are tracked and its kQuitHeartBeat is awaited
  void AddNode(uint32_t id);
  void SetRecoveringNodeId(int id) { recovering_node_id_ = id; }
  int GetRecoveringNodeId() const { return recovering_node_id_; }
  bool AllQuit() const { return all_quit_; }
  // exposed for tests: handle one message as the actor would
  void Handle(Message& msg);

 protected:
  void Main() override;

 private:
  Master* master_;
  std::mutex mu_;
  std::map<uint32_t, time_t> heartbeats_;
  size_t serving_nodes_ = 0;
  size_t quit_count_ = 0;
  std::atomic<int> recovering_node_id_{-1};
  std::atomic<bool> all_quit_{false};
};

class HeartBeatCheckThread {
 public:
  HeartBeatCheckThread(MasterThread* master_thread, int interval_s, std::string relaunch_cmd,
                       std::function<int(const std::string&)> launcher = nullptr)
      : master_thread_(master_thread), interval_(interval_s), relaunch_cmd_(std::move(relaunch_cmd)),
        launcher_(std::move(launcher)) {}
  ~HeartBeatCheckThread() { Stop(); }
  void Start();
  void Stop();
  // one scan: returns the ids declared failed in this pass (tests call it directly)
  std::vector<uint32_t> CheckOnce(time_t now);

 private:
  MasterThread* master_thread_;
  int interval_;
  std::string relaunch_cmd_;
  std::function<int(const std::string&)> launcher_;
  std::thread th_;
  std::atomic<bool> running_{false};
};

// ControlAgent — the host control plane of a rank whose data plane lives elsewhere (the device backend:
// shards in HBM, Get / Add / Clock as kernels).  It is the part of an Engine such a rank still needs:
// the mailbox endpoint, the heartbeat to the master (driver/engine.cpp:114-148) and the rollback /
// force-quit notifications (comm/mailbox.cpp:159-190), without server or worker threads.
class ControlAgent {
 public:
  ControlAgent(const Node& node, const std::vector<Node>& nodes, const Node& master_node);
  ~ControlAgent();
  void Start();                       // mailbox up, connected to every node and the master, heartbeat running
  void Stop(bool quit_heartbeat = true);
  bool NeedRollBack() const { return engine_ && engine_->IsNeedRollBack(); }
  void ClearRollBack() { if (engine_) engine_->SetNeedRollBack(false); }
  int RollBackCount() const { return engine_ ? engine_->GetRollBackSignals() : 0; }
  // announce this node to the master as a scale-out node (driver/engine.hpp:81-90; the node first wrote
  // its id:host:port into the scale file, lr_example.cpp:396-409)
  void Scale();
  Engine* GetEngine() { return engine_.get(); }

 private:
  Node node_, master_node_;
  std::vector<Node> nodes_;
  std::unique_ptr<Engine> engine_;
  bool started_ = false;
};

class Master {
 public:
  Master(const Node& master_node, const std::vector<Node>& nodes);
  ~Master();
  // called on the master thread right before kRollBack is broadcast (the device backend bumps the
  // rendezvous generation here so that every rank that receives the message finds the new one)
  void SetRollBackHook(std::function<void(uint32_t)> f) { rollback_hook_ = std::move(f); }
  // called on the master thread when a node announced itself with kScale, after the node was added to the
  // master's node list and before kScaleRollback is broadcast
  void SetScaleHook(std::function<void(uint32_t)> f) { scale_hook_ = std::move(f); }
  void StartMaster();                 // bring-up: id mapper + mailbox + sender + the two actors
  void StopMaster();                  // blocks until every node quit (or Shutdown was called)
  void Shutdown();                    // tear down without waiting for the nodes
  void RollBack(uint32_t failed_node_id);
  void ScaleRollBack(uint32_t scale_node_id);
  MasterThread* GetMasterThread() { return master_thread_.get(); }
  Engine* GetEngine() { return engine_.get(); }

 private:
  Node master_node_;
  std::vector<Node> nodes_;
  std::unique_ptr<Engine> engine_;
  std::unique_ptr<MasterThread> master_thread_;
  std::unique_ptr<HeartBeatCheckThread> heartbeat_check_;
  std::function<void(uint32_t)> rollback_hook_;
  std::function<void(uint32_t)> scale_hook_;
  bool started_ = false;
};

}  // namespace minips
