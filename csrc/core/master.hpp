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

class Master {
 public:
  Master(const Node& master_node, const std::vector<Node>& nodes);
  ~Master();
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
  bool started_ = false;
};

}  // namespace minips
