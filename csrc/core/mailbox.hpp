// mailbox.hpp — control-plane transport of the host runtime.
//
// In the reference the Mailbox is both data plane and control plane: a ZeroMQ ROUTER bound on
// the node's port plus one DEALER per peer (comm/mailbox.hpp:22-56, comm/mailbox.cpp:40-412),
// a receiving thread that dispatches by flag, a barrier built from all-to-all kBarrier
// messages, and a Sender thread that drains a queue into Send (comm/sender.cpp:7-30).  In this
// framework the GPU data plane is peer memory; the mailbox remains as the control plane and
// as the transport of the CPU ("host threads") backend.  libzmq headers are not available,
// so the wire is plain TCP with the same framing idea: [header+Meta][frame sizes][frames].
//
// Layering: the reference's mailbox includes engine.hpp and calls up into the Engine; here
// the upcalls go through the MailboxListener interface.
#pragma once

#include <atomic>
#include <condition_variable>
#include <map>
#include <mutex>
#include <thread>
#include <unordered_map>
#include <vector>

#include "core/base.hpp"
#include "core/partition.hpp"

namespace minips {

class AbstractMailbox {
 public:
  virtual ~AbstractMailbox() = default;
  virtual int Send(const Message& msg) = 0;
};

class MailboxListener {
 public:
  virtual ~MailboxListener() = default;
  virtual void OnForceQuit(uint32_t /*node_id*/) {}
  virtual void OnRollBack(int /*failed_node_id*/) {}
  virtual void OnScaleRollBack(int /*scale_node_id*/) {}
  virtual void OnCheckpointRequest() {}  // before a kCheckpoint is forwarded to a server thread
};

class Mailbox : public AbstractMailbox {
 public:
  Mailbox(const Node& node, const std::vector<Node>& nodes, AbstractIdMapper* id_mapper,
          MailboxListener* listener = nullptr, const Node* master_node = nullptr);
  ~Mailbox() override;

  void RegisterQueue(uint32_t queue_id, ThreadsafeQueue<Message>* queue);
  void DeregisterQueue(uint32_t queue_id);
  size_t GetQueueMapSize() const;

  void Start();                   // bind, start receiving, connect to every node (incl. self / master)
  void Stop(bool barrier = true);  // [barrier] -> stop receiver -> kExit into every registered queue
  int Send(const Message& msg) override;
  void Barrier();
  void ForceQuit(uint32_t node_id);  // announce that `node_id` (this node) leaves before training
  void ConnectTo(const Node& n);      // scale-out: add a peer at run time
  void SetListener(MailboxListener* l) { listener_ = l; }
  void UpdateNodes(const std::vector<Node>& nodes);
  const std::vector<Node>& GetNodes() const { return nodes_; }
  void SetHasScaleNode(bool v) { has_scale_node_ = v; }

  // low-level pieces, exposed for the ported mailbox tests
  void BindAndListen();
  void StartReceiving();
  void StopReceiving();
  void ConnectAll();
  void CloseSockets();
  int GetBarrierCount() const { return barrier_count_; }

 private:
  void Receiving();
  void Dispatch(Message& msg);
  bool RecvMessage(int fd, Message* msg);
  int SendTo(uint32_t node_id, const Message& msg);
  int ConnectWithRetry(const Node& n);
  bool HasNodeId(uint32_t id) const;

  Node node_;
  std::vector<Node> nodes_;
  Node master_node_;
  bool has_master_ = false;
  bool has_scale_node_ = false;
  AbstractIdMapper* id_mapper_;
  MailboxListener* listener_;

  mutable std::mutex queue_mu_;
  std::map<uint32_t, ThreadsafeQueue<Message>*> queue_map_;

  std::mutex send_mu_;
  std::unordered_map<uint32_t, int> senders_;  // node id -> connected fd

  int listen_fd_ = -1;
  int wake_pipe_[2] = {-1, -1};
  std::thread receiver_thread_;
  std::atomic<bool> receiving_{false};

  std::mutex barrier_mu_;
  std::condition_variable barrier_cv_;
  int barrier_count_ = 0;
};

class AbstractSender {
 public:
  virtual ~AbstractSender() = default;
  virtual void Start() = 0;
  virtual void Stop() = 0;
  virtual ThreadsafeQueue<Message>* GetMessageQueue() = 0;
};

class Sender : public AbstractSender {
 public:
  explicit Sender(AbstractMailbox* mailbox) : mailbox_(mailbox) {}
  ~Sender() override { Stop(); }
  void Start() override;
  void Stop() override;  // pushes kExit and joins
  ThreadsafeQueue<Message>* GetMessageQueue() override { return &send_queue_; }

 private:
  void Run();
  AbstractMailbox* mailbox_;
  ThreadsafeQueue<Message> send_queue_;
  std::thread th_;
};

}  // namespace minips
