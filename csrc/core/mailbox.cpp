#include "core/mailbox.hpp"

#include <arpa/inet.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/uio.h>
#include <unistd.h>

#include <cerrno>
#include <chrono>

namespace minips {

namespace {

constexpr uint32_t kMagic = 0x4d505331;  // "MPS1"

struct WireHeader {
  uint32_t magic;
  int32_t sender, recver, model_id, failed_node_id, flag;
  uint32_t nframes;
};

bool WriteAll(int fd, const void* buf, size_t n) {
  const char* p = static_cast<const char*>(buf);
  while (n > 0) {
    ssize_t w = ::send(fd, p, n, MSG_NOSIGNAL);
    if (w < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    p += w;
    n -= static_cast<size_t>(w);
  }
  return true;
}

bool ReadAll(int fd, void* buf, size_t n) {
  char* p = static_cast<char*>(buf);
  while (n > 0) {
    ssize_t r = ::recv(fd, p, n, 0);
    if (r == 0) return false;
    if (r < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    p += r;
    n -= static_cast<size_t>(r);
  }
  return true;
}

}  // namespace

Mailbox::Mailbox(const Node& node, const std::vector<Node>& nodes, AbstractIdMapper* id_mapper,
                 MailboxListener* listener, const Node* master_node)
    : node_(node), nodes_(nodes), id_mapper_(id_mapper), listener_(listener) {
  if (master_node && master_node->is_master) {
    master_node_ = *master_node;
    has_master_ = true;
  }
}

Mailbox::~Mailbox() {
  if (receiving_) {
    try {
      StopReceiving();
    } catch (...) {
    }
  }
  CloseSockets();
}

void Mailbox::RegisterQueue(uint32_t queue_id, ThreadsafeQueue<Message>* queue) {
  std::lock_guard<std::mutex> lk(queue_mu_);
  MINIPS_CHECK(queue_map_.find(queue_id) == queue_map_.end()) << "queue " << queue_id << " registered twice";
  queue_map_[queue_id] = queue;
}

void Mailbox::DeregisterQueue(uint32_t queue_id) {
  std::lock_guard<std::mutex> lk(queue_mu_);
  MINIPS_CHECK(queue_map_.find(queue_id) != queue_map_.end()) << "queue " << queue_id << " not registered";
  queue_map_.erase(queue_id);
}

size_t Mailbox::GetQueueMapSize() const {
  std::lock_guard<std::mutex> lk(queue_mu_);
  return queue_map_.size();
}

void Mailbox::BindAndListen() {
  listen_fd_ = ::socket(AF_INET, SOCK_STREAM, 0);
  MINIPS_CHECK(listen_fd_ >= 0) << "socket(): " << strerror(errno);
  int one = 1;
  ::setsockopt(listen_fd_, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_addr.s_addr = htonl(INADDR_ANY);
  addr.sin_port = htons(static_cast<uint16_t>(node_.port));
  MINIPS_CHECK(::bind(listen_fd_, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) == 0)
      << "bind to *:" << node_.port << " failed: " << strerror(errno);
  MINIPS_CHECK(::listen(listen_fd_, 256) == 0) << "listen: " << strerror(errno);
  MINIPS_CHECK(::pipe(wake_pipe_) == 0) << "pipe: " << strerror(errno);
}

int Mailbox::ConnectWithRetry(const Node& n) {
  addrinfo hints{};
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  addrinfo* res = nullptr;
  const std::string host = (n.hostname == "localhost") ? "127.0.0.1" : n.hostname;
  const std::string port = std::to_string(n.port);
  MINIPS_CHECK(::getaddrinfo(host.c_str(), port.c_str(), &hints, &res) == 0 && res)
      << "cannot resolve " << n.hostname;
  int fd = -1;
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(60);
  while (true) {
    fd = ::socket(AF_INET, SOCK_STREAM, 0);
    MINIPS_CHECK(fd >= 0) << "socket(): " << strerror(errno);
    if (::connect(fd, res->ai_addr, res->ai_addrlen) == 0) break;
    ::close(fd);
    fd = -1;
    if (std::chrono::steady_clock::now() > deadline) break;
    std::this_thread::sleep_for(std::chrono::milliseconds(20));
  }
  ::freeaddrinfo(res);
  MINIPS_CHECK(fd >= 0) << "connect to node " << n.id << " (" << n.hostname << ":" << n.port << ") timed out";
  int one = 1;
  ::setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
  return fd;
}

void Mailbox::ConnectTo(const Node& n) {
  int fd = ConnectWithRetry(n);
  std::lock_guard<std::mutex> lk(send_mu_);
  auto it = senders_.find(n.id);
  if (it != senders_.end()) ::close(it->second);
  senders_[n.id] = fd;
}

void Mailbox::ConnectAll() {
  // every node, including this one: local traffic takes the same path as remote traffic
  // (comm/mailbox.cpp:51-53)
  for (const auto& n : nodes_) ConnectTo(n);
  if (has_master_ && !HasNodeId(master_node_.id)) ConnectTo(master_node_);
}

bool Mailbox::HasNodeId(uint32_t id) const {
  for (const auto& n : nodes_)
    if (n.id == id) return true;
  return false;
}

void Mailbox::StartReceiving() {
  receiving_ = true;
  receiver_thread_ = std::thread([this] { Receiving(); });
}

void Mailbox::StopReceiving() {
  if (!receiving_) return;
  // a local kExit stops the receive loop, exactly like the reference's self-addressed message
  Message exit_msg;
  exit_msg.meta.flag = Flag::kExit;
  exit_msg.meta.sender = static_cast<int>(node_.id);
  exit_msg.meta.recver = static_cast<int>(node_.id);
  bool sent = false;
  try {
    sent = SendTo(node_.id, exit_msg) == 0;
  } catch (...) {
  }
  if (!sent && wake_pipe_[1] >= 0) {
    char c = 'x';
    (void)!::write(wake_pipe_[1], &c, 1);
  }
  if (receiver_thread_.joinable()) receiver_thread_.join();
  receiving_ = false;
}

void Mailbox::Start() {
  BindAndListen();
  StartReceiving();
  ConnectAll();
}

void Mailbox::Stop(bool barrier) {
  if (barrier) Barrier();
  StopReceiving();
  {
    std::lock_guard<std::mutex> lk(queue_mu_);
    for (auto& kv : queue_map_) {
      Message m;
      m.meta.flag = Flag::kExit;
      m.meta.recver = static_cast<int>(kv.first);
      kv.second->Push(m);
    }
  }
  CloseSockets();
}

void Mailbox::CloseSockets() {
  std::lock_guard<std::mutex> lk(send_mu_);
  for (auto& kv : senders_) ::close(kv.second);
  senders_.clear();
  if (listen_fd_ >= 0) {
    ::close(listen_fd_);
    listen_fd_ = -1;
  }
  for (int i = 0; i < 2; ++i)
    if (wake_pipe_[i] >= 0) {
      ::close(wake_pipe_[i]);
      wake_pipe_[i] = -1;
    }
}

int Mailbox::Send(const Message& msg) {
  // data flags address a thread id, control flags a node id (comm/mailbox.cpp:235-243)
  uint32_t node_id;
  if (IsControlFlag(msg.meta.flag)) node_id = static_cast<uint32_t>(msg.meta.recver);
  else node_id = id_mapper_->GetNodeIdForThread(static_cast<uint32_t>(msg.meta.recver));
  return SendTo(node_id, msg);
}

int Mailbox::SendTo(uint32_t node_id, const Message& msg) {
  std::lock_guard<std::mutex> lk(send_mu_);
  auto it = senders_.find(node_id);
  if (it == senders_.end()) {
    MINIPS_LOG(0) << "Mailbox::Send: no connection to node " << node_id << " for " << msg.DebugString();
    return -1;
  }
  WireHeader h{kMagic,
               msg.meta.sender,
               msg.meta.recver,
               msg.meta.model_id,
               msg.meta.failed_node_id,
               static_cast<int32_t>(msg.meta.flag),
               static_cast<uint32_t>(msg.data.size())};
  std::vector<uint64_t> sizes(msg.data.size());
  for (size_t i = 0; i < msg.data.size(); ++i) sizes[i] = msg.data[i].size();
  const int fd = it->second;
  bool ok = WriteAll(fd, &h, sizeof(h));
  if (ok && !sizes.empty()) ok = WriteAll(fd, sizes.data(), sizes.size() * sizeof(uint64_t));
  for (size_t i = 0; ok && i < msg.data.size(); ++i)
    if (sizes[i]) ok = WriteAll(fd, msg.data[i].data(), sizes[i]);  // payload goes out straight from the SArray
  return ok ? 0 : -1;
}

bool Mailbox::RecvMessage(int fd, Message* msg) {
  WireHeader h;
  if (!ReadAll(fd, &h, sizeof(h))) return false;
  if (h.magic != kMagic) return false;
  msg->meta.sender = h.sender;
  msg->meta.recver = h.recver;
  msg->meta.model_id = h.model_id;
  msg->meta.failed_node_id = h.failed_node_id;  // (the reference drops this field on receive)
  msg->meta.flag = static_cast<Flag>(h.flag);
  msg->data.clear();
  std::vector<uint64_t> sizes(h.nframes);
  if (h.nframes && !ReadAll(fd, sizes.data(), sizes.size() * sizeof(uint64_t))) return false;
  for (uint32_t i = 0; i < h.nframes; ++i) {
    char* buf = sizes[i] ? new char[sizes[i]] : nullptr;
    if (sizes[i] && !ReadAll(fd, buf, sizes[i])) {
      delete[] buf;
      return false;
    }
    if (buf) msg->data.emplace_back(buf, sizes[i], [](char* p) { delete[] p; });
    else msg->data.emplace_back();
  }
  return true;
}

void Mailbox::Dispatch(Message& msg) {
  switch (msg.meta.flag) {
    case Flag::kBarrier: {
      std::lock_guard<std::mutex> lk(barrier_mu_);
      ++barrier_count_;
      barrier_cv_.notify_all();
      return;
    }
    case Flag::kForceQuit: {
      const uint32_t gone = static_cast<uint32_t>(msg.meta.sender);
      bool removed = false;
      {
        // nodes_ is read by Barrier() on user threads: it changes only under barrier_mu_
        std::lock_guard<std::mutex> lk(barrier_mu_);
        for (auto it = nodes_.begin(); it != nodes_.end(); ++it)
          if (it->id == gone) {
            nodes_.erase(it);
            removed = true;
            break;
          }
        if (removed) barrier_cv_.notify_all();  // the barrier target just shrank
      }
      if (removed) {
        if (listener_) listener_->OnForceQuit(gone);
      }
      return;
    }
    case Flag::kRollBack:
      if (listener_) listener_->OnRollBack(msg.meta.failed_node_id >= 0 ? msg.meta.failed_node_id : msg.meta.sender);
      return;
    case Flag::kScaleRollback:
      if (listener_) listener_->OnScaleRollBack(msg.meta.failed_node_id >= 0 ? msg.meta.failed_node_id
          : msg.meta.sender);
      return;
    case Flag::kCheckpoint: {
      // requests travel worker -> server thread; acks travel back to a user thread id
      const uint32_t local_tid = static_cast<uint32_t>(msg.meta.recver) % 1000;
      if (listener_ && local_tid < 50) listener_->OnCheckpointRequest();
      break;
    }
    default:
      break;
  }
  ThreadsafeQueue<Message>* q = nullptr;
  {
    std::lock_guard<std::mutex> lk(queue_mu_);
    auto it = queue_map_.find(static_cast<uint32_t>(msg.meta.recver));
    if (it != queue_map_.end()) q = it->second;
  }
  if (q) q->Push(std::move(msg));
  else MINIPS_LOG(0) << "Mailbox: no queue for recver " << msg.meta.recver << ", dropped " << msg.DebugString();
}

void Mailbox::Receiving() {
  std::vector<int> conns;
  bool running = true;
  while (running) {
    std::vector<pollfd> fds;
    fds.push_back({listen_fd_, POLLIN, 0});
    fds.push_back({wake_pipe_[0], POLLIN, 0});
    for (int c : conns) fds.push_back({c, POLLIN, 0});
    const int rc = ::poll(fds.data(), fds.size(), 1000);
    if (rc < 0) {
      if (errno == EINTR) continue;
      break;
    }
    if (rc == 0) continue;
    if (fds[1].revents & POLLIN) break;
    if (fds[0].revents & POLLIN) {
      const int c = ::accept(listen_fd_, nullptr, nullptr);
      if (c >= 0) {
        int one = 1;
        ::setsockopt(c, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
        conns.push_back(c);
      }
    }
    for (size_t i = 2; i < fds.size() && running; ++i) {
      if (!(fds[i].revents & (POLLIN | POLLHUP | POLLERR))) continue;
      Message msg;
      if (!RecvMessage(fds[i].fd, &msg)) {
        ::close(fds[i].fd);
        for (auto it = conns.begin(); it != conns.end(); ++it)
          if (*it == fds[i].fd) {
            conns.erase(it);
            break;
          }
        continue;
      }
      if (msg.meta.flag == Flag::kExit) {
        running = false;
        break;
      }
      Dispatch(msg);
    }
  }
  for (int c : conns) ::close(c);
}

void Mailbox::Barrier() {
  Message m;
  m.meta.flag = Flag::kBarrier;
  m.meta.sender = static_cast<int>(node_.id);
  std::vector<Node> peers;
  {
    std::lock_guard<std::mutex> lk(barrier_mu_);
    peers = nodes_;
  }
  for (const auto& n : peers) {
    m.meta.recver = static_cast<int>(n.id);
    SendTo(n.id, m);
  }
  std::unique_lock<std::mutex> lk(barrier_mu_);
  // nodes_ may shrink while we wait (ForceQuit); re-evaluate the target on every wake-up
  barrier_cv_.wait(lk, [this] { return barrier_count_ >= static_cast<int>(nodes_.size()) + (has_scale_node_ ? 1
      : 0); });
  barrier_count_ -= static_cast<int>(nodes_.size()) + (has_scale_node_ ? 1 : 0);
  if (barrier_count_ < 0) barrier_count_ = 0;
}

void Mailbox::ForceQuit(uint32_t node_id) {
  Message m;
  m.meta.flag = Flag::kForceQuit;
  m.meta.sender = static_cast<int>(node_id);
  std::vector<Node> peers;
  {
    std::lock_guard<std::mutex> lk(barrier_mu_);
    peers = nodes_;
  }
  for (const auto& n : peers) {
    m.meta.recver = static_cast<int>(n.id);
    SendTo(n.id, m);
  }
}

void Mailbox::UpdateNodes(const std::vector<Node>& nodes) {
  std::lock_guard<std::mutex> lk(barrier_mu_);
  nodes_ = nodes;
  barrier_cv_.notify_all();
}

// ------------------------------------------------------------------ Sender
void Sender::Start() {
  th_ = std::thread([this] { Run(); });
}

void Sender::Run() {
  while (true) {
    Message m;
    send_queue_.WaitAndPop(&m);
    if (m.meta.flag == Flag::kExit) break;
    mailbox_->Send(m);
  }
}

void Sender::Stop() {
  if (!th_.joinable()) return;
  Message m;
  m.meta.flag = Flag::kExit;
  send_queue_.Push(m);
  th_.join();
}

}  // namespace minips
