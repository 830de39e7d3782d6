#include "core/io.hpp"

#include <arpa/inet.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <cerrno>
#include <chrono>

#include "core/file_io.hpp"

namespace minips {

namespace {

bool WriteAll(int fd, const void* buf, size_t n) {
  const char* p = static_cast<const char*>(buf);
  while (n) {
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
  while (n) {
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
// frame: [int32 type][uint64 len][len bytes]
bool SendFrame(int fd, int type, const BinStream& body) {
  const int32_t t = type;
  const uint64_t n = body.size();
  return WriteAll(fd, &t, sizeof(t)) && WriteAll(fd, &n, sizeof(n)) && (n == 0 || WriteAll(fd,
      body.get_remained_buffer(), n));
}
bool RecvFrame(int fd, int* type, BinStream* body) {
  int32_t t = 0;
  uint64_t n = 0;
  if (!ReadAll(fd, &t, sizeof(t)) || !ReadAll(fd, &n, sizeof(n))) return false;
  std::vector<char> buf(n);
  if (n && !ReadAll(fd, buf.data(), n)) return false;
  *type = t;
  *body = BinStream(std::move(buf));
  return true;
}

}  // namespace

// ------------------------------------------------------------------ Coordinator
Coordinator::Coordinator(int proc_id, std::string hostname, std::string master_host, int master_port)
    : proc_id_(proc_id),
      hostname_(std::move(hostname)),
      master_host_(std::move(master_host)),
      master_port_(master_port) {}

Coordinator::~Coordinator() {
  if (fd_ >= 0) ::close(fd_);
}

void Coordinator::serve() {
  if (fd_ >= 0) return;
  addrinfo hints{};
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  addrinfo* res = nullptr;
  const std::string host = master_host_ == "localhost" ? "127.0.0.1" : master_host_;
  MINIPS_CHECK(::getaddrinfo(host.c_str(), std::to_string(master_port_).c_str(), &hints, &res) == 0 && res)
      << "cannot resolve assigner host " << master_host_;
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(60);
  while (true) {
    fd_ = ::socket(AF_INET, SOCK_STREAM, 0);
    if (::connect(fd_, res->ai_addr, res->ai_addrlen) == 0) break;
    ::close(fd_);
    fd_ = -1;
    MINIPS_CHECK(std::chrono::steady_clock::now() < deadline)
        << "block assigner at " << master_host_ << ":" << master_port_ << " unreachable";
    std::this_thread::sleep_for(std::chrono::milliseconds(20));
  }
  ::freeaddrinfo(res);
  int one = 1;
  ::setsockopt(fd_, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
}

BinStream Coordinator::ask_master(BinStream& question, int type) {
  std::lock_guard<std::mutex> lk(mu_);
  MINIPS_CHECK(fd_ >= 0) << "Coordinator::serve() first";
  MINIPS_CHECK(SendFrame(fd_, type, question)) << "assigner connection lost";
  int rtype = 0;
  BinStream answer;
  MINIPS_CHECK(RecvFrame(fd_, &rtype, &answer)) << "assigner connection lost";
  return answer;
}

void Coordinator::notify_master(BinStream& message, int type) {
  std::lock_guard<std::mutex> lk(mu_);
  MINIPS_CHECK(fd_ >= 0) << "Coordinator::serve() first";
  MINIPS_CHECK(SendFrame(fd_, type, message)) << "assigner connection lost";
}

// ------------------------------------------------------------------ BlockAssigner
BlockAssigner::BlockAssigner(int port, int expected_loaders, size_t block_size)
    : port_(port), expected_loaders_(expected_loaders), block_size_(block_size) {
  listen_fd_ = ::socket(AF_INET, SOCK_STREAM, 0);
  MINIPS_CHECK(listen_fd_ >= 0);
  int one = 1;
  ::setsockopt(listen_fd_, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_addr.s_addr = htonl(INADDR_ANY);
  addr.sin_port = htons(static_cast<uint16_t>(port));
  MINIPS_CHECK(::bind(listen_fd_, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) == 0)
      << "block assigner cannot bind port " << port << ": " << strerror(errno);
  MINIPS_CHECK(::listen(listen_fd_, 128) == 0);
}

BlockAssigner::~BlockAssigner() {
  Join();
  if (listen_fd_ >= 0) ::close(listen_fd_);
}

void BlockAssigner::Scan(const std::string& url) {
  // every block of every file under url; the "replica hosts" of block i rotate over the hosts
  // seen so far, which keeps the locality preference of the reference meaningful on one box
  std::vector<Block>& blocks = pending_[url];
  size_t idx = 0;
  for (const auto& file : io::ListDirectory(url)) {
    const int64_t size = io::FileSize(file);
    for (int64_t off = 0; off < size; off += static_cast<int64_t>(block_size_)) {
      Block b{file, static_cast<size_t>(off), {}};
      if (!known_hosts_.empty()) b.hosts.push_back(known_hosts_[idx % known_hosts_.size()]);
      blocks.push_back(std::move(b));
      ++idx;
    }
  }
  scanned_[url] = true;
}

std::pair<std::string, size_t> BlockAssigner::Answer(const std::string& url, const std::string& host) {
  std::lock_guard<std::mutex> lk(mu_);
  if (std::find(known_hosts_.begin(), known_hosts_.end(), host) == known_hosts_.end()) known_hosts_.push_back(host);
  if (!scanned_[url]) Scan(url);
  auto& blocks = pending_[url];
  if (blocks.empty()) return {"", 0};
  // prefer a block whose replica lives on the asking host
  size_t pick = 0;
  for (size_t i = 0; i < blocks.size(); ++i)
    if (std::find(blocks[i].hosts.begin(), blocks[i].hosts.end(), host) != blocks[i].hosts.end()) {
      pick = i;
      break;
    }
  auto out = std::make_pair(blocks[pick].file, blocks[pick].offset);
  blocks.erase(blocks.begin() + static_cast<long>(pick));
  return out;
}

void BlockAssigner::Serve() {
  std::vector<int> conns;
  int finished = 0;
  while (finished < expected_loaders_) {
    std::vector<pollfd> fds;
    fds.push_back({listen_fd_, POLLIN, 0});
    for (int c : conns) fds.push_back({c, POLLIN, 0});
    const int rc = ::poll(fds.data(), fds.size(), 1000);
    if (rc < 0 && errno != EINTR) break;
    if (rc <= 0) continue;
    if (fds[0].revents & POLLIN) {
      const int c = ::accept(listen_fd_, nullptr, nullptr);
      if (c >= 0) conns.push_back(c);
    }
    for (size_t i = 1; i < fds.size(); ++i) {
      if (!(fds[i].revents & (POLLIN | POLLHUP | POLLERR))) continue;
      int type = 0;
      BinStream body;
      if (!RecvFrame(fds[i].fd, &type, &body)) {
        ::close(fds[i].fd);
        conns.erase(std::find(conns.begin(), conns.end(), fds[i].fd));
        continue;
      }
      if (type == kAssignerBlockRequest) {
        std::string url, host;
        int num_threads = 0, id = 0;
        body >> url >> host >> num_threads >> id;
        auto ans = Answer(url, host);
        BinStream reply;
        reply << ans.first << static_cast<uint64_t>(ans.second);
        SendFrame(fds[i].fd, type, reply);
      } else if (type == kAssignerExit) {
        ++finished;
      }
    }
  }
  for (int c : conns) ::close(c);
}

void BlockAssigner::StartAsync() {
  th_ = std::thread([this] { Serve(); });
}

void BlockAssigner::Join() {
  if (th_.joinable()) th_.join();
}

// ------------------------------------------------------------------ FileSplitter
FileSplitter::FileSplitter(Coordinator* coordinator, std::string url, int num_threads, size_t block_size)
    : coordinator_(coordinator), url_(std::move(url)), num_threads_(num_threads), block_size_(block_size) {}

size_t FileSplitter::read_block(const std::string& file, size_t offset) {
  const int fd = ::open(file.c_str(), O_RDONLY);
  if (fd < 0) return 0;
  buf_.resize(block_size_);
  size_t got = 0;
  while (got < block_size_) {
    const ssize_t r = ::pread(fd, buf_.data() + got, block_size_ - got, static_cast<off_t>(offset + got));
    if (r <= 0) break;
    got += static_cast<size_t>(r);
  }
  ::close(fd);
  return got;
}

std::pair<const char*, size_t> FileSplitter::fetch_block(bool is_next) {
  if (is_next) {
    offset_ += block_size_;
  } else {
    BinStream q;
    q << url_ << coordinator_->get_hostname() << num_threads_ << coordinator_->get_proc_id();
    BinStream a = coordinator_->ask_master(q, kAssignerBlockRequest);
    uint64_t off = 0;
    a >> file_ >> off;
    offset_ = off;
    if (file_.empty()) return {nullptr, 0};
  }
  const size_t n = read_block(file_, offset_);
  if (n == 0) return {nullptr, 0};
  return {buf_.data(), n};
}

// ------------------------------------------------------------------ LineInputFormat
LineInputFormat::LineInputFormat(Coordinator* coordinator, const std::string& url, int num_threads, size_t block_size)
    : splitter_(coordinator, url, num_threads, block_size), block_size_(block_size) {}

bool LineInputFormat::fetch_new_block() {
  auto blk = splitter_.fetch_block(false);
  data_ = blk.first;
  len_ = blk.second;
  pos_ = 0;
  if (!data_) return false;
  if (splitter_.get_offset() != 0) {
    // the previous block's reader owns the line that straddles into this block — unless the
    // previous block ended exactly on a newline, which only its reader can know; convention:
    // a reader always consumes through the first newline at or after its block's end.
    while (pos_ < len_ && data_[pos_] != '\n') ++pos_;
    if (pos_ < len_) ++pos_;
    else return fetch_new_block();  // no newline at all: the whole block is the middle of someone else's line
  }
  return true;
}

// Ownership convention (the usual one for line-oriented splits): the reader of a block that
// does not start at offset 0 discards everything through the first newline; every reader
// keeps going while a line STARTS at or before its block's end, finishing that last line from
// the following block(s).  Each line is therefore produced exactly once.
bool LineInputFormat::read_tail(std::string* line) {
  bool any = false;
  while (true) {
    auto nb = splitter_.fetch_block(true);
    if (!nb.first) break;
    any = true;
    const char* e = static_cast<const char*>(memchr(nb.first, '\n', nb.second));
    if (e) {
      line->append(nb.first, static_cast<size_t>(e - nb.first));
      break;
    }
    line->append(nb.first, nb.second);
    if (nb.second < block_size_) break;
  }
  return any;
}

bool LineInputFormat::next(std::string* line) {
  while (true) {
    if (!data_ && !fetch_new_block()) return false;
    line->clear();
    if (pos_ < len_) {
      const char* start = data_ + pos_;
      const char* nl = static_cast<const char*>(memchr(start, '\n', len_ - pos_));
      if (nl) {
        line->assign(start, static_cast<size_t>(nl - start));
        pos_ += static_cast<size_t>(nl - start) + 1;
      } else {
        // the line runs past the end of the block: finish it, then this block is exhausted
        line->assign(start, len_ - pos_);
        const bool full = (len_ == block_size_);
        data_ = nullptr;
        if (full) read_tail(line);
      }
    } else {
      // consumed exactly to the end: a line starting on the boundary of a full block is ours
      const bool full = (len_ == block_size_);
      data_ = nullptr;
      if (!full || !read_tail(line)) continue;
    }
    if (!line->empty() && line->back() == '\r') line->pop_back();
    if (line->empty()) continue;
    return true;
  }
}

// ------------------------------------------------------------------ DataManager
DataManager::DataManager(const Node& node, const std::vector<Node>& nodes, const Config& config)
    : node_(node), nodes_(nodes), config_(config) {}

void DataManager::Start() {
  if (node_.id != nodes_.front().id) return;
  const int loaders = static_cast<int>(nodes_.size()) * config_.num_local_load_thread;
  assigner_.reset(new BlockAssigner(config_.master_port, loaders, config_.block_size));
  assigner_->StartAsync();
}

void DataManager::Run(const std::function<void(const std::string&, int)>& fn) {
  std::vector<std::thread> threads;
  std::vector<std::string> errors;
  std::mutex err_mu;
  for (int t = 0; t < config_.num_local_load_thread; ++t) {
    threads.emplace_back([&, t] {
      try {
        Coordinator coord(static_cast<int>(node_.id) * 1000 + t, config_.worker_host + "#" + std::to_string(node_.id),
                          config_.master_host, config_.master_port);
        coord.serve();
        LineInputFormat in(&coord, config_.url, config_.num_local_load_thread, config_.block_size);
        std::string line;
        while (in.next(&line))
          if (!line.empty()) fn(line, t);
        BinStream bye;
        bye << coord.get_hostname();
        coord.notify_master(bye, kAssignerExit);
      } catch (const std::exception& e) {
        std::lock_guard<std::mutex> lk(err_mu);
        errors.push_back(e.what());
      }
    });
  }
  for (auto& th : threads) th.join();
  MINIPS_CHECK(errors.empty()) << "data loading failed: " << errors.front();
}

void DataManager::Stop() {
  if (assigner_) assigner_->Join();
  assigner_.reset();
}

}  // namespace minips
