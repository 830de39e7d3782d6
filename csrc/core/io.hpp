// io.hpp — parallel, dynamically balanced loading of line-oriented input.
//
// The reference loads libsvm text from HDFS with a second ZMQ channel (io/coordinator.{hpp,cpp},
// io/hdfs_assigner.{hpp,cpp}, io/hdfs_file_splitter.{hpp,cpp}, io/line_input_format.hpp,
// io/hdfs_manager.{hpp,cpp}): node 0 runs a block assigner that hands out (file, offset)
// pairs on request, preferring blocks stored on the asking host; loader threads fetch a block,
// skip the partial first line, and finish the last line from the following block.  Here the
// same protocol (request 301 = give me a block, 300 = this loader is done) runs over plain
// TCP against files on the local file system (or the hdfs:// -> local mapping of file_io.hpp),
// with a configurable block size.
#pragma once

#include <atomic>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "core/base.hpp"
#include "core/bin_stream.hpp"

namespace minips {

constexpr int kAssignerExit = 300;
constexpr int kAssignerBlockRequest = 301;

// client side of the assigner channel (io/coordinator.hpp:43-70)
class Coordinator {
 public:
  Coordinator(int proc_id, std::string hostname, std::string master_host, int master_port);
  ~Coordinator();
  void serve();  // connect (retries until the assigner is up)
  BinStream ask_master(BinStream& question, int type);
  void notify_master(BinStream& message, int type);
  int get_proc_id() const { return proc_id_; }
  const std::string& get_hostname() const { return hostname_; }

 private:
  int proc_id_;
  std::string hostname_, master_host_;
  int master_port_;
  int fd_ = -1;
  std::mutex mu_;
};

// node-0 service (io/hdfs_assigner.hpp:19-79)
class BlockAssigner {
 public:
  BlockAssigner(int port, int expected_loaders, size_t block_size = 64u << 20);
  ~BlockAssigner();
  void Serve();       // blocking: returns once every loader said kAssignerExit
  void StartAsync();  // Serve on an own thread
  void Join();
  int port() const { return port_; }
  // assignment logic, directly testable: next (file, offset) for `host`; ("", 0) when exhausted
  std::pair<std::string, size_t> Answer(const std::string& url, const std::string& host);

 private:
  struct Block {
    std::string file;
    size_t offset;
    std::vector<std::string> hosts;
  };
  void Scan(const std::string& url);
  int port_;
  int expected_loaders_;
  size_t block_size_;
  int listen_fd_ = -1;
  std::thread th_;
  std::mutex mu_;
  std::map<std::string, std::vector<Block>> pending_;  // url -> unassigned blocks
  std::map<std::string, bool> scanned_;
  std::vector<std::string> known_hosts_;
};

// fetches blocks for one loader thread (io/hdfs_file_splitter.hpp)
class FileSplitter {
 public:
  FileSplitter(Coordinator* coordinator, std::string url, int num_threads, size_t block_size);
  // is_next=false: ask the assigner for a fresh block; true: read the block that follows the
  // current one in the same file (to finish a straddling line).  Empty view = no more data.
  std::pair<const char*, size_t> fetch_block(bool is_next);
  size_t get_offset() const { return offset_; }
  const std::string& current_file() const { return file_; }

 private:
  size_t read_block(const std::string& file, size_t offset);
  Coordinator* coordinator_;
  std::string url_;
  int num_threads_;
  size_t block_size_;
  std::string file_;
  size_t offset_ = 0;
  std::vector<char> buf_;
};

// line iterator over the blocks of one loader (io/line_input_format.hpp:43-121)
class LineInputFormat {
 public:
  LineInputFormat(Coordinator* coordinator, const std::string& url, int num_threads, size_t block_size);
  bool next(std::string* line);

 private:
  bool fetch_new_block();
  bool read_tail(std::string* line);
  FileSplitter splitter_;
  const char* data_ = nullptr;
  size_t len_ = 0, pos_ = 0;
  size_t block_size_;
};

// HDFSManager equivalent (io/hdfs_manager.hpp:21-55)
class DataManager {
 public:
  struct Config {
    std::string master_host = "localhost";  // where the assigner runs (node 0)
    int master_port = 19201;                // assigner_master_port
    std::string worker_host = "localhost";
    int worker_port = 0;                    // unused over TCP, kept for flag parity
    std::string hdfs_namenode = "localhost";
    int hdfs_namenode_port = 9000;
    std::string url;                        // file or directory (hdfs:// allowed)
    int num_local_load_thread = 2;
    size_t block_size = 64u << 20;
  };
  DataManager(const Node& node, const std::vector<Node>& nodes, const Config& config);
  void Start();                                               // node 0: start the assigner
  void Run(const std::function<void(const std::string& line, int thread)>& fn);
  void Stop();                                                // node 0: join the assigner

 private:
  Node node_;
  std::vector<Node> nodes_;
  Config config_;
  std::unique_ptr<BlockAssigner> assigner_;
};

}  // namespace minips
