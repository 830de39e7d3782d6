// file_io.hpp — one stream type for "a path that may be local or hdfs://".
//
// The reference routes every hostfile / checkpoint / report read and write through
// petuum::io::{ifstream,ofstream} (base/third_party/general_fstream.hpp:110-111, .cpp:124-172),
// which opens hdfs://[host:port]/path through libhdfs3 and anything else as a local file.
// There is no HDFS on an 8-GPU box: hdfs:// URLs are resolved under a local root directory
// ($MINIPS_HDFS_ROOT, default /tmp/minips_hdfs) so scripts and flags written for the
// reference keep working, and everything else is a plain file with a 16 MiB buffer.
#pragma once

#include <fstream>
#include <memory>
#include <string>
#include <vector>

namespace minips {
namespace io {

// hdfs://host:port/a/b -> <root>/a/b ; hdfs:///a/b -> <root>/a/b ; other paths unchanged
std::string ResolvePath(const std::string& path);
bool IsHdfsPath(const std::string& path);
bool Exists(const std::string& path);
void MakeDirsFor(const std::string& file_path);           // mkdir -p dirname(file_path)
std::vector<std::string> ListDirectory(const std::string& path);  // files (sorted), path itself if a file
int64_t FileSize(const std::string& path);

class ifstream : public std::ifstream {
 public:
  explicit ifstream(const std::string& path, std::ios_base::openmode mode = std::ios_base::in);
  ~ifstream() override {
    if (is_open()) close();
  }

 private:
  std::vector<char> buf_;
};

class ofstream : public std::ofstream {
 public:
  explicit ofstream(const std::string& path, std::ios_base::openmode mode = std::ios_base::out);
  // flush while the buffer below is still alive (members die before the base class)
  ~ofstream() override {
    if (is_open()) close();
  }

 private:
  std::vector<char> buf_;
};

}  // namespace io
}  // namespace minips
