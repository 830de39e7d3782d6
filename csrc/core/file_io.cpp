#include "core/file_io.hpp"

#include <dirent.h>
#include <sys/stat.h>
#include <sys/types.h>

#include <algorithm>
#include <cstdlib>

namespace minips {
namespace io {

namespace {
constexpr size_t kBufferSize = 16u << 20;  // general_fstream.hpp:15

std::string HdfsRoot() {
  const char* e = std::getenv("MINIPS_HDFS_ROOT");
  return e && *e ? std::string(e) : std::string("/tmp/minips_hdfs");
}
}  // namespace

bool IsHdfsPath(const std::string& path) { return path.compare(0, 7, "hdfs://") == 0; }

std::string ResolvePath(const std::string& path) {
  if (!IsHdfsPath(path)) return path;
  std::string rest = path.substr(7);           // host:port/a/b  or  /a/b
  const size_t slash = rest.find('/');
  std::string tail = slash == std::string::npos ? std::string("/") : rest.substr(slash);
  return HdfsRoot() + tail;
}

bool Exists(const std::string& path) {
  struct stat st;
  return ::stat(ResolvePath(path).c_str(), &st) == 0;
}

int64_t FileSize(const std::string& path) {
  struct stat st;
  if (::stat(ResolvePath(path).c_str(), &st) != 0) return -1;
  return static_cast<int64_t>(st.st_size);
}

void MakeDirsFor(const std::string& file_path) {
  const std::string p = ResolvePath(file_path);
  for (size_t i = 1; i < p.size(); ++i) {
    if (p[i] == '/') {
      const std::string dir = p.substr(0, i);
      ::mkdir(dir.c_str(), 0755);
    }
  }
}

std::vector<std::string> ListDirectory(const std::string& path) {
  std::vector<std::string> out;
  const std::string p = ResolvePath(path);
  struct stat st;
  if (::stat(p.c_str(), &st) != 0) return out;
  if (!S_ISDIR(st.st_mode)) {
    out.push_back(p);
    return out;
  }
  DIR* d = ::opendir(p.c_str());
  if (!d) return out;
  while (dirent* e = ::readdir(d)) {
    const std::string name = e->d_name;
    if (name == "." || name == "..") continue;
    const std::string full = p + (p.back() == '/' ? "" : "/") + name;
    struct stat fs;
    if (::stat(full.c_str(), &fs) == 0 && S_ISREG(fs.st_mode)) out.push_back(full);
  }
  ::closedir(d);
  std::sort(out.begin(), out.end());
  return out;
}

ifstream::ifstream(const std::string& path, std::ios_base::openmode mode) : buf_(kBufferSize) {
  rdbuf()->pubsetbuf(buf_.data(), static_cast<std::streamsize>(buf_.size()));
  open(ResolvePath(path), mode | std::ios_base::in);
}

ofstream::ofstream(const std::string& path, std::ios_base::openmode mode) : buf_(1u << 20) {
  rdbuf()->pubsetbuf(buf_.data(), static_cast<std::streamsize>(buf_.size()));
  if (IsHdfsPath(path)) MakeDirsFor(path);
  open(ResolvePath(path), mode | std::ios_base::out);
}

}  // namespace io
}  // namespace minips
