/* hdfs/hdfs.h shim — the libhdfs3 calls the reference uses, implemented on the LOCAL file system.
 *
 * There is no HDFS on an 8-GPU box; this lets the reference's loader / checkpoint code run
 * unmodified.  An hdfs path "/a/b" (the reference strips scheme and host) is looked up under
 * $MINIPS_REF_HDFS_ROOT (default: the path as given).  Files are presented as sequences of
 * "blocks" of $MINIPS_REF_HDFS_BLOCK bytes (default 64 MiB); every block reports the single
 * host "localhost", so the block assigner's locality logic sees one data node. */
#ifndef MINIPS_SHIM_HDFS_H_
#define MINIPS_SHIM_HDFS_H_
#include <dirent.h>
#include <errno.h>
#include <fcntl.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <sys/types.h>
#include <time.h>
#include <unistd.h>

#ifdef __cplusplus
extern "C++" {  /* the reference includes this header inside extern "C" */
#include <algorithm>
#include <string>
#include <vector>
#endif

typedef int32_t tSize;
typedef time_t tTime;
typedef int64_t tOffset;
typedef uint16_t tPort;
typedef enum tObjectKind { kObjectKindFile = 'F', kObjectKindDirectory = 'D' } tObjectKind;

struct hdfsBuilder {
  std::string nn;
  int port;
};
struct HdfsFileSystemInternalWrapper {
  int dummy;
};
struct HdfsFileInternalWrapper {
  int fd;
  bool write;
};
typedef struct HdfsFileSystemInternalWrapper* hdfsFS;
typedef struct HdfsFileInternalWrapper* hdfsFile;

typedef struct {
  tObjectKind mKind;
  char* mName;
  tTime mLastMod;
  tOffset mSize;
  short mReplication;
  tOffset mBlockSize;
  char* mOwner;
  char* mGroup;
  short mPermissions;
  tTime mLastAccess;
} hdfsFileInfo;

typedef struct BlockLocation {
  int corrupt;
  int numOfNodes;
  char** hosts;
  char** names;
  char** topologyPaths;
  tOffset length;
  tOffset offset;
} BlockLocation;

namespace minips_hdfs_shim {
inline std::string Resolve(const char* path) {
  std::string p = path ? path : "";
  while (!p.empty() && p.back() == '\0') p.pop_back();
  const char* root = getenv("MINIPS_REF_HDFS_ROOT");
  if (root && *root) return std::string(root) + p;
  return p;
}
inline tOffset BlockSize() {
  const char* e = getenv("MINIPS_REF_HDFS_BLOCK");
  return e ? atoll(e) : (64ll << 20);
}
inline void Fill(hdfsFileInfo* info, const std::string& shown, const struct stat& st) {
  memset(info, 0, sizeof(*info));
  info->mKind = S_ISDIR(st.st_mode) ? kObjectKindDirectory : kObjectKindFile;
  info->mName = strdup(shown.c_str());
  info->mSize = st.st_size;
  info->mBlockSize = BlockSize();
  info->mReplication = 1;
  info->mLastMod = st.st_mtime;
}
}  // namespace minips_hdfs_shim

inline struct hdfsBuilder* hdfsNewBuilder(void) { return new hdfsBuilder{"", 0}; }
inline void hdfsBuilderSetNameNode(struct hdfsBuilder* b, const char* nn) { b->nn = nn ? nn : ""; }
inline void hdfsBuilderSetNameNodePort(struct hdfsBuilder* b, tPort port) { b->port = port; }
inline hdfsFS hdfsBuilderConnect(struct hdfsBuilder*) { return new HdfsFileSystemInternalWrapper{0}; }
inline void hdfsFreeBuilder(struct hdfsBuilder* b) { delete b; }
inline hdfsFS hdfsConnect(const char*, tPort) { return new HdfsFileSystemInternalWrapper{0}; }
inline int hdfsDisconnect(hdfsFS fs) {
  delete fs;
  return 0;
}

inline int hdfsExists(hdfsFS, const char* path) {
  struct stat st;
  return ::stat(minips_hdfs_shim::Resolve(path).c_str(), &st) == 0 ? 0 : -1;
}

inline hdfsFileInfo* hdfsGetPathInfo(hdfsFS, const char* path) {
  struct stat st;
  if (::stat(minips_hdfs_shim::Resolve(path).c_str(), &st) != 0) return NULL;
  hdfsFileInfo* info = static_cast<hdfsFileInfo*>(malloc(sizeof(hdfsFileInfo)));
  minips_hdfs_shim::Fill(info, path, st);
  return info;
}

inline hdfsFileInfo* hdfsListDirectory(hdfsFS fs, const char* path, int* num) {
  *num = 0;
  const std::string real = minips_hdfs_shim::Resolve(path);
  struct stat st;
  if (::stat(real.c_str(), &st) != 0) return NULL;
  if (!S_ISDIR(st.st_mode)) {
    *num = 1;
    return hdfsGetPathInfo(fs, path);
  }
  std::vector<std::string> names;
  DIR* d = ::opendir(real.c_str());
  if (!d) return NULL;
  while (dirent* e = ::readdir(d)) {
    if (!strcmp(e->d_name, ".") || !strcmp(e->d_name, "..")) continue;
    names.push_back(e->d_name);
  }
  ::closedir(d);
  std::sort(names.begin(), names.end());
  hdfsFileInfo* out = static_cast<hdfsFileInfo*>(malloc(sizeof(hdfsFileInfo) * (names.size() + 1)));
  std::string base = path;
  while (!base.empty() && base.back() == '\0') base.pop_back();
  if (!base.empty() && base.back() != '/') base += "/";
  for (size_t i = 0; i < names.size(); ++i) {
    struct stat fs2;
    ::stat((real + "/" + names[i]).c_str(), &fs2);
    minips_hdfs_shim::Fill(out + i, base + names[i], fs2);
  }
  *num = static_cast<int>(names.size());
  return out;
}

inline void hdfsFreeFileInfo(hdfsFileInfo* info, int num) {
  if (!info) return;
  for (int i = 0; i < num; ++i) free(info[i].mName);
  free(info);
}

inline BlockLocation* hdfsGetFileBlockLocations(hdfsFS, const char*, tOffset start, tOffset length, int* num) {
  static char host[] = "localhost";
  static char* hosts[] = {host};
  BlockLocation* b = static_cast<BlockLocation*>(calloc(1, sizeof(BlockLocation)));
  b->numOfNodes = 1;
  b->hosts = hosts;
  b->names = hosts;
  b->topologyPaths = hosts;
  b->offset = start;
  b->length = length;
  if (num) *num = 1;
  return b;
}
inline void hdfsFreeFileBlockLocations(BlockLocation* b, int) { free(b); }

inline hdfsFile hdfsOpenFile(hdfsFS, const char* path, int flags, int, short, tOffset) {
  const std::string real = minips_hdfs_shim::Resolve(path);
  const bool wr = (flags & O_WRONLY) || (flags & O_RDWR);
  int fd = wr ? ::open(real.c_str(), O_WRONLY | O_CREAT | ((flags & O_APPEND) ? O_APPEND : O_TRUNC), 0644)
              : ::open(real.c_str(), O_RDONLY);
  if (fd < 0) return NULL;
  return new HdfsFileInternalWrapper{fd, wr};
}
inline int hdfsCloseFile(hdfsFS, hdfsFile f) {
  if (!f) return -1;
  ::close(f->fd);
  delete f;
  return 0;
}
inline int hdfsFileIsOpenForWrite(hdfsFile f) { return f && f->write; }
inline int hdfsFileIsOpenForRead(hdfsFile f) { return f && !f->write; }
inline int hdfsSeek(hdfsFS, hdfsFile f, tOffset pos) { return ::lseek(f->fd, pos, SEEK_SET) < 0 ? -1 : 0; }
inline tOffset hdfsTell(hdfsFS, hdfsFile f) { return ::lseek(f->fd, 0, SEEK_CUR); }
inline tSize hdfsRead(hdfsFS, hdfsFile f, void* buf, tSize len) {
  /* libhdfs may return short reads; the reference loops until the block is complete */
  return static_cast<tSize>(::read(f->fd, buf, static_cast<size_t>(len)));
}
inline tSize hdfsWrite(hdfsFS, hdfsFile f, const void* buf, tSize len) {
  return static_cast<tSize>(::write(f->fd, buf, static_cast<size_t>(len)));
}
inline int hdfsFlush(hdfsFS, hdfsFile) { return 0; }
inline const char* hdfsGetLastError(void) { return strerror(errno); }

#ifdef __cplusplus
}
#endif
#endif
