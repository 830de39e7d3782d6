// Synthetic libsvm writer for the reference arm of bench.py (our own tool, not reference code).
//
//   gen_libsvm <out> <rows> <features> <nnz_per_row> <seed> [threads]
//
// Each row: "<+1|-1> i1:v1 i2:v2 ...\n" with 1-based, strictly increasing feature ids (one per
// stratum of the feature range, so they are sorted and unique without a sort) and values in
// (0,1).  The label comes from a planted linear model so the loss can fall.  Rows are produced in
// per-thread chunks and written in order.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

static inline uint64_t mix(uint64_t& s) {
  s += 0x9E3779B97F4A7C15ull;
  uint64_t z = s;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

static inline char* put_uint(char* p, uint32_t v) {
  char tmp[12];
  int n = 0;
  do {
    tmp[n++] = static_cast<char>('0' + v % 10);
    v /= 10;
  } while (v);
  while (n) *p++ = tmp[--n];
  return p;
}

static void make_rows(int64_t first, int64_t count, int64_t features, int nnz, uint64_t seed, std::string* out) {
  out->clear();
  out->reserve(static_cast<size_t>(count) * (4 + nnz * 16));
  std::vector<char> line(64 + nnz * 24);
  const int64_t stratum = features / nnz;
  for (int64_t r = first; r < first + count; ++r) {
    uint64_t s = seed * 0x100000001B3ull + static_cast<uint64_t>(r) * 0xD6E8FEB86659FD93ull;
    char* p = line.data() + 3;  // room for the label written afterwards
    double margin = 0.0;
    for (int k = 0; k < nnz; ++k) {
      const uint64_t h = mix(s);
      const int64_t lo = k * stratum;
      const int64_t width = (k == nnz - 1) ? features - lo : stratum;
      const uint32_t idx = static_cast<uint32_t>(lo + static_cast<int64_t>(h % static_cast<uint64_t>(width)));
      const uint32_t frac = 1 + static_cast<uint32_t>((h >> 40) % 9998);  // value = 0.frac (4 digits)
      // planted weight: +-1 by a hash of the feature id
      uint64_t ws = idx * 0x9E3779B97F4A7C15ull + 12345;
      const double w = (mix(ws) & 1) ? 1.0 : -1.0;
      margin += w * (frac / 10000.0);
      *p++ = ' ';
      p = put_uint(p, idx + 1);
      *p++ = ':';
      *p++ = '0';
      *p++ = '.';
      *p++ = static_cast<char>('0' + frac / 1000);
      *p++ = static_cast<char>('0' + frac / 100 % 10);
      *p++ = static_cast<char>('0' + frac / 10 % 10);
      *p++ = static_cast<char>('0' + frac % 10);
    }
    *p++ = '\n';
    char* b = line.data() + 1;
    b[0] = margin >= 0 ? '+' : '-';
    b[1] = '1';
    out->append(b, static_cast<size_t>(p - b));
  }
}

int main(int argc, char** argv) {
  if (argc < 6) {
    std::fprintf(stderr, "usage: %s <out> <rows> <features> <nnz_per_row> <seed> [threads]\n", argv[0]);
    return 2;
  }
  const char* path = argv[1];
  const int64_t rows = std::atoll(argv[2]), features = std::atoll(argv[3]);
  const int nnz = std::atoi(argv[4]);
  const uint64_t seed = std::strtoull(argv[5], nullptr, 10);
  int threads = argc > 6 ? std::atoi(argv[6]) : static_cast<int>(std::thread::hardware_concurrency());
  if (threads < 1) threads = 1;
  if (threads > 32) threads = 32;
  if (nnz < 1 || features < nnz || rows < 1) {
    std::fprintf(stderr, "bad shape\n");
    return 2;
  }
  FILE* f = std::fopen(path, "wb");
  if (!f) {
    std::perror(path);
    return 1;
  }
  const int64_t chunk = 1 << 16;
  std::vector<std::string> bufs(threads);
  for (int64_t base = 0; base < rows; base += chunk * threads) {
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) {
      const int64_t first = base + t * chunk;
      bufs[t].clear();
      if (first >= rows) continue;
      const int64_t count = first + chunk <= rows ? chunk : rows - first;
      pool.emplace_back(make_rows, first, count, features, nnz, seed, &bufs[t]);
    }
    for (auto& th : pool) th.join();
    for (int t = 0; t < threads; ++t)
      if (!bufs[t].empty() && std::fwrite(bufs[t].data(), 1, bufs[t].size(), f) != bufs[t].size()) {
        std::perror("write");
        return 1;
      }
  }
  std::fclose(f);
  return 0;
}
