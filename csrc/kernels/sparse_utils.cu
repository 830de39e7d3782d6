// sparse_utils.cu — device versions of the reference's host-side sparse helpers.
//
//   batch_unique_keys   lib/batch_data_sampler.cpp:50-71 builds the sorted unique feature-id set
//                       of a batch with std::set (and lr_example.cpp:231-235 precomputes it for
//                       every iteration).  Device form: bitmap over the key space + popcount
//                       scan -> ascending unique keys, plus the nnz -> slot map used to
//                       pre-aggregate gradients per key.
//   libsvm_to_csr       lib/parser.hpp:17-47 (strtok_r per line) + io/line_input_format.hpp line
//                       splitting.  Device form: newline scan -> line table, ':' count -> row_ptr,
//                       then one thread per line parses "label idx:val ..." into CSR.
//
// Both are built on one exclusive-scan primitive (reduce / scan-of-block-sums / downsweep).
#include "common.cuh"

namespace mps {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 16;                          // per thread
constexpr int kScanTile = kScanThreads * kScanItems;    // 4096 items per CTA

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total) {
  __shared__ uint32_t warp_sums[kScanThreads / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  uint32_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) warp_sums[wid] = incl;
  __syncthreads();
  if (wid == 0) {
    uint32_t w = lane < kScanThreads / 32 ? warp_sums[lane] : 0;
    uint32_t wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += t;
    }
    if (lane < kScanThreads / 32) warp_sums[lane] = wi - w;  // exclusive warp offsets
    if (lane == kScanThreads / 32 - 1) *total = wi;
  }
  __syncthreads();
  const uint32_t r = warp_sums[wid] + incl - v;
  __syncthreads();
  return r;
}

// F maps an item index to its value (uint32).  Phase 1: per-tile totals.
template <typename F>
__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(F f, int64_t n, uint32_t* __restrict__ tile_sums) {
  __shared__ uint32_t total;
  const int64_t base = static_cast<int64_t>(blockIdx.x) * kScanTile + threadIdx.x * kScanItems;
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i)
    if (base + i < n) s += f(base + i);
  block_exclusive_scan(s, &total);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// Phase 2: exclusive scan of the tile totals in place by ONE CTA; tile_sums[n_tiles] = grand total.
__global__ void __launch_bounds__(kScanThreads) scan_tiles_kernel(uint32_t* __restrict__ tile_sums, int64_t n_tiles) {
  __shared__ uint32_t total;
  uint32_t carry = 0;
  for (int64_t base = 0; base < n_tiles; base += kScanThreads) {
    const int64_t i = base + threadIdx.x;
    const uint32_t v = i < n_tiles ? tile_sums[i] : 0;
    const uint32_t ex = block_exclusive_scan(v, &total);
    if (i < n_tiles) tile_sums[i] = carry + ex;
    carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) tile_sums[n_tiles] = carry;
}

// Phase 3: G consumes (index, exclusive prefix, value) for every item.
template <typename F, typename G>
__global__ void __launch_bounds__(kScanThreads)
scan_apply_kernel(F f, G g, int64_t n, const uint32_t* __restrict__ tile_sums) {
  __shared__ uint32_t total;
  const int64_t base = static_cast<int64_t>(blockIdx.x) * kScanTile + threadIdx.x * kScanItems;
  uint32_t vals[kScanItems];
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    vals[i] = (base + i < n) ? f(base + i) : 0;
    s += vals[i];
  }
  uint32_t prefix = tile_sums[blockIdx.x] + block_exclusive_scan(s, &total);
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i < n) g(base + i, prefix, vals[i]);
    prefix += vals[i];
  }
}

template <typename F, typename G>
static int exclusive_scan(F f, G g, int64_t n, uint32_t* tile_sums, cudaStream_t st) {
  if (n <= 0) return 0;
  const int64_t n_tiles = (n + kScanTile - 1) / kScanTile;
  scan_reduce_kernel<<<static_cast<unsigned>(n_tiles), kScanThreads, 0, st>>>(f, n, tile_sums);
  scan_tiles_kernel<<<1, kScanThreads, 0, st>>>(tile_sums, n_tiles);
  scan_apply_kernel<<<static_cast<unsigned>(n_tiles), kScanThreads, 0, st>>>(f, g, n, tile_sums);
  MPS_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------ batch unique keys
__global__ void __launch_bounds__(256)
mark_keys_kernel(const __grid_constant__ MpsCsr csr, int64_t row_begin, int64_t batch, uint32_t num_dims,
                 uint32_t* __restrict__ bitmap) {
  // one warp per row, like the training kernels
  const int lane = threadIdx.x & 31;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r = warp; r < batch; r += nwarps) {
    int64_t ridx = row_begin + r;
    if (ridx >= csr.n_rows) ridx %= csr.n_rows;
    const int64_t beg = csr.row_ptr[ridx], end = csr.row_ptr[ridx + 1];
    for (int64_t j = beg + lane; j < end; j += 32) {
      const uint32_t k = csr.col[j];
      if (k < num_dims) atomicOr(bitmap + (k >> 5), 1u << (k & 31));
    }
  }
}

__global__ void __launch_bounds__(256)
map_slots_kernel(const __grid_constant__ MpsCsr csr, int64_t row_begin, int64_t batch, uint32_t num_dims,
                 const uint32_t* __restrict__ bitmap, const uint32_t* __restrict__ word_prefix,
                 int64_t nnz_base, int32_t* __restrict__ slots) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r = warp; r < batch; r += nwarps) {
    const int64_t ridx = row_begin + r;  // slot map is defined for non-wrapping batches only
    if (ridx >= csr.n_rows) continue;
    const int64_t beg = csr.row_ptr[ridx], end = csr.row_ptr[ridx + 1];
    for (int64_t j = beg + lane; j < end; j += 32) {
      const uint32_t k = csr.col[j];
      int32_t slot = -1;
      if (k < num_dims) {
        const uint32_t w = bitmap[k >> 5];
        slot = static_cast<int32_t>(word_prefix[k >> 5] + __popc(w & ((1u << (k & 31)) - 1u)));
      }
      slots[j - nnz_base] = slot;
    }
  }
}

struct PopcWord {
  const uint32_t* bitmap;
  __device__ uint32_t operator()(int64_t i) const { return __popc(bitmap[i]); }
};
struct EmitKeys {
  const uint32_t* bitmap;
  uint32_t* word_prefix;
  uint32_t* keys;
  __device__ void operator()(int64_t i, uint32_t prefix, uint32_t) const {
    word_prefix[i] = prefix;
    uint32_t w = bitmap[i];
    uint32_t out = prefix;
    while (w) {
      const int b = __ffs(w) - 1;
      keys[out++] = static_cast<uint32_t>(i) * 32u + static_cast<uint32_t>(b);
      w &= w - 1;
    }
  }
};

// ------------------------------------------------------------------ libsvm text -> CSR
__device__ __forceinline__ bool is_space(char c) { return c == ' ' || c == '\t' || c == '\r'; }

// minimal strtod: [+-]digits[.digits][(e|E)[+-]digits]; advances p; returns false if no digits
__device__ bool parse_number(const char*& p, const char* end, double* out) {
  const char* s = p;
  bool neg = false;
  if (s < end && (*s == '+' || *s == '-')) { neg = (*s == '-'); ++s; }
  double v = 0.0;
  int digits = 0;
  while (s < end && *s >= '0' && *s <= '9') { v = v * 10.0 + (*s - '0'); ++s; ++digits; }
  if (s < end && *s == '.') {
    ++s;
    double scale = 0.1;
    while (s < end && *s >= '0' && *s <= '9') { v += (*s - '0') * scale; scale *= 0.1; ++s; ++digits; }
  }
  if (digits == 0) return false;
  if (s < end && (*s == 'e' || *s == 'E')) {
    const char* e = s + 1;
    bool eneg = false;
    if (e < end && (*e == '+' || *e == '-')) { eneg = (*e == '-'); ++e; }
    int ex = 0, ed = 0;
    while (e < end && *e >= '0' && *e <= '9') { ex = ex * 10 + (*e - '0'); ++e; ++ed; }
    if (ed) {
      v *= pow(10.0, eneg ? -ex : ex);
      s = e;
    }
  }
  *out = neg ? -v : v;
  p = s;
  return true;
}

struct NewlineFlag {
  const char* text;
  int64_t n;
  // position n is a virtual newline when the text does not end with one
  __device__ uint32_t operator()(int64_t i) const { return (i == n) ? (text[n - 1] != '\n') : (text[i] == '\n'); }
};
struct EmitLineEnds {
  int64_t* line_end;
  __device__ void operator()(int64_t i, uint32_t prefix, uint32_t v) const {
    if (v) line_end[prefix] = i;
  }
};

// per line: valid flag (first non-blank char is not '#' and the line is not empty) and nnz (':' count)
__global__ void __launch_bounds__(256)
line_stats_kernel(const char* __restrict__ text, const int64_t* __restrict__ line_end, int64_t n_lines,
                  uint32_t* __restrict__ valid, uint32_t* __restrict__ nnz) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n_lines) return;
  const char* p = text + (k ? line_end[k - 1] + 1 : 0);
  const char* end = text + line_end[k];
  while (p < end && is_space(*p)) ++p;
  const bool ok = p < end && *p != '#';
  uint32_t c = 0;
  if (ok)
    for (const char* q = p; q < end && *q != '#'; ++q) c += (*q == ':');
  valid[k] = ok;
  nnz[k] = ok ? c : 0;
}

struct ValidFlag {
  const uint32_t* valid;
  __device__ uint32_t operator()(int64_t i) const { return valid[i]; }
};
struct EmitRowOfLine {
  int64_t* row_of_line;
  __device__ void operator()(int64_t i, uint32_t prefix, uint32_t v) const {
    row_of_line[i] = v ? static_cast<int64_t>(prefix) : static_cast<int64_t>(-1);
  }
};
struct NnzOfLine {
  const uint32_t* nnz;
  __device__ uint32_t operator()(int64_t i) const { return nnz[i]; }
};
struct EmitRowPtr {
  const int64_t* row_of_line;
  int64_t* row_ptr;
  int64_t* nnz_of_line_prefix;
  __device__ void operator()(int64_t i, uint32_t prefix, uint32_t) const {
    nnz_of_line_prefix[i] = prefix;
    if (row_of_line[i] >= 0) row_ptr[row_of_line[i]] = prefix;
  }
};

__global__ void __launch_bounds__(256)
parse_lines_kernel(const char* __restrict__ text, const int64_t* __restrict__ line_end, int64_t n_lines,
                   const int64_t* __restrict__ row_of_line, const int64_t* __restrict__ nnz_prefix,
                   uint32_t* __restrict__ col, float* __restrict__ val, float* __restrict__ label,
                   int map_negative_label_to_zero) {
  const int64_t k = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (k >= n_lines) return;
  const int64_t row = row_of_line[k];
  if (row < 0) return;
  const char* p = text + (k ? line_end[k - 1] + 1 : 0);
  const char* end = text + line_end[k];
  while (p < end && is_space(*p)) ++p;
  double y = 0.0;
  parse_number(p, end, &y);
  if (map_negative_label_to_zero && y < 0) y = 0.0;  // lr_example.cpp:296-297
  label[row] = static_cast<float>(y);
  int64_t out = nnz_prefix[k];
  while (p < end) {
    while (p < end && is_space(*p)) ++p;
    if (p >= end || *p == '#') break;
    double idx = 0.0, v = 0.0;
    if (!parse_number(p, end, &idx) || p >= end || *p != ':') break;
    ++p;
    if (!parse_number(p, end, &v)) break;
    col[out] = static_cast<uint32_t>(static_cast<int64_t>(idx) - 1);  // 1-based in the file (lib/parser.hpp:33)
    val[out] = static_cast<float>(v);
    ++out;
  }
}

// ------------------------------------------------------------------ compact batch format
// A batch travels host -> device as row lengths (u8), feature ids bit-packed at ceil(log2(features)) bits
// (16 / 20 / 24), fp32 values and u8 labels: 15-20 % fewer bytes than int64 row_ptr + int32 ids + fp32
// labels, lossless for < 256 non-zeros per row.  These kernels rebuild the CSR arrays the training kernels read.
struct RowLenU8 {
  const uint8_t* len;
  __device__ uint32_t operator()(int64_t i) const { return len[i]; }
};
struct EmitRowPtr64 {
  int64_t* row_ptr;
  int64_t n;
  __device__ void operator()(int64_t i, uint32_t prefix, uint32_t v) const {
    row_ptr[i] = prefix;
    if (i == n - 1) row_ptr[n] = static_cast<int64_t>(prefix) + v;
  }
};

__global__ void __launch_bounds__(256)
csr_unpack_kernel(const uint32_t* __restrict__ words, int id_bits, int bit_off, int64_t nnz, uint32_t* __restrict__ col,
                  const uint8_t* __restrict__ label_u8, int64_t rows, float* __restrict__ label) {
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const uint32_t mask = id_bits >= 32 ? 0xffffffffu : ((1u << id_bits) - 1u);
  // id i occupies bits [bit_off + i * id_bits, +id_bits) of the little-endian stream: two aligned words and a
  // funnel shift (the staging buffer starts on a 16-byte boundary and has a word of slack at the end)
  auto id_at = [&](int64_t i) {
    const int64_t p = bit_off + i * id_bits;
    const uint32_t lo = words[p >> 5], hi = words[(p >> 5) + 1];
    return __funnelshift_r(lo, hi, static_cast<uint32_t>(p & 31)) & mask;
  };
  const int64_t quads = nnz >> 2;
  for (int64_t q = tid; q < quads; q += stride) {
    uint4 o;
    o.x = id_at(4 * q);
    o.y = id_at(4 * q + 1);
    o.z = id_at(4 * q + 2);
    o.w = id_at(4 * q + 3);
    *reinterpret_cast<uint4*>(col + 4 * q) = o;
  }
  for (int64_t i = (quads << 2) + tid; i < nnz; i += stride) col[i] = id_at(i);
  for (int64_t i = tid; i < rows; i += stride) label[i] = static_cast<float>(label_u8[i]);
}

}  // namespace mps

extern "C" {

// scratch_u32 needs 2 * ceil(num_dims / 32) + ceil(words / 4096) + 2 words.  keys_out holds up
// to min(batch nnz, num_dims) keys (ascending); count_out[0] receives their number.
// slots_out (optional, sized like the batch's nnz range) maps every nnz to its key's slot.
int mps_batch_unique_keys(const MpsCsr* csr, int64_t row_begin, int64_t batch, uint32_t num_dims,
                          uint32_t* scratch_u32, uint32_t* keys_out, uint32_t* count_out, int32_t* slots_out,
                          int64_t nnz_base, void* stream) {
  auto st = static_cast<cudaStream_t>(stream);
  const int64_t words = (static_cast<int64_t>(num_dims) + 31) / 32;
  uint32_t* bitmap = scratch_u32;
  uint32_t* word_prefix = scratch_u32 + words;
  uint32_t* tile_sums = word_prefix + words;
  const int64_t n_tiles = (words + mps::kScanTile - 1) / mps::kScanTile;
  MPS_CHECK(cudaMemsetAsync(bitmap, 0, words * sizeof(uint32_t), st));
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int64_t want = (batch + 7) / 8;
  const int grid = static_cast<int>(want < 1 ? 1 : (want < sms * 8 ? want : sms * 8));
  mps::mark_keys_kernel<<<grid, 256, 0, st>>>(*csr, row_begin, batch, num_dims, bitmap);
  MPS_LAUNCH_CHECK();
  int rc = mps::exclusive_scan(mps::PopcWord{bitmap}, mps::EmitKeys{bitmap, word_prefix, keys_out}, words, tile_sums,
      st);
  if (rc) return rc;
  MPS_CHECK(cudaMemcpyAsync(count_out, tile_sums + n_tiles, sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
  if (slots_out) {
    mps::map_slots_kernel<<<grid, 256, 0, st>>>(*csr, row_begin, batch, num_dims, bitmap, word_prefix, nnz_base,
                                                slots_out);
    MPS_LAUNCH_CHECK();
  }
  return 0;
}

// Pass 1 of the parser: line table.  line_end [max_lines] receives the byte offset of every
// line terminator; counts_out[0] = number of lines.  scratch_u32: ceil((nbytes+1)/4096) + 2 words.
int mps_libsvm_scan_lines(const char* text, int64_t nbytes, int64_t* line_end, uint32_t* scratch_u32,
                          uint32_t* counts_out, void* stream) {
  if (nbytes <= 0) return 0;
  auto st = static_cast<cudaStream_t>(stream);
  const int64_t n = nbytes + 1;
  const int64_t n_tiles = (n + mps::kScanTile - 1) / mps::kScanTile;
  int rc = mps::exclusive_scan(mps::NewlineFlag{text, nbytes}, mps::EmitLineEnds{line_end}, n, scratch_u32, st);
  if (rc) return rc;
  MPS_CHECK(cudaMemcpyAsync(counts_out, scratch_u32 + n_tiles, sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
  return 0;
}

// Pass 2: rows and row_ptr.  Work arrays (all [n_lines] unless noted): valid, nnz (u32),
// row_of_line, nnz_prefix (i64); row_ptr [n_lines + 1]; counts_out[1] = rows, counts_out[2] = nnz.
int mps_libsvm_index_lines(const char* text, const int64_t* line_end, int64_t n_lines, uint32_t* valid,
                           uint32_t* nnz, int64_t* row_of_line, int64_t* nnz_prefix, int64_t* row_ptr,
                           uint32_t* scratch_u32, uint32_t* counts_out, void* stream) {
  if (n_lines <= 0) return 0;
  auto st = static_cast<cudaStream_t>(stream);
  const int grid = static_cast<int>((n_lines + 255) / 256);
  mps::line_stats_kernel<<<grid, 256, 0, st>>>(text, line_end, n_lines, valid, nnz);
  MPS_LAUNCH_CHECK();
  const int64_t n_tiles = (n_lines + mps::kScanTile - 1) / mps::kScanTile;
  int rc = mps::exclusive_scan(mps::ValidFlag{valid}, mps::EmitRowOfLine{row_of_line}, n_lines, scratch_u32, st);
  if (rc) return rc;
  MPS_CHECK(cudaMemcpyAsync(counts_out + 1, scratch_u32 + n_tiles, sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
  rc = mps::exclusive_scan(mps::NnzOfLine{nnz}, mps::EmitRowPtr{row_of_line, row_ptr, nnz_prefix}, n_lines,
                           scratch_u32, st);
  if (rc) return rc;
  MPS_CHECK(cudaMemcpyAsync(counts_out + 2, scratch_u32 + n_tiles, sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
  return 0;
}

// Pass 3: fill col / val / label (row_ptr[rows] = nnz is written by the caller).
int mps_libsvm_parse(const char* text, const int64_t* line_end, int64_t n_lines, const int64_t* row_of_line,
                     const int64_t* nnz_prefix, uint32_t* col, float* val, float* label,
                     int map_negative_label_to_zero, void* stream) {
  if (n_lines <= 0) return 0;
  const int grid = static_cast<int>((n_lines + 255) / 256);
  mps::parse_lines_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      text, line_end, n_lines, row_of_line, nnz_prefix, col, val, label, map_negative_label_to_zero);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_csr_unpack(const uint8_t* row_len, const void* packed_ids, int id_bits, int bit_off, const uint8_t* label_u8,
                   int64_t rows, int64_t nnz, int64_t* row_ptr, uint32_t* col, float* label, uint32_t* scratch_u32,
                   void* stream) {
  if (rows <= 0) return 0;
  if ((reinterpret_cast<uintptr_t>(packed_ids) & 3) || (reinterpret_cast<uintptr_t>(col) & 15) || id_bits < 1 ||
      id_bits > 32 || bit_off < 0 || bit_off > 31)
    return static_cast<int>(cudaErrorInvalidValue);
  auto st = static_cast<cudaStream_t>(stream);
  int rc = mps::exclusive_scan(mps::RowLenU8{row_len}, mps::EmitRowPtr64{row_ptr, rows}, rows, scratch_u32, st);
  if (rc) return rc;
  const int64_t work = (nnz >> 2) > rows ? (nnz >> 2) : rows;
  int64_t grid = (work + 255) / 256;
  if (grid > 148 * 8) grid = 148 * 8;
  if (grid < 1) grid = 1;
  mps::csr_unpack_kernel<<<static_cast<unsigned>(grid), 256, 0, st>>>(static_cast<const uint32_t*>(packed_ids), id_bits,
                                                                     bit_off, nnz, col, label_u8, rows, label);
  MPS_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
