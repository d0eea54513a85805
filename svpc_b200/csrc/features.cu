// Host side of the input pipeline: clip features go from the per-video `.npy` files straight into the
// pinned input arena.  Replaces, for one training step, the reference's
//   np.concatenate([np.load(<v>_resnet.npy), np.load(<v>_bn.npy)], axis=1)          recursive_caption_dataset.py:187-189
//   feat = np.zeros((max_v_len + max_t_len, D)); feat[1:valid_l+1] = raw_feat[rows]  :389-416 (float64 scratch)
//   feat.astype(np.float32); default_collate stack; .cuda()                          :332, train.py:91-112
// which materialises every row four times on the host.  Here each needed row is read once from the page cache (mmap) and
// written once, as fp32, at its final place in the (B, L, D) clip block of the arena; rows the clip does not cover are
// zero-filled.  No device code in this file: these are the only entry points of the ABI whose pointers are HOST pointers.
#include "common.cuh"
#include <fcntl.h>
#include <stdint.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <atomic>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include <cuda_fp16.h>

namespace {

struct NpyInfo {
  long long rows = 0, cols = 0, data_offset = 0;
  int dtype = -1;  // 0 '<f4', 1 '<f8', 2 '<f2'
};

// Parses the header of a NumPy .npy file, format versions 1.0 - 3.0 (magic, version, little-endian header length, a
// Python dict literal with 'descr', 'fortran_order', 'shape').  Returns false (with the error set) for anything the
// feature files of the reference cannot be: non-C order, rank other than 2, a dtype other than little-endian float.
bool parse_npy_header(const unsigned char* p, size_t size, const char* path, NpyInfo* out) {
  if (size < 10 || memcmp(p, "\x93NUMPY", 6) != 0) {
    svpc_set_error("npy: %s is not a .npy file", path);
    return false;
  }
  const int major = p[6];
  size_t hlen, hoff;
  if (major == 1) {
    hlen = (size_t)p[8] | ((size_t)p[9] << 8);
    hoff = 10;
  } else if (major == 2 || major == 3) {
    if (size < 12) {
      svpc_set_error("npy: %s truncated header", path);
      return false;
    }
    hlen = (size_t)p[8] | ((size_t)p[9] << 8) | ((size_t)p[10] << 16) | ((size_t)p[11] << 24);
    hoff = 12;
  } else {
    svpc_set_error("npy: %s has unsupported format version %d", path, major);
    return false;
  }
  if (hoff + hlen > size) {
    svpc_set_error("npy: %s truncated header", path);
    return false;
  }
  const std::string h((const char*)p + hoff, hlen);
  auto value_after = [&](const char* key) -> size_t {
    size_t k = h.find(key);
    if (k == std::string::npos) return std::string::npos;
    k = h.find(':', k);
    if (k == std::string::npos) return std::string::npos;
    ++k;
    while (k < h.size() && h[k] == ' ') ++k;
    return k;
  };
  size_t k = value_after("'descr'");
  if (k == std::string::npos || k + 1 >= h.size() || (h[k] != '\'' && h[k] != '"')) {
    svpc_set_error("npy: %s has no simple 'descr'", path);
    return false;
  }
  const size_t e = h.find(h[k], k + 1);
  const std::string descr = h.substr(k + 1, e == std::string::npos ? 0 : e - k - 1);
  if (descr == "<f4") out->dtype = 0;
  else if (descr == "<f8") out->dtype = 1;
  else if (descr == "<f2") out->dtype = 2;
  else {
    svpc_set_error("npy: %s has dtype '%s' (need <f4, <f8 or <f2)", path, descr.c_str());
    return false;
  }
  k = value_after("'fortran_order'");
  if (k == std::string::npos || h.compare(k, 5, "False") != 0) {
    svpc_set_error("npy: %s is not C-ordered", path);
    return false;
  }
  k = value_after("'shape'");
  if (k == std::string::npos || h[k] != '(') {
    svpc_set_error("npy: %s has no 'shape'", path);
    return false;
  }
  std::vector<long long> dims;
  ++k;
  while (k < h.size() && h[k] != ')') {
    if (h[k] >= '0' && h[k] <= '9') {
      long long v = 0;
      while (k < h.size() && h[k] >= '0' && h[k] <= '9') v = v * 10 + (h[k++] - '0');
      dims.push_back(v);
    } else {
      ++k;
    }
  }
  if (dims.size() != 2) {
    svpc_set_error("npy: %s has rank %d (need 2)", path, (int)dims.size());
    return false;
  }
  out->rows = dims[0];
  out->cols = dims[1];
  out->data_offset = (long long)(hoff + hlen);
  static const int esz[3] = {4, 8, 2};
  if ((unsigned long long)out->data_offset + (unsigned long long)out->rows * out->cols * esz[out->dtype] > size) {
    svpc_set_error("npy: %s is shorter than its header says", path);
    return false;
  }
  return true;
}

struct Mapped {
  const unsigned char* base = nullptr;
  size_t size = 0;
  NpyInfo info;
};

bool map_file(const char* path, Mapped* m) {
  const int fd = open(path, O_RDONLY);
  if (fd < 0) {
    svpc_set_error("npy: cannot open %s", path);
    return false;
  }
  struct stat st;
  if (fstat(fd, &st) != 0 || st.st_size <= 0) {
    close(fd);
    svpc_set_error("npy: cannot stat %s", path);
    return false;
  }
  void* p = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
  close(fd);
  if (p == MAP_FAILED) {
    svpc_set_error("npy: cannot map %s", path);
    return false;
  }
  m->base = (const unsigned char*)p;
  m->size = (size_t)st.st_size;
  if (!parse_npy_header(m->base, m->size, path, &m->info)) {
    munmap(p, m->size);
    m->base = nullptr;
    return false;
  }
  return true;
}

// fp32 -> bf16, round to nearest even (what a device-side cast of the reference's fp32 block gives)
inline unsigned short f32_to_bf16(float f) {
  unsigned int u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (unsigned short)((u >> 16) | 0x40);     // NaN stays NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return (unsigned short)(u >> 16);
}

inline void convert_row_bf16(const unsigned char* src, int dtype, long long cols, unsigned short* dst) {
  if (dtype == 0) {
    const float* s = (const float*)src;
    for (long long c = 0; c < cols; ++c) dst[c] = f32_to_bf16(s[c]);
  } else if (dtype == 1) {
    const double* s = (const double*)src;
    for (long long c = 0; c < cols; ++c) dst[c] = f32_to_bf16((float)s[c]);
  } else {
    const __half* s = (const __half*)src;
    for (long long c = 0; c < cols; ++c) dst[c] = f32_to_bf16(__half2float(s[c]));
  }
}

inline void convert_row(const unsigned char* src, int dtype, long long cols, float* dst) {
  if (dtype == 0) {
    memcpy(dst, src, (size_t)cols * 4);
  } else if (dtype == 1) {
    const double* s = (const double*)src;
    for (long long c = 0; c < cols; ++c) dst[c] = (float)s[c];           // == ndarray.astype(np.float32): round to nearest
  } else {
    const __half* s = (const __half*)src;
    for (long long c = 0; c < cols; ++c) dst[c] = __half2float(s[c]);
  }
}

}  // namespace

extern "C" int svpc_npy_info(const char* path, long long* rows, long long* cols, int* dtype, long long* data_offset) {
  Mapped m;
  if (!path || !map_file(path, &m)) return SVPC_ERR_ARG;
  if (rows) *rows = m.info.rows;
  if (cols) *cols = m.info.cols;
  if (dtype) *dtype = m.info.dtype;
  if (data_offset) *data_offset = m.info.data_offset;
  munmap((void*)m.base, m.size);
  return SVPC_OK;
}

// One job = one (file, clip): rows[row_start[j] .. row_start[j+1]) of file paths[j] go to rows first_row.. of the clip
// block dst[j] (block_rows rows of dst_pitch floats), columns [col_off[j], col_off[j] + file cols); every other row of
// that column range in the block is zero-filled.  Jobs are spread over n_threads host threads; files named by several
// jobs (the clips of one video) are mapped once.
extern "C" int svpc_features_gather(int n_jobs, const char* const* paths, const long long* row_start, const long long* rows,
                                    void* const* dst, int dst_dtype, long long dst_pitch, const long long* col_off,
                                    int first_row, int block_rows, int n_threads) {
  if (n_jobs <= 0) return SVPC_OK;
  if (!paths || !row_start || !rows || !dst || !col_off || dst_pitch <= 0 || first_row < 0 || block_rows <= 0 ||
      (dst_dtype != SVPC_F32 && dst_dtype != SVPC_BF16)) {
    svpc_set_error("features_gather: bad argument");
    return SVPC_ERR_ARG;
  }
  std::map<std::string, int> index;
  std::vector<Mapped> files;
  std::vector<int> job_file(n_jobs);
  int rc = SVPC_OK;
  for (int j = 0; j < n_jobs && rc == SVPC_OK; ++j) {
    auto it = index.find(paths[j]);
    if (it == index.end()) {
      Mapped m;
      if (!map_file(paths[j], &m)) {
        rc = SVPC_ERR_ARG;
        break;
      }
      it = index.emplace(paths[j], (int)files.size()).first;
      files.push_back(m);
    }
    job_file[j] = it->second;
    const Mapped& m = files[it->second];
    const long long n = row_start[j + 1] - row_start[j];
    if (n < 0 || first_row + n > block_rows || col_off[j] < 0 || col_off[j] + m.info.cols > dst_pitch) {
      svpc_set_error("features_gather: job %d does not fit its block (%lld rows from row %d of %d, cols %lld+%lld of %lld)", j, n,
                     first_row, block_rows, col_off[j], m.info.cols, dst_pitch);
      rc = SVPC_ERR_ARG;
      break;
    }
    for (long long i = row_start[j]; i < row_start[j + 1]; ++i)
      if (rows[i] < 0 || rows[i] >= m.info.rows) {
        svpc_set_error("features_gather: row %lld out of range for %s (%lld rows)", rows[i], paths[j], m.info.rows);
        rc = SVPC_ERR_ARG;
        break;
      }
  }
  if (rc == SVPC_OK) {
    static const int esz[3] = {4, 8, 2};
    std::atomic<int> next{0};
    auto work = [&]() {
      for (;;) {
        const int j = next.fetch_add(1, std::memory_order_relaxed);
        if (j >= n_jobs) break;
        const Mapped& m = files[job_file[j]];
        const long long cols = m.info.cols, n = row_start[j + 1] - row_start[j];
        const size_t row_bytes = (size_t)cols * esz[m.info.dtype];
        const unsigned char* data = m.base + m.info.data_offset;
        if (dst_dtype == SVPC_F32) {
          float* blk = (float*)dst[j] + col_off[j];
          for (int r = 0; r < first_row; ++r) memset(blk + (long long)r * dst_pitch, 0, (size_t)cols * 4);
          for (long long i = 0; i < n; ++i)
            convert_row(data + (size_t)rows[row_start[j] + i] * row_bytes, m.info.dtype, cols, blk + (first_row + i) * dst_pitch);
          for (long long r = first_row + n; r < block_rows; ++r) memset(blk + r * dst_pitch, 0, (size_t)cols * 4);
        } else {
          unsigned short* blk = (unsigned short*)dst[j] + col_off[j];
          for (int r = 0; r < first_row; ++r) memset(blk + (long long)r * dst_pitch, 0, (size_t)cols * 2);
          for (long long i = 0; i < n; ++i)
            convert_row_bf16(data + (size_t)rows[row_start[j] + i] * row_bytes, m.info.dtype, cols, blk + (first_row + i) * dst_pitch);
          for (long long r = first_row + n; r < block_rows; ++r) memset(blk + r * dst_pitch, 0, (size_t)cols * 2);
        }
      }
    };
    const int nt = n_threads < 1 ? 1 : (n_threads > n_jobs ? n_jobs : n_threads);
    if (nt == 1) {
      work();
    } else {
      std::vector<std::thread> pool;
      for (int t = 1; t < nt; ++t) pool.emplace_back(work);
      work();
      for (auto& t : pool) t.join();
    }
  }
  for (auto& m : files) munmap((void*)m.base, m.size);
  return rc;
}
