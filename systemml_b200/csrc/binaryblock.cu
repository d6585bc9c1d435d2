// binaryblock.cu -- SystemML "binaryblock" matrices (Hadoop SequenceFile of MatrixIndexes -> MatrixBlock records, 1000 x 1000
// blocks) straight into row-sharded device buffers.
//
// Reference: ReaderBinaryBlock.readBinaryBlockMatrixFromSequenceFile (J/runtime/io/ReaderBinaryBlock.java:124-207) and
// ReaderBinaryBlockParallel read the records into a JVM MatrixBlock (double[] on the heap, capped at 2^31 cells per dense
// block array), which GPUObject.copyFromHostToDevice later uploads in one synchronous cudaMemcpy (GPUObject.java:755-870).
// The record layouts are MatrixIndexes.write (J/runtime/matrix/data/MatrixIndexes.java:117-130: two big-endian longs,
// 1-based block row / column) and MatrixBlock.write (MatrixBlock.java:1924-2180): int rlen, int clen, byte type, then
//   DENSE_BLOCK (3)        rlen*clen big-endian doubles, row-major
//   SPARSE_BLOCK (2)       nnz (int, or long when rlen*clen > 2^31-1), then per row: int count, count x (int col, double val)
//   ULTRA_SPARSE_BLOCK (1) int nnz, then nnz x (int row, int col, double val)  [clen == 1: (int row, double val)]
//   EMPTY_BLOCK (0)        nothing
// The container is Hadoop's SequenceFile, version 6, uncompressed, as SequenceFile.Writer(fs, job, path, MatrixIndexes.class,
// MatrixBlock.class) produces it (WriterBinaryBlock.java:101-125).  Hadoop itself is a dependency that is not part of
// the reference tree (org.apache.hadoop:hadoop-common 2.6.0, pom.xml:64); its published format (the class comment of
// org.apache.hadoop.io.SequenceFile) is restated here: "SEQ" 6, key and value class names as Text (vint length + UTF-8),
// two booleans (compression, block compression), metadata (int count + Text pairs), a 16-byte sync marker; then records
// [int recordLength][int keyLength][key][value], with [int -1][sync marker] escapes in between.
//
// Here the file never becomes a host matrix.  The raw bytes of a dense block travel as they are: read() into a pinned
// staging slot, one asynchronous host-to-device copy, and a kernel that swaps the byte order and places the block into the
// row-major device matrix of this rank's row range -- the CPU touches no double.  Slots form a ring (file read of block i+1
// overlaps the copy and the placement of block i).  Sparse, ultra-sparse and empty blocks are expanded on the host into
// the slot (native byte order) and take the same copy + placement path without the swap.  With a callback per completed
// block row, the first pass of an algorithm starts on the rows that have arrived while the rest is still in flight.
#include <dirent.h>
#include <errno.h>
#include <stdlib.h>
#include <sys/stat.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "common.cuh"

extern "C" int sysml_b200_mmchain_dense(sysml_b200_ctx* ctx, const double* X, const double* v, const double* w, double* out, int64_t n,
                                        int64_t m, int chain_type);
int sysml_reduce_partial_rows(sysml_b200_ctx* ctx, const double* partial, int P, int m, double* out);

namespace {

const char* KEY_CLASS = "org.apache.sysml.runtime.matrix.data.MatrixIndexes";
const char* VAL_CLASS = "org.apache.sysml.runtime.matrix.data.MatrixBlock";
enum { BT_EMPTY = 0, BT_ULTRA = 1, BT_SPARSE = 2, BT_DENSE = 3 };
constexpr i64 INT_MAX64 = 0x7fffffffLL;

// ---------------- big-endian primitives (java.io.DataOutput / DataInput) ----------------
inline void put_be32(std::vector<unsigned char>& b, uint32_t v) {
  for (int s = 24; s >= 0; s -= 8) b.push_back((unsigned char)(v >> s));
}
inline void put_be64(std::vector<unsigned char>& b, uint64_t v) {
  for (int s = 56; s >= 0; s -= 8) b.push_back((unsigned char)(v >> s));
}
inline uint32_t get_be32(const unsigned char* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
inline uint64_t get_be64(const unsigned char* p) { return ((uint64_t)get_be32(p) << 32) | get_be32(p + 4); }
inline double be_double(const unsigned char* p) {
  uint64_t u = get_be64(p);
  double d;
  memcpy(&d, &u, 8);
  return d;
}
inline void put_double(std::vector<unsigned char>& b, double d) {
  uint64_t u;
  memcpy(&u, &d, 8);
  put_be64(b, u);
}
// org.apache.hadoop.io.WritableUtils.writeVLong / readVLong
inline void put_vlong(std::vector<unsigned char>& b, i64 i) {
  if (i >= -112 && i <= 127) {
    b.push_back((unsigned char)(signed char)i);
    return;
  }
  int len = -112;
  if (i < 0) {
    i ^= -1LL;
    len = -120;
  }
  for (i64 tmp = i; tmp != 0; tmp >>= 8) len--;
  b.push_back((unsigned char)(signed char)len);
  len = (len < -120) ? -(len + 120) : -(len + 112);
  for (int idx = len; idx != 0; idx--) b.push_back((unsigned char)((i >> ((idx - 1) * 8)) & 0xff));
}
inline void put_text(std::vector<unsigned char>& b, const char* s) {
  const size_t n = strlen(s);
  put_vlong(b, (i64)n);
  b.insert(b.end(), s, s + n);
}

struct Reader {
  FILE* f = nullptr;
  std::string err;
  bool read(void* dst, size_t n) {
    if (n == 0) return true;
    if (fread(dst, 1, n, f) != n) {
      err = "unexpected end of file";
      return false;
    }
    return true;
  }
  bool vlong(i64* out) {
    signed char first;
    if (!read(&first, 1)) return false;
    if (first >= -112) {
      *out = first;
      return true;
    }
    const int len = (first < -120) ? (-119 - first) : (-111 - first);  // total size including the first byte
    i64 v = 0;
    for (int i = 0; i < len - 1; i++) {
      unsigned char c;
      if (!read(&c, 1)) return false;
      v = (v << 8) | c;
    }
    const bool neg = first < -120;
    *out = neg ? (v ^ -1LL) : v;
    return true;
  }
  bool text(std::string* out) {
    i64 n;
    if (!vlong(&n) || n < 0 || n > 4096) {
      if (err.empty()) err = "bad Text length";
      return false;
    }
    out->resize((size_t)n);
    return read(&(*out)[0], (size_t)n);
  }
};

// MatrixBlock.evalSparseFormatOnDisk / isUltraSparseSerialize (MatrixBlock.java:897-899, 1017-1031, 2387-2422)
inline void on_disk_format(i64 rlen, i64 clen, i64 nnz, int* type) {
  if (nnz == 0) {
    *type = BT_EMPTY;
    return;
  }
  const double sp = ((double)nnz / (double)rlen) / (double)clen;
  const i64 sizeDense = 9 + rlen * clen * 8;
  const i64 sizeSparse = 9 + ((rlen * clen > INT_MAX64) ? 8 : 4) + rlen * 4 + nnz * 12;
  const i64 sizeUltra = 9 + 4 + nnz * (clen > 1 ? 16 : 12);
  const bool sparseDst = sp < 0.4 && (sizeSparse < sizeDense || sizeUltra < sizeDense);
  if (nnz < rlen && sparseDst) *type = BT_ULTRA;
  else if (sparseDst) *type = BT_SPARSE;
  else *type = BT_DENSE;
}

// ---------------- device side: raw block -> its place in the row-major matrix ----------------
// src: rlen*clen doubles, big-endian (swap = 1) or native (swap = 0); dst: &X[(row0 - row_begin) * cols + col0], pitch cols
__global__ void bb_place_kernel(const unsigned long long* __restrict__ src, double* __restrict__ dst, int rlen, int clen, i64 pitch, int r_lo,
                                int r_hi, int swap) {
  const i64 total = (i64)(r_hi - r_lo) * clen;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int r = (int)(i / clen) + r_lo, c = (int)(i % clen);
    unsigned long long u = src[(i64)r * clen + c];
    if (swap) {
      const unsigned lo = (unsigned)u, hi = (unsigned)(u >> 32);
      u = ((unsigned long long)__byte_perm(lo, 0, 0x0123) << 32) | __byte_perm(hi, 0, 0x0123);
    }
    dst[(i64)(r - r_lo) * pitch + c] = __longlong_as_double((long long)u);
  }
}

struct Slot {
  unsigned char* host = nullptr;  // pinned
  unsigned char* dev = nullptr;
  cudaEvent_t done = nullptr;     // the placement kernel that read `dev` (and the copy that read `host`) has finished
  bool used = false;
};

// what the reader threads share: per block row of the range, the blocks still to arrive; the callback runs under the lock
// (the context is not thread-safe, and the stripes must be handed over one at a time)
struct Progress {
  std::mutex mu;
  std::vector<int> blocks_left;
};

struct Loader {
  sysml_b200_ctx* ctx;
  i64 rows, cols, row_begin, row_end;
  int brlen, bclen;
  double* dX;
  Progress* prog = nullptr;
  std::vector<Slot> slots;
  size_t slot_bytes = 0;
  int next = 0;
  cudaStream_t copy_stream = nullptr;
  i64 nblocks_seen = 0;
  std::string err;

  int init(int nslots) {
    slot_bytes = (size_t)brlen * bclen * 8;
    slots.resize(nslots);
    if (cudaStreamCreateWithFlags(&copy_stream, cudaStreamNonBlocking) != cudaSuccess) return SYSML_E_CUDA;
    for (auto& s : slots) {
      if (cudaMallocHost((void**)&s.host, slot_bytes) != cudaSuccess) return SYSML_E_OOM;
      if (cudaMalloc((void**)&s.dev, slot_bytes) != cudaSuccess) return SYSML_E_OOM;
      if (cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming) != cudaSuccess) return SYSML_E_CUDA;
    }
    return SYSML_OK;
  }
  void destroy() {
    if (copy_stream) cudaStreamSynchronize(copy_stream);
    for (auto& s : slots) {
      if (s.done) cudaEventDestroy(s.done);
      if (s.host) cudaFreeHost(s.host);
      if (s.dev) cudaFree(s.dev);
    }
    if (copy_stream) cudaStreamDestroy(copy_stream);
  }
  Slot* acquire() {
    Slot* s = &slots[next];
    next = (next + 1) % (int)slots.size();
    if (s->used) cudaEventSynchronize(s->done);  // its previous block has left the slot
    s->used = true;
    return s;
  }
  // the slot holds rlen*clen doubles (big-endian when swap); place rows that fall into [row_begin, row_end)
  int submit(Slot* s, i64 bi, i64 bj, int rlen, int clen, int swap) {
    const i64 row0 = bi * brlen, col0 = bj * bclen;
    const i64 lo = std::max(row0, row_begin), hi = std::min(row0 + rlen, row_end);
    if (hi <= lo) return SYSML_OK;
    const size_t bytes = (size_t)rlen * clen * 8;
    if (cudaMemcpyAsync(s->dev, s->host, bytes, cudaMemcpyHostToDevice, copy_stream) != cudaSuccess) return SYSML_E_CUDA;
    const i64 total = (hi - lo) * clen;
    i64 blocks = (total + 255) / 256;
    if (blocks > (i64)ctx->sm_count * 8) blocks = (i64)ctx->sm_count * 8;
    bb_place_kernel<<<(unsigned)blocks, 256, 0, copy_stream>>>((const unsigned long long*)s->dev, dX + (lo - row_begin) * cols + col0, rlen, clen, cols,
                                                               (int)(lo - row0), (int)(hi - row0), swap);
    if (cudaGetLastError() != cudaSuccess) return SYSML_E_CUDA;
    __sync_fetch_and_add(&ctx->launches, 1);  // reader threads share the context's counter
    if (cudaEventRecord(s->done, copy_stream) != cudaSuccess) return SYSML_E_CUDA;
    return SYSML_OK;
  }
};

// complete: the last block of that block row; otherwise only an ordering point (make the compute stream wait on ready_on)
typedef int (*stripe_cb)(void* user, i64 first_row, i64 n_rows, cudaStream_t ready_on, bool complete);

// reads one SequenceFile; every block that intersects the row range is placed on the device
int read_file(Loader& L, const std::string& path, stripe_cb cb, void* user) {
  Reader R;
  R.f = fopen(path.c_str(), "rb");
  if (!R.f) {
    L.err = "cannot open " + path + ": " + strerror(errno);
    return SYSML_E_INVALID;
  }
  struct Closer {
    FILE* f;
    ~Closer() { fclose(f); }
  } closer{R.f};
  unsigned char hdr[4];
  std::string kc, vc;
  unsigned char flags[2], sync[16], cnt4[4];
  if (!R.read(hdr, 4) || hdr[0] != 'S' || hdr[1] != 'E' || hdr[2] != 'Q') {
    L.err = path + ": not a SequenceFile";
    return SYSML_E_INVALID;
  }
  if (hdr[3] != 6) {
    L.err = path + ": SequenceFile version " + std::to_string((int)hdr[3]) + " (only version 6 is understood)";
    return SYSML_E_UNSUPPORTED;
  }
  if (!R.text(&kc) || !R.text(&vc) || !R.read(flags, 2)) {
    L.err = path + ": " + R.err;
    return SYSML_E_INVALID;
  }
  if (kc != KEY_CLASS || vc != VAL_CLASS) {
    L.err = path + ": key/value classes " + kc + " / " + vc + " are not MatrixIndexes / MatrixBlock";
    return SYSML_E_INVALID;
  }
  if (flags[0] || flags[1]) {
    L.err = path + ": compressed SequenceFiles are not supported";
    return SYSML_E_UNSUPPORTED;
  }
  if (!R.read(cnt4, 4)) return SYSML_E_INVALID;
  for (uint32_t i = 0, n = get_be32(cnt4); i < n; i++) {
    std::string a, b;
    if (!R.text(&a) || !R.text(&b)) return SYSML_E_INVALID;
  }
  if (!R.read(sync, 16)) return SYSML_E_INVALID;
  std::vector<unsigned char> small;  // non-dense payloads
  for (;;) {
    unsigned char len4[4];
    if (fread(len4, 1, 4, R.f) != 4) break;  // clean end of file
    int32_t reclen = (int32_t)get_be32(len4);
    if (reclen == -1) {
      unsigned char s2[16];
      if (!R.read(s2, 16) || memcmp(s2, sync, 16) != 0) {
        L.err = path + ": corrupt sync marker";
        return SYSML_E_INVALID;
      }
      continue;
    }
    unsigned char kl4[4], key[16], bh[9];
    if (!R.read(kl4, 4) || get_be32(kl4) != 16 || !R.read(key, 16) || !R.read(bh, 9)) {
      L.err = path + ": truncated record";
      return SYSML_E_INVALID;
    }
    const i64 bi = (i64)get_be64(key) - 1, bj = (i64)get_be64(key + 8) - 1;
    const int rlen = (int)get_be32(bh), clen = (int)get_be32(bh + 4), type = bh[8];
    const i64 payload = (i64)reclen - 16 - 9;
    if (bi < 0 || bj < 0 || rlen < 0 || clen < 0 || rlen > L.brlen || clen > L.bclen || bi * L.brlen + rlen > L.rows || bj * L.bclen + clen > L.cols ||
        payload < 0) {
      L.err = path + ": block (" + std::to_string(bi + 1) + "," + std::to_string(bj + 1) + ") of " + std::to_string(rlen) + " x " + std::to_string(clen) +
              " does not fit the declared matrix";
      return SYSML_E_DIM;
    }
    const i64 row0 = bi * L.brlen;
    const bool wanted = row0 < L.row_end && row0 + rlen > L.row_begin;
    if (!wanted) {
      if (fseeko(R.f, (off_t)payload, SEEK_CUR) != 0) return SYSML_E_INVALID;
      continue;
    }
    Slot* s = L.acquire();
    int swap = 0;
    if (type == BT_DENSE) {
      if (payload != (i64)rlen * clen * 8 || !R.read(s->host, (size_t)payload)) {
        L.err = path + ": dense block with a wrong payload size";
        return SYSML_E_INVALID;
      }
      swap = 1;
    } else {
      small.resize((size_t)payload);
      if (!R.read(small.data(), (size_t)payload)) return SYSML_E_INVALID;
      double* d = (double*)s->host;
      memset(d, 0, (size_t)rlen * clen * 8);
      const unsigned char* p = small.data();
      const unsigned char* end = p + payload;
      if (type == BT_SPARSE) {
        p += ((i64)rlen * clen > INT_MAX64) ? 8 : 4;  // nnz: not needed to expand
        for (int r = 0; r < rlen; r++) {
          if (p + 4 > end) return SYSML_E_INVALID;
          const int nr = (int)get_be32(p);
          p += 4;
          if (nr < 0 || p + (i64)nr * 12 > end) return SYSML_E_INVALID;
          for (int j = 0; j < nr; j++, p += 12) {
            const int c = (int)get_be32(p);
            if (c < 0 || c >= clen) return SYSML_E_INVALID;
            d[(size_t)r * clen + c] = be_double(p + 4);
          }
        }
      } else if (type == BT_ULTRA) {
        if (p + 4 > end) return SYSML_E_INVALID;
        const int nnz = (int)get_be32(p);
        p += 4;
        const int step = clen > 1 ? 16 : 12;
        if (nnz < 0 || p + (i64)nnz * step > end) return SYSML_E_INVALID;
        for (int q = 0; q < nnz; q++, p += step) {
          const int r = (int)get_be32(p), c = clen > 1 ? (int)get_be32(p + 4) : 0;
          if (r < 0 || r >= rlen || c < 0 || c >= clen) return SYSML_E_INVALID;
          d[(size_t)r * clen + c] = be_double(p + (clen > 1 ? 8 : 4));
        }
      } else if (type != BT_EMPTY) {
        L.err = path + ": invalid block format " + std::to_string(type);
        return SYSML_E_INVALID;
      }
    }
    int rc = L.submit(s, bi, bj, rlen, clen, swap);
    if (rc) return rc;
    L.nblocks_seen++;
    // a whole block row of the range has been queued: hand it to the caller (the work it queues waits on the copy stream).
    // Blocks of one block row may come from different reader threads: each thread's stream is waited for as its last
    // block of the row is counted, so by the time the count reaches zero every contributor has been ordered.
    const i64 idx = bi - L.row_begin / L.brlen;
    if (idx >= 0 && idx < (i64)L.prog->blocks_left.size()) {
      std::lock_guard<std::mutex> lock(L.prog->mu);
      const int left = --L.prog->blocks_left[(size_t)idx];
      if (cb) {
        const i64 lo = std::max(row0, L.row_begin), hi = std::min(row0 + rlen, L.row_end);
        rc = cb(user, lo - L.row_begin, hi - lo, L.copy_stream, left == 0);
        if (rc) return rc;
      }
    }
  }
  return SYSML_OK;
}

int list_parts(const std::string& path, std::vector<std::string>* out, std::string* err) {
  struct stat st;
  if (stat(path.c_str(), &st) != 0) {
    *err = "cannot stat " + path + ": " + strerror(errno);
    return SYSML_E_INVALID;
  }
  if (!S_ISDIR(st.st_mode)) {
    out->push_back(path);
    return SYSML_OK;
  }
  DIR* d = opendir(path.c_str());
  if (!d) return SYSML_E_INVALID;
  while (struct dirent* e = readdir(d)) {  // IOUtilFunctions.getSequenceFilePaths: every file except _SUCCESS, _logs, .crc and hidden ones
    const std::string n = e->d_name;
    if (n.empty() || n[0] == '.' || n[0] == '_') continue;
    out->push_back(path + "/" + n);
  }
  closedir(d);
  std::sort(out->begin(), out->end());
  return SYSML_OK;
}

int load(sysml_b200_ctx* ctx, const char* path, i64 rows, i64 cols, int brlen, int bclen, i64 row_begin, i64 row_end, double* dX, stripe_cb cb,
         void* user) {
  if (!path || rows < 0 || cols < 0 || brlen < 1 || bclen < 1 || row_begin < 0 || row_end > rows || row_begin > row_end)
    return sysml_set_error(ctx, SYSML_E_INVALID, "binaryblock: bad arguments");
  if ((i64)brlen * bclen > (1 << 26)) return sysml_set_error(ctx, SYSML_E_LIMIT, "binaryblock: blocks of more than 2^26 cells");
  if (row_begin % brlen != 0) return sysml_set_error(ctx, SYSML_E_INVALID, "binaryblock: the row range must start at a block boundary");
  if (row_end > row_begin && cols > 0 && !dX) return sysml_set_error(ctx, SYSML_E_INVALID, "binaryblock: null destination");
  if (row_end == row_begin || cols == 0) return SYSML_OK;
  std::vector<std::string> parts;
  std::string err;
  int rc = list_parts(path, &parts, &err);
  if (rc) return sysml_set_error(ctx, rc, "binaryblock: %s", err.c_str());
  Progress prog;
  {
    const i64 br0 = row_begin / brlen, br1 = (row_end + brlen - 1) / brlen;
    prog.blocks_left.assign((size_t)std::max<i64>(br1 - br0, 0), (int)((cols + bclen - 1) / bclen));
  }
  // blocks that never appear in the file are empty blocks (writers may skip them): the destination starts at zero
  SYSML_CUDA(ctx, cudaMemsetAsync(dX, 0, (size_t)(row_end - row_begin) * cols * sizeof(double), ctx->stream));
  SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  // part files are read in parallel (ReaderBinaryBlockParallel reads them with one task each), every reader with its own
  // staging ring and copy stream
  // (at most 8 readers: measured on a 16-CPU host, 16 part files, 4 GB: 8 readers 11.4 GB/s, 16 readers 5.2 GB/s)
  int cap = 8;
  if (const char* e = getenv("SYSML_B200_BB_READERS")) cap = atoi(e);  // A/B switch for measurements
  cap = cap < 1 ? 1 : (cap > 16 ? 16 : cap);
  const int nthreads = (int)std::min<size_t>(parts.size(), (size_t)cap);
  std::vector<int> trc((size_t)nthreads, SYSML_OK);
  std::vector<std::string> terr((size_t)nthreads);
  auto work = [&](int t) {
    cudaSetDevice(ctx->device);
    Loader L{ctx, rows, cols, row_begin, row_end, brlen, bclen, dX};
    L.prog = &prog;
    int r = L.init(4);
    if (r) L.err = "cannot set up the staging ring";
    for (size_t i = (size_t)t; !r && i < parts.size(); i += (size_t)nthreads) r = read_file(L, parts[i], cb, user);
    if (L.copy_stream) {
      cudaError_t e = cudaStreamSynchronize(L.copy_stream);
      if (!r && e != cudaSuccess) {
        r = SYSML_E_CUDA;
        L.err = cudaGetErrorString(e);
      }
    }
    trc[(size_t)t] = r;
    terr[(size_t)t] = L.err;
    L.destroy();
  };
  if (nthreads <= 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++) th.emplace_back(work, t);
    for (auto& x : th) x.join();
  }
  sysml_set_device(ctx);
  for (int t = 0; t < nthreads; t++)
    if (trc[(size_t)t]) return sysml_set_error(ctx, trc[(size_t)t], "binaryblock: %s", terr[(size_t)t].empty() ? "read failed" : terr[(size_t)t].c_str());
  // block rows with skipped (empty) blocks never reached zero above: report them now, in order
  if (cb) {
    const i64 br0 = row_begin / brlen;
    for (size_t i = 0; i < prog.blocks_left.size(); i++)
      if (prog.blocks_left[i] > 0) {
        const i64 lo = std::max((br0 + (i64)i) * brlen, row_begin), hi = std::min((br0 + (i64)i + 1) * brlen, row_end);
        rc = cb(user, lo - row_begin, hi - lo, ctx->stream, true);
        if (rc) return rc;
      }
  }
  // later work on the context's stream sees the matrix
  SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SYSML_OK;
}

// The first pass overlapped with the upload: block rows are grouped into fixed row ranges of ~64 MB; when the last block row
// of a group has arrived, mmchain runs on the group and writes ITS partial result into the group's own slot.  The slots
// are added in group order at the end (fixed-order compensated reduction), so the result does not depend on which reader
// thread was faster.
struct ChainState {
  sysml_b200_ctx* ctx;
  const double* dX;
  const double* v;
  const double* w;
  double* partials;  // [ngroups][cols]
  i64 cols, nrows;
  int ct, brlen;
  i64 group_rows;    // multiple of brlen
  std::vector<int> stripes_left;
  cudaEvent_t ev;
};

int chain_cb(void* user, i64 first_row, i64 n_rows, cudaStream_t ready_on, bool complete) {
  ChainState* st = (ChainState*)user;
  // this reader's copies and placements of the stripe are queued on its stream: the compute stream waits for them
  if (ready_on != st->ctx->stream) {
    if (cudaEventRecord(st->ev, ready_on) != cudaSuccess || cudaStreamWaitEvent(st->ctx->stream, st->ev, 0) != cudaSuccess) return SYSML_E_CUDA;
  }
  if (!complete) return SYSML_OK;
  const i64 g = first_row / st->group_rows;
  if (--st->stripes_left[(size_t)g] > 0) return SYSML_OK;
  const i64 lo = g * st->group_rows, hi = std::min(lo + st->group_rows, st->nrows);
  return sysml_b200_mmchain_dense(st->ctx, st->dX + lo * st->cols, st->v, st->w ? st->w + lo : nullptr, st->partials + (size_t)g * st->cols, hi - lo,
                                  st->cols, st->ct);
}

}  // namespace

extern "C" {

int sysml_b200_binaryblock_write(const char* path, const double* h_X, int64_t rows, int64_t cols, int brlen, int bclen, int dense_only,
                                 int64_t first_block_row) {
  if (!path || rows < 0 || cols < 0 || brlen < 1 || bclen < 1 || first_block_row < 0 || (rows * cols > 0 && !h_X))
    return sysml_set_error(nullptr, SYSML_E_INVALID, "binaryblock_write: bad arguments");
  FILE* f = fopen(path, "wb");
  if (!f) return sysml_set_error(nullptr, SYSML_E_INVALID, "binaryblock_write: cannot create %s: %s", path, strerror(errno));
  std::vector<unsigned char> b;
  b.push_back('S'); b.push_back('E'); b.push_back('Q'); b.push_back(6);
  put_text(b, KEY_CLASS);
  put_text(b, VAL_CLASS);
  b.push_back(0);
  b.push_back(0);
  put_be32(b, 0);  // no metadata
  unsigned char sync[16];
  for (int i = 0; i < 16; i++) sync[i] = (unsigned char)(0x5a ^ (i * 37 + (int)(rows % 251) + (int)(cols % 241)));
  b.insert(b.end(), sync, sync + 16);
  bool ok = fwrite(b.data(), 1, b.size(), f) == b.size();
  i64 pos = (i64)b.size(), last_sync = pos;
  const i64 nbr = (rows + brlen - 1) / brlen, nbc = (cols + bclen - 1) / bclen;
  for (i64 bi = 0; ok && bi < nbr; bi++)
    for (i64 bj = 0; ok && bj < nbc; bj++) {
      const int rlen = (int)std::min<i64>(brlen, rows - bi * brlen), clen = (int)std::min<i64>(bclen, cols - bj * bclen);
      const double* src = h_X + bi * brlen * cols + bj * bclen;
      i64 nnz = 0;
      for (int r = 0; r < rlen; r++)
        for (int c = 0; c < clen; c++) nnz += src[(i64)r * cols + c] != 0.0;
      int type = BT_DENSE;
      if (!dense_only) on_disk_format(rlen, clen, nnz, &type);
      b.clear();
      put_be64(b, (uint64_t)(first_block_row + bi + 1));
      put_be64(b, (uint64_t)(bj + 1));
      put_be32(b, (uint32_t)rlen);
      put_be32(b, (uint32_t)clen);
      b.push_back((unsigned char)type);
      if (type == BT_DENSE) {
        const size_t off = b.size();
        b.resize(off + (size_t)rlen * clen * 8);
        unsigned char* q = b.data() + off;
        for (int r = 0; r < rlen; r++) {
          const double* row = src + (i64)r * cols;
          for (int c = 0; c < clen; c++, q += 8) {
            uint64_t u;
            memcpy(&u, row + c, 8);
            u = __builtin_bswap64(u);
            memcpy(q, &u, 8);
          }
        }
      } else if (type == BT_SPARSE) {
        if ((i64)rlen * clen > INT_MAX64) put_be64(b, (uint64_t)nnz);
        else put_be32(b, (uint32_t)nnz);
        for (int r = 0; r < rlen; r++) {
          int nr = 0;
          for (int c = 0; c < clen; c++) nr += src[(i64)r * cols + c] != 0.0;
          put_be32(b, (uint32_t)nr);
          for (int c = 0; c < clen; c++)
            if (src[(i64)r * cols + c] != 0.0) {
              put_be32(b, (uint32_t)c);
              put_double(b, src[(i64)r * cols + c]);
            }
        }
      } else if (type == BT_ULTRA) {
        put_be32(b, (uint32_t)nnz);
        for (int r = 0; r < rlen; r++)
          for (int c = 0; c < clen; c++)
            if (src[(i64)r * cols + c] != 0.0) {
              put_be32(b, (uint32_t)r);
              if (clen > 1) put_be32(b, (uint32_t)c);
              put_double(b, src[(i64)r * cols + c]);
            }
      }
      std::vector<unsigned char> head;
      if (pos >= last_sync + 2000) {  // SequenceFile.Writer.checkAndWriteSync: SYNC_INTERVAL = 100 * SYNC_SIZE
        put_be32(head, 0xffffffffu);
        head.insert(head.end(), sync, sync + 16);
        last_sync = pos + 20;
      }
      put_be32(head, (uint32_t)b.size());  // record length = key + value
      put_be32(head, 16);                  // key length
      ok = fwrite(head.data(), 1, head.size(), f) == head.size() && fwrite(b.data(), 1, b.size(), f) == b.size();
      pos += (i64)head.size() + (i64)b.size();
    }
  ok = (fclose(f) == 0) && ok;
  if (!ok) return sysml_set_error(nullptr, SYSML_E_INVALID, "binaryblock_write: write to %s failed", path);
  return SYSML_OK;
}

int sysml_b200_binaryblock_read_dense(sysml_b200_ctx* ctx, const char* path, int64_t rows, int64_t cols, int brlen, int bclen, int64_t row_begin,
                                      int64_t row_end, double* d_X) {
  SYSML_ENTER(ctx);
  return load(ctx, path, rows, cols, brlen, bclen, row_begin, row_end, d_X, nullptr, nullptr);
}

int sysml_b200_binaryblock_read_mmchain(sysml_b200_ctx* ctx, const char* path, int64_t rows, int64_t cols, int brlen, int bclen, int64_t row_begin,
                                        int64_t row_end, double* d_X, const double* v, const double* w, double* out, int chain_type) {
  SYSML_ENTER(ctx);
  if (cols < 0 || (cols > 0 && (!v || !out))) return sysml_set_error(ctx, SYSML_E_INVALID, "binaryblock_read_mmchain: null vector");
  if (chain_type < 0 || chain_type > 2 || (chain_type != SYSML_CHAIN_XtXv && row_end > row_begin && !w))
    return sysml_set_error(ctx, SYSML_E_INVALID, "binaryblock_read_mmchain: bad chain type / missing w");
  if (cols == 0) return SYSML_OK;
  const i64 nrows = row_end - row_begin;
  if (nrows == 0) {
    SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, (size_t)cols * sizeof(double), ctx->stream));
    return SYSML_OK;
  }
  ChainState st{};
  st.ctx = ctx; st.dX = d_X; st.v = v; st.w = w; st.cols = cols; st.nrows = nrows; st.ct = chain_type; st.brlen = brlen;
  const i64 want = std::max<i64>(((i64)64 << 20) / (8 * std::max<i64>(cols, 1)), 1);
  st.group_rows = ((want + brlen - 1) / brlen) * brlen;
  const i64 ngroups = (nrows + st.group_rows - 1) / st.group_rows;
  if (ngroups > 0x7fffffffLL) return sysml_set_error(ctx, SYSML_E_LIMIT, "binaryblock_read_mmchain: too many row groups");
  st.stripes_left.resize((size_t)ngroups);
  for (i64 g = 0; g < ngroups; g++) {
    const i64 lo = g * st.group_rows, hi = std::min(lo + st.group_rows, nrows);
    st.stripes_left[(size_t)g] = (int)((hi - lo + brlen - 1) / brlen);
  }
  SYSML_CUDA(ctx, cudaMalloc((void**)&st.partials, (size_t)ngroups * cols * sizeof(double)));
  if (cudaEventCreateWithFlags(&st.ev, cudaEventDisableTiming) != cudaSuccess) {
    cudaFree(st.partials);
    return sysml_set_error(ctx, SYSML_E_CUDA, "binaryblock_read_mmchain: cannot create an event");
  }
  int rc = load(ctx, path, rows, cols, brlen, bclen, row_begin, row_end, d_X, chain_cb, &st);
  if (!rc) rc = sysml_reduce_partial_rows(ctx, st.partials, (int)ngroups, (int)cols, out);
  cudaStreamSynchronize(ctx->stream);
  cudaEventDestroy(st.ev);
  cudaFree(st.partials);
  return rc;
}

}  // extern "C"
