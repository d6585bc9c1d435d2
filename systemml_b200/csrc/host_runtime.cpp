// host_runtime.cpp -- C++ mirror of SystemML's GPU execution layer (see include/sysml_b200_host.h).
//
//   GPUInstructionParser, *GPUInstruction.processInstruction   J/runtime/instructions/{GPUInstructionParser.java, gpu/*.java}
//   ExecutionContext.*ForGPUInstruction                        J/runtime/controlprogram/context/ExecutionContext.java:304-452
//   GPUObject                                                  J/runtime/instructions/gpu/context/GPUObject.java
//   GPUMemoryManager                                           J/runtime/instructions/gpu/context/GPUMemoryManager.java
//
// Host logic only: every numeric operation is a call into the kernels of include/sysml_b200.h.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <memory>
#include <set>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/sysml_b200.h"
#include "../../include/sysml_b200_host.h"

namespace {

typedef long long i64;

struct DMLRuntimeException : std::runtime_error {
  int code;
  DMLRuntimeException(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};
[[noreturn]] void fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  throw DMLRuntimeException(code, buf);
}

struct GPUObject;

// ------------------------------------------------------------------------------------------------
// GPUMemoryManager: size-keyed free list of "rmvar-ed" pointers, eviction of unlocked matrices when the budget is
// exhausted (GPUMemoryManager.java:157-262) in the order the configured policy gives: LRU / LFU / MRU through the
// timestamp every release updates (GPUObject.updateReleaseLocks, GPUObject.java:650-667), MIN_EVICT by size
// against the request (GPUComparator, GPUMemoryManager.java:531-571).
// Deviation: no cudaMemset after malloc (every kernel writes all of its output cells).
// ------------------------------------------------------------------------------------------------
enum EvictionPolicy { LRU = 0, LFU = 1, MIN_EVICT = 2, MRU = 3 };  // DMLScript.EvictionPolicy (api/DMLScript.java:110)

struct GPUMemoryManager {
  sysml_b200_ctx* ctx = nullptr;
  EvictionPolicy policy = LRU;
  double util_factor = 0.9;
  bool eager_free = false;
  i64 memory_limit = 0;
  std::map<void*, i64> allocated;                 // allocatedGPUPointers
  std::map<i64, std::set<void*>> rmvar;           // rmvarGPUPointers
  std::vector<GPUObject*> objects;                // allocatedGPUObjects
  i64 n_malloc = 0, n_reuse = 0, n_free = 0, n_evict = 0;

  void check(int rc) {
    if (rc) fail(rc, "GPU : %s", sysml_b200_last_error(ctx));
  }
  i64 allocated_bytes() const {
    i64 s = 0;
    for (auto& kv : allocated) s += kv.second;
    return s;
  }
  i64 rmvar_bytes() const {
    i64 s = 0;
    for (auto& kv : rmvar) s += kv.first * (i64)kv.second.size();
    return s;
  }
  i64 getAvailableMemory() {  // GPUMemoryManager.java:521-526 (cudaMemGetInfo * GPU_MEMORY_UTILIZATION_FACTOR)
    if (memory_limit > 0) return memory_limit - allocated_bytes();
    int64_t f = 0, t = 0;
    check(sysml_b200_mem_info(ctx, &f, &t));
    return (i64)(f * util_factor);
  }
  void* getRmvarPointer(i64 size) {
    auto it = rmvar.find(size);
    if (it == rmvar.end() || it->second.empty()) return nullptr;
    void* p = *it->second.begin();
    it->second.erase(it->second.begin());
    if (it->second.empty()) rmvar.erase(it);
    return p;
  }
  void* cudaMallocWarnIfFails(i64 size) {
    void* p = nullptr;
    int rc = sysml_b200_malloc(ctx, size, &p);
    if (rc || !p) return nullptr;
    allocated[p] = size;
    n_malloc++;
    return p;
  }
  void guardedCudaFree(void* p) {
    if (!p) return;
    auto it = allocated.find(p);
    if (it == allocated.end()) fail(SYSML_E_INVALID, "Attempting to free an unaccounted pointer");
    i64 size = it->second;
    allocated.erase(it);
    auto r = rmvar.find(size);
    if (r != rmvar.end()) {
      r->second.erase(p);
      if (r->second.empty()) rmvar.erase(r);
    }
    check(sysml_b200_free(ctx, p));
    n_free++;
  }
  void* malloc(const char* opcode, i64 size);  // needs GPUObject (below)
  void free(void* p, bool eager) {             // GPUMemoryManager.java:290-318
    if (!p) return;
    if (eager) {
      guardedCudaFree(p);
      return;
    }
    auto it = allocated.find(p);
    if (it == allocated.end()) fail(SYSML_E_INVALID, "ERROR : Internal state corrupted, cache block size map is not aware of a block it trying to free up");
    auto& fl = rmvar[it->second];
    if (fl.count(p)) fail(SYSML_E_INVALID, "GPU : Internal state corrupted, double free");
    fl.insert(p);
  }
  void clearMemory() {
    std::vector<void*> all;
    for (auto& kv : allocated) all.push_back(kv.first);
    for (void* p : all) guardedCudaFree(p);
    rmvar.clear();
  }
};

// MatrixBlock.evalSparseFormatInMemory (MatrixBlock.java:1001-1015) with the MCSR estimate it uses
// (DEFAULT_SPARSEBLOCK, MatrixBlock.java:107; SparseBlockMCSR.estimateMemory, SparseBlockMCSR.java:101-116;
// SparseRowVector.initialCapacity = 4) and estimateSizeDenseInMemory (:2339-2349).
bool evalSparseFormatInMemory(i64 nrows, i64 ncols, i64 nnz) {
  if (nrows <= 0 || ncols <= 0) return false;
  const double lsparsity = (double)nnz / (double)nrows / (double)ncols;
  const bool lsparse = lsparsity < 0.4;  // SPARSITY_TURN_POINT
  const double cnnz = std::max(4.0, ceil(lsparsity * (double)ncols));
  const double rlen = std::min((double)nrows, ceil(lsparsity * (double)nrows * (double)ncols));
  const double sizeSparse = 44.0 + 16.0 + 24.0 + (double)nrows * 8.0 + rlen * (76.0 + cnnz * 12.0);
  const double sizeDense = 44.0 + 8.0 * (double)nrows * (double)ncols;
  return lsparse && sizeSparse < sizeDense;
}

// MatrixObject + its host block (dense row-major or CSR) -- the part of CacheableData this layer needs.
struct MatrixObject {
  std::string name;
  i64 rows = 0, cols = 0, nnz = -1;
  bool host_valid = false, host_sparse = false;
  std::vector<double> h_dense;
  std::vector<int32_t> h_rowptr, h_colind;
  std::vector<double> h_val;
  std::unique_ptr<GPUObject> gpu;

  // MatrixBlock.recomputeNonZeros + examSparsity (MatrixBlock.java:963-990) on the host block, what LibMatrixMult does
  // to every matmult result (LibMatrixMult.java:113,149,168,225): a dense block whose exact nnz says "sparse in memory"
  // is rewritten as CSR (denseToSparse keeps row order and drops zeros; -0.0 counts as a zero).
  void recomputeNonZeros() {
    if (!host_valid) return;
    i64 c = 0;
    if (host_sparse) {
      for (double v : h_val) c += (v != 0.0);
    } else {
      for (double v : h_dense) c += (v != 0.0);
    }
    nnz = c;
  }
  void examSparsity() {
    if (!host_valid || host_sparse) return;
    recomputeNonZeros();
    if (!evalSparseFormatInMemory(rows, cols, nnz)) return;
    h_rowptr.assign(rows + 1, 0);
    h_colind.clear();
    h_val.clear();
    h_colind.reserve(nnz);
    h_val.reserve(nnz);
    for (i64 r = 0; r < rows; r++) {
      for (i64 j = 0; j < cols; j++) {
        const double v = h_dense[r * cols + j];
        if (v != 0.0) {
          h_colind.push_back((int32_t)j);
          h_val.push_back(v);
        }
      }
      h_rowptr[r + 1] = (int32_t)h_val.size();
    }
    std::vector<double>().swap(h_dense);
    host_sparse = true;
  }
};

// GPUObject: device state of one matrix (GPUObject.java:52-99): dense pointer or CSR triple, dirty bit,
// read-lock count, write-lock flag, LRU timestamp.
struct GPUObject {
  GPUMemoryManager* mm;
  MatrixObject* mat;
  double* dense = nullptr;
  int32_t *rowptr = nullptr, *colind = nullptr;
  double* val = nullptr;
  i64 csr_nnz = 0;
  bool dirty = false;
  bool borrowed = false;  // device block owned by the caller: never freed, never evicted
  int readLocks = 0;
  bool writeLock = false;
  bool uploading = false;  // its device arrays are being allocated right now: not an eviction candidate
  i64 timestamp = 0;
  i64* h2d_counter;
  i64* d2h_counter;
  i64* clock;

  bool isAllocated() const { return dense != nullptr || rowptr != nullptr; }
  bool isSparse() const { return rowptr != nullptr; }
  bool isLocked() const { return writeLock || readLocks > 0 || uploading; }
  i64 getSizeOnDevice() const {  // GPUObject.getSizeOnDevice (GPUObject.java:741-753)
    if (isSparse()) return (mat->rows + 1) * 4 + std::max<i64>(csr_nnz, 1) * 12;
    return dense ? std::max<i64>(mat->rows * mat->cols, 1) * 8 : 0;
  }
  void updateReleaseLocks() {  // GPUObject.java:650-667
    switch (mm->policy) {
      case LRU: timestamp = ++*clock; break;
      case LFU: timestamp += 1; break;
      case MIN_EVICT: break;
      case MRU: timestamp = -(++*clock); break;
    }
  }

  void copyFromHostToDevice(const char* opcode) {  // GPUObject.java:755-870
    if (!mat->host_valid) fail(SYSML_E_INVALID, "GPU : no host copy of variable %s to upload", mat->name.c_str());
    if (mat->host_sparse) {
      // The three arrays are allocated into locals and published together (CSRPointer.allocateEmpty finishes before
      // GPUObject sees the pointer, CSRPointer.java:271-287): a later malloc of this very upload may have to evict,
      // and a half-built object must not be its victim.
      const i64 nnz = mat->h_rowptr.empty() ? 0 : mat->h_rowptr.back();
      struct Guard {
        GPUObject* o;
        GPUMemoryManager* mm;
        void* p[3] = {nullptr, nullptr, nullptr};
        bool done = false;
        ~Guard() {
          o->uploading = false;
          if (!done)
            for (void* q : p)
              if (q) {
                try { mm->free(q, true); } catch (...) {}
              }
        }
      } gd{this, mm};
      uploading = true;
      gd.p[0] = mm->malloc(opcode, (mat->rows + 1) * 4);
      gd.p[1] = mm->malloc(opcode, std::max<i64>(nnz, 1) * 4);
      gd.p[2] = mm->malloc(opcode, std::max<i64>(nnz, 1) * 8);
      mm->check(sysml_b200_h2d(mm->ctx, gd.p[0], mat->h_rowptr.data(), (mat->rows + 1) * 4));
      if (nnz) {
        mm->check(sysml_b200_h2d(mm->ctx, gd.p[1], mat->h_colind.data(), nnz * 4));
        mm->check(sysml_b200_h2d(mm->ctx, gd.p[2], mat->h_val.data(), nnz * 8));
      }
      rowptr = (int32_t*)gd.p[0];
      colind = (int32_t*)gd.p[1];
      val = (double*)gd.p[2];
      csr_nnz = nnz;
      gd.done = true;
    } else {
      const i64 cells = mat->rows * mat->cols;
      dense = (double*)mm->malloc(opcode, std::max<i64>(cells, 1) * 8);
      if (cells) mm->check(sysml_b200_h2d(mm->ctx, dense, mat->h_dense.data(), cells * 8));
    }
    (*h2d_counter)++;
    dirty = false;
  }
  void copyFromDeviceToHost() {  // GPUObject.java:879-946
    if (isSparse()) {
      mat->h_rowptr.resize(mat->rows + 1);
      mat->h_colind.resize(csr_nnz);
      mat->h_val.resize(csr_nnz);
      mm->check(sysml_b200_d2h(mm->ctx, mat->h_rowptr.data(), rowptr, (mat->rows + 1) * 4));
      if (csr_nnz) {
        mm->check(sysml_b200_d2h(mm->ctx, mat->h_colind.data(), colind, csr_nnz * 4));
        mm->check(sysml_b200_d2h(mm->ctx, mat->h_val.data(), val, csr_nnz * 8));
      }
      mat->host_sparse = true;
    } else {
      mat->h_dense.resize(mat->rows * mat->cols);
      if (mat->rows * mat->cols) mm->check(sysml_b200_d2h(mm->ctx, mat->h_dense.data(), dense, mat->rows * mat->cols * 8));
      mat->host_sparse = false;
      mat->host_valid = true;
      mat->recomputeNonZeros();  // tmp.recomputeNonZeros() (GPUObject.java:896); the host block stays dense
    }
    mat->host_valid = true;
    (*d2h_counter)++;
    dirty = false;
  }
  void acquireDeviceRead(const char* opcode) {  // GPUObject.java:528-545
    if (!isAllocated()) copyFromHostToDevice(opcode);
    readLocks++;
  }
  void acquireDeviceModifyDense(const char* opcode) {  // GPUObject.java:547-565
    if (!isAllocated()) {
      uploading = true;  // registered but without a block yet: the malloc below must not pick it
      struct Reset { bool& f; ~Reset() { f = false; } } rs{uploading};
      dense = (double*)mm->malloc(opcode, std::max<i64>(mat->rows * mat->cols, 1) * 8);
    }
    if (isSparse()) fail(SYSML_E_INVALID, "GPU : output %s is held in sparse format", mat->name.c_str());
    dirty = true;
    writeLock = true;
  }
  // GPUObject.acquireDeviceModifySparse (GPUObject.java:567-585).  The row pointer comes from the operator's symbolic pass
  // (as CSRPointer.allocateForDgeam / allocateForMatrixMultiply build it, CSRPointer.java:227-399) and is adopted; the two
  // nnz-sized arrays are taken from the memory manager here.
  void acquireDeviceModifySparse(const char* opcode, int32_t* adopted_rowptr, i64 nnz) {
    if (isAllocated()) fail(SYSML_E_INVALID, "GPU : sparse output %s is already allocated", mat->name.c_str());
    uploading = true;
    struct Reset { bool& f; ~Reset() { f = false; } } rs{uploading};
    int32_t* ci = (int32_t*)mm->malloc(opcode, std::max<i64>(nnz, 1) * 4);
    double* va = nullptr;
    try {
      va = (double*)mm->malloc(opcode, std::max<i64>(nnz, 1) * 8);
    } catch (...) {
      mm->free(ci, true);
      throw;
    }
    rowptr = adopted_rowptr;
    colind = ci;
    val = va;
    csr_nnz = nnz;
    dirty = true;
    writeLock = true;
  }
  void releaseInput() {  // GPUObject.java:609-624
    if (readLocks <= 0) fail(SYSML_E_INVALID, "GPU : releaseInput of %s without a read lock", mat->name.c_str());
    readLocks--;
    updateReleaseLocks();
  }
  void releaseOutput() {  // GPUObject.java:626-645
    if (!writeLock) fail(SYSML_E_INVALID, "GPU : releaseOutput of %s without a write lock", mat->name.c_str());
    writeLock = false;
    updateReleaseLocks();
    if (!isAllocated()) fail(SYSML_E_INVALID, "GPU : attempting to release an output that was not allocated");
  }
  // GPUObject.sparseToDense (GPUObject.java:356-404): the CSR arrays are replaced by a dense block, for operators that
  // have no sparse form (LibMatrixCUDA.getDensePointer, LibMatrixCUDA.java:655-664, 1382-1383)
  void sparseToDense(const char* opcode) {
    if (!isSparse()) return;
    const bool was = uploading;
    uploading = true;
    double* d = nullptr;
    try {
      d = (double*)mm->malloc(opcode, std::max<i64>(mat->rows * mat->cols, 1) * 8);
    } catch (...) {
      uploading = was;
      throw;
    }
    uploading = was;
    int rc = sysml_b200_csr_to_dense(mm->ctx, rowptr, colind, val, mat->rows, mat->cols, d);
    if (rc) {
      mm->free(d, true);
      mm->check(rc);
    }
    mm->free(rowptr, mm->eager_free);
    mm->free(colind, mm->eager_free);
    mm->free(val, mm->eager_free);
    rowptr = colind = nullptr;
    val = nullptr;
    csr_nnz = 0;
    dense = d;
  }
  void acquireHostRead() {  // GPUObject.java:593-607
    if (isAllocated() && dirty) copyFromDeviceToHost();
  }
  void clearData(bool eager) {  // GPUObject.java:948-970
    if (borrowed) {
      dense = nullptr;
      dirty = false;
      return;
    }
    mm->free(dense, eager);
    mm->free(rowptr, eager);
    mm->free(colind, eager);
    mm->free(val, eager);
    dense = nullptr;
    rowptr = colind = nullptr;
    val = nullptr;
    csr_nnz = 0;
    dirty = false;
  }
};

void* GPUMemoryManager::malloc(const char* opcode, i64 size) {
  if (size < 0) fail(SYSML_E_INVALID, "Cannot allocate memory of size %lld", size);
  // Step 1: exact match in the free list
  void* A = getRmvarPointer(size);
  if (A) n_reuse++;
  // Step 2: new allocation while memory is available
  if (!A && size <= getAvailableMemory()) A = cudaMallocWarnIfFails(size);
  // Step 3: free the smallest larger free-list entry, then allocate
  if (!A) {
    i64 key = -1;
    for (auto& kv : rmvar)
      if (kv.first > size && !kv.second.empty()) {
        key = kv.first;
        break;
      }
    if (key >= 0) {
      void* p = getRmvarPointer(key);
      guardedCudaFree(p);
      A = cudaMallocWarnIfFails(size);
    }
  }
  // Step 4: eagerly free the whole free list
  if (!A) {
    std::vector<void*> toFree;
    for (auto& kv : rmvar)
      for (void* p : kv.second) toFree.push_back(p);
    for (void* p : toFree) guardedCudaFree(p);
    if (size <= getAvailableMemory()) A = cudaMallocWarnIfFails(size);
  }
  // Step 5: evict unlocked matrices in the order of the policy; dirty ones go back to the host first.
  // The reference sorts allocatedGPUObjects with GPUComparator(size) and evicts from the END of the list
  // (GPUMemoryManager.java:229-246): descending timestamp, i.e. smallest timestamp first (LRU: oldest release; LFU:
  // fewest releases; MRU: the negated clock makes the latest release the smallest); MIN_EVICT: the smallest block
  // that is at least as large as the request first, then larger ones, then the too-small ones from the largest down.
  if (!A) {
    std::vector<GPUObject*> cand;
    for (GPUObject* o : objects)
      if (o->isAllocated() && !o->isLocked() && !o->borrowed) cand.push_back(o);
    if (policy == MIN_EVICT) {
      std::stable_sort(cand.begin(), cand.end(), [size](GPUObject* a, GPUObject* b) {
        const i64 da = a->getSizeOnDevice() - size, db = b->getSizeOnDevice() - size;
        if (da >= 0 && db >= 0) return da < db;
        if (da >= 0 || db >= 0) return da >= 0;  // blocks that are large enough go before those that are not
        return da > db;
      });
    } else {
      std::stable_sort(cand.begin(), cand.end(), [](GPUObject* a, GPUObject* b) { return a->timestamp < b->timestamp; });
    }
    for (GPUObject* o : cand) {
      if (size <= getAvailableMemory()) break;
      if (o->dirty) o->copyFromDeviceToHost();
      o->clearData(true);
      n_evict++;
    }
    if (size <= getAvailableMemory()) A = cudaMallocWarnIfFails(size);
  }
  if (!A)
    fail(SYSML_E_OOM, "There is not enough memory on device for this matrix, request (%lld) for %s. allocated=%lld freelist=%lld",
         size, opcode, allocated_bytes(), rmvar_bytes());
  return A;
}

// ------------------------------------------------------------------------------------------------
// Instruction strings (lops/Lop.java:82-86, instructions/cp/CPOperand.java:40-43, InstructionUtils.java:181-190)
// ------------------------------------------------------------------------------------------------
const char* OPERAND_DELIM = "\xC2\xB0";  // U+00B0
const char* VALUETYPE_PREFIX = "\xC2\xB7";  // U+00B7

std::vector<std::string> split(const std::string& s, const std::string& d) {
  std::vector<std::string> out;
  size_t pos = 0;
  while (true) {
    size_t q = s.find(d, pos);
    if (q == std::string::npos) {
      out.push_back(s.substr(pos));
      break;
    }
    out.push_back(s.substr(pos, q - pos));
    pos = q + d.size();
  }
  return out;
}

struct CPOperand {
  std::string name, dataType, valueType;
  bool literal = false;
  bool isMatrix() const { return dataType == "MATRIX"; }
  bool isScalar() const { return dataType == "SCALAR"; }
};
CPOperand parseOperand(const std::string& str, const std::string& vt_delim) {
  std::vector<std::string> p = split(str, vt_delim);
  CPOperand o;
  o.name = p[0];
  if (p.size() >= 3) {
    o.dataType = p[1];
    o.valueType = p[2];
    if (p.size() >= 4) o.literal = (p[3] == "true");
  } else if (p.size() == 2) {  // InstructionUtils tolerates name + value type
    o.dataType = "MATRIX";
    o.valueType = p[1];
  } else {
    o.dataType = "UNKNOWN";
  }
  return o;
}

enum GPUType { AggregateBinary, MMTSJ, MMChain, AggregateUnary, ArithmeticBinary, RelationalBinary, BuiltinBinary, BuiltinUnary, Reorg };

struct OpcodeMap {
  std::map<std::string, GPUType> m;
  OpcodeMap() {  // GPUInstructionParser.java:42-145 (hot-path subset) + the entries this library adds
    m["ba+*"] = AggregateBinary;
    m["tsmm"] = MMTSJ;
    m["mmchain"] = MMChain;  // new on GPU
    m["r'"] = Reorg;
    for (const char* o : {"+", "-", "*", "/", "%%", "%/%", "^", "1-*", "^2", "*2", "-nz", "+*", "-*"}) m[o] = ArithmeticBinary;
    for (const char* o : {"==", "!=", "<", ">", "<=", ">="}) m[o] = RelationalBinary;
    for (const char* o : {"min", "max"}) m[o] = BuiltinBinary;
    for (const char* o : {"exp", "log", "abs", "sqrt", "round", "floor", "ceil", "sin", "cos", "tan", "tanh", "sign", "sigmoid"}) m[o] = BuiltinUnary;
    for (const char* o : {"ua+", "uak+", "uar+", "uark+", "uac+", "uack+", "ua*", "uar*", "uac*", "uamean", "uarmean", "uacmean", "uamax",
                          "uarmax", "uacmax", "uamin", "uarmin", "uacmin", "uasqk+", "uarsqk+", "uacsqk+", "uavar", "uarvar", "uacvar",
                          "uarimax", "uarimin"})  // uarimax/uarimin: new on GPU
      m[o] = AggregateUnary;
  }
};
const OpcodeMap& opcodes() {
  static OpcodeMap o;
  return o;
}

struct ParsedInstruction {
  GPUType type;
  std::string opcode;
  std::vector<CPOperand> ops;       // operands in instruction order
  std::vector<std::string> extras;  // trailing non-operand fields (flags, LEFT/RIGHT, chain type, k)
};

void checkNumFields(const std::vector<std::string>& parts, size_t expected) {  // InstructionUtils.checkNumFields
  if (parts.size() - 1 != expected) fail(SYSML_E_INVALID, "checkNumFields() for (%s) -- expected number (%zu) != is not equal to actual number (%zu).", parts[0].c_str(), expected, parts.size() - 1);
}

ParsedInstruction parseSingleInstruction(const std::string& str) {
  if (str.empty()) fail(SYSML_E_INVALID, "Unable to parse instruction: empty string");
  std::string od = OPERAND_DELIM, vd = VALUETYPE_PREFIX;
  if (str.find(od) == std::string::npos) od = ";";
  if (str.find(vd) == std::string::npos) vd = ":";
  std::vector<std::string> all = split(str, od);
  if (all.size() < 2) fail(SYSML_E_INVALID, "Unable to parse instruction: %s", str.c_str());
  if (all[0] != "GPU") fail(SYSML_E_INVALID, "Not a GPU instruction (exec type %s): %s", all[0].c_str(), str.c_str());
  std::vector<std::string> parts(all.begin() + 1, all.end());  // getInstructionPartsWithValueType
  ParsedInstruction pi;
  pi.opcode = parts[0];
  auto it = opcodes().m.find(pi.opcode);
  if (it == opcodes().m.end()) fail(SYSML_E_UNSUPPORTED, "Invalid GPU Instruction Type: opcode %s", pi.opcode.c_str());
  pi.type = it->second;
  auto op = [&](size_t i) { return parseOperand(parts[i], vd); };
  switch (pi.type) {
    case AggregateBinary:  // in1, in2, out, leftT, rightT (AggregateBinaryGPUInstruction.java:56-71)
      checkNumFields(parts, 5);
      pi.ops = {op(1), op(2), op(3)};
      pi.extras = {parts[4], parts[5]};
      break;
    case MMTSJ:  // in, out, LEFT|RIGHT (MMTSJGPUInstruction.java:61-75)
      checkNumFields(parts, 3);
      pi.ops = {op(1), op(2)};
      pi.extras = {parts[3]};
      if (parts[3] != "LEFT" && parts[3] != "RIGHT") fail(SYSML_E_INVALID, "Unknown MMTSJType %s", parts[3].c_str());
      break;
    case MMChain: {  // X, v, [w,] out, type [, k]  (MMChainCPInstruction.java:45-70; k is ignored on the GPU)
      size_t n = parts.size() - 1;
      bool has_k = n >= 5 && !parts[n].empty() && isdigit((unsigned char)parts[n][0]);
      size_t nf = has_k ? n - 1 : n;
      if (nf == 4) {
        pi.ops = {op(1), op(2), op(3)};
        pi.extras = {parts[4]};
      } else if (nf == 5) {
        pi.ops = {op(1), op(2), op(3), op(4)};
        pi.extras = {parts[5]};
      } else
        fail(SYSML_E_INVALID, "mmchain expects 4 or 5 fields (+ optional k), got %zu", n);
      if (pi.extras[0] != "XtXv" && pi.extras[0] != "XtwXv" && pi.extras[0] != "XtXvy") fail(SYSML_E_INVALID, "Unknown ChainType %s", pi.extras[0].c_str());
      if ((pi.extras[0] == "XtXv") != (pi.ops.size() == 3)) fail(SYSML_E_INVALID, "ChainType %s does not match the number of inputs", pi.extras[0].c_str());
      break;
    }
    case AggregateUnary:
    case BuiltinUnary:
    case Reorg:  // in, out
      if (parts.size() < 3) fail(SYSML_E_INVALID, "%s expects an input and an output", pi.opcode.c_str());
      pi.ops = {op(1), op(2)};
      break;
    case ArithmeticBinary:
      if (pi.opcode == "+*" || pi.opcode == "-*") {  // in1, scalar, in2, out (MatrixMatrixAxpyGPUInstruction.java:44-75)
        checkNumFields(parts, 4);
        pi.ops = {op(1), op(2), op(3), op(4)};
        if (!pi.ops[1].isScalar()) fail(SYSML_E_INVALID, "Expected second operand to be a scalar");
        break;
      }
      // fallthrough
    case RelationalBinary:
    case BuiltinBinary:  // in1, in2, out (ArithmeticBinaryGPUInstruction.java:42-67)
      checkNumFields(parts, 3);
      pi.ops = {op(1), op(2), op(3)};
      break;
  }
  return pi;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// ExecutionContext
// ------------------------------------------------------------------------------------------------
struct sysml_b200_ec {
  sysml_b200_ctx* ctx = nullptr;
  GPUMemoryManager mm;
  std::map<std::string, std::unique_ptr<MatrixObject>> matrices;
  std::map<std::string, double> scalars;
  i64 clock = 0;
  i64 n_inst = 0, n_h2d = 0, n_d2h = 0;
  std::string err;
  void* workspace = nullptr;  // the kernels' scratch block, allocated through (and accounted by) the memory manager

  // (Re)allocates the scratch block at `bytes` through the memory manager -- evicting matrices if that is what it takes --
  // and hands it to the kernels.  The old block goes back first, so the manager can reuse or release it.
  void growWorkspace(i64 bytes) {
    check(sysml_b200_ctx_set_workspace(ctx, nullptr, 0));  // (synchronises) the library owns nothing in between: it holds no block
    if (workspace) {
      mm.free(workspace, true);
      workspace = nullptr;
    }
    const i64 want = std::max<i64>(((bytes + bytes / 4) + 255) / 256 * 256, 1 << 20);
    void* p = nullptr;
    try {
      p = mm.malloc("workspace", want);
    } catch (const DMLRuntimeException&) {
      p = mm.malloc("workspace", (bytes + 255) / 256 * 256);  // without the head-room, then
    }
    workspace = p;
    auto it = mm.allocated.find(p);
    check(sysml_b200_ctx_set_workspace(ctx, p, it != mm.allocated.end() ? it->second : bytes));
  }

  MatrixObject* getMatrixObject(const std::string& name) {
    auto it = matrices.find(name);
    if (it == matrices.end()) fail(SYSML_E_INVALID, "Variable %s does not exist or is not a matrix", name.c_str());
    return it->second.get();
  }
  GPUObject* gpuObject(MatrixObject* mo) {
    if (!mo->gpu) {
      mo->gpu.reset(new GPUObject());
      mo->gpu->mm = &mm;
      mo->gpu->mat = mo;
      mo->gpu->h2d_counter = &n_h2d;
      mo->gpu->d2h_counter = &n_d2h;
      mo->gpu->clock = &clock;
      mm.objects.push_back(mo->gpu.get());
    }
    return mo->gpu.get();
  }
  // ExecutionContext.getMatrixInputForGPUInstruction (:367-380)
  MatrixObject* getMatrixInputForGPUInstruction(const std::string& name, const char* opcode) {
    MatrixObject* mo = getMatrixObject(name);
    gpuObject(mo)->acquireDeviceRead(opcode);
    return mo;
  }
  // ExecutionContext.getDenseMatrixOutputForGPUInstruction (:304-364): (re)creates the variable with the
  // given metadata, allocates its device block and takes the write lock
  MatrixObject* getDenseMatrixOutputForGPUInstruction(const std::string& name, i64 rows, i64 cols, const char* opcode) {
    auto it = matrices.find(name);
    if (it != matrices.end() && it->second->gpu && it->second->gpu->isLocked())
      fail(SYSML_E_INVALID, "GPU : output variable %s is an input of the same instruction", name.c_str());
    removeMatrix(name);
    std::unique_ptr<MatrixObject> mo(new MatrixObject());
    mo->name = name;
    mo->rows = rows;
    mo->cols = cols;
    MatrixObject* p = mo.get();
    matrices[name] = std::move(mo);
    gpuObject(p)->acquireDeviceModifyDense(opcode);
    return p;
  }
  // ExecutionContext.getSparseMatrixOutputForGPUInstruction (:321-326): as above with a CSR device block of nnz cells
  MatrixObject* getSparseMatrixOutputForGPUInstruction(const std::string& name, i64 rows, i64 cols, int32_t* rowptr, i64 nnz,
                                                       const char* opcode) {
    auto it = matrices.find(name);
    if (it != matrices.end() && it->second->gpu && it->second->gpu->isLocked())
      fail(SYSML_E_INVALID, "GPU : output variable %s is an input of the same instruction", name.c_str());
    removeMatrix(name);
    std::unique_ptr<MatrixObject> mo(new MatrixObject());
    mo->name = name;
    mo->rows = rows;
    mo->cols = cols;
    mo->nnz = nnz;
    MatrixObject* p = mo.get();
    matrices[name] = std::move(mo);
    gpuObject(p)->acquireDeviceModifySparse(opcode, rowptr, nnz);
    return p;
  }
  void releaseMatrixInputForGPUInstruction(const std::string& name) { getMatrixObject(name)->gpu->releaseInput(); }
  void releaseMatrixOutputForGPUInstruction(const std::string& name) {
    MatrixObject* mo = getMatrixObject(name);
    mo->gpu->releaseOutput();
    mo->host_valid = false;  // the device copy is the only valid one now (dirty)
  }
  double getScalarInput(const CPOperand& o) {  // ExecutionContext.getScalarInput: literal or variable
    if (o.literal) {
      char* end = nullptr;
      double v = strtod(o.name.c_str(), &end);
      if (end == o.name.c_str()) {
        if (o.name == "TRUE" || o.name == "true") return 1.0;
        if (o.name == "FALSE" || o.name == "false") return 0.0;
        fail(SYSML_E_INVALID, "Cannot parse scalar literal '%s'", o.name.c_str());
      }
      return v;
    }
    auto it = scalars.find(o.name);
    if (it == scalars.end()) fail(SYSML_E_INVALID, "Scalar variable %s does not exist", o.name.c_str());
    return it->second;
  }
  void removeMatrix(const std::string& name) {
    auto it = matrices.find(name);
    if (it == matrices.end()) return;
    if (it->second->gpu) {
      GPUObject* g = it->second->gpu.get();
      if (g->isLocked()) fail(SYSML_E_INVALID, "GPU : cannot remove locked variable %s", name.c_str());
      g->clearData(mm.eager_free);
      mm.objects.erase(std::remove(mm.objects.begin(), mm.objects.end(), g), mm.objects.end());
    }
    matrices.erase(it);
  }
  void check(int rc) { mm.check(rc); }

  void processInstruction(const ParsedInstruction& pi);
};

namespace {

int binaryOpcode(const std::string& o) {  // LibMatrixCUDA.getBinaryOp (LibMatrixCUDA.java:1551-1575)
  static const std::map<std::string, int> m = {{"+", 0},  {"-", 1},  {"*", 2},  {"/", 3},    {"^", 4},    {"<", 5},   {"<=", 6},
                                               {">", 7},  {">=", 8}, {"==", 9}, {"!=", 10},  {"min", 11}, {"max", 12}, {"&&", 13},
                                               {"||", 14}, {"1-*", 15}, {"-nz", 16}, {"%%", 17}, {"%/%", 18}, {"*2", 2}, {"^2", 4}};
  auto it = m.find(o);
  if (it == m.end()) fail(SYSML_E_UNSUPPORTED, "The given value function is not supported: %s", o.c_str());
  return it->second;
}
int unaryOpcode(const std::string& o) {
  static const std::map<std::string, int> m = {{"exp", 0},  {"sqrt", 1}, {"round", 2}, {"abs", 3},     {"log", 4}, {"floor", 5}, {"ceil", 6},
                                               {"sign", 7}, {"sigmoid", 8}, {"sin", 9}, {"cos", 10}, {"tan", 11}, {"tanh", 12}};
  auto it = m.find(o);
  if (it == m.end()) fail(SYSML_E_UNSUPPORTED, "Unsupported unary builtin: %s", o.c_str());
  return it->second;
}
// Which cellwise operators keep a sparse operand sparse (the CPU back-end's rules; the reference's GPU operators have no
// sparse form and densify, LibMatrixCUDA.java:1360-1383):
//  - scalar: ScalarOperator's constructor (ScalarOperator.java:53-59, isSparseSafeStatic :84-89) with the alternatives of
//    RightScalarOperator (:42-49) and LeftScalarOperator (:41-46).  One deliberate difference: Power with a NEGATIVE constant
//    is sparse-safe there ("Power && cst != 0") although 0^c is infinite; it takes the dense form here.
//  - matrix-matrix: BinaryOperator.sparseSafe (BinaryOperator.java:60-66), Divide by an operand without zeros
//    (LibMatrixBincell.isSparseSafeDivide :185-189)
//  - unary: UnaryOperator's constructor (UnaryOperator.java:38-47)
bool scalarSparseSafe(const std::string& o, double c, bool scalarLeft) {
  if (o == "*" || o == "*2" || o == "^2" || o == "&&" || o == "-nz") return true;
  if ((o == "!=" && c == 0) || (o == "==" && c != 0) || (o == "-" && c == 0) || (o == "max" && c <= 0) || (o == "min" && c >= 0)) return true;
  if (!scalarLeft) return (o == ">" && c >= 0) || (o == ">=" && c > 0) || (o == "<" && c <= 0) || (o == "<=" && c < 0) || (o == "/" && c != 0) || (o == "^" && c > 0);
  return (o == ">" && c <= 0) || (o == ">=" && c < 0) || (o == "<" && c >= 0) || (o == "<=" && c > 0);
}
bool binarySparseSafe(const std::string& o) { return o == "+" || o == "-" || o == "*" || o == "&&" || o == "||"; }
bool unarySparseSafe(const std::string& o) { return o == "sin" || o == "tan" || o == "tanh" || o == "round" || o == "abs" || o == "sqrt" || o == "sign"; }
// OptimizerUtils.getBinaryOpSparsity, worst case (OptimizerUtils.java:1079-1103)
double binaryOpSparsity(const std::string& o, double sp1, double sp2) {
  if (o == "+" || o == "-" || o == "||" || o == "<" || o == ">" || o == "!=" || o == "min" || o == "max") return std::min(1.0, sp1 + sp2);
  if (o == "*" || o == "&&") return std::min(sp1, sp2);
  if (o == "/") return std::min(1.0, sp1 + (1.0 - sp2));
  if (o == "%%" || o == "^" || o == "-nz") return sp1;
  return 1.0;
}
// InstructionUtils.parseBasicAggregateUnaryOperator: opcode -> (aggregate, direction)
void aggOpcode(const std::string& o, int* op, int* dir) {
  std::string s = o;
  *dir = SYSML_DIR_ALL;
  if (s.compare(0, 3, "uar") == 0 && s != "uar" ) { *dir = SYSML_DIR_ROW; s = "ua" + s.substr(3); }
  else if (s.compare(0, 3, "uac") == 0) { *dir = SYSML_DIR_COL; s = "ua" + s.substr(3); }
  if (s == "ua+" || s == "uak+") *op = SYSML_AGG_SUM;
  else if (s == "uasqk+") *op = SYSML_AGG_SUMSQ;
  else if (s == "uamax") *op = SYSML_AGG_MAX;
  else if (s == "uamin") *op = SYSML_AGG_MIN;
  else if (s == "uaimax") *op = SYSML_AGG_MAXINDEX;
  else if (s == "uaimin") *op = SYSML_AGG_MININDEX;
  else if (s == "uamean") *op = SYSML_AGG_MEAN;
  else if (s == "ua*") *op = SYSML_AGG_PROD;
  else if (s == "uavar") *op = SYSML_AGG_VAR;
  else fail(SYSML_E_UNSUPPORTED, "Unknown aggregate opcode %s", o.c_str());
}

}  // namespace

void sysml_b200_ec::processInstruction(const ParsedInstruction& pi) {
  n_inst++;  // GPUStatistics.incrementNoOfExecutedGPUInst
  const char* opc = pi.opcode.c_str();
  // Inputs are read-locked and outputs write-locked for the duration of the instruction and released when
  // it ends, also when it fails (ec.releaseMatrixInput/OutputForGPUInstruction, ExecutionContext.java:401-452).
  struct Locks {
    sysml_b200_ec* ec;
    std::vector<std::string> ins, outs;
    MatrixObject* in(const std::string& name, const char* opcode) {
      MatrixObject* mo = ec->getMatrixInputForGPUInstruction(name, opcode);
      ins.push_back(name);
      return mo;
    }
    MatrixObject* out(const std::string& name, i64 rows, i64 cols, const char* opcode) {
      MatrixObject* mo = ec->getDenseMatrixOutputForGPUInstruction(name, rows, cols, opcode);
      outs.push_back(name);
      return mo;
    }
    MatrixObject* outSparse(const std::string& name, i64 rows, i64 cols, int32_t* rowptr, i64 nnz, const char* opcode) {
      MatrixObject* mo = ec->getSparseMatrixOutputForGPUInstruction(name, rows, cols, rowptr, nnz, opcode);
      outs.push_back(name);
      return mo;
    }
    void releaseAll() {
      for (auto& n : ins) {
        try { ec->releaseMatrixInputForGPUInstruction(n); } catch (...) {}
      }
      for (auto& n : outs) {
        try { ec->releaseMatrixOutputForGPUInstruction(n); } catch (...) {}
      }
      ins.clear();
      outs.clear();
    }
    ~Locks() { releaseAll(); }
  } L{this, {}, {}};
  // device temporaries of this instruction, taken from and returned to the memory manager like the reference's
  // (gCtx.allocate / gCtx.cudaFreeHelper inside the operator, LibMatrixCUDA.java:711,732,926,947)
  struct Temps {
    sysml_b200_ec* ec;
    std::vector<void*> ptrs;
    void* get(i64 bytes, const char* opcode) {
      void* p = ec->mm.malloc(opcode, bytes);
      ptrs.push_back(p);
      return p;
    }
    ~Temps() {
      if (ptrs.empty()) return;
      sysml_b200_sync(ec->ctx);  // the kernels that read them may still be queued
      for (void* p : ptrs) {
        try { ec->mm.free(p, ec->mm.eager_free); } catch (...) {}
      }
    }
  } T{this, {}};
  // operators without a sparse form convert the input in place, as LibMatrixCUDA.getDensePointer does
  // (LibMatrixCUDA.java:655-664, 1382-1383 -> GPUObject.sparseToDense)
  auto densify = [&](MatrixObject* mo) {
    if (mo->gpu->isSparse()) mo->gpu->sparseToDense(opc);
  };
  // A cellwise operator in its sparse form: the symbolic pass sizes the result into a row pointer taken from the memory
  // manager, the output variable adopts it and allocates the nnz-sized arrays, the numeric pass fills them
  // (sysml_b200.h, "two-call protocol").  Returns false -- with nothing changed -- when the library finds the operator not
  // sparse-safe for these operands (SYSML_E_UNSUPPORTED): the caller then takes the dense form.
  auto sparseCellwise = [&](const std::string& outName, i64 rows, i64 cols, auto call) -> bool {
    int32_t* rp = (int32_t*)mm.malloc(opc, (rows + 1) * 4);
    i64 nnz = 0;
    int rc = call(rp, (int32_t*)nullptr, (double*)nullptr, &nnz);
    if (rc) {
      mm.free(rp, true);
      if (rc == SYSML_E_UNSUPPORTED) return false;
      check(rc);
    }
    MatrixObject* out = nullptr;
    try {
      out = L.outSparse(outName, rows, cols, rp, nnz, opc);
    } catch (...) {
      mm.free(rp, true);
      throw;
    }
    check(call(out->gpu->rowptr, out->gpu->colind, out->gpu->val, &nnz));
    return true;
  };
  // exact sparsity of an operand (MatrixBlock.getSparsity); a dense device block whose nnz is not known is counted
  auto sparsityOf = [&](MatrixObject* mo) -> double {
    const double cells = (double)mo->rows * (double)mo->cols;
    if (cells == 0) return 0.0;
    if (mo->gpu->isSparse()) return (double)mo->gpu->csr_nnz / cells;
    if (mo->nnz < 0) {
      int64_t c = 0;
      check(sysml_b200_count_nnz(ctx, mo->gpu->dense, mo->rows * mo->cols, &c));
      mo->nnz = c;
    }
    return (double)mo->nnz / cells;
  };
  switch (pi.type) {
    case AggregateBinary: {  // AggregateBinaryGPUInstruction.processInstruction (:73-89) -> LibMatrixCuMatMult.matmult (:139-229)
      MatrixObject* m1 = L.in(pi.ops[0].name, opc);
      MatrixObject* m2 = L.in(pi.ops[1].name, opc);
      bool lt = pi.extras[0] == "true", rt = pi.extras[1] == "true";
      const i64 rlen = lt ? m1->cols : m1->rows, clen = rt ? m2->rows : m2->cols;
      if ((lt ? m1->rows : m1->cols) != (rt ? m2->cols : m2->rows))
        fail(SYSML_E_DIM, "GPU : incompatible dimensions for ba+*: %lldx%lld%s and %lldx%lld%s", m1->rows, m1->cols, lt ? "'" : "", m2->rows, m2->cols, rt ? "'" : "");
      GPUObject *g1 = m1->gpu.get(), *g2 = m2->gpu.get();
      // Operand views as the kernels take them.  The reference passes the transpose flags on to cuSPARSE / cuBLAS for
      // every sparsity pair (LibMatrixCuMatMult.java:139-229); here a flag the kernel of that pair has no form for is
      // resolved by materialising the transposed operand in a temporary (freed when the instruction ends).
      struct View {
        bool sparse;
        const double* dense;
        const int32_t *rp, *ci;
        const double* va;
        i64 rows, cols;
      };
      View A{g1->isSparse(), g1->dense, g1->rowptr, g1->colind, g1->val, m1->rows, m1->cols};
      View B{g2->isSparse(), g2->dense, g2->rowptr, g2->colind, g2->val, m2->rows, m2->cols};
      auto transposeCsr = [&](View& V, i64 nnz) {
        int32_t* rp = (int32_t*)T.get((V.cols + 1) * 4, opc);
        int32_t* ci = (int32_t*)T.get(std::max<i64>(nnz, 1) * 4, opc);
        double* va = (double*)T.get(std::max<i64>(nnz, 1) * 8, opc);
        check(sysml_b200_csr_transpose(ctx, V.rp, V.ci, V.va, V.rows, V.cols, rp, ci, va));
        V = View{true, nullptr, rp, ci, va, V.cols, V.rows};
      };
      auto transposeDense = [&](View& V) {
        double* d = (double*)T.get(std::max<i64>(V.rows * V.cols, 1) * 8, opc);
        check(sysml_b200_transpose(ctx, V.dense, d, V.rows, V.cols));
        V = View{false, d, nullptr, nullptr, nullptr, V.cols, V.rows};
      };
      // LibMatrixMult.isUltraSparseMatrixMult (LibMatrixMult.java:3761-3777) -> matrixMultUltraSparse with a sparse result
      // (:135, 1502-1592).  Taken here for its left branch with a dense right operand and no transpose flags (a selection
      // or permutation matrix, a few cells, times X); the other forms go through the general kernels, which compute the
      // same cells in the same order into a dense block.  (The reference's permutation-matrix clause, :3769-3770, needs a scan
      // of the row lengths and is covered here only as far as the sparsity clauses reach.)
      if (A.sparse && !B.sparse && !lt && !rt && m2->cols > 1) {
        const double cells1 = (double)m1->rows * (double)m1->cols;
        const double spA = cells1 > 0 ? (double)g1->csr_nnz / cells1 : 0.0;
        const bool us1 = spA < 0.00004 && g1->csr_nnz < 40;       // MatrixBlock.isUltraSparse (MatrixBlock.java:876-885)
        const bool us1NoNnz = spA < 0.00004;
        const double spB = sparsityOf(m2);
        // OptimizerUtils.getMatMultSparsity, worst case (OptimizerUtils.java:985-991)
        const double outSp = std::min(1.0, spA * (double)m1->cols) * std::min(1.0, spB * (double)m2->rows);
        const bool ultra = us1 || (us1NoNnz && outSp < 0.0004) ||
                           (spA < 0.0004 && g1->csr_nnz < 40 && cells1 + (double)m2->rows * (double)m2->cols < (double)m1->rows * (double)m2->cols && outSp < 0.4);
        if (ultra) {
          const bool done = sparseCellwise(pi.ops[2].name, rlen, clen, [&](int32_t* rp, int32_t* ci, double* va, i64* nnz) {
            return sysml_b200_matmult_ultra_sparse_left(ctx, A.rp, A.ci, A.va, B.dense, A.rows, A.cols, B.cols, rp, ci, va, (int64_t*)nnz);
          });
          if (done) break;
        }
      }
      if (B.sparse && rt) {  // no kernel takes t(B) of a CSR right operand
        transposeCsr(B, g2->csr_nnz);
        rt = false;
      }
      if (A.sparse && B.sparse && lt) {  // sparse x sparse has no flags at all
        transposeCsr(A, g1->csr_nnz);
        lt = false;
      }
      if (A.sparse && !B.sparse && rt) {  // CSR x dense takes the left flag only
        transposeDense(B);
        rt = false;
      }
      MatrixObject* out = L.out(pi.ops[2].name, rlen, clen, opc);
      int rc;
      if (A.sparse && B.sparse) {
        rc = sysml_b200_matmult_csr_csr(ctx, A.rp, A.ci, A.va, B.rp, B.ci, B.va, out->gpu->dense, A.rows, A.cols, B.rows, B.cols);
      } else if (B.sparse) {
        rc = sysml_b200_matmult_dense_csr(ctx, A.dense, B.rp, B.ci, B.va, out->gpu->dense, A.rows, A.cols, B.rows, B.cols, lt);
      } else if (A.sparse) {
        rc = sysml_b200_matmult_csr_dense(ctx, A.rp, A.ci, A.va, B.dense, out->gpu->dense, A.rows, A.cols, B.rows, B.cols, lt);
      } else {
        rc = sysml_b200_matmult_dense(ctx, A.dense, B.dense, out->gpu->dense, A.rows, A.cols, B.rows, B.cols, lt, rt);
      }
      check(rc);
      // ret.recomputeNonZeros() after every matrixMult (LibMatrixMult.java:112,148,167,224): exact count on the device
      int64_t nnz = 0;
      check(sysml_b200_count_nnz(ctx, out->gpu->dense, rlen * clen, &nnz));
      out->nnz = nnz;
      break;
    }
    case MMTSJ: {  // MMTSJGPUInstruction.processInstruction (:78-89)
      MatrixObject* mat = L.in(pi.ops[0].name, opc);
      const bool left = pi.extras[0] == "LEFT";
      const i64 dim = left ? mat->cols : mat->rows;
      MatrixObject* out = L.out(pi.ops[1].name, dim, dim, opc);
      int64_t nnz = 0;
      int rc;
      if (mat->gpu->isSparse()) {
        rc = left ? sysml_b200_tsmm_csr_left(ctx, mat->gpu->rowptr, mat->gpu->colind, mat->gpu->val, out->gpu->dense, mat->rows, mat->cols, &nnz)
                  : sysml_b200_tsmm_csr_right(ctx, mat->gpu->rowptr, mat->gpu->colind, mat->gpu->val, out->gpu->dense, mat->rows, mat->cols, &nnz);
      } else {
        rc = sysml_b200_tsmm_dense(ctx, mat->gpu->dense, out->gpu->dense, mat->rows, mat->cols, left ? 1 : 0, &nnz);
      }
      out->nnz = nnz;  // ret.setNonZeros(copyUpperToLowerTriangle(ret)), LibMatrixMult.java:352-353
      check(rc);
      break;
    }
    case MMChain: {  // MMChainCPInstruction.processInstruction (cp/MMChainCPInstruction.java:69-83) on the GPU
      const bool hasW = pi.ops.size() == 4;
      MatrixObject* X = L.in(pi.ops[0].name, opc);
      MatrixObject* v = L.in(pi.ops[1].name, opc);
      MatrixObject* w = hasW ? L.in(pi.ops[2].name, opc) : nullptr;
      const std::string& outName = pi.ops[hasW ? 3 : 2].name;
      const int ct = pi.extras[0] == "XtXv" ? SYSML_CHAIN_XtXv : (pi.extras[0] == "XtwXv" ? SYSML_CHAIN_XtwXv : SYSML_CHAIN_XtXvy);
      int rc = SYSML_OK;
      std::string msg;
      if (v->rows * v->cols != X->cols) { rc = SYSML_E_DIM; msg = "mmchain: v must have ncol(X) cells"; }
      if (hasW && w->rows * w->cols != X->rows) { rc = SYSML_E_DIM; msg = "mmchain: w must have nrow(X) cells"; }
      densify(v);
      if (w) densify(w);
      MatrixObject* out = L.out(outName, X->cols, 1, opc);
      if (!rc) {
        if (X->gpu->isSparse())
          rc = sysml_b200_mmchain_csr(ctx, X->gpu->rowptr, X->gpu->colind, X->gpu->val, v->gpu->dense, w ? w->gpu->dense : nullptr, out->gpu->dense, X->rows, X->cols, ct);
        else
          rc = sysml_b200_mmchain_dense(ctx, X->gpu->dense, v->gpu->dense, w ? w->gpu->dense : nullptr, out->gpu->dense, X->rows, X->cols, ct);
      }
      (void)0;
      if (!msg.empty()) fail(rc, "GPU : %s", msg.c_str());
      check(rc);
      break;
    }
    case AggregateUnary: {  // AggregateUnaryGPUInstruction.processInstruction (:67-101)
      int op, dir;
      aggOpcode(pi.opcode, &op, &dir);
      MatrixObject* in1 = L.in(pi.ops[0].name, opc);
      const i64 orows = dir == SYSML_DIR_ROW ? in1->rows : 1, ocols = dir == SYSML_DIR_COL ? in1->cols : 1;
      const bool scalarOut = dir == SYSML_DIR_ALL;
      const std::string outName = scalarOut ? std::string("__scalar_tmp__") : pi.ops[1].name;
      MatrixObject* out = L.out(outName, orows, ocols, opc);
      int rc;
      if (in1->gpu->isSparse())
        rc = sysml_b200_agg_unary_csr(ctx, in1->gpu->rowptr, in1->gpu->colind, in1->gpu->val, in1->rows, in1->cols, op, dir, out->gpu->dense);
      else
        rc = sysml_b200_agg_unary_dense(ctx, in1->gpu->dense, in1->rows, in1->cols, op, dir, out->gpu->dense);
      L.releaseAll();
      if (!rc && scalarOut) {  // scalar: D2H of 8 bytes + ec.setScalarOutput (LibMatrixCUDA.java:692-693)
        double val = 0;
        rc = sysml_b200_d2h(ctx, &val, out->gpu->dense, 8);
        n_d2h++;
        scalars[pi.ops[1].name] = val;
      }
      if (scalarOut) removeMatrix(outName);
      check(rc);
      break;
    }
    case BuiltinUnary:
    case Reorg: {
      MatrixObject* in1 = L.in(pi.ops[0].name, opc);
      const bool tr = pi.type == Reorg;
      if (!tr && in1->gpu->isSparse() && unarySparseSafe(pi.opcode)) {  // MatrixBlock.unaryOperations: sp = sparse && op.sparseSafe (:2594)
        GPUObject* g = in1->gpu.get();
        const int code = unaryOpcode(pi.opcode);
        if (sparseCellwise(pi.ops[1].name, in1->rows, in1->cols, [&](int32_t* rp, int32_t* ci, double* va, i64* nnz) {
              return sysml_b200_csr_unary_op(ctx, g->rowptr, g->colind, g->val, in1->rows, in1->cols, code, rp, ci, va, (int64_t*)nnz);
            }))
          break;
      }
      densify(in1);
      MatrixObject* out = L.out(pi.ops[1].name, tr ? in1->cols : in1->rows, tr ? in1->rows : in1->cols, opc);
      int rc = tr ? sysml_b200_transpose(ctx, in1->gpu->dense, out->gpu->dense, in1->rows, in1->cols)
                  : sysml_b200_unary(ctx, in1->gpu->dense, out->gpu->dense, in1->rows * in1->cols, unaryOpcode(pi.opcode));
      check(rc);
      break;
    }
    case ArithmeticBinary:
    case RelationalBinary:
    case BuiltinBinary: {
      if (pi.opcode == "+*" || pi.opcode == "-*") {  // MatrixMatrixAxpyGPUInstruction.processInstruction (:78-101)
        MatrixObject* in1 = L.in(pi.ops[0].name, opc);
        MatrixObject* in2 = L.in(pi.ops[2].name, opc);
        const double scalar = getScalarInput(pi.ops[1]) * (pi.opcode == "-*" ? -1.0 : 1.0);
        int bcast = -1;
        if (in1->rows == in2->rows && in1->cols == in2->cols) bcast = SYSML_BCAST_NONE;
        else if (in1->rows == in2->rows && in2->cols == 1) bcast = SYSML_BCAST_COL;
        else if (in2->rows == 1 && in1->cols == in2->cols) bcast = SYSML_BCAST_ROW;
        if (bcast < 0)
          fail(SYSML_E_DIM, "Incorrect dimensions of inputs in GPU axpy operation. input1:%lld X %lld and input2:%lld X %lld", in1->rows, in1->cols, in2->rows, in2->cols);
        densify(in1);
        densify(in2);
        MatrixObject* out = L.out(pi.ops[3].name, in1->rows, in1->cols, opc);
        int rc = sysml_b200_axpy(ctx, in1->gpu->dense, in2->gpu->dense, scalar, out->gpu->dense, in1->rows, in1->cols, bcast);
        check(rc);
        break;
      }
      const CPOperand &o1 = pi.ops[0], &o2 = pi.ops[1], &o3 = pi.ops[2];
      if (o1.isMatrix() && o2.isMatrix()) {  // MatrixMatrixArithmeticGPUInstruction.processInstruction (:38-71)
        MatrixObject* in1 = L.in(o1.name, opc);
        MatrixObject* in2 = L.in(o2.name, opc);
        // LibMatrixBincell.getBinaryAccessType (LibMatrixBincell.java:145-163)
        int bcast = -1;
        i64 orows = in1->rows, ocols = in1->cols;
        if (in1->rows == in2->rows && in1->cols == in2->cols) bcast = SYSML_BCAST_NONE;
        else if (in1->cols > 1 && in2->cols == 1 && in1->rows == in2->rows) bcast = SYSML_BCAST_COL;
        else if (in1->rows > 1 && in1->cols > 1 && in2->rows == 1 && in1->cols == in2->cols) bcast = SYSML_BCAST_ROW;
        else if (in1->cols == 1 && in2->rows == 1) {  // OUTER_VECTOR_VECTOR: n x 1 op 1 x m -> n x m
          bcast = SYSML_BCAST_OUTER;
          ocols = in2->cols;
        }
        if (bcast < 0) fail(SYSML_E_DIM, "GPU : incompatible dimensions %lldx%lld vs %lldx%lld for %s", in1->rows, in1->cols, in2->rows, in2->cols, opc);
        // Sparse form (MatrixBlock.binaryOperations -> estimateSparsityOnBinary, MatrixBlock.java:2440-2485, 2741-2766): taken
        // when the operator is sparse-safe (or Divide by an operand without zeros) and the estimated result is sparse in memory
        const bool sp1 = in1->gpu->isSparse(), sp2 = in2->gpu->isSparse();
        if ((sp1 || sp2) && bcast != SYSML_BCAST_OUTER) {
          const std::string& o = pi.opcode;
          const int code = binaryOpcode(o);
          bool safe = binarySparseSafe(o);
          double s1 = sparsityOf(in1), s2 = sparsityOf(in2);
          if (!safe && o == "/" && sp1 && s2 == 1.0) safe = true;
          // a vector operand counts once per row / column it is broadcast over (nz2 * n or nz2 * m, :2468-2471): its own sparsity
          const i64 estnnz = (i64)std::llround(binaryOpSparsity(o, s1, s2) * (double)orows * (double)ocols);
          if (safe && evalSparseFormatInMemory(orows, ocols, estnnz)) {
            GPUObject *g1 = in1->gpu.get(), *g2 = in2->gpu.get();
            bool done = false;
            if (sp1 && sp2 && bcast == SYSML_BCAST_NONE) {
              done = sparseCellwise(o3.name, orows, ocols, [&](int32_t* rp, int32_t* ci, double* va, i64* nnz) {
                return sysml_b200_csr_csr_op(ctx, g1->rowptr, g1->colind, g1->val, g2->rowptr, g2->colind, g2->val, orows, ocols, code, rp, ci, va,
                                             (int64_t*)nnz);
              });
            } else if (sp1 && !sp2) {
              done = sparseCellwise(o3.name, orows, ocols, [&](int32_t* rp, int32_t* ci, double* va, i64* nnz) {
                return sysml_b200_csr_dense_op(ctx, g1->rowptr, g1->colind, g1->val, orows, ocols, g2->dense, bcast, code, 0, rp, ci, va, (int64_t*)nnz);
              });
            } else if (!sp1 && sp2 && bcast == SYSML_BCAST_NONE && (o == "*" || o == "&&")) {
              done = sparseCellwise(o3.name, orows, ocols, [&](int32_t* rp, int32_t* ci, double* va, i64* nnz) {
                return sysml_b200_csr_dense_op(ctx, g2->rowptr, g2->colind, g2->val, orows, ocols, g1->dense, bcast, code, 1, rp, ci, va, (int64_t*)nnz);
              });
            }
            if (done) break;
          }
        }
        densify(in1);
        densify(in2);
        MatrixObject* out = L.out(o3.name, orows, ocols, opc);
        int rc = sysml_b200_binary_mm(ctx, in1->gpu->dense, in2->gpu->dense, out->gpu->dense, orows, ocols, bcast, binaryOpcode(pi.opcode));
        check(rc);
      } else if ((o1.isScalar() && o2.isMatrix()) || (o1.isMatrix() && o2.isScalar())) {
        // ScalarMatrixArithmeticGPUInstruction.processInstruction (:40-61)
        const bool scalarLeft = o1.isScalar();
        const CPOperand& mo = scalarLeft ? o2 : o1;
        const double scalar = getScalarInput(scalarLeft ? o1 : o2);
        MatrixObject* in = L.in(mo.name, opc);
        double s = scalar;
        int code = binaryOpcode(pi.opcode);
        if (pi.opcode == "*2") s = 2.0;      // Multiply2 / Power2 (LibMatrixCUDA.java:1563-1564)
        if (pi.opcode == "^2") s = 2.0;
        // MatrixBlock.scalarOperations: the result is sparse iff the input is and the operator is sparse-safe for this constant
        if (in->gpu->isSparse() && scalarSparseSafe(pi.opcode, s, scalarLeft)) {
          GPUObject* g = in->gpu.get();
          if (sparseCellwise(o3.name, in->rows, in->cols, [&](int32_t* rp, int32_t* ci, double* va, i64* nnz) {
                return sysml_b200_csr_scalar_op(ctx, g->rowptr, g->colind, g->val, in->rows, in->cols, s, code, scalarLeft ? 1 : 0, rp, ci, va,
                                                (int64_t*)nnz);
              }))
            break;
        }
        densify(in);
        MatrixObject* out = L.out(o3.name, in->rows, in->cols, opc);
        int rc = sysml_b200_binary_ms(ctx, in->gpu->dense, s, out->gpu->dense, in->rows * in->cols, code, scalarLeft ? 1 : 0);
        check(rc);
      } else {
        fail(SYSML_E_UNSUPPORTED, "Unsupported GPU ArithmeticInstruction.");
      }
      break;
    }
  }
  // GPUInstruction.postprocessInstruction: sysml.gpu.sync.postProcess = true (GPUInstruction.java:202-208)
  check(sysml_b200_sync(ctx));
}

// ------------------------------------------------------------------------------------------------
// C entry points
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_host_err;

#define EC_TRY(ec, ...)                          \
  try {                                          \
    __VA_ARGS__;                                 \
    return SYSML_OK;                             \
  } catch (const DMLRuntimeException& e) {       \
    if (ec) (ec)->err = e.what();                \
    g_host_err = e.what();                       \
    return e.code;                               \
  } catch (const std::exception& e) {            \
    if (ec) (ec)->err = e.what();                \
    g_host_err = e.what();                       \
    return SYSML_E_INVALID;                      \
  }

extern "C" {

int sysml_b200_ec_create(int device, double mem_util_factor, int eager_cuda_free, int64_t memory_limit, sysml_b200_ec** out) {
  if (!out) return SYSML_E_INVALID;
  *out = nullptr;
  sysml_b200_ctx* ctx = nullptr;
  int rc = sysml_b200_ctx_create(device, &ctx);
  if (rc) {
    g_host_err = sysml_b200_last_error(nullptr);
    return rc;
  }
  sysml_b200_ec* ec = new sysml_b200_ec();
  ec->ctx = ctx;
  ec->mm.ctx = ctx;
  ec->mm.util_factor = mem_util_factor > 0 ? mem_util_factor : 0.9;
  ec->mm.eager_free = eager_cuda_free != 0;
  ec->mm.memory_limit = memory_limit;
  try {
    ec->growWorkspace(1 << 20);
  } catch (const std::exception& e) {
    g_host_err = e.what();
    sysml_b200_ctx_destroy(ctx);
    delete ec;
    return SYSML_E_OOM;
  }
  *out = ec;
  return SYSML_OK;
}

int sysml_b200_ec_destroy(sysml_b200_ec* ec) {
  if (!ec) return SYSML_E_INVALID;
  try {
    sysml_b200_ctx_set_workspace(ec->ctx, nullptr, 0);
    ec->matrices.clear();
    ec->mm.clearMemory();
  } catch (...) {
  }
  sysml_b200_ctx_destroy(ec->ctx);
  delete ec;
  return SYSML_OK;
}

const char* sysml_b200_ec_last_error(sysml_b200_ec* ec) { return ec ? ec->err.c_str() : g_host_err.c_str(); }

int sysml_b200_ec_set_matrix_dense(sysml_b200_ec* ec, const char* name, const double* h_data, int64_t rows, int64_t cols) {
  if (!ec || !name || rows < 0 || cols < 0 || (rows * cols > 0 && !h_data)) return SYSML_E_INVALID;
  EC_TRY(ec, {
    ec->removeMatrix(name);
    std::unique_ptr<MatrixObject> mo(new MatrixObject());
    mo->name = name;
    mo->rows = rows;
    mo->cols = cols;
    mo->h_dense.assign(h_data, h_data + rows * cols);
    mo->host_valid = true;
    ec->matrices[name] = std::move(mo);
  })
}

int sysml_b200_ec_set_matrix_csr(sysml_b200_ec* ec, const char* name, const int32_t* rp, const int32_t* ci, const double* va, int64_t rows, int64_t cols) {
  if (!ec || !name || rows < 0 || cols < 0 || !rp) return SYSML_E_INVALID;
  EC_TRY(ec, {
    ec->removeMatrix(name);
    std::unique_ptr<MatrixObject> mo(new MatrixObject());
    mo->name = name;
    mo->rows = rows;
    mo->cols = cols;
    mo->h_rowptr.assign(rp, rp + rows + 1);
    const i64 nnz = rp[rows];
    mo->h_colind.assign(ci, ci + nnz);
    mo->h_val.assign(va, va + nnz);
    mo->nnz = nnz;
    mo->host_sparse = true;
    mo->host_valid = true;
    ec->matrices[name] = std::move(mo);
  })
}

int sysml_b200_ec_set_matrix_device(sysml_b200_ec* ec, const char* name, double* d_data, int64_t rows, int64_t cols) {
  if (!ec || !name || rows < 0 || cols < 0 || (rows * cols > 0 && !d_data)) return SYSML_E_INVALID;
  EC_TRY(ec, {
    ec->removeMatrix(name);
    std::unique_ptr<MatrixObject> mo(new MatrixObject());
    mo->name = name;
    mo->rows = rows;
    mo->cols = cols;
    MatrixObject* p = mo.get();
    ec->matrices[name] = std::move(mo);
    GPUObject* g = ec->gpuObject(p);
    g->dense = d_data;
    g->borrowed = true;
    g->dirty = true;  // the device copy is the only one
  })
}

int sysml_b200_ec_set_scalar(sysml_b200_ec* ec, const char* name, double value) {
  if (!ec || !name) return SYSML_E_INVALID;
  ec->scalars[name] = value;
  return SYSML_OK;
}

int sysml_b200_ec_remove(sysml_b200_ec* ec, const char* name) {
  if (!ec || !name) return SYSML_E_INVALID;
  EC_TRY(ec, {
    ec->removeMatrix(name);
    ec->scalars.erase(name);
  })
}

int sysml_b200_ec_execute(sysml_b200_ec* ec, const char* instruction) {
  if (!ec || !instruction) return SYSML_E_INVALID;
  EC_TRY(ec, {
    ParsedInstruction pi = parseSingleInstruction(instruction);
    // The kernels' scratch is a block of the memory manager, like every other device allocation of this context.  An
    // instruction that needs more than the block holds fails without side effects (SYSML_E_WORKSPACE): the block is
    // re-allocated at the requested size -- through the same malloc ladder, evictions included -- and the instruction
    // runs again (it overwrites its outputs, so that is safe).
    for (int attempt = 0;; attempt++) {
      try {
        ec->processInstruction(pi);
        break;
      } catch (const DMLRuntimeException& e) {
        if (e.code != SYSML_E_WORKSPACE || attempt >= 3) throw;
        ec->n_inst--;
        ec->growWorkspace(sysml_b200_ctx_workspace_needed(ec->ctx));
      }
    }
  })
}

int sysml_b200_parse_instruction(const char* instruction, char* buf, int buflen) {
  if (!instruction || !buf || buflen <= 0) return SYSML_E_INVALID;
  sysml_b200_ec* none = nullptr;
  EC_TRY(none, {
    ParsedInstruction pi = parseSingleInstruction(instruction);
    static const char* names[] = {"AggregateBinary", "MMTSJ", "MMChain", "AggregateUnary", "ArithmeticBinary", "RelationalBinary", "BuiltinBinary", "BuiltinUnary", "Reorg"};
    std::string s = std::string(names[pi.type]) + " " + pi.opcode;
    for (auto& o : pi.ops) s += " " + o.name + ":" + o.dataType + (o.literal ? ":lit" : "");
    for (auto& e : pi.extras) s += " [" + e + "]";
    snprintf(buf, buflen, "%s", s.c_str());
  })
}

int sysml_b200_ec_get_matrix_dims(sysml_b200_ec* ec, const char* name, int64_t* rows, int64_t* cols, int64_t* nnz) {
  if (!ec || !name) return SYSML_E_INVALID;
  EC_TRY(ec, {
    MatrixObject* mo = ec->getMatrixObject(name);
    if (rows) *rows = mo->rows;
    if (cols) *cols = mo->cols;
    if (nnz) *nnz = mo->nnz;
  })
}

int sysml_b200_ec_get_matrix_dense(sysml_b200_ec* ec, const char* name, double* h_out) {
  if (!ec || !name || !h_out) return SYSML_E_INVALID;
  EC_TRY(ec, {
    MatrixObject* mo = ec->getMatrixObject(name);
    if (mo->gpu) mo->gpu->acquireHostRead();
    if (!mo->host_valid) fail(SYSML_E_INVALID, "variable %s has no valid copy", name);
    if (mo->host_sparse) {
      std::fill(h_out, h_out + mo->rows * mo->cols, 0.0);
      for (i64 r = 0; r < mo->rows; r++)
        for (int k = mo->h_rowptr[r]; k < mo->h_rowptr[r + 1]; k++) h_out[r * mo->cols + mo->h_colind[k]] = mo->h_val[k];
    } else {
      std::copy(mo->h_dense.begin(), mo->h_dense.end(), h_out);
    }
  })
}

int sysml_b200_ec_is_sparse_on_device(sysml_b200_ec* ec, const char* name, int* is_sparse) {
  if (!ec || !name || !is_sparse) return SYSML_E_INVALID;
  EC_TRY(ec, {
    MatrixObject* mo = ec->getMatrixObject(name);
    *is_sparse = (mo->gpu && mo->gpu->isAllocated()) ? (mo->gpu->isSparse() ? 1 : 0) : -1;
  })
}

int sysml_b200_ec_get_scalar(sysml_b200_ec* ec, const char* name, double* value) {
  if (!ec || !name || !value) return SYSML_E_INVALID;
  auto it = ec->scalars.find(name);
  if (it == ec->scalars.end()) {
    ec->err = std::string("Scalar variable ") + name + " does not exist";
    return SYSML_E_INVALID;
  }
  *value = it->second;
  return SYSML_OK;
}

int sysml_b200_ec_set_eviction_policy(sysml_b200_ec* ec, const char* policy) {
  if (!ec || !policy) return SYSML_E_INVALID;
  EC_TRY(ec, {
    std::string p = policy;
    for (auto& c : p) c = (char)tolower((unsigned char)c);
    if (p == "lru") ec->mm.policy = LRU;
    else if (p == "lfu") ec->mm.policy = LFU;
    else if (p == "min_evict") ec->mm.policy = MIN_EVICT;
    else if (p == "mru") ec->mm.policy = MRU;
    else fail(SYSML_E_INVALID, "The eviction policy is not supported:%s", policy);
  })
}

int sysml_b200_cellwise_sparse_safe(const char* opcode, int kind, double constant, int scalar_left) {
  if (!opcode) return -1;
  const std::string o = opcode;
  switch (kind) {
    case 0: return scalarSparseSafe(o, constant, scalar_left != 0) ? 1 : 0;
    case 1: return binarySparseSafe(o) ? 1 : 0;
    case 2: return unarySparseSafe(o) ? 1 : 0;
  }
  return -1;
}

int sysml_b200_eval_sparse_format_in_memory(int64_t rows, int64_t cols, int64_t nnz) {
  return evalSparseFormatInMemory(rows, cols, nnz) ? 1 : 0;
}

int sysml_b200_ec_exam_sparsity(sysml_b200_ec* ec, const char* name, int* is_sparse) {
  if (!ec || !name) return SYSML_E_INVALID;
  EC_TRY(ec, {
    MatrixObject* mo = ec->getMatrixObject(name);
    if (mo->gpu) mo->gpu->acquireHostRead();
    if (!mo->host_valid) fail(SYSML_E_INVALID, "variable %s has no valid copy", name);
    mo->examSparsity();
    if (is_sparse) *is_sparse = mo->host_sparse ? 1 : 0;
  })
}

int sysml_b200_ec_stats(sysml_b200_ec* ec, int64_t* out9) {
  if (!ec || !out9) return SYSML_E_INVALID;
  out9[0] = ec->n_inst;
  out9[1] = ec->n_h2d;
  out9[2] = ec->n_d2h;
  out9[3] = ec->mm.n_malloc;
  out9[4] = ec->mm.n_reuse;
  out9[5] = ec->mm.n_free;
  out9[6] = ec->mm.n_evict;
  out9[7] = ec->mm.allocated_bytes();
  out9[8] = ec->mm.rmvar_bytes();
  return SYSML_OK;
}

}  // extern "C"
