// kmeans_pipes_gpu.cc -- the K-means GPU map task as a Hadoop Pipes program (the -gpubin).
//
// Submitted exactly like the reference expects (pipes/Submitter.java:419-420, 349-379):
//   bin/hadoop pipes -cpubin kmeans-pipes-cpu -gpubin kmeans-pipes-gpu -input ... -output ...
//        -jobconf kmeans.k=128,kmeans.d=2,kmeans.centroids=centroids.bin
// (no -inputformat / -reader: hadoop.pipes.java.recordreader is then absent, which Java reads as false,
// pipes/Submitter.java:137-139 -- the C++ reader route below)
// and started by pipes/Application.java:156-194 as `kmeans-pipes-gpu <GPUDeviceId>`.
// It is written against the Pipes C++ API only (hadoop/Pipes.hh), so it builds against the
// reference's libhadooppipes.a as well as against this repository's runtime, and reaches
// CUDA only through the C ABI of libhgpumap.so (include/hgpumap.h).
//
// Same record contract as the CPU program of the job (JobConf keys kmeans.k, kmeans.d,
// kmeans.centroids, kmeans.input.format; map output key = decimal cluster id, value =
// "count s0 s1 ..."), with one difference the protocol allows: the mapper buffers its
// input, maps whole batches on the GPU and emits ALREADY COMBINED records from close(), so
// createCombiner() returns NULL and the runtime's CombineRunner (HadoopPipes.cc:660-706)
// never buffers a pair.  Records leave in the byte order of their keys, like a combiner
// spill would emit them.
//
// Input routes:
//   piped (hadoop.pipes.java.recordreader=true):  MAP_ITEM values are appended to a pinned
//       batch; kmeans.input.format = binary (rows of d LE float32) | binary-be | text (one
//       TextInputFormat line per record, decoded on the device: kernels_text.cu) | path
//       (value names a file [\t offset \t length] of kmeans.file.format rows, the shipped
//       NLineInputFormat config, conf/mapred-site.xml:13-22);
//   C++ reader (=false): the FileSplit (mapred/FileSplit.java:81-85) is handed to
//       hgm_map_file and read straight into pinned memory -- no per-record traffic at all
//       (text: framed like LineRecordReader, mapred/LineRecordReader.java:79-95; sequencefile: like
//       SequenceFileRecordReader, records walked on the device between sync entries).
// There is no CPU fallback: any hgm_* failure throws HadoopUtils::Error and the task fails.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "hadoop/Pipes.hh"
#include "hadoop/SerialUtils.hh"
#include "hadoop/StringUtils.hh"
#include "hadoop/TemplateFactory.hh"
#include "hgpumap.h"
#include "../worker/worker_client.hh"

#include <dlfcn.h>
#include <unistd.h>

static int g_device = 0;

// libhgpumap.so is loaded on first use, not linked: a task that hands its splits to the slot's worker never
// touches CUDA, and loading the library (9 MB of device code to register) was 1.8 of the 2.6 ms such a child
// takes to start and exit.  Looked for beside this executable, where the reference-runtime build keeps it, then
// on the loader's path.  There is still no CPU fallback: no library, no task.
namespace gpulib {
decltype(&::hgm_open) fn_hgm_open = nullptr;
decltype(&::hgm_close) fn_hgm_close = nullptr;
decltype(&::hgm_last_error) fn_hgm_last_error = nullptr;
decltype(&::hgm_set_centroids) fn_hgm_set_centroids = nullptr;
decltype(&::hgm_map_split) fn_hgm_map_split = nullptr;
decltype(&::hgm_map_file) fn_hgm_map_file = nullptr;
decltype(&::hgm_host_alloc) fn_hgm_host_alloc = nullptr;
decltype(&::hgm_host_free) fn_hgm_host_free = nullptr;
decltype(&::hgm_last_map_ms) fn_hgm_last_map_ms = nullptr;
decltype(&::hgm_last_rechecks) fn_hgm_last_rechecks = nullptr;
decltype(&::hgm_staged_bytes) fn_hgm_staged_bytes = nullptr;
void load() {
  if (fn_hgm_open) return;
  char exe[4096];
  const ssize_t n = readlink("/proc/self/exe", exe, sizeof exe - 1);
  std::string dir = n > 0 ? std::string(exe, (size_t)n) : std::string(".");
  dir = dir.substr(0, dir.find_last_of('/'));
  const std::string tries[] = {dir + "/libhgpumap.so", dir + "/../../hadoop-1.0.3-gpu_b200/libhgpumap.so", "libhgpumap.so"};
  void *lib = NULL;
  std::string why;
  for (const std::string &path : tries) {
    lib = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (lib) break;
    why += std::string(dlerror()) + "; ";
  }
  if (!lib) throw HadoopUtils::Error("cannot load libhgpumap.so (there is no CPU fallback in a GPU map slot): " + why);
  fn_hgm_close = reinterpret_cast<decltype(fn_hgm_close)>(dlsym(lib, "hgm_close"));
  if (!fn_hgm_close) throw HadoopUtils::Error("libhgpumap.so lacks hgm_close");
  fn_hgm_last_error = reinterpret_cast<decltype(fn_hgm_last_error)>(dlsym(lib, "hgm_last_error"));
  if (!fn_hgm_last_error) throw HadoopUtils::Error("libhgpumap.so lacks hgm_last_error");
  fn_hgm_set_centroids = reinterpret_cast<decltype(fn_hgm_set_centroids)>(dlsym(lib, "hgm_set_centroids"));
  if (!fn_hgm_set_centroids) throw HadoopUtils::Error("libhgpumap.so lacks hgm_set_centroids");
  fn_hgm_map_split = reinterpret_cast<decltype(fn_hgm_map_split)>(dlsym(lib, "hgm_map_split"));
  if (!fn_hgm_map_split) throw HadoopUtils::Error("libhgpumap.so lacks hgm_map_split");
  fn_hgm_map_file = reinterpret_cast<decltype(fn_hgm_map_file)>(dlsym(lib, "hgm_map_file"));
  if (!fn_hgm_map_file) throw HadoopUtils::Error("libhgpumap.so lacks hgm_map_file");
  fn_hgm_host_alloc = reinterpret_cast<decltype(fn_hgm_host_alloc)>(dlsym(lib, "hgm_host_alloc"));
  if (!fn_hgm_host_alloc) throw HadoopUtils::Error("libhgpumap.so lacks hgm_host_alloc");
  fn_hgm_host_free = reinterpret_cast<decltype(fn_hgm_host_free)>(dlsym(lib, "hgm_host_free"));
  if (!fn_hgm_host_free) throw HadoopUtils::Error("libhgpumap.so lacks hgm_host_free");
  fn_hgm_last_map_ms = reinterpret_cast<decltype(fn_hgm_last_map_ms)>(dlsym(lib, "hgm_last_map_ms"));
  if (!fn_hgm_last_map_ms) throw HadoopUtils::Error("libhgpumap.so lacks hgm_last_map_ms");
  fn_hgm_last_rechecks = reinterpret_cast<decltype(fn_hgm_last_rechecks)>(dlsym(lib, "hgm_last_rechecks"));
  if (!fn_hgm_last_rechecks) throw HadoopUtils::Error("libhgpumap.so lacks hgm_last_rechecks");
  fn_hgm_staged_bytes = reinterpret_cast<decltype(fn_hgm_staged_bytes)>(dlsym(lib, "hgm_staged_bytes"));
  if (!fn_hgm_staged_bytes) throw HadoopUtils::Error("libhgpumap.so lacks hgm_staged_bytes");
  fn_hgm_open = reinterpret_cast<decltype(fn_hgm_open)>(dlsym(lib, "hgm_open"));
  if (!fn_hgm_open) throw HadoopUtils::Error("libhgpumap.so lacks hgm_open");
}
}  // namespace gpulib
#define hgm_open gpulib::fn_hgm_open
#define hgm_close gpulib::fn_hgm_close
#define hgm_last_error gpulib::fn_hgm_last_error
#define hgm_set_centroids gpulib::fn_hgm_set_centroids
#define hgm_map_split gpulib::fn_hgm_map_split
#define hgm_map_file gpulib::fn_hgm_map_file
#define hgm_host_alloc gpulib::fn_hgm_host_alloc
#define hgm_host_free gpulib::fn_hgm_host_free
#define hgm_last_map_ms gpulib::fn_hgm_last_map_ms
#define hgm_last_rechecks gpulib::fn_hgm_last_rechecks
#define hgm_staged_bytes gpulib::fn_hgm_staged_bytes

namespace {

const std::string KMEANS = "KMEANS";

void check(hgm_ctx *ctx, int rc, const char *what) {
  if (rc != HGM_OK) throw HadoopUtils::Error(std::string(what) + ": " + hgm_last_error(ctx));
}

std::string partialToString(long long count, const double *sums, int d) {
  char buf[64];
  snprintf(buf, sizeof buf, "%lld", count);
  std::string out(buf);
  for (int j = 0; j < d; ++j) {
    snprintf(buf, sizeof buf, " %.17g", sums[j]);
    out += buf;
  }
  return out;
}

// FileSplit as Java serialises it: UTF8.writeString(path) = 2-byte BE length + bytes, then
// two big-endian longs (mapred/FileSplit.java:81-85).  Also accepts the vint-length form the
// reference's test fixture uses (test/.../pipes/WordCountInputFormat.java:42-44).
struct Split {
  std::string path;
  int64_t start = 0, length = -1;
};
Split parseSplit(const std::string &raw) {
  Split s;
  const unsigned char *p = reinterpret_cast<const unsigned char *>(raw.data());
  const size_t n = raw.size();
  if (n >= 2) {
    const size_t len = ((size_t)p[0] << 8) | p[1];
    if (2 + len + 16 == n) {
      s.path.assign(raw, 2, len);
      auto be64 = [](const unsigned char *q) { int64_t v = 0; for (int i = 0; i < 8; ++i) v = (v << 8) | q[i]; return v; };
      s.start = be64(p + 2 + len);
      s.length = be64(p + 2 + len + 8);
      return s;
    }
  }
  HadoopUtils::StringInStream in(raw);
  HadoopUtils::deserializeString(s.path, in);
  return s;
}
std::string stripScheme(const std::string &uri) {  // "file:/x" -> "/x"
  if (uri.compare(0, 5, "file:") == 0) {
    std::string rest = uri.substr(5);
    while (rest.size() > 1 && rest[0] == '/' && rest[1] == '/') rest.erase(0, 1);
    return rest;
  }
  return uri;
}

}  // namespace

static int formatCode(const std::string &f) {
  if (f == "binary") return HGM_FMT_F32LE;
  if (f == "binary-be") return HGM_FMT_F32BE;
  if (f == "text") return HGM_FMT_TEXT;
  if (f == "sequencefile") return HGM_FMT_SEQ_F32BE;     // BytesWritable rows as DataOutput.writeFloat wrote them
  HADOOP_ASSERT(f == "sequencefile-le", "record format " + f + " is not one of binary, binary-be, text, sequencefile, sequencefile-le");
  return HGM_FMT_SEQ_F32LE;
}

class KMeansGpuMap : public HadoopPipes::Mapper {
  HadoopPipes::MapContext *ctx_;
  hgm_ctx *gpu_ = nullptr;             // own CUDA context ...
  hgmw::Client *worker_ = nullptr;     // ... or the slot's persistent worker (kmeans.gpu.worker.socket)
  bool worker_cache_ = false;          // kmeans.gpu.worker.cache: keep file splits in the worker's HBM cache
  HadoopPipes::TaskContext::Counter *cacheHits_ = nullptr;
  int k_, d_;
  enum { BINARY, PATH, SPLIT } format_ = BINARY;
  int fmt_code_ = HGM_FMT_F32LE;
  std::vector<double> sums_, part_sums_;
  std::vector<int64_t> counts_, part_counts_;
  unsigned char *batch_ = nullptr;  // pinned
  size_t batch_cap_ = 0, batch_used_ = 0;
  int64_t points_ = 0;
  HadoopPipes::TaskContext::Counter *pointsIn_, *gpuMicros_, *batches_, *rechecks_, *staged_;
  int64_t staged_seen_ = 0;

  void fold() {
    for (size_t i = 0; i < sums_.size(); ++i) sums_[i] += part_sums_[i];
    for (size_t i = 0; i < counts_.size(); ++i) counts_[i] += part_counts_[i];
  }
  void account(int64_t n, const hgmw::MapReply *from_worker = nullptr) {
    float ms = 0;
    int32_t launches = 0;
    int64_t rechecks = 0, staged = 0;
    if (from_worker) {
      ms = from_worker->map_ms;
      rechecks = from_worker->rechecks;
      staged = from_worker->staged_bytes;
      if (from_worker->cache_hit) ctx_->incrementCounter(cacheHits_, (uint64_t)from_worker->cache_hit);
    } else {
      hgm_last_map_ms(gpu_, &ms, &launches);
      hgm_last_rechecks(gpu_, &rechecks);
      int64_t total = 0;
      hgm_staged_bytes(gpu_, &total);
      staged = total - staged_seen_;
      staged_seen_ = total;
    }
    if (rechecks) ctx_->incrementCounter(rechecks_, (uint64_t)rechecks);
    if (staged) ctx_->incrementCounter(staged_, (uint64_t)staged);
    points_ += n;
    ctx_->incrementCounter(pointsIn_, (uint64_t)n);
    ctx_->incrementCounter(gpuMicros_, (uint64_t)(ms * 1000.0f));
    ctx_->incrementCounter(batches_, 1);
  }
  void flush() {
    if (batch_used_ == 0) return;
    int64_t n = 0;
    if (worker_) {
      hgmw::MapReply info;
      try {
        n = worker_->map_bytes(batch_, batch_used_, fmt_code_, part_sums_.data(), part_counts_.data(), &info);
      } catch (const std::exception &e) {
        throw HadoopUtils::Error(e.what());
      }
      fold();
      account(n, &info);
      batch_used_ = 0;
      return;
    }
    check(gpu_, hgm_map_split(gpu_, batch_, batch_used_, fmt_code_, part_sums_.data(), part_counts_.data(), nullptr, &n),
          "hgm_map_split");
    fold();
    account(n);
    batch_used_ = 0;
  }
  std::vector<hgmw::Client::FileSplit> worker_files_;  // path records of this task, sent to the worker in one request

  void flushFiles() {
    if (worker_files_.empty()) return;
    hgmw::MapReply info;
    int64_t n = 0;
    try {
      n = worker_->map_files(worker_files_, fmt_code_, worker_cache_, part_sums_.data(), part_counts_.data(), &info);
    } catch (const std::exception &e) {
      throw HadoopUtils::Error(e.what());
    }
    worker_files_.clear();
    fold();
    account(n, &info);
  }

  void mapFile(const std::string &path, int64_t off, int64_t len) {
    int64_t n = 0;
    if (worker_ && format_ == PATH) {  // a batch of splits per task: one round trip to the worker for all of them
      worker_files_.push_back(hgmw::Client::FileSplit{path, off, len});
      if (worker_files_.size() >= 4096) flushFiles();
      return;
    }
    if (worker_) {
      hgmw::MapReply info;
      try {
        n = worker_->map_file(path, off, len, fmt_code_, worker_cache_, part_sums_.data(), part_counts_.data(), &info);
      } catch (const std::exception &e) {
        throw HadoopUtils::Error(e.what());
      }
      fold();
      account(n, &info);
      return;
    }
    check(gpu_, hgm_map_file(gpu_, path.c_str(), off, len, fmt_code_, part_sums_.data(), part_counts_.data(), nullptr, &n),
          "hgm_map_file");
    fold();
    account(n);
  }

 public:
  explicit KMeansGpuMap(HadoopPipes::MapContext &context) : ctx_(&context) {
    const HadoopPipes::JobConf *conf = context.getJobConf();
    k_ = conf->getInt("kmeans.k");
    d_ = conf->getInt("kmeans.d");
    HADOOP_ASSERT(k_ > 0 && d_ > 0, "kmeans.k and kmeans.d must be positive");
    if (conf->hasKey("kmeans.input.format")) {
      const std::string &f = conf->get("kmeans.input.format");
      if (f == "path") {
        format_ = PATH;
        if (conf->hasKey("kmeans.file.format")) fmt_code_ = formatCode(conf->get("kmeans.file.format"));
      } else {
        fmt_code_ = formatCode(f);
      }
    }
    // absent = false, as Submitter.getIsJavaRecordReader reads it (pipes/Submitter.java:137-139): Java then sends
    // RUN_MAP with pipedInput = false and no MAP_ITEMs, and the runtime asks the factory for a reader
    if (!(conf->hasKey("hadoop.pipes.java.recordreader") && conf->getBoolean("hadoop.pipes.java.recordreader")))
      format_ = SPLIT;
    // a SequenceFile is decoded from the file (the sync marker is in its header); piped through Java's
    // SequenceFileRecordReader its BytesWritable values arrive as plain bytes: that is "binary-be"
    HADOOP_ASSERT(format_ != BINARY || (fmt_code_ != HGM_FMT_SEQ_F32BE && fmt_code_ != HGM_FMT_SEQ_F32LE),
                  "sequencefile input needs hadoop.pipes.java.recordreader=false or kmeans.input.format=path");
    if (conf->hasKey("kmeans.gpu.device")) g_device = conf->getInt("kmeans.gpu.device");

    const std::string &cpath = conf->get("kmeans.centroids");
    FILE *f = fopen(cpath.c_str(), "rb");
    HADOOP_ASSERT(f != NULL, "cannot open centroid file " + cpath);
    std::vector<float> cent((size_t)k_ * d_);
    const size_t got = fread(cent.data(), sizeof(float), cent.size(), f);
    fclose(f);
    HADOOP_ASSERT(got == cent.size(), "centroid file " + cpath + " is shorter than k*d floats");

    if (conf->hasKey("kmeans.gpu.worker.socket")) {
      // the slot's persistent worker owns the CUDA context (csrc/worker/): no ~1 s of context creation
      // per task, and file splits can stay in its HBM cache from one iteration's job to the next
      std::string sock = conf->get("kmeans.gpu.worker.socket");
      const size_t at = sock.find("%d");
      if (at != std::string::npos) sock.replace(at, 2, HadoopUtils::toString(g_device));
      worker_cache_ = conf->hasKey("kmeans.gpu.worker.cache") && conf->getBoolean("kmeans.gpu.worker.cache");
      try {
        worker_ = new hgmw::Client(sock);
        if (conf->hasKey("kmeans.gpu.worker.cache.epoch")) worker_->cache_epoch(conf->get("kmeans.gpu.worker.cache.epoch"));
        worker_->set_centroids(cent.data(), k_, d_);
      } catch (const std::exception &e) {
        throw HadoopUtils::Error(e.what());  // no worker, no task: the scheduler re-queues the split
      }
    } else {
      gpulib::load();
      check(nullptr, hgm_open(g_device, &gpu_), "hgm_open");
      check(gpu_, hgm_set_centroids(gpu_, cent.data(), k_, d_), "hgm_set_centroids");
    }
    sums_.assign((size_t)k_ * d_, 0.0);
    part_sums_.assign((size_t)k_ * d_, 0.0);
    counts_.assign(k_, 0);
    part_counts_.assign(k_, 0);
    if (format_ == BINARY) {
      const size_t row = 4 * (size_t)d_;
      batch_cap_ = fmt_code_ == HGM_FMT_TEXT ? (size_t(64) << 20) : std::max<size_t>(row, ((size_t(64) << 20) / row) * row);
      if (worker_) {  // the batch crosses a socket, not PCIe: ordinary memory
        batch_ = static_cast<unsigned char *>(malloc(batch_cap_));
        HADOOP_ASSERT(batch_ != NULL, "cannot allocate the record batch");
      } else {
        void *p = nullptr;
        check(gpu_, hgm_host_alloc(gpu_, batch_cap_, &p), "hgm_host_alloc");
        batch_ = static_cast<unsigned char *>(p);
      }
    }
    cacheHits_ = context.getCounter(KMEANS, "GPU_CACHE_HITS");
    pointsIn_ = context.getCounter(KMEANS, "POINTS");
    gpuMicros_ = context.getCounter(KMEANS, "GPU_MAP_MICROS");
    batches_ = context.getCounter(KMEANS, "GPU_BATCHES");
    rechecks_ = context.getCounter(KMEANS, "GPU_RECHECKS");      // points the tensor-core path re-scanned exactly
    staged_ = context.getCounter(KMEANS, "GPU_BYTES_STAGED");    // split bytes copied host -> device
  }

  ~KMeansGpuMap() {
    if (gpu_) {
      if (batch_) hgm_host_free(gpu_, batch_);
      hgm_close(gpu_);
    } else if (worker_) {
      free(batch_);
      delete worker_;
    }
  }

  void map(HadoopPipes::MapContext &context) {
    const std::string &v = context.getInputValue();
    if (format_ == BINARY && fmt_code_ == HGM_FMT_TEXT) {
      // one TextInputFormat line per record: batch the lines, the device splits and parses them
      HADOOP_ASSERT(v.find_first_of("\r\n") == std::string::npos, "text record holds a line end");
      HADOOP_ASSERT(v.size() + 1 <= batch_cap_, "text record longer than the staging batch");
      if (batch_used_ + v.size() + 1 > batch_cap_) flush();
      memcpy(batch_ + batch_used_, v.data(), v.size());
      batch_used_ += v.size();
      batch_[batch_used_++] = '\n';
    } else if (format_ == BINARY) {
      HADOOP_ASSERT(v.size() % (4 * (size_t)d_) == 0, "binary record is not a whole number of rows");
      size_t off = 0;
      while (off < v.size()) {
        if (batch_used_ == batch_cap_) flush();
        const size_t take = std::min(v.size() - off, batch_cap_ - batch_used_);
        memcpy(batch_ + batch_used_, v.data() + off, take);
        batch_used_ += take;
        off += take;
      }
    } else if (format_ == PATH) {
      std::vector<std::string> parts = HadoopUtils::splitString(v, "\t");
      HADOOP_ASSERT(parts.size() == 1 || parts.size() == 3, "path record must be file[\\toffset\\tlength]: " + v);
      mapFile(stripScheme(parts[0]), parts.size() == 3 ? atoll(parts[1].c_str()) : 0,
              parts.size() == 3 ? atoll(parts[2].c_str()) : -1);
    } else {  // SPLIT: the reader hands over the split descriptor as the key
      const Split s = parseSplit(context.getInputKey());
      mapFile(stripScheme(s.path), s.start, s.length);
    }
  }

  // emit the combined output of the whole task (HadoopPipes.cc:1007-1020 calls close() on
  // the mapper before the uplink says DONE)
  void close() {
    flush();
    flushFiles();
    std::vector<std::string> keys;
    for (int c = 0; c < k_; ++c)
      if (counts_[c] != 0) keys.push_back(HadoopUtils::toString(c));
    std::sort(keys.begin(), keys.end());  // byte order of std::map<string,...> in CombineRunner
    for (const std::string &key : keys) {
      const int c = atoi(key.c_str());
      ctx_->emit(key, partialToString((long long)counts_[c], &sums_[(size_t)c * d_], d_));
    }
  }
};

// Hands the mapper one record: the serialised split itself (C++ reader route,
// HadoopPipes.cc:793-799, 874-882).
class WholeSplitReader : public HadoopPipes::RecordReader {
  std::string split_;
  bool given_ = false;

 public:
  explicit WholeSplitReader(HadoopPipes::MapContext &context) : split_(context.getInputSplit()) {}
  bool next(std::string &key, std::string &value) {
    if (given_) return false;
    key = split_;
    value.clear();
    given_ = true;
    return true;
  }
  float getProgress() { return given_ ? 1.0f : 0.0f; }
};

// Reduce side = centroid update; identical to the CPU program's (it normally runs from the
// -cpubin, pipes/PipesReducer.java:59-93 starts the child with runOnGPU=false).
class KMeansReduce : public HadoopPipes::Reducer {
  int d_;

 public:
  explicit KMeansReduce(HadoopPipes::TaskContext &context) : d_(context.getJobConf()->getInt("kmeans.d")) {}
  void reduce(HadoopPipes::ReduceContext &context) {
    long long count = 0;
    std::vector<double> sums(d_, 0.0);
    while (context.nextValue()) {
      const char *s = context.getInputValue().c_str();
      char *end;
      count += strtoll(s, &end, 10);
      for (int j = 0; j < d_; ++j) {
        s = end;
        sums[j] += strtod(s, &end);
        HADOOP_ASSERT(end != s, "malformed partial record " + context.getInputValue());
      }
    }
    char buf[64];
    snprintf(buf, sizeof buf, "%lld", count);
    std::string out(buf);
    for (int j = 0; j < d_; ++j) {
      snprintf(buf, sizeof buf, " %.9g", (float)(sums[j] / (double)count));
      out += buf;
    }
    context.emit(context.getInputKey(), out);
  }
};

class KMeansPartitioner : public HadoopPipes::Partitioner {
 public:
  explicit KMeansPartitioner(HadoopPipes::TaskContext &) {}
  int partition(const std::string &key, int numOfReduces) { return HadoopUtils::toInt(key) % numOfReduces; }
};

// The runtime creates the C++ reader only when Java says the input is not piped.
class GpuFactory : public HadoopPipes::Factory {
 public:
  HadoopPipes::Mapper *createMapper(HadoopPipes::MapContext &c) const { return new KMeansGpuMap(c); }
  HadoopPipes::Reducer *createReducer(HadoopPipes::ReduceContext &c) const { return new KMeansReduce(c); }
  HadoopPipes::Partitioner *createPartitioner(HadoopPipes::MapContext &c) const { return new KMeansPartitioner(c); }
  HadoopPipes::RecordReader *createRecordReader(HadoopPipes::MapContext &c) const {
    // The runtime asks on every map task and insists that a reader exists exactly when RUN_MAP said the input is
    // not piped (HadoopPipes.cc:793-799).  The factory cannot see that flag, only the key Java derived it from:
    // absent = false = not piped (pipes/Submitter.java:137-139).
    const HadoopPipes::JobConf *conf = c.getJobConf();
    if (conf->hasKey("hadoop.pipes.java.recordreader") && conf->getBoolean("hadoop.pipes.java.recordreader")) return NULL;
    return new WholeSplitReader(c);
  }
};

int main(int argc, char **argv) {
  // pipes/Application.java:179-181 appends the GPU device id to the command line
  if (argc > 1) g_device = atoi(argv[argc - 1]);
  else if (const char *env = getenv("HGM_DEVICE")) g_device = atoi(env);
  // As in the shipped examples the bool is the exit status (wordcount-simple.cc:63-66);
  // Hadoop ignores it and waits for DONE.
  return HadoopPipes::runTask(GpuFactory());
}
