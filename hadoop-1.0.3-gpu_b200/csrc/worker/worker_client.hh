// worker_client.hh -- the map task's side of worker_proto.hh: what kmeans_pipes_gpu.cc uses instead of
// its own hgm_ctx when the job names a worker socket (kmeans.gpu.worker.socket, "%d" = device id).
// Errors throw std::runtime_error with the worker's message; the caller turns them into a failed task.
#pragma once
#include <errno.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <stdexcept>
#include <string>
#include <vector>

#include "worker_proto.hh"

namespace hgmw {

class Client {
  int fd_ = -1;
  int k_ = 0, d_ = 0;

  void put(const void *buf, size_t n) {
    const char *p = static_cast<const char *>(buf);
    while (n) {
      const ssize_t w = write(fd_, p, n);
      if (w < 0) { if (errno == EINTR) continue; throw std::runtime_error(std::string("GPU worker write: ") + strerror(errno)); }
      p += w;
      n -= (size_t)w;
    }
  }
  void get(void *buf, size_t n) {
    char *p = static_cast<char *>(buf);
    while (n) {
      const ssize_t r = read(fd_, p, n);
      if (r == 0) throw std::runtime_error("GPU worker closed the connection");
      if (r < 0) { if (errno == EINTR) continue; throw std::runtime_error(std::string("GPU worker read: ") + strerror(errno)); }
      p += r;
      n -= (size_t)r;
    }
  }
  std::vector<char> reply(uint32_t op) {
    Header h;
    get(&h, sizeof h);
    if (h.magic != kMagic || h.op != op) throw std::runtime_error("GPU worker protocol error");
    std::vector<char> p((size_t)h.payload);
    if (h.payload) get(p.data(), p.size());
    if (h.status != 0) throw std::runtime_error("GPU worker: " + std::string(p.begin(), p.end()));
    return p;
  }
  int64_t map_reply(uint32_t op, double *sums, int64_t *counts, MapReply *info) {
    const std::vector<char> p = reply(op);
    const size_t want = sizeof(MapReply) + (size_t)k_ * d_ * 8 + (size_t)k_ * 8;
    if (p.size() != want) throw std::runtime_error("GPU worker: map reply of the wrong size");
    MapReply r;
    memcpy(&r, p.data(), sizeof r);
    memcpy(sums, p.data() + sizeof r, (size_t)k_ * d_ * 8);
    memcpy(counts, p.data() + sizeof r + (size_t)k_ * d_ * 8, (size_t)k_ * 8);
    if (info) *info = r;
    return r.n_points;
  }

 public:
  explicit Client(const std::string &path) {
    sockaddr_un addr{};
    if (path.size() >= sizeof addr.sun_path) throw std::runtime_error("GPU worker socket path too long: " + path);
    fd_ = socket(AF_UNIX, SOCK_STREAM, 0);
    addr.sun_family = AF_UNIX;
    memcpy(addr.sun_path, path.c_str(), path.size() + 1);
    if (fd_ < 0 || connect(fd_, reinterpret_cast<sockaddr *>(&addr), sizeof addr) != 0) {
      const std::string why = strerror(errno);
      if (fd_ >= 0) close(fd_);
      fd_ = -1;
      throw std::runtime_error("cannot reach the GPU worker at " + path + ": " + why);
    }
    const uint32_t version = kVersion;
    Header h{kMagic, kHello, 0, 0, sizeof version};
    put(&h, sizeof h);
    put(&version, sizeof version);
    const std::vector<char> p = reply(kHello);
    uint32_t theirs = 0;
    if (p.size() >= 4) memcpy(&theirs, p.data(), 4);
    if (theirs != kVersion) throw std::runtime_error("GPU worker speaks another protocol version");
  }
  ~Client() { if (fd_ >= 0) close(fd_); }
  Client(const Client &) = delete;
  Client &operator=(const Client &) = delete;

  void set_centroids(const float *c, int k, int d) {
    const int32_t kd[2] = {k, d};
    Header h{kMagic, kSetCentroids, 0, 0, sizeof kd + (uint64_t)k * d * 4};
    put(&h, sizeof h);
    put(kd, sizeof kd);
    put(c, (size_t)k * d * 4);
    reply(kSetCentroids);
    k_ = k;
    d_ = d;
  }
  int64_t map_bytes(const void *bytes, size_t nbytes, int32_t fmt, double *sums, int64_t *counts, MapReply *info = nullptr) {
    Header h{kMagic, kMapBytes, 0, 0, 4 + (uint64_t)nbytes};
    put(&h, sizeof h);
    put(&fmt, 4);
    put(bytes, nbytes);
    return map_reply(kMapBytes, sums, counts, info);
  }
  int64_t map_file(const std::string &path, int64_t off, int64_t len, int32_t fmt, bool use_cache, double *sums,
                   int64_t *counts, MapReply *info = nullptr) {
    const int32_t head[2] = {fmt, use_cache ? 1 : 0};
    const int64_t range[2] = {off, len};
    Header h{kMagic, kMapFile, 0, 0, sizeof head + sizeof range + path.size()};
    put(&h, sizeof h);
    put(head, sizeof head);
    put(range, sizeof range);
    put(path.data(), path.size());
    return map_reply(kMapFile, sums, counts, info);
  }
  struct FileSplit { std::string path; int64_t offset, length; };
  // a task's batch of file splits in one request (kMapFiles); info->cache_hit = how many were resident
  int64_t map_files(const std::vector<FileSplit> &splits, int32_t fmt, bool use_cache, double *sums, int64_t *counts,
                    MapReply *info = nullptr) {
    std::vector<char> body;
    const int32_t head[3] = {fmt, use_cache ? 1 : 0, (int32_t)splits.size()};
    body.insert(body.end(), reinterpret_cast<const char *>(head), reinterpret_cast<const char *>(head) + sizeof head);
    for (const FileSplit &f : splits) {
      const int64_t range[2] = {f.offset, f.length};
      const int32_t plen = (int32_t)f.path.size();
      body.insert(body.end(), reinterpret_cast<const char *>(range), reinterpret_cast<const char *>(range) + sizeof range);
      body.insert(body.end(), reinterpret_cast<const char *>(&plen), reinterpret_cast<const char *>(&plen) + 4);
      body.insert(body.end(), f.path.begin(), f.path.end());
    }
    Header h{kMagic, kMapFiles, 0, 0, body.size()};
    put(&h, sizeof h);
    put(body.data(), body.size());
    return map_reply(kMapFiles, sums, counts, info);
  }
  // kmeans.gpu.worker.cache.epoch: the cache holds splits of ONE data set; a task of another one clears it
  void cache_epoch(const std::string &epoch) {
    Header h{kMagic, kDropCache, 0, 0, epoch.size()};
    put(&h, sizeof h);
    if (!epoch.empty()) put(epoch.data(), epoch.size());
    reply(kDropCache);
  }
  WorkerStats stats() {
    Header h{kMagic, kStats, 0, 0, 0};
    put(&h, sizeof h);
    const std::vector<char> p = reply(kStats);
    WorkerStats s{};
    if (p.size() == sizeof s) memcpy(&s, p.data(), sizeof s);
    return s;
  }
  void shutdown() {
    Header h{kMagic, kShutdown, 0, 0, 0};
    put(&h, sizeof h);
    reply(kShutdown);
  }
};

}  // namespace hgmw
