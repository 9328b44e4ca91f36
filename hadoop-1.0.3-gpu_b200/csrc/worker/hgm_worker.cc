// hgm_worker.cc -- the persistent GPU worker of one map slot (see worker_proto.hh).
//
//   hgm-worker --device 0 --socket /tmp/hgm-gpu0.sock [--max-batch-mib 1024]
//
// One process per device, started with the TaskTracker.  It reaches CUDA only
// through the C ABI of libhgpumap.so; one client (= one map task) is served at a time, which is
// exactly the slot discipline of the reference (one task per GPU slot, mapred/TaskTracker.java:
// 2526-2562).  File splits mapped with use_cache=1 stay in the HBM split cache under
// (path, offset, length, format, mtime, size): the next iteration's task for the same split maps
// from HBM without touching the file (BASELINE config 3).  The cache is bounded by HBM: when a split does not
// fit, stale versions of the same split (the file changed) and then the least recently used splits are dropped
// until it does; kDropCache empties it (a new job's first task).  A MAP_BYTES batch is pinned host memory:
// batches over --max-batch-mib are refused.
#include <errno.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/un.h>
#include <unistd.h>

#include <map>
#include <string>
#include <vector>

#include "hgpumap.h"
#include "worker_proto.hh"

namespace {

bool read_all(int fd, void *buf, size_t n) {
  char *p = static_cast<char *>(buf);
  while (n) {
    const ssize_t got = read(fd, p, n);
    if (got == 0) return false;
    if (got < 0) { if (errno == EINTR) continue; return false; }
    p += got;
    n -= (size_t)got;
  }
  return true;
}
bool write_all(int fd, const void *buf, size_t n) {
  const char *p = static_cast<const char *>(buf);
  while (n) {
    const ssize_t put = write(fd, p, n);
    if (put < 0) { if (errno == EINTR) continue; return false; }
    p += put;
    n -= (size_t)put;
  }
  return true;
}

struct Worker {
  hgm_ctx *gpu = nullptr;
  int device = 0;
  int k = 0, d = 0;
  std::vector<float> centroids;
  void *pinned = nullptr;
  size_t pinned_cap = 0;
  struct Cached { int64_t id; uint64_t last_use; std::string split; };  // split = key without mtime and size
  std::map<std::string, Cached> cache_ids;  // key: path|offset|length|format|mtime|size
  int64_t next_id = 0;
  uint64_t tick = 0;
  uint64_t request_tick = 0;  // value of `tick` when the current request began: entries used since are in use by it
  std::string epoch;  // the data set the cached splits belong to (kDropCache)
  std::string stat_error;
  size_t max_batch = size_t(1024) << 20;
  int64_t staged_before = 0;
  hgmw::WorkerStats stats{};

  bool reply(int fd, uint32_t op, int status, const void *payload, size_t n) {
    hgmw::Header h{hgmw::kMagic, op, status, 0, n};
    return write_all(fd, &h, sizeof h) && (n == 0 || write_all(fd, payload, n));
  }
  bool reply_error(int fd, uint32_t op, int code, const std::string &msg) {
    fprintf(stderr, "hgm-worker: %s\n", msg.c_str());
    return reply(fd, op, code ? code : HGM_EINVAL, msg.data(), msg.size());
  }
  bool reply_gpu_error(int fd, uint32_t op, int code) { return reply_error(fd, op, code, hgm_last_error(gpu)); }

  bool reply_map(int fd, uint32_t op, int64_t n_points, int cache_hit, const std::vector<double> &sums,
                 const std::vector<int64_t> &counts) {
    hgmw::MapReply r{};
    r.n_points = n_points;
    r.cache_hit = cache_hit;
    hgm_last_map_ms(gpu, &r.map_ms, &r.launches);
    hgm_last_rechecks(gpu, &r.rechecks);
    int64_t staged = 0;
    hgm_staged_bytes(gpu, &staged);
    r.staged_bytes = staged - staged_before;
    staged_before = staged;
    std::vector<char> out(sizeof r + sums.size() * 8 + counts.size() * 8);
    memcpy(out.data(), &r, sizeof r);
    memcpy(out.data() + sizeof r, sums.data(), sums.size() * 8);
    memcpy(out.data() + sizeof r + sums.size() * 8, counts.data(), counts.size() * 8);
    stats.map_calls += 1;
    return reply(fd, op, 0, out.data(), out.size());
  }

  bool handle_set_centroids(int fd, const std::vector<char> &p) {
    if (p.size() < 8) return reply_error(fd, hgmw::kSetCentroids, HGM_EINVAL, "short SET_CENTROIDS");
    int32_t nk, nd;
    memcpy(&nk, p.data(), 4);
    memcpy(&nd, p.data() + 4, 4);
    if (nk <= 0 || nd <= 0 || p.size() != 8 + (size_t)nk * nd * 4)
      return reply_error(fd, hgmw::kSetCentroids, HGM_EINVAL, "SET_CENTROIDS payload does not match k*d");
    // every task of an iteration brings the same table: the index built for the first one serves the rest
    if (nk == k && nd == d && memcmp(centroids.data(), p.data() + 8, centroids.size() * 4) == 0)
      return reply(fd, hgmw::kSetCentroids, 0, nullptr, 0);
    if (nd != d) cache_ids.clear();  // the library drops cached rows of another d
    centroids.assign((size_t)nk * nd, 0.f);
    memcpy(centroids.data(), p.data() + 8, centroids.size() * 4);
    const int rc = hgm_set_centroids(gpu, centroids.data(), nk, nd);
    if (rc != HGM_OK) return reply_gpu_error(fd, hgmw::kSetCentroids, rc);
    k = nk;
    d = nd;
    return reply(fd, hgmw::kSetCentroids, 0, nullptr, 0);
  }

  // the batch is read from the socket straight into pinned memory, so hgm_map_split copies from it by DMA
  bool handle_map_bytes(int fd, uint64_t payload) {
    int32_t fmt = 0;
    if (payload < 4 || !read_all(fd, &fmt, 4)) return false;
    const size_t nbytes = (size_t)(payload - 4);
    auto skip_payload = [&]() {  // keep the stream in step, then report
      std::vector<char> sink(1 << 20);
      for (size_t left = nbytes; left;) {
        const size_t m = left < sink.size() ? left : sink.size();
        if (!read_all(fd, sink.data(), m)) return false;
        left -= m;
      }
      return true;
    };
    if (nbytes > max_batch) {
      if (!skip_payload()) return false;
      char msg[160];
      snprintf(msg, sizeof msg, "MAP_BYTES batch of %zu bytes is over the worker's --max-batch-mib %zu", nbytes, max_batch >> 20);
      return reply_error(fd, hgmw::kMapBytes, HGM_EINVAL, msg);
    }
    if (nbytes > pinned_cap) {
      if (pinned) hgm_host_free(gpu, pinned);
      pinned = nullptr;
      pinned_cap = 0;
      if (hgm_host_alloc(gpu, nbytes, &pinned) != HGM_OK) {
        if (!skip_payload()) return false;
        return reply_gpu_error(fd, hgmw::kMapBytes, HGM_ENOMEM);
      }
      pinned_cap = nbytes;
    }
    if (nbytes && !read_all(fd, pinned, nbytes)) return false;
    if (k == 0) return reply_error(fd, hgmw::kMapBytes, HGM_EINVAL, "MAP_BYTES before SET_CENTROIDS");
    std::vector<double> sums((size_t)k * d);
    std::vector<int64_t> counts(k);
    int64_t n = 0;
    const int rc = hgm_map_split(gpu, pinned, nbytes, fmt, sums.data(), counts.data(), nullptr, &n);
    if (rc != HGM_OK) return reply_gpu_error(fd, hgmw::kMapBytes, rc);
    return reply_map(fd, hgmw::kMapBytes, n, 0, sums, counts);
  }

  // The split in the HBM cache: its id, or -1 if it does not fit even after evictions (the caller streams it).
  // *hit = it was there already.  A non-zero *rc (with -1) is an error to report.
  int64_t ensure_cached(const std::string &path, int64_t off, int64_t len, int32_t fmt, int *hit, int *rc) {
    *rc = HGM_OK;
    *hit = 0;
    struct stat st{};
    if (stat(path.c_str(), &st) != 0) {
      stat_error = "cannot stat " + path + ": " + strerror(errno);
      *rc = HGM_EIO;
      return -1;
    }
    char key[64], version[64];
    snprintf(key, sizeof key, "|%lld|%lld|%d", (long long)off, (long long)len, fmt);
    snprintf(version, sizeof version, "|%lld|%lld", (long long)st.st_mtime, (long long)st.st_size);
    const std::string split = path + key, id_key = split + version;
    auto it = cache_ids.find(id_key);
    if (it != cache_ids.end()) {
      *hit = 1;
      it->second.last_use = ++tick;
      return it->second.id;
    }
    const int64_t id = next_id++;
    // other versions of this split are dead weight: the file has changed under them
    for (auto v = cache_ids.begin(); v != cache_ids.end();) {
      if (v->second.split == split) { hgm_cache_drop(gpu, v->second.id); v = cache_ids.erase(v); }
      else ++v;
    }
    int put = hgm_cache_put_file(gpu, id, path.c_str(), off, len, fmt);
    // HBM is full: make room, least recently used first (arena blocks are returned to the driver as they
    // empty, so a few drops may be needed before a block frees up)
    while (put == HGM_ENOMEM) {
      // never a split the request being served has already been promised (MAP_FILES maps them all at its end)
      auto lru = cache_ids.end();
      for (auto v = cache_ids.begin(); v != cache_ids.end(); ++v)
        if (v->second.last_use <= request_tick && (lru == cache_ids.end() || v->second.last_use < lru->second.last_use)) lru = v;
      if (lru == cache_ids.end()) break;  // everything left belongs to this request: stream the split instead
      hgm_cache_drop(gpu, lru->second.id);
      cache_ids.erase(lru);
      stats.cache_evictions += 1;
      put = hgm_cache_put_file(gpu, id, path.c_str(), off, len, fmt);
    }
    if (put == HGM_OK) {
      cache_ids[id_key] = Cached{id, ++tick, split};
      return id;
    }
    if (put != HGM_ENOMEM) *rc = put;  // did not fit: the cache is only a cache
    return -1;
  }

  bool handle_map_file(int fd, const std::vector<char> &p) {
    request_tick = tick;
    if (p.size() < 24) return reply_error(fd, hgmw::kMapFile, HGM_EINVAL, "short MAP_FILE");
    int32_t fmt, use_cache;
    int64_t off, len;
    memcpy(&fmt, p.data(), 4);
    memcpy(&use_cache, p.data() + 4, 4);
    memcpy(&off, p.data() + 8, 8);
    memcpy(&len, p.data() + 16, 8);
    const std::string path(p.data() + 24, p.size() - 24);
    if (k == 0) return reply_error(fd, hgmw::kMapFile, HGM_EINVAL, "MAP_FILE before SET_CENTROIDS");
    std::vector<double> sums((size_t)k * d);
    std::vector<int64_t> counts(k);
    int64_t n = 0;
    int rc;
    if (use_cache) {
      int hit = 0;
      const int64_t id = ensure_cached(path, off, len, fmt, &hit, &rc);
      if (id < 0 && rc == HGM_EIO) return reply_error(fd, hgmw::kMapFile, HGM_EIO, stat_error);
      if (id < 0 && rc != HGM_OK) return reply_gpu_error(fd, hgmw::kMapFile, rc);
      if (id >= 0) {
        rc = hgm_reset_partials(gpu);
        if (rc == HGM_OK) rc = hgm_cache_map(gpu, id, nullptr, &n);
        if (rc == HGM_OK) rc = hgm_read_partials(gpu, sums.data(), counts.data());
        if (rc != HGM_OK) return reply_gpu_error(fd, hgmw::kMapFile, rc);
        (hit ? stats.cache_hits : stats.cache_misses) += 1;
        return reply_map(fd, hgmw::kMapFile, n, hit, sums, counts);
      }
      // did not fit: stream the split
    }
    rc = hgm_map_file(gpu, path.c_str(), off, len, fmt, sums.data(), counts.data(), nullptr, &n);
    if (rc != HGM_OK) return reply_gpu_error(fd, hgmw::kMapFile, rc);
    stats.cache_misses += use_cache ? 1 : 0;
    return reply_map(fd, hgmw::kMapFile, n, 0, sums, counts);
  }

  // A task's whole batch of file splits in one request: the cached ones are mapped back to back into the device
  // partials and read once; a per-split round trip (request, launch, synchronise, 3 KB reply: ~70 us) was most of a
  // batched task's time once its splits were resident.
  bool handle_map_files(int fd, const std::vector<char> &p) {
    request_tick = tick;
    if (p.size() < 12) return reply_error(fd, hgmw::kMapFiles, HGM_EINVAL, "short MAP_FILES");
    int32_t fmt, use_cache, count;
    memcpy(&fmt, p.data(), 4);
    memcpy(&use_cache, p.data() + 4, 4);
    memcpy(&count, p.data() + 8, 4);
    if (k == 0) return reply_error(fd, hgmw::kMapFiles, HGM_EINVAL, "MAP_FILES before SET_CENTROIDS");
    std::vector<double> sums((size_t)k * d, 0.0), part((size_t)k * d);
    std::vector<int64_t> counts(k, 0), part_counts(k);
    std::vector<int64_t> resident;
    int64_t n_total = 0;
    int hits = 0;
    size_t at = 12;
    for (int32_t i = 0; i < count; ++i) {
      if (at + 20 > p.size()) return reply_error(fd, hgmw::kMapFiles, HGM_EINVAL, "truncated MAP_FILES entry");
      int64_t off, len;
      int32_t plen;
      memcpy(&off, p.data() + at, 8);
      memcpy(&len, p.data() + at + 8, 8);
      memcpy(&plen, p.data() + at + 16, 4);
      at += 20;
      if (plen < 0 || at + (size_t)plen > p.size()) return reply_error(fd, hgmw::kMapFiles, HGM_EINVAL, "truncated MAP_FILES path");
      const std::string path(p.data() + at, (size_t)plen);
      at += (size_t)plen;
      int rc = HGM_OK, hit = 0;
      const int64_t id = use_cache ? ensure_cached(path, off, len, fmt, &hit, &rc) : -1;
      if (id < 0 && rc == HGM_EIO) return reply_error(fd, hgmw::kMapFiles, HGM_EIO, stat_error);
      if (id < 0 && rc != HGM_OK) return reply_gpu_error(fd, hgmw::kMapFiles, rc);
      if (id >= 0) {
        resident.push_back(id);
        hits += hit;
        (hit ? stats.cache_hits : stats.cache_misses) += 1;
        continue;
      }
      int64_t n = 0;  // not cacheable: streamed on its own, folded on the host
      rc = hgm_map_file(gpu, path.c_str(), off, len, fmt, part.data(), part_counts.data(), nullptr, &n);
      if (rc != HGM_OK) return reply_gpu_error(fd, hgmw::kMapFiles, rc);
      stats.cache_misses += use_cache ? 1 : 0;
      for (size_t j = 0; j < sums.size(); ++j) sums[j] += part[j];
      for (size_t j = 0; j < counts.size(); ++j) counts[j] += part_counts[j];
      n_total += n;
    }
    if (!resident.empty()) {
      int rc = hgm_reset_partials(gpu);
      int64_t n = 0;
      // neighbours in the cache's arena (splits staged one after the other, as a task's splits are) share a launch
      if (rc == HGM_OK) rc = hgm_cache_map_many(gpu, resident.data(), (int32_t)resident.size(), &n, nullptr);
      n_total += n;
      if (rc == HGM_OK) rc = hgm_read_partials(gpu, part.data(), part_counts.data());
      if (rc != HGM_OK) return reply_gpu_error(fd, hgmw::kMapFiles, rc);
      for (size_t j = 0; j < sums.size(); ++j) sums[j] += part[j];
      for (size_t j = 0; j < counts.size(); ++j) counts[j] += part_counts[j];
    }
    return reply_map(fd, hgmw::kMapFiles, n_total, hits, sums, counts);
  }

  // returns false when the worker should exit
  bool serve(int fd) {
    stats.tasks += 1;
    for (;;) {
      hgmw::Header h;
      if (!read_all(fd, &h, sizeof h)) return true;  // the task is done (or died): next client
      if (h.magic != hgmw::kMagic) { fprintf(stderr, "hgm-worker: bad magic, dropping the client\n"); return true; }
      if (h.op == hgmw::kMapBytes) {
        if (!handle_map_bytes(fd, h.payload)) return true;
        continue;
      }
      if (h.payload > (uint64_t(1) << 30)) { fprintf(stderr, "hgm-worker: oversized control message\n"); return true; }
      std::vector<char> p((size_t)h.payload);
      if (h.payload && !read_all(fd, p.data(), p.size())) return true;
      bool ok = true;
      switch (h.op) {
        case hgmw::kHello: {
          struct { uint32_t version; int32_t device; } out{hgmw::kVersion, device};
          ok = reply(fd, h.op, 0, &out, sizeof out);
          break;
        }
        case hgmw::kSetCentroids: ok = handle_set_centroids(fd, p); break;
        case hgmw::kMapFile: ok = handle_map_file(fd, p); break;
        case hgmw::kMapFiles: ok = handle_map_files(fd, p); break;
        case hgmw::kStats: {
          hgm_cache_info(gpu, &stats.cached_splits, &stats.cached_bytes);
          hgm_launch_count(gpu, &stats.kernel_launches);
          ok = reply(fd, h.op, 0, &stats, sizeof stats);
          break;
        }
        case hgmw::kDropCache: {
          const std::string asked(p.begin(), p.end());
          if (asked.empty() || asked != epoch) {
            for (auto &v : cache_ids) hgm_cache_drop(gpu, v.second.id);
            cache_ids.clear();
            epoch = asked;
          }
          ok = reply(fd, h.op, 0, nullptr, 0);
          break;
        }
        case hgmw::kShutdown:
          reply(fd, h.op, 0, nullptr, 0);
          return false;
        default: ok = reply_error(fd, h.op, HGM_EINVAL, "unknown request");
      }
      if (!ok) return true;
    }
  }
};

}  // namespace

int main(int argc, char **argv) {
  int device = 0;
  long max_batch_mib = 1024;
  std::string path;
  for (int i = 1; i + 1 < argc; i += 2) {
    if (!strcmp(argv[i], "--device")) device = atoi(argv[i + 1]);
    else if (!strcmp(argv[i], "--socket")) path = argv[i + 1];
    else if (!strcmp(argv[i], "--max-batch-mib")) max_batch_mib = atol(argv[i + 1]);
    else { fprintf(stderr, "usage: hgm-worker --device N --socket PATH [--max-batch-mib M]\n"); return 2; }
  }
  if (max_batch_mib < 1) { fprintf(stderr, "hgm-worker: --max-batch-mib must be at least 1\n"); return 2; }
  if (path.empty() || path.size() >= sizeof(sockaddr_un{}.sun_path)) {
    fprintf(stderr, "usage: hgm-worker --device N --socket PATH (PATH shorter than %zu bytes)\n", sizeof(sockaddr_un{}.sun_path));
    return 2;
  }
  signal(SIGPIPE, SIG_IGN);
  Worker w;
  w.device = device;
  w.max_batch = (size_t)max_batch_mib << 20;
  if (hgm_open(device, &w.gpu) != HGM_OK) {  // no device, no worker: tasks that need it fail loudly
    fprintf(stderr, "hgm-worker: hgm_open(%d): %s\n", device, hgm_last_error(nullptr));
    return 1;
  }
  const int ls = socket(AF_UNIX, SOCK_STREAM, 0);
  sockaddr_un addr{};
  addr.sun_family = AF_UNIX;
  snprintf(addr.sun_path, sizeof addr.sun_path, "%s", path.c_str());
  unlink(path.c_str());
  if (ls < 0 || bind(ls, reinterpret_cast<sockaddr *>(&addr), sizeof addr) != 0 || listen(ls, 8) != 0) {
    fprintf(stderr, "hgm-worker: cannot listen on %s: %s\n", path.c_str(), strerror(errno));
    return 1;
  }
  chmod(path.c_str(), 0600);  // the slot belongs to the user the TaskTracker runs tasks as
  fprintf(stderr, "hgm-worker: device %d ready on %s\n", device, path.c_str());
  for (bool run = true; run;) {
    const int fd = accept(ls, nullptr, nullptr);
    if (fd < 0) { if (errno == EINTR) continue; break; }
    run = w.serve(fd);
    close(fd);
  }
  close(ls);
  unlink(path.c_str());
  if (w.pinned) hgm_host_free(w.gpu, w.pinned);
  hgm_close(w.gpu);
  return 0;
}
