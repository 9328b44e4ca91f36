// kmeans_pipes_cpu.cc -- the CPU K-means Pipes program (the "-cpubin").
// TEST INFRASTRUCTURE ONLY: it is the oracle and the timed CPU baseline; the
// product under hadoop-1.0.3-gpu_b200/ never links or executes it.
//
// PARITY UNPINNED for the arithmetic (the reference ships no K-means program;
// see kmeans_oracle.c).  What is pinned is everything AROUND the arithmetic:
// this file is written against the reference's own API header
//   /root/reference/hadoop-1.0.3/src/c++/pipes/api/hadoop/Pipes.hh:149-256
// in the shape of the reference's shipped example
//   src/examples/pipes/impl/wordcount-part.cc:27-75  (mapper + combiner +
//   partitioner through TemplateFactory, counters via getCounter), and
//   src/examples/pipes/impl/sort.cc:35-44 (reading JobConf keys),
// and it is linked with the reference's own runtime (HadoopPipes.cc,
// SerialUtils.cc, StringUtils.cc compiled from where they lie, see Makefile).
// So the record pump (HadoopPipes.cc:865-890), CombineRunner buffering and
// key order (:660-706), vint framing and the DONE handshake are the
// reference's, byte for byte.
//
// Record contract (this project's):
//   JobConf   kmeans.k, kmeans.d, kmeans.centroids (local path of k*d
//             little-endian float32), kmeans.input.format = binary|text|path
//   map in    value = one record: binary -> m*d LE float32 (m >= 1 points);
//             text -> decimals separated by " ,\t" (toFloat semantics);
//             path -> "file[\toffset\tlength]" naming binary rows to read
//   map out   key = decimal cluster id, value = "1 x0 x1 ..." (%.17g)
//   combine   key = cluster id, value = "count s0 s1 ..." (count decimal
//             int64, sums %.17g FP64) -- same type as map out, so the
//             combiner may run zero or more times
//   partition cluster id mod numReduces
//   reduce    key = cluster id, value = "count c0 c1 ..." new centroid
//             coordinates (%.9g of FP32(sum/count))
#include "hadoop/Pipes.hh"
#include "hadoop/TemplateFactory.hh"
#include "hadoop/StringUtils.hh"
#include "hadoop/SerialUtils.hh"

#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>
#include <string>
#include <vector>

extern "C" int32_t hgo_nearest(const float *x, const float *cent, int32_t k, int32_t d);
extern "C" int32_t hgo_decode_text_record(const char *rec, int64_t len, float *out, int32_t cap);

static const std::string KMEANS = "KMEANS";

// Wall-clock of the map work itself, written to $HGM_ORACLE_STATS on exit:
// the reference's runTask joins a ping thread that sleeps 5 s
// (HadoopPipes.cc:1049,1147-1151), which would otherwise swamp short runs.
static struct Stats {
  double t_first_map, t_last_emit;
  uint64_t points;
  Stats() : t_first_map(0), t_last_emit(0), points(0) {}
  static double now() {
    struct timeval tv; gettimeofday(&tv, NULL);
    return tv.tv_sec + 1e-6 * tv.tv_usec;
  }
  ~Stats() {
    const char *path = getenv("HGM_ORACLE_STATS");
    if (!path) return;
    FILE *f = fopen(path, "w");
    if (!f) return;
    fprintf(f, "{\"points\": %llu, \"map_seconds\": %.6f}\n", (unsigned long long)points,
            t_last_emit - t_first_map);
    fclose(f);
  }
} g_stats;

static std::vector<float> loadCentroids(const HadoopPipes::JobConf *conf, int k, int d) {
  const std::string &path = conf->get("kmeans.centroids");
  FILE *f = fopen(path.c_str(), "rb");
  HADOOP_ASSERT(f != NULL, "cannot open centroid file " + path);
  std::vector<float> c((size_t)k * d);
  size_t got = fread(c.data(), sizeof(float), c.size(), f);
  fclose(f);
  HADOOP_ASSERT(got == c.size(), "centroid file " + path + " is shorter than k*d floats");
  return c;
}

static std::string partialToString(long long count, const double *sums, int d) {
  char buf[64];
  snprintf(buf, sizeof buf, "%lld", count);
  std::string out(buf);
  for (int j = 0; j < d; ++j) {
    snprintf(buf, sizeof buf, " %.17g", sums[j]);
    out += buf;
  }
  return out;
}

class KMeansMap : public HadoopPipes::Mapper {
  int k, d;
  enum { BINARY, TEXT, PATH } format;
  std::vector<float> centroids;
  std::vector<float> row;
  HadoopPipes::TaskContext::Counter *pointsIn;

  void mapPoint(HadoopPipes::MapContext &context, const float *x) {
    int32_t c = hgo_nearest(x, centroids.data(), k, d);
    double coords[64];
    std::vector<double> big;
    double *p = coords;
    if (d > 64) { big.resize(d); p = big.data(); }
    for (int j = 0; j < d; ++j) p[j] = (double)x[j];
    context.emit(HadoopUtils::toString(c), partialToString(1, p, d));
    g_stats.points += 1;
  }

public:
  KMeansMap(HadoopPipes::TaskContext &context) {
    const HadoopPipes::JobConf *conf = context.getJobConf();
    k = conf->getInt("kmeans.k");
    d = conf->getInt("kmeans.d");
    HADOOP_ASSERT(k > 0 && d > 0, "kmeans.k and kmeans.d must be positive");
    format = BINARY;
    if (conf->hasKey("kmeans.input.format")) {
      const std::string &f = conf->get("kmeans.input.format");
      if (f == "text") format = TEXT;
      else if (f == "path") format = PATH;
      else HADOOP_ASSERT(f == "binary", "unknown kmeans.input.format " + f);
    }
    centroids = loadCentroids(conf, k, d);
    row.resize(d);
    pointsIn = context.getCounter(KMEANS, "POINTS");
  }

  void mapBinary(HadoopPipes::MapContext &context, const char *bytes, size_t len) {
    HADOOP_ASSERT(len % (4 * (size_t)d) == 0, "binary record is not a whole number of rows");
    size_t m = len / (4 * (size_t)d);
    for (size_t i = 0; i < m; ++i) {
      memcpy(row.data(), bytes + i * 4 * d, 4 * (size_t)d);  // rows are little-endian float32
      mapPoint(context, row.data());
    }
    context.incrementCounter(pointsIn, m);
  }

  void map(HadoopPipes::MapContext &context) {
    if (g_stats.t_first_map == 0) g_stats.t_first_map = Stats::now();
    const std::string &v = context.getInputValue();
    if (format == BINARY) {
      mapBinary(context, v.data(), v.size());
    } else if (format == TEXT) {
      std::vector<float> vals(v.size() / 2 + 1);
      int n = hgo_decode_text_record(v.data(), (int64_t)v.size(), vals.data(), (int32_t)vals.size());
      HADOOP_ASSERT(n >= 0 && n % d == 0, "text record does not hold a whole number of points: " + v);
      for (int i = 0; i < n / d; ++i) mapPoint(context, &vals[(size_t)i * d]);
      context.incrementCounter(pointsIn, n / d);
    } else {
      std::vector<std::string> parts = HadoopUtils::splitString(v, "\t");
      HADOOP_ASSERT(parts.size() == 1 || parts.size() == 3, "path record must be file[\\toffset\\tlength]: " + v);
      FILE *f = fopen(parts[0].c_str(), "rb");
      HADOOP_ASSERT(f != NULL, "cannot open " + parts[0]);
      long long off = 0, len = -1;
      if (parts.size() == 3) { off = atoll(parts[1].c_str()); len = atoll(parts[2].c_str()); }
      if (len < 0) { fseek(f, 0, SEEK_END); len = ftell(f) - off; }
      fseek(f, off, SEEK_SET);
      std::vector<char> buf(1 << 20);
      size_t rowBytes = 4 * (size_t)d, chunk = (buf.size() / rowBytes) * rowBytes;
      while (len > 0) {
        size_t want = (size_t)len < chunk ? (size_t)len : chunk;
        size_t got = fread(buf.data(), 1, want, f);
        HADOOP_ASSERT(got == want, "short read on " + parts[0]);
        mapBinary(context, buf.data(), got);
        len -= (long long)got;
      }
      fclose(f);
    }
  }
};

// Combiner and (first half of) reducer: add partial records "count s0 s1 ...".
class KMeansCombine : public HadoopPipes::Reducer {
protected:
  int d;
  std::vector<double> sums;
  long long count;

  void accumulate(HadoopPipes::ReduceContext &context) {
    count = 0;
    sums.assign(d, 0.0);
    while (context.nextValue()) {
      const char *s = context.getInputValue().c_str();
      char *end;
      count += strtoll(s, &end, 10);
      for (int j = 0; j < d; ++j) {
        s = end;
        sums[j] += strtod(s, &end);
        HADOOP_ASSERT(end != s, "malformed partial record " + context.getInputValue());
      }
    }
  }

public:
  KMeansCombine(HadoopPipes::TaskContext &context) {
    d = context.getJobConf()->getInt("kmeans.d");
    count = 0;
  }
  void reduce(HadoopPipes::ReduceContext &context) {
    accumulate(context);
    context.emit(context.getInputKey(), partialToString(count, sums.data(), d));
    g_stats.t_last_emit = Stats::now();
  }
};

// Reducer: centroid update, new centre = FP32(sum / count).
class KMeansReduce : public KMeansCombine {
public:
  KMeansReduce(HadoopPipes::TaskContext &context) : KMeansCombine(context) {}
  void reduce(HadoopPipes::ReduceContext &context) {
    accumulate(context);
    char buf[64];
    snprintf(buf, sizeof buf, "%lld", count);
    std::string out(buf);
    for (int j = 0; j < d; ++j) {
      float c = (float)(sums[j] / (double)count);
      snprintf(buf, sizeof buf, " %.9g", c);
      out += buf;
    }
    context.emit(context.getInputKey(), out);
  }
};

class KMeansPartitioner : public HadoopPipes::Partitioner {
public:
  KMeansPartitioner(HadoopPipes::TaskContext &) {}
  int partition(const std::string &key, int numOfReduces) {
    return HadoopUtils::toInt(key) % numOfReduces;
  }
};

int main(int, char **) {
  // As in the shipped examples (wordcount-simple.cc:63-66) the bool from runTask
  // is the exit status, so 1 means success; Hadoop waits for DONE instead.
  return HadoopPipes::runTask(
      HadoopPipes::TemplateFactory<KMeansMap, KMeansReduce, KMeansPartitioner, KMeansCombine>());
}
