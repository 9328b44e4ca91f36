/* kmeans_oracle.c -- CPU oracle for the K-means map path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this file's shared object.  Nothing under
 * hadoop-1.0.3-gpu_b200/ links, imports or calls it.
 *
 * PARITY UNPINNED: /root/reference ships no K-means mapper, combiner, golden
 * vector or fixture; the historical
 * mapper lives in koichi626/hadoop-gpu, which is not vendored and has no pinned
 * version.  What the reference does pin is the slot this code fills and the
 * wire encodings around it, and those are what is restated and cited here:
 *
 *   - the map body slot   HadoopPipes::Mapper::map(MapContext&)
 *       src/c++/pipes/api/hadoop/Pipes.hh:158-161, invoked from
 *       src/c++/pipes/impl/HadoopPipes.cc:884-885
 *   - the combiner slot   HadoopPipes::Reducer::reduce(ReduceContext&)
 *       Pipes.hh:166-169, driven by CombineRunner::spillAll HadoopPipes.cc:697-705
 *   - Hadoop vint/vlong   src/c++/utils/impl/SerialUtils.cc:182-244
 *   - text -> float       src/c++/utils/impl/StringUtils.cc:49-56 (sscanf "%f%c")
 *   - XDR float           SerialUtils.cc:246-262 (big-endian IEEE-754)
 *
 * The arithmetic contract is this project's own:
 *   distance  = FP32 sum_j (x_j - c_j)^2, j ascending, every operation rounded
 *               separately (no FMA contraction: build with -ffp-contract=off);
 *   argmin    = strict '<' scan over ascending centroid index starting from
 *               +inf / index 0, so the lowest index wins ties and a NaN or
 *               +inf distance never wins;
 *   combiner  = int64 counts (exact) and FP64 sums of the FP32 coordinates.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---- a1: nearest centroid (the map body) -------------------------------- */

int32_t hgo_nearest(const float *x, const float *cent, int32_t k, int32_t d) {
  float best = INFINITY;
  int32_t arg = 0;
  for (int32_t c = 0; c < k; ++c) {
    const float *cc = cent + (size_t)c * d;
    float acc = 0.0f;
    for (int32_t j = 0; j < d; ++j) {
      float diff = x[j] - cc[j];
      float sq = diff * diff;
      acc = acc + sq;
    }
    if (acc < best) { best = acc; arg = c; }
  }
  return arg;
}

void hgo_assign(const float *pts, int64_t n, int32_t d, const float *cent, int32_t k,
                int32_t *assign) {
  for (int64_t i = 0; i < n; ++i) assign[i] = hgo_nearest(pts + (size_t)i * d, cent, k, d);
}

/* ---- a1+a2: map + combine, as one map task would leave it --------------- *
 * sums[k*d] FP64, counts[k] int64, abs_sums[k*d] (sum |x|, for the tolerance
 * floor; may be NULL), assign[n] (may be NULL).  Outputs are overwritten. */
void hgo_map_combine(const float *pts, int64_t n, int32_t d, const float *cent, int32_t k,
                     double *sums, int64_t *counts, double *abs_sums, int32_t *assign) {
  memset(sums, 0, sizeof(double) * (size_t)k * d);
  memset(counts, 0, sizeof(int64_t) * (size_t)k);
  if (abs_sums) memset(abs_sums, 0, sizeof(double) * (size_t)k * d);
  for (int64_t i = 0; i < n; ++i) {
    const float *x = pts + (size_t)i * d;
    int32_t a = hgo_nearest(x, cent, k, d);
    if (assign) assign[i] = a;
    counts[a] += 1;
    for (int32_t j = 0; j < d; ++j) {
      sums[(size_t)a * d + j] += (double)x[j];
      if (abs_sums) abs_sums[(size_t)a * d + j] += fabs((double)x[j]);
    }
  }
}

/* Accumulate-only variant: given assignments, rebuild sums/counts (used to
 * cross-check a device assignment array independently of its device sums). */
void hgo_combine_assigned(const float *pts, int64_t n, int32_t d, const int32_t *assign,
                          int32_t k, double *sums, int64_t *counts) {
  memset(sums, 0, sizeof(double) * (size_t)k * d);
  memset(counts, 0, sizeof(int64_t) * (size_t)k);
  for (int64_t i = 0; i < n; ++i) {
    int32_t a = assign[i];
    counts[a] += 1;
    for (int32_t j = 0; j < d; ++j) sums[(size_t)a * d + j] += (double)pts[(size_t)i * d + j];
  }
}

/* ---- a6/a5: Hadoop vint (SerialUtils.cc:182-244) ------------------------- */

int32_t hgo_vint_encode(int64_t t, uint8_t *out) {
  if (t >= -112 && t <= 127) { out[0] = (uint8_t)(int8_t)t; return 1; }
  int8_t len = -112;
  if (t < 0) { t ^= -1ll; len = -120; }
  uint64_t tmp = (uint64_t)t;
  while (tmp != 0) { tmp >>= 8; len--; }
  out[0] = (uint8_t)len;
  int nbytes = (len < -120) ? -(len + 120) : -(len + 112);
  for (int idx = nbytes; idx != 0; idx--) {
    int shift = (idx - 1) * 8;
    out[1 + (nbytes - idx)] = (uint8_t)(((uint64_t)t >> shift) & 0xFF);
  }
  return 1 + nbytes;
}

/* returns bytes consumed, or -1 if the buffer is too short */
int32_t hgo_vint_decode(const uint8_t *buf, int64_t avail, int64_t *value) {
  if (avail < 1) return -1;
  int8_t b = (int8_t)buf[0];
  if (b >= -112) { *value = b; return 1; }
  int negative = b < -120;
  int len = negative ? (-120 - b) : (-112 - b);
  if (avail < 1 + len) return -1;
  int64_t t = 0;
  for (int i = 0; i < len; ++i) t = (int64_t)(((uint64_t)t << 8) | buf[1 + i]);
  if (negative) t ^= -1ll;
  *value = t;
  return 1 + len;
}

/* ---- a7: text field -> float (StringUtils.cc:49-56) ----------------------- */

int32_t hgo_to_float(const char *s, float *out) {
  char trash;
  return sscanf(s, "%f%c", out, &trash) == 1 ? 0 : -1;
}

/* Decode one text record: fields separated by any of " ,\t" (splitString
 * semantics, StringUtils.cc:80-95: empty fields are skipped). Returns the
 * number of floats written, or -1 on a malformed field / overflow of cap. */
int32_t hgo_decode_text_record(const char *rec, int64_t len, float *out, int32_t cap) {
  int32_t n = 0;
  int64_t i = 0;
  char field[128];
  while (i < len) {
    while (i < len && (rec[i] == ' ' || rec[i] == ',' || rec[i] == '\t')) ++i;
    int64_t s = i;
    while (i < len && !(rec[i] == ' ' || rec[i] == ',' || rec[i] == '\t')) ++i;
    if (i > s) {
      if (i - s >= (int64_t)sizeof(field) || n >= cap) return -1;
      memcpy(field, rec + s, (size_t)(i - s));
      field[i - s] = 0;
      if (hgo_to_float(field, &out[n]) != 0) return -1;
      ++n;
    }
  }
  return n;
}

/* a7 in bulk: field i is buf[off[i] .. off[i+1]) (at most 127 bytes); rc[i] = 0 or -1 */
void hgo_to_float_many(const char *buf, const int64_t *off, int64_t n, float *out, int32_t *rc) {
  char field[128];
  for (int64_t i = 0; i < n; ++i) {
    int64_t len = off[i + 1] - off[i];
    if (len < 0 || len >= (int64_t)sizeof(field)) { rc[i] = -1; continue; }
    memcpy(field, buf + off[i], (size_t)len);
    field[len] = 0;
    out[i] = 0.0f;
    rc[i] = hgo_to_float(field, &out[i]);
  }
}

/* ---- section 8(f)3: TextInputFormat framing --------------------------------
 * LineReader.readLine (core/org/apache/hadoop/util/LineReader.java:106-167) over
 * a file held in memory: returns the bytes consumed from pos (content plus the
 * terminator LF, CR or CR LF; a CR is only joined with an LF right behind it),
 * *content = bytes before the terminator. 0 at end of file. */
int64_t hgo_read_line(const char *file, int64_t size, int64_t pos, int64_t *content) {
  int64_t i = pos;
  while (i < size && file[i] != '\n' && file[i] != '\r') ++i;
  *content = i - pos;
  if (i >= size) return i - pos;
  if (file[i] == '\n') return i - pos + 1;
  return (i + 1 < size && file[i + 1] == '\n') ? i - pos + 2 : i - pos + 1;
}

/* LineRecordReader for FileSplit(off, len)
 * (mapred/org/apache/hadoop/mapred/LineRecordReader.java:79-95 constructor,
 * :137-157 next): a split that does not start at byte 0 backs up one byte and
 * throws away one line; records are the lines that START before off+len.
 * Every record is decoded as a K-means text record (hgo_decode_text_record) of
 * whole d-dimensional points, as the CPU mapper does. Returns the number of
 * floats written, -1 - (record index) if that record is malformed or ragged,
 * INT64_MIN if out is too small. *n_records = records seen. */
int64_t hgo_text_split_decode(const char *file, int64_t size, int64_t off, int64_t len, int32_t d,
                              float *out, int64_t cap, int64_t *n_records) {
  int64_t start = off, end = off + len, content = 0, total = 0, rec = 0;
  if (start != 0) {
    --start;
    start += hgo_read_line(file, size, start, &content);
  }
  int64_t pos = start;
  while (pos < end) {
    int64_t used = hgo_read_line(file, size, pos, &content);
    if (used == 0) break;
    if (cap - total < content) { /* a record of c bytes holds fewer than c numbers */
      if (cap - total < (content + 1) / 2) return INT64_MIN;
    }
    int64_t room = cap - total > 0x7fffffff ? 0x7fffffff : cap - total;
    int32_t n = hgo_decode_text_record(file + pos, content, out + total, (int32_t)room);
    if (n < 0 || n % d != 0) { if (n_records) *n_records = rec; return -1 - rec; }
    total += n;
    pos += used;
    ++rec;
  }
  if (n_records) *n_records = rec;
  return total;
}

/* ---- section 8(f)3: SequenceFile framing -------------------------------------
 * An uncompressed version-6 SequenceFile of BytesWritable values held in memory,
 * read the way SequenceFileRecordReader reads FileSplit(off, len)
 * (mapred/org/apache/hadoop/mapred/SequenceFileRecordReader.java:40-52 constructor,
 * :71-84 next) on top of SequenceFile.Reader (core/org/apache/hadoop/io/
 * SequenceFile.java: init :1506-1571, readRecordLength :1949-1969, sync :2221-2246).
 * No SequenceFile fixture exists in the reference tree, so this restatement is
 * pinned by the source only; tests/seqfile.py writes files the way
 * SequenceFile.Writer does (:985-990 sync, :1025-1030 checkAndWriteSync, append). */

static int64_t hgo_be32(const unsigned char *p) {
  return (int32_t)(((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]);
}

/* Text.readString: vint length + bytes. Returns bytes consumed or -1. */
static int64_t hgo_skip_text(const unsigned char *p, int64_t avail, char *copy, int64_t copy_cap) {
  int64_t len = 0;
  int32_t used = hgo_vint_decode(p, avail, &len);
  if (used <= 0 || len < 0 || used + len > avail) return -1;
  if (copy) {
    int64_t m = len < copy_cap - 1 ? len : copy_cap - 1;
    memcpy(copy, p + used, (size_t)m);
    copy[m] = 0;
  }
  return used + len;
}

/* Parses the header; returns its length (position of the first record) or -1 if the
 * file is not an uncompressed v6 SequenceFile of BytesWritable values. sync16 <- marker. */
int64_t hgo_seq_header(const unsigned char *file, int64_t size, unsigned char *sync16) {
  char cls[128];
  int64_t pos = 4, n;
  if (size < 4 || file[0] != 'S' || file[1] != 'E' || file[2] != 'Q' || file[3] != 6) return -1;
  if ((n = hgo_skip_text(file + pos, size - pos, NULL, 0)) < 0) return -1;       /* key class */
  pos += n;
  if ((n = hgo_skip_text(file + pos, size - pos, cls, sizeof cls)) < 0) return -1; /* value class */
  pos += n;
  if (strcmp(cls, "org.apache.hadoop.io.BytesWritable") != 0) return -1;
  if (pos + 2 > size || file[pos] != 0 || file[pos + 1] != 0) return -1;          /* compressed / block-compressed */
  pos += 2;
  if (pos + 4 > size) return -1;
  int64_t meta = hgo_be32(file + pos);
  pos += 4;
  for (int64_t i = 0; i < 2 * meta; ++i) {
    if ((n = hgo_skip_text(file + pos, size - pos, NULL, 0)) < 0) return -1;
    pos += n;
  }
  if (pos + 16 > size) return -1;
  memcpy(sync16, file + pos, 16);
  return pos + 16;
}

/* SequenceFile.Reader.sync(position): the position of the first sync entry (escape + marker)
 * that starts at or after position; the end of the file if there is none. */
static int64_t hgo_seq_sync(const unsigned char *file, int64_t size, const unsigned char *sync16, int64_t position) {
  if (position + 20 >= size) return size;
  for (int64_t h = position + 4; h + 16 <= size; ++h)
    if (memcmp(file + h, sync16, 16) == 0) return h - 4;
  return size;
}

/* Decodes the BytesWritable values of FileSplit(off, len) as rows of d float32 (big-endian
 * if be != 0). Returns the number of floats, -1 - (record index) on a malformed record,
 * INT64_MIN if the header is not acceptable or out is too small. */
int64_t hgo_seq_split_decode(const unsigned char *file, int64_t size, int64_t off, int64_t len, int32_t d,
                             int32_t be, float *out, int64_t cap, int64_t *n_records) {
  unsigned char sync16[16];
  int64_t pos = hgo_seq_header(file, size, sync16);
  int64_t end = off + len, total = 0, rec = 0;
  if (pos < 0) return INT64_MIN;
  if (off > pos) pos = hgo_seq_sync(file, size, sync16, off);
  int more = pos < end;
  while (more) {
    int64_t before = pos;
    int sync_seen = 0;
    if (pos >= size) break;                       /* readRecordLength: -1 at end of file */
    if (pos + 4 > size) return -1 - rec;
    int64_t length = hgo_be32(file + pos);
    pos += 4;
    if (length == -1) {
      if (pos + 16 > size || memcmp(file + pos, sync16, 16) != 0) return -1 - rec; /* "File is corrupt!" */
      pos += 16;
      sync_seen = 1;
      if (pos >= size) break;
      if (pos + 4 > size) return -1 - rec;
      length = hgo_be32(file + pos);
      pos += 4;
    }
    if (before >= end && sync_seen) break;        /* the record belongs to the next split */
    if (pos + 4 > size) return -1 - rec;
    int64_t key_len = hgo_be32(file + pos);
    pos += 4;
    if (length < 0 || key_len < 0 || key_len > length || pos + length > size) return -1 - rec;
    const unsigned char *val = file + pos + key_len;
    int64_t val_len = length - key_len;
    pos += length;
    if (val_len < 4) return -1 - rec;
    int64_t bytes = hgo_be32(val);
    if (bytes != val_len - 4 || bytes % (4 * (int64_t)d) != 0) return -1 - rec;
    if (total + bytes / 4 > cap) return INT64_MIN;
    for (int64_t i = 0; i < bytes / 4; ++i) {
      const unsigned char *q = val + 4 + 4 * i;
      uint32_t u = be ? (((uint32_t)q[0] << 24) | ((uint32_t)q[1] << 16) | ((uint32_t)q[2] << 8) | q[3])
                      : (((uint32_t)q[3] << 24) | ((uint32_t)q[2] << 16) | ((uint32_t)q[1] << 8) | q[0]);
      memcpy(&out[total + i], &u, 4);
    }
    total += bytes / 4;
    ++rec;
  }
  if (n_records) *n_records = rec;
  return total;
}

/* XDR float as on the wire (progress messages): 4 bytes, big-endian */
void hgo_xdr_float(float f, uint8_t out[4]) {
  uint32_t bits;
  memcpy(&bits, &f, 4);
  out[0] = (uint8_t)(bits >> 24); out[1] = (uint8_t)(bits >> 16);
  out[2] = (uint8_t)(bits >> 8);  out[3] = (uint8_t)bits;
}

/* ---- synthetic workload; integer-only randomness so
 * that C, numpy and CUDA produce identical bits from the same (seed, index). */

static inline uint64_t hgo_mix64(uint64_t seed, uint64_t ctr) {
  uint64_t z = seed + (ctr + 1) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

/* sum of eight 16-bit uniforms, centred: Irwin-Hall(8), |v| <= 262140 */
static inline int32_t hgo_ih8(uint64_t a, uint64_t b) {
  int32_t s = 0;
  for (int i = 0; i < 4; ++i) { s += (int32_t)((a >> (16 * i)) & 0xFFFF); s += (int32_t)((b >> (16 * i)) & 0xFFFF); }
  return s - 262140;
}

#define HGO_IH8_SCALE 0x1.3988e2p-16f /* 1/sqrt(8*(65536^2-1)/12) = 1.8688124e-05 as FP32 */

void hgo_synth_centres(uint64_t seed, int32_t k, int32_t d, float *out) {
  for (int64_t i = 0; i < (int64_t)k * d; ++i) {
    uint32_t u = (uint32_t)(hgo_mix64(seed, (uint64_t)i) >> 40); /* 24 bits */
    float f = (float)u * 0x1.9p-17f;                              /* 200 * 2^-24 */
    out[i] = f - 100.0f;
  }
}

/* centroids for the timed iteration: centres + 0.1 * noise */
void hgo_synth_perturb(uint64_t seed, int32_t k, int32_t d, const float *centres, float *out) {
  for (int64_t i = 0; i < (int64_t)k * d; ++i) {
    int32_t v = hgo_ih8(hgo_mix64(seed, 2 * (uint64_t)i), hgo_mix64(seed, 2 * (uint64_t)i + 1));
    float noise = (float)v * HGO_IH8_SCALE;
    float scaled = noise * 0.1f;
    out[i] = centres[i] + scaled;
  }
}

void hgo_synth_points(uint64_t seed, int64_t first, int64_t n, int32_t d, int32_t k,
                      const float *centres, float *out) {
  const uint64_t stride = 2 * (uint64_t)d + 1;
  for (int64_t r = 0; r < n; ++r) {
    uint64_t i = (uint64_t)(first + r);
    uint64_t h = hgo_mix64(seed, i * stride);
    uint32_t comp = (uint32_t)(((h >> 32) * (uint64_t)k) >> 32);
    for (int32_t j = 0; j < d; ++j) {
      int32_t v = hgo_ih8(hgo_mix64(seed, i * stride + 1 + 2 * (uint64_t)j),
                          hgo_mix64(seed, i * stride + 2 + 2 * (uint64_t)j));
      float noise = (float)v * HGO_IH8_SCALE;
      out[(size_t)r * d + j] = centres[(size_t)comp * d + j] + noise;
    }
  }
}
