// pipes_runtime.cc -- child side of the Hadoop Pipes protocol, written from scratch.
//
// What the Java task expects from the process it forks (pipes/Application.java:130-215):
// connect to 127.0.0.1:$hadoop.pipes.command.port, answer the HMAC-SHA1 challenge, then obey
// the downlink (START, SET_JOB_CONF, SET_INPUT_TYPES, RUN_MAP, MAP_ITEM..., CLOSE | RUN_REDUCE,
// REDUCE_KEY, REDUCE_VALUE..., CLOSE | ABORT) and report on the uplink (OUTPUT,
// PARTITIONED_OUTPUT, STATUS, PROGRESS, REGISTER/INCREMENT_COUNTER, DONE).  Codes and
// framing: pipes/BinaryProtocol.java:66-84 and :262-368.  The reference's two debugging modes are here
// as well: replay of a downlink file ($hadoop.pipes.command.file -> <file>.out, HadoopPipes.cc:1130-1136)
// and, with neither variable set, the line-oriented text protocol on stdin/stdout (:181-295, 1137-1139),
// byte for byte including its quirks (tests/test_pipes_runtime.py).  Behaviour restated from
// c++/pipes/impl/HadoopPipes.cc (cited per function); none of its code is used.
//
// Differences that matter for a GPU child, all invisible on the wire:
//  - input is pulled from the descriptor in 1 MiB reads and decoded in place: no stdio,
//    no per-record fread, no 64 KiB bounce copy (HadoopPipes.cc:519-524, SerialUtils.cc:272-292);
//  - the keep-alive thread sleeps on a condition variable, so the process ends when the task
//    ends instead of up to 5 s later (HadoopPipes.cc:1049, 1151);
//  - nothing is left unflushed at exit.
#include <arpa/inet.h>
#include <errno.h>
#include <fcntl.h>
#include <netinet/in.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <unistd.h>

#include <map>
#include <memory>
#include <string>
#include <vector>


#include "hgm/pipes_api.hh"

using std::string;
using std::vector;

// =========================================================================================
// HadoopUtils
// =========================================================================================
namespace HadoopUtils {

Error::Error(const string &msg, const string &file, int line, const string &function)
    : text_(msg + " at " + file + ":" + toString(line) + " in " + function) {}

string toString(int32_t x) { return std::to_string(x); }

// StringUtils.cc:40-56: the whole field must be a number ("%d%c" / "%f%c" may match once)
int toInt(const string &val) {
  int result;
  char trailing;
  HADOOP_ASSERT(sscanf(val.c_str(), "%d%c", &result, &trailing) == 1, "Problem converting " + val + " to integer.");
  return result;
}
float toFloat(const string &val) {
  float result;
  char trailing;
  HADOOP_ASSERT(sscanf(val.c_str(), "%f%c", &result, &trailing) == 1, "Problem converting " + val + " to float.");
  return result;
}
bool toBool(const string &val) {
  if (val == "true") return true;
  HADOOP_ASSERT(val == "false", "Problem converting " + val + " to boolean.");
  return false;
}
uint64_t getCurrentMillis() {
  struct timeval tv;
  gettimeofday(&tv, nullptr);
  return (uint64_t)tv.tv_sec * 1000 + tv.tv_usec / 1000;
}
// StringUtils.cc:80-95: split on any separator character, dropping empty fields
vector<string> splitString(const string &str, const char *separator) {
  vector<string> out;
  size_t start = str.find_first_not_of(separator);
  while (start != string::npos) {
    size_t end = str.find_first_of(separator, start);
    out.push_back(str.substr(start, end == string::npos ? string::npos : end - start));
    start = end == string::npos ? end : str.find_first_not_of(separator, end);
  }
  return out;
}
// StringUtils.cc:97-130: backslash-escape unprintables, backslash and the given delimiters
string quoteString(const string &str, const char *deliminators) {
  string out;
  out.reserve(str.size());
  for (unsigned char ch : str) {
    if (isprint(ch) && ch != '\\' && !strchr(deliminators, ch)) {
      out += (char)ch;
    } else if (ch == '\\') {
      out += "\\\\";
    } else if (ch == '\t') {
      out += "\\t";
    } else if (ch == '\n') {
      out += "\\n";
    } else if (ch == ' ') {
      out += "\\s";
    } else {
      char hex[8];
      snprintf(hex, sizeof hex, "\\%02x", ch);
      out += hex;
    }
  }
  return out;
}
string unquoteString(const string &str) {
  string out;
  for (size_t i = 0; i < str.size(); ++i) {
    if (str[i] != '\\') { out += str[i]; continue; }
    HADOOP_ASSERT(i + 1 < str.size(), "trailing \\ in '" + str + "'");
    const char n = str[++i];
    if (isxdigit((unsigned char)n)) {
      HADOOP_ASSERT(i + 1 < str.size() && isxdigit((unsigned char)str[i + 1]),
                    "escape pattern \\<hex><hex> is missing second digit in '" + str + "'");
      out += (char)strtol(str.substr(i, 2).c_str(), nullptr, 16);
      ++i;
    } else if (n == '\\') out += '\\';
    else if (n == 't') out += '\t';
    else if (n == 'n') out += '\n';
    else if (n == 's') out += ' ';
    else HADOOP_ASSERT(false, string("unknown escape character '") + n + "' found in '" + str + "'");
  }
  return out;
}

void StringInStream::read(void *buf, size_t len) {
  HADOOP_ASSERT(pos_ + len <= buffer_.size(), "unexpected end of string reached");
  memcpy(buf, buffer_.data() + pos_, len);
  pos_ += len;
}

// Hadoop vlong (WritableUtils.writeVLong; SerialUtils.cc:182-244).  One byte for
// [-112, 127]; otherwise a tag byte (-113..-120 positive, -121..-128 negative, magnitude
// one's-complemented) followed by 1..8 big-endian bytes.
static int encodeVLong(int64_t v, unsigned char out[9]) {
  if (v >= -112 && v <= 127) { out[0] = (unsigned char)(int8_t)v; return 1; }
  int tag = -112;
  uint64_t mag = (uint64_t)v;
  if (v < 0) { mag = ~mag; tag = -120; }
  int n = 0;
  for (uint64_t t = mag; t; t >>= 8) ++n;
  out[0] = (unsigned char)(int8_t)(tag - n);
  for (int i = 0; i < n; ++i) out[1 + i] = (unsigned char)(mag >> (8 * (n - 1 - i)));
  return 1 + n;
}
void serializeLong(int64_t t, OutStream &stream) {
  unsigned char buf[9];
  stream.write(buf, encodeVLong(t, buf));
}
void serializeInt(int32_t t, OutStream &stream) { serializeLong(t, stream); }
int64_t deserializeLong(InStream &stream) {
  int8_t first;
  stream.read(&first, 1);
  if (first >= -112) return first;
  const bool negative = first < -120;
  const int n = negative ? -120 - first : -112 - first;
  unsigned char raw[8];
  stream.read(raw, n);
  uint64_t mag = 0;
  for (int i = 0; i < n; ++i) mag = (mag << 8) | raw[i];
  return (int64_t)(negative ? ~mag : mag);
}
int32_t deserializeInt(InStream &stream) { return (int32_t)deserializeLong(stream); }
// XDR float = IEEE-754 big-endian (SerialUtils.cc:246-262)
void serializeFloat(float t, OutStream &stream) {
  uint32_t bits;
  memcpy(&bits, &t, 4);
  bits = htonl(bits);
  stream.write(&bits, 4);
}
void deserializeFloat(float &t, InStream &stream) {
  uint32_t bits;
  stream.read(&bits, 4);
  bits = ntohl(bits);
  memcpy(&t, &bits, 4);
}
void serializeString(const string &t, OutStream &stream) {
  serializeInt((int32_t)t.size(), stream);
  if (!t.empty()) stream.write(t.data(), t.size());
}
void deserializeString(string &t, InStream &stream) {
  const int32_t len = deserializeInt(stream);
  if (len <= 0) { t.clear(); return; }
  t.resize(len);
  stream.read(&t[0], len);
}

}  // namespace HadoopUtils

// =========================================================================================
// HadoopPipes runtime
// =========================================================================================
namespace HadoopPipes {
using namespace HadoopUtils;

namespace {

enum Msg : int32_t {  // BinaryProtocol.java:66-84
  START = 0, SET_JOB_CONF, SET_INPUT_TYPES, RUN_MAP, MAP_ITEM, RUN_REDUCE, REDUCE_KEY, REDUCE_VALUE,
  CLOSE, ABORT, AUTHENTICATION_REQ,
  OUTPUT = 50, PARTITIONED_OUTPUT, STATUS, PROGRESS, DONE, REGISTER_COUNTER, INCREMENT_COUNTER,
  AUTHENTICATION_RESP
};

// ---- descriptor streams -------------------------------------------------------------------
class FdIn : public InStream {
  int fd_;
  vector<unsigned char> buf_;
  size_t lo_ = 0, hi_ = 0;

 public:
  explicit FdIn(int fd) : fd_(fd), buf_(1 << 20) {}
  // one byte, or -1 at end of file (the text protocol reads with getc, HadoopPipes.cc:189-199)
  int getByte() {
    while (lo_ == hi_) {
      const ssize_t got = ::read(fd_, buf_.data(), buf_.size());
      if (got < 0 && errno == EINTR) continue;
      HADOOP_ASSERT(got >= 0, string("read error on command stream: ") + strerror(errno));
      if (got == 0) return -1;
      lo_ = 0; hi_ = (size_t)got;
    }
    return buf_[lo_++];
  }
  void read(void *dst, size_t len) override {
    unsigned char *out = static_cast<unsigned char *>(dst);
    while (len) {
      if (lo_ == hi_) {
        if (len >= buf_.size()) {  // large payload: straight into the caller's buffer
          const ssize_t got = ::read(fd_, out, len);
          if (got < 0 && errno == EINTR) continue;
          HADOOP_ASSERT(got > 0, got == 0 ? string("end of file") : string("read error on command stream: ") + strerror(errno));
          out += got; len -= (size_t)got;
          continue;
        }
        const ssize_t got = ::read(fd_, buf_.data(), buf_.size());
        if (got < 0 && errno == EINTR) continue;
        HADOOP_ASSERT(got > 0, got == 0 ? string("end of file") : string("read error on command stream: ") + strerror(errno));
        lo_ = 0; hi_ = (size_t)got;
      }
      const size_t take = std::min(len, hi_ - lo_);
      memcpy(out, buf_.data() + lo_, take);
      lo_ += take; out += take; len -= take;
    }
  }
};

class FdOut : public OutStream {
  int fd_;
  vector<unsigned char> buf_;
  size_t used_ = 0;
  void drain(const unsigned char *p, size_t n) {
    while (n) {
      const ssize_t put = ::write(fd_, p, n);
      if (put < 0 && errno == EINTR) continue;
      HADOOP_ASSERT(put > 0, string("write error to file: ") + strerror(errno));
      p += put; n -= (size_t)put;
    }
  }

 public:
  explicit FdOut(int fd) : fd_(fd), buf_(128 * 1024) {}  // BinaryProtocol.java:54 BUFFER_SIZE
  void write(const void *src, size_t len) override {
    if (used_ + len > buf_.size()) flush();
    if (len >= buf_.size()) { drain(static_cast<const unsigned char *>(src), len); return; }
    memcpy(buf_.data() + used_, src, len);
    used_ += len;
  }
  void flush() override {
    if (used_) drain(buf_.data(), used_);
    used_ = 0;
  }
};

// ---- uplink: binary (HadoopPipes.cc:302-366) or, for a child started by hand with neither a command
// port nor a command file, the text lines of TextUpwardProtocol (HadoopPipes.cc:122-179) ------------------
class Uplink {
  FdOut out_;
  bool text_ = false;

  void put(const string &s) { out_.write(s.data(), s.size()); }
  // keys and values are quoted (tab, newline, backslash, unprintables); the reference passes the quoted
  // string to fprintf as the FORMAT (HadoopPipes.cc:128-130), so a '%' in a key is undefined there --
  // it is written as it is here
  void putQuoted(const string &s) { put(quoteString(s, "\t\n")); }
  template <class... A>
  void putf(const char *fmt, A... a) {
    char buf[96];
    const int n = snprintf(buf, sizeof buf, fmt, a...);
    out_.write(buf, (size_t)n);
  }

 public:
  explicit Uplink(int fd) : out_(fd) {}
  void setText() { text_ = true; }
  void authenticate(const string &digest) {
    serializeInt(AUTHENTICATION_RESP, out_);
    serializeString(digest, out_);
    out_.flush();
  }
  void output(const string &k, const string &v) {
    if (text_) { put("output\t"); putQuoted(k); put("\t"); putQuoted(v); put("\n"); return; }
    serializeInt(OUTPUT, out_);
    serializeString(k, out_);
    serializeString(v, out_);
  }
  void partitionedOutput(int reduce, const string &k, const string &v) {
    if (text_) {  // the word is misspelt in the reference (HadoopPipes.cc:145) and a reader of it expects that
      putf("parititionedOutput\t%d\t", reduce); putQuoted(k); put("\t"); putQuoted(v); put("\n");
      return;
    }
    serializeInt(PARTITIONED_OUTPUT, out_);
    serializeInt(reduce, out_);
    serializeString(k, out_);
    serializeString(v, out_);
  }
  void status(const string &msg) {
    if (text_) { put("status\t" + msg + "\n"); return; }
    serializeInt(STATUS, out_);
    serializeString(msg, out_);
  }
  void progress(float p) {
    if (text_) { putf("progress\t%f\n", (double)p); out_.flush(); return; }
    serializeInt(PROGRESS, out_);
    serializeFloat(p, out_);
    out_.flush();
  }
  void done() {
    if (text_) put("done\n");
    else serializeInt(DONE, out_);
    out_.flush();
  }
  void registerCounter(int id, const string &group, const string &name) {
    if (text_) { putf("registerCounter\t%d\t", id); put(group + "\t" + name + "\n"); return; }
    serializeInt(REGISTER_COUNTER, out_);
    serializeInt(id, out_);
    serializeString(group, out_);
    serializeString(name, out_);
  }
  void incrementCounter(int id, uint64_t amount) {
    if (text_) { putf("incrCounter\t%d\t%ld\n", id, (long)amount); return; }
    serializeInt(INCREMENT_COUNTER, out_);
    serializeInt(id, out_);
    serializeLong((int64_t)amount, out_);
  }
  void flush() { out_.flush(); }
};

class Conf : public JobConf {  // HadoopPipes.cc:48-83
  std::map<string, string> kv_;

 public:
  void set(const string &k, const string &v) { kv_[k] = v; }
  bool hasKey(const string &key) const override { return kv_.count(key) != 0; }
  const string &get(const string &key) const override {
    auto it = kv_.find(key);
    if (it == kv_.end()) throw Error("Key " + key + " not found in JobConf");
    return it->second;
  }
  int getInt(const string &key) const override { return toInt(get(key)); }
  float getFloat(const string &key) const override { return toFloat(get(key)); }
  bool getBoolean(const string &key) const override { return toBool(get(key)); }
};

// SHA-1 (FIPS 180-4) and HMAC (RFC 2104), here rather than from libcrypto: a map task lives for milliseconds and
// loading OpenSSL was a third of what it costs to start one.  Checked against RFC 2202's vectors and against the
// Java side's digests (tests/test_pipes_runtime.py).
void sha1(const unsigned char *data, size_t len, unsigned char out[20]) {
  uint32_t h[5] = {0x67452301u, 0xEFCDAB89u, 0x98BADCFEu, 0x10325476u, 0xC3D2E1F0u};
  auto rol = [](uint32_t v, int n) { return (v << n) | (v >> (32 - n)); };
  auto block = [&](const unsigned char *p) {
    uint32_t w[80];
    for (int i = 0; i < 16; ++i)
      w[i] = (uint32_t)p[4 * i] << 24 | (uint32_t)p[4 * i + 1] << 16 | (uint32_t)p[4 * i + 2] << 8 | p[4 * i + 3];
    for (int i = 16; i < 80; ++i) w[i] = rol(w[i - 3] ^ w[i - 8] ^ w[i - 14] ^ w[i - 16], 1);
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4];
    for (int i = 0; i < 80; ++i) {
      uint32_t f, k;
      if (i < 20) { f = (b & c) | (~b & d); k = 0x5A827999u; }
      else if (i < 40) { f = b ^ c ^ d; k = 0x6ED9EBA1u; }
      else if (i < 60) { f = (b & c) | (b & d) | (c & d); k = 0x8F1BBCDCu; }
      else { f = b ^ c ^ d; k = 0xCA62C1D6u; }
      const uint32_t t = rol(a, 5) + f + e + k + w[i];
      e = d; d = c; c = rol(b, 30); b = a; a = t;
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e;
  };
  size_t done = 0;
  for (; done + 64 <= len; done += 64) block(data + done);
  unsigned char tail[128] = {0};
  const size_t rest = len - done;
  memcpy(tail, data + done, rest);
  tail[rest] = 0x80;
  const size_t padded = rest + 9 <= 64 ? 64 : 128;
  const uint64_t bits = (uint64_t)len * 8;
  for (int i = 0; i < 8; ++i) tail[padded - 1 - i] = (unsigned char)(bits >> (8 * i));
  block(tail);
  if (padded == 128) block(tail + 64);
  for (int i = 0; i < 5; ++i)
    for (int j = 0; j < 4; ++j) out[4 * i + j] = (unsigned char)(h[i] >> (24 - 8 * j));
}

void hmacSha1(const string &key, const string &msg, unsigned char mac[20]) {
  unsigned char k[64] = {0};
  if (key.size() > 64) sha1(reinterpret_cast<const unsigned char *>(key.data()), key.size(), k);
  else memcpy(k, key.data(), key.size());
  string inner(64, '\0'), outer(64, '\0');
  for (int i = 0; i < 64; ++i) { inner[i] = (char)(k[i] ^ 0x36); outer[i] = (char)(k[i] ^ 0x5c); }
  inner += msg;
  unsigned char digest[20];
  sha1(reinterpret_cast<const unsigned char *>(inner.data()), inner.size(), digest);
  outer.append(reinterpret_cast<const char *>(digest), 20);
  sha1(reinterpret_cast<const unsigned char *>(outer.data()), outer.size(), mac);
}

// base64(HMAC-SHA1(key, msg)) (HadoopPipes.cc:421-448; Java: SecureShuffleUtils.hashFromString)
string hmacDigest(const string &key, const string &msg) {
  unsigned char mac[20];
  hmacSha1(key, msg, mac);
  static const char alphabet[] = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
  string out;
  for (int i = 0; i < 20; i += 3) {
    const int n = 20 - i >= 3 ? 3 : 20 - i;
    const uint32_t v = (uint32_t)mac[i] << 16 | (uint32_t)(n > 1 ? mac[i + 1] : 0) << 8 | (uint32_t)(n > 2 ? mac[i + 2] : 0);
    out += alphabet[v >> 18 & 63];
    out += alphabet[v >> 12 & 63];
    out += n > 1 ? alphabet[v >> 6 & 63] : '=';
    out += n > 2 ? alphabet[v & 63] : '=';
  }
  return out;
}

class Task;

// ---- the combiner buffer (HadoopPipes.cc:564-706) -----------------------------------------------
// Map output is held per key, keys in byte order, and handed to the combiner when the
// buffered key+value bytes reach io.sort.mb MiB and once more at close.
class CombineBuffer : public RecordWriter {
  std::map<string, vector<string>> held_;
  int64_t bytes_ = 0, limit_;
  Task *task_;
  Reducer *combiner_;

 public:
  CombineBuffer(int64_t limit, Task *task, Reducer *combiner) : limit_(limit), task_(task), combiner_(combiner) {}
  void emit(const string &key, const string &value) override {
    bytes_ += (int64_t)(key.size() + value.size());
    held_[key].push_back(value);
    if (bytes_ >= limit_) spill();
  }
  void close() override { spill(); }
  void spill();
};

// ---- the task: context object + downlink interpreter (HadoopPipes.cc:708-1036) -----------------
class Task : public MapContext, public ReduceContext {
 public:
  const Factory &factory;
  FdIn in;
  Uplink up;
  std::unique_ptr<Conf> conf;
  string password;
  bool authenticated = false;
  string keyClass, valueClass, split;
  string key, value;          // current record
  string pendingKey;          // key announced by REDUCE_KEY / MAP_ITEM, not yet taken
  bool havePendingKey = false, haveNewValue = false, hasTask = false;
  bool done = false;
  pthread_mutex_t doneLock = PTHREAD_MUTEX_INITIALIZER;
  pthread_cond_t doneCond = PTHREAD_COND_INITIALIZER;
  std::unique_ptr<RecordReader> reader;
  std::unique_ptr<Mapper> mapper;
  std::unique_ptr<Reducer> reducer;  // combiner in a map task
  std::unique_ptr<RecordWriter> writer;
  std::unique_ptr<Partitioner> partitioner;
  int numReduces = 0;
  float progressValue = 0.f;
  uint64_t lastProgressMs = 0;
  string statusText;
  bool statusSet = false;
  int nextCounterId = 0;

  bool text = false;  // TextProtocol on stdin/stdout (HadoopPipes.cc:181-295, 1137-1139)

  Task(const Factory &f, int inFd, int outFd, bool textProtocol = false)
      : factory(f), in(inFd), up(outFd), text(textProtocol) {
    if (text) up.setText();
    if (const char *path = getenv("hadoop.pipes.shared.secret.location")) {  // HadoopPipes.cc:375-393
      if (FILE *fp = fopen(path, "rb")) {
        char buf[512];
        const size_t n = fread(buf, 1, sizeof buf, fp);
        fclose(fp);
        password.assign(buf, n);
      } else {
        fprintf(stderr, "Could not open the password file\n");
      }
    }
  }

  bool isDone() {
    pthread_mutex_lock(&doneLock);
    const bool d = done;
    pthread_mutex_unlock(&doneLock);
    return d;
  }
  void markDone() {
    pthread_mutex_lock(&doneLock);
    done = true;
    pthread_cond_broadcast(&doneCond);
    pthread_mutex_unlock(&doneLock);
  }

  // text protocol: bytes up to a tab or newline; returns the separator, -1 at end of file.  Nothing is
  // unquoted on the way in (HadoopPipes.cc:189-199), only the uplink quotes.
  int readUpto(string &buffer) {
    buffer.clear();
    int ch;
    while ((ch = in.getByte()) != -1) {
      if (ch == '\t' || ch == '\n') return ch;
      buffer += (char)ch;
    }
    return -1;
  }
  void wantSep(int sep, int wanted, const string &command) {
    if (sep != wanted) throw Error(string(wanted == '\t' ? "Short" : "Long") + " text protocol command " + command);
  }
  // one line of the text downlink (HadoopPipes.cc:214-289): no authentication in this mode
  void nextTextEvent() {
    string command, arg;
    int sep = readUpto(command);
    if (command == "mapItem") {
      wantSep(sep, '\t', command);
      sep = readUpto(pendingKey);
      wantSep(sep, '\t', command);
      sep = readUpto(value);
      wantSep(sep, '\n', command);
      havePendingKey = true;
    } else if (command == "reduceValue") {
      wantSep(sep, '\t', command);
      sep = readUpto(value);
      wantSep(sep, '\n', command);
      haveNewValue = true;
    } else if (command == "reduceKey") {
      wantSep(sep, '\t', command);
      sep = readUpto(pendingKey);
      wantSep(sep, '\n', command);
      havePendingKey = true;
    } else if (command == "start") {
      wantSep(sep, '\t', command);
      sep = readUpto(arg);
      wantSep(sep, '\n', command);
      const int version = toInt(arg);
      if (version != 0) throw Error("Protocol version " + toString(version) + " not supported");
    } else if (command == "setJobConf") {
      wantSep(sep, '\t', command);
      sep = readUpto(arg);
      const int len = toInt(arg);
      // as the reference builds it: `len` empty strings, then the `len` values read (HadoopPipes.cc:244-250),
      // paired off two by two -- so an odd `len` pairs ("", first) and shifts the rest
      vector<string> values((size_t)std::max(len, 0));
      for (int i = 0; i < len; ++i) {
        wantSep(sep, '\t', command);
        sep = readUpto(arg);
        values.push_back(arg);
      }
      wantSep(sep, '\n', command);
      conf.reset(new Conf());
      for (size_t i = 0; i + 1 < values.size(); i += 2) conf->set(values[i], values[i + 1]);
    } else if (command == "setInputTypes") {
      wantSep(sep, '\t', command);
      sep = readUpto(keyClass);
      wantSep(sep, '\t', command);
      sep = readUpto(valueClass);
      wantSep(sep, '\n', command);
    } else if (command == "runMap") {
      string reduces;
      wantSep(sep, '\t', command);
      sep = readUpto(split);
      wantSep(sep, '\t', command);
      sep = readUpto(reduces);
      wantSep(sep, '\t', command);
      sep = readUpto(arg);
      wantSep(sep, '\n', command);
      numReduces = toInt(reduces);
      startMap(toBool(arg));
    } else if (command == "runReduce") {
      string piped;
      wantSep(sep, '\t', command);
      sep = readUpto(arg);
      wantSep(sep, '\t', command);
      sep = readUpto(piped);
      wantSep(sep, '\n', command);
      toInt(arg);  // partition number: informational
      startReduce(toBool(piped));
    } else if (command == "abort") {
      wantSep(sep, '\n', command);
      throw Error("Aborted by driver");
    } else if (command == "close") {
      wantSep(sep, '\n', command);
      markDone();
    } else {
      throw Error("Illegal text protocol command " + command);
    }
  }

  void startReduce(bool piped) {  // HadoopPipes.cc:817-826
    reducer.reset(factory.createReducer(*this));
    writer.reset(factory.createRecordWriter(*this));
    HADOOP_ASSERT((writer == nullptr) == piped,
                  piped ? "RecordWriter defined when not needed." : "RecordWriter not defined");
    hasTask = true;
  }

  // one downlink message (HadoopPipes.cc:464-552)
  void nextEvent() {
    if (text) return nextTextEvent();
    const int32_t cmd = deserializeInt(in);
    if (!authenticated && cmd != AUTHENTICATION_REQ) {
      fprintf(stderr, "Command:%d received before authentication. Exiting..\n", cmd);
      exit(-1);
    }
    switch (cmd) {
      case AUTHENTICATION_REQ: {
        string digest, challenge;
        deserializeString(digest, in);
        deserializeString(challenge, in);
        if (!password.empty()) {  // empty password = command-file debugging (HadoopPipes.cc:396-401)
          if (hmacDigest(password, challenge) != digest) {
            fprintf(stderr, "Server failed to authenticate. Exiting\n");
            exit(-1);
          }
          up.authenticate(hmacDigest(password, digest));
        }
        authenticated = true;
        break;
      }
      case START: {
        const int32_t version = deserializeInt(in);
        if (version != 0) throw Error("Protocol version " + toString(version) + " not supported");
        break;
      }
      case SET_JOB_CONF: {
        const int32_t entries = deserializeInt(in);
        HADOOP_ASSERT(entries % 2 == 0, "Odd length of job conf values");
        conf.reset(new Conf());
        string k, v;
        for (int i = 0; i < entries; i += 2) {
          deserializeString(k, in);
          deserializeString(v, in);
          conf->set(k, v);
        }
        break;
      }
      case SET_INPUT_TYPES:
        deserializeString(keyClass, in);
        deserializeString(valueClass, in);
        break;
      case RUN_MAP: {
        deserializeString(split, in);
        numReduces = deserializeInt(in);
        const bool piped = deserializeInt(in) != 0;
        startMap(piped);
        break;
      }
      case MAP_ITEM:
        deserializeString(pendingKey, in);
        deserializeString(value, in);
        havePendingKey = true;
        break;
      case RUN_REDUCE: {
        deserializeInt(in);  // partition number: informational
        const bool piped = deserializeInt(in) != 0;
        startReduce(piped);
        break;
      }
      case REDUCE_KEY:
        deserializeString(pendingKey, in);
        havePendingKey = true;
        break;
      case REDUCE_VALUE:
        deserializeString(value, in);
        haveNewValue = true;
        break;
      case CLOSE:
        markDone();
        break;
      case ABORT:
        throw Error("Aborted by driver");
      default:
        HADOOP_ASSERT(false, "Unknown binary command " + toString(cmd));
    }
  }

  // HadoopPipes.cc:791-815
  void startMap(bool pipedInput) {
    HADOOP_ASSERT(conf != nullptr, "RUN_MAP before SET_JOB_CONF");
    reader.reset(factory.createRecordReader(*this));
    HADOOP_ASSERT((reader == nullptr) == pipedInput,
                  pipedInput ? "RecordReader defined when not needed." : "RecordReader not defined");
    mapper.reset(factory.createMapper(*this));
    if (numReduces != 0) {
      reducer.reset(factory.createCombiner(*this));
      partitioner.reset(factory.createPartitioner(*this));
    }
    if (reducer) {
      int64_t mb = 100;
      if (conf->hasKey("io.sort.mb")) mb = conf->getInt("io.sort.mb");
      writer.reset(new CombineBuffer(mb * 1024 * 1024, this, reducer.get()));
    }
    hasTask = true;
  }

  void waitForTask() {
    while (!isDone() && !hasTask) nextEvent();
  }

  // advance to the next key and run user code on it (HadoopPipes.cc:865-890)
  bool nextKey() {
    if (!reader) {
      while (!havePendingKey) {
        nextValue();
        if (isDone()) return false;
      }
      key.swap(pendingKey);
    } else {
      if (!reader->next(key, value)) {
        markDone();
        return false;
      }
      progressValue = reader->getProgress();
    }
    havePendingKey = false;
    if (mapper) mapper->map(*this);
    else reducer->reduce(*this);
    return true;
  }

  // ---- ReduceContext / TaskContext -----------------------------------------------------------
  bool nextValue() override {  // HadoopPipes.cc:895-903
    if (havePendingKey || isDone()) return false;
    haveNewValue = false;
    progress();
    nextEvent();
    return haveNewValue;
  }
  const JobConf *getJobConf() override { return conf.get(); }
  const string &getInputKey() override { return key; }
  const string &getInputValue() override { return value; }
  const string &getInputKeyClass() override { return keyClass; }
  const string &getInputValueClass() override { return valueClass; }
  const string &getInputSplit() override { return split; }

  void progress() override {  // at most one PROGRESS (and pending STATUS) per second, HadoopPipes.cc:933-945
    const uint64_t now = getCurrentMillis();
    if (now - lastProgressMs > 1000) {
      lastProgressMs = now;
      if (statusSet) { up.status(statusText); statusSet = false; }
      up.progress(progressValue);
    }
  }
  void setStatus(const string &s) override {
    statusText = s;
    statusSet = true;
    progress();
  }
  void emit(const string &k, const string &v) override {  // HadoopPipes.cc:977-987
    progress();
    if (writer) writer->emit(k, v);
    else send(k, v);
  }
  // partition (if any) and put on the wire; also where the combiner's output goes
  void send(const string &k, const string &v) {
    if (partitioner) up.partitionedOutput(partitioner->partition(k, numReduces), k, v);
    else up.output(k, v);
  }
  Counter *getCounter(const string &group, const string &name) override {  // HadoopPipes.cc:992-998
    const int id = nextCounterId++;
    up.registerCounter(id, group, name);
    return new Counter(id);
  }
  void incrementCounter(const Counter *counter, uint64_t amount) override { up.incrementCounter(counter->getId(), amount); }

  void closeAll() {  // HadoopPipes.cc:1007-1020
    if (reader) reader->close();
    if (mapper) mapper->close();
    if (reducer) reducer->close();
    if (writer) writer->close();
  }
};

// the view of the buffered map output a combiner iterates (HadoopPipes.cc:564-654)
class CombineContext : public ReduceContext {
  Task *task_;
  std::map<string, vector<string>>::iterator key_, end_;
  size_t idx_ = 0;
  bool first_ = true;

 public:
  CombineContext(Task *t, std::map<string, vector<string>> &data) : task_(t), key_(data.begin()), end_(data.end()) {}
  bool more() const { return key_ != end_; }
  void advance() { ++key_; first_ = true; idx_ = 0; }
  const JobConf *getJobConf() override { return task_->getJobConf(); }
  const string &getInputKey() override { return key_->first; }
  const string &getInputValue() override { return key_->second[idx_]; }
  void emit(const string &k, const string &v) override { task_->send(k, v); }
  void progress() override { task_->progress(); }
  void setStatus(const string &s) override { task_->setStatus(s); }
  bool nextValue() override {
    if (first_) first_ = false;
    else ++idx_;
    return idx_ < key_->second.size();
  }
  Counter *getCounter(const string &g, const string &n) override { return task_->getCounter(g, n); }
  void incrementCounter(const Counter *c, uint64_t a) override { task_->incrementCounter(c, a); }
};

void CombineBuffer::spill() {
  CombineContext ctx(task_, held_);
  for (; ctx.more(); ctx.advance()) combiner_->reduce(ctx);
  held_.clear();
  bytes_ = 0;
}

// Keep-alive: if the parent's command port stops accepting connections three times in a
// row, the task JVM is gone and so should we be (HadoopPipes.cc:1042-1085).
void *keepAlive(void *arg) {
  Task *task = static_cast<Task *>(arg);
  const char *portStr = getenv("hadoop.pipes.command.port");
  int failures = 0;
  pthread_mutex_lock(&task->doneLock);
  while (!task->done) {
    struct timespec until;
    clock_gettime(CLOCK_REALTIME, &until);
    until.tv_sec += 5;
    pthread_cond_timedwait(&task->doneCond, &task->doneLock, &until);
    if (task->done || !portStr) continue;
    pthread_mutex_unlock(&task->doneLock);
    bool ok = false;
    const int sock = socket(PF_INET, SOCK_STREAM, 0);
    if (sock >= 0) {
      sockaddr_in addr{};
      addr.sin_family = AF_INET;
      addr.sin_port = htons((uint16_t)atoi(portStr));
      addr.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
      ok = connect(sock, reinterpret_cast<sockaddr *>(&addr), sizeof addr) == 0;
      if (ok) shutdown(sock, SHUT_RDWR);
      close(sock);
    }
    pthread_mutex_lock(&task->doneLock);
    if (ok) failures = 0;
    else if (!task->done && ++failures == 3) {
      fprintf(stderr, "Hadoop Pipes Exception: in ping: parent command port is gone\n");
      exit(1);
    }
  }
  pthread_mutex_unlock(&task->doneLock);
  return nullptr;
}

}  // namespace

// HadoopPipes.cc:1093-1179
bool runTask(const Factory &factory) {
  int inFd = -1, outFd = -1, sock = -1;
  bool textProtocol = false;
  try {
    if (const char *portStr = getenv("hadoop.pipes.command.port")) {
      sock = socket(PF_INET, SOCK_STREAM, 0);
      HADOOP_ASSERT(sock != -1, string("problem creating socket: ") + strerror(errno));
      sockaddr_in addr{};
      addr.sin_family = AF_INET;
      addr.sin_port = htons((uint16_t)toInt(portStr));
      addr.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
      HADOOP_ASSERT(connect(sock, reinterpret_cast<sockaddr *>(&addr), sizeof addr) == 0,
                    string("problem connecting command socket: ") + strerror(errno));
      inFd = outFd = sock;
    } else if (const char *file = getenv("hadoop.pipes.command.file")) {
      inFd = open(file, O_RDONLY);
      HADOOP_ASSERT(inFd >= 0, string("cannot open command file ") + file + ": " + strerror(errno));
      const string outName = string(file) + ".out";
      outFd = open(outName.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
      HADOOP_ASSERT(outFd >= 0, "cannot create " + outName + ": " + strerror(errno));
    } else {  // started by hand: the text protocol on stdin/stdout (HadoopPipes.cc:1137-1139)
      inFd = STDIN_FILENO;
      outFd = STDOUT_FILENO;
      textProtocol = true;
    }
    Task task(factory, inFd, outFd, textProtocol);
    pthread_t pinger;
    pthread_create(&pinger, nullptr, keepAlive, &task);
    bool ok = true;
    string failure;
    try {
      task.waitForTask();
      while (!task.isDone()) task.nextKey();
      task.closeAll();
      task.up.done();
    } catch (Error &err) {
      ok = false;
      failure = err.getMessage();
    }
    task.markDone();
    pthread_join(pinger, nullptr);
    if (!ok) throw Error(failure);
    if (sock != -1) {
      shutdown(sock, SHUT_RDWR);
      close(sock);
    } else if (!textProtocol) {
      close(inFd);
      close(outFd);
    }
    return true;
  } catch (Error &err) {
    fprintf(stderr, "Hadoop Pipes Exception: %s\n", err.getMessage().c_str());
    return false;
  }
}

}  // namespace HadoopPipes
