// pipes_api.hh -- the C++ surface a Hadoop Pipes program is written against, restated.
//
// A Pipes executable (the -cpubin / -gpubin of this fork, pipes/Submitter.java:419-420) is
// user code compiled against the reference's c++/pipes/api/hadoop/Pipes.hh:36-256 and
// c++/utils/api/hadoop/{StringUtils,SerialUtils}.hh.  This header declares the same
// namespaces, class names, virtuals and free functions with the same meaning, so such a
// program compiles unchanged against either runtime (tests build the reference's own
// wordcount examples against this one and diff the uplink bytes).  The implementation
// behind it (pipes_runtime.cc) is this project's own.
#pragma once

#include <stdint.h>

#include <string>
#include <type_traits>
#include <vector>

namespace HadoopUtils {

// Error type thrown by the runtime and by user code (SerialUtils.hh:29-63).
class Error {
  std::string text_;

 public:
  explicit Error(const std::string &msg) : text_(msg) {}
  Error(const std::string &msg, const std::string &file, int line, const std::string &function);
  const std::string &getMessage() const { return text_; }
};

#define HADOOP_ASSERT(CONDITION, MESSAGE)                                          \
  do {                                                                             \
    if (!(CONDITION)) throw HadoopUtils::Error((MESSAGE), __FILE__, __LINE__, __func__); \
  } while (0)

// StringUtils.hh:34-78
std::string toString(int32_t x);
int toInt(const std::string &val);
float toFloat(const std::string &val);
bool toBool(const std::string &val);
uint64_t getCurrentMillis();
std::vector<std::string> splitString(const std::string &str, const char *separator);
std::string quoteString(const std::string &str, const char *deliminators);
std::string unquoteString(const std::string &str);

// SerialUtils.hh:51-168: byte streams and the Hadoop Writable encodings
class InStream {
 public:
  virtual void read(void *buf, size_t len) = 0;
  virtual ~InStream() {}
};
class OutStream {
 public:
  virtual void write(const void *buf, size_t len) = 0;
  virtual void flush() = 0;
  virtual ~OutStream() {}
};
class StringInStream : public InStream {
  const std::string &buffer_;
  size_t pos_ = 0;

 public:
  explicit StringInStream(const std::string &str) : buffer_(str) {}
  void read(void *buf, size_t len) override;
};
void serializeInt(int32_t t, OutStream &stream);
void serializeLong(int64_t t, OutStream &stream);
void serializeFloat(float t, OutStream &stream);
void serializeString(const std::string &t, OutStream &stream);
int32_t deserializeInt(InStream &stream);
int64_t deserializeLong(InStream &stream);
void deserializeFloat(float &t, InStream &stream);
void deserializeString(std::string &t, InStream &stream);

}  // namespace HadoopUtils

namespace HadoopPipes {

class JobConf {  // Pipes.hh:46-54; get() throws Error for a missing key
 public:
  virtual bool hasKey(const std::string &key) const = 0;
  virtual const std::string &get(const std::string &key) const = 0;
  virtual int getInt(const std::string &key) const = 0;
  virtual float getFloat(const std::string &key) const = 0;
  virtual bool getBoolean(const std::string &key) const = 0;
  virtual ~JobConf() {}
};

class TaskContext {  // Pipes.hh:59-118
 public:
  class Counter {
    int id_;

   public:
    explicit Counter(int counterId) : id_(counterId) {}
    int getId() const { return id_; }
  };
  virtual const JobConf *getJobConf() = 0;
  virtual const std::string &getInputKey() = 0;
  virtual const std::string &getInputValue() = 0;
  virtual void emit(const std::string &key, const std::string &value) = 0;
  virtual void progress() = 0;
  virtual void setStatus(const std::string &status) = 0;
  virtual Counter *getCounter(const std::string &group, const std::string &name) = 0;
  virtual void incrementCounter(const Counter *counter, uint64_t amount) = 0;
  virtual ~TaskContext() {}
};

class MapContext : public TaskContext {  // Pipes.hh:120-138
 public:
  virtual const std::string &getInputSplit() = 0;
  virtual const std::string &getInputKeyClass() = 0;
  virtual const std::string &getInputValueClass() = 0;
};

class ReduceContext : public TaskContext {  // Pipes.hh:140-146
 public:
  virtual bool nextValue() = 0;
};

class Closable {
 public:
  virtual void close() {}
  virtual ~Closable() {}
};
class Mapper : public Closable {  // Pipes.hh:158-161
 public:
  virtual void map(MapContext &context) = 0;
};
class Reducer : public Closable {  // Pipes.hh:166-169
 public:
  virtual void reduce(ReduceContext &context) = 0;
};
class Partitioner {  // Pipes.hh:174-178
 public:
  virtual int partition(const std::string &key, int numOfReduces) = 0;
  virtual ~Partitioner() {}
};
class RecordReader : public Closable {  // Pipes.hh:184-193
 public:
  virtual bool next(std::string &key, std::string &value) = 0;
  virtual float getProgress() = 0;
};
class RecordWriter : public Closable {  // Pipes.hh:198-202
 public:
  virtual void emit(const std::string &key, const std::string &value) = 0;
};

class Factory {  // Pipes.hh:207-247; the optional creators default to "use Java's"
 public:
  virtual Mapper *createMapper(MapContext &context) const = 0;
  virtual Reducer *createReducer(ReduceContext &context) const = 0;
  virtual Reducer *createCombiner(MapContext &) const { return nullptr; }
  virtual Partitioner *createPartitioner(MapContext &) const { return nullptr; }
  virtual RecordReader *createRecordReader(MapContext &) const { return nullptr; }
  virtual RecordWriter *createRecordWriter(ReduceContext &) const { return nullptr; }
  virtual ~Factory() {}
};

// Run the task this process was started for; true on success (Pipes.hh:255).
bool runTask(const Factory &factory);

// TemplateFactory.hh:76-92: build each object from its context; `void` = not supplied.
template <class M, class R, class P = void, class C = void, class RR = void, class RW = void>
class TemplateFactory : public Factory {
  template <class T, class Base, class Ctx>
  static Base *make(Ctx &context) {
    if constexpr (std::is_void<T>::value) {
      return nullptr;
    } else {
      return new T(context);
    }
  }

 public:
  Mapper *createMapper(MapContext &c) const override { return new M(c); }
  Reducer *createReducer(ReduceContext &c) const override { return new R(c); }
  Reducer *createCombiner(MapContext &c) const override { return make<C, Reducer>(c); }
  Partitioner *createPartitioner(MapContext &c) const override { return make<P, Partitioner>(c); }
  RecordReader *createRecordReader(MapContext &c) const override { return make<RR, RecordReader>(c); }
  RecordWriter *createRecordWriter(ReduceContext &c) const override { return make<RW, RecordWriter>(c); }
};

}  // namespace HadoopPipes
