// TensorBoard events file writer: TFRecord framing around hand-encoded Event protos.
//
// The reference attaches tf.summary.scalar("loss_<task_id>") / ("accuracy_<task_id>") to the graph
// (distribute_training.py:128-132) and MonitoredTrainingSession's chief writes them — plus global_step/sec —
// every 100 steps into events.out.tfevents.* in the checkpoint directory.
//
//   record = uint64le length | uint32le masked_crc32c(length bytes) | data | uint32le masked_crc32c(data)
//   Event  { double wall_time = 1; int64 step = 2; oneof { string file_version = 3; Summary summary = 5 } }
//   Summary{ repeated Value value = 1 { string tag = 1; float simple_value = 2 } }
#include <errno.h>
#include <stdio.h>
#include <string.h>
#include <unistd.h>

#include <mutex>
#include <string>

#include "../../include/ttd_io.h"
#include "coding.h"

namespace ttdio {

struct EventsWriter {
  FILE* f = nullptr;
  std::mutex mu;
  std::string path;
};

static void put_double(std::string* s, double v) {
  uint64_t bits;
  memcpy(&bits, &v, 8);
  put_fixed64(s, bits);
}

static std::string event_header(double wall_time, int64_t step) {
  std::string e;
  e.push_back(char((1 << 3) | 1));  // wall_time, fixed64
  put_double(&e, wall_time);
  if (step != 0) {
    e.push_back(char((2 << 3) | 0));
    put_varint64(&e, static_cast<uint64_t>(step));
  }
  return e;
}

static bool write_record(EventsWriter* w, const std::string& data) {
  std::string rec;
  put_fixed64(&rec, data.size());
  put_fixed32(&rec, crc_mask(crc32c(rec.data(), 8)));
  rec += data;
  put_fixed32(&rec, crc_mask(crc32c(data.data(), data.size())));
  std::lock_guard<std::mutex> lock(w->mu);
  if (fwrite(rec.data(), 1, rec.size(), w->f) != rec.size()) {
    set_error("events: short write to " + w->path);
    return false;
  }
  return true;
}

}  // namespace ttdio

using namespace ttdio;

extern "C" void* ttdio_events_open(const char* logdir, double wall_time, const char* filename_suffix) {
  char host[256] = "localhost";
  gethostname(host, sizeof(host) - 1);
  char name[1024];
  snprintf(name, sizeof(name), "%s/events.out.tfevents.%010lld.%s%s", logdir, static_cast<long long>(wall_time), host,
           filename_suffix ? filename_suffix : "");
  EventsWriter* w = new EventsWriter;
  w->path = name;
  w->f = fopen(name, "ab");
  if (!w->f) {
    set_error(std::string("events: cannot open ") + name + ": " + strerror(errno));
    delete w;
    return nullptr;
  }
  std::string e = event_header(wall_time, 0);
  const char* ver = "brain.Event:2";
  e.push_back(char((3 << 3) | 2));
  put_varint32(&e, static_cast<uint32_t>(strlen(ver)));
  e += ver;
  if (!write_record(w, e)) {
    fclose(w->f);
    delete w;
    return nullptr;
  }
  fflush(w->f);
  return w;
}

extern "C" const char* ttdio_events_path(void* h) { return static_cast<EventsWriter*>(h)->path.c_str(); }

extern "C" int ttdio_events_add_scalar(void* h, const char* tag, float value, int64_t step, double wall_time) {
  std::string val;  // Summary.Value
  val.push_back(char((1 << 3) | 2));
  put_varint32(&val, static_cast<uint32_t>(strlen(tag)));
  val += tag;
  val.push_back(char((2 << 3) | 5));
  uint32_t bits;
  memcpy(&bits, &value, 4);
  put_fixed32(&val, bits);
  std::string summary;
  summary.push_back(char((1 << 3) | 2));
  put_varint32(&summary, static_cast<uint32_t>(val.size()));
  summary += val;
  std::string e = event_header(wall_time, step);
  e.push_back(char((5 << 3) | 2));
  put_varint32(&e, static_cast<uint32_t>(summary.size()));
  e += summary;
  return write_record(static_cast<EventsWriter*>(h), e) ? 0 : -2;
}

extern "C" int ttdio_events_flush(void* h) { return fflush(static_cast<EventsWriter*>(h)->f) == 0 ? 0 : -2; }

extern "C" void ttdio_events_close(void* h) {
  EventsWriter* w = static_cast<EventsWriter*>(h);
  if (w->f) fclose(w->f);
  delete w;
}
