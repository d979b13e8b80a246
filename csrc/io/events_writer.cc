// TensorBoard event-file writer and TFRecord reader.
//
// File name:  <prefix>.out.tfevents.<10-digit unix time>.<hostname>
// Record:     u64 length | u32 masked_crc32c(length) | bytes | u32 masked_crc32c(bytes)
// First record is Event{wall_time, file_version: "brain.Event:2"}.
// Event proto fields: wall_time = 1 (double), step = 2 (int64), file_version = 3 (string),
// summary = 5 (Summary{ value = 1 repeated { tag = 1 (string), simple_value = 2 (float) } }).
// This is what pywrap_tensorflow.EventsWriter does for the reference's logger.py:148-165.
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "../../include/dtf_b200_io.h"
#include "io_common.h"

namespace dtfio {

thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

}  // namespace dtfio

using namespace dtfio;

extern "C" const char* dtf_io_last_error(void) { return g_err; }

struct dtf_events_writer {
  FILE* f = nullptr;
  std::string filename;
};

static bool write_record(FILE* f, const void* data, size_t n) {
  uint8_t header[12];
  uint64_t len = n;
  memcpy(header, &len, 8);
  uint32_t lcrc = dtf_crc32c_mask(dtf_crc32c(header, 8));
  memcpy(header + 8, &lcrc, 4);
  uint32_t dcrc = dtf_crc32c_mask(dtf_crc32c(data, n));
  if (fwrite(header, 1, 12, f) != 12) return false;
  if (n && fwrite(data, 1, n, f) != n) return false;
  if (fwrite(&dcrc, 1, 4, f) != 4) return false;
  return true;
}

extern "C" dtf_events_writer* dtf_events_open(const char* path_prefix, const char* hostname) {
  char host[256];
  if (hostname && *hostname) {
    snprintf(host, sizeof(host), "%s", hostname);
  } else if (gethostname(host, sizeof(host)) != 0) {
    snprintf(host, sizeof(host), "localhost");
  }
  host[sizeof(host) - 1] = 0;
  double now = static_cast<double>(time(nullptr));
  char name[4096];
  snprintf(name, sizeof(name), "%s.out.tfevents.%010lld.%s", path_prefix,
           static_cast<long long>(now), host);
  FILE* f = fopen(name, "wb");
  if (!f) {
    set_error("cannot open %s for writing", name);
    return nullptr;
  }
  auto* w = new dtf_events_writer();
  w->f = f;
  w->filename = name;
  std::string ev;
  put_double(&ev, 1, now);
  put_string(&ev, 3, "brain.Event:2");
  if (!write_record(f, ev.data(), ev.size())) {
    set_error("short write to %s", name);
    fclose(f);
    delete w;
    return nullptr;
  }
  fflush(f);
  return w;
}

extern "C" const char* dtf_events_filename(dtf_events_writer* w) { return w ? w->filename.c_str() : ""; }

extern "C" int dtf_events_write_scalars(dtf_events_writer* w, double wall_time, int64_t step, int n,
                                        const char* const* tags, const float* values) {
  if (!w || !w->f) { set_error("events writer is closed"); return -1; }
  std::string summary;
  for (int i = 0; i < n; ++i) {
    std::string v;
    put_string(&v, 1, tags[i]);
    put_float(&v, 2, values[i]);
    put_bytes(&summary, 1, v);
  }
  std::string ev;
  put_double(&ev, 1, wall_time);
  put_varint_field(&ev, 2, static_cast<uint64_t>(step));
  put_bytes(&ev, 5, summary);
  if (!write_record(w->f, ev.data(), ev.size())) { set_error("short write"); return -1; }
  return 0;
}

extern "C" int dtf_events_write_raw(dtf_events_writer* w, const void* event_proto, size_t n) {
  if (!w || !w->f) { set_error("events writer is closed"); return -1; }
  if (!write_record(w->f, event_proto, n)) { set_error("short write"); return -1; }
  return 0;
}

extern "C" int dtf_events_flush(dtf_events_writer* w) {
  if (!w || !w->f) return 0;
  return fflush(w->f) == 0 ? 0 : -1;
}

extern "C" int dtf_events_close(dtf_events_writer* w) {
  if (!w) return 0;
  int rc = 0;
  if (w->f) rc = fclose(w->f) == 0 ? 0 : -1;
  delete w;
  return rc;
}

struct dtf_tfrecord_reader {
  FILE* f = nullptr;
  std::vector<uint8_t> buf;
};

extern "C" dtf_tfrecord_reader* dtf_tfrecord_open(const char* path) {
  FILE* f = fopen(path, "rb");
  if (!f) { set_error("cannot open %s", path); return nullptr; }
  auto* r = new dtf_tfrecord_reader();
  r->f = f;
  return r;
}

extern "C" int dtf_tfrecord_next(dtf_tfrecord_reader* r, const void** data, size_t* n) {
  uint8_t header[12];
  size_t got = fread(header, 1, 12, r->f);
  if (got == 0) return 0;
  if (got != 12) { set_error("truncated record header"); return -1; }
  uint64_t len;
  uint32_t lcrc;
  memcpy(&len, header, 8);
  memcpy(&lcrc, header + 8, 4);
  if (dtf_crc32c_mask(dtf_crc32c(header, 8)) != lcrc) { set_error("corrupted record length"); return -2; }
  if (len > (1ull << 31)) { set_error("record too large"); return -2; }
  r->buf.resize(len + 4);
  if (fread(r->buf.data(), 1, len + 4, r->f) != len + 4) { set_error("truncated record"); return -1; }
  uint32_t dcrc;
  memcpy(&dcrc, r->buf.data() + len, 4);
  if (dtf_crc32c_mask(dtf_crc32c(r->buf.data(), len)) != dcrc) { set_error("corrupted record data"); return -2; }
  *data = r->buf.data();
  *n = len;
  return 1;
}

extern "C" int dtf_tfrecord_close(dtf_tfrecord_reader* r) {
  if (!r) return 0;
  if (r->f) fclose(r->f);
  delete r;
  return 0;
}
