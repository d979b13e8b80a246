// Self-test driver for the host I/O library, meant to be built with
// -fsanitize=address,undefined (tests/test_host_sanitized.py): every public entry point of
// include/dtf_b200_io.h is driven through a write -> read-back cycle in a scratch directory so
// the sanitizers see the SSTable builder/reader, the bundle writer/merger/reader, the TFRecord
// framing and crc32c on realistic inputs. Exit code 0 = all checks passed.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../../include/dtf_b200_io.h"

static int g_fail = 0;
#define CHECK(cond)                                                                  \
  do {                                                                               \
    if (!(cond)) {                                                                   \
      std::fprintf(stderr, "selftest: %s:%d: CHECK(%s) failed: %s\n", __FILE__, __LINE__, #cond, \
                   dtf_io_last_error());                                             \
      ++g_fail;                                                                      \
    }                                                                                \
  } while (0)

static void test_crc() {
  // RFC 3720 B.4 vectors
  std::vector<unsigned char> zeros(32, 0), ones(32, 0xff), inc(32);
  for (int i = 0; i < 32; ++i) inc[i] = static_cast<unsigned char>(i);
  CHECK(dtf_crc32c(zeros.data(), 32) == 0x8a9136aau);
  CHECK(dtf_crc32c(ones.data(), 32) == 0x62a8ab43u);
  CHECK(dtf_crc32c(inc.data(), 32) == 0x46dd794eu);
  CHECK(dtf_crc32c("123456789", 9) == 0xe3069283u);
  const uint32_t whole = dtf_crc32c(inc.data(), 32);
  // unaligned split: extend must compose
  for (int cut = 0; cut <= 32; ++cut)
    CHECK(dtf_crc32c_extend(dtf_crc32c(inc.data(), cut), inc.data() + cut, 32 - cut) == whole);
  CHECK(dtf_crc32c_unmask(dtf_crc32c_mask(whole)) == whole);
}

static void test_events(const std::string& dir) {
  const std::string prefix = dir + "/events";
  dtf_events_writer* w = dtf_events_open(prefix.c_str(), "selftest-host");
  CHECK(w != nullptr);
  if (!w) return;
  const std::string fname = dtf_events_filename(w);
  const char* tags[2] = {"Global Step", "Loss"};
  for (int step = 0; step < 50; ++step) {
    const float vals[2] = {static_cast<float>(step), 2.5f / (1 + step)};
    CHECK(dtf_events_write_scalars(w, 1.5e9 + step, step, 2, tags, vals) == 0);
  }
  CHECK(dtf_events_flush(w) == 0);
  CHECK(dtf_events_close(w) == 0);
  dtf_tfrecord_reader* r = dtf_tfrecord_open(fname.c_str());
  CHECK(r != nullptr);
  if (!r) return;
  int n = 0, rc;
  const void* data;
  size_t len;
  while ((rc = dtf_tfrecord_next(r, &data, &len)) == 1) {
    CHECK(len > 0);
    ++n;
  }
  CHECK(rc == 0);
  CHECK(n == 51);  // file_version record + 50 scalar events
  CHECK(dtf_tfrecord_close(r) == 0);
}

static void test_table(const std::string& dir) {
  const std::string path = dir + "/table.sst";
  dtf_table_builder* b = dtf_table_builder_open(path.c_str(), 256);  // small blocks: many of them
  CHECK(b != nullptr);
  if (!b) return;
  char key[64], val[200];
  for (int i = 0; i < 500; ++i) {
    std::snprintf(key, sizeof key, "layer_%04d/kernel", i);
    std::memset(val, 'a' + i % 26, sizeof val);
    CHECK(dtf_table_builder_add(b, key, std::strlen(key), val, 1 + i % 199) == 0);
  }
  // out-of-order key must be rejected, not corrupt the table
  CHECK(dtf_table_builder_add(b, "a", 1, "x", 1) != 0);
  CHECK(dtf_table_builder_finish(b) == 0);
}

static void test_bundle(const std::string& dir) {
  const int kShards = 3;
  std::vector<std::string> parts;
  std::vector<std::vector<float>> payload;
  for (int s = 0; s < kShards; ++s) {
    const std::string prefix = dir + "/tmp_part" + std::to_string(s) + "/part";
    const std::string mk = "mkdir -p " + dir + "/tmp_part" + std::to_string(s);
    CHECK(std::system(mk.c_str()) == 0);
    parts.push_back(prefix);
    dtf_bundle_writer* w = dtf_bundle_writer_open(prefix.c_str());
    CHECK(w != nullptr);
    if (!w) return;
    for (int v = 0; v < 4; ++v) {
      std::vector<float> t(static_cast<size_t>(1000 * (v + 1) + s));
      for (size_t i = 0; i < t.size(); ++i) t[i] = std::sin(0.01f * i + s + v);
      const int64_t dims[2] = {static_cast<int64_t>(t.size()), 1};
      const std::string name = "shard" + std::to_string(s) + "/var" + std::to_string(v);
      CHECK(dtf_bundle_writer_add(w, name.c_str(), 1, 2, dims, t.data(), t.size() * 4) == 0);
      payload.push_back(std::move(t));
    }
    if (s == 0) {  // a scalar int64 like global_step
      const int64_t step = 1000;
      CHECK(dtf_bundle_writer_add(w, "global_step", 9, 0, nullptr, &step, 8) == 0);
    }
    CHECK(dtf_bundle_writer_finish(w) == 0);
  }
  std::vector<const char*> cparts;
  for (auto& p : parts) cparts.push_back(p.c_str());
  const std::string merged = dir + "/model.ckpt-1000";
  CHECK(dtf_bundle_merge(cparts.data(), kShards, merged.c_str(), 1) == 0);

  dtf_bundle_reader* r = dtf_bundle_reader_open(merged.c_str());
  CHECK(r != nullptr);
  if (!r) return;
  CHECK(dtf_bundle_reader_num_shards(r) == kShards);
  CHECK(dtf_bundle_reader_num_entries(r) == kShards * 4 + 1);
  for (int i = 0; i < dtf_bundle_reader_num_entries(r); ++i) {
    char name[256];
    int dtype, ndim, shard;
    int64_t dims[DTF_BUNDLE_MAX_DIMS], nbytes;
    CHECK(dtf_bundle_reader_entry(r, i, name, sizeof name, &dtype, &ndim, dims, &nbytes, &shard) == 0);
    std::vector<char> buf(static_cast<size_t>(nbytes));
    CHECK(dtf_bundle_reader_read(r, name, buf.data(), buf.size(), /*verify_crc=*/1) == 0);
    if (std::strcmp(name, "global_step") == 0) {
      int64_t step;
      std::memcpy(&step, buf.data(), 8);
      CHECK(dtype == 9 && ndim == 0 && step == 1000);
    } else {
      int s = -1, v = -1;
      CHECK(std::sscanf(name, "shard%d/var%d", &s, &v) == 2);
      CHECK(shard == s && dtype == 1 && ndim == 2);
      const std::vector<float>& want = payload[static_cast<size_t>(s * 4 + v)];
      CHECK(static_cast<size_t>(nbytes) == want.size() * 4);
      CHECK(std::memcmp(buf.data(), want.data(), buf.size()) == 0);
    }
  }
  // wrong size and unknown name are errors, not overruns
  char small[4];
  CHECK(dtf_bundle_reader_read(r, "shard0/var0", small, sizeof small, 1) != 0);
  CHECK(dtf_bundle_reader_read(r, "no/such/tensor", small, sizeof small, 1) != 0);
  CHECK(dtf_bundle_reader_close(r) == 0);
  CHECK(dtf_bundle_reader_open((dir + "/does_not_exist").c_str()) == nullptr);
}

int main(int argc, char** argv) {
  if (argc < 2) {
    std::fprintf(stderr, "usage: %s <scratch dir>\n", argv[0]);
    return 2;
  }
  const std::string dir = argv[1];
  test_crc();
  test_events(dir);
  test_table(dir);
  test_bundle(dir);
  if (g_fail == 0) std::printf("selftest: all checks passed\n");
  return g_fail == 0 ? 0 : 1;
}
