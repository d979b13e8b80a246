// TF1 "tensor bundle" checkpoints (Saver V2): writer, reader and MergeV2Checkpoints (SURVEY T12,
// K14). On disk:
//   <prefix>.index                      LevelDB-format table: "" -> BundleHeaderProto,
//                                       <tensor name> -> BundleEntryProto
//   <prefix>.data-0000i-of-0000N        raw little-endian tensor bytes, back to back
// BundleHeaderProto { num_shards = 1 (int32); endianness = 2 (LITTLE = 0);
//                     version = 3 { producer = 1 (int32) } }
// BundleEntryProto  { dtype = 1; shape = 2 { dim = 2 { size = 1 } }; shard_id = 3; offset = 4;
//                     size = 5; crc32c = 6 (fixed32, masked crc of the bytes); slices = 7 }
// The reference reaches this code through MonitoredTrainingSession(checkpoint_dir=...)
// (templates/00_between_graph_replication_async_mnist.py:43-45) and tf.train.Supervisor(logdir)
// (templates/00_mnist_replica.py:196-211). With the default sharded saver every PS task writes
// its own data shard and the chief merges the indices; dtf_bundle_merge is that merge.
#include <errno.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/dtf_b200_io.h"
#include "io_common.h"
#include "table.h"

using namespace dtfio;

namespace {

struct Entry {
  int dtype = 0;
  std::vector<int64_t> dims;
  int shard_id = 0;
  int64_t offset = 0;
  int64_t size = 0;
  uint32_t crc = 0;
};

std::string data_filename(const std::string& prefix, int shard, int num_shards) {
  char buf[64];
  snprintf(buf, sizeof(buf), ".data-%05d-of-%05d", shard, num_shards);
  return prefix + buf;
}

std::string encode_header(int num_shards) {
  std::string version;
  put_varint_field(&version, 1, 1);  // producer = kTensorBundleVersion
  std::string h;
  put_varint_field(&h, 1, static_cast<uint64_t>(num_shards));
  put_bytes(&h, 3, version);
  return h;
}

std::string encode_entry(const Entry& e) {
  std::string shape;
  for (int64_t d : e.dims) {
    std::string dim;
    put_varint_field(&dim, 1, static_cast<uint64_t>(d));
    put_bytes(&shape, 2, dim);
  }
  std::string out;
  put_varint_field(&out, 1, static_cast<uint64_t>(e.dtype));
  put_bytes(&out, 2, shape);
  if (e.shard_id) put_varint_field(&out, 3, static_cast<uint64_t>(e.shard_id));
  if (e.offset) put_varint_field(&out, 4, static_cast<uint64_t>(e.offset));
  put_varint_field(&out, 5, static_cast<uint64_t>(e.size));
  put_fixed32_field(&out, 6, e.crc);
  return out;
}

bool skip_field(const uint8_t*& p, const uint8_t* end, int wt) {
  uint64_t v;
  switch (wt) {
    case 0: return get_varint(p, end, &v);
    case 1: if (end - p < 8) return false; p += 8; return true;
    case 2: if (!get_varint(p, end, &v) || static_cast<uint64_t>(end - p) < v) return false; p += v; return true;
    case 5: if (end - p < 4) return false; p += 4; return true;
    default: return false;
  }
}

bool decode_shape(const uint8_t* p, const uint8_t* end, std::vector<int64_t>* dims) {
  while (p < end) {
    uint64_t key;
    if (!get_varint(p, end, &key)) return false;
    const int field = static_cast<int>(key >> 3), wt = static_cast<int>(key & 7);
    if (field == 2 && wt == 2) {
      uint64_t len;
      if (!get_varint(p, end, &len) || static_cast<uint64_t>(end - p) < len) return false;
      const uint8_t* q = p;
      const uint8_t* qend = p + len;
      int64_t size = 0;
      while (q < qend) {
        uint64_t k2;
        if (!get_varint(q, qend, &k2)) return false;
        if ((k2 >> 3) == 1 && (k2 & 7) == 0) {
          uint64_t v;
          if (!get_varint(q, qend, &v)) return false;
          size = static_cast<int64_t>(v);
        } else if (!skip_field(q, qend, static_cast<int>(k2 & 7))) {
          return false;
        }
      }
      dims->push_back(size);
      p += len;
    } else if (!skip_field(p, end, wt)) {
      return false;
    }
  }
  return true;
}

bool decode_entry(const std::string& s, Entry* e) {
  const uint8_t* p = reinterpret_cast<const uint8_t*>(s.data());
  const uint8_t* end = p + s.size();
  while (p < end) {
    uint64_t key, v;
    if (!get_varint(p, end, &key)) return false;
    const int field = static_cast<int>(key >> 3), wt = static_cast<int>(key & 7);
    if (wt == 0 && field >= 1 && field <= 5) {
      if (!get_varint(p, end, &v)) return false;
      if (field == 1) e->dtype = static_cast<int>(v);
      if (field == 3) e->shard_id = static_cast<int>(v);
      if (field == 4) e->offset = static_cast<int64_t>(v);
      if (field == 5) e->size = static_cast<int64_t>(v);
    } else if (field == 2 && wt == 2) {
      if (!get_varint(p, end, &v) || static_cast<uint64_t>(end - p) < v) return false;
      if (!decode_shape(p, p + v, &e->dims)) return false;
      p += v;
    } else if (field == 6 && wt == 5) {
      if (end - p < 4) return false;
      memcpy(&e->crc, p, 4);
      p += 4;
    } else if (!skip_field(p, end, wt)) {
      return false;
    }
  }
  return true;
}

bool decode_header(const std::string& s, int* num_shards) {
  const uint8_t* p = reinterpret_cast<const uint8_t*>(s.data());
  const uint8_t* end = p + s.size();
  *num_shards = 0;
  while (p < end) {
    uint64_t key, v;
    if (!get_varint(p, end, &key)) return false;
    if ((key >> 3) == 1 && (key & 7) == 0) {
      if (!get_varint(p, end, &v)) return false;
      *num_shards = static_cast<int>(v);
    } else if ((key >> 3) == 2 && (key & 7) == 0) {
      if (!get_varint(p, end, &v)) return false;
      if (v != 0) {
        set_error("big-endian bundles are not supported");
        return false;
      }
    } else if (!skip_field(p, end, static_cast<int>(key & 7))) {
      return false;
    }
  }
  return true;
}

bool write_index(const std::string& prefix, int num_shards, const std::map<std::string, Entry>& entries) {
  const std::string path = prefix + ".index";
  const std::string tmp = path + ".tempstate";
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) {
    set_error("cannot open %s", tmp.c_str());
    return false;
  }
  TableBuilder tb(f, 262144);
  bool ok = tb.Add("", encode_header(num_shards));
  for (auto& kv : entries) ok = ok && tb.Add(kv.first, encode_entry(kv.second));
  ok = ok && tb.Finish();
  fclose(f);
  if (!ok) return false;
  if (rename(tmp.c_str(), path.c_str()) != 0) {
    set_error("rename %s failed: %s", tmp.c_str(), strerror(errno));
    return false;
  }
  return true;
}

}  // namespace

struct dtf_bundle_writer {
  std::string prefix;
  FILE* data = nullptr;
  int64_t offset = 0;
  std::map<std::string, Entry> entries;
};

extern "C" dtf_bundle_writer* dtf_bundle_writer_open(const char* prefix) {
  auto* w = new dtf_bundle_writer();
  w->prefix = prefix;
  const std::string path = data_filename(w->prefix, 0, 1);
  w->data = fopen(path.c_str(), "wb");
  if (!w->data) {
    set_error("cannot open %s", path.c_str());
    delete w;
    return nullptr;
  }
  return w;
}

extern "C" int dtf_bundle_writer_add(dtf_bundle_writer* w, const char* name, int dtype, int ndim,
                                     const int64_t* dims, const void* data, size_t nbytes) {
  if (w->entries.count(name)) {
    set_error("tensor %s added twice", name);
    return -1;
  }
  Entry e;
  e.dtype = dtype;
  e.dims.assign(dims, dims + ndim);
  e.shard_id = 0;
  e.offset = w->offset;
  e.size = static_cast<int64_t>(nbytes);
  e.crc = dtf_crc32c_mask(dtf_crc32c(data, nbytes));
  if (nbytes && fwrite(data, 1, nbytes, w->data) != nbytes) {
    set_error("short write to data shard");
    return -1;
  }
  w->offset += static_cast<int64_t>(nbytes);
  w->entries[name] = e;
  return 0;
}

extern "C" int dtf_bundle_writer_finish(dtf_bundle_writer* w) {
  int rc = 0;
  if (fclose(w->data) != 0) rc = -1;
  if (rc == 0 && !write_index(w->prefix, 1, w->entries)) rc = -1;
  delete w;
  return rc;
}

// Index-only writer: the data shards were written elsewhere (every PS task pwrite()s the bytes it
// owns straight into the final data files, in parallel); the chief only records where each tensor
// lives. crc32c is the UNMASKED checksum of the whole tensor (see dtf_crc32c_combine).
struct dtf_bundle_index_writer {
  std::string prefix;
  int num_shards = 1;
  std::map<std::string, Entry> entries;
};

extern "C" dtf_bundle_index_writer* dtf_bundle_index_open(const char* prefix, int num_shards) {
  if (num_shards < 1) {
    set_error("bundle index: %d data shards", num_shards);
    return nullptr;
  }
  auto* w = new dtf_bundle_index_writer();
  w->prefix = prefix;
  w->num_shards = num_shards;
  return w;
}

extern "C" int dtf_bundle_index_add(dtf_bundle_index_writer* w, const char* name, int dtype, int ndim,
                                    const int64_t* dims, int shard_id, int64_t offset, int64_t size,
                                    uint32_t crc32c) {
  if (w->entries.count(name)) {
    set_error("tensor %s added twice", name);
    return -1;
  }
  if (shard_id < 0 || shard_id >= w->num_shards || offset < 0 || size < 0) {
    set_error("tensor %s: shard %d / offset %lld / size %lld out of range", name, shard_id,
              static_cast<long long>(offset), static_cast<long long>(size));
    return -1;
  }
  Entry e;
  e.dtype = dtype;
  e.dims.assign(dims, dims + ndim);
  e.shard_id = shard_id;
  e.offset = offset;
  e.size = size;
  e.crc = dtf_crc32c_mask(crc32c);
  w->entries[name] = e;
  return 0;
}

extern "C" int dtf_bundle_index_data_path(dtf_bundle_index_writer* w, int shard_id, char* buf, size_t cap) {
  const std::string p = data_filename(w->prefix, shard_id, w->num_shards);
  if (p.size() + 1 > cap) {
    set_error("data file name does not fit");
    return -1;
  }
  memcpy(buf, p.c_str(), p.size() + 1);
  return 0;
}

extern "C" int dtf_bundle_index_finish(dtf_bundle_index_writer* w) {
  const int rc = write_index(w->prefix, w->num_shards, w->entries) ? 0 : -1;
  delete w;
  return rc;
}

struct dtf_bundle_reader {
  std::string prefix;
  int num_shards = 0;
  std::vector<std::pair<std::string, Entry>> entries;
  std::map<std::string, size_t> by_name;
  std::vector<FILE*> files;
};

extern "C" dtf_bundle_reader* dtf_bundle_reader_open(const char* prefix) {
  std::vector<std::pair<std::string, std::string>> kv;
  if (!ReadTable(std::string(prefix) + ".index", &kv)) return nullptr;
  if (kv.empty() || !kv[0].first.empty()) {
    set_error("%s.index has no bundle header", prefix);
    return nullptr;
  }
  auto* r = new dtf_bundle_reader();
  r->prefix = prefix;
  if (!decode_header(kv[0].second, &r->num_shards) || r->num_shards < 1) {
    if (!g_err[0]) set_error("bad bundle header");
    delete r;
    return nullptr;
  }
  for (size_t i = 1; i < kv.size(); ++i) {
    Entry e;
    if (!decode_entry(kv[i].second, &e)) {
      set_error("bad bundle entry for %s", kv[i].first.c_str());
      delete r;
      return nullptr;
    }
    r->by_name[kv[i].first] = r->entries.size();
    r->entries.emplace_back(kv[i].first, e);
  }
  r->files.assign(r->num_shards, nullptr);
  return r;
}

extern "C" int dtf_bundle_reader_num_shards(dtf_bundle_reader* r) { return r->num_shards; }
extern "C" int dtf_bundle_reader_num_entries(dtf_bundle_reader* r) { return static_cast<int>(r->entries.size()); }

extern "C" int dtf_bundle_reader_entry(dtf_bundle_reader* r, int index, char* name_buf,
                                       size_t name_cap, int* dtype, int* ndim, int64_t* dims,
                                       int64_t* nbytes, int* shard_id) {
  if (index < 0 || index >= static_cast<int>(r->entries.size())) {
    set_error("entry index out of range");
    return -1;
  }
  const auto& kv = r->entries[index];
  if (kv.first.size() + 1 > name_cap || kv.second.dims.size() > DTF_BUNDLE_MAX_DIMS) {
    set_error("entry %s does not fit the caller's buffers", kv.first.c_str());
    return -1;
  }
  memcpy(name_buf, kv.first.c_str(), kv.first.size() + 1);
  *dtype = kv.second.dtype;
  *ndim = static_cast<int>(kv.second.dims.size());
  for (size_t i = 0; i < kv.second.dims.size(); ++i) dims[i] = kv.second.dims[i];
  *nbytes = kv.second.size;
  *shard_id = kv.second.shard_id;
  return 0;
}

extern "C" int dtf_bundle_reader_read(dtf_bundle_reader* r, const char* name, void* dst,
                                      size_t nbytes, int verify_crc) {
  auto it = r->by_name.find(name);
  if (it == r->by_name.end()) {
    set_error("tensor %s not found in checkpoint %s", name, r->prefix.c_str());
    return -1;
  }
  const Entry& e = r->entries[it->second].second;
  if (static_cast<int64_t>(nbytes) != e.size) {
    set_error("tensor %s has %lld bytes, caller asked for %zu", name, static_cast<long long>(e.size), nbytes);
    return -1;
  }
  if (e.shard_id < 0 || e.shard_id >= r->num_shards) {
    set_error("tensor %s refers to shard %d of %d", name, e.shard_id, r->num_shards);
    return -1;
  }
  FILE*& f = r->files[e.shard_id];
  if (!f) {
    const std::string path = data_filename(r->prefix, e.shard_id, r->num_shards);
    f = fopen(path.c_str(), "rb");
    if (!f) {
      set_error("cannot open %s", path.c_str());
      return -1;
    }
  }
  if (fseeko(f, e.offset, SEEK_SET) != 0 || (nbytes && fread(dst, 1, nbytes, f) != nbytes)) {
    set_error("short read of tensor %s", name);
    return -1;
  }
  if (verify_crc && dtf_crc32c_mask(dtf_crc32c(dst, nbytes)) != e.crc) {
    set_error("checksum mismatch for tensor %s", name);
    return -2;
  }
  return 0;
}

extern "C" int dtf_bundle_reader_close(dtf_bundle_reader* r) {
  if (!r) return 0;
  for (FILE* f : r->files)
    if (f) fclose(f);
  delete r;
  return 0;
}

extern "C" int dtf_bundle_merge(const char* const* part_prefixes, int n, const char* merged_prefix,
                                int delete_old_dirs) {
  std::map<std::string, Entry> merged;
  for (int i = 0; i < n; ++i) {
    std::vector<std::pair<std::string, std::string>> kv;
    if (!ReadTable(std::string(part_prefixes[i]) + ".index", &kv)) return -1;
    int shards = 0;
    if (kv.empty() || !kv[0].first.empty() || !decode_header(kv[0].second, &shards) || shards != 1) {
      set_error("%s is not a single-shard bundle", part_prefixes[i]);
      return -1;
    }
    for (size_t k = 1; k < kv.size(); ++k) {
      Entry e;
      if (!decode_entry(kv[k].second, &e)) {
        set_error("bad entry %s in %s", kv[k].first.c_str(), part_prefixes[i]);
        return -1;
      }
      if (merged.count(kv[k].first)) {
        set_error("tensor %s appears in more than one shard", kv[k].first.c_str());
        return -1;
      }
      e.shard_id = i;
      merged[kv[k].first] = e;
    }
  }
  for (int i = 0; i < n; ++i) {
    const std::string from = data_filename(part_prefixes[i], 0, 1);
    const std::string to = data_filename(merged_prefix, i, n);
    if (rename(from.c_str(), to.c_str()) != 0) {
      set_error("rename %s -> %s failed: %s", from.c_str(), to.c_str(), strerror(errno));
      return -1;
    }
    unlink((std::string(part_prefixes[i]) + ".index").c_str());
  }
  if (!write_index(merged_prefix, n, merged)) return -1;
  if (delete_old_dirs) {
    for (int i = 0; i < n; ++i) {
      std::string p = part_prefixes[i];
      const size_t slash = p.find_last_of('/');
      if (slash != std::string::npos) rmdir(p.substr(0, slash).c_str());
    }
  }
  return 0;
}
