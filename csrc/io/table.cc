// LevelDB-format table builder + reader (see table.h). Blocks are written uncompressed
// (type byte 0), which is what TensorFlow's BundleWriter selects for checkpoint indices.
#include "table.h"

#include <string.h>

#include "../../include/dtf_b200_io.h"
#include "io_common.h"

namespace dtfio {

void BlockBuilder::Reset() {
  buf_.clear();
  restarts_.clear();
  restarts_.push_back(0);
  counter_ = 0;
  last_key_.clear();
}

void BlockBuilder::Add(const std::string& key, const std::string& value) {
  size_t shared = 0;
  if (counter_ < interval_) {
    const size_t m = std::min(last_key_.size(), key.size());
    while (shared < m && last_key_[shared] == key[shared]) ++shared;
  } else {
    restarts_.push_back(static_cast<uint32_t>(buf_.size()));
    counter_ = 0;
  }
  const size_t non_shared = key.size() - shared;
  put_varint(&buf_, shared);
  put_varint(&buf_, non_shared);
  put_varint(&buf_, value.size());
  buf_.append(key.data() + shared, non_shared);
  buf_.append(value);
  last_key_ = key;
  ++counter_;
}

std::string BlockBuilder::Finish() {
  std::string out = buf_;
  for (uint32_t r : restarts_) put_fixed32(&out, r);
  put_fixed32(&out, static_cast<uint32_t>(restarts_.size()));
  return out;
}

static void encode_handle(std::string* s, uint64_t off, uint64_t size) {
  put_varint(s, off);
  put_varint(s, size);
}

bool TableBuilder::WriteBlock(const std::string& contents, uint64_t* off, uint64_t* size) {
  *off = offset_;
  *size = contents.size();
  uint8_t trailer[kBlockTrailerSize];
  trailer[0] = 0;  // kNoCompression
  uint32_t crc = dtf_crc32c_extend(dtf_crc32c(contents.data(), contents.size()), trailer, 1);
  crc = dtf_crc32c_mask(crc);
  memcpy(trailer + 1, &crc, 4);
  if (fwrite(contents.data(), 1, contents.size(), f_) != contents.size()) return false;
  if (fwrite(trailer, 1, kBlockTrailerSize, f_) != kBlockTrailerSize) return false;
  offset_ += contents.size() + kBlockTrailerSize;
  return true;
}

bool TableBuilder::Flush() {
  if (data_.empty()) return true;
  const std::string contents = data_.Finish();
  if (!WriteBlock(contents, &pending_off_, &pending_size_)) return false;
  data_.Reset();
  pending_index_ = true;
  return true;
}

bool TableBuilder::Add(const std::string& key, const std::string& value) {
  if (has_keys_ && !(last_key_ < key)) {
    set_error("table keys must be strictly increasing");
    return false;
  }
  if (pending_index_) {
    // index key: any string >= last key of the finished block and < this key; the last key
    // itself is the simplest valid separator
    std::string h;
    encode_handle(&h, pending_off_, pending_size_);
    index_.Add(last_key_, h);
    pending_index_ = false;
  }
  last_key_ = key;
  has_keys_ = true;
  data_.Add(key, value);
  if (static_cast<int>(data_.CurrentSizeEstimate()) >= block_size_) return Flush();
  return true;
}

bool TableBuilder::Finish() {
  if (!Flush()) return false;
  // metaindex block (no filter, no meta entries)
  BlockBuilder meta(1);
  uint64_t meta_off, meta_size, idx_off, idx_size;
  if (!WriteBlock(meta.Finish(), &meta_off, &meta_size)) return false;
  if (pending_index_) {
    std::string h;
    encode_handle(&h, pending_off_, pending_size_);
    index_.Add(last_key_, h);
    pending_index_ = false;
  }
  if (!WriteBlock(index_.Finish(), &idx_off, &idx_size)) return false;
  std::string footer;
  encode_handle(&footer, meta_off, meta_size);
  encode_handle(&footer, idx_off, idx_size);
  footer.resize(kFooterSize - 8, 0);
  put_fixed64(&footer, kTableMagic);
  if (fwrite(footer.data(), 1, footer.size(), f_) != footer.size()) return false;
  offset_ += footer.size();
  return fflush(f_) == 0;
}

// ------------------------------------------------------------------------------------------------
// reader
// ------------------------------------------------------------------------------------------------
static bool read_block(const std::string& file, uint64_t off, uint64_t size, std::string* out) {
  if (off + size + kBlockTrailerSize > file.size()) {
    set_error("block handle out of range");
    return false;
  }
  const uint8_t* p = reinterpret_cast<const uint8_t*>(file.data()) + off;
  const uint8_t type = p[size];
  uint32_t stored;
  memcpy(&stored, p + size + 1, 4);
  const uint32_t actual = dtf_crc32c_mask(dtf_crc32c(p, size + 1));
  if (stored != actual) {
    set_error("block checksum mismatch");
    return false;
  }
  if (type != 0) {
    set_error("compressed table blocks (type %d) are not supported", static_cast<int>(type));
    return false;
  }
  out->assign(reinterpret_cast<const char*>(p), size);
  return true;
}

static bool parse_block(const std::string& block,
                        std::vector<std::pair<std::string, std::string>>* out) {
  if (block.size() < 4) {
    set_error("block too small");
    return false;
  }
  uint32_t num_restarts;
  memcpy(&num_restarts, block.data() + block.size() - 4, 4);
  const size_t restart_bytes = static_cast<size_t>(num_restarts) * 4 + 4;
  if (restart_bytes > block.size()) {
    set_error("bad restart array");
    return false;
  }
  const uint8_t* p = reinterpret_cast<const uint8_t*>(block.data());
  const uint8_t* end = p + block.size() - restart_bytes;
  std::string key;
  while (p < end) {
    uint64_t shared, non_shared, vlen;
    if (!get_varint(p, end, &shared) || !get_varint(p, end, &non_shared) ||
        !get_varint(p, end, &vlen) || shared > key.size() ||
        static_cast<uint64_t>(end - p) < non_shared + vlen) {
      set_error("corrupted block entry");
      return false;
    }
    key.resize(shared);
    key.append(reinterpret_cast<const char*>(p), non_shared);
    p += non_shared;
    out->emplace_back(key, std::string(reinterpret_cast<const char*>(p), vlen));
    p += vlen;
  }
  return true;
}

bool ReadTable(const std::string& path, std::vector<std::pair<std::string, std::string>>* out) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) {
    set_error("cannot open %s", path.c_str());
    return false;
  }
  std::string file;
  char buf[1 << 16];
  size_t n;
  while ((n = fread(buf, 1, sizeof(buf), f)) > 0) file.append(buf, n);
  fclose(f);
  if (file.size() < static_cast<size_t>(kFooterSize)) {
    set_error("%s is too short to be a table", path.c_str());
    return false;
  }
  const uint8_t* fp = reinterpret_cast<const uint8_t*>(file.data()) + file.size() - kFooterSize;
  uint64_t magic;
  memcpy(&magic, fp + kFooterSize - 8, 8);
  if (magic != kTableMagic) {
    set_error("%s: bad table magic", path.c_str());
    return false;
  }
  const uint8_t* p = fp;
  const uint8_t* end = fp + kFooterSize - 8;
  uint64_t meta_off, meta_size, idx_off, idx_size;
  if (!get_varint(p, end, &meta_off) || !get_varint(p, end, &meta_size) ||
      !get_varint(p, end, &idx_off) || !get_varint(p, end, &idx_size)) {
    set_error("bad footer");
    return false;
  }
  std::string index_block;
  if (!read_block(file, idx_off, idx_size, &index_block)) return false;
  std::vector<std::pair<std::string, std::string>> index_entries;
  if (!parse_block(index_block, &index_entries)) return false;
  for (auto& ie : index_entries) {
    const uint8_t* hp = reinterpret_cast<const uint8_t*>(ie.second.data());
    const uint8_t* hend = hp + ie.second.size();
    uint64_t off, size;
    if (!get_varint(hp, hend, &off) || !get_varint(hp, hend, &size)) {
      set_error("bad block handle");
      return false;
    }
    std::string block;
    if (!read_block(file, off, size, &block)) return false;
    if (!parse_block(block, out)) return false;
  }
  return true;
}

}  // namespace dtfio

using namespace dtfio;

struct dtf_table_builder {
  FILE* f;
  TableBuilder* b;
};

extern "C" dtf_table_builder* dtf_table_builder_open(const char* path, int block_size) {
  FILE* f = fopen(path, "wb");
  if (!f) {
    set_error("cannot open %s", path);
    return nullptr;
  }
  auto* t = new dtf_table_builder();
  t->f = f;
  t->b = new TableBuilder(f, block_size > 0 ? block_size : 262144);
  return t;
}

extern "C" int dtf_table_builder_add(dtf_table_builder* t, const void* key, size_t klen,
                                     const void* value, size_t vlen) {
  return t->b->Add(std::string(static_cast<const char*>(key), klen),
                   std::string(static_cast<const char*>(value), vlen))
             ? 0
             : -1;
}

extern "C" int dtf_table_builder_finish(dtf_table_builder* t) {
  const bool ok = t->b->Finish();
  fclose(t->f);
  delete t->b;
  delete t;
  if (!ok && !g_err[0]) set_error("table write failed");
  return ok ? 0 : -1;
}
