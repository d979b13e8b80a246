// LevelDB-format sorted string table (the container of a TF1 checkpoint's ".index" file):
// prefix-compressed data blocks with restart points, per-block trailer (type byte + masked
// crc32c), metaindex block, index block, 48-byte footer with magic 0xdb4775248b80fb57.
// Equivalent of tensorflow/core/lib/io/{table_builder,block_builder,format,table}.cc.
#pragma once
#include <stdint.h>
#include <stdio.h>

#include <string>
#include <utility>
#include <vector>

namespace dtfio {

constexpr uint64_t kTableMagic = 0xdb4775248b80fb57ull;
constexpr int kFooterSize = 48;
constexpr int kBlockTrailerSize = 5;

class BlockBuilder {
 public:
  explicit BlockBuilder(int restart_interval) : interval_(restart_interval) { Reset(); }
  void Reset();
  void Add(const std::string& key, const std::string& value);
  std::string Finish();
  size_t CurrentSizeEstimate() const { return buf_.size() + restarts_.size() * 4 + 4; }
  bool empty() const { return buf_.empty(); }

 private:
  int interval_;
  std::string buf_;
  std::vector<uint32_t> restarts_;
  int counter_ = 0;
  std::string last_key_;
};

class TableBuilder {
 public:
  TableBuilder(FILE* f, int block_size) : f_(f), block_size_(block_size), data_(16), index_(1) {}
  bool Add(const std::string& key, const std::string& value);  // keys strictly increasing
  bool Finish();
  uint64_t FileSize() const { return offset_; }

 private:
  bool Flush();
  bool WriteBlock(const std::string& contents, uint64_t* off, uint64_t* size);
  FILE* f_;
  int block_size_;
  BlockBuilder data_, index_;
  uint64_t offset_ = 0;
  std::string last_key_;
  bool pending_index_ = false;
  uint64_t pending_off_ = 0, pending_size_ = 0;
  bool has_keys_ = false;
};

// Reads a whole table into memory and returns its entries in key order.
bool ReadTable(const std::string& path, std::vector<std::pair<std::string, std::string>>* out);

}  // namespace dtfio
