// Immutable sorted string table in the LevelDB / TensorFlow `lib/io/table` format:
//   [data block]... [metaindex block] [index block] [48-byte footer]
// Blocks are prefix-compressed with restart points and carry a 5-byte trailer (type + masked crc32c).
#pragma once
#include <stdint.h>

#include <string>
#include <utility>
#include <vector>

namespace ttdio {

constexpr uint64_t kTableMagic = 0xdb4775248b80fb57ull;
constexpr size_t kFooterSize = 48;
constexpr size_t kBlockTrailerSize = 5;

class BlockBuilder {
 public:
  explicit BlockBuilder(int restart_interval) : interval_(restart_interval) { reset(); }
  void reset();
  void add(const std::string& key, const std::string& value);  // keys must arrive in sorted order
  std::string finish();                                        // appends the restart array
  size_t size_estimate() const { return buf_.size() + restarts_.size() * 4 + 4; }
  bool empty() const { return buf_.empty(); }

 private:
  int interval_;
  int counter_ = 0;
  std::string buf_, last_key_;
  std::vector<uint32_t> restarts_;
};

class TableBuilder {
 public:
  explicit TableBuilder(size_t block_size = 262144) : block_size_(block_size), data_(16), index_(1) {}
  void add(const std::string& key, const std::string& value);
  std::string finish();  // returns the whole file image

 private:
  void flush_data_block();
  std::pair<uint64_t, uint64_t> write_block(const std::string& contents);
  size_t block_size_;
  BlockBuilder data_, index_;
  std::string file_, last_key_;
  bool pending_index_ = false;
  std::pair<uint64_t, uint64_t> pending_handle_{0, 0};
};

// Parses a whole table image; verifies every block checksum; accepts snappy-compressed blocks.
bool read_table(const std::string& file, std::vector<std::pair<std::string, std::string>>* entries);

}  // namespace ttdio
