#include "sstable.h"

#include <algorithm>

#include "coding.h"

namespace ttdio {

// ------------------------------------------------------------------------------------------------
// block builder
// ------------------------------------------------------------------------------------------------
void BlockBuilder::reset() {
  buf_.clear();
  last_key_.clear();
  restarts_.clear();
  restarts_.push_back(0);
  counter_ = 0;
}

void BlockBuilder::add(const std::string& key, const std::string& value) {
  size_t shared = 0;
  if (counter_ < interval_) {
    const size_t lim = std::min(last_key_.size(), key.size());
    while (shared < lim && last_key_[shared] == key[shared]) ++shared;
  } else {
    restarts_.push_back(static_cast<uint32_t>(buf_.size()));
    counter_ = 0;
  }
  put_varint32(&buf_, static_cast<uint32_t>(shared));
  put_varint32(&buf_, static_cast<uint32_t>(key.size() - shared));
  put_varint32(&buf_, static_cast<uint32_t>(value.size()));
  buf_.append(key.data() + shared, key.size() - shared);
  buf_.append(value);
  last_key_ = key;
  ++counter_;
}

std::string BlockBuilder::finish() {
  std::string out = buf_;
  for (uint32_t r : restarts_) put_fixed32(&out, r);
  put_fixed32(&out, static_cast<uint32_t>(restarts_.size()));
  return out;
}

// ------------------------------------------------------------------------------------------------
// table builder
// ------------------------------------------------------------------------------------------------
static void shortest_separator(std::string* start, const std::string& limit) {
  // LevelDB BytewiseComparator::FindShortestSeparator
  const size_t min_len = std::min(start->size(), limit.size());
  size_t diff = 0;
  while (diff < min_len && (*start)[diff] == limit[diff]) ++diff;
  if (diff >= min_len) return;  // one is a prefix of the other
  const uint8_t b = static_cast<uint8_t>((*start)[diff]);
  if (b < 0xff && b + 1 < static_cast<uint8_t>(limit[diff])) {
    (*start)[diff] = static_cast<char>(b + 1);
    start->resize(diff + 1);
  }
}

static void short_successor(std::string* key) {
  for (size_t i = 0; i < key->size(); ++i) {
    const uint8_t b = static_cast<uint8_t>((*key)[i]);
    if (b != 0xff) {
      (*key)[i] = static_cast<char>(b + 1);
      key->resize(i + 1);
      return;
    }
  }
}

static std::string encode_handle(const std::pair<uint64_t, uint64_t>& h) {
  std::string s;
  put_varint64(&s, h.first);
  put_varint64(&s, h.second);
  return s;
}

std::pair<uint64_t, uint64_t> TableBuilder::write_block(const std::string& contents) {
  const std::pair<uint64_t, uint64_t> handle(file_.size(), contents.size());
  file_.append(contents);
  const char type = 0;  // kNoCompression
  uint32_t crc = crc32c(contents.data(), contents.size());
  crc = crc32c_extend(crc, &type, 1);
  file_.push_back(type);
  put_fixed32(&file_, crc_mask(crc));
  return handle;
}

void TableBuilder::flush_data_block() {
  if (data_.empty()) return;
  pending_handle_ = write_block(data_.finish());
  data_.reset();
  pending_index_ = true;
}

void TableBuilder::add(const std::string& key, const std::string& value) {
  if (pending_index_) {
    shortest_separator(&last_key_, key);
    index_.add(last_key_, encode_handle(pending_handle_));
    pending_index_ = false;
  }
  last_key_ = key;
  data_.add(key, value);
  if (data_.size_estimate() >= block_size_) flush_data_block();
}

std::string TableBuilder::finish() {
  flush_data_block();
  BlockBuilder meta(16);
  const auto meta_handle = write_block(meta.finish());
  if (pending_index_) {
    short_successor(&last_key_);
    index_.add(last_key_, encode_handle(pending_handle_));
    pending_index_ = false;
  }
  const auto index_handle = write_block(index_.finish());
  std::string footer = encode_handle(meta_handle) + encode_handle(index_handle);
  footer.resize(40, '\0');
  put_fixed64(&footer, kTableMagic);
  file_.append(footer);
  return file_;
}

// ------------------------------------------------------------------------------------------------
// reader
// ------------------------------------------------------------------------------------------------
static bool read_block(const std::string& file, uint64_t off, uint64_t size, std::string* contents) {
  // (written so that a corrupt handle with off/size near 2^64 cannot wrap the sum around)
  if (file.size() < kBlockTrailerSize || size > file.size() - kBlockTrailerSize ||
      off > file.size() - kBlockTrailerSize - size) {
    set_error("sstable: block handle points outside the file");
    return false;
  }
  const char* p = file.data() + off;
  const char type = p[size];
  const uint32_t stored = crc_unmask(get_fixed32(p + size + 1));
  const uint32_t actual = crc32c(p, size + 1);
  if (stored != actual) {
    set_error("sstable: block checksum mismatch");
    return false;
  }
  if (type == 0) {
    contents->assign(p, size);
    return true;
  }
  if (type == 1) {
    if (!snappy_uncompress(p, size, contents)) {
      set_error("sstable: corrupt snappy block");
      return false;
    }
    return true;
  }
  set_error("sstable: unknown block compression type");
  return false;
}

static bool parse_block(const std::string& block, std::vector<std::pair<std::string, std::string>>* out) {
  if (block.size() < 4) {
    set_error("sstable: block too small");
    return false;
  }
  const uint32_t n_restarts = get_fixed32(block.data() + block.size() - 4);
  if (static_cast<uint64_t>(n_restarts) * 4 + 4 > block.size()) {
    set_error("sstable: bad restart count");
    return false;
  }
  const char* p = block.data();
  const char* limit = block.data() + block.size() - 4 - 4 * static_cast<size_t>(n_restarts);
  std::string key;
  while (p < limit) {
    uint32_t shared, non_shared, vlen;
    p = get_varint32(p, limit, &shared);
    if (p) p = get_varint32(p, limit, &non_shared);
    if (p) p = get_varint32(p, limit, &vlen);
    if (p == nullptr || shared > key.size() || p + non_shared + vlen > limit) {
      set_error("sstable: corrupt block entry");
      return false;
    }
    key.resize(shared);
    key.append(p, non_shared);
    p += non_shared;
    out->emplace_back(key, std::string(p, vlen));
    p += vlen;
  }
  return true;
}

bool read_table(const std::string& file, std::vector<std::pair<std::string, std::string>>* entries) {
  if (file.size() < kFooterSize) {
    set_error("sstable: file shorter than a footer");
    return false;
  }
  const char* footer = file.data() + file.size() - kFooterSize;
  if (get_fixed64(footer + 40) != kTableMagic) {
    set_error("sstable: bad magic number (not a TF checkpoint index?)");
    return false;
  }
  uint64_t moff, msize, ioff, isize;
  const char* p = footer;
  const char* lim = footer + 40;
  p = get_varint64(p, lim, &moff);
  if (p) p = get_varint64(p, lim, &msize);
  if (p) p = get_varint64(p, lim, &ioff);
  if (p) p = get_varint64(p, lim, &isize);
  if (p == nullptr) {
    set_error("sstable: corrupt footer");
    return false;
  }
  std::string index;
  if (!read_block(file, ioff, isize, &index)) return false;
  std::vector<std::pair<std::string, std::string>> handles;
  if (!parse_block(index, &handles)) return false;
  for (const auto& h : handles) {
    uint64_t off, size;
    const char* q = h.second.data();
    const char* ql = q + h.second.size();
    q = get_varint64(q, ql, &off);
    if (q) q = get_varint64(q, ql, &size);
    if (q == nullptr) {
      set_error("sstable: corrupt block handle");
      return false;
    }
    std::string block;
    if (!read_block(file, off, size, &block)) return false;
    if (!parse_block(block, entries)) return false;
  }
  return true;
}

}  // namespace ttdio
