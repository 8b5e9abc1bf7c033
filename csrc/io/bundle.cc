// TF1 "V2 checkpoint" (tensor bundle) writer / reader — SURVEY A.4.
//
// What the reference gets from MonitoredTrainingSession(checkpoint_dir=..., save_checkpoint_secs=60)
// (distribute_training.py:209-215) is TensorFlow's Saver -> SaveV2 -> BundleWriter -> table::TableBuilder
// chain, with one data shard per ps device and MergeV2Checkpoints joining the per-shard indices. This file
// is that chain without TensorFlow or protobuf: the two protos involved are tiny and are encoded by hand.
//
//   <prefix>.data-0000k-of-0000N   raw little-endian tensor bytes of shard k, concatenated in key order
//   <prefix>.index                 SSTable: "" -> BundleHeaderProto, <tensor name> -> BundleEntryProto
//
// Writing is two-phase like TF's: every shard owner calls write_shard (data file + a partial index under
// <prefix>_temp/), then one process calls merge to produce <prefix>.index.
#include <errno.h>
#include <stdio.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <map>
#include <string>
#include <vector>

#include "../../include/ttd_io.h"
#include "coding.h"
#include "sstable.h"

namespace ttdio {

struct Entry {
  int32_t dtype = 0;
  std::vector<int64_t> shape;
  int32_t shard_id = 0;
  int64_t offset = 0;
  int64_t size = 0;
  uint32_t crc = 0;  // masked crc32c of the tensor bytes
};

// ---- protobuf wire helpers -----------------------------------------------------------------------
static void put_tag(std::string* s, int field, int wire) { put_varint32(s, static_cast<uint32_t>((field << 3) | wire)); }

static std::string encode_header(int num_shards) {
  // BundleHeaderProto { int32 num_shards = 1; Endianness endianness = 2 (LITTLE = 0, omitted); VersionDef version = 3 }
  std::string v;  // VersionDef { int32 producer = 1 }
  put_tag(&v, 1, 0);
  put_varint32(&v, 1);
  std::string s;
  put_tag(&s, 1, 0);
  put_varint32(&s, static_cast<uint32_t>(num_shards));
  put_tag(&s, 3, 2);
  put_varint32(&s, static_cast<uint32_t>(v.size()));
  s += v;
  return s;
}

static std::string encode_entry(const Entry& e) {
  // BundleEntryProto { dtype=1; TensorShapeProto shape=2; shard_id=3; offset=4; size=5; fixed32 crc32c=6 }
  std::string shape;  // TensorShapeProto { repeated Dim dim = 2 { int64 size = 1 } }
  for (int64_t d : e.shape) {
    std::string dim;
    if (d != 0) {
      put_tag(&dim, 1, 0);
      put_varint64(&dim, static_cast<uint64_t>(d));
    }
    put_tag(&shape, 2, 2);
    put_varint32(&shape, static_cast<uint32_t>(dim.size()));
    shape += dim;
  }
  std::string s;
  if (e.dtype != 0) {
    put_tag(&s, 1, 0);
    put_varint32(&s, static_cast<uint32_t>(e.dtype));
  }
  put_tag(&s, 2, 2);  // mutable_shape() marks the field present even for scalars
  put_varint32(&s, static_cast<uint32_t>(shape.size()));
  s += shape;
  if (e.shard_id != 0) {
    put_tag(&s, 3, 0);
    put_varint32(&s, static_cast<uint32_t>(e.shard_id));
  }
  if (e.offset != 0) {
    put_tag(&s, 4, 0);
    put_varint64(&s, static_cast<uint64_t>(e.offset));
  }
  if (e.size != 0) {
    put_tag(&s, 5, 0);
    put_varint64(&s, static_cast<uint64_t>(e.size));
  }
  if (e.crc != 0) {
    put_tag(&s, 6, 5);
    put_fixed32(&s, e.crc);
  }
  return s;
}

// generic field walker: calls f(field, wire, varint value, bytes ptr, bytes len)
template <class F>
static bool walk_proto(const char* p, const char* end, F f) {
  while (p < end) {
    uint32_t tag;
    p = get_varint32(p, end, &tag);
    if (!p) return false;
    const int field = static_cast<int>(tag >> 3), wire = static_cast<int>(tag & 7);
    if (wire == 0) {
      uint64_t v;
      p = get_varint64(p, end, &v);
      if (!p) return false;
      f(field, wire, v, nullptr, 0);
    } else if (wire == 2) {
      uint32_t len;
      p = get_varint32(p, end, &len);
      if (!p || p + len > end) return false;
      f(field, wire, 0, p, len);
      p += len;
    } else if (wire == 5) {
      if (p + 4 > end) return false;
      f(field, wire, get_fixed32(p), nullptr, 0);
      p += 4;
    } else if (wire == 1) {
      if (p + 8 > end) return false;
      f(field, wire, get_fixed64(p), nullptr, 0);
      p += 8;
    } else {
      return false;
    }
  }
  return true;
}

static bool decode_entry(const std::string& s, Entry* e) {
  return walk_proto(s.data(), s.data() + s.size(), [&](int field, int, uint64_t v, const char* b, size_t n) {
    switch (field) {
      case 1: e->dtype = static_cast<int32_t>(v); break;
      case 2:
        walk_proto(b, b + n, [&](int f2, int w2, uint64_t, const char* b2, size_t n2) {
          if (f2 == 2 && w2 == 2) {
            int64_t size = 0;
            walk_proto(b2, b2 + n2, [&](int f3, int, uint64_t v3, const char*, size_t) {
              if (f3 == 1) size = static_cast<int64_t>(v3);
            });
            e->shape.push_back(size);
          }
        });
        break;
      case 3: e->shard_id = static_cast<int32_t>(v); break;
      case 4: e->offset = static_cast<int64_t>(v); break;
      case 5: e->size = static_cast<int64_t>(v); break;
      case 6: e->crc = static_cast<uint32_t>(v); break;
      default: break;
    }
  });
}

static bool decode_header(const std::string& s, int* num_shards, int* endianness) {
  *num_shards = 0;
  *endianness = 0;
  return walk_proto(s.data(), s.data() + s.size(), [&](int field, int, uint64_t v, const char*, size_t) {
    if (field == 1) *num_shards = static_cast<int>(v);
    if (field == 2) *endianness = static_cast<int>(v);
  });
}

// ---- file helpers ------------------------------------------------------------------------------------
static bool write_file_atomic(const std::string& path, const std::string& data) {
  const std::string tmp = path + ".tempstate" + std::to_string(static_cast<long long>(getpid()));
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) {
    set_error("cannot open " + tmp + " for writing: " + strerror(errno));
    return false;
  }
  const bool ok = fwrite(data.data(), 1, data.size(), f) == data.size();
  const bool closed = fclose(f) == 0;
  if (!ok || !closed) {
    set_error("short write to " + tmp);
    unlink(tmp.c_str());
    return false;
  }
  if (rename(tmp.c_str(), path.c_str()) != 0) {
    set_error("cannot rename " + tmp + " to " + path + ": " + strerror(errno));
    return false;
  }
  return true;
}

static bool read_file(const std::string& path, std::string* out) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) {
    set_error("cannot open " + path + ": " + strerror(errno));
    return false;
  }
  out->clear();
  char buf[1 << 16];
  size_t n;
  while ((n = fread(buf, 1, sizeof(buf), f)) > 0) out->append(buf, n);
  fclose(f);
  return true;
}

static std::string shard_name(const std::string& prefix, int k, int n) {
  char buf[64];
  snprintf(buf, sizeof(buf), ".data-%05d-of-%05d", k, n);
  return prefix + buf;
}
static std::string part_name(const std::string& prefix, int k, int n) {
  char buf[64];
  snprintf(buf, sizeof(buf), "_temp/part-%05d-of-%05d.index", k, n);
  return prefix + buf;
}

static std::string build_index(int num_shards, const std::map<std::string, Entry>& entries) {
  TableBuilder tb;
  tb.add("", encode_header(num_shards));
  for (const auto& kv : entries) tb.add(kv.first, encode_entry(kv.second));
  return tb.finish();
}

static bool parse_index(const std::string& file, int* num_shards, std::map<std::string, Entry>* entries) {
  std::vector<std::pair<std::string, std::string>> kv;
  if (!read_table(file, &kv)) return false;
  bool saw_header = false;
  for (const auto& e : kv) {
    if (e.first.empty()) {
      int endian = 0;
      if (!decode_header(e.second, num_shards, &endian)) {
        set_error("bundle: corrupt header proto");
        return false;
      }
      if (endian != 0) {
        set_error("bundle: big-endian bundles are not supported");
        return false;
      }
      saw_header = true;
    } else {
      Entry en;
      if (!decode_entry(e.second, &en)) {
        set_error("bundle: corrupt entry proto for " + e.first);
        return false;
      }
      (*entries)[e.first] = en;
    }
  }
  if (!saw_header) {
    set_error("bundle: index has no header entry");
    return false;
  }
  return true;
}

struct Reader {
  std::string prefix;
  int num_shards = 0;
  std::map<std::string, Entry> entries;
  std::vector<std::string> names;
};

}  // namespace ttdio

using namespace ttdio;

// ============================================ C ABI ============================================

extern "C" int ttdio_bundle_write_shard(const char* prefix, int shard_id, int num_shards, int n_tensors,
                                        const char* const* names, const int32_t* dtypes, const int32_t* ndims,
                                        const int64_t* const* shapes, const void* const* data, const int64_t* nbytes) {
  if (num_shards < 1 || shard_id < 0 || shard_id >= num_shards) {
    set_error("bundle: bad shard id " + std::to_string(shard_id) + " of " + std::to_string(num_shards));
    return -1;
  }
  const std::string pfx(prefix);
  std::vector<int> order(n_tensors);
  for (int i = 0; i < n_tensors; ++i) order[i] = i;
  std::sort(order.begin(), order.end(), [&](int a, int b) { return strcmp(names[a], names[b]) < 0; });
  std::map<std::string, Entry> entries;
  std::string blob;
  for (int i : order) {
    Entry e;
    e.dtype = dtypes[i];
    e.shape.assign(shapes[i], shapes[i] + ndims[i]);
    e.shard_id = shard_id;
    e.offset = static_cast<int64_t>(blob.size());
    e.size = nbytes[i];
    e.crc = crc_mask(crc32c(data[i], static_cast<size_t>(nbytes[i])));
    if (nbytes[i] > 0) blob.append(static_cast<const char*>(data[i]), static_cast<size_t>(nbytes[i]));
    if (!entries.emplace(names[i], e).second) {
      set_error(std::string("bundle: duplicate tensor name ") + names[i]);
      return -1;
    }
  }
  if (!write_file_atomic(shard_name(pfx, shard_id, num_shards), blob)) return -2;
  const std::string temp_dir = pfx + "_temp";
  if (mkdir(temp_dir.c_str(), 0755) != 0 && errno != EEXIST) {
    set_error("cannot create " + temp_dir + ": " + strerror(errno));
    return -2;
  }
  if (!write_file_atomic(part_name(pfx, shard_id, num_shards), build_index(1, entries))) return -2;
  return 0;
}

extern "C" int ttdio_bundle_merge(const char* prefix, int num_shards) {
  const std::string pfx(prefix);
  std::map<std::string, Entry> all;
  for (int k = 0; k < num_shards; ++k) {
    std::string file;
    if (!read_file(part_name(pfx, k, num_shards), &file)) return -2;
    int ns = 0;
    std::map<std::string, Entry> part;
    if (!parse_index(file, &ns, &part)) return -3;
    for (auto& kv : part) {
      if (!all.emplace(kv.first, kv.second).second) {
        set_error("bundle: tensor " + kv.first + " appears in more than one shard");
        return -3;
      }
    }
  }
  if (!write_file_atomic(pfx + ".index", build_index(num_shards, all))) return -2;
  for (int k = 0; k < num_shards; ++k) unlink(part_name(pfx, k, num_shards).c_str());
  rmdir((pfx + "_temp").c_str());
  return 0;
}

extern "C" void* ttdio_bundle_open(const char* prefix) {
  Reader* r = new Reader;
  r->prefix = prefix;
  std::string file;
  if (!read_file(r->prefix + ".index", &file) || !parse_index(file, &r->num_shards, &r->entries)) {
    delete r;
    return nullptr;
  }
  for (const auto& kv : r->entries) r->names.push_back(kv.first);
  return r;
}

extern "C" void ttdio_bundle_close(void* h) { delete static_cast<Reader*>(h); }
extern "C" int ttdio_bundle_num_entries(void* h) { return static_cast<int>(static_cast<Reader*>(h)->names.size()); }
extern "C" int ttdio_bundle_num_shards(void* h) { return static_cast<Reader*>(h)->num_shards; }

extern "C" int ttdio_bundle_entry(void* h, int i, const char** name, int32_t* dtype, int32_t* ndim, int64_t* shape8,
                                  int32_t* shard_id, int64_t* offset, int64_t* size, uint32_t* crc_masked) {
  Reader* r = static_cast<Reader*>(h);
  if (i < 0 || i >= static_cast<int>(r->names.size())) {
    set_error("bundle: entry index out of range");
    return -1;
  }
  const Entry& e = r->entries[r->names[i]];
  if (e.shape.size() > 8) {
    set_error("bundle: tensors of rank > 8 are not supported");
    return -1;
  }
  *name = r->names[i].c_str();
  *dtype = e.dtype;
  *ndim = static_cast<int32_t>(e.shape.size());
  for (size_t d = 0; d < e.shape.size(); ++d) shape8[d] = e.shape[d];
  *shard_id = e.shard_id;
  *offset = e.offset;
  *size = e.size;
  *crc_masked = e.crc;
  return 0;
}

extern "C" int ttdio_bundle_read(void* h, const char* name, void* out, int64_t nbytes) {
  Reader* r = static_cast<Reader*>(h);
  auto it = r->entries.find(name);
  if (it == r->entries.end()) {
    set_error(std::string("bundle: no tensor named ") + name);
    return -1;
  }
  const Entry& e = it->second;
  if (nbytes != e.size) {
    set_error(std::string("bundle: ") + name + " holds " + std::to_string(e.size) + " bytes, caller expects " +
              std::to_string(nbytes));
    return -1;
  }
  const std::string path = shard_name(r->prefix, e.shard_id, r->num_shards);
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) {
    set_error("cannot open " + path + ": " + strerror(errno));
    return -2;
  }
  bool ok = fseek(f, static_cast<long>(e.offset), SEEK_SET) == 0 &&
            fread(out, 1, static_cast<size_t>(e.size), f) == static_cast<size_t>(e.size);
  fclose(f);
  if (!ok) {
    set_error("short read from " + path);
    return -2;
  }
  if (crc_mask(crc32c(out, static_cast<size_t>(e.size))) != e.crc) {
    set_error(std::string("bundle: checksum mismatch for tensor ") + name);
    return -3;
  }
  return 0;
}
