// Host map collectives on (String key, primitive value) maps: allreduceMap / reduceMap / reduceScatterMap of
// the reference (PCS:2053-2081, 1490-1516, 1299-1327: hash-partition by String.hashCode() % p, reduce-scatter
// with recv-and-merge DOP:225-257, allgather / gather, union), natively.
//
// The binding hands over the map as one key blob (NUL-separated UTF-8) plus one value array. The reduce-scatter
// is a direct exchange over the persistent mesh: at step s every rank sends rank (r + s)'s partition to it and
// receives its own partition from rank (r - s) - p - 1 full-duplex steps like the reference's ring, but every
// message is an original partition (the ring forwards partial results that grow on the way) and all merging for
// a key happens in ONE open-addressing table on its owner (64-bit hash + entry index per slot, keys compared only
// on a hash match). A message is [u64 n][u32 key lengths][key bytes][values]: no per-entry (de)serialisation
// objects and no boxing, which is where the reference (Kryo + HashMap<String, Double>) and a pure-Python ring
// spend their time.
#include <string.h>

#include <string>
#include <string_view>
#include <unordered_map>
#include <vector>

#include "../../include/mp4j.h"
#include "process_comm.h"

namespace mp4j {

namespace {

struct Part {
  std::vector<uint32_t> lens;
  std::string keys;         // concatenated key bytes
  std::vector<char> vals;   // n * es bytes
  size_t n() const { return lens.size(); }
};

std::string pack(const Part& p) {
  std::string out;
  const uint64_t n = p.n();
  out.reserve(8 + n * 4 + p.keys.size() + p.vals.size());
  out.append((const char*)&n, 8);
  out.append((const char*)p.lens.data(), n * 4);
  out.append(p.keys);
  out.append(p.vals.data(), p.vals.size());
  return out;
}

// views into `buf`, which must outlive them
struct PartView {
  uint64_t n = 0;
  const uint32_t* lens = nullptr;
  const char* keys = nullptr;
  const char* vals = nullptr;
};

PartView view(const std::string& buf, size_t es) {
  PartView v;
  if (buf.size() < 8) throw Error("map collective: short message");
  memcpy(&v.n, buf.data(), 8);
  v.lens = (const uint32_t*)(buf.data() + 8);
  size_t kb = 0;
  if (buf.size() < 8 + v.n * 4) throw Error("map collective: truncated message");
  for (uint64_t i = 0; i < v.n; ++i) kb += v.lens[i];
  if (buf.size() != 8 + v.n * 4 + kb + v.n * es) throw Error("map collective: malformed message");
  v.keys = buf.data() + 8 + v.n * 4;
  v.vals = v.keys + kb;
  return v;
}

inline uint64_t hash_bytes(const char* s, size_t n) {
  uint64_t h = 0x9e3779b97f4a7c15ull ^ (n * 0xff51afd7ed558ccdull);
  while (n >= 8) {
    uint64_t w;
    memcpy(&w, s, 8);
    h = (h ^ w) * 0xff51afd7ed558ccdull;
    h ^= h >> 32;
    s += 8;
    n -= 8;
  }
  if (n) {
    uint64_t w = 0;
    memcpy(&w, s, n);
    h = (h ^ w) * 0xc4ceb9fe1a85ec53ull;
    h ^= h >> 29;
  }
  h *= 0x9e3779b97f4a7c15ull;
  return h ^ (h >> 32);
}

// Index of the keys of ONE Part: open addressing; a slot holds the key's 64-bit hash, where its bytes start, its
// length and its entry number, so that a probe touches one slot, then (on a hash match) the key bytes and the value.
class KeyTable {
 public:
  explicit KeyTable(Part* part, size_t expect) : part_(part) {
    size_t cap = 1024;
    while (cap < 2 * expect) cap <<= 1;
    slots_.assign(cap, Slot{});
    size_t off = 0;
    for (size_t i = 0; i < part->n(); ++i) {
      place(Slot{hash_bytes(part->keys.data() + off, part->lens[i]), off, (uint32_t)i + 1, part->lens[i]});
      off += part->lens[i];
    }
  }
  // part (+)= incoming message: op(local value, incoming value) on common keys, insert otherwise (DOP:225-257)
  void merge(const std::string& incoming_buf, int dtype, int op, size_t es) {
    const PartView in = view(incoming_buf, es);
    Part& pt = *part_;
    // hashes first, so that the probe loop can prefetch the home slot of the key 16 places ahead
    std::vector<uint64_t> hs((size_t)in.n);
    {
      size_t off = 0;
      for (uint64_t i = 0; i < in.n; ++i) {
        hs[(size_t)i] = hash_bytes(in.keys + off, in.lens[i]);
        off += in.lens[i];
      }
    }
    constexpr uint64_t kAhead = 16;
    size_t off = 0;
    for (uint64_t i = 0; i < in.n; ++i) {
      if (i + kAhead < in.n) __builtin_prefetch(&slots_[hs[(size_t)(i + kAhead)] & (slots_.size() - 1)]);
      const char* k = in.keys + off;
      const uint32_t len = in.lens[i];
      off += len;
      const uint64_t h = hs[(size_t)i];
      size_t slot = h & (slots_.size() - 1);
      bool found = false;
      while (slots_[slot].ent) {
        const Slot& sl = slots_[slot];
        if (sl.hash == h && sl.len == len && memcmp(pt.keys.data() + sl.off, k, len) == 0) {
          host_reduce_into(dtype, op, pt.vals.data() + (size_t)(sl.ent - 1) * es, in.vals + i * es, 1);
          found = true;
          break;
        }
        slot = (slot + 1) & (slots_.size() - 1);
      }
      if (found) continue;
      if (pt.n() >= 0xfffffffeull) throw Error("map collective: more than 2^32 keys in one partition");
      slots_[slot] = Slot{h, pt.keys.size(), (uint32_t)pt.n() + 1, len};
      pt.lens.push_back(len);
      pt.keys.append(k, len);
      pt.vals.insert(pt.vals.end(), in.vals + i * es, in.vals + (i + 1) * es);
      if (2 * pt.n() > slots_.size()) grow();
    }
  }

 private:
  struct Slot {
    uint64_t hash = 0;
    uint64_t off = 0;   // byte offset of the key in part_->keys
    uint32_t ent = 0;   // entry number + 1; 0 = free
    uint32_t len = 0;
  };
  void place(const Slot& s) {
    size_t slot = s.hash & (slots_.size() - 1);
    while (slots_[slot].ent) slot = (slot + 1) & (slots_.size() - 1);
    slots_[slot] = s;
  }
  void grow() {  // rehash from the stored hashes: no key bytes are touched
    std::vector<Slot> old;
    old.swap(slots_);
    slots_.assign(old.size() * 2, Slot{});
    for (const Slot& s : old)
      if (s.ent) place(s);
  }
  Part* part_;
  std::vector<Slot> slots_;
};

// The reduce-scatter step of all three collectives: parts[q] goes to rank q, parts[rank] absorbs what the others
// hold for this rank. p - 1 full-duplex exchanges over the mesh.
void exchange_and_merge(ProcessComm* pc, std::vector<Part>* parts, int rank, int p, int dtype, int op, size_t es) {
  Part& mine = (*parts)[(size_t)rank];
  KeyTable table(&mine, mine.n() * 2 + 16);
  for (int s = 1; s < p; ++s) {
    const int dst = (rank + s) % p, src = (rank - s + p) % p;
    const std::string out = pack((*parts)[(size_t)dst]);
    (*parts)[(size_t)dst] = Part();  // consumed
    const std::string in = pc->sendrecv_bytes(dst, out.data(), out.size(), src);
    table.merge(in, dtype, op, es);
  }
}

void append(Part* dst, const std::string& buf, size_t es) {
  PartView in = view(buf, es);
  size_t kb = 0;
  for (uint64_t i = 0; i < in.n; ++i) kb += in.lens[i];
  dst->lens.insert(dst->lens.end(), in.lens, in.lens + in.n);
  dst->keys.append(in.keys, kb);
  dst->vals.insert(dst->vals.end(), in.vals, in.vals + in.n * es);
}

}  // namespace

// mode 0: allreduceMap (every rank gets the merged map); mode 1: reduceMap (only `root` does, others get
// an empty map). Output: NUL-separated keys + value array, same layout as the input.
void ProcessComm::map_reduce(int mode, int root, const char* blob, size_t blob_len, const void* vals, size_t n,
                             int dtype, int op, std::string* out_blob, std::string* out_vals) {
  const size_t es = dtype_size(dtype);
  if (es == 0 || !op_valid_for(dtype, op) || op == MP4J_COPY) throw Error("invalid operator for dtype");
  if (mode == 1 && (root < 0 || root >= n_)) throw Error("root rank out of range");
  const int p = n_;
  // ---- hash-partition (String.hashCode() % p, negatives wrapped: PCS:2063-2071)
  std::vector<Part> parts((size_t)p);
  {
    const char* s = blob;
    const char* end = blob + blob_len;
    for (size_t i = 0; i < n; ++i) {
      const char* e = (const char*)memchr(s, 0, (size_t)(end - s));
      if (!e) e = end;
      const size_t len = (size_t)(e - s);
      int q = 0;
      if (p > 1) {
        const int32_t h = mp4j_java_string_hash(s, len);
        q = (int)(h % p);
        if (q < 0) q += p;
      }
      Part& pt = parts[(size_t)q];
      pt.lens.push_back((uint32_t)len);
      pt.keys.append(s, len);
      pt.vals.insert(pt.vals.end(), (const char*)vals + i * es, (const char*)vals + (i + 1) * es);
      s = e < end ? e + 1 : end;
    }
  }
  Part result;
  if (p == 1) {
    result = std::move(parts[0]);
  } else {
    std::lock_guard<std::mutex> lk(op_mu_);
    const int next = (rank_ + 1) % p, prev = (rank_ - 1 + p) % p;
    exchange_and_merge(this, &parts, rank_, p, dtype, op, es);
    Part& mine = parts[(size_t)rank_];
    if (mode == 0) {
      // ---- ring allgather of the owned partitions (disjoint key sets: plain concatenation)
      std::vector<std::string> blocks((size_t)p);
      blocks[(size_t)rank_] = pack(mine);
      for (int s = 0; s < p - 1; ++s) {
        const int sseg = ((rank_ - s) % p + p) % p, rseg = ((rank_ - 1 - s) % p + p) % p;
        blocks[(size_t)rseg] = sendrecv_bytes(next, blocks[(size_t)sseg].data(), blocks[(size_t)sseg].size(), prev);
      }
      for (int q = 0; q < p; ++q) append(&result, blocks[(size_t)q], es);
    } else if (rank_ == root) {
      result = std::move(mine);
      for (int q = 0; q < p; ++q)
        if (q != root) append(&result, recv_bytes(q), es);
    } else {
      const std::string out = pack(mine);
      send_bytes(root, out.data(), out.size());
    }
  }
  // ---- back to the binding's layout
  out_blob->clear();
  out_blob->reserve(result.keys.size() + result.n());
  size_t off = 0;
  for (size_t i = 0; i < result.n(); ++i) {
    if (i) out_blob->push_back('\0');
    out_blob->append(result.keys.data() + off, result.lens[i]);
    off += result.lens[i];
  }
  out_vals->assign(result.vals.data(), result.vals.size());
}

// reduceScatterMap (PCS:1299-1327): the caller addresses rank q with its q-th map. blob / vals hold the p maps back
// to back (counts[q] entries each); the ring reduce-scatter merges natively; the result is this rank's map.
void ProcessComm::map_reduce_scatter(const char* blob, size_t blob_len, const void* vals, const int64_t* counts,
                                     int dtype, int op, std::string* out_blob, std::string* out_vals) {
  const size_t es = dtype_size(dtype);
  if (es == 0 || !op_valid_for(dtype, op) || op == MP4J_COPY) throw Error("invalid operator for dtype");
  const int p = n_;
  std::vector<Part> parts((size_t)p);
  {
    const char* s = blob;
    const char* end = blob + blob_len;
    size_t i = 0;
    for (int q = 0; q < p; ++q) {
      Part& pt = parts[(size_t)q];
      for (int64_t k = 0; k < counts[q]; ++k, ++i) {
        const char* e = (const char*)memchr(s, 0, (size_t)(end - s));
        if (!e) e = end;
        const size_t len = (size_t)(e - s);
        pt.lens.push_back((uint32_t)len);
        pt.keys.append(s, len);
        pt.vals.insert(pt.vals.end(), (const char*)vals + i * es, (const char*)vals + (i + 1) * es);
        s = e < end ? e + 1 : end;
      }
    }
  }
  if (p > 1) {
    std::lock_guard<std::mutex> lk(op_mu_);
    exchange_and_merge(this, &parts, rank_, p, dtype, op, es);
  }
  const Part& result = parts[(size_t)rank_];
  out_blob->clear();
  out_blob->reserve(result.keys.size() + result.n());
  size_t off = 0;
  for (size_t i = 0; i < result.n(); ++i) {
    if (i) out_blob->push_back('\0');
    out_blob->append(result.keys.data() + off, result.lens[i]);
    off += result.lens[i];
  }
  out_vals->assign(result.vals.data(), result.vals.size());
}

}  // namespace mp4j
