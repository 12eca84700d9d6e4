// Host map collectives on (String key, primitive value) maps: allreduceMap / reduceMap of the reference
// (PCS:2053-2081, 1490-1516: hash-partition by String.hashCode() % p, ring reduce-scatter with
// recv-and-merge DOP:225-257, allgather / gather, union), natively.
//
// The binding hands over the map as one key blob (NUL-separated UTF-8) plus one value array; every ring
// step ships a partition as [u64 n][u32 key lengths][key bytes][values] over the byte transport and the
// receiver merges it into a hash table of string views - no per-entry (de)serialisation objects and no
// boxing, which is where the reference (Kryo + HashMap<String, Double>) and a pure-Python ring spend
// their time.
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

// local = local (+) incoming: op(local value, incoming value) on common keys, insert otherwise (DOP:225-257)
void merge(Part* local, const std::string& incoming_buf, int dtype, int op, size_t es) {
  PartView in = view(incoming_buf, es);
  std::unordered_map<std::string_view, size_t> index;
  index.reserve((size_t)(local->n() + in.n) * 2);
  // string_views into local->keys would dangle when it grows: index the local keys by offset first,
  // append new keys into a side buffer and splice once at the end
  {
    size_t off = 0;
    for (size_t i = 0; i < local->n(); ++i) {
      index.emplace(std::string_view(local->keys.data() + off, local->lens[i]), i);
      off += local->lens[i];
    }
  }
  std::string new_keys;
  std::vector<uint32_t> new_lens;
  std::vector<char> new_vals;
  // (an incoming message is a map: its keys are distinct, so new keys need no index of their own)
  size_t off = 0;
  for (uint64_t i = 0; i < in.n; ++i) {
    std::string_view k(in.keys + off, in.lens[i]);
    off += in.lens[i];
    auto it = index.find(k);
    if (it != index.end()) {
      host_reduce_into(dtype, op, local->vals.data() + it->second * es, in.vals + i * es, 1);
    } else {
      new_lens.push_back(in.lens[i]);
      new_keys.append(k.data(), k.size());
      new_vals.insert(new_vals.end(), in.vals + i * es, in.vals + (i + 1) * es);
    }
  }
  local->lens.insert(local->lens.end(), new_lens.begin(), new_lens.end());
  local->keys.append(new_keys);
  local->vals.insert(local->vals.end(), new_vals.begin(), new_vals.end());
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
    // ---- ring reduce-scatter: step s sends partition (rank-1-s), merges what arrives into (rank-2-s)
    for (int s = 0; s < p - 1; ++s) {
      const int sseg = ((rank_ - 1 - s) % p + p) % p, rseg = ((rank_ - 2 - s) % p + p) % p;
      const std::string out = pack(parts[(size_t)sseg]);
      const std::string in = sendrecv_bytes(next, out.data(), out.size(), prev);
      merge(&parts[(size_t)rseg], in, dtype, op, es);
      parts[(size_t)sseg] = Part();  // consumed
    }
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
    const int next = (rank_ + 1) % p, prev = (rank_ - 1 + p) % p;
    for (int s = 0; s < p - 1; ++s) {
      const int sseg = ((rank_ - 1 - s) % p + p) % p, rseg = ((rank_ - 2 - s) % p + p) % p;
      const std::string out = pack(parts[(size_t)sseg]);
      const std::string in = sendrecv_bytes(next, out.data(), out.size(), prev);
      merge(&parts[(size_t)rseg], in, dtype, op, es);
      parts[(size_t)sseg] = Part();
    }
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
